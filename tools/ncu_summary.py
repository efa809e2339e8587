#!/usr/bin/env python
"""Summarise gpurun_out/prof_<capture>_raw.csv (ncu --page raw --csv) into profiles/<tag>_ncu_full_summary.md,
copy the raw pages, and refresh profiles/ncu_traffic.json (dram bytes per launch of each stage's k=11 convs)."""
import csv, json, shutil, sys

tag, mb = sys.argv[1], int(sys.argv[2])
names = {'s0k11': 'stage0', 's1k11': 'stage1', 's2k11': 'stage2', 's3k11': 'stage3', 'ups': 'ups1'}
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_sector_hit_rate.pct"]


def num(v, u):
    x = float(v.replace(',', ''))
    return x * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1, 'us': 1, 'ms': 1e3, 'ns': 1e-3}.get(u, 1)


lines = [f"# {tag}: `ncu --set full --clock-control none --import-source on` of conv_tc_kernel launches", "",
         f"Command: `tools/gpu_round.sh` (one warm-up forward, then one profiled forward of V1, {mb} utterances x 862 frames, fp16).",
         "Two launches per capture = (convs1[1] d=3, convs2[1] d=1 with residual) of the k=11 ResBlock1 of the stage; `ups` = ups[1] (256->128, stride 8).",
         f"Raw pages: `{tag}_ncu_full_<capture>_raw.csv`. Durations are cold-cache and serialised (compare shares, not absolutes).", "",
         "| capture | instantiation <T,C_IN,N,MB,KS,KBW,TPS,NST,NSL,NSTG,CG> | grid | us | DRAM rd MB | DRAM wr MB | DRAM % | tensor pipe % | smem->TC wavefronts % | LSU smem wavefronts % | L2->SM GB | L2 hit % |",
         "|---|---|---|---|---|---|---|---|---|---|---|---|"]
traffic = {}
for n, st in names.items():
    src = f'gpurun_out/prof_{n}_raw.csv'
    shutil.copy(src, f'profiles/{tag}_ncu_full_{n}_raw.csv')
    rows = list(csv.reader(open(src)))
    hdr, units = rows[0], rows[1]
    tot = []
    for r in rows[2:]:
        d, u = dict(zip(hdr, r)), dict(zip(hdr, units))
        g = lambda k: num(d[k], u[k])
        kn = d['Kernel Name']
        inst = kn[kn.index('<') + 1:kn.index('>(')].replace('(int)', '') if '>(' in kn else kn
        lines.append(f"| {n} | {inst} | {d['Grid Size']} | {g(want[0]):.1f} | {g(want[1])/1e6:.1f} | {g(want[2])/1e6:.1f} | {g(want[3]):.1f} | {g(want[4]):.1f} | "
                     f"{g(want[5]):.1f} | {g(want[6]):.1f} | {g(want[7])/1e9:.2f} | {g(want[8]):.1f} |")
        tot.append(g(want[1]) + g(want[2]))
    traffic[st] = sum(tot) / len(tot)
open(f'profiles/{tag}_ncu_full_summary.md', 'w').write("\n".join(lines) + "\n")
json.dump({"v1_b64x10s": {k: {"dram_bytes_per_launch": v, "micro_batch": mb,
                              "source": f"profiles/{tag}_ncu_full_summary.md (mean of the two captured k=11 launches: convs1 without / convs2 with residual)"}
                          for k, v in traffic.items() if k.startswith('stage')}}, open('profiles/ncu_traffic.json', 'w'), indent=1)
print("\n".join(lines))
