#!/usr/bin/env python
"""bench.py -- HiFi-GAN V1 generator synthesis throughput (audio-seconds / second) on B200.

Contract (driver): `python bench.py --gpus N --steps K --warmup W [--impl reference]`; for N>1 it
is launched by torch.distributed.run with one rank per GPU.  One "step" = one pass of the hot path
(Generator.call) over one batch of synthetic mels.  Workload at any N: BASELINE.json configs[1]
(V1, batch 64 x 10 s utterances = 862 mel frames -> 220,672 samples each) PER GPU: utterances are
sharded across ranks with no data-path collective (weak scaling).

  value    : whole-job audio-s/s, mels already resident in HBM, CUDA-event timed on the launch stream
  e2e      : same metric through Generator.call on HOST (pinned) buffers: H2D of the mels and D2H of
             the waveforms inside the timed region
  roofline : the tcgen05 conv kernel family IS the step (>95 % of device time); achieved = algorithmic
             FLOPs of the forward (2*MAC, dense taps) / CUDA-event step time
  cpu_baseline / --impl reference : the oracle's torch-CPU fp32 restatement of the reference
             (TensorFlow is not installable here: "port"), on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SR = 22050
T_10S = 862            # ceil(10 * 22050 / 256)
CONFIGS = {            # name -> (config version, per-GPU batch, frames)
    "v1_b64x10s": ("1", 64, T_10S),
    "v1_b1x256": ("1", 1, 256),
    "v2_b256x10s": ("2", 256, T_10S),
    "v3_b64x10s": ("3", 64, T_10S),
    "v1_b512x10s": ("1", 512, T_10S),
}


emit = None


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "tflops_burst": p["bf16_tflops"], "tflops_sustained": p["bf16_tflops_sustained"],
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "tflops_burst": 1590.0, "tflops_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


def ncu_traffic(stage_label, workload, mb):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
    `ncu --set full` capture (profiles/ncu_traffic.json); None when no capture matches this workload."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(path):
        return None
    try:
        t = json.load(open(path)).get(workload, {}).get(stage_label)
        if t:      # the capture's launches process `micro_batch` utterances; traffic is linear in the utterance count
            return t["dram_bytes_per_launch"] * mb / t["micro_batch"]
    except Exception:
        pass
    return None


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw,power.limit")

    def __init__(self, index):
        self.index, self.samples, self.stop_flag, self.th = index, [], False, None

    def _run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop_flag = True
        self.th.join(timeout=6)

    def summary(self):
        sm = sorted(int(s[0]) for s in self.samples if s and s[0].isdigit())
        mx = [int(s[1]) for s in self.samples if len(s) > 1 and s[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples for n, v in zip(names, s[2:6]) if v.lower().startswith("active")})
        def fl(v):
            try:
                return float(v)
            except Exception:
                return None
        pw = sorted(p for p in (fl(s[6]) for s in self.samples if len(s) > 6) if p is not None)
        lim = [p for p in (fl(s[7]) for s in self.samples if len(s) > 7) if p is not None]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples),
                "power_w": pw[len(pw) // 2] if pw else None, "power_limit_w": max(lim) if lim else None}


def cpu_port_audio_s_per_s(version, B, T, repeats, threads=None):
    """Times the oracle's torch-CPU fp32 restatement (kind 'port')."""
    import numpy as np
    import torch
    from neuraltexttospeech_b200.hparams import HParams
    from neuraltexttospeech_b200.synthetic import synthetic_mel, synthetic_weights
    from oracle import hifigan_oracle as O
    # all the host threads this process may use (torchrun exports OMP_NUM_THREADS=1; override it at run time)
    torch.set_num_threads(threads or len(os.sched_getaffinity(0)))
    h = HParams.builtin(version)
    w = synthetic_weights(h, "A")
    mel = synthetic_mel(B, T, h.num_mels)
    O.generator_forward(h, w, mel[:, : min(T, 32)], "torch32")          # warm the thread pool
    ts = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        out = O.generator_forward(h, w, mel, "torch32")
        ts.append(time.perf_counter() - t0)
    secs = out.shape[0] * out.shape[1] / SR
    return secs / min(ts), secs / (sum(ts) / len(ts)), torch.get_num_threads(), ts


def metric_name(version):
    return f"HiFi-GAN V{version} audio-seconds/sec"


def tf_pin():
    """when a box carries tensorflow + tensorflow_addons and a checkout of the reference's HiFiGAN_TF, run the
    REAL reference on a tiny case and record how far the oracle is from it; otherwise say why not (oracle/tf_probe.py)."""
    try:
        from oracle import tf_probe, hifigan_oracle as O
        from neuraltexttospeech_b200.hparams import HParams
        from neuraltexttospeech_b200.synthetic import synthetic_mel, synthetic_weights
        ok, why = tf_probe.available()
        if not ok:
            return {"kind": "port", "tf": "unavailable: " + why}
        h = HParams.builtin("1")
        w = synthetic_weights(h, "B")
        mel = synthetic_mel(1, 16, h.num_mels)
        return tf_probe.pin_oracle(h, w, mel, O.generator_forward(h, w, mel, "torch32"))
    except Exception as e:                      # a probe must never take the bench down
        return {"kind": "port", "tf": f"probe failed: {type(e).__name__}: {e}"}


def run_reference(args, rank, world):
    if rank != 0:
        return
    version, B, T = CONFIGS[args.config]
    # Bounded sample of the workload: `sb` utterances of the batch per step (same frames, same weights), the largest
    # power of two (<= 16, <= B) whose timed steps fit ~60 s.  The CPU port's rate is nearly flat in the batch size
    # (profiles/r2_cpu_port_batch_effect.json: audio-s/s at B = 1, 4, 16), so the rate of the sample stands for the batch.
    best1, _, threads, ts1 = cpu_port_audio_s_per_s(version, 1, T, 1)
    sb = 1
    while sb * 2 <= min(B, 16) and ts1[0] * (sb * 2) * (args.steps + 1) <= 60.0:
        sb *= 2
    best, mean, threads, ts = cpu_port_audio_s_per_s(version, sb, T, max(1, args.warmup // 3) + args.steps)
    ts = ts[-args.steps:]
    ms = 1000 * sum(ts) / len(ts)
    val = (sb * T * 256 / SR) / (sum(ts) / len(ts))
    pin = tf_pin()
    line = {
        "impl": "reference", "metric": metric_name(version), "value": val, "unit": "audio-s/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": {"workload": args.config, "generator": f"config_v{version}.json", "per_gpu_batch": B, "frames": T,
                   "sample_per_step": f"{sb} utterance(s) x {T} frames ({sb}/{B} of the batch)",
                   "one_utterance_rate": (T * 256 / SR) / ts1[0]},
        "cpu_baseline": {"value": val, "unit": "audio-s/s", "cores": threads, "kind": "port",
                         "sample": f"{sb} utterance(s) x {T} frames per step; torch-CPU fp32 restatement of HiFiGAN_TF/model.py "
                                   "(TensorFlow/tfa not installable on this image)", "tf_reference": pin},
        "e2e": {"value": val, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="v1_b64x10s", choices=sorted(CONFIGS))
    ap.add_argument("--precision", default="fp16", choices=["fp32", "fp16", "bf16", "tf32"])
    ap.add_argument("--micro-batch", type=int, default=0, help="utterances per forward; 0 = the library's own rule")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-adjacent", action="store_true", help="skip the PCM16 / mel front-end side measurements")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    # stdout carries exactly ONE JSON line: anything libraries print there (NCCL's version banner ...) goes to stderr
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    global emit
    emit = lambda line: (real_stdout.write(json.dumps(line) + "\n"), real_stdout.flush())

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from neuraltexttospeech_b200 import Generator, HParams
    from neuraltexttospeech_b200.synthetic import synthetic_mel, synthetic_weights
    from neuraltexttospeech_b200.topology import boundary_bytes_per_frame, count_macs_per_frame, per_stage_work

    version, B, T = CONFIGS[args.config]
    strong = args.config == "v1_b512x10s"      # BASELINE configs[4]: 512 utterances in total, sharded 512/N per GPU
    if strong:
        B = B // world
    h = HParams.builtin(version)
    gen = Generator(h, precision=args.precision, device=local_rank).load_weights(synthetic_weights(h, "A"))
    mel_np = synthetic_mel(B, T, h.num_mels, seed=1234 + rank)          # each rank: its own shard of utterances
    L = T * 256
    audio_s_per_step = B * L / SR

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- device-resident measurement ("value")
    mel_dev = torch.from_numpy(mel_np).cuda()
    out_dev = torch.empty((B, L, 1), dtype=torch.float32, device="cuda")
    # utterances per forward call (hfg_auto_micro_batch: one activation buffer <= 1 GiB)
    mb = max(1, min(args.micro_batch, B)) if args.micro_batch > 0 else gen.auto_micro_batch(B, T)
    stream = torch.cuda.current_stream()

    n_forwards = (B + mb - 1) // mb
    mean_fwd_batch = B / n_forwards        # stage stamps are averaged over forwards: a ragged tail (512 = 6 x 76 + 56) counts

    def step_device():
        n = 0
        for b0 in range(0, B, mb):
            gen._call_device(mel_dev[b0:b0 + mb], out_dev[b0:b0 + mb])
            n += gen.last_launch_count()
        return n

    for _ in range(args.warmup):
        launches_per_step = step_device()
    barrier()
    # stage stamps (7 cudaEventRecord per forward on the launch stream) stay on during the timed region so that
    # the dominant kernel's launch time is measured live, not in a separate pass
    gen.set_option("stage_timing", 1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clk:
        e0.record(stream)
        for _ in range(args.steps):
            step_device()
        e1.record(stream)
        barrier()
    dev_ms = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    clocks = clk.summary()
    stage_ms = gen.stage_times_ms()          # per forward of `mb` utterances, averaged over the timed forwards
    stage_launches = gen.stage_launches()    # kernels per stage as launched (a fused ResBlock iteration is ONE launch)
    gen.set_option("stage_timing", 0)

    # ---------------- end-to-end through Generator.call on pinned host buffers ("e2e")
    e2e = None
    adjacent = {}
    if not args.no_e2e:
        # The call a user makes: `audio = generator(mel)` on an ordinary (pageable) numpy array, result allocated by the
        # call.  Inside: the mels are staged to pinned memory and uploaded per micro-batch, the waveforms land in a
        # page-locked numpy array from the generator's pool (recycled when the previous result is dropped).
        mel_user = np.array(mel_np, copy=True)                               # plain numpy, not pinned
        gen.set_option("micro_batch", args.micro_batch if args.micro_batch > 0 else 0)   # 0: the library's own split (copies overlap compute)
        mb_host = args.micro_batch if args.micro_batch > 0 else gen.auto_micro_batch(B, T, host_path=True)
        audio = None
        for _ in range(3):
            audio = gen(mel_user)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            audio = gen(mel_user)                 # H2D + forward + D2H + sync, every step; `audio` is read back on the host
            checksum = float(audio[0, 0, 0]) + float(audio[-1, -1, 0])
        barrier()
        e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1000) / args.steps
        assert audio.shape == (B, L, 1) and audio.dtype == np.float32 and np.isfinite(checksum)
        e2e = {"value": world * audio_s_per_step / (e2e_ms / 1000), "unit": "audio-s/s",
               "h2d_bytes_per_step": int(mel_np.nbytes), "d2h_bytes_per_step": int(audio.nbytes),
               "ms_per_step": e2e_ms, "micro_batch": mb_host,
               "api": "audio = Generator.__call__(mel: pageable numpy [B,T,80]) -> numpy [B,L,1] allocated by the call (hfg_forward_host)"}
        del audio
        # same public call with the 16-bit PCM epilogue: half the D2H bytes
        for _ in range(2):
            pcm = gen.synthesize_pcm16(mel_user)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pcm = gen.synthesize_pcm16(mel_user)
        barrier()
        pcm_ms = max_over_ranks((time.perf_counter() - t0) * 1000) / args.steps
        adjacent["e2e_pcm16"] = {"value": world * audio_s_per_step / (pcm_ms / 1000), "unit": "audio-s/s", "ms_per_step": pcm_ms,
                                 "h2d_bytes_per_step": int(mel_np.nbytes), "d2h_bytes_per_step": int(pcm.nbytes),
                                 "api": "Generator.synthesize_pcm16(numpy) -> hfg_forward_host_pcm16"}
        del pcm
    # mel front-end kernel on the waveforms this step produced: HBM-bound, bytes = wav in + mel out
    if not args.no_adjacent:
        from neuraltexttospeech_b200 import STFT
        fe = STFT(device=local_rank)
        wav = out_dev.view(B, L)
        Tm = fe.num_frames(L)
        for _ in range(3):
            fe.mel_spectrogram(wav)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        f0.record(stream)
        for _ in range(10):
            fe.mel_spectrogram(wav)
        f1.record(stream)
        torch.cuda.synchronize()
        fe_ms = f0.elapsed_time(f1) / 10
        fe_bytes = B * L * 4 + B * Tm * 80 * 4
        adjacent["mel_frontend"] = {"ms": fe_ms, "utterances": B, "frames": Tm, "audio_s_per_s": audio_s_per_step / (fe_ms / 1000),
                                    "bound": "hbm", "achieved_gbs": fe_bytes / (fe_ms * 1e-3) / 1e9, "peak_gbs": peaks()["hbm_gbs"],
                                    "frac": fe_bytes / (fe_ms * 1e-3) / 1e9 / peaks()["hbm_gbs"],
                                    "note": "wav [B, L] fp32 resident in HBM (one step's output, L2-sized per utterance) -> log-mel [B, T, 80]"}
        fe.close()
        # discriminator forward passes at the reference's training-step shape (config_v1.json: batch_size
        # 16, segment_size 8192; gan.py scores y and y_hat -> 32 waveforms), cut from the audio this step produced
        from neuraltexttospeech_b200.discriminators import MultiPeriodDiscriminator, MultiScaleDiscriminator, disc_flops
        from neuraltexttospeech_b200.synthetic import synthetic_disc_weights
        nb_d, seg = min(B, 32), min(L, 8192)
        seg_wav = wav[:nb_d, :seg].contiguous()
        # every conv but the first and conv_post runs as 3xTF32 (three tcgen05 kind::tf32 products per fp32 product):
        # the bound is a third of the TF32 tensor peak, itself taken as half of the measured bf16 peak (not measured)
        drow = {"waveforms": nb_d, "samples": seg, "dtype": "f32 (3xTF32 on the tensor cores)", "bound": "tensor",
                "peak_tflops": peaks()["tflops_sustained"] / 2 / 3,
                "peak_source": "measured bf16 sustained peak / 2 (TF32) / 3 (three products per fp32 product); TF32 peak itself not measured"}
        for kind, cls in (("mpd", MultiPeriodDiscriminator), ("msd", MultiScaleDiscriminator)):
            dm = cls(device=local_rank).load_weights(synthetic_disc_weights(kind, "A"))
            for _ in range(3):
                dm.forward(seg_wav)
            torch.cuda.synchronize()
            f0.record(stream)
            for _ in range(5):
                dm.forward(seg_wav)
            f1.record(stream)
            torch.cuda.synchronize()
            d_ms = f0.elapsed_time(f1) / 5
            fl, by = disc_flops(dm, nb_d, seg)
            drow[kind] = {"ms": d_ms, "gflop": fl / 1e9, "tflops": fl / d_ms / 1e9, "frac": fl / d_ms / 1e9 / drow["peak_tflops"],
                          "feature_map_mbytes": by / 1e6, "launches": 75 if kind == "mpd" else 62}
            del dm
        adjacent["discriminators"] = drow

    if rank == 0:
        pk = peaks()
        act_b = 2 if args.precision in ("fp16", "bf16") else 4
        flops = 2.0 * count_macs_per_frame(h) * T * B
        bytes_boundary = boundary_bytes_per_frame(h, act_b) * T * B
        tensor_bound = args.precision != "fp32"
        p_tensor, bw = pk["tflops_sustained"] * 1e12, pk["hbm_gbs"] * 1e9
        if args.precision == "tf32":          # no measured TF32 GEMM peak on this pool: half of the measured bf16 peak (the nominal ratio)
            p_tensor *= 0.5
            pk = dict(pk, tflops_sustained=pk["tflops_sustained"] * 0.5, source=pk["source"] + "; TF32 = 0.5 x measured bf16 (nominal ratio, not measured)")
        # per-stage roofline: t_roof = max(F / P_tensor, B / BW_hbm) per forward of `mb` utterances
        stages, t_roof_sum = [], 0.0
        for w, ms, nl in zip(per_stage_work(h, act_b), stage_ms, stage_launches):
            w["launches"] = nl
            f_st, b_st = 2.0 * w["macs"] * T * mean_fwd_batch, float(w["bytes"]) * T * mean_fwd_batch
            t_t, t_h = f_st / p_tensor * 1e3, b_st / bw * 1e3
            t_roof = max(t_t, t_h) if tensor_bound else t_h
            t_roof_sum += t_roof
            stages.append({"label": w["label"], "ms": round(ms, 4), "launches": nl, "bound": "tensor" if (tensor_bound and t_t >= t_h) else "hbm",
                           "roof_ms": round(t_roof, 4), "frac": round(t_roof / ms, 4) if ms > 0 else None,
                           "tflops": round(f_st / (ms * 1e-3) / 1e12, 1) if ms > 0 else None,
                           "gbs": round(b_st / (ms * 1e-3) / 1e9, 1) if ms > 0 else None})
        # dominant kernel: the stage with the largest measured time; all of its launches are conv_tc_kernel instances
        dom = max(range(len(stages)), key=lambda i: stages[i]["ms"])
        dw, dst = dict(per_stage_work(h, act_b)[dom], launches=stage_launches[dom]), stages[dom]
        dom_flops_launch = 2.0 * dw["macs"] * T * mean_fwd_batch / dw["launches"]
        dom_ms_launch = dst["ms"] / dw["launches"]
        dom_bytes_launch = float(dw["bytes"]) * T * mean_fwd_batch / dw["launches"]
        dom_tensor = dst["bound"] == "tensor"              # the dominant stage's own binding roof (V2 / V3: HBM)
        achieved = (dom_flops_launch / (dom_ms_launch * 1e-3) / 1e12) if dom_tensor else (dom_bytes_launch / (dom_ms_launch * 1e-3) / 1e9)
        dom_peak = pk["tflops_sustained"] if dom_tensor else pk["hbm_gbs"]
        traffic = ncu_traffic(dst["label"], args.config, mean_fwd_batch)
        line = {
            "metric": metric_name(version), "value": world * audio_s_per_step / (dev_ms / 1000), "unit": "audio-s/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms, "higher_is_better": True,
            "scaling": "strong" if strong else "weak", "vs_baseline": None,
            "dtype": {"fp16": "fp16 (fp32 accumulate)", "bf16": "bf16 (fp32 accumulate)", "fp32": "fp32",
                      "tf32": "tf32 operands, fp32 storage (fp32 accumulate)"}[args.precision],
            "data": "synthetic",
            "config": {"workload": args.config, "generator": f"config_v{version}.json", "per_gpu_batch": B, "frames": T,
                       "samples_per_utterance": L, "micro_batch": mb, "weights": "random-init (Glorot-uniform v, g=||v||, b=0)",
                       "l2": "per-step working set (GBs of activations) exceeds the 126 MB L2; no flush needed",
                       "sharding": "utterances sharded across ranks, no collective on the data path"},
            "e2e": e2e, "gpu_launches": int(launches_per_step * args.steps), "clocks": clocks,
            "roofline": {"bound": "tensor" if dom_tensor else "hbm", "achieved": achieved, "peak": dom_peak,
                         "unit": "TFLOP/s" if dom_tensor else "GB/s", "frac": achieved / dom_peak, "traffic": traffic,
                         "kernel": (f"resblock_tc_kernel (fused ResBlock1 iterations, tcgen05) + conv_tc_kernel (ups), {dst['label']}: "
                                    if stage_launches[dom] < per_stage_work(h, act_b)[dom]["launches"] else
                                    f"conv_tc_kernel (tcgen05 implicit GEMM), {dst['label']}: ") + f"{dw['launches']} launches per forward of {mb} utterances",
                         "flops_per_launch": dom_flops_launch, "bytes_per_launch": dom_bytes_launch, "ms_per_launch": dom_ms_launch,
                         "timing": "CUDA events on the launch stream around the stage, inside the timed region, averaged over its forwards",
                         "peak_source": pk["source"] + (" bf16 sustained (fp16 runs at the same tcgen05 rate; kernel timed inside a long step)" if dom_tensor
                                                        else " HBM copy bandwidth, sustained (layer-boundary bytes of the stage, SURVEY 8d)")},
            "roofline_step": {"achieved_tflops": flops / (dev_ms / 1000) / 1e12, "frac_of_tensor_peak": flops / (dev_ms / 1000) / p_tensor,
                              "per_stage_roof_ms_per_step": t_roof_sum * n_forwards, "frac_of_per_stage_roofline": t_roof_sum * n_forwards / dev_ms,
                              "algorithmic_flops_per_step": flops, "layer_boundary_bytes_per_step": bytes_boundary,
                              "hbm_floor_ms": bytes_boundary / bw * 1000, "tensor_floor_ms": flops / p_tensor * 1000},
            "stages": {"batch": mean_fwd_batch, "forwards_per_step": n_forwards, "rows": stages},
            "adjacent": adjacent,
        }
        if not args.no_cpu_baseline and world == 1:
            # CPU baseline on rank 0: BASELINE.json configs[0] (V1, batch 1, 256 frames), oracle port
            best, mean, threads, ts = cpu_port_audio_s_per_s("1", 1, 256, 5)
            line["cpu_baseline"] = {"value": best, "unit": "audio-s/s", "cores": threads, "kind": "port", "tf_reference": tf_pin(),
                                    "sample": "V1 batch 1 x 256 frames (configs[0]), best of 5; torch-CPU fp32 restatement of "
                                              "HiFiGAN_TF/model.py (TensorFlow not installable on this image)"}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
