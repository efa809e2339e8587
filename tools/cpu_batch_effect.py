"""Batch effect of the CPU arm: the torch-CPU fp32 port (oracle/hifigan_oracle.py, what `bench.py --impl reference` times)
at B = 1, 4, 16 utterances of 862 frames on all host threads of the box it runs on.  Writes the table as JSON to stdout;
the round's copy is profiles/r2d_cpu_batch_effect.json."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

rows = []
for version in (sys.argv[1:] or ["1"]):
    for B in (1, 4, 16):
        best, mean, threads, ts = bench.cpu_port_audio_s_per_s(version, B, 862, 2)
        rows.append({"config": f"V{version}", "utterances": B, "frames": 862, "threads": threads, "audio_s_per_s_best": round(best, 2),
                     "audio_s_per_s_mean": round(mean, 2), "seconds_per_forward": [round(t, 2) for t in ts]})
        print(rows[-1], file=sys.stderr, flush=True)
print(json.dumps({"what": "torch-CPU fp32 port of HiFiGAN_TF/model.py Generator.call, batch effect", "rows": rows}, indent=1))
