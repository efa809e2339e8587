import sys, time; sys.path.insert(0, "/root/repo")
import torch
from neuraltexttospeech_b200 import Generator, HParams
from neuraltexttospeech_b200.synthetic import synthetic_mel, synthetic_weights
h = HParams.builtin("1"); gen = Generator(h).load_weights(synthetic_weights(h, "A"))
for B, T in ((1, 256), (1, 862), (4, 862)):
    mel = torch.from_numpy(synthetic_mel(B, T, 80)).cuda()
    out = torch.empty((B, T * 256, 1), device="cuda")
    for _ in range(5): gen._call_device(mel, out)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(100): gen._call_device(mel, out)
    t_issue = (time.perf_counter() - t0) / 100
    torch.cuda.synchronize()
    t_total = (time.perf_counter() - t0) / 100
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        gen._call_device(mel, out)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            gen._call_device(mel, out)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(100): g.replay()
    torch.cuda.synchronize()
    t_graph = (time.perf_counter() - t0) / 100
    print(f"B {B} T {T}: cpu issue {t_issue*1e3:.3f} ms, plain total {t_total*1e3:.3f} ms, graph replay {t_graph*1e3:.3f} ms")
