"""Times the mel front-end kernel alone: B utterances of L samples resident on the device (tools/mel_bench.py [B] [L])."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from neuraltexttospeech_b200 import STFT

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
L = int(sys.argv[2]) if len(sys.argv) > 2 else 220672
fe = STFT()
wav = (torch.rand(B, L, device="cuda") - 0.5)
T = fe.num_frames(L)
for _ in range(3):
    fe.mel_spectrogram(wav)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for _ in range(20):
    fe.mel_spectrogram(wav)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
byts = B * L * 4 + B * T * 80 * 4
print(f"mel_frontend B={B} L={L} T={T}: {ms * 1000:.1f} us, {byts / ms / 1e6:.0f} GB/s, {B * T / 2 / ms / 1e3:.2f} M FFT/s")
