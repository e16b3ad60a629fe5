"""K1x timings (tac_stft_ex): the SGMSE+ front end (n_fft 510, hop 128), the notebook's dB front end (n_fft 2046, hop 512) and
torchaudio's default power spectrogram (n_fft 400, hop 200), 512 x 1 s and 64 x 10 s clips (CUDA events).  python tools/time_stft_ex.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")); sys.path.insert(0, ROOT)
import torch
import tac_spectral as ts
dev = torch.device("cuda", 0)
def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
for B, T in ((512, 16000), (64, 160000)):
    x = 0.1 * torch.randn(B, T, device=dev)
    fe = ts.SpecsFrontEnd()
    sp = ts.Spectrogram()
    cases = (("sgmse n=510 hop=128 (chirp-z, complex out, pad to x64)", lambda: fe(x)),
             ("notebook n=2046 hop=512 mag+phase+dB (chirp-z)", lambda: ts.create_complex_spectrogram(x, 16000, n_fft=2046, hop_length=512, win_length=2046, db_scale=True)),
             ("torchaudio default n=400 hop=200 power (chirp-z)", lambda: sp(x)),
             ("pow2 reference point n=512 hop=128 power", lambda: ts.Spectrogram(n_fft=512, hop_length=128)(x)))
    for name, fn in cases:
        out = fn()
        nbytes = 4 * B * T + sum(o.numel() * o.element_size() for o in (out.values() if isinstance(out, dict) else [out]) if torch.is_tensor(o))
        ms = timeit(fn)
        print(json.dumps({"case": f"{name}, {B} x {T / 16000:.0f} s", "ms": ms, "bytes": nbytes, "gbs": nbytes / ms / 1e6}))
