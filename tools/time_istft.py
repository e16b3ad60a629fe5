"""Time K4 (tac_istft) on the decoder-output shape of the reference, [B, 2, 1025, 600] -> [B, 306688] (CUDA events).
python tools/time_istft.py [B]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")); sys.path.insert(0, ROOT)
import torch
import tac_spectral as ts
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
peak = 6538.0
try: peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception: pass
dev = torch.device("cuda", 0)
def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
for (n_fft, hop, K, F) in [(2048, 512, 1025, 600), (2048, 512, 1025, 32), (1024, 256, 513, 600), (512, 128, 257, 1200)]:
    rows = B if F > 100 else 512
    dec = torch.randn(rows, 2, K, F, device=dev)
    dec[:, 0].abs_()
    ms = timeit(lambda: ts.batch_reconstruct_waveform(dec[:, 0], dec[:, 1], 16000, n_fft, hop, n_fft))
    L = hop * (F - 1)
    byts = rows * (8 * K * F + 4 * L)
    print(json.dumps({"case": f"istft polar n={n_fft} hop={hop} rows={rows} F={F}", "ms": ms, "bytes": byts, "gbs": byts / ms / 1e6,
                      "frac_of_measured_hbm": byts / ms / 1e6 / peak, "audio_s_per_s": rows * L / 16000 / (ms / 1e3)}))
    spec = torch.view_as_complex(torch.randn(rows, K, F, 2, device=dev))
    ms = timeit(lambda: ts.istft(spec, n_fft, hop))
    print(json.dumps({"case": f"istft complex n={n_fft} hop={hop} rows={rows} F={F}", "ms": ms, "gbs": byts / ms / 1e6,
                      "frac_of_measured_hbm": byts / ms / 1e6 / peak}))
    if F == 600 and n_fft == 2048:
        w = torch.hann_window(n_fft, device=dev)
        def ref():
            c = dec[:, 0] * torch.exp(1j * dec[:, 1])
            return torch.stack([torch.istft(c[i], n_fft=n_fft, hop_length=hop, win_length=n_fft, window=w, return_complex=False) for i in range(rows)], 0)
        print(json.dumps({"case": "reference path on this GPU (torch.istft per item, train_pkl_file_task_aware.py:183-208)", "ms": timeit(ref, 3)}))
    del dec, spec
