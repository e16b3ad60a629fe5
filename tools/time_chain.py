"""Chain decoder spectrogram -> waveform -> discriminator inputs (tac_istft_disc_inputs) against the two separate calls, on the
reference's decoder shape [B, 2, 1025, 600] (CUDA events; inputs rotate over a ring larger than L2).
python tools/time_chain.py [B] [once]   ('once': a single call of each form, for an ncu --metrics dram__bytes pass)"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")); sys.path.insert(0, ROOT)
import torch
import tac_spectral as ts
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
once = len(sys.argv) > 2 and sys.argv[2] == "once"
dev = torch.device("cuda", 0)
decs = []
for i in range(1 if once else 2):                      # 2 x 630 MB of decoder output: every call reads cold data
    d = torch.randn(B, 2, 1025, 600, device=dev); d[:, 0].abs_(); decs.append(d)
def timeit(fn, n=6):
    for i in range(2): fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(n): fn(i)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
for n_ffts, hops in (((1024, 2048, 512), (256, 512, 128)), ((2048, 1024, 512, 256, 128), (512, 256, 128, 64, 32))):
    ms_f = ts.MultiScaleSTFT(n_ffts, hops, n_ffts)
    def separate(i):
        d = decs[i % len(decs)]
        w = ts.batch_reconstruct_waveform(d[:, 0], d[:, 1], 16000)
        return w, ms_f(w.unsqueeze(1))
    def chain(i, g=0):
        d = decs[i % len(decs)]
        return ts.reconstruct_and_disc_inputs(d[:, 0], d[:, 1], 16000, disc_n_ffts=n_ffts, disc_hop_lengths=hops, disc_win_lengths=n_ffts, rows_per_group=g)
    if once:
        separate(0); chain(0); torch.cuda.synchronize(); continue
    w1, z1 = separate(0); w2, z2 = chain(0)
    same = torch.equal(w1, w2) and all(torch.equal(a, b) for a, b in zip(z1, z2))
    del w1, z1, w2, z2
    T = 512 * 599
    out_bytes = B * (4 * T + sum(8 * (1 + (T - n) // h) * (n // 2 + 1) for n, h in zip(n_ffts, hops)))
    rec = {"case": f"chain B={B} scales={list(n_ffts)}", "bit_identical": same, "separate_ms": timeit(separate),
           "chain_ms": timeit(chain), "alg_bytes": B * 8 * 1025 * 600 + out_bytes}
    for g in (8, 16, 64):
        rec[f"chain_g{g}_ms"] = timeit(lambda i: chain(i, g))
    rec["chain_gbs"] = rec["alg_bytes"] / rec["chain_ms"] / 1e6
    print(json.dumps(rec))
