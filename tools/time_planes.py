"""K1 planes kernel (MS-STFT discriminator inputs) on cfg4 (5 scales, B=512 x 10 s) and cfg2 (3 scales, 512 x 1 s): ms, GB/s of
algorithmic bytes (4*B*T + 8*B*sum F_s*K_s), fraction of MEASURED_PEAKS hbm_gbs.  python tools/time_planes.py [B4]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")); sys.path.insert(0, ROOT)
import torch
from tac_spectral import _lib as L
from tac_spectral.stft import num_frames
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
lib = L.lib()
dev = torch.device("cuda", 0)
def run(name, B, T, n_ffts, reps=5):
    hops = [n // 4 for n in n_ffts]
    g = torch.Generator(device="cpu").manual_seed(1)
    x = (0.1 * torch.randn(min(B, 64), T, generator=g)).repeat((B + 63) // 64, 1)[:B].contiguous().to(dev)
    outs = [torch.empty((B, 2, num_frames(T, n, h, False), n // 2 + 1), device=dev, dtype=torch.float32) for n, h in zip(n_ffts, hops)]
    for n in n_ffts: L.check(lib.tac_prepare_fft(n, n, 1))
    nf, hp, ptrs = L.int_array(n_ffts), L.int_array(hops), L.ptr_array(outs)
    st = torch.cuda.current_stream().cuda_stream
    call = lambda: L.check(lib.tac_stft_multi(x.data_ptr(), B, 1, T, len(n_ffts), nf, hp, nf, 0, 0, 1, L.OUT_PLANES, 0, ptrs, st))
    for _ in range(3): call()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): call()
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    nbytes = 4 * B * T + 8 * B * sum(o.shape[2] * o.shape[3] for o in outs)
    gbs = nbytes / ms / 1e6
    print(json.dumps({"case": name, "B": B, "T": T, "n_ffts": list(n_ffts), "ms": ms, "bytes": nbytes, "gbs": gbs, "frac_of_measured_hbm": gbs / peak,
                      "audio_s_per_s": B * T / 16000 / (ms * 1e-3)}))
    del outs, x
    torch.cuda.empty_cache()
print("lib", os.environ.get("TAC_SPECTRAL_LIB", "default"))
B4 = int(sys.argv[1]) if len(sys.argv) > 1 else 512
run("cfg4", B4, 160000, (2048, 1024, 512, 256, 128))
run("cfg2-disc", 512, 16000, (1024, 2048, 512), reps=20)
run("cfg3-disc", 4096, 16000, (1024, 2048, 512), reps=5)
for n in (() if os.environ.get("TAC_PLANES_SHORT") else (2048, 1024, 512, 256, 128)):
    run(f"cfg4-only-{n}", B4, 160000, (n,), reps=3)
