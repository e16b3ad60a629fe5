#!/usr/bin/env python
"""bench.py -- audio-seconds/sec through the STFT-loss + DPCA step (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload cfg2|cfg3|cfg1]

One "step" = one pass of the spectral hot path over one batch of synthetic 16 kHz audio
("Step metric"): K1 front end (mag/phase spectrogram of every clip), K2 fused
multi-resolution STFT loss forward + backward on (target, decoded), K3 DPCA (statistics, eigen-solve,
global top-dim_p selection, projection) of the per-view latents.  Workload at N=1 = BASELINE.json
configs[1] ("SpectralResE2D2 latent=32: two correlated views, batch 256"); for N>1 every rank holds
the same per-GPU shard (weak scaling) and the loss sums + DPCA statistics are all-reduced once per
step over NCCL.

Prints ONE JSON line (rank 0).  `value` = whole-job audio-s/s with inputs resident in HBM (CUDA
graph replay of the step, rotating input sets larger than L2); `e2e` = the same through the public
API with pinned HOST buffers, H2D copies and the D2H read of the loss inside the timed region;
`roofline` = the dominant kernel against MEASURED_PEAKS.json; `cpu_baseline` = the CPU oracle
(torch CPU restatement of the reference call sites) on a bounded sample of the same workload.

`--impl reference` times that CPU path alone, with all host threads, on the same workload/metric.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "task-aware-audio-coding-perceptual_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

SR = 16000
WORKLOADS = {
    # name: clips per rank = batch * views, 1 s clips; DPCA views of [batch, d]; dim_p = total kept components
    "cfg1": dict(desc="SpectralResE1D1 latent=8: 1 view, batch 16, 1 s clips", batch=16, views=1, T=16000, d=128, dim_p=8),
    "cfg2": dict(desc="SpectralResE2D2 latent=32: 2 correlated views, batch 256, 1 s clips, 3-scale loss fwd+bwd + "
                      "front end + DPCA d=128/view dim_p=32", batch=256, views=2, T=16000, d=128, dim_p=32),
    "cfg3": dict(desc="SpectralResE4D1 latent=128: 4 correlated views, batch 128 per GPU, 1 s clips, DPCA d=128/view "
                      "dim_p=128", batch=128, views=4, T=16000, d=128, dim_p=128),
}
N_FFTS = (1024, 2048, 512)          # msstftd.py:174
HOPS = (256, 512, 128)              # msstftd.py:175


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of CUDA-graph replay")
    ap.add_argument("--no-overlap", action="store_true", help="run every leg on one stream")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the cfg 3 / cfg 4 / cfg 5 extra block")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU-baseline budget (bounded sample)")
    return ap.parse_args()


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------ synthetic data
def synth_host(batch, views, T, d, seed):
    """Correlated multi-view clips in [-1, 1], "decoded" = view + noise, latents with a decaying spectrum
    (the recipe of SURVEY.md 8d).  Returns CPU tensors x [batch*views, T], y, zs [views][batch, d]."""
    import math
    import torch
    g = torch.Generator(device="cpu").manual_seed(seed)
    n = 0.1 * torch.randn(batch, T + 1, generator=g)
    src = 0.5 * (n[:, 1:] + n[:, :-1])
    t = torch.arange(T, dtype=torch.float32) / SR
    for f0 in (220.0, 880.0, 3520.0):
        ph = 2 * math.pi * torch.rand(batch, 1, generator=g)
        src = src + 0.05 * torch.sin(2 * math.pi * f0 * t[None, :] + ph)
    amps = (1.0, 0.9, 0.8, 0.7)
    xs, ys = [], []
    for i in range(views):
        v = amps[i % 4] * src + 0.05 * torch.randn(batch, T, generator=g)
        xs.append(v.clamp(-1, 1))
        ys.append((v + 0.02 * torch.randn(batch, T, generator=g)).clamp(-1, 1))
    x = torch.stack(xs, 1).reshape(batch * views, T).contiguous()     # views of a clip stay adjacent
    y = torch.stack(ys, 1).reshape(batch * views, T).contiguous()
    sigma = 0.97 ** torch.arange(d, dtype=torch.float32)
    shared = torch.randn(batch, d, generator=g)
    zs = []
    for i in range(views):
        G = (shared + torch.randn(batch, d, generator=g)) / math.sqrt(2.0)
        Q, _ = torch.linalg.qr(torch.randn(d, d, generator=g))
        zs.append((((G * sigma) @ Q.T) * (1.0 + 0.03 * i) + 0.3).contiguous())
    return x, y, zs


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock and throttle reasons with NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nv = None

    def _loop(self):
        nv = self._nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.02)

    def __enter__(self):
        if self._nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


# ------------------------------------------------------------------------------------------ CPU path (oracle)
class CpuPath:
    """The CPU restatement of the reference call sites (oracle/spectral_oracle.py: torch.stft on the msstftd.py
    convention, autograd backward, torch.linalg.svd DPCA) on a BOUNDED sample of the workload: `clips` items (all
    views) per audio pass, scaled linearly to the batch; DPCA on the workload's full latent batch (it is tiny)."""

    def __init__(self, w, clips, threads):
        import torch
        from oracle import spectral_oracle as O
        torch.set_num_threads(threads)
        self.O, self.w, self.clips, self.threads = O, w, clips, threads
        self.x, self.y, _ = synth_host(clips, w["views"], w["T"], w["d"], seed=77)
        self.zfull = synth_host(w["batch"], w["views"], 64, w["d"], seed=78)[2]

    def audio_pass(self):
        O = self.O
        mag, ph = O.frontend_ref(self.x)                                # K1
        O.stack_mag_phase_ref(mag, ph)
        yr = self.y.clone().requires_grad_(True)                        # K2 fwd + bwd
        loss = O.mrstft_loss_ref(self.x, yr, N_FFTS, HOPS)
        loss.backward()
        return float(loss.detach())

    def dpca_pass(self):
        self.O.dpca_project_ref(self.zfull, self.w["dim_p"])            # K3

    def step_seconds(self):
        """One bounded-sample step: seconds the FULL workload step would take (audio pass scaled)."""
        t0 = time.perf_counter(); self.audio_pass(); t1 = time.perf_counter(); self.dpca_pass(); t2 = time.perf_counter()
        return (t1 - t0) * (self.w["batch"] / self.clips) + (t2 - t1)

    def sample(self, n):
        w = self.w
        return (f"{n} timed passes of {self.clips} items x {w['views']} view(s) x {w['T'] / SR:g} s (front end + 3-scale loss fwd+bwd by "
                f"torch autograd), scaled x{w['batch'] / self.clips:g}, + DPCA (torch.linalg.svd) on the full [{w['batch']}, {w['d']}] x "
                f"{w['views']} latents; {self.threads} torch threads")


def cpu_baseline(w, budget_s, threads):
    cp = CpuPath(w, min(w["batch"], 16), threads)
    cp.step_seconds()                                                   # warm-up
    ts, t_end = [], time.perf_counter() + budget_s
    while len(ts) < 3 or (time.perf_counter() < t_end and len(ts) < 200):
        ts.append(cp.step_seconds())
    t_step = sum(ts) / len(ts)
    return (w["batch"] * w["views"] * w["T"] / SR) / t_step, t_step, cp.sample(len(ts))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    w = WORKLOADS[args.workload]
    threads = os.cpu_count() or 1
    cp = CpuPath(w, min(w["batch"], 16), threads)
    for _ in range(args.warmup):
        cp.step_seconds()
    ts = [cp.step_seconds() for _ in range(args.steps)]
    t_step = sum(ts) / len(ts)
    value = (w["batch"] * w["views"] * w["T"] / SR) / t_step
    line = {"impl": "reference", "metric": "audio-seconds/sec through STFT-loss+DPCA step", "value": value,
            "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {w['desc']}", "note": "reference CPU path (the torch CPU calls the reference's "
                       "call sites make, restated in oracle/); rank 0's host cores only -- the reference has no multi-GPU path, so "
                       "the value does not grow with --gpus"},
            "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": threads, "kind": "port", "sample": cp.sample(args.steps),
                             "torch": torch.__version__},
            "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ stock torch on the same GPU
def torch_gpu_baseline(x, y, zs, dim_p, iters=10):
    """The same step through the library calls the reference's call sites make (torch.stft -> cuFFT, autograd,
    torch.svd -> cuSOLVER, boolean-mask selection, cuBLAS matmuls), on the same B200: BASELINE.md's "kernel to beat".
    Written out here (not imported from oracle/): msstftd.py:125-127,151 convention for the loss scales, the notebook
    front end (spectrogram.ipynb cell 10), data_pca + select + project of pkl_file_model_updated.py:37-42,489-511."""
    import torch
    dev = x.device
    wins = {n: torch.hann_window(n, device=dev) for n in set(N_FFTS) | {2048}}
    norm = {n: wins[n].pow(2).sum().sqrt() for n in N_FFTS}

    def step():
        D = torch.stft(x, 2048, 512, 2048, wins[2048], center=True, pad_mode="constant", return_complex=True)
        fe = torch.stack([D.abs(), D.angle()], 1)
        yr = y.detach().requires_grad_(True)
        loss = 0.0
        for n, h in zip(N_FFTS, HOPS):
            X = torch.stft(x, n, h, n, wins[n], center=False, return_complex=True).abs() / norm[n]
            Y = torch.stft(yr, n, h, n, wins[n], center=False, return_complex=True).abs() / norm[n]
            loss = loss + (X - Y).abs().mean() + (X.clamp_min(1e-5).log() - Y.clamp_min(1e-5).log()).abs().mean()
        (loss / len(N_FFTS)).backward()
        pcs = []
        for z in zs:
            mu = z.mean(0)
            _, sv, v = torch.svd(z - mu)
            pcs.append((sv, v, mu))
        ind = torch.argsort(torch.cat([p[0] for p in pcs]), descending=True)[:dim_p]
        outs, off = [], 0
        for z, (sv, v, mu) in zip(zs, pcs):
            sel = ind[(ind >= off) & (ind < off + sv.numel())] - off
            off += sv.numel()
            zp = torch.matmul(z - mu, v[:, sel])
            outs.append(torch.matmul(zp, v[:, sel].T) + mu)
        return fe, yr.grad, outs, loss

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        step()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


# ------------------------------------------------------------------------------------------ host placement
def bind_to_gpu_numa(index):
    """Pin this process to the CPUs of its GPU's NUMA node (NVML affinity mask) BEFORE pinned host buffers are
    allocated, so that the e2e H2D copies do not cross sockets.  Best effort; returns what was done."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * i + b for i, wd in enumerate(mask) for b in range(64) if (int(wd) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if not cpus:
            return {"bound": False, "why": "NVML affinity mask empty inside this cpuset"}
        os.sched_setaffinity(0, cpus)
        out = {"bound": True, "cpus": len(cpus), "of": len(allowed)}
        out.update(prefer_gpu_memory_node(pynvml, h))
        return out
    except Exception as e:
        return {"bound": False, "why": f"{type(e).__name__}: {e}"}


def prefer_gpu_memory_node(pynvml, h):
    """set_mempolicy(MPOL_PREFERRED, node of the GPU's PCIe slot) for this process, so that the pinned host buffers of the e2e
    leg are allocated on the GPU's own socket even when the cpuset spans (or sits on) another one: with 8 ranks pulling
    66 MB per step each from ONE node's DRAM the copies ran at 25 GB/s per GPU instead of 55.  Best effort."""
    try:
        import ctypes
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        cand = [bus.lower(), bus.lower()[4:] if len(bus) > 12 else bus.lower()]       # NVML pads the domain to 8 digits
        node = None
        for c in cand:
            f = f"/sys/bus/pci/devices/{c}/numa_node"
            if os.path.exists(f):
                node = int(open(f).read().strip())
                break
        if node is None or node < 0:
            return {"mem_node": None}
        nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()]
        if len(nodes) < 2:
            return {"mem_node": node, "mem_policy": "single node"}
        libc = ctypes.CDLL(None, use_errno=True)
        mask = ctypes.c_ulong(1 << node)
        MPOL_PREFERRED, SYS_set_mempolicy = 1, 238                                    # x86_64
        rc = libc.syscall(SYS_set_mempolicy, MPOL_PREFERRED, ctypes.byref(mask), ctypes.c_ulong(8 * ctypes.sizeof(mask)))
        return {"mem_node": node, "mem_policy": "preferred" if rc == 0 else f"set_mempolicy errno {ctypes.get_errno()}"}
    except Exception as e:
        return {"mem_node": None, "mem_policy": f"{type(e).__name__}: {e}"}


# ------------------------------------------------------------------------------------------ N-rank parity on hardware
def check_parity_n(step, dset, w, dev, rank, world, rerun=True):
    """The sharded step (p2p exchange or NCCL, as benchmarked) against the SAME kernels run on the whole batch by one
    GPU: every rank runs its shard once more, the shards are all-gathered, rank 0 runs the 1-rank step on the
    concatenation (the reference's semantics: loss means and PCA over the whole batch on one device,
    train_pkl_file_task_aware.py:597-629, pkl_file_model_updated.py:489-511) and compares loss, the per-view selection
    (cnt / idx), the projected latents and the loss gradient.  Tolerances: loss 1e-9 relative (float64 sums in a
    different order), latents 1e-4 of their largest entry (north star), gradient 1e-6 of its largest entry."""
    import torch
    import torch.distributed as dist
    from tac_spectral.step import SpectralStep
    x, y, zs = dset
    if rerun:                                                           # False: judge the results already in `step` (graph replay)
        step.run(x, y, zs)
    torch.cuda.synchronize()

    def gather(t):
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t.contiguous())
        return torch.cat(out, 0)

    X, Y = gather(x), gather(y)
    Zs = [gather(z) for z in zs]
    G = gather(step.grad_y)
    Zo = [gather(z) for z in step.z_out]
    cnts = gather(step.cnt[None])                                       # [world, nv]: must agree across ranks
    losses = gather(step.loss.reshape(1))
    if rank != 0:
        return None
    one = SpectralStep(X.shape[0], X.shape[1], len(Zs), Zs[0].shape[0], Zs[0].shape[1], w["dim_p"], N_FFTS, HOPS, device=dev)
    one.run(X, Y, Zs)
    torch.cuda.synchronize()
    res = {"loss_rel": abs(float(losses[0]) - float(one.loss)) / abs(float(one.loss)),
           "loss_rank_spread": float((losses - losses[0]).abs().max()),
           "cnt_equal": bool((cnts == one.cnt[None]).all()),
           "idx_equal": bool(all(torch.equal(step.idx[v, :int(one.cnt[v])], one.idx[v, :int(one.cnt[v])]) for v in range(len(Zs)))),
           "z_out_rel": max(float((a - b).abs().max() / b.abs().max()) for a, b in zip(Zo, one.z_out)),
           "grad_rel": float((G - one.grad_y).abs().max() / one.grad_y.abs().max())}
    ok = (res["loss_rel"] <= 1e-9 and res["loss_rank_spread"] == 0.0 and res["cnt_equal"] and res["idx_equal"]
          and res["z_out_rel"] <= 1e-4 and res["grad_rel"] <= 1e-6)
    res["verdict"] = "ok" if ok else "MISMATCH"
    return res


# ------------------------------------------------------------------------------------------ cfg 3 / cfg 4 / cfg 5
def _time_ms(fn, iters, warm=2):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def run_extras(args, dev, rank, world, group, peak, barrier):
    """The other BASELINE.json configs, measured in the same run so that they are driver-visible:
    cfg 4 (K1 discriminator inputs, 5 scales, 512 x 10 s -- the one genuinely HBM-bound config), cfg 5 (60 s clips,
    dim_p sweep 8..256 over 4 views of d = 128) and, at 8 ranks, the cfg 3 step (4 views, batch 1024 global,
    dim_p = 128).  Device-timed with CUDA events; every rank runs them (the exchange needs all ranks), rank 0 reports."""
    import torch
    import torch.distributed as dist
    import tac_spectral as ts
    from tac_spectral.step import SpectralStep
    out = {}
    g = torch.Generator(device="cpu").manual_seed(4321 + rank)

    # ---- cfg 4: msstftd front end, n_fft 2048..128, hop n/4, 512 clips x 160 000 samples -> [512, 2, F_s, K_s] planes
    try:
        n4, h4 = [2048, 1024, 512, 256, 128], [512, 256, 128, 64, 32]
        B4, T4 = 512, 160000
        x4 = (0.1 * torch.randn(B4, 1, T4, generator=g)).to(dev)
        ms4 = ts.MultiScaleSTFT(n4, h4, n4)
        outs = None

        def k1():
            nonlocal outs
            outs = None                                                  # 6.6 GB of planes: one set alive at a time
            outs = ms4(x4)
        t4 = _time_ms(k1, iters=5, warm=2)
        bytes4 = 4 * B4 * T4 + 8 * B4 * sum((1 + (T4 - n) // h) * (n // 2 + 1) for n, h in zip(n4, h4))
        out["cfg4_k1_planes"] = {"workload": "msstftd front end, 5 scales 2048..128 hop n/4, 512 x 10 s, planes [B,2,F,K]",
                                 "ms": t4, "bytes": bytes4, "gbs": bytes4 / (t4 * 1e-3) / 1e9, "frac": bytes4 / (t4 * 1e-3) / 1e9 / peak,
                                 "audio_s_per_s": B4 * T4 / SR / (t4 * 1e-3), "includes": "output allocation by the caching allocator"}
        # the same five scales as ONE magnitude plane per scale from the same single launch (tac_stft_multi_epi): half the
        # bytes written; algorithmic bytes 4 B T + 4 B sum F_s K_s
        outs = None
        torch.cuda.empty_cache()
        try:
            def k1m():
                nonlocal outs
                outs = None
                outs = ms4.magnitudes(x4)
            t4m = _time_ms(k1m, iters=5, warm=2)
            bytes4m = 4 * B4 * T4 + 4 * B4 * sum((1 + (T4 - n) // h) * (n // 2 + 1) for n, h in zip(n4, h4))
            out["cfg4_k1_mag"] = {"workload": "same 5 scales, 512 x 10 s, one magnitude plane per scale [B,1,F,K] (stft_planes_mag_kernel)",
                                  "ms": t4m, "bytes": bytes4m, "gbs": bytes4m / (t4m * 1e-3) / 1e9, "frac": bytes4m / (t4m * 1e-3) / 1e9 / peak,
                                  "audio_s_per_s": B4 * T4 / SR / (t4m * 1e-3), "includes": "output allocation by the caching allocator"}
        except Exception as e:
            out["cfg4_k1_mag"] = {"error": f"{type(e).__name__}: {e}"}
        del outs, x4
        outs = None
        torch.cuda.empty_cache()
    except Exception as e:
        out["cfg4_k1_planes"] = {"error": f"{type(e).__name__}: {e}"}

    # ---- next rows (rank-local, a few ms): single-scale planes in the persistent form, the iSTFT on the decoder shape, the
    # chain iSTFT -> discriminator inputs, the non-power-of-two front ends (chirp-z on the warp core)
    try:
        rows_ = {}
        xs = (0.1 * torch.randn(512, 1, 160000, generator=g)).to(dev)
        for n in (2048, 512, 128):
            sp = ts.MultiScaleSTFT([n], [n // 4], [n])
            o = None

            def one():
                nonlocal o
                o = None
                o = sp(xs)
            t = _time_ms(one, iters=5, warm=2)
            nb = 4 * 512 * 160000 + 8 * 512 * (1 + (160000 - n) // (n // 4)) * (n // 2 + 1)
            rows_[f"planes_single_scale_{n}"] = {"ms": t, "gbs": nb / (t * 1e-3) / 1e9, "frac": nb / (t * 1e-3) / 1e9 / peak}
            o = None
        del xs
        torch.cuda.empty_cache()
        dec = torch.randn(128, 2, 1025, 600, generator=g).to(dev)
        dec[:, 0].abs_()
        t = _time_ms(lambda: ts.batch_reconstruct_waveform(dec[:, 0], dec[:, 1], SR), iters=5, warm=2)
        nb = 128 * (8 * 1025 * 600 + 4 * 512 * 599)
        rows_["istft_128x2x1025x600"] = {"ms": t, "gbs": nb / (t * 1e-3) / 1e9, "frac": nb / (t * 1e-3) / 1e9 / peak}
        t = _time_ms(lambda: ts.reconstruct_and_disc_inputs(dec[:, 0], dec[:, 1], SR), iters=5, warm=2)
        rows_["chain_istft_disc_inputs_3_scales"] = {"ms": t}
        del dec
        torch.cuda.empty_cache()
        x1 = (0.1 * torch.randn(512, 16000, generator=g)).to(dev)
        fe510, sp400 = ts.SpecsFrontEnd(), ts.Spectrogram()
        rows_["sgmse_front_end_n510_512x1s"] = {"ms": _time_ms(lambda: fe510(x1), iters=10, warm=2)}
        rows_["power_spectrogram_n400_512x1s"] = {"ms": _time_ms(lambda: sp400(x1), iters=10, warm=2)}
        rows_["notebook_db_front_end_n2046_512x1s"] = {"ms": _time_ms(lambda: ts.create_complex_spectrogram(
            x1, SR, n_fft=2046, hop_length=512, win_length=2046, db_scale=True), iters=10, warm=2)}
        del x1
        torch.cuda.empty_cache()
        out["next_rows"] = rows_
    except Exception as e:
        out["next_rows"] = {"error": f"{type(e).__name__}: {e}"}

    # ---- cfg 5: E4D1 latent=256 sweep: 16 items x 4 views x 60 s per GPU, DPCA d = 128 x 4, dim_p 8..256
    try:
        b5, v5, T5, d5 = 16, 4, 960000, 128
        x5 = (0.1 * torch.randn(b5 * v5, T5, generator=g)).to(dev)
        y5 = x5 + 0.02 * torch.randn(b5 * v5, T5, generator=g).to(dev)
        z5 = [z.to(dev) for z in synth_host(b5, v5, 64, d5, seed=555 + rank)[2]]
        sweep = {}
        for dim_p in (8, 16, 32, 64, 128, 256):
            st5 = SpectralStep(b5 * v5, T5, v5, b5, d5, dim_p, N_FFTS, HOPS, device=dev, group=group)
            t_step = _time_ms(lambda: st5.run(x5, y5, z5), iters=3, warm=1)
            t_dpca = _time_ms(lambda: (st5.leg_dpca_stats(z5), st5.leg_dpca_solve(), st5.leg_dpca_project(z5)), iters=5, warm=1)
            sweep[str(dim_p)] = {"step_ms": t_step, "dpca_chain_ms_local": t_dpca, "cnt": st5.cnt.tolist(),
                                 "audio_s_per_s": world * b5 * v5 * T5 / SR / (t_step * 1e-3)}
            st5.close()
            st5 = None
        out["cfg5_dim_p_sweep"] = {"workload": f"{b5} items x {v5} views x 60 s per GPU, 3-scale loss fwd+bwd + front end + DPCA "
                                               f"d={d5} x {v5} views (global batch {b5 * world})", "per_dim_p": sweep}
        del x5, y5, z5
        torch.cuda.empty_cache()
    except Exception as e:
        out["cfg5_dim_p_sweep"] = {"error": f"{type(e).__name__}: {e}"}
        if world > 1:
            raise                                                        # ranks must not diverge inside an exchange

    # ---- cfg 3 (the north star's 8-GPU config): 4 views, batch 128 per GPU (1024 global at 8), dim_p = 128
    if world == 8 or os.environ.get("TAC_BENCH_CFG3"):
        w3 = WORKLOADS["cfg3"]
        hx, hy, hz = synth_host(w3["batch"], w3["views"], w3["T"], w3["d"], seed=3000 + rank)
        x3, y3, z3 = hx.to(dev), hy.to(dev), [z.to(dev) for z in hz]
        st3 = SpectralStep(w3["batch"] * w3["views"], w3["T"], w3["views"], w3["batch"], w3["d"], w3["dim_p"], N_FFTS, HOPS,
                           device=dev, group=group)
        st3.run(x3, y3, z3)
        torch.cuda.synchronize()
        g3 = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g3):
            st3.run(x3, y3, z3)
        barrier()
        t3 = _time_ms(g3.replay, iters=max(10, args.steps), warm=3)
        tt = torch.tensor([t3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t3 = float(tt.item())
        out["cfg3_step"] = {"workload": w3["desc"], "n_gpus": world, "ms_per_step": t3, "loss": float(st3.loss.item()),
                            "audio_s_per_s": world * w3["batch"] * w3["views"] * w3["T"] / SR / (t3 * 1e-3), "cnt": st3.cnt.tolist(),
                            "note": "inputs L2-resident (one input set); device time, max over ranks"}
        g3 = None
        st3.close()
    return out


# ------------------------------------------------------------------------------------------ native
def run_native(args):
    import torch
    import torch.distributed as dist
    import tac_spectral as ts
    from tac_spectral import _lib as L
    from tac_spectral.step import SpectralStep

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (native) needs a CUDA device: the spectral hot path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL carries the rendezvous, the barriers and the max-over-ranks of the timing (and the step's exchange only
        # with TAC_STEP_EXCHANGE=nccl); NCCL_DEBUG is left as the caller set it
        opts = None
        if os.environ.get("TAC_NCCL_HIPRI", "1") != "0":
            try:
                opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
            except Exception:
                opts = None
        dist.init_process_group("nccl", device_id=dev, pg_options=opts)
        group = dist.group.WORLD
    lib = L.lib()
    numa = bind_to_gpu_numa(local)                    # before any pinned allocation: e2e copies stay on the GPU's socket

    w = WORKLOADS[args.workload]
    batch, views, T, d, dim_p = w["batch"], w["views"], w["T"], w["d"], w["dim_p"]
    rows = batch * views
    step = SpectralStep(rows, T, views, batch, d, dim_p, N_FFTS, HOPS, device=dev, group=group, overlap=not args.no_overlap)
    alg = step.algorithmic_bytes()
    exch_desc = {"none": "none (1 GPU)", "nccl": "NCCL all-reduce of one packed float64 buffer, two slices",
                 "p2p": f"one-shot NVLink peer stores + flags fused into the producing / consuming kernels "
                        f"({step.xchg.backend if step.xchg else '-'} peer memory), rank-ordered float64 sums"}[step.exchange]

    # ---- inputs resident in HBM: a ring of input sets larger than L2 (126 MB) so no step finds its inputs cached
    set_bytes = 2 * rows * T * 4 + views * batch * d * 4
    nsets = max(2, -(-400_000_000 // set_bytes))
    host_sets = [synth_host(batch, views, T, d, seed=1000 * rank + i) for i in range(min(nsets, 2))]
    dsets = []
    for i in range(nsets):
        hx, hy, hz = host_sets[i % len(host_sets)]
        jitter = 1.0 - 0.001 * (i // len(host_sets))                     # distinct values per set, same statistics
        dsets.append(((hx * jitter).to(dev), (hy * jitter).to(dev), [(z * jitter).to(dev) for z in hz]))
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    # ---- eager warm-up (also builds tables / sets attributes before any capture)
    for i in range(2):
        step.run(*dsets[i % nsets])
    torch.cuda.synchronize()

    graphs = None
    launches_per_step = None
    if not args.no_graph:
        try:
            graphs = []
            for i in range(nsets):
                g = torch.cuda.CUDAGraph()
                c0 = int(lib.tac_launch_count())
                with torch.cuda.graph(g):
                    step.run(*dsets[i])
                launches_per_step = int(lib.tac_launch_count()) - c0
                graphs.append(g)
        except Exception as e:                                           # capture unsupported -> eager timing, said so in config
            if rank == 0:
                print(f"[bench] CUDA-graph capture failed ({type(e).__name__}: {e}); timing eager launches", file=sys.stderr)
            graphs = None
            torch.cuda.synchronize()
    launch_label = "cuda-graph replay" if graphs is not None else "eager"     # fixed NOW: `graphs` is dropped before the extras
    if launches_per_step is None:
        c0 = int(lib.tac_launch_count())
        step.run(*dsets[0])
        launches_per_step = int(lib.tac_launch_count()) - c0
        torch.cuda.synchronize()

    def do_step(i):
        if graphs is not None:
            graphs[i % nsets].replay()
        else:
            step.run(*dsets[i % nsets])

    # ---- timed region: W warm-up steps, then exactly K steps between barrier + synchronize
    for i in range(args.warmup):
        do_step(i)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    step_evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    with ClockSampler(local) as clk:                  # started BEFORE the barrier: spawning the sampler takes milliseconds and
        barrier()                                     # would otherwise skew the ranks' entry into the timed region
        # N > 1: the hosts leave the barrier tens of microseconds to milliseconds apart; a GPU-side rendezvous in front of the
        # start event makes every rank's device enter step 1 together (else the first step measures that skew: its DPCA
        # chain waits for the slowest peer's statistics)
        step.device_barrier()
        e0.record()
        for i in range(args.steps):
            do_step(args.warmup + i)
            step_evs[i].record()                      # per-step spread (p50 / p95 below); the headline stays total / steps
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()
    ms_total = e0.elapsed_time(e1)
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    per_step = sorted(a.elapsed_time(b) for a, b in zip([e0] + step_evs[:-1], step_evs))
    pct = lambda q: per_step[min(len(per_step) - 1, int(q * len(per_step)))]
    spread = {"ms_per_step_p50": pct(0.50), "ms_per_step_p95": pct(0.95), "ms_per_step_max": per_step[-1]}
    audio_s_step = world * rows * T / SR
    value = audio_s_step / (ms_step * 1e-3)
    loss_val = float(step.loss.item())
    if world > 1:
        step.check_exchange()                         # a pop that gave up waiting would have produced garbage: fail loudly

    # ---- the same K steps on the same ring, EAGER (SURVEY 8d asks for both): ctypes marshalling + launches on the host's
    # clock; what a drop-in user who does not capture the step pays
    ms_eager = None
    if graphs is not None:
        for i in range(args.warmup):
            step.run(*dsets[i % nsets])
        barrier()
        step.device_barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for i in range(args.steps):
            step.run(*dsets[(args.warmup + i) % nsets])
        a1.record()
        torch.cuda.synchronize()
        te = torch.tensor([a0.elapsed_time(a1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        ms_eager = float(te.item()) / args.steps
    else:
        ms_eager = ms_step
    # public autograd API, eager, per call: tac_spectral.mrstft_loss(x, y).backward() (workspace allocation, ctypes, launches)
    api_ms = None
    try:
        import tac_spectral as _ts
        xa, ya = dsets[0][0], dsets[0][1].clone().requires_grad_(True)
        for _ in range(3):
            ya.grad = None
            _ts.mrstft_loss(xa, ya, N_FFTS).backward()
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        nrep = max(5, min(args.steps, 20))
        b0.record()
        for _ in range(nrep):
            ya.grad = None
            _ts.mrstft_loss(xa, ya, N_FFTS).backward()
        b1.record()
        torch.cuda.synchronize()
        api_ms = b0.elapsed_time(b1) / nrep
        del xa, ya
    except Exception as e:
        api_ms = f"{type(e).__name__}: {e}"

    # ---- attribution pass: the same K steps, eager, CUDA events around each leg on its launching stream
    legs = ["frontend", "loss_fused", "dpca_stats", "dpca_solve", "dpca_project"]
    acc = {k: 0.0 for k in legs}
    evs = []
    for i in range(args.steps):
        x, y, zs = dsets[(args.warmup + i) % nsets]
        rec = {}
        for name, fn in (("loss_fused", lambda: step.leg_loss_fused(x, y)), ("frontend", lambda: step.leg_frontend(x)),
                         ("dpca_stats", lambda: step.leg_dpca_stats(zs)),
                         ("dpca_solve", step.leg_dpca_solve), ("dpca_project", lambda: step.leg_dpca_project(zs))):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            rec[name] = (a, b)                         # rank-local legs: no push / pop, no collective
        evs.append(rec)
    torch.cuda.synchronize()
    for rec in evs:
        for k, (a, b) in rec.items():
            acc[k] += a.elapsed_time(b)
    leg_ms = {k: v / args.steps for k, v in acc.items()}
    peak, peak_src = load_peaks()
    byte_legs = {"frontend": alg.get("frontend", 0), "loss_fused": alg["loss_fused"],
                 "dpca_stats": alg["dpca_stats"], "dpca_project": alg["dpca_project"]}
    dom = max(("frontend", "loss_fused"), key=lambda k: leg_ms[k])
    kern_name = {"frontend": "stft_frontend_kernel<Plan2048>", "loss_fused": "mrstft_stream_kernel<true> + mrstft_stream_small_kernel<true> + mrstft_combine_kernel + mrstft_stream_sums_kernel (sums + grad_y in one pass over x, y)"}[dom]
    achieved = byte_legs[dom] / (leg_ms[dom] * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": kern_name, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": byte_legs[dom],
                "kernel_ms": leg_ms[dom],
                "per_leg": {k: {"ms": leg_ms[k], "bytes": byte_legs.get(k), "gbs": (byte_legs[k] / (leg_ms[k] * 1e-3) / 1e9) if byte_legs.get(k) else None}
                            for k in legs},
                "step_bytes": sum(byte_legs.values()), "step_gbs": sum(byte_legs.values()) / (ms_step * 1e-3) / 1e9,
                "step_frac": sum(byte_legs.values()) / (ms_step * 1e-3) / 1e9 / peak}
    # per-launch DRAM traffic and warp-instruction count of the dominant kernel, from the committed ncu capture
    # (profiles/traffic.json: {leg: {"traffic": dram bytes read + written, "inst": smsp__inst_executed.sum, "source": file}})
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof):
        try:
            ent = json.load(open(prof)).get(dom) or {}
            roofline["traffic"] = ent.get("traffic")
            if ent.get("inst"):
                sm_hz = (clk.summary().get("sm_mhz") or 1965) * 1e6
                n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
                # the kernel is FP32-issue bound, not HBM bound: issued warp-instructions / (4 per SM per clock)
                roofline["issue_frac"] = ent["inst"] / (leg_ms[dom] * 1e-3 * sm_hz * n_sm * 4)
                roofline["issue_note"] = "warp instructions per launch (ncu) / (4 schedulers x SMs x SM clock x kernel time)"
            roofline["traffic_source"] = ent.get("source")
            if ent.get("traffic"):
                roofline["hbm_traffic_frac"] = ent["traffic"] / (leg_ms[dom] * 1e-3) / 1e9 / peak
        except Exception:
            pass

    # "bound" keeps the contract's vocabulary (the kernel's NOMINAL roof: it moves 12 B per sample pair and no tensor work);
    # "binding_roof" names the roof that actually limits it: FP32 issue slots when the issue fraction exceeds both HBM fractions
    hbm_best = max(roofline["frac"], roofline.get("hbm_traffic_frac") or 0.0)
    roofline["binding_roof"] = "fp32-issue" if (roofline.get("issue_frac") or 0.0) > hbm_best else "hbm"
    if dom == "loss_fused":
        # SURVEY 8d counts fwd + bwd as two passes (20 B per sample); the fused kernel reads x, y once and writes grad_y (12 B):
        # `frac` is on the 12 B the kernel must move (of record), `frac_8d` on the survey's 20 B
        roofline["frac_8d"] = (20 * rows * T) / (leg_ms[dom] * 1e-3) / 1e9 / peak

    # ---- N > 1: the sharded, exchanged step against the whole batch on ONE GPU (rank 0), on hardware
    parity_n = None
    parity_cfg = "n/a (1 GPU)"
    if world > 1:
        parity_n = check_parity_n(step, dsets[0], w, dev, rank, world)
        if parity_n is not None:
            parity_cfg = (f"{parity_n['verdict']} vs whole batch on 1 GPU: loss_rel {parity_n['loss_rel']:.1e} rank_spread "
                          f"{parity_n['loss_rank_spread']:.1e} cnt_equal {parity_n['cnt_equal']} idx_equal {parity_n['idx_equal']} "
                          f"z_out_rel {parity_n['z_out_rel']:.1e} grad_rel {parity_n['grad_rel']:.1e}")

    # ---- e2e: public API, pinned HOST buffers; every step's inputs cross PCIe (H2D) and its loss is read back (D2H).
    # Two device input sets + two graphs: the copy of step i+1 (copy stream) overlaps the compute of step i, as a
    # training loop with a prefetching loader does; every step still pays its own H2D and D2H inside the timed region.
    hx, hy, hz = host_sets[0]
    px, py = hx.pin_memory(), hy.pin_memory()
    pz = [z.pin_memory() for z in hz]
    nbuf = 2
    # real data in the warm-up / capture pass (uninitialised memory may hold NaN bit patterns)
    dbuf = [(dsets[0][0].clone(), dsets[0][1].clone(), [z.clone() for z in dsets[0][2]]) for _ in range(nbuf)]
    g_e2e = [None] * nbuf
    if graphs is not None:
        for i in range(nbuf):
            step.run(*dbuf[i]); torch.cuda.synchronize()
            g_e2e[i] = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g_e2e[i]):
                step.run(*dbuf[i])
    copy_stream = torch.cuda.Stream(device=dev)
    main_stream = torch.cuda.current_stream(dev)
    h_loss = [torch.empty((), dtype=torch.float64).pin_memory() for _ in range(nbuf)]
    ev_copied = [torch.cuda.Event() for _ in range(nbuf)]
    ev_done = [torch.cuda.Event() for _ in range(nbuf)]
    ev_loss = [torch.cuda.Event() for _ in range(nbuf)]

    # two copy streams (target audio on one, decoded audio + latents on the other): one stream alone does not saturate the
    # link (tools/h2d_probe.py on this pool: 51 GB/s on one stream, 55 GB/s on two)
    copy_stream2 = torch.cuda.Stream(device=dev)
    ev_copied2 = [torch.cuda.Event() for _ in range(nbuf)]

    def enqueue_copy(i):
        b = i % nbuf
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(ev_done[b])                       # the step that last used this buffer has finished
            dbuf[b][0].copy_(px, non_blocking=True)
            ev_copied[b].record(copy_stream)
        with torch.cuda.stream(copy_stream2):
            copy_stream2.wait_event(ev_done[b])
            dbuf[b][1].copy_(py, non_blocking=True)
            for a_, b_ in zip(dbuf[b][2], pz):
                a_.copy_(b_, non_blocking=True)
            ev_copied2[b].record(copy_stream2)

    def enqueue_step(i):
        b = i % nbuf
        main_stream.wait_event(ev_copied[b])
        main_stream.wait_event(ev_copied2[b])
        if g_e2e[b] is not None:
            g_e2e[b].replay()
        else:
            step.run(*dbuf[b])
        h_loss[b].copy_(step.loss, non_blocking=True)
        ev_done[b].record(main_stream)
        ev_loss[b].record(main_stream)

    def e2e_loop(n):
        vals = []
        for b in range(nbuf):
            ev_done[b].record(main_stream)
        enqueue_copy(0)
        for i in range(n):
            if i + 1 < n:
                enqueue_copy(i + 1)
            enqueue_step(i)
            if i >= 1:                                               # read the previous step's loss on the host
                ev_loss[(i - 1) % nbuf].synchronize()
                vals.append(float(h_loss[(i - 1) % nbuf]))
        ev_loss[(n - 1) % nbuf].synchronize()
        vals.append(float(h_loss[(n - 1) % nbuf]))
        return vals

    e2e_loop(max(3, args.warmup))
    barrier()
    t0 = time.perf_counter()
    e2e_vals = e2e_loop(args.steps)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_val = audio_s_step / (float(dt.item()) / args.steps)
    h2d = 2 * rows * T * 4 + views * batch * d * 4
    e2e = {"value": e2e_val, "unit": "audio-s/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8,
           "ms_per_step": 1e3 * float(dt.item()) / args.steps, "pipeline": "double-buffered: H2D of step i+1 (two copy streams) overlaps compute of step i",
           "loss_last": e2e_vals[-1]}

    tgpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            t_ms = torch_gpu_baseline(dsets[0][0], dsets[0][1], dsets[0][2], dim_p)
            tgpu = {"value": audio_s_step / (t_ms * 1e-3), "unit": "audio-s/s", "ms_per_step": t_ms,
                    "what": "same step via torch.stft (cuFFT) + autograd + torch.svd (cuSOLVER) + cuBLAS on this GPU, inputs resident, eager"}
        except Exception as e:                                           # reported comparator only
            tgpu = {"error": f"{type(e).__name__}: {e}"}
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, t_step, sample = cpu_baseline(w, args.cpu_seconds, threads)
        cpu = {"value": v, "unit": "audio-s/s", "cores": threads, "kind": "port", "sample": sample, "ms_per_step": 1e3 * t_step}

    extra = None
    if not args.no_extra:
        graphs = g_e2e = None                                            # free the main workload's buffers first
        step.close()
        step = dsets = dbuf = None
        torch.cuda.empty_cache()
        extra = run_extras(args, dev, rank, world, group, peak, barrier)

    if rank == 0:
        line = {"metric": "audio-seconds/sec through STFT-loss+DPCA step", "value": value, "unit": "audio-s/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"{args.workload}: {w['desc']}", "clips_per_gpu": rows, "samples_per_clip": T,
                           "loss_scales": list(N_FFTS), "loss_kernel": "fused fwd+bwd (tac_mrstft_loss_fwd_bwd): reads x, y; writes grad_y", "front_end": "n_fft 2048 hop 512 centred, mag+phase [rows,2,1025,32] (unpadded)",
                           "dpca": {"views": views, "batch": batch, "d": d, "dim_p": dim_p, "gram": "tcgen05 3xTF32"},
                           "parallelism": f"dp{world} (batch shards; exchange of loss sums + DPCA statistics: {exch_desc})",
                           "launch": launch_label, "parity_n": parity_cfg,
                           "overlap": not args.no_overlap,
                           "l2": f"ring of {nsets} input sets ({nsets * set_bytes / 1e6:.0f} MB) > 126 MB L2"},
                "roofline": roofline, "cpu_baseline": cpu, "torch_gpu_baseline": tgpu, "e2e": e2e, "gpu_launches": launches_per_step * args.steps,
                "launches_per_step": launches_per_step, "clocks": clk.summary(), "loss": loss_val, **spread,
                "ms_per_step_eager": ms_eager, "eager_public_api": {"mrstft_loss_fwd_bwd_ms": api_ms, "what": "tac_spectral.mrstft_loss(x, y).backward(), eager, per call, cfg rows"},
                "parity_n": parity_n, "numa": numa, "extra": extra}
        print(json.dumps(line), flush=True)
    sys.stdout.flush()
    parity_failed = parity_n is not None and parity_n.get("verdict") != "ok"
    if parity_failed:
        print(f"[bench] N-rank parity MISMATCH: {parity_n}", file=sys.stderr, flush=True)
    if world > 1:
        # CUDA graphs that captured NCCL kernels must go before the communicator; never let teardown hang the run
        graphs = g_e2e = None
        torch.cuda.synchronize()
        killer = threading.Timer(20.0, lambda: os._exit(3 if parity_failed else 0))
        killer.daemon = True
        killer.start()
        try:
            dist.barrier(device_ids=[local])
            dist.destroy_process_group()
        except Exception:
            pass
        killer.cancel()
        os._exit(3 if parity_failed else 0)


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
