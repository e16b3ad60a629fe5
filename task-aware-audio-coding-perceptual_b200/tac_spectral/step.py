"""One training-step pass of the spectral hot path with every buffer pre-allocated (K1 + K2 + K3).

``SpectralStep`` is the allocation-free, host-sync-free form of what the reference does per batch in
``train_spectral_ae`` (``train_pkl_file_task_aware.py:597-629``) on this path:

* K1  front end  ``create_complex_spectrogram`` (``spectrogram.ipynb`` cell 10 L26-52) for every clip,
* K2  multi-resolution STFT loss of (target, decoded) and its gradient w.r.t. the decoded signal,
* K3  DPCA of the per-view latents (``pkl_file_model_updated.py:489-511`` / ``:398-427``),

launched through the C ABI on the caller's current stream, so that the whole step can be captured
in a CUDA graph (``torch.cuda.graph``) and replayed.  Multi-GPU (``group``): the per-scale loss sums
and the DPCA sufficient statistics are the step's only exchange; everything after it
(eigen-solve, selection, projection) runs redundantly on every rank on the global statistics, and the
input gradient needs no collective because the global normaliser is known from shapes.

``exchange="auto"`` (default) picks ``"p2p"`` when every rank of the group can map every other rank's buffer (one node,
at most 8 ranks) and ``"nccl"`` otherwise, with a warning.  ``exchange="p2p"``: the exchange is ONE-SHOT over NVLink peer memory and fused into the
kernels on either side of it (``exchange.py``, ``csrc/xchg.cuh``): the kernel that completes the statistics
peer-stores them into every rank's slot and raises a flag, the one-CTA-per-view tridiagonalisation spins on the
flags itself and adds the slots in rank order; the loss sums travel the same way inside the single-CTA kernel
that forms the loss scalar.  No kernel of the step waits for another rank except those two consumers, so rank
skew and collective latency stay off the spectral chain.  ``exchange="nccl"``: the packed float64 buffer is
all-reduced by NCCL in two slices (round-1 form, kept for A/B and as the fallback without peer access).

The loss backward and the front end do not depend on the all-reduce; with ``overlap=True`` they run
on a side stream next to the latency-bound DPCA chain (one CTA per view in the eigen-solve).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import os
import warnings

import torch

from . import _lib as L
from .loss import LOG_FLOOR
from .plan import StepPlan
from .stft import num_frames


class SpectralStep:
    def __init__(self, rows: int, T: int, n_views: int, batch: int, d: int, dim_p: int,
                 n_ffts: Sequence[int] = (1024, 2048, 512), hops: Optional[Sequence[int]] = None,
                 eps: float = LOG_FLOOR, frontend: bool = True, fe_n_fft: int = 2048, fe_hop: int = 512,
                 f_pad: int = 0, device="cuda", group=None, gram_algo: Optional[str] = None, overlap: bool = True,
                 exchange: Optional[str] = None):
        self.rows, self.T, self.nv, self.B, self.d, self.dim_p = int(rows), int(T), int(n_views), int(batch), int(d), int(dim_p)
        self.n_ffts = list(n_ffts)
        self.hops = [n // 4 for n in self.n_ffts] if hops is None else list(hops)
        self.S = len(self.n_ffts)
        self.eps = float(eps)
        self.frontend, self.fe_n_fft, self.fe_hop, self.f_pad = bool(frontend), int(fe_n_fft), int(fe_hop), int(f_pad)
        self.group, self.overlap = group, bool(overlap)
        self.device = torch.device(device)
        self.world = 1
        if group is not None:
            import torch.distributed as dist
            self.world = dist.get_world_size(group)
        exchange = exchange or os.environ.get("TAC_STEP_EXCHANGE", "auto")
        if exchange not in ("auto", "p2p", "nccl"):
            raise ValueError(f"unknown exchange {exchange!r} (auto | p2p | nccl)")
        # "auto": the peer exchange where it can exist (one node, <= XCHG_MAX_WORLD ranks, peer mapping succeeds on EVERY
        # rank -- PeerBuffer votes), NCCL otherwise, with one warning; an explicit "p2p" keeps the hard error
        self._exchange_auto = exchange == "auto"
        if self._exchange_auto:
            exchange = "p2p"
            if self.world > L.XCHG_MAX_WORLD:
                warnings.warn(f"SpectralStep: {self.world} ranks exceed the peer exchange's {L.XCHG_MAX_WORLD}; using the NCCL "
                              f"all-reduce for the step's exchange")
                exchange = "nccl"
        self.exchange = exchange if self.world > 1 else "none"
        self.xchg = None
        from .dpca import _algo_code
        self.algo = _algo_code(gram_algo, self.d)
        lib = L.lib()
        dev = self.device
        S, nv, d = self.S, self.nv, self.d
        self._nf, self._hp = L.int_array(self.n_ffts), L.int_array(self.hops)
        # ---- packed all-reduce payload: [2S loss sums | nv * (1 + d + d*d) DPCA statistics], float64 (plan.py)
        global_rows = self.rows * self.world
        if self.world > 1:
            # shards need not be equal (plan.shard_bounds hands out uneven ones): the loss normaliser uses the TRUE global
            # clip count, gathered once here; the DPCA count n travels in the statistics themselves
            t = torch.tensor([self.rows], dtype=torch.int64, device=dev)
            dist.all_reduce(t, group=group)
            global_rows = int(t.item())
        self.plan = StepPlan(self.rows, self.T, self.n_ffts, self.hops, nv, d, self.world, global_rows)
        self.stat_stride = self.plan.stat_stride
        self.packed = torch.zeros(self.plan.packed_len, dtype=torch.float64, device=dev)
        self.sums, self.stats = self.plan.views(self.packed)
        inv = self.plan.inv_counts()
        self.inv64 = torch.tensor(inv, dtype=torch.float64, device=dev)        # loss = dot(sums, inv64) (tac_loss_finalize)
        self.coef = torch.tensor(inv, dtype=torch.float32, device=dev)         # d loss / d sums (constant)
        self.loss = torch.zeros((), dtype=torch.float64, device=dev)
        # ---- workspaces and outputs
        self.ws_loss_bytes = int(lib.tac_mrstft_workspace_bytes(self.rows, T, S, self._nf, self._hp))
        self.ws_loss = torch.empty(max(self.ws_loss_bytes, 16), dtype=torch.uint8, device=dev)
        self.ws_dpca_bytes = int(lib.tac_dpca_workspace_bytes(nv, self.B, d))
        self.ws_dpca = torch.empty(max(self.ws_dpca_bytes, 256), dtype=torch.uint8, device=dev)
        self.grad_y = torch.empty((self.rows, T), dtype=torch.float32, device=dev)
        self.s = torch.empty((nv, d), dtype=torch.float32, device=dev)
        self.v = torch.empty((nv, d, d), dtype=torch.float32, device=dev)
        self.mu = torch.empty((nv, d), dtype=torch.float32, device=dev)
        self.rank = torch.empty(nv, dtype=torch.int32, device=dev)
        self.idx = torch.zeros((nv, d), dtype=torch.int32, device=dev)
        self.cnt = torch.zeros(nv, dtype=torch.int32, device=dev)
        self.z_out = [torch.empty((self.B, d), dtype=torch.float32, device=dev) for _ in range(nv)]
        self._zout_ptrs = L.ptr_array(self.z_out)
        if self.frontend:
            self.fe_F = num_frames(T, self.fe_n_fft, self.fe_hop, True)
            Fp = self.f_pad if self.f_pad > 0 else self.fe_F
            self.fe_out = torch.empty((self.rows, 2, self.fe_n_fft // 2 + 1, Fp), dtype=torch.float32, device=dev)
            self._fe_ptr = L.ptr_array([self.fe_out])
            self._fe_nf, self._fe_hp = L.int_array([self.fe_n_fft]), L.int_array([self.fe_hop])
        with torch.cuda.device(dev):
            for n in set(self.n_ffts):
                L.check(lib.tac_prepare_fft(n, n, 1))
            if self.frontend:
                L.check(lib.tac_prepare_fft(self.fe_n_fft, self.fe_n_fft, 0))
            # the DPCA chain is a string of small, latency-bound kernels: it gets a HIGH-PRIORITY stream so that its CTAs
            # are placed ahead of the thousands of pending loss-kernel CTAs whenever an SM slot frees up
            self.side = torch.cuda.Stream(device=dev, priority=-2) if self.overlap else None
            # the loss chain gets its own stream one priority level above the caller's: the front end (caller's stream) starts
            # first, under the statistics + collective; once the loss kernel is launched its CTAs take every SM slot that
            # frees up, and the rest of the short-CTA front end fills the loss kernel's last, partial wave
            self.fill = self.overlap and self.frontend and os.environ.get("TAC_STEP_FILL", "1") != "0"
            self.mid = torch.cuda.Stream(device=dev, priority=-1) if self.fill else None
            self._ev_stats = torch.cuda.Event()
        if self.exchange == "p2p":
            # last in the constructor: the rendezvous synchronises the ranks, so the first step's pops find peers that
            # have finished building their tables
            from .exchange import Exchange
            try:
                self.xchg = Exchange({"stats": nv * self.stat_stride, "sums": 2 * S, "sync": 2}, group, dev)
                self._sync_buf = torch.zeros(4, dtype=torch.float64, device=dev)
            except RuntimeError as e:                                # raised on ALL ranks or on none (PeerBuffer votes)
                if not self._exchange_auto:
                    raise
                warnings.warn(f"SpectralStep: no peer-mapped memory across the group ({e}); using the NCCL all-reduce for the "
                              f"step's exchange")
                self.exchange = "nccl"

    def device_barrier(self):
        """Enqueue a GPU-side rendezvous of all ranks on the current stream (one push + one pop of a two-value channel of the
        step's exchange): work enqueued behind it starts within microseconds of each other on every rank, however far apart
        the hosts left their last ``dist.barrier()``.  No-op without the peer exchange."""
        if self.xchg is None:
            return
        from .exchange import push, pop
        with torch.cuda.device(self.device):
            push(self.xchg.desc("sync"), self._sync_buf[:2])
            pop(self.xchg.desc("sync"), self._sync_buf[2:])

    def check_exchange(self):
        """Raise ``RuntimeError`` if a kernel of the peer exchange gave up waiting for another rank (``Exchange.check``;
        synchronises the device).  The wait is bounded by ``TAC_XCHG_TIMEOUT_S`` (default 600 s): ranks that pause for
        longer between steps -- rank-0 checkpointing, validation, a stalled loader -- should meet in ``dist.barrier()``
        (or ``device_barrier()`` behind one) before the next step.  No-op with the NCCL exchange."""
        if self.xchg is not None:
            self.xchg.check()

    def close(self):
        if self.xchg is not None:
            self.xchg.close()
            self.xchg = None

    # ------------------------------------------------------------------ individual legs (current stream)
    def _check(self, x, y, zs):
        if tuple(x.shape) != (self.rows, self.T) or tuple(y.shape) != (self.rows, self.T):
            raise ValueError(f"Shape mismatch: expected x, y [{self.rows}, {self.T}], got {tuple(x.shape)}, {tuple(y.shape)}")
        if len(zs) != self.nv or any(tuple(z.shape) != (self.B, self.d) for z in zs):
            raise ValueError(f"expected {self.nv} latent views of shape [{self.B}, {self.d}]")
        for t in (x, y, *zs):
            L.require_cuda(t, "input")
            if t.dtype != torch.float32 or not t.is_contiguous():
                raise ValueError("inputs must be contiguous float32")

    def leg_frontend(self, x):
        st = torch.cuda.current_stream(self.device).cuda_stream
        L.check(L.lib().tac_stft_multi(x.data_ptr(), self.rows, 1, self.T, 1, self._fe_nf, self._fe_hp, self._fe_nf, 1,
                                       L.PAD_ZERO, 0, L.OUT_MAGPHASE, self.f_pad, self._fe_ptr, st))

    def leg_loss_fwd(self, x, y):
        st = torch.cuda.current_stream(self.device).cuda_stream
        L.check(L.lib().tac_mrstft_loss_fwd(x.data_ptr(), y.data_ptr(), self.rows, self.T, self.S, self._nf, self._hp, self._nf,
                                            self.eps, self.sums.data_ptr(), self.ws_loss.data_ptr(), self.ws_loss_bytes, st))

    def leg_loss_bwd(self, x, y):
        st = torch.cuda.current_stream(self.device).cuda_stream
        L.check(L.lib().tac_mrstft_loss_bwd(x.data_ptr(), y.data_ptr(), self.rows, self.T, self.S, self._nf, self._hp, self._nf,
                                            self.eps, self.coef.data_ptr(), self.grad_y.data_ptr(), None, 0, st))

    def leg_loss_fused(self, x, y):
        """Sums AND gradient in one pass (``tac_mrstft_loss_fwd_bwd``): d loss / d sums is the constant ``coef``."""
        st = torch.cuda.current_stream(self.device).cuda_stream
        L.check(L.lib().tac_mrstft_loss_fwd_bwd(x.data_ptr(), y.data_ptr(), self.rows, self.T, self.S, self._nf, self._hp, self._nf,
                                                self.eps, self.coef.data_ptr(), self.sums.data_ptr(), self.grad_y.data_ptr(),
                                                self.ws_loss.data_ptr(), self.ws_loss_bytes, st))

    def _x(self, name, on):
        return self.xchg.desc(name) if (on and self.xchg is not None) else None

    def leg_dpca_stats(self, zs, push=False):
        """Local statistics; ``push``: the finishing kernel also peer-stores them to every rank (p2p exchange)."""
        st = torch.cuda.current_stream(self.device).cuda_stream
        L.check(L.lib().tac_dpca_stats_push(L.ptr_array(zs), self.nv, self.B, self.d, self.algo, self.stats.data_ptr(),
                                            self.ws_dpca.data_ptr(), self.ws_dpca_bytes, self._x("stats", push), st))

    def leg_allreduce(self, what="all"):
        """The step's exchange: the packed float64 buffer [loss sums | DPCA statistics].  It is reduced in two
        slices so that the eigen-solve (needs the statistics only) overlaps the loss kernel."""
        if self.group is not None and self.world > 1:
            self.plan.allreduce(self.packed, self.group, what)

    def leg_dpca_solve(self, pop=False):
        """Eigen-solve + selection; ``pop``: the tridiagonalisation kernel first waits for and sums the world's pushes."""
        st = torch.cuda.current_stream(self.device).cuda_stream
        lib = L.lib()
        L.check(lib.tac_dpca_eig_pop(self.stats.data_ptr(), self.nv, self.d, self.s.data_ptr(), self.v.data_ptr(), self.mu.data_ptr(),
                                     self.rank.data_ptr(), self.ws_dpca.data_ptr(), self.ws_dpca_bytes, self._x("stats", pop), st))
        L.check(lib.tac_dpca_select(self.s.data_ptr(), self.rank.data_ptr(), self.nv, self.d, self.dim_p, self.idx.data_ptr(),
                                    self.cnt.data_ptr(), st))

    def leg_dpca_project(self, zs):
        st = torch.cuda.current_stream(self.device).cuda_stream
        L.check(L.lib().tac_dpca_project(L.ptr_array(zs), self.nv, self.B, self.d, self.v.data_ptr(), self.mu.data_ptr(),
                                         self.idx.data_ptr(), self.cnt.data_ptr(), self._zout_ptrs, self.ws_dpca.data_ptr(),
                                         self.ws_dpca_bytes, st))

    def leg_loss_value(self, exchange=False):
        """loss = sums . inv in one single-CTA kernel; ``exchange``: the sums of all ranks are exchanged in it first."""
        st = torch.cuda.current_stream(self.device).cuda_stream
        L.check(L.lib().tac_loss_finalize(self.sums.data_ptr(), 2 * self.S, self.inv64.data_ptr(), self._x("sums", exchange),
                                          self.loss.data_ptr(), st))

    # ------------------------------------------------------------------ the step
    def run(self, x: torch.Tensor, y: torch.Tensor, zs: List[torch.Tensor]) -> torch.Tensor:
        """Enqueue one step; returns the (device, float64) loss scalar of the global batch.  Results:
        ``self.loss``, ``self.grad_y``, ``self.fe_out``, ``self.z_out``, ``self.cnt`` / ``self.idx``.

        Schedule: statistics (+ push) run first, on the high-priority side stream, with the short-CTA front-end
        kernel underneath; then the eigen-solve chain (1 CTA per view) runs NEXT TO the loss kernel that fills the
        other SMs.  With the p2p exchange the loss kernel waits for the LOCAL statistics only: the tridiagonalisation
        CTAs are resident by then and do the waiting for the peers themselves."""
        self._check(x, y, zs)
        p2p = self.exchange == "p2p"
        with torch.cuda.device(self.device):
            main = torch.cuda.current_stream(self.device)
            if self.overlap:
                self.side.wait_stream(main)
                with torch.cuda.stream(self.side):
                    self.leg_dpca_stats(zs, push=p2p)
                    if not p2p:
                        self.leg_allreduce("stats")
                    self._ev_stats.record(self.side)
                    self.leg_dpca_solve(pop=p2p)
                    self.leg_dpca_project(zs)
                # the front end's CTAs are short (~12 us): it runs UNDER the statistics + collective, whose kernels get the
                # next free SMs through the side stream's priority; the long-CTA loss kernel must not flood the SMs before
                # the collective has run
                if self.fill:
                    self.mid.wait_stream(main)
                    self.leg_frontend(x)
                    with torch.cuda.stream(self.mid):
                        # the loss kernel waits for the statistics (+ collective) even at one GPU: launched at t = 0 its CTAs
                        # keep the one-CTA-per-SM tridiagonalisation off the SMs until its tail (measured 0.567 -> 0.620 ms)
                        self.mid.wait_event(self._ev_stats)
                        self._spectral_chain(x, y, frontend=False)
                    main.wait_stream(self.mid)
                else:
                    if self.frontend:
                        self.leg_frontend(x)
                    main.wait_event(self._ev_stats)
                    self._spectral_chain(x, y, frontend=False)
                main.wait_stream(self.side)
            else:
                self._dpca_chain(zs)
                self._spectral_chain(x, y)
        return self.loss

    def _spectral_chain(self, x, y, frontend=True):
        self.leg_loss_fused(x, y)
        if self.frontend and frontend:
            self.leg_frontend(x)
        p2p = self.exchange == "p2p"
        if not p2p:
            self.leg_allreduce("sums")
        self.leg_loss_value(exchange=p2p)

    def _dpca_chain(self, zs):
        p2p = self.exchange == "p2p"
        self.leg_dpca_stats(zs, push=p2p)
        if not p2p:
            self.leg_allreduce("stats")
        self.leg_dpca_solve(pop=p2p)
        self.leg_dpca_project(zs)

    def algorithmic_bytes(self) -> dict:
        """SURVEY.md 8d byte counts per leg (per rank)."""
        r, T, B, d, nv = self.rows, self.T, self.B, self.d, self.nv
        out = {"loss_fwd": 8 * r * T, "loss_bwd": 12 * r * T, "loss_fused": 12 * r * T,
               "dpca_stats": 4 * B * d * nv, "dpca_project": 8 * B * d * nv + 4 * d * self.dim_p}
        if self.frontend:
            Fp = self.f_pad if self.f_pad > 0 else self.fe_F
            out["frontend"] = 4 * r * T + 8 * r * (self.fe_n_fft // 2 + 1) * Fp
        return out
