"""Host-side mirror of the DTAC distributed-PCA (DPCA) latent projection (K3).

Same names, argument meaning and results as the reference call sites:

* :func:`data_pca` -- ``pkl_file_model_updated.py:37-42`` / ``pkl_file_models.py:51-56``: returns
  ``(s[r], v[d, r], mean[d])`` with ``r = min(B, d)``, ``s`` descending, singular values of the
  centred batch (no ``1/(B-1)``).
* :func:`dpca_project` -- the select / project block of ``ResE2D1.forward``
  (``pkl_file_model_updated.py:489-511``) and ``ResE4D1.forward`` (``:398-427``) as a function of
  the per-view latents and ``dim_p`` (the host draw ``torch.randint(8, num_features/2)`` of
  ``:489`` / ``:398`` stays in the caller so the RNG stream is unchanged).
* :func:`apply_pca` -- ``pkl_file_models.py:224-228``;  :func:`var_topk` -- ``:901-907``.

Device flow (no host synchronisation anywhere): ``tac_dpca_stats`` (sufficient statistics
``[n, sum z, sum z z^T]`` per view, float64) -> optional all-reduce over the data-parallel group
-> ``tac_dpca_eig`` (Jacobi, every rank redundantly) -> ``tac_dpca_select`` -> ``tac_dpca_project``.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib as L

MAX_D = 256
MAX_VIEWS = 8
TC_MAX_D = 128


def _algo_code(algo: Optional[str], d: int) -> int:
    if algo is None:
        return L.GRAM_TF32X3 if d <= TC_MAX_D else L.GRAM_FP64
    if algo == "tf32x3":
        return L.GRAM_TF32X3
    if algo == "fp64":
        return L.GRAM_FP64
    raise ValueError(f"unknown Gram algorithm {algo!r} (tf32x3 | fp64)")


def _check_views(zs: Sequence[torch.Tensor]) -> Tuple[int, int, int]:
    if not 1 <= len(zs) <= MAX_VIEWS:
        raise ValueError(f"need 1..{MAX_VIEWS} views, got {len(zs)}")
    B, d = zs[0].shape
    for i, z in enumerate(zs):
        L.require_cuda(z, f"z[{i}]")
        if z.dim() != 2 or tuple(z.shape) != (B, d):
            raise ValueError(f"view {i}: expected [{B}, {d}], got {tuple(z.shape)}")
    if d > MAX_D:
        raise ValueError(f"latent width {d} > {MAX_D} unsupported")
    return len(zs), B, d


def _workspace(nv: int, B: int, d: int, device) -> Tuple[torch.Tensor, int]:
    n = int(L.lib().tac_dpca_workspace_bytes(nv, B, d))
    return torch.empty(max(n, 256), dtype=torch.uint8, device=device), n


def dpca_stats(zs: Sequence[torch.Tensor], algo: Optional[str] = None) -> torch.Tensor:
    """Sufficient statistics ``[nv, 1 + d + d*d]`` float64 of the local shard (sum-reducible)."""
    nv, B, d = _check_views(zs)
    zc = [z.detach().contiguous().float() for z in zs]
    stats = torch.empty((nv, 1 + d + d * d), dtype=torch.float64, device=zc[0].device)
    ws, n = _workspace(nv, B, d, zc[0].device)
    with L.device_of(zc[0]):
        L.check(L.lib().tac_dpca_stats(L.ptr_array(zc), nv, B, d, _algo_code(algo, d), stats.data_ptr(),
                                       ws.data_ptr(), n, L.stream_of(zc[0])))
    return stats


def dpca_eig(stats: torch.Tensor, d: int):
    """``(s [nv, d], v [nv, d, d], mu [nv, d], rank [nv])`` from (all-reduced) statistics."""
    nv = stats.shape[0]
    dev = stats.device
    s = torch.empty((nv, d), dtype=torch.float32, device=dev)
    v = torch.empty((nv, d, d), dtype=torch.float32, device=dev)
    mu = torch.empty((nv, d), dtype=torch.float32, device=dev)
    rank = torch.empty(nv, dtype=torch.int32, device=dev)
    ws, n = _workspace(nv, 1, d, dev)
    with L.device_of(stats):
        L.check(L.lib().tac_dpca_eig(stats.data_ptr(), nv, d, s.data_ptr(), v.data_ptr(), mu.data_ptr(),
                                     rank.data_ptr(), ws.data_ptr(), n, L.stream_of(stats)))
    return s, v, mu, rank


def _allreduce_stats(stats: torch.Tensor, group) -> torch.Tensor:
    if group is not None:
        import torch.distributed as dist
        dist.all_reduce(stats, group=group)          # NCCL over NVLink: <= 0.53 MB of float64 per step
    return stats


def data_pca(z: torch.Tensor, algo: Optional[str] = None, group=None):
    """Drop-in for the reference's ``data_pca(z)``: ``(s, v, mean)``.

    The returned triple is DETACHED (the reference's is differentiable through ``torch.svd``): use :func:`dpca_project`
    / :func:`apply_pca`, whose backward carries the basis term, when gradients must reach ``z``.  A ``z`` that requires
    grad therefore raises a warning here instead of silently cutting the graph."""
    _, B, d = _check_views([z])
    if torch.is_grad_enabled() and z.requires_grad:
        import warnings
        warnings.warn("tac_spectral.data_pca returns detached (s, v, mean); gradients through the PCA basis are "
                      "implemented in dpca_project / apply_pca (basis_grad=True)", stacklevel=2)
    stats = _allreduce_stats(dpca_stats([z], algo), group)
    s, v, mu, _ = dpca_eig(stats, d)
    n = B
    if group is not None:
        n = int(stats[0, 0].item())                     # global batch (shards need not be equal); one host read
    r = min(n, d)                                       # torch.svd(some=True) returns min(B, d) columns
    return s[0, :r], v[0, :, :r], mu[0]


def basis_grad_term(zc: torch.Tensor, A: torch.Tensor, V: torch.Tensor, s: torch.Tensor, sel: torch.Tensor) -> torch.Tensor:
    """The part of ``d loss / d z`` that flows through the PCA BASIS (the reference back-propagates through
    ``torch.svd`` in ``data_pca``, ``pkl_file_model_updated.py:37-42,503-511``): with ``C = zc^T zc = V diag(s^2) V^T``
    and the spectral projector ``P = sum_{i in sel} v_i v_i^T``,

        dP = V (K o (V^T dC V)) V^T,   K_ij = (m_i - m_j) / (s_i^2 - s_j^2)  (m = selection mask; 0 where m_i == m_j)

    so for ``A = zc^T g`` (``g`` = upstream gradient; summed over ranks when the batch is sharded) ``Cbar = V (K o V^T A V)
    V^T`` and the term is ``zc (Cbar + Cbar^T)``.  Null directions of a rank-deficient batch (zero columns of ``V``, as
    ``tac_dpca_eig`` emits them) enter through their projector ``N = I - V V^T``: ``Cbar += W A N + N A W`` with
    ``W = V_sel diag(1 / s^2) V_sel^T``.  Plain torch algebra on ``d x d`` matrices (autograd glue, float64)."""
    dt = torch.float64
    Vd, Ad, zd = V.to(dt), A.to(dt), zc.to(dt)
    lam = s.to(dt) ** 2
    m = sel.to(dt)
    valid = (Vd * Vd).sum(0) > 0.25                                   # real eigenvectors (null columns are zero)
    dm = m[:, None] - m[None, :]
    diff = lam[:, None] - lam[None, :]
    pair = (dm != 0) & valid[:, None] & valid[None, :]
    K = torch.where(pair, dm / torch.where(pair, diff, torch.ones_like(diff)), torch.zeros_like(diff))
    M = Vd.T @ Ad @ Vd
    Cbar = Vd @ (K * M) @ Vd.T
    Vv = Vd * valid.to(dt)
    N = torch.eye(Vd.shape[0], dtype=dt, device=V.device) - Vv @ Vv.T
    w = torch.where((m > 0) & valid & (lam > 0), 1.0 / lam.clamp_min(1e-300), torch.zeros_like(lam))
    W = (Vd * w) @ Vd.T
    Cbar = Cbar + W @ Ad @ N + N @ Ad @ W
    return (zd @ (Cbar + Cbar.T)).to(zc.dtype)


class _Project(torch.autograd.Function):
    """``z_out_i = (z_i - mu_i) P_i + mu_i`` for all views in one launch chain.  Backward:
    ``grad_z = (g - mean g) P + mean g`` (same kernel; the dependence of the mean on ``z`` is kept) plus, with
    ``basis_grad`` (default, the reference's semantics), the term through the eigenvectors (:func:`basis_grad_term`)."""

    @staticmethod
    def forward(ctx, s, v, mu, idx, cnt, group, basis_grad, *zs):
        nv, B, d = len(zs), zs[0].shape[0], zs[0].shape[1]
        zc = [z.contiguous().float() for z in zs]
        outs = [torch.empty_like(z) for z in zc]
        ws, n = _workspace(nv, B, d, zc[0].device)
        with L.device_of(zc[0]):
            L.check(L.lib().tac_dpca_project(L.ptr_array(zc), nv, B, d, v.data_ptr(), mu.data_ptr(), idx.data_ptr(),
                                             cnt.data_ptr(), L.ptr_array(outs), ws.data_ptr(), n, L.stream_of(zc[0])))
        ctx.basis_grad = bool(basis_grad)
        ctx.save_for_backward(s, v, mu, idx, cnt, *(zc if basis_grad else ()))
        ctx.group = group
        ctx.dtypes = [z.dtype for z in zs]
        return tuple(outs)

    @staticmethod
    def backward(ctx, *gs):
        s, v, mu, idx, cnt, *zsaved = ctx.saved_tensors
        nv = len(gs)
        gc = [g.contiguous().float() for g in gs]
        B, d = gc[0].shape
        gsum = torch.stack([g.sum(0) for g in gc])
        n = torch.tensor(float(B), device=gsum.device)
        if ctx.group is not None:
            import torch.distributed as dist
            dist.all_reduce(gsum, group=ctx.group)
            n = n * dist.get_world_size(ctx.group)
        gmean = (gsum / n).contiguous()
        outs = [torch.empty_like(g) for g in gc]
        ws, nb = _workspace(nv, B, d, gc[0].device)
        with L.device_of(gc[0]):
            L.check(L.lib().tac_dpca_project(L.ptr_array(gc), nv, B, d, v.data_ptr(), gmean.data_ptr(), idx.data_ptr(),
                                             cnt.data_ptr(), L.ptr_array(outs), ws.data_ptr(), nb, L.stream_of(gc[0])))
        if ctx.basis_grad:
            A = torch.stack([(z - mu[i]).T @ g for i, (z, g) in enumerate(zip(zsaved, gc))])      # [nv, d, d]
            if ctx.group is not None:
                import torch.distributed as dist
                dist.all_reduce(A, group=ctx.group)
            ar = torch.arange(d, device=idx.device)
            for i in range(nv):
                sel = torch.zeros(d, dtype=torch.float32, device=idx.device)
                sel.scatter_add_(0, idx[i].long(), (ar < cnt[i]).float())              # selection mask, no host sync
                outs[i] = outs[i] + basis_grad_term(zsaved[i] - mu[i], A[i], v[i], s[i], sel)
        return (None, None, None, None, None, None, None) + tuple(o.to(dt) for o, dt in zip(outs, ctx.dtypes))


def dpca_select(s: torch.Tensor, rank: torch.Tensor, dim_p: int):
    """Global top-``dim_p`` across the concatenated singular values; ``(idx [nv, d] int32, cnt [nv] int32)``."""
    nv, d = s.shape
    idx = torch.zeros((nv, d), dtype=torch.int32, device=s.device)
    cnt = torch.empty(nv, dtype=torch.int32, device=s.device)
    with L.device_of(s):
        L.check(L.lib().tac_dpca_select(s.data_ptr(), rank.data_ptr(), nv, d, int(dim_p), idx.data_ptr(), cnt.data_ptr(),
                                        L.stream_of(s)))
    return idx, cnt


def dpca_project(zs: Sequence[torch.Tensor], dim_p: int, algo: Optional[str] = None, group=None,
                 return_info: bool = False, basis_grad: bool = True):
    """The DPCA block: per-view PCA, global top-``dim_p`` selection over all views' singular values,
    per-view projection onto the selected components and back (``pkl_file_model_updated.py:489-511``).

    With ``group`` the statistics are all-reduced so that all ranks use the PCA of the global batch.
    ``basis_grad`` (default True): gradients flow through the PCA basis as in the reference, which back-propagates
    through ``torch.svd``; ``False`` = stop-gradient on the projector (cheaper; only the centring is differentiated)."""
    nv, B, d = _check_views(zs)
    if dim_p < 0:
        raise ValueError("dim_p must be >= 0")
    stats = _allreduce_stats(dpca_stats(zs, algo), group)
    s, v, mu, rank = dpca_eig(stats, d)
    idx, cnt = dpca_select(s, rank, dim_p)
    outs = list(_Project.apply(s, v, mu, idx, cnt, group, basis_grad, *zs))
    if return_info:
        return outs, {"s": s, "v": v, "mu": mu, "rank": rank, "idx": idx, "cnt": cnt, "stats": stats}
    return outs


def apply_pca(z: torch.Tensor, dim_p: int, algo: Optional[str] = None, group=None, basis_grad: bool = True) -> torch.Tensor:
    """``apply_pca`` (``pkl_file_models.py:224-228``): rank-``dim_p`` reconstruction of one view."""
    return dpca_project([z], dim_p, algo, group, basis_grad=basis_grad)[0]


def var_topk(z: torch.Tensor, k: int, group=None):
    """E1D1 bottleneck (``pkl_file_models.py:901-907``): ``(z[:, idx], idx)`` with ``idx`` the ``k`` features of
    largest unbiased variance over the batch (the diagonal of the DPCA covariance / (n - 1))."""
    _, B, d = _check_views([z])
    if not 1 <= k <= d:
        raise ValueError(f"k={k} must be in [1, {d}]")
    stats = _allreduce_stats(dpca_stats([z], "fp64"), group)
    var = torch.empty(d, dtype=torch.float32, device=z.device)
    top = torch.empty(k, dtype=torch.int32, device=z.device)
    with L.device_of(z):
        L.check(L.lib().tac_var_topk(stats.data_ptr(), d, int(k), var.data_ptr(), top.data_ptr(), L.stream_of(z)))
    idx = top.long()
    return z.index_select(1, idx), idx


class _Nuclear(torch.autograd.Function):
    """``|zc / |zc|_F|_* / B`` with ``zc = z - mean``: forward from the eigen-solve of the Gram matrix, backward
    ``grad_z = g * (z - mean) Gm`` with ``Gm = (V diag(1/s) V^T - (sum s / |zc|_F^2) I) / (|zc|_F B)``."""

    @staticmethod
    def forward(ctx, z, algo, group):
        zc = z.contiguous().float()
        B, d = zc.shape
        stats = _allreduce_stats(dpca_stats([zc], algo), group)
        s, v, mu, rank = dpca_eig(stats, d)
        out = torch.empty(1, dtype=torch.float32, device=zc.device)
        aux = torch.empty(3, dtype=torch.float64, device=zc.device)
        with L.device_of(zc):
            L.check(L.lib().tac_nuclear_from_s(s.data_ptr(), rank.data_ptr(), stats.data_ptr(), 1, d, out.data_ptr(),
                                               aux.data_ptr(), L.stream_of(zc)))
        ctx.save_for_backward(zc, s, v, mu, rank, aux)
        ctx.meta = (z.shape, z.dtype)
        return out[0]

    @staticmethod
    def backward(ctx, g):
        zc, s, v, mu, rank, aux = ctx.saved_tensors
        B, d = zc.shape
        gm = torch.empty((1, d, d), dtype=torch.float32, device=zc.device)
        grad = torch.empty_like(zc)
        lib = L.lib()
        with L.device_of(zc):
            st = L.stream_of(zc)
            L.check(lib.tac_nuclear_grad_matrix(s.data_ptr(), v.data_ptr(), rank.data_ptr(), aux.data_ptr(), 1, d, gm.data_ptr(), st))
            L.check(lib.tac_dpca_apply(L.ptr_array([zc]), 1, B, d, gm.data_ptr(), mu.data_ptr(), 0, L.ptr_array([grad]), st))
        shape, dtype = ctx.meta
        return (grad * g.to(grad.dtype)).reshape(shape).to(dtype), None, None


def nuclear_loss(z_sample: torch.Tensor, algo: Optional[str] = None, group=None) -> torch.Tensor:
    """The reference's latent regulariser (``pkl_file_model_updated.py:526-532``, ``pkl_file_models.py:943-947``):
    centre, divide by the Frobenius norm, nuclear norm / batch size -- from the Gram eigen-solve that DPCA already
    runs (``sum sqrt(eig) / sqrt(sum eig) / B``) instead of a second SVD.  ``z_sample`` is ``[B, d]``, ``d <= 256``
    (the concatenated views of ``torch.cat((z1, z2), dim=1)``); with ``group`` the batch of all ranks is used."""
    L.require_cuda(z_sample, "z_sample")
    if z_sample.dim() != 2:
        raise ValueError(f"expected z_sample of shape [B, d], got {tuple(z_sample.shape)}")
    if z_sample.shape[1] > 256:
        raise ValueError("nuclear_loss supports d <= 256")
    return _Nuclear.apply(z_sample, algo, group)


class _CosinePairs(torch.autograd.Function):
    """``mean_b sum_{i<j} cos(z_i[b], z_j[b])``: one pass over the views forward, one pass backward."""

    @staticmethod
    def forward(ctx, eps, group, *zs):
        zc = [z.contiguous().float() for z in zs]
        nv, B, d = _check_views(zc)
        dev = zc[0].device
        lib = L.lib()
        ws_bytes = int(lib.tac_cosine_pairs_workspace_bytes(B))
        ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=dev)
        acc = torch.zeros(2, dtype=torch.float64, device=dev)             # [sum of cosines, local batch]
        acc[1] = float(B)
        with L.device_of(zc[0]):
            L.check(lib.tac_cosine_pairs_fwd(L.ptr_array(zc), nv, B, d, float(eps), acc.data_ptr(), ws.data_ptr(), ws_bytes,
                                             L.stream_of(zc[0])))
        acc = _allreduce_stats(acc, group)                                 # batch shards: both entries are sums
        ctx.save_for_backward(acc, *zc)
        ctx.eps = float(eps)
        ctx.meta = [(z.shape, z.dtype) for z in zs]
        return (acc[0] / acc[1]).float()

    @staticmethod
    def backward(ctx, g):
        acc, *zc = ctx.saved_tensors
        nv, B, d = len(zc), zc[0].shape[0], zc[0].shape[1]
        grads = [torch.empty_like(z) for z in zc]
        gs = (g.float() / acc[1].float()).reshape(1).contiguous()          # g / global batch, stays on the device
        with L.device_of(zc[0]):
            L.check(L.lib().tac_cosine_pairs_bwd(L.ptr_array(zc), nv, B, d, ctx.eps, gs.data_ptr(), 1.0, L.ptr_array(grads),
                                                 L.stream_of(zc[0])))
        return (None, None) + tuple(gr.reshape(sh).to(dt) for gr, (sh, dt) in zip(grads, ctx.meta))


def pairwise_cosine_loss(zs: Sequence[torch.Tensor], eps: float = 1e-8, group=None) -> torch.Tensor:
    """``torch.mean(cos_sim(z1, z2) + cos_sim(z1, z3) + ...)`` over all view pairs with ``cos_sim =
    torch.nn.CosineSimilarity()`` (``pkl_file_models.py:390-391,761-763``; ``pkl_file_model_updated.py:429-430,513-514``)
    for 2, 3 or 4 views ``[B, d]``; with ``group`` the mean runs over the batch of all ranks."""
    for i, z in enumerate(zs):
        L.require_cuda(z, f"zs[{i}]")
    if len(zs) not in (2, 3, 4):
        raise ValueError(f"pairwise_cosine_loss takes 2, 3 or 4 views, got {len(zs)}")
    return _CosinePairs.apply(eps, group, *zs)
