"""CPU restatement (numpy, float64) of the algebra the packed-complex FFT core relies on
(`csrc/fft_core_pk.cuh`): the three-instruction twiddled DIT butterfly in both factorisations, the in-place
bit-reversed-input network, the factorised inter-pass twiddles, and the real-FFT post-processing with conjugated odd
slots together with its adjoint.  The CUDA kernels are checked against the oracle in the `-m gpu` tests; this file pins
the formulas their comments cite, so that a change of the algebra fails here first."""
import numpy as np
import pytest


def rotm(v):            # -j v = (im, -re)
    return np.complex128(complex(v.imag, -v.real))


def brev(k, n):
    bits = n.bit_length() - 1
    return int(format(k, "0%db" % bits)[::-1], 2) if bits else 0


def dit_bfly(a, b, N, K, inv):
    """a' = a + W b, b' = a - W b, W = exp(-+ 2 pi j K / N), as fft_core_pk.cuh:dit_bfly forms it."""
    if K == 0:
        return a + b, a - b
    if 4 * K == N:
        r = -rotm(b) if inv else rotm(b)
        return a + r, a - r
    c, s = np.cos(2 * np.pi * K / N), np.sin(2 * np.pi * K / N)
    sg = -1.0 if inv else 1.0
    if abs(c) >= s:
        t = sg * s / c
        u = b + t * rotm(b)                      # fma2(rot(b), t, b)
        return a + c * u, a - c * u              # fma2(u, +-c, a)
    r = -sg * c / s
    u = b + r * rotm(b)
    ru = rotm(u)
    return a + (sg * s) * ru, a - (sg * s) * ru


def dit(z, inv=False):
    """In-place network: slot brev(i) holds x[i] on entry, slot k holds X[k] on return."""
    z = list(z)
    n = len(z)

    def rec(off, N):
        if N < 2:
            return
        rec(off, N // 2)
        rec(off + N // 2, N // 2)
        for k in range(N // 2):
            z[off + k], z[off + k + N // 2] = dit_bfly(z[off + k], z[off + k + N // 2], N, k, inv)
    rec(0, n)
    return np.array(z)


@pytest.mark.parametrize("n", [2, 4, 8, 16, 32])
def test_dit_network_matches_fft(n):
    rng = np.random.default_rng(n)
    x = rng.standard_normal(n) + 1j * rng.standard_normal(n)
    slots = np.empty(n, dtype=np.complex128)
    for i in range(n):
        slots[brev(i, n)] = x[i]
    assert np.allclose(dit(slots), np.fft.fft(x), rtol=1e-12, atol=1e-12)
    assert np.allclose(dit(slots, inv=True), np.fft.ifft(x) * n, rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("A,BQ", [(32, 32), (16, 32), (16, 16), (8, 16), (8, 8), (4, 8)])
def test_factorised_interpass_twiddles(A, BQ):
    """W_M^(b k1) with k1 = 8 q + r equals W_M^(b r) W_M^(8 b q): the 7 + A/8 - 1 table reads of warp_fft_fwd_pk; the adjoint
    uses the same split on b."""
    M = A * BQ
    tw = np.exp(-2j * np.pi * (np.arange(A)[:, None] * np.arange(BQ)[None, :] % M) / M)      # tables.cu: tw[k1][b]
    for k1 in range(A):
        q, r = divmod(k1, 8)
        assert np.allclose(tw[r] * tw[8 * q] if q else tw[r], tw[k1], atol=1e-14)
    for b in range(BQ):
        q, r = divmod(b, 8)
        assert np.allclose(tw[:, r] * tw[:, 8 * q] if q else tw[:, r], tw[:, b], atol=1e-14)


def two_pass_fft(zp, A, BQ):
    """M = A x BQ decomposition of the core: pass 1 over a (stride BQ), twiddle, pass 2 over b -> Z[k1 + A k2]."""
    M = A * BQ
    Z = np.empty(M, dtype=np.complex128)
    t = np.empty((A, BQ), dtype=np.complex128)
    for b in range(BQ):
        col = np.array([zp[BQ * a + b] for a in range(A)])
        t[:, b] = np.fft.fft(col) * np.exp(-2j * np.pi * b * np.arange(A) / M)
    for k1 in range(A):
        Z[k1 + A * np.arange(BQ)] = np.fft.fft(t[k1])
    return Z


def post(Z, n):
    """warp_fft_post_pk: even slots X[k], odd slots conj(X[M - k]), 0.5 folded into the window table."""
    M = n // 2
    X = np.empty(M + 1, dtype=np.complex128)
    for k in range(1, (M + 1) // 2 + (M % 2 == 0)):
        if k == M - k:
            continue
        m, p = Z[k], np.conj(Z[M - k])
        S, D = m + p, m - p
        cs = np.exp(1j * 2 * np.pi * k / n)                   # (cos, sin)(2 pi k / n)
        T = cs.real * rotm(D) - cs.imag * D                   # -j W_n^k D
        X[k] = S + T
        X[M - k] = np.conj(S - T)                             # the kernel hands out S - T itself (odd slot)
    m, p = Z[0], np.conj(Z[0])
    S, D = m + p, m - p
    X[0] = S + rotm(D)
    X[M] = np.conj(S - rotm(D))
    if M % 2 == 0:
        X[M // 2] = 2.0 * np.conj(Z[M // 2])
    return X


@pytest.mark.parametrize("A,BQ", [(16, 16), (8, 8), (4, 8)])
def test_real_fft_from_packed_transform(A, BQ):
    n = 2 * A * BQ
    rng = np.random.default_rng(n)
    x = rng.standard_normal(n)
    zp = 0.5 * (x[0::2] + 1j * x[1::2])                      # the 0.5 lives in win2 (tables.cu)
    X = post(two_pass_fft(zp, A, BQ), n)
    assert np.allclose(X, np.fft.rfft(x), rtol=1e-11, atol=1e-11)


def test_three_pass_partner_chunks():
    """DESIGN.md 'what comes after' (1): with k = k1 + 32 k2 + 128 k3 (lane k1, chunk k2, radix-8 over k3) the real-FFT partner
    M - k of every bin of a chunk lies in ONE chunk of ONE lane: lane (32 - k1) % 32, chunk 3 - k2 (k1 != 0) or (4 - k2) % 4."""
    M = 1024
    for k1 in range(32):
        for k2 in range(4):
            ks = [k1 + 32 * k2 + 128 * k3 for k3 in range(8)]
            partners = {(M - k) % M for k in ks}
            p1 = (32 - k1) % 32
            p2 = (3 - k2) if k1 else (4 - k2) % 4
            assert partners == {p1 + 32 * p2 + 128 * k3 for k3 in range(8)}
