/* tac_spectral.h -- C ABI of the B200-native spectral hot path (libtac_spectral.so).
 *
 * Drop-in boundary for the one data-parallel path of ahmd-mohsin/Task-Aware-audio-coding-perceptual
 * that BASELINE.json:north_star names.  The reference has no FFI of its own on this path: it calls
 * torch / torchaudio / librosa library routines from Python.  Each entry point below therefore
 * names the reference *call site* (file:line under the reference checkout) whose arithmetic it
 * replaces; INTEGRATION.md shows the ctypes stub a maintainer adds at that call site.
 *
 * Conventions
 *   - every function returns 0 on success, a TAC_E* code otherwise; tac_last_error() returns a
 *     thread-local, human readable message for the last failure (never throws across the ABI);
 *   - all pointers are DEVICE pointers unless the parameter is documented "host"; the caller
 *     allocates every input, output and workspace; float/double buffers must be 16-byte aligned
 *     where stated; the library owns only immutable per-device tables (window*norm, twiddles),
 *     built by tac_prepare_fft() (or lazily on first use) and freed by tac_shutdown();
 *   - all work is enqueued on `stream` (a cudaStream_t); no entry point synchronises the host,
 *     allocates, or creates streams once the tables for its n_fft exist, so calls are CUDA-graph
 *     capturable after tac_prepare_fft();
 *   - tac_stft_multi, the MR-STFT loss and tac_istft take n_fft in {64, 128, 256, 512, 1024, 2048} (periodic Hann,
 *     win_length <= n_fft, even hop); tac_stft_ex takes ANY n_fft with n_fft + n_fft/2 <= 4096 (chirp-z for the
 *     non-powers of two: 2046, 510, 400 ...; ask tac_stft_ex_supported).
 */
#ifndef TAC_SPECTRAL_H_
#define TAC_SPECTRAL_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define TAC_API __attribute__((visibility("default")))
#else
#define TAC_API
#endif

#define TAC_OK 0
#define TAC_EINVAL 1      /* bad argument (shape, unsupported n_fft, null pointer)        */
#define TAC_EALIGN 2      /* pointer not aligned as documented                             */
#define TAC_ECUDA 3       /* CUDA runtime error (message carries cudaGetErrorString)      */
#define TAC_EWORKSPACE 4  /* workspace too small (see tac_*_workspace_bytes)               */

#define TAC_MAX_SCALES 8

/* pad_mode for center != 0 */
#define TAC_PAD_ZERO 0     /* librosa >= 0.10 ('constant')           */
#define TAC_PAD_REFLECT 1  /* torch.stft / torchaudio default         */

/* out_mode of tac_stft_multi */
#define TAC_OUT_COMPLEX 0   /* complex64 [B, C, K, F]   -- torchaudio Spectrogram(power=None) layout  */
#define TAC_OUT_PLANES 1    /* float32  [B, 2C, F, K]   -- msstftd.py:152-153 (cat(real, imag) + rearrange) */
#define TAC_OUT_MAGPHASE 2  /* float32  [B*C, 2, K, f_pad] -- notebook front end + pad_tensor + stack  */

typedef void* tac_stream_t; /* cudaStream_t */

TAC_API const char* tac_last_error(void);
TAC_API int tac_version(void);
/* Number of CUDA kernels this library has enqueued so far in this process (launches recorded into a
 * CUDA graph during stream capture count once, at capture time).  bench.py's `gpu_launches`. */
TAC_API uint64_t tac_launch_count(void);

/* Build (idempotent) the device tables for one (n_fft, win_length, normalized) on the current
 * device.  Synchronous (cudaMalloc + cudaMemcpy); call once outside any stream capture. */
TAC_API int tac_prepare_fft(int n_fft, int win_length, int normalized);
/* Free every cached table on every device. */
TAC_API int tac_shutdown(void);

/* Number of frames torch.stft produces: 1 + (T + (center ? 2*(n_fft/2) : 0) - n_fft) / hop, or 0. */
TAC_API int64_t tac_stft_num_frames(int64_t T, int n_fft, int hop, int center);

/* K1 -- batched multi-scale STFT, all scales in ONE launch.
 * Replaces: torchaudio.transforms.Spectrogram(...)(x) at msstftd.py:125-127,151, the
 * cat(real, imag) + rearrange of msstftd.py:152-153, the per-scale loop of msstftd.py:188-191, and
 * librosa.stft + np.abs + np.angle of spectrogram.ipynb cell 10 L26-52 (raw :4120-4134) together
 * with pad_tensor / torch.stack (train_pkl_file_task_aware.py:64-68, pkl_file_models.py:889-892).
 *   x          [B*C, T] float32, row-major (16-byte aligned)
 *   n_fft/hop/win  host arrays of n_scales entries
 *   center     0: frames start at f*hop (msstftd.py:127); 1: signal padded by n_fft/2 each side
 *   normalized 1: divide by sqrt(sum(window^2)) (torchaudio 'window' normalisation); 0: raw
 *   out        host array of n_scales device pointers, layout per out_mode above, F = frames of
 *              that scale; for TAC_OUT_MAGPHASE the time axis has f_pad >= F entries (f_pad = 0
 *              means f_pad = F) and columns [F, f_pad) are zero-filled (pad_tensor semantics).
 */
TAC_API int tac_stft_multi(const float* x, int64_t B, int64_t C, int64_t T, int n_scales, const int* n_fft,
                   const int* hop, const int* win, int center, int pad_mode, int normalized,
                   int out_mode, int64_t f_pad, void* const* out, tac_stream_t stream);

/* K1x -- single-scale STFT for ANY frame length (8 <= n_fft, n_fft + n_fft/2 <= 4096), output epilogue fused into the
 * write-out.  Frame lengths of the list above run on the warp FFT core, every other one on a Bluestein (chirp-z)
 * transform in shared memory.  Replaces:
 *   - librosa.stft(n_fft=2046, hop 512, win 2046) + np.abs + np.angle + librosa.amplitude_to_db of spectrogram.ipynb
 *     cell 9 (raw :2894-2913)                                  -> TAC_EPI_MAGPHASE (+ rowmax, tac_db_from_rowmax) / TAC_EPI_MAGDB
 *   - torch.stft(n_fft=510, hop 128, hann, center=True) + spec_fwd (abs(z)^0.5 * exp(1j angle z) * 0.15) + pad_spec of
 *     sgmse/sgmse/data_module.py:109-116,162-169,212-214 and sgmse/sgmse/util/other.py:76-90   -> TAC_EPI_COMPRESS, f_pad
 *   - torchaudio.transforms.Spectrogram() (n_fft=400, hop 200, power=2) and 10*log10(S + 1e-10) of
 *     testing_scripts.ipynb:246-249                            -> TAC_EPI_POWER / TAC_EPI_LOGPOWER
 *   - the north star's mel alternative (melscale_fbanks(K, 0, 8000, 80, 16000, None, 'htk')) -> TAC_EPI_MEL
 * Every output is time-innermost with f_pad >= F entries per row (f_pad = 0: f_pad = F); frames [F, f_pad) are
 * zero-filled, f_pad < F crops. */
#define TAC_EPI_COMPLEX 0   /* complex64 [rows, K, f_pad]      z                                                    */
#define TAC_EPI_COMPRESS 1  /* complex64 [rows, K, f_pad]      |z|^p0 * exp(j angle z) * p1                         */
#define TAC_EPI_MAGPHASE 2  /* float32   [rows, 2, K, f_pad]   |z|, angle z                                         */
#define TAC_EPI_MAGDB 3     /* float32   [rows, 2, K, f_pad]   |z|, max(20 log10(max(|z|, p0)) - p1, p2)            */
#define TAC_EPI_POWER 4     /* float32   [rows, K, f_pad]      |z|^p0                                               */
#define TAC_EPI_LOGPOWER 5  /* float32   [rows, K, f_pad]      p1 * log10(|z|^p0 + p2)                              */
#define TAC_EPI_DB 6        /* float32   [rows, K, f_pad]      max(20 log10(max(|z|, p0)) - p1, p2)   (p0 = amin,
                                                               p1 = 20 log10(max(amin, ref)), p2 = floor in dB)     */
#define TAC_EPI_MEL 7       /* float32   [rows, n_mels, f_pad] sum_k fb[k][m] |z_k|^p0 ; p1 != 0: p1 * log10(. + p2) */
typedef struct tac_epilogue {
    int kind;
    float p0, p1, p2;
    int n_mels;             /* TAC_EPI_MEL */
    const float* mel_fb;    /* device [K][n_mels] float32 -- the layout of torchaudio.functional.melscale_fbanks */
    const int* mel_range;   /* device [n_mels][2]: first and one-past-last frequency row of each filter that is non-zero */
    unsigned* rowmax;       /* optional device [rows]: running maximum of |z| per row as float bits (atomicMax); the caller
                               zero-fills it.  Kinds that form |z|: MAGPHASE, MAGDB, DB, POWER with p0 = 1. */
} tac_epilogue;
/*   x       [rows, T] float32;  window: 0 periodic Hann, 1 its square root (data_module.py:13-17), centred in n_fft when
 *           win < n_fft;  center / pad_mode / normalized as for tac_stft_multi. */
/* K1, one real plane per scale from the ONE-PASS multi-scale kernel (north star kernel 1: "framing, windowing, real-FFT,
 * magnitude and log ... for several window/hop sizes per pass"): the per-scale loop of msstftd.py:185-192 with the
 * magnitude taken in the kernel, and testing_scripts.ipynb:246-249 (Spectrogram() then 10*log10(x + 1e-10)) /
 * spectrogram.ipynb:2894-2913 (magnitude) for several resolutions at once.  Arguments as tac_stft_multi; epi->kind is
 *   TAC_EPI_POWER     out[s] float32 [B, C, F_s, K_s] = |z|^p0              (p0 = 1 magnitude, 2 power)
 *   TAC_EPI_LOGPOWER  out[s] float32 [B, C, F_s, K_s] = p1 * log10(|z|^p0 + p2)
 * Frequency is innermost (the planes layout with one channel instead of the re / im pair): half of tac_stft_multi's
 * TAC_OUT_PLANES bytes.  Mel, dB-against-the-row-maximum and phase stay with tac_stft_ex (one n_fft per call). */
TAC_API int tac_stft_multi_epi(const float* x, int64_t B, int64_t C, int64_t T, int n_scales, const int* n_fft,
                               const int* hop, const int* win, int center, int pad_mode, int normalized,
                               const tac_epilogue* epi, void* const* out, tac_stream_t stream);
TAC_API int tac_stft_ex_supported(int n_fft);
/* Builds (once per device) the tables tac_stft_ex needs for this configuration -- the only allocating, synchronising part of
 * that path -- so that the first tac_stft_ex call may sit inside a CUDA-graph capture or on a non-blocking stream.  window: 0
 * periodic Hann, 1 its square root.  A tac_stft_ex call that finds its tables missing DURING a capture returns TAC_EINVAL
 * instead of allocating.  The reference builds its windows in the constructors (msstftd.py:125-127, data_module.py:109-118);
 * call this from the same place. */
TAC_API int tac_prepare_stft_ex(int n_fft, int win, int normalized, int window);
TAC_API int tac_stft_ex(const float* x, int64_t rows, int64_t T, int n_fft, int hop, int win, int window, int center,
                        int pad_mode, int normalized, const tac_epilogue* epi, int64_t f_pad, void* out, tac_stream_t stream);
/* out[rows, K, F] = max(20 log10(max(mag, amin)) - 20 log10(max(rowmax, amin)), -top_db): the second phase of
 * librosa.amplitude_to_db(S, ref=np.max) when the row maxima came out of the STFT epilogue (tac_epilogue.rowmax). */
TAC_API int tac_db_from_rowmax(const float* mag, int64_t rows, int64_t row_stride, int K, int F, int64_t ld, const unsigned* rowmax,
                               float amin, float top_db, float* out, tac_stream_t stream);

/* K2 -- fused multi-resolution STFT loss (no spectrogram is written to memory).
 * Definition of record (the reference holds no such function): per scale s, on the
 * msstftd.py:125-127 STFT convention, X = |STFT_s(x)|, Y = |STFT_s(y)|,
 *   sums[2s]   = sum |X - Y|,   sums[2s+1] = sum |ln max(X, eps) - ln max(Y, eps)|
 * over all B*K_s*F_s elements (float64, deterministic two-level reduction).
 *   workspace  tac_mrstft_workspace_bytes(...) bytes, 16-byte aligned.
 */
TAC_API size_t tac_mrstft_workspace_bytes(int64_t B, int64_t T, int n_scales, const int* n_fft, const int* hop);
TAC_API int tac_mrstft_loss_fwd(const float* x, const float* y, int64_t B, int64_t T, int n_scales,
                        const int* n_fft, const int* hop, const int* win, float eps, double* sums,
                        void* workspace, size_t workspace_bytes, tac_stream_t stream);
/* Backward of the above w.r.t. y:  grad_y[B, T] = sum_s coef[2s] * d(sum|X-Y|)/dy
 *                                               + coef[2s+1] * d(sum|ln X - ln Y|)/dy.
 * coef is a DEVICE array [2*n_scales] float32 (the caller folds grad_output / (S * count_s) into
 * it, so no host sync is needed).  grad_y is fully overwritten. */
TAC_API int tac_mrstft_loss_bwd(const float* x, const float* y, int64_t B, int64_t T, int n_scales,
                        const int* n_fft, const int* hop, const int* win, float eps, const float* coef,
                        float* grad_y, void* workspace, size_t workspace_bytes, tac_stream_t stream);

/* Fused training-step form of K2: the per-scale sums AND grad_y in ONE pass (three transforms per frame
 * instead of five).  Possible because d loss / d sums is known from shapes alone (coef), so the backward
 * does not have to wait for the forward value.  Same workspace size as the forward. */
TAC_API int tac_mrstft_loss_fwd_bwd(const float* x, const float* y, int64_t B, int64_t T, int n_scales,
                            const int* n_fft, const int* hop, const int* win, float eps, const float* coef,
                            double* sums, float* grad_y, void* workspace, size_t workspace_bytes,
                            tac_stream_t stream);

/* K2' -- the reference's spectral reductions in one pass over (obs, dec), both [B, 2, K, F]:
 * pkl_file_models.py:924,933-940 / pkl_file_model_updated.py:520-540.
 *   out5       float32 [5]: magnitude_loss, phase_loss, total_loss, spec_snr (dB), mean(mse)
 *   per_sample float32 [B]: 0.5 * mean((obs - dec)^2) per item (may be NULL)
 *   workspace  tac_spec_losses_workspace_bytes(B) bytes
 */
TAC_API size_t tac_spec_losses_workspace_bytes(int64_t B);
TAC_API int tac_spec_losses(const float* obs, const float* dec, int64_t B, int64_t K, int64_t F, float* out5,
                    float* per_sample, void* workspace, size_t workspace_bytes, tac_stream_t stream);
/* K2' for callers that cannot allocate device memory.  No counterpart in the reference (which lets torch allocate
 * whatever it likes); same arithmetic, same kernels and bit-identical results to tac_spec_losses, minus the workspace
 * argument.  Meant for hosts that own pointers to obs/dec/out5 but cannot get scratch at the call site: a plugin living
 * inside another framework's memory pool that it must not bypass, a real-time or callback thread where cudaMalloc's
 * implicit synchronisation and driver locks are forbidden, a process with a device-memory budget fixed at start-up.
 *   - The partial sums live in a static __device__ array inside libtac_spectral: 1 MiB on EVERY GPU that loads the
 *     library, held until the process exits, sized for B <= TAC_SPEC_LOSSES_STATIC_MAX_B.  Larger B -> TAC_EINVAL
 *     (split the batch, or use tac_spec_losses with your own workspace).
 *   - The call itself runs no cudaMalloc / cudaFree / allocating library routine.  The FIRST call on each device
 *     (with that device current) resolves the array's device address with cudaGetSymbolAddress and caches it; make
 *     that first call outside any stream capture.
 *   - All calls on one device share the one array.  Calls that follow each other on a stream (or are otherwise ordered)
 *     are fine -- every slot a call reads it has written first, nothing needs clearing -- but two calls whose kernels
 *     can overlap on the GPU (different streams, no event between them) race on it and return garbage.
 *   out5, per_sample as for tac_spec_losses (per_sample may be NULL). */
#define TAC_SPEC_LOSSES_STATIC_MAX_B 4096
TAC_API int tac_spec_losses_static(const float* obs, const float* dec, int64_t B, int64_t K, int64_t F, float* out5,
                           float* per_sample, tac_stream_t stream);
/* Backward of K2' w.r.t. dec: grad_dec[b, c, :, :] = w[b, c] * (dec - obs), w a DEVICE float32 [B, 2]
 * table into which the caller folds the upstream gradients and normalisers of all five outputs
 * (autograd of pkl_file_models.py:933-940 in one pass instead of ~12 elementwise kernels). */
TAC_API int tac_spec_losses_bwd(const float* obs, const float* dec, int64_t B, int64_t K, int64_t F, const float* w,
                        float* grad_dec, tac_stream_t stream);

/* K3 -- DPCA.  Replaces data_pca (pkl_file_model_updated.py:37-42, pkl_file_models.py:51-56) and the
 * select / project block (pkl_file_model_updated.py:489-511 two views, :398-427 four views;
 * apply_pca pkl_file_models.py:224-228).
 *
 * tac_dpca_stats: sufficient statistics of nv views in one launch chain.
 *   z      host array of nv device pointers, each [B, d] float32 row-major (ld = d)
 *   stats  device float64 [nv][1 + d + d*d]: n, sum_b z, sum_b z z^T   (sum-reducible across
 *          ranks: all-reduce this buffer, then call tac_dpca_eig on every rank)
 *   algo   TAC_GRAM_TF32X3: tcgen05 tensor-core Gram of the locally centred batch, 3xTF32 split,
 *          fp32 accumulation in TMEM, float64 re-assembly;  TAC_GRAM_FP64: SIMT float64.
 */
#define TAC_GRAM_TF32X3 0
#define TAC_GRAM_FP64 1
TAC_API size_t tac_dpca_workspace_bytes(int nv, int64_t B, int d);
TAC_API int tac_dpca_stats(const float* const* z, int nv, int64_t B, int d, int algo, double* stats,
                   void* workspace, size_t workspace_bytes, tac_stream_t stream);
/* tac_dpca_eig: per view, C = sum zz^T - (sum z)(sum z)^T / n, symmetric eigen-decomposition
 * in float64, s = sqrt(max(eig, 0)) descending.  d <= 128: Householder tridiagonalisation (one CTA per view), then
 * one warp per eigenpair on d/8 CTAs per view (multisection on the Sturm count, twisted factorisation,
 * back-transformation), then ordering -- three launches on `stream`; 128 < d <= 256: one-sided Jacobi, one CTA per view.
 * workspace: tac_dpca_workspace_bytes(nv, B, d) bytes (T, the reflectors and the float64 eigenvectors live there).
 *   s [nv][d] float32 (entries >= rank are 0), v [nv][d][d] float32 with eigenvector j in COLUMN j
 *   (row-major [d, d], i.e. data_pca's v), mu [nv][d] float32, rank [nv] int32 = min(n, d).
 */
TAC_API int tac_dpca_eig(const double* stats, int nv, int d, float* s, float* v, float* mu, int* rank,
                 void* workspace, size_t workspace_bytes, tac_stream_t stream);
/* tac_dpca_select: global top-dim_p over the concatenated per-view singular values
 * (pkl_file_model_updated.py:496-500 / :404-412).  sel [nv][d] uint8 mask-free index lists:
 *   idx [nv][d] int32 (first cnt[v] entries valid, in global descending order), cnt [nv] int32. */
TAC_API int tac_dpca_select(const float* s, const int* rank, int nv, int d, int dim_p, int* idx, int* cnt,
                    tac_stream_t stream);
/* tac_dpca_project: z_out = ((z - mu) V_sel) V_sel^T + mu for every view
 * (pkl_file_model_updated.py:503-511).  z_out host array of nv device pointers [B, d]. */
TAC_API int tac_dpca_project(const float* const* z, int nv, int64_t B, int d, const float* v, const float* mu,
                     const int* idx, const int* cnt, float* const* z_out, void* workspace,
                     size_t workspace_bytes, tac_stream_t stream);
/* E1D1 bottleneck (pkl_file_models.py:901-907): unbiased column variances from stats and the
 * indices of the k largest; var [d] float32, top_idx [k] int32 (descending variance). */
TAC_API int tac_var_topk(const double* stats, int d, int k, float* var, int* top_idx, tac_stream_t stream);

/* dB channel of the notebook front end: librosa.amplitude_to_db(S, ref=np.max) per utterance (spectrogram.ipynb cell 9
 * L31-63, raw :2894-2913; amin 1e-5, top_db 80 are librosa's defaults).
 *   mag  row r = magnitude plane [K][ld] float32 at mag + r * row_stride (only the first F columns of a row count: the
 *        tac_stft_multi TAC_OUT_MAGPHASE output with row_stride = 2*K*f_pad, ld = f_pad is accepted as is)
 *   out  [rows][K][F] float32;  workspace rows * 4 bytes. */
TAC_API int tac_amplitude_to_db(const float* mag, int64_t rows, int64_t row_stride, int K, int F, int64_t ld, float amin,
                        float top_db, float* out, void* workspace, size_t workspace_bytes, tac_stream_t stream);

/* Latent regulariser of pkl_file_models.py:943-947 / pkl_file_model_updated.py:526-532 (z -= mean; z /= |z|_F;
 * nuc = |z|_* / B) from the singular values tac_dpca_eig already produced: no second SVD.
 *   out [nv] float32 = sum_i s_i / (sqrt(sum_i s_i^2) * n);  aux [nv][3] float64 = {sum s, |z|_F, n} (may be NULL). */
TAC_API int tac_nuclear_from_s(const float* s, const int* rank, const double* stats, int nv, int d, float* out, double* aux,
                       tac_stream_t stream);
/* Backward helper: gm [nv][d][d] float32 with  d nuc / d z = (z - mean) gm  (V diag(1/s) V^T form of the nuclear-norm
 * sub-gradient, null directions excluded); apply it with tac_dpca_apply(add_mean = 0). */
TAC_API int tac_nuclear_grad_matrix(const float* s, const float* v, const int* rank, const double* aux, int nv, int d, float* gm,
                            tac_stream_t stream);
/* z_out_v = (z_v - mu_v) P_v (+ mu_v if add_mean) for nv views, P [nv][d][d] float32 row-major: the GEMM of
 * tac_dpca_project with a caller-supplied matrix. */
TAC_API int tac_dpca_apply(const float* const* z, int nv, int64_t B, int d, const float* P, const float* mu, int add_mean,
                   float* const* z_out, tac_stream_t stream);

/* ---- next row: batched inverse STFT ------------------------------------------------------------
 * Replaces batch_reconstruct_waveform, train_pkl_file_task_aware.py:174-208 (magnitude * exp(1j * phase), then a Python
 * list comprehension of torch.istft(n_fft=2048, hop_length=512, win_length=2048, hann) per item, then torch.stack), and
 * reconstruct_waveform (:210-236).  One launch for the whole batch: polar -> complex, one-sided inverse real FFT,
 * synthesis window, overlap-add and division by the window envelope fused; semantics of torch.istft(center, normalized,
 * onesided=True, length).
 *   in_mode  TAC_IN_POLAR    a = magnitude, b = phase   (float32 planes)
 *            TAC_IN_PLANES   a = real,      b = imaginary
 *            TAC_IN_COMPLEX  a = interleaved complex64, b ignored (strides in complex elements)
 *   element (row r, bin k, frame f) lives at a[r * row_stride + k * ld + f]; K = n_fft/2 + 1 bins, F frames -- the two
 *   planes of the [B, 2, 1025, 600] decoder output are passed as a = base, b = base + K*ld, row_stride = 2*K*ld, ld = 600
 *   out      [rows][out_len] float32; tac_istft_length() samples are produced per row (hop*(F-1) when centred),
 *            cropped to out_len; a larger out_len continues into the padded tail of the overlap-added signal and is
 *            zero-filled behind it, exactly like torch.istft(length=)
 * hop must lie in [n_fft/8, n_fft]; the call fails with TAC_EINVAL (as torch.istft raises) when the overlap-added squared
 * window vanishes inside the kept range.  No workspace. */
#define TAC_IN_POLAR 0
#define TAC_IN_PLANES 1
#define TAC_IN_COMPLEX 2
TAC_API int64_t tac_istft_length(int64_t F, int n_fft, int hop, int center);
TAC_API int tac_istft(const float* a, const float* b, int in_mode, int64_t rows, int64_t row_stride, int64_t ld, int64_t F,
              int n_fft, int hop, int win, int center, int normalized, float* out, int64_t out_len, tac_stream_t stream);

/* ---- chain: decoder spectrogram -> waveform -> discriminator STFT inputs (VERDICT r1 "missing" 6) --------
 * Replaces the sequence train_pkl_file_task_aware.py:290 (batch_reconstruct_waveform(clean_mag, clean_phase)) -> :346-353
 * (unsqueeze(1), disc(ground_truth_clean_waveforms)) -> msstftd.py:188-191 -> :151-153 (spec_transform, cat(real, imag),
 * 'b c w t -> b c t w') for one batch: both kernels enqueued back to back on `stream`, no host synchronisation, every
 * output written once (the waveform is an output too: it feeds the MSE of :333).  rows_per_group = 0 runs the whole batch as
 * one group (fastest measured on B200); > 0 walks the batch in groups of that many rows, tac_istft then tac_stft_multi per
 * group -- with tac_chain_rows_per_group's value the group's waveform stays inside a 40 MiB budget of the 126 MB L2 between
 * the two kernels (at the price of two partial last waves per group).  Bit-identical to the two calls on the whole batch.
 *   a, b, in_mode, rows, row_stride, ld, F, n_fft, hop, win, center, normalized: as tac_istft
 *   wave    [rows][out_len] float32
 *   d_n_fft / d_hop / d_win (n_scales host entries), d_normalized: the discriminator scales (center = 0 framing)
 *   planes  host array of n_scales device pointers, each [rows, 2, F_s, K_s] float32 (one channel), F_s =
 *           tac_stft_num_frames(out_len, d_n_fft[s], d_hop[s], 0) */
TAC_API int64_t tac_chain_rows_per_group(int64_t rows, int64_t out_len, int64_t l2_budget_bytes);
TAC_API int tac_istft_disc_inputs(const float* a, const float* b, int in_mode, int64_t rows, int64_t row_stride, int64_t ld,
                          int64_t F, int n_fft, int hop, int win, int center, int normalized, float* wave, int64_t out_len,
                          int n_scales, const int* d_n_fft, const int* d_hop, const int* d_win, int d_normalized,
                          void* const* planes, int64_t rows_per_group, tac_stream_t stream);

/* Adjoint of the analysis STFT of tac_stft_multi (center = 0 framing): the backward of the Spectrogram module the
 * reference instantiates at msstftd.py:125-127 -- torchaudio's transform is differentiable and adversarial / feature
 * losses are back-propagated through the discriminator input to the generated waveform.
 *   out[r][t] = scale * sum_f w[t - f*hop] * sum_{k=0}^{n/2} Re((a + i b)[r][k][f] * e^{+2 pi i k (t - f*hop) / n})
 * for t in [0, out_len) (zero behind n_fft + hop*(F-1)); a, b = dL/dRe X, dL/dIm X in the layouts of tac_istft
 * (TAC_IN_PLANES or TAC_IN_COMPLEX); scale = 1 / sqrt(sum w^2) for normalized spectrograms, else 1.  Same kernel as
 * tac_istft without the one-sided weights and without the envelope division. */
TAC_API int tac_stft_adjoint(const float* a, const float* b, int in_mode, int64_t rows, int64_t row_stride, int64_t ld, int64_t F,
                     int n_fft, int hop, int win, float scale, float* out, int64_t out_len, tac_stream_t stream);

/* ---- next row (second half): pairwise cosine loss of the view latents --------------------------
 * Replaces  cos_sim = torch.nn.CosineSimilarity(); cos_loss = torch.mean(cos_sim(z1, z2) [+ ... all six pairs])
 * at pkl_file_models.py:390-391, 761-763 and pkl_file_model_updated.py:429-430, 513-514.
 *   z      host array of nv (2, 3 or 4) device pointers, each [B][d] float32 row-major
 *   sum    [1] float64 (device) = sum_b sum_{i<j} cos(z_i[b], z_j[b]) -- SUM-reducible across batch shards; the caller
 *          divides by the global batch.  eps: torch's clamp on each norm (1e-8).
 *   bwd    grad_z[i] [B][d] = scale * (*gscale if gscale else 1) * d(sum)/dz_i  (scale = 1 / global batch). */
TAC_API size_t tac_cosine_pairs_workspace_bytes(int64_t B);
TAC_API int tac_cosine_pairs_fwd(const float* const* z, int nv, int64_t B, int d, float eps, double* sum, void* workspace,
                         size_t workspace_bytes, tac_stream_t stream);
TAC_API int tac_cosine_pairs_bwd(const float* const* z, int nv, int64_t B, int d, float eps, const float* gscale, float scale,
                         float* const* grad_z, tac_stream_t stream);

/* ---- the bottleneck that is live in the SpectralRes models ----------------------------------------
 * Replaces the split / alpha-mix / truncate / cat block of pkl_file_models.py:1108-1127 (SpectralResE2D2), :692-718
 * (SpectralResE4D1) and :335-354 (SpectralResE2D1): view i's latent z[i] [B, d] is split at column `split` into a private
 * and a shared part; with alpha[p] != NULL (random_bottle_neck) the shared parts of the view pair (2p, 2p+1) are replaced
 * by alpha[p][b] * share_{2p} + (1 - alpha[p][b]) * share_{2p+1}; both parts are cut to `trunc` columns and concatenated:
 *   out [B, nv * w],  w = tac_bottleneck_width(d, split, trunc) = min(trunc, min(split, d)) + min(trunc, max(d - split, 0)),
 * view i in columns [i w, (i+1) w).  (E4D1 passes split = sum of all views' widths (:633, :692-695), which makes the shared
 * part empty -- reproduced as is.)  alpha: host array of (nv+1)/2 device pointers to [B] float32, or NULL.
 * tac_bottleneck_mix_bwd is the adjoint: grad_z[i] [B, d] is fully written (zeros outside the kept columns). */
TAC_API int tac_bottleneck_width(int d, int split, int trunc);
TAC_API int tac_bottleneck_mix(const float* const* z, int nv, int64_t B, int d, int split, int trunc, const float* const* alpha,
                               float* out, tac_stream_t stream);
TAC_API int tac_bottleneck_mix_bwd(const float* grad_out, int nv, int64_t B, int d, int split, int trunc, const float* const* alpha,
                                   float* const* grad_z, tac_stream_t stream);

/* ---- the loss tail of task_aware() -- train_pkl_file_task_aware.py:307-364 -----------------------
 * tac_interp_mse: F.mse_loss(a, F.interpolate(b, size=La, mode='linear', align_corners=False)) when Lb < La (and with the
 * roles swapped when La < Lb; plain MSE when equal) -- :311-333 -- without materialising the resampled waveform.
 *   a [B, La], b [B, Lb] float32;  out[0] float64;  workspace tac_interp_mse_workspace_bytes() bytes.
 * tac_multi_reduce: out[0] = sum_s mean_s, out[1 + s] = mean_s over nseg <= 48 tensor pairs of any sizes, with
 *   mode[s] = 0: mean |a_s - b_s|  (F.l1_loss of the feature-map pairs, :352-354)
 *             1: mean (a_s - b_s)^2
 *             2: mean (a_s - c[s])^2  (F.mse_loss(y, ones_like / zeros_like), :357-359; b_s ignored)
 * a, b: host arrays of device pointers; n, mode, c: host arrays.  Deterministic (fixed-order float64 partials). */
TAC_API size_t tac_interp_mse_workspace_bytes(void);
TAC_API int tac_interp_mse(const float* a, int64_t La, const float* b, int64_t Lb, int64_t B, double* out, void* workspace,
                           size_t workspace_bytes, tac_stream_t stream);
TAC_API size_t tac_multi_reduce_workspace_bytes(int nseg, const int64_t* n);
TAC_API int tac_multi_reduce(int nseg, const float* const* a, const float* const* b, const int64_t* n, const int* mode, const float* c,
                             double* out, void* workspace, size_t workspace_bytes, tac_stream_t stream);

/* ---- multi-GPU exchange: one-shot sum over NVLink peer memory ------------------------------------------
 * Replaces the packed ncclAllReduce(sum) of [loss sums | per view n, sum z, sum z z^T] that SURVEY.md 8e prescribes for
 * the batch-sharded form of train_pkl_file_task_aware.py:597-629 / pkl_file_model_updated.py:489-511 (the reference
 * itself computes both over the whole batch on one device).  Every rank owns ONE buffer of tac_xchg_bytes() in
 * peer-mappable memory (torch.distributed._symmetric_memory, or tac_ipc_alloc below) that every process of the node has
 * mapped; `peer[r]` is rank r's buffer as mapped in THIS process (peer[rank] = the local one).  A push peer-stores the
 * local payload into slot[rank] of every rank and raises a flag there; a pop spins on the local flags and adds the slots
 * in RANK ORDER (bit-identical float64 sums on all ranks).  Sequence counters live in the buffer, so push / pop pairs are
 * CUDA-graph replayable; every rank must execute the same sequence of pushes and pops.  The buffer must be zero-filled
 * (and the ranks synchronised) once before first use.
 * Waiting is bounded: a pop whose peers have not pushed within the timeout (tac_xchg_set_timeout, else the environment
 * variable TAC_XCHG_TIMEOUT_S, else 600 s; 0 = wait for ever) stops waiting, records the failed sequence number in its
 * buffer and finishes on undefined data -- it does NOT trap, so the CUDA context survives and the host can report the
 * failure: check tac_xchg_status after synchronising (any non-zero value is fatal for the exchange; tear it down).
 * Ranks that pause for longer than the timeout between steps (checkpoint, validation) should meet in a host barrier
 * before the next step. */
#define TAC_XCHG_MAX_WORLD 8
typedef struct tac_xchg {
    int world, rank;
    int64_t payload;                    /* float64 values per rank slot */
    void* peer[TAC_XCHG_MAX_WORLD];
} tac_xchg;
TAC_API size_t tac_xchg_bytes(int world, int64_t payload);
/* Process-wide wait bound of the pops enqueued from now on, in seconds; 0: unbounded; negative: back to the default
 * (TAC_XCHG_TIMEOUT_S or 600 s).  No reference counterpart (NCCL's watchdog plays this role for SURVEY 8e's all-reduce). */
TAC_API int tac_xchg_set_timeout(double seconds);
/* *timed_out = 0, or 1 + the sequence number of the latest pop of this descriptor's channel that gave up waiting.
 * Synchronous (one 4-byte device-to-host copy): call it outside the step, after a stream / device synchronise. */
TAC_API int tac_xchg_status(const tac_xchg* x, unsigned* timed_out);
/* src [payload] float64 (device) -> slot[rank] on all ranks, then the flags. */
TAC_API int tac_xchg_push(const tac_xchg* x, const double* src, tac_stream_t stream);
/* dst [payload] float64 = sum over ranks (rank order) of the slots of the matching push. */
TAC_API int tac_xchg_pop(const tac_xchg* x, double* dst, tac_stream_t stream);
/* tac_dpca_stats whose last kernel pushes the statistics (x->payload = nv * (1 + d + d*d)); x NULL / world 1: plain. */
TAC_API int tac_dpca_stats_push(const float* const* z, int nv, int64_t B, int d, int algo, double* stats, void* workspace,
                        size_t workspace_bytes, const tac_xchg* x, tac_stream_t stream);
/* tac_dpca_eig (d <= 128) whose first kernel pops: the one-CTA-per-view tridiagonalisation waits for the peers' flags
 * itself, adds the slots into `stats` (overwritten with the GLOBAL statistics) and goes on -- no kernel in between. */
TAC_API int tac_dpca_eig_pop(double* stats, int nv, int d, float* s, float* v, float* mu, int* rank, void* workspace,
                     size_t workspace_bytes, const tac_xchg* x, tac_stream_t stream);
/* Tail of the loss chain in one single-CTA kernel: push sums[n] (n <= 64), pop the global sums back into `sums`, and
 * loss[0] = sum_j sums[j] * inv[j] (float64; inv = 1 / (S * global count), the constant d loss / d sums).  x NULL / world
 * 1: the dot product alone.  Replaces the second all-reduce + torch.dot of the step. */
TAC_API int tac_loss_finalize(double* sums, int n, const double* inv, const tac_xchg* x, double* loss, tac_stream_t stream);
/* Peer-mappable device memory through CUDA IPC for callers without a symmetric-memory allocator: the ONE place the
 * library allocates on behalf of the caller (memory from a caching allocator cannot be exported).  handle64: 64 bytes
 * (cudaIpcMemHandle_t) to hand to the other processes of the node, which map it with tac_ipc_open. */
TAC_API int tac_ipc_alloc(size_t bytes, void** ptr, unsigned char* handle64);
TAC_API int tac_ipc_open(const unsigned char* handle64, void** ptr);
TAC_API int tac_ipc_close(void* ptr);
TAC_API int tac_ipc_free(void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* TAC_SPECTRAL_H_ */
