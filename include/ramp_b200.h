/* ramp_b200 — C ABI of the B200-native RAMP reverse-diffusion hot path.
 *
 * Plain pointers and sizes only.  Every pointer is a DEVICE pointer unless its name ends in
 * `_host`.  `stream` is a cudaStream_t passed as void*.  Every function returns 0 on success and
 * a negative status otherwise; ramp_last_error() then returns a description (thread-local).
 * Nothing here allocates after *_create; callers own all buffers they pass.
 *
 * The reference (wondmgezahu/RAMP) has no FFI: its seam is Python class lookup
 * (mpd/trainer/train_loaders.py:11-25).  Each entry point below cites the reference code whose
 * arithmetic it replaces; the Python classes in ramp_b200/ that mirror the reference API bind
 * these symbols with ctypes.
 */
#ifndef RAMP_B200_H
#define RAMP_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* ramp_last_error(void);
int ramp_version(void);

/* ---- low-level dense contraction (building block of the score network; exposed for tests) ----
 * D[(n,l), :] = sum_t A[n, par_t + a_pstride*(l + joff_t), 0:K] . W[wtap_t, :, 0:K]^T (+bias +bias_tab[row2scene[n]] +residual)
 * A / W are bf16 hi/lo planes (value = hi + lo); precision 3 = hi*hi + hi*lo + lo*hi, 1 = hi*hi.
 * taps_host: ntaps triples (par, joff, wtap).  Output row of (n,l) = n*o_L + l*o_stride + o_off.
 * Replaces torch F.linear / F.conv1d / F.conv_transpose1d calls of mpd/models/layers/layers.py:262-297
 * and mpd/models/layers/layers_attention_mini.py:41,58-62,92-99 and their autograd input-gradients. */
int ramp_gemm(const void* a_hi, const void* a_lo, int n_samples, int L_in, int ldA, int K, int a_pstride, int L_out,
              int ntaps, const int* taps_host, const void* w_hi, const void* w_lo, int Nout, int Kw, int n_wtaps,
              const float* bias, const float* bias_tab, const int* row2scene, const float* residual, int ldR,
              float* out_f32, int ldO, void* out_hi, void* out_lo, int ldS, int o_L, int o_stride, int o_off,
              int precision, void* stream);
/* fp32 -> bf16 hi/lo planes */
int ramp_split_bf16(const float* x, size_t n, void* hi, void* lo, void* stream);

#ifdef __cplusplus
}
#endif
#endif
