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

/* ---- score network: TemporalUnetInference energy gradient ------------------------------------
 * Replaces TemporalUnetInference.forward / forward_no_energy / EnergyGradFunction
 * (mpd/models/diffusion_models/UnetInference.py:19-37,157-224): eps = d/dx 0.5*sum f(x)^2, with the
 * backward sweep written out analytically (no autograd). */
typedef struct ramp_unet_config {
  int state_dim;   /* D: 4 (2-D models) or 6 (3-D) */
  int horizon;     /* H: multiple of 8, 8..64 */
  int ctx_dim;     /* 320 (ObstacleEncoderSet) or 256 (ObstacleEncoder) */
  int precision;   /* 3 = split-bf16 x3 (fp32-equivalent, the parity mode); 1 = plain bf16 */
  int max_rows;    /* most network rows per call */
  int max_scenes;  /* most distinct conditioning latents */
} ramp_unet_config;
int ramp_unet_create(const ramp_unet_config* cfg, void** handle);
int ramp_unet_destroy(void* handle);
/* name = reference state-dict key without the leading "model." (e.g. "downs.0.0.blocks.0.block.0.weight");
 * data = fp32, host or device, copied. */
int ramp_unet_set_param(void* handle, const char* name, const float* data, size_t numel);
int ramp_unet_finalize(void* handle, void* stream);      /* packs weights, allocates the workspace */
size_t ramp_unet_workspace_bytes(void* handle);
/* latents [n_scenes, ctx_dim]: cached scene-encoder outputs (UnetInference.py:146-156); an all-zero row is
 * the unconditional branch (UnetInference.py:190-197). */
int ramp_unet_set_scenes(void* handle, const float* latents, int n_scenes, void* stream);
/* x [n_rows, H, D]; t = the (batch-uniform) diffusion timestep; row2scene [n_rows] int32;
 * eps_out / f_out [n_rows, H, D], either may be NULL (f_out = forward_no_energy output). */
int ramp_unet_energy_grad(void* handle, const float* x, int n_rows, float t, const int32_t* row2scene, float* eps_out,
                          float* f_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif
