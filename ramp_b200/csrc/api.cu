// C-ABI entry points (include/ramp_b200.h).
#include "../../include/ramp_b200.h"
#include "common.cuh"
#include "unet_ops.cuh"
#include "unet_engine.cuh"

namespace ramp {
static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
static unsigned long long g_launches = 0;
void count_launch() { ++g_launches; }

__global__ void split_kernel(const float* __restrict__ x, size_t n, bf16* hi, bf16* lo) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    bf16 h, l;
    split_bf16(x[i], h, l);
    hi[i] = h;
    lo[i] = l;
  }
}
}  // namespace ramp

using namespace ramp;

extern "C" {

const char* ramp_last_error(void) { return g_err.c_str(); }
int ramp_version(void) { return 100; }
int ramp_plane_format(void) { return RAMP_PLANES_FP16 ? 2 : 0; }
unsigned long long ramp_launch_count(void) { return g_launches; }
void ramp_launch_count_add(unsigned long long n) { g_launches += n; }

int ramp_split_bf16(const float* x, size_t n, void* hi, void* lo, void* stream) {
  size_t b = (n + 255) / 256;
  if (b > 148 * 16) b = 148 * 16;
  if (n == 0) return 0;
  split_kernel<<<(int)b, 256, 0, (cudaStream_t)stream>>>(x, n, (bf16*)hi, (bf16*)lo);
  RAMP_LAUNCH_OK();
  return 0;
}

int ramp_gemm(const void* a_hi, const void* a_lo, int n_samples, int L_in, int ldA, int K, int a_pstride, int L_out,
              int ntaps, const int* taps_host, const void* w_hi, const void* w_lo, int Nout, int Kw, int n_wtaps,
              const float* bias, const float* bias_tab, const int* row2scene, const float* residual, int ldR,
              float* out_f32, int ldO, void* out_hi, void* out_lo, int ldS, int o_L, int o_stride, int o_off,
              int precision, void* stream) {
  RAMP_CHECK(ntaps >= 1 && ntaps <= 5, "ramp_gemm: ntaps must be 1..5");
  GemmDesc d{};
  d.a_hi = (const bf16*)a_hi; d.a_lo = (const bf16*)a_lo;
  d.n_samples = n_samples; d.L_in = L_in; d.ldA = ldA; d.K = K; d.a_pstride = a_pstride; d.L_out = L_out;
  d.ntaps = ntaps;
  for (int t = 0; t < ntaps; ++t) d.taps[t] = GemmTap{taps_host[3 * t], taps_host[3 * t + 1], taps_host[3 * t + 2]};
  d.w_hi = (const bf16*)w_hi; d.w_lo = (const bf16*)w_lo; d.Nout = Nout; d.Kw = Kw; d.n_wtaps = n_wtaps;
  d.bias = bias; d.bias_tab = bias_tab; d.row2scene = row2scene; d.residual = residual; d.ldR = ldR;
  d.out_f32 = out_f32; d.ldO = ldO; d.out_hi = (bf16*)out_hi; d.out_lo = (bf16*)out_lo; d.ldS = ldS;
  d.o_L = o_L; d.o_stride = o_stride; d.o_off = o_off; d.precision = precision;
  return gemm_launch(d, (cudaStream_t)stream);
}

int ramp_unet_create(const ramp_unet_config* cfg, void** handle) {
  RAMP_CHECK(cfg && handle, "ramp_unet_create: null argument");
  UnetConfig c{cfg->state_dim, cfg->horizon, cfg->ctx_dim, cfg->precision, cfg->max_rows, cfg->max_scenes};
  *handle = new UnetEngine(c);
  return 0;
}
int ramp_unet_destroy(void* handle) {
  delete (UnetEngine*)handle;
  return 0;
}
int ramp_unet_set_param(void* handle, const char* name, const float* data, size_t numel) {
  RAMP_CHECK(handle && name && data, "ramp_unet_set_param: null argument");
  return ((UnetEngine*)handle)->set_param(name, data, numel);
}
int ramp_unet_finalize(void* handle, void* stream) {
  RAMP_CHECK(handle, "ramp_unet_finalize: null handle");
  return ((UnetEngine*)handle)->finalize((cudaStream_t)stream);
}
size_t ramp_unet_workspace_bytes(void* handle) { return handle ? ((UnetEngine*)handle)->workspace_bytes() : 0; }
int ramp_unet_set_scenes(void* handle, const float* latents, int n_scenes, void* stream) {
  RAMP_CHECK(handle && latents, "ramp_unet_set_scenes: null argument");
  return ((UnetEngine*)handle)->set_scenes(latents, n_scenes, (cudaStream_t)stream);
}
int ramp_unet_energy_grad(void* handle, const float* x, int n_rows, int x_repeat, float t, const int32_t* row2scene,
                          float* eps_out, float* f_out, void* stream) {
  RAMP_CHECK(handle && x, "ramp_unet_energy_grad: null argument");
  return ((UnetEngine*)handle)->energy_grad(x, n_rows, x_repeat, t, row2scene, eps_out, f_out, (cudaStream_t)stream);
}

int ramp_unet_profile(void* handle, int enable, float* ms4_host, int* launches4_host) {
  RAMP_CHECK(handle, "ramp_unet_profile: null handle");
  return ((UnetEngine*)handle)->profile(enable, ms4_host, launches4_host);
}

int ramp_unet_profile_dump(void* handle, const char* path) {
  RAMP_CHECK(handle && path, "ramp_unet_profile_dump: null argument");
  return ((UnetEngine*)handle)->profile_dump(path);
}

}  // extern "C"
