// Shared device/host helpers for the ramp_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>
#include <cstdint>
#include <cstdio>
#include <string>

namespace ramp {

// ---- error plumbing (C-ABI returns int status + ramp_last_error()) -------------------------
void set_error(const std::string& msg);
#define RAMP_CUDA_OK(expr)                                                                      \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      ramp::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));                      \
      return -1;                                                                                \
    }                                                                                           \
  } while (0)
#define RAMP_CHECK(cond, msg)                                                                   \
  do {                                                                                          \
    if (!(cond)) {                                                                              \
      ramp::set_error(std::string(msg));                                                        \
      return -2;                                                                                \
    }                                                                                           \
  } while (0)

// every kernel launch of this library is followed by RAMP_LAUNCH_OK(): counts it and checks the launch
void count_launch();
#define RAMP_LAUNCH_OK()                                                                        \
  do {                                                                                          \
    ramp::count_launch();                                                                       \
    RAMP_CUDA_OK(cudaGetLastError());                                                           \
  } while (0)

typedef __nv_bfloat16 bf16;

// Split-precision activation / weight view: value = hi + lo (two bf16 planes, same layout).
struct Split {
  bf16* hi;
  bf16* lo;
};

// ---- bf16 hi/lo split ------------------------------------------------------------------------
__device__ __forceinline__ void split_bf16(float v, bf16& hi, bf16& lo) {
  hi = __float2bfloat16_rn(v);
  lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Mish(x) = x * tanh(softplus(x)), and its derivative (PyTorch softplus threshold 20).
__device__ __forceinline__ float mish_f(float x) {
  float sp = (x > 20.f) ? x : log1pf(expf(x));
  return x * tanhf(sp);
}
__device__ __forceinline__ float mish_grad_f(float x) {
  float sp = (x > 20.f) ? x : log1pf(expf(x));
  float th = tanhf(sp);
  float sg = 1.f / (1.f + expf(-x));  // d softplus / dx
  return th + x * (1.f - th * th) * sg;
}
// exact GELU (erf form, torch F.gelu default) and derivative
__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_grad_f(float x) {
  float cdf = 0.5f * (1.f + erff(x * 0.70710678118654752f));
  float pdf = 0.39894228040143268f * expf(-0.5f * x * x);
  return cdf + x * pdf;
}

// ---- tcgen05 GEMM host interface (gemm_tc.cu) ----------------------------------------------------
// D[rows, Nout] = sum_taps A_tap[rows, K] * W_tap[Nout, K]^T  (+bias, +residual), rows = n_samples * L_out.
// A is a channels-last activation [n_samples, L_in, ldA] given as hi/lo bf16 planes; tap t reads the
// sequence position  par[t] + a_pstride * (l + joff[t])  (zero outside [0, L_in)) — this covers k5
// convolutions, stride-2 convolutions, transposed convolutions and their input gradients.
struct GemmTap {
  int par;    // parity / phase inside the a_pstride-periodic view of the sequence
  int joff;   // offset (in units of a_pstride positions) added to the output position
  int wtap;   // which [Nout, K] slice of the weight tensor
};
struct GemmDesc {
  // A operand
  const bf16* a_hi; const bf16* a_lo;
  int n_samples; int L_in; int ldA; int K;         // K channels are read from column 0..K-1
  int a_pstride;                                    // 1 (dense) or 2 (stride-2 view)
  int L_out;                                        // rows per sample produced by this GEMM (<= 128)
  int ntaps; GemmTap taps[5];
  // B operand (weights), [n_wtaps, Nout, Kw] bf16 hi/lo, Kw = weight row length (>= K, multiple of 8)
  const bf16* w_hi; const bf16* w_lo; int Nout; int Kw; int n_wtaps;
  // epilogue; output row of (sample n, position l) = n * o_L + l * o_stride + o_off
  const float* bias;            // [Nout] or null
  const float* bias_tab;        // [n_scenes, Nout] or null, added per row via row2scene[sample]
  const int* row2scene;
  const float* residual; int ldR;
  const float* residual2; int ldR2;   // optional second fp32 addend (same row mapping)
  float* out_f32; int ldO;
  bf16* out_hi; bf16* out_lo; int ldS;
  int o_L, o_stride, o_off;
  int precision;                // 1: hi*hi only; 3: hi*hi + hi*lo + lo*hi
};
int gemm_launch(const GemmDesc& d, cudaStream_t stream);
int gemm_init();   // resolves cuTensorMapEncodeTiled, sets smem attributes

}  // namespace ramp
