// Shared device/host helpers for the ramp_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
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

// ---- programmatic dependent launch (RAMP_PDL bit mask: 1 = GEMM, 2 = GroupNorm, 4 = attention) ----------------------
// A kernel launched with the attribute may start while the previous kernel of the stream is still draining; it must
// call pdl_wait() before touching anything that kernel wrote (without the attribute pdl_wait() is a no-op).
#ifndef RAMP_PDL_DEFAULT
#define RAMP_PDL_DEFAULT 0
#endif
inline int pdl_mode() {
  static const int m = [] { const char* e = getenv("RAMP_PDL"); return e ? atoi(e) : RAMP_PDL_DEFAULT; }();
  return m;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(int bit, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                              Args&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = (pdl_mode() & bit) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif

// Split-precision activation / weight view: value = hi + lo (two bf16 planes, same layout).
struct Split {
  bf16* hi;
  bf16* lo;
};

// ---- hi/lo split ------------------------------------------------------------------------------
// Every GEMM / attention operand is stored as two 16-bit planes, value = hi + lo.  RAMP_PLANES_FP16 picks the format:
//   1 (default): fp16 + fp16.  hi = fp16(v) (11 bits, clamped to the fp16 range), lo = fp16(v - hi): 22 bits for |v| >= 2^-3,
//      an absolute error <= 3e-8 (half an fp16 subnormal step) below that.  The backward sweep is loss-scaled by
//      RAMP_GRAD_SCALE (a power of two: exact) so that gradients sit in the well-resolved range too.  CPU study
//      (scripts/precision_study.py, profiles/r2_precision_study_cpu.txt): free-run deviation 4-6e-4 -> 4-7e-5 (= fp32 noise).
//   0: bf16 + bf16 (8 + 8 bits, round 1's format), built as a second library by `python -m ramp_b200.build --planes-bf16`.
// A mixed pair (fp16 hi, bf16 residual) is NOT an option: tcgen05.mma kind::f16 with A = f16, B = bf16 raises
// "illegal instruction" on B200 (measured, round 2) — both operands of a product must share one format.
#ifndef RAMP_PLANES_FP16
#define RAMP_PLANES_FP16 1
#endif
#if RAMP_PLANES_FP16
#define RAMP_GRAD_SCALE 256.0f
__device__ __forceinline__ void split_bf16(float v, bf16& hi, bf16& lo) {
  const __half h = __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));
  hi = __ushort_as_bfloat16(__half_as_ushort(h));                    // fp16 bits in a 16-bit plane element
  lo = __ushort_as_bfloat16(__half_as_ushort(__float2half_rn(v - __half2float(h))));
}
// two values at once: packed f16x2 planes (element 0 in the low half)
__device__ __forceinline__ void split2_bf16(float v0, float v1, uint32_t& hi, uint32_t& lo) {
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(v1), "f"(v0));      // saturates at +-65504 (one F2FP)
  float h0, h1;
  asm("{.reg .b16 a, b; mov.b32 {a, b}, %2; cvt.f32.f16 %0, a; cvt.f32.f16 %1, b;}" : "=f"(h0), "=f"(h1) : "r"(hi));
  const float r0 = v0 - h0, r1 = v1 - h1;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(r1), "f"(r0));
}
#else
#define RAMP_GRAD_SCALE 1.0f
__device__ __forceinline__ void split_bf16(float v, bf16& hi, bf16& lo) {
  hi = __float2bfloat16_rn(v);
  lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}
// two values at once: packed bf16x2 planes (element 0 in the low half), 6 instructions per pair
__device__ __forceinline__ void split2_bf16(float v0, float v1, uint32_t& hi, uint32_t& lo) {
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(v1), "f"(v0));
  const float r0 = v0 - __uint_as_float(hi << 16), r1 = v1 - __uint_as_float(hi & 0xffff0000u);
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(r1), "f"(r0));
}
#endif

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

// Phi(x) and phi(x) of the standard normal from ONE exponential: erf by Abramowitz-Stegun 7.1.26 (|err| < 1.5e-7,
// no cancellation on the negative side), phi(x) = exp(-x^2/2) / sqrt(2 pi) shares the exponential.  Used by the fused
// GEGLU epilogue, where the few epilogue warps are instruction-issue bound (erff + expf cost 2.5x as much).
__device__ __forceinline__ void normal_cdf_pdf(float x, float& cdf, float& pdf) {
  // z = |x| / sqrt 2 never materialises: its factor is folded into the rational's slope and into the exponent
  // (exp(-z^2) = 2^(x^2 * (-0.5 log2 e))), the 0.5 of Phi into the polynomial's coefficients
  float t, e;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f * 0.70710678118654752f, fabsf(x), 1.f)));   // one MUFU, 1 ulp
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * x * -0.72134752044448170f));   // relative error ~2^-22 < the polynomial's 1.5e-7
  const float poly = t * fmaf(t, fmaf(t, fmaf(t, fmaf(t, 0.5f * 1.061405429f, 0.5f * -1.453152027f), 0.5f * 1.421413741f),
                                      0.5f * -0.284496736f), 0.5f * 0.254829592f);
  const float half = poly * e;
  cdf = x < 0.f ? half : 1.f - half;
  pdf = 0.39894228040143268f * e;
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
  // fused epilogues (EPI_*): the thread that owns an accumulator row also owns the whole output row when
  // Nout == BN == 256, so LayerNorm and its backward are thread-local; GEGLU pairs (a, gate) live in one tile
  // because FF1's output channels are packed tile-interleaved (geglu_perm()).
  int epi;                      // GemmEpi
  const float* ln_gamma; const float* ln_beta;   // [256]
  float* ln_stats;              // [rows, 2] (mean, rstd): written by EPI_LN_FWD, read by EPI_LN_BWD
  const float* ln_x;            // EPI_LN_BWD: the LayerNorm input, fp32 [rows, 256]
  const float* u;               // EPI_GEGLU_BWD: saved FF1 output, fp32 [rows, 2048] (tile-interleaved)
  bf16* out2_hi; bf16* out2_lo; int ldS2;   // EPI_LN_FWD: LN(out) planes; EPI_GEGLU_FWD: a * gelu(gate) planes
};
enum GemmEpi {
  EPI_PLAIN = 0,
  EPI_GEGLU_FWD = 1,   // acc (+bias) = (a | gate) per 256-col tile -> out2 = a * gelu(gate) [rows, Nout/2] and, for the
                       // backward sweep, out_f32 = u = (gelu(gate) | a * gelu'(gate)) in the same tile-interleaved columns
  EPI_GEGLU_BWD = 2,   // acc = d = d(a*gelu(gate)) [rows, 1024] -> out_hi/lo = (d * u_s | d * u_q) [rows, 2048]
  EPI_LN_FWD = 3,      // v = acc + bias + bias_tab + residual -> out_f32 = v, out2 = LN(v), ln_stats
  EPI_LN_BWD = 4       // acc = dLN-output -> out = residual + LN_backward(ln_x, acc)
};
// tile-interleaved position of FF1 output channel o (0..2047; a = 0..1023, gate = 1024..2047)
__host__ __device__ inline int geglu_perm(int o) {
  const int part = o >> 10, c = o & 1023;
  return (c >> 7) * 256 + part * 128 + (c & 127);
}
int gemm_launch(const GemmDesc& d, cudaStream_t stream);
int gemm_init();   // resolves cuTensorMapEncodeTiled, sets smem attributes

}  // namespace ramp
