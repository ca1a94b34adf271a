// Self-attention core of BasicTransformerBlock.attn1 (layers_attention_mini.py:100-126): 4 heads x 64,
// softmax(q k^T / 8) v over L <= 64 tokens, and its input gradient (what autograd evaluates at
// UnetInference.py:23-27).  FP32 CUDA-core kernels, register tiled:
//   * a (sample, head) problem is owned by 2L threads; a 128-thread CTA packs 64/L problems;
//   * every inner product runs as a 4 x 8 (or 4 x L/8) register micro-tile fed by three 128-bit shared
//     loads per 32 FMAs; the operand that is indexed by the reduction index first is kept transposed in
//     shared memory so those loads are contiguous;
//   * softmax and the dS row reduction stay in registers (a score row lives in 8 adjacent lanes);
//   * the backward re-stages Q, K, dO in natural layout once the transposed copies are dead, and the
//     L x L matrices P, dS, dS^T take turns in ONE shared buffer, so two CTAs fit per SM.
#include "unet_ops.cuh"

namespace ramp {

static constexpr int HD = 64;        // head dim
static constexpr int NS = HD + 4;    // natural row stride  X[l][d]
__device__ __forceinline__ constexpr int tstride(int L) { return L + 4; }   // transposed row stride  Xt[d][l]

template <int N> struct VecLd;
template <> struct VecLd<8> { static __device__ __forceinline__ void ld(const float* p, float* v) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w; } };
template <> struct VecLd<4> { static __device__ __forceinline__ void ld(const float* p, float* v) {
  const float4 a = *reinterpret_cast<const float4*>(p); v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; } };
template <> struct VecLd<2> { static __device__ __forceinline__ void ld(const float* p, float* v) {
  const float2 a = *reinterpret_cast<const float2*>(p); v[0] = a.x; v[1] = a.y; } };
template <> struct VecLd<1> { static __device__ __forceinline__ void ld(const float* p, float* v) { v[0] = p[0]; } };

// global [L rows, ld floats] (64 used) -> natural smem X[l][d], row stride NS
template <int L>
__device__ __forceinline__ void stage_natural(const float* __restrict__ g, int ld, float* dst, int t) {
  for (int idx = t; idx < L * 16; idx += 2 * L) {
    const int l = idx >> 4, c4 = idx & 15;
    *reinterpret_cast<float4*>(dst + l * NS + c4 * 4) = *reinterpret_cast<const float4*>(g + (size_t)l * ld + c4 * 4);
  }
}
// global -> transposed smem Xt[d][l], row stride L + 4 (lanes walk l, so the scalar stores are conflict free)
template <int L>
__device__ __forceinline__ void stage_transposed(const float* __restrict__ g, int ld, float* dst, int t) {
  constexpr int TS = tstride(L);
  for (int idx = t; idx < L * 16; idx += 2 * L) {
    const int l = idx % L, c4 = idx / L;
    const float4 v = *reinterpret_cast<const float4*>(g + (size_t)l * ld + c4 * 4);
    float* d = dst + (c4 * 4) * TS + l;
    d[0] = v.x; d[TS] = v.y; d[2 * TS] = v.z; d[3 * TS] = v.w;
  }
}

// acc[4][CW] += sum_d At[d][i0..i0+3] * Bt[d][j0..j0+CW-1]      (both operands transposed, reduction over HD)
template <int L, int CW>
__device__ __forceinline__ void mm_tt(const float* At, const float* Bt, int i0, int j0, float (&acc)[4][CW]) {
  constexpr int TS = tstride(L);
#pragma unroll 8
  for (int d = 0; d < HD; ++d) {
    float a[4], b[CW];
    VecLd<4>::ld(At + d * TS + i0, a);
    VecLd<CW>::ld(Bt + d * TS + j0, b);
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < CW; ++c) acc[r][c] = fmaf(a[r], b[c], acc[r][c]);
  }
}
// acc[4][8] = sum_k Mt[k][i0..i0+3] * X[k][d0..d0+7]            (Mt: L x L with stride L+4, X natural)
template <int L>
__device__ __forceinline__ void mm_tn(const float* Mt, const float* X, int i0, int d0, float (&acc)[4][8]) {
  constexpr int TS = tstride(L);
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int c = 0; c < 8; ++c) acc[r][c] = 0.f;
#pragma unroll 4
  for (int k = 0; k < L; ++k) {
    float a[4], b[8];
    VecLd<4>::ld(Mt + k * TS + i0, a);
    VecLd<8>::ld(X + k * NS + d0, b);
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[r][c] = fmaf(a[r], b[c], acc[r][c]);
  }
}
__device__ __forceinline__ float grp8_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
  return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 4));
}
__device__ __forceinline__ float grp8_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  return v + __shfl_xor_sync(0xffffffffu, v, 4);
}
// rows i0..i0+3 of P = softmax(scale * S) for the thread's CW columns (a row spans the 8 lanes of its group)
template <int CW>
__device__ __forceinline__ void softmax_rows(float (&s)[4][CW]) {
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    float m = -INFINITY;
#pragma unroll
    for (int c = 0; c < CW; ++c) { s[r][c] *= 0.125f; m = fmaxf(m, s[r][c]); }
    m = grp8_max(m);
    float sum = 0.f;
#pragma unroll
    for (int c = 0; c < CW; ++c) { s[r][c] = expf(s[r][c] - m); sum += s[r][c]; }
    const float inv = 1.f / grp8_sum(sum);
#pragma unroll
    for (int c = 0; c < CW; ++c) s[r][c] *= inv;
  }
}
// 4 x 8 tile -> hi/lo planes at row (row0 + r), columns col0 .. col0 + 7
__device__ __forceinline__ void store_tile(bf16* hi, bf16* lo, size_t row0, int ld, int col0, const float (&acc)[4][8]) {
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    __align__(16) bf16 h[8];
    __align__(16) bf16 l[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) split_bf16(acc[r][c], h[c], l[c]);
    const size_t o = (row0 + r) * ld + col0;
    *reinterpret_cast<uint4*>(hi + o) = *reinterpret_cast<const uint4*>(h);
    *reinterpret_cast<uint4*>(lo + o) = *reinterpret_cast<const uint4*>(l);
  }
}

template <int L> struct AttnShape {
  static constexpr int CW = L / 8;             // score columns per thread
  static constexpr int TP = 2 * L;             // threads per problem
  static constexpr int NP = 128 / TP;          // problems per CTA
  static constexpr int TS = L + 4;
  static constexpr int FWD_FLOATS = 2 * HD * TS + L * NS + L * TS;             // Qt, Kt, V, Pt
  static constexpr int BWD_FLOATS = 4 * (HD * TS > L * NS ? HD * TS : L * NS) + L * TS;   // 4 operand slots + one LxL
  static constexpr int SLOT = (HD * TS > L * NS ? HD * TS : L * NS);
};

template <int L>
__global__ void __launch_bounds__(128) attn_fwd_tiled(const float* __restrict__ qkv, int n_problems, bf16* hi, bf16* lo) {
  using S = AttnShape<L>;
  extern __shared__ __align__(16) float sm[];
  const int p = threadIdx.x / S::TP, t = threadIdx.x % S::TP;
  const int prob = blockIdx.x * S::NP + p;
  const bool live = prob < n_problems;
  const int n = prob >> 2, h = prob & 3;
  float* Qt = sm + p * S::FWD_FLOATS; float* Kt = Qt + HD * S::TS; float* V = Kt + HD * S::TS; float* Pt = V + L * NS;
  const float* base = qkv + (size_t)n * L * 768 + h * HD;
  if (live) {
    stage_transposed<L>(base, 768, Qt, t);
    stage_transposed<L>(base + 256, 768, Kt, t);
    stage_natural<L>(base + 512, 768, V, t);
  }
  __syncthreads();
  const int ty = t >> 3, tx = t & 7, i0 = ty * 4;
  if (live) {
    float s[4][S::CW];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < S::CW; ++c) s[r][c] = 0.f;
    mm_tt<L, S::CW>(Qt, Kt, i0, tx * S::CW, s);
    softmax_rows<S::CW>(s);
#pragma unroll
    for (int c = 0; c < S::CW; ++c)
      *reinterpret_cast<float4*>(Pt + (tx * S::CW + c) * S::TS + i0) = make_float4(s[0][c], s[1][c], s[2][c], s[3][c]);
  }
  __syncthreads();
  if (live) {
    float o[4][8];
    mm_tn<L>(Pt, V, i0, tx * 8, o);
    store_tile(hi, lo, (size_t)n * L + i0, 256, h * HD + tx * 8, o);
  }
}

template <int L>
__global__ void __launch_bounds__(128) attn_bwd_tiled(const float* __restrict__ qkv, const float* __restrict__ d_out,
                                                      int n_problems, bf16* hi, bf16* lo) {
  using S = AttnShape<L>;
  extern __shared__ __align__(16) float sm[];
  const int p = threadIdx.x / S::TP, t = threadIdx.x % S::TP;
  const int prob = blockIdx.x * S::NP + p;
  const bool live = prob < n_problems;
  const int n = prob >> 2, h = prob & 3;
  float* A = sm + p * S::BWD_FLOATS; float* B = A + S::SLOT; float* C = B + S::SLOT; float* D = C + S::SLOT;
  float* M = D + S::SLOT;                              // the one L x L buffer (stride TS)
  const float* base = qkv + (size_t)n * L * 768 + h * HD;
  const float* dob = d_out + (size_t)n * L * 256 + h * HD;
  if (live) {
    stage_transposed<L>(base, 768, A, t);              // Qt
    stage_transposed<L>(base + 256, 768, B, t);        // Kt
    stage_transposed<L>(base + 512, 768, C, t);        // Vt
    stage_transposed<L>(dob, 256, D, t);               // dOt
  }
  __syncthreads();
  const int ty = t >> 3, tx = t & 7, i0 = ty * 4, j0 = tx * S::CW;
  float P[4][S::CW], dS[4][S::CW];
  if (live) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < S::CW; ++c) { P[r][c] = 0.f; dS[r][c] = 0.f; }
    mm_tt<L, S::CW>(A, B, i0, j0, P);                  // S = Q K^T
    softmax_rows<S::CW>(P);
    mm_tt<L, S::CW>(D, C, i0, j0, dS);                 // dP = dO V^T
#pragma unroll
    for (int r = 0; r < 4; ++r) {                      // dS = P * (dP - rowsum(P * dP)) * scale
      float acc = 0.f;
#pragma unroll
      for (int c = 0; c < S::CW; ++c) acc += P[r][c] * dS[r][c];
      acc = grp8_sum(acc);
#pragma unroll
      for (int c = 0; c < S::CW; ++c) dS[r][c] = P[r][c] * (dS[r][c] - acc) * 0.125f;
    }
  }
  __syncthreads();                                      // transposed operands are dead
  if (live) {
    stage_natural<L>(base, 768, A, t);                 // Q
    stage_natural<L>(base + 256, 768, B, t);           // K
    stage_natural<L>(dob, 256, D, t);                  // dO
    // M = P (row-major [i][j]): dV[j][d] = sum_i P[i][j] dO[i][d]
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < S::CW; ++c) M[(i0 + r) * S::TS + j0 + c] = P[r][c];
  }
  __syncthreads();
  const size_t row0 = (size_t)n * L + i0;              // i0 doubles as the key-row tile for dK / dV
  float acc[4][8];
  if (live) { mm_tn<L>(M, D, i0, tx * 8, acc); store_tile(hi, lo, row0, 768, 512 + h * HD + tx * 8, acc); }
  __syncthreads();
  if (live) {                                           // M = dS (row-major): dK[j][d] = sum_i dS[i][j] Q[i][d]
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < S::CW; ++c) M[(i0 + r) * S::TS + j0 + c] = dS[r][c];
  }
  __syncthreads();
  if (live) { mm_tn<L>(M, A, i0, tx * 8, acc); store_tile(hi, lo, row0, 768, 256 + h * HD + tx * 8, acc); }
  __syncthreads();
  if (live) {                                           // M = dS^T ([j][i]): dQ[i][d] = sum_j dS[i][j] K[j][d]
#pragma unroll
    for (int c = 0; c < S::CW; ++c)
      *reinterpret_cast<float4*>(M + (j0 + c) * S::TS + i0) = make_float4(dS[0][c], dS[1][c], dS[2][c], dS[3][c]);
  }
  __syncthreads();
  if (live) { mm_tn<L>(M, B, i0, tx * 8, acc); store_tile(hi, lo, row0, 768, h * HD + tx * 8, acc); }
}

template <int L>
static int launch_fwd(const float* qkv, int n_samples, bf16* hi, bf16* lo, cudaStream_t s) {
  using S = AttnShape<L>;
  static bool attr = false;
  const int smem = S::NP * S::FWD_FLOATS * (int)sizeof(float);
  if (!attr) { RAMP_CUDA_OK(cudaFuncSetAttribute(attn_fwd_tiled<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); attr = true; }
  const int np = n_samples * 4;
  attn_fwd_tiled<L><<<(np + S::NP - 1) / S::NP, 128, smem, s>>>(qkv, np, hi, lo);
  RAMP_LAUNCH_OK();
  return 0;
}
template <int L>
static int launch_bwd(const float* qkv, const float* d_out, int n_samples, bf16* hi, bf16* lo, cudaStream_t s) {
  using S = AttnShape<L>;
  static bool attr = false;
  const int smem = S::NP * S::BWD_FLOATS * (int)sizeof(float);
  if (!attr) { RAMP_CUDA_OK(cudaFuncSetAttribute(attn_bwd_tiled<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); attr = true; }
  const int np = n_samples * 4;
  attn_bwd_tiled<L><<<(np + S::NP - 1) / S::NP, 128, smem, s>>>(qkv, d_out, np, hi, lo);
  RAMP_LAUNCH_OK();
  return 0;
}

int attn_forward(const float* qkv, int n_samples, int L, bf16* out_hi, bf16* out_lo, cudaStream_t s) {
  if ((n_samples * 4) % (L >= 8 && L <= 64 ? 64 / L : 1) != 0) return attn_forward_generic(qkv, n_samples, L, out_hi, out_lo, s);
  switch (L) {
    case 64: return launch_fwd<64>(qkv, n_samples, out_hi, out_lo, s);
    case 32: return launch_fwd<32>(qkv, n_samples, out_hi, out_lo, s);
    case 16: return launch_fwd<16>(qkv, n_samples, out_hi, out_lo, s);
    case 8: return launch_fwd<8>(qkv, n_samples, out_hi, out_lo, s);
    default: return attn_forward_generic(qkv, n_samples, L, out_hi, out_lo, s);
  }
}
int attn_backward(const float* qkv, const float* d_out, int n_samples, int L, bf16* dq_hi, bf16* dq_lo, cudaStream_t s) {
  if ((n_samples * 4) % (L >= 8 && L <= 64 ? 64 / L : 1) != 0) return attn_backward_generic(qkv, d_out, n_samples, L, dq_hi, dq_lo, s);
  switch (L) {
    case 64: return launch_bwd<64>(qkv, d_out, n_samples, dq_hi, dq_lo, s);
    case 32: return launch_bwd<32>(qkv, d_out, n_samples, dq_hi, dq_lo, s);
    case 16: return launch_bwd<16>(qkv, d_out, n_samples, dq_hi, dq_lo, s);
    case 8: return launch_bwd<8>(qkv, d_out, n_samples, dq_hi, dq_lo, s);
    default: return attn_backward_generic(qkv, d_out, n_samples, L, dq_hi, dq_lo, s);
  }
}

}  // namespace ramp
