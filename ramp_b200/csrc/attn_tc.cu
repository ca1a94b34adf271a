// Tensor-core self-attention core (layers_attention_mini.py:100-126) for L = 64 / 32 / 16 / 8 tokens (8 runs zero-padded to 16 with masked keys), 4 heads x 64, and its
// input gradient (autograd at UnetInference.py:23-27), on mma.sync.m16n8k16 (f16 or bf16, the plane format of
// common.cuh) with split-precision operands: every product A.B is issued as A_lo.B_hi + A_hi.B_lo + A_hi.B_hi into one
// fp32 accumulator, the same scheme as the tcgen05 GEMM.  tcgen05 does not fit here: a (sample, head) problem is a 64x64x64 batch item and
// the softmax sits between the two products, so the FlashAttention-2 register pipeline is used instead —
//   S = Q K^T (ldmatrix operands) -> softmax on the accumulator fragments -> P re-packed in registers as the A
//   operand of O = P V (V through ldmatrix.trans).
// Backward: warp-per-16-query-rows computes P, dP = dO V^T, dS and dQ = dS K from registers; P and dS are parked in
// shared memory as bf16 hi/lo tiles, and after one barrier each warp owns 16 key rows for dV = P^T dO, dK = dS^T Q
// (transposed operands come for free from ldmatrix.trans).
// Inputs are the bf16 hi/lo planes the QKV / out-projection GEMMs write; outputs are hi/lo planes for the next GEMM.
#include "unet_ops.cuh"

#ifndef RAMP_ATTN_ALIAS
#define RAMP_ATTN_ALIAS 1      // 0: round 1's layout (P / dS in their own 4 PT bytes) for same-box A/B runs
#endif

namespace ramp {

__device__ __forceinline__ uint32_t s_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ldsm4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm4t(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
#if RAMP_PLANES_FP16
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
#else
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
#endif
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// c += (ah + al) . (bh + bl) without the lo.lo term
__device__ __forceinline__ void mma3(float (&c)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4], uint32_t bh0,
                                     uint32_t bh1, uint32_t bl0, uint32_t bl1) {
  mma16816(c, al, bh0, bh1);
  mma16816(c, ah, bl0, bl1);
  mma16816(c, ah, bh0, bh1);
}

// bf16 tile [rows][NCH * 8] in shared memory, 16-byte chunks XOR-swizzled so that the 8 row addresses of an
// ldmatrix 8x8 block fall into 8 different bank groups (NCH = 8: 128-byte rows; 4: 64-byte rows; 2: 32-byte rows).
template <int NCH>
__device__ __forceinline__ uint32_t tile_off(int r, int c) {
  return (uint32_t)(r * NCH * 16 + ((c ^ (NCH == 8 ? (r & 7) : (NCH == 4 ? ((r >> 1) & 3) : ((r >> 2) & 1)))) << 4));
}
// global bf16 [LV rows, ld] (64 columns used) -> tile of L >= LV rows (rows >= LV zero-filled), by the `nth` threads of
// one problem (cp.async, 16 B each)
template <int L, int LV>
__device__ __forceinline__ void load_tile(uint32_t tile, const bf16* __restrict__ g, int ld, int t, int nth) {
  for (int idx = t; idx < L * 8; idx += nth) {
    const int r = idx >> 3, c = idx & 7;
    if (r < LV)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(tile + tile_off<8>(r, c)), "l"(g + (size_t)r * ld + c * 8));
    else
      asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(tile + tile_off<8>(r, c)), "r"(0));
  }
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}
// A fragments (16 rows m0.., 16 k-columns starting at chunk kc) of a row-major tile
template <int NCH>
__device__ __forceinline__ void a_frag(uint32_t tile, int m0, int kc, int lane, uint32_t (&r)[4]) {
  ldsm4(tile + tile_off<NCH>(m0 + (lane & 7) + ((lane >> 3) & 1) * 8, kc + (lane >> 4)), r);
}
// A fragments of the TRANSPOSE of a stored tile X[k][m]: A[m0 + .][k0 + .] = X[k0 + .][m0 + .]
template <int NCH>
__device__ __forceinline__ void a_frag_t(uint32_t tile, int m0, int k0, int lane, uint32_t (&r)[4]) {
  const int i = lane >> 3;
  ldsm4t(tile + tile_off<NCH>(k0 + (lane & 7) + (i >> 1) * 8, (m0 >> 3) + (i & 1)), r);
}
// B fragments of two adjacent n-tiles (n0 .. n0+15) from storage Y[n][k], 16 k-columns at chunk kc:
// r[0], r[1] = (b0, b1) of tile n0; r[2], r[3] = (b0, b1) of tile n0 + 8
template <int NCH>
__device__ __forceinline__ void b_frag_nk(uint32_t tile, int n0, int kc, int lane, uint32_t (&r)[4]) {
  const int i = lane >> 3;
  ldsm4(tile + tile_off<NCH>(n0 + (lane & 7) + (i >> 1) * 8, kc + (i & 1)), r);
}
// same from storage Z[k][n] (k rows k0 .. k0+15, n columns at chunks nc, nc+1)
template <int NCH>
__device__ __forceinline__ void b_frag_kn(uint32_t tile, int k0, int nc, int lane, uint32_t (&r)[4]) {
  const int i = lane >> 3;
  ldsm4t(tile + tile_off<NCH>(k0 + (lane & 7) + (i & 1) * 8, nc + (i >> 1)), r);
}
__device__ __forceinline__ float quad_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
__device__ __forceinline__ float quad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

// S[16 x L] = Q[m0.., :] K^T for one warp (X = K) — also dP = dO V^T (A tile = dO, X = V)
template <int L>
__device__ __forceinline__ void qk_product(uint32_t Ah, uint32_t Al, uint32_t Xh, uint32_t Xl, int m0, int lane,
                                           float (&s)[L / 8][4]) {
#pragma unroll
  for (int nt = 0; nt < L / 8; ++nt) { s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f; }
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
    uint32_t ah[4], al[4];
    a_frag<8>(Ah, m0, 2 * kk, lane, ah);
    a_frag<8>(Al, m0, 2 * kk, lane, al);
#pragma unroll
    for (int n2 = 0; n2 < L / 16; ++n2) {
      uint32_t bh[4], bl[4];
      b_frag_nk<8>(Xh, 16 * n2, 2 * kk, lane, bh);
      b_frag_nk<8>(Xl, 16 * n2, 2 * kk, lane, bl);
      mma3(s[2 * n2], ah, al, bh[0], bh[1], bl[0], bl[1]);
      mma3(s[2 * n2 + 1], ah, al, bh[2], bh[3], bl[2], bl[3]);
    }
  }
}
// rows g and g + 8 of the warp's 16 x L block: p = softmax(scale * s)
template <int L, int LV>
__device__ __forceinline__ void softmax_frag(float (&s)[L / 8][4]) {
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    float m = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < L / 8; ++nt) {
      if (8 * nt >= LV) { s[nt][2 * h] = -INFINITY; s[nt][2 * h + 1] = -INFINITY; }     // padded keys
      s[nt][2 * h] *= 0.125f; s[nt][2 * h + 1] *= 0.125f;
      m = fmaxf(m, fmaxf(s[nt][2 * h], s[nt][2 * h + 1]));
    }
    m = quad_max(m);
    float sum = 0.f;
#pragma unroll
    for (int nt = 0; nt < L / 8; ++nt) {
      s[nt][2 * h] = expf(s[nt][2 * h] - m); s[nt][2 * h + 1] = expf(s[nt][2 * h + 1] - m);
      sum += s[nt][2 * h] + s[nt][2 * h + 1];
    }
    const float inv = 1.f / quad_sum(sum);
#pragma unroll
    for (int nt = 0; nt < L / 8; ++nt) { s[nt][2 * h] *= inv; s[nt][2 * h + 1] *= inv; }
  }
}
// accumulator fragments of two adjacent n-tiles -> hi / lo A fragments of the 16-wide k-step they cover
__device__ __forceinline__ void acc_to_a(const float (&c0)[4], const float (&c1)[4], uint32_t (&ah)[4], uint32_t (&al)[4]) {
  split2_bf16(c0[0], c0[1], ah[0], al[0]);        // register operands use the plane format too
  split2_bf16(c0[2], c0[3], ah[1], al[1]);
  split2_bf16(c1[0], c1[1], ah[2], al[2]);
  split2_bf16(c1[2], c1[3], ah[3], al[3]);
}
// out[16 x 64] = M[16 x L] (register fragments) . X[L x 64]  (X stored [k][n]: V for O = P V, K for dQ = dS K)
template <int L>
__device__ __forceinline__ void pv_product(const float (&p)[L / 8][4], uint32_t Xh, uint32_t Xl, int lane, float (&o)[8][4]) {
#pragma unroll
  for (int dn = 0; dn < 8; ++dn) { o[dn][0] = o[dn][1] = o[dn][2] = o[dn][3] = 0.f; }
#pragma unroll
  for (int kk = 0; kk < L / 16; ++kk) {
    uint32_t ah[4], al[4];
    acc_to_a(p[2 * kk], p[2 * kk + 1], ah, al);
#pragma unroll
    for (int d2 = 0; d2 < 4; ++d2) {
      uint32_t bh[4], bl[4];
      b_frag_kn<8>(Xh, 16 * kk, 2 * d2, lane, bh);
      b_frag_kn<8>(Xl, 16 * kk, 2 * d2, lane, bl);
      mma3(o[2 * d2], ah, al, bh[0], bh[1], bl[0], bl[1]);
      mma3(o[2 * d2 + 1], ah, al, bh[2], bh[3], bl[2], bl[3]);
    }
  }
}
// 16 x 64 accumulator block -> hi / lo planes, rows row0 + {g, g + 8}, columns col0 + 8 dn + 2 t
__device__ __forceinline__ void store_frag(const float (&o)[8][4], bf16* hi, bf16* lo, size_t row0, int ld, int col0, int lane,
                                           bool second_half = true) {
  const int g = lane >> 2, t = lane & 3;
#pragma unroll
  for (int dn = 0; dn < 8; ++dn) {
    uint32_t h0, l0, h1, l1;
    split2_bf16(o[dn][0], o[dn][1], h0, l0);
    split2_bf16(o[dn][2], o[dn][3], h1, l1);
    const size_t a = (row0 + g) * ld + col0 + 8 * dn + 2 * t, b = a + (size_t)8 * ld;
    *reinterpret_cast<uint32_t*>(hi + a) = h0; *reinterpret_cast<uint32_t*>(lo + a) = l0;
    if (second_half) { *reinterpret_cast<uint32_t*>(hi + b) = h1; *reinterpret_cast<uint32_t*>(lo + b) = l1; }
  }
}

template <int L> struct TcShape {
  static constexpr int W = L / 16;                 // warps per (sample, head) problem
  static constexpr int NP = 4 / W;                 // problems per 128-thread CTA
  static constexpr int TILE = L * 128;             // bytes of one [L][64] bf16 tile
  static constexpr int PT = L * L * 2;             // bytes of one [L][L] bf16 tile
  static constexpr int FWD_BYTES = 6 * TILE;       // Q, K, V hi/lo
  // Q, K, V, dO hi/lo.  P and dS (hi/lo, 4 PT <= 4 TILE) are parked over K and V once every warp of the problem is done
  // with them: 64 KB instead of 96 KB per L = 64 problem, i.e. three CTAs per SM instead of two (ncu r2: the backward at
  // L = 64 ran at sm__warps_active 17.7 %, 2.3x its HBM floor, with its load / compute / store phases serialised per CTA)
  static constexpr int BWD_BYTES = RAMP_ATTN_ALIAS ? 8 * TILE : 8 * TILE + 4 * PT;
};

template <int L, int LV>     // L = tile rows (multiple of 16), LV = tokens (LV < L: zero-padded queries / masked keys)
__global__ void __launch_bounds__(128) attn_fwd_tc(const bf16* __restrict__ qkv_hi, const bf16* __restrict__ qkv_lo,
                                                   int n_problems, bf16* __restrict__ out_hi, bf16* __restrict__ out_lo) {
  using S = TcShape<L>;
  extern __shared__ __align__(128) unsigned char sm_tc[];
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pl = warp / S::W, wp = warp % S::W;              // problem within the CTA, warp within the problem
  const int prob = blockIdx.x * S::NP + pl;
  const int n = prob >> 2, h = prob & 3;
  const uint32_t base = s_addr(sm_tc) + pl * S::FWD_BYTES;
  const uint32_t Qh = base, Ql = base + S::TILE, Kh = base + 2 * S::TILE, Kl = base + 3 * S::TILE, Vh = base + 4 * S::TILE,
                 Vl = base + 5 * S::TILE;
  if (prob < n_problems) {
    const size_t off = (size_t)n * LV * 768 + h * 64;
    const int t = wp * 32 + lane, nth = S::W * 32;
    load_tile<L, LV>(Qh, qkv_hi + off, 768, t, nth);        load_tile<L, LV>(Ql, qkv_lo + off, 768, t, nth);
    load_tile<L, LV>(Kh, qkv_hi + off + 256, 768, t, nth);  load_tile<L, LV>(Kl, qkv_lo + off + 256, 768, t, nth);
    load_tile<L, LV>(Vh, qkv_hi + off + 512, 768, t, nth);  load_tile<L, LV>(Vl, qkv_lo + off + 512, 768, t, nth);
  }
  cp_async_wait_all();
  __syncthreads();
  if (prob >= n_problems) return;
  const int m0 = 16 * wp;
  float s[L / 8][4];
  qk_product<L>(Qh, Ql, Kh, Kl, m0, lane, s);
  softmax_frag<L, LV>(s);
  float o[8][4];
  pv_product<L>(s, Vh, Vl, lane, o);
  store_frag(o, out_hi, out_lo, (size_t)n * LV + m0, 256, h * 64, lane, m0 + 8 < LV);
}

// (row g | g + 8, columns 8 nt + 2 t) of a 16 x L fragment block -> bf16 hi / lo tiles [L][L] at rows m0 + .
template <int L>
__device__ __forceinline__ void park_frag(const float (&p)[L / 8][4], uint32_t Th, uint32_t Tl, int m0, int lane) {
  const int g = lane >> 2, t = lane & 3;
#pragma unroll
  for (int nt = 0; nt < L / 8; ++nt) {
    uint32_t h0, l0, h1, l1;
    split2_bf16(p[nt][0], p[nt][1], h0, l0);
    split2_bf16(p[nt][2], p[nt][3], h1, l1);
    const uint32_t a = tile_off<L / 8>(m0 + g, nt) + 4 * t, b = tile_off<L / 8>(m0 + g + 8, nt) + 4 * t;
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(Th + a), "r"(h0));
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(Tl + a), "r"(l0));
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(Th + b), "r"(h1));
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(Tl + b), "r"(l1));
  }
}
// out[16 key rows j0.. x 64] = M^T . X, M stored [i][j] as hi/lo tiles (P or dS), X stored [i][d] (dO or Q)
template <int L>
__device__ __forceinline__ void tn_product(uint32_t Mh, uint32_t Ml, uint32_t Xh, uint32_t Xl, int j0, int lane, float (&o)[8][4]) {
#pragma unroll
  for (int dn = 0; dn < 8; ++dn) { o[dn][0] = o[dn][1] = o[dn][2] = o[dn][3] = 0.f; }
#pragma unroll
  for (int kk = 0; kk < L / 16; ++kk) {
    uint32_t ah[4], al[4];
    a_frag_t<L / 8>(Mh, j0, 16 * kk, lane, ah);
    a_frag_t<L / 8>(Ml, j0, 16 * kk, lane, al);
#pragma unroll
    for (int d2 = 0; d2 < 4; ++d2) {
      uint32_t bh[4], bl[4];
      b_frag_kn<8>(Xh, 16 * kk, 2 * d2, lane, bh);
      b_frag_kn<8>(Xl, 16 * kk, 2 * d2, lane, bl);
      mma3(o[2 * d2], ah, al, bh[0], bh[1], bl[0], bl[1]);
      mma3(o[2 * d2 + 1], ah, al, bh[2], bh[3], bl[2], bl[3]);
    }
  }
}

template <int L, int LV>
__global__ void __launch_bounds__(128) attn_bwd_tc(const bf16* __restrict__ qkv_hi, const bf16* __restrict__ qkv_lo,
                                                   const bf16* __restrict__ do_hi, const bf16* __restrict__ do_lo,
                                                   int n_problems, bf16* __restrict__ dq_hi, bf16* __restrict__ dq_lo) {
  using S = TcShape<L>;
  extern __shared__ __align__(128) unsigned char sm_tc[];
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pl = warp / S::W, wp = warp % S::W;
  const int prob = blockIdx.x * S::NP + pl;
  const bool live = prob < n_problems;
  const int n = prob >> 2, h = prob & 3;
  const uint32_t base = s_addr(sm_tc) + pl * S::BWD_BYTES;
  const uint32_t Qh = base, Ql = base + S::TILE, Kh = base + 2 * S::TILE, Kl = base + 3 * S::TILE, Vh = base + 4 * S::TILE,
                 Vl = base + 5 * S::TILE, Dh = base + 6 * S::TILE, Dl = base + 7 * S::TILE;
  const uint32_t Ph = RAMP_ATTN_ALIAS ? Kh : base + 8 * S::TILE, Pl = Ph + S::PT, Gh = Pl + S::PT, Gl = Gh + S::PT;     // G = dS; aliases K / V (see TcShape)
  if (live) {
    const size_t off = (size_t)n * LV * 768 + h * 64, offd = (size_t)n * LV * 256 + h * 64;
    const int t = wp * 32 + lane, nth = S::W * 32;
    load_tile<L, LV>(Qh, qkv_hi + off, 768, t, nth);        load_tile<L, LV>(Ql, qkv_lo + off, 768, t, nth);
    load_tile<L, LV>(Kh, qkv_hi + off + 256, 768, t, nth);  load_tile<L, LV>(Kl, qkv_lo + off + 256, 768, t, nth);
    load_tile<L, LV>(Vh, qkv_hi + off + 512, 768, t, nth);  load_tile<L, LV>(Vl, qkv_lo + off + 512, 768, t, nth);
    load_tile<L, LV>(Dh, do_hi + offd, 256, t, nth);        load_tile<L, LV>(Dl, do_lo + offd, 256, t, nth);
  }
  cp_async_wait_all();
  __syncthreads();
  const int m0 = 16 * wp;
  const size_t row0 = (size_t)n * LV + m0;
  const bool half2 = m0 + 8 < LV;
  float p[L / 8][4], dp[L / 8][4];
  if (live) {
    qk_product<L>(Qh, Ql, Kh, Kl, m0, lane, p);
    softmax_frag<L, LV>(p);
    qk_product<L>(Dh, Dl, Vh, Vl, m0, lane, dp);           // dP = dO V^T
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {                        // dS = P * (dP - rowsum(P * dP)) * scale
      float acc = 0.f;
#pragma unroll
      for (int nt = 0; nt < L / 8; ++nt) acc += p[nt][2 * hh] * dp[nt][2 * hh] + p[nt][2 * hh + 1] * dp[nt][2 * hh + 1];
      acc = quad_sum(acc);
#pragma unroll
      for (int nt = 0; nt < L / 8; ++nt) {
        dp[nt][2 * hh] = p[nt][2 * hh] * (dp[nt][2 * hh] - acc) * 0.125f;
        dp[nt][2 * hh + 1] = p[nt][2 * hh + 1] * (dp[nt][2 * hh + 1] - acc) * 0.125f;
      }
    }
    float o[8][4];
    pv_product<L>(dp, Kh, Kl, lane, o);                     // dQ = dS K
    store_frag(o, dq_hi, dq_lo, row0, 768, h * 64, lane, half2);
  }
  __syncthreads();                                          // K and V are dead from here on
  if (live) {
    park_frag<L>(p, Ph, Pl, m0, lane);
    park_frag<L>(dp, Gh, Gl, m0, lane);
  }
  __syncthreads();
  if (live) {
    float o[8][4];
    tn_product<L>(Ph, Pl, Dh, Dl, m0, lane, o);             // dV = P^T dO   (m0 doubles as the key-row block)
    store_frag(o, dq_hi, dq_lo, row0, 768, 512 + h * 64, lane, half2);
    tn_product<L>(Gh, Gl, Qh, Ql, m0, lane, o);             // dK = dS^T Q
    store_frag(o, dq_hi, dq_lo, row0, 768, 256 + h * 64, lane, half2);
  }
}

template <int L, int LV>
static int launch_fwd_tc(const bf16* qh, const bf16* ql, int n_samples, bf16* oh, bf16* ol, cudaStream_t s) {
  using S = TcShape<L>;
  static bool attr = false;
  const int smem = S::NP * S::FWD_BYTES;
  if (!attr) { RAMP_CUDA_OK(cudaFuncSetAttribute(attn_fwd_tc<L, LV>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); attr = true; }
  const int np = n_samples * 4;
  RAMP_CUDA_OK(launch_pdl(4, attn_fwd_tc<L, LV>, dim3((np + S::NP - 1) / S::NP), dim3(128), smem, s, qh, ql, np, oh, ol));
  RAMP_LAUNCH_OK();
  return 0;
}
template <int L, int LV>
static int launch_bwd_tc(const bf16* qh, const bf16* ql, const bf16* dh, const bf16* dl, int n_samples, bf16* oh, bf16* ol,
                         cudaStream_t s) {
  using S = TcShape<L>;
  static bool attr = false;
  const int smem = S::NP * S::BWD_BYTES;
  if (!attr) { RAMP_CUDA_OK(cudaFuncSetAttribute(attn_bwd_tc<L, LV>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); attr = true; }
  const int np = n_samples * 4;
  RAMP_CUDA_OK(launch_pdl(4, attn_bwd_tc<L, LV>, dim3((np + S::NP - 1) / S::NP), dim3(128), smem, s, qh, ql, dh, dl, np, oh, ol));
  RAMP_LAUNCH_OK();
  return 0;
}

bool attn_tc_supported(int L) { return L == 64 || L == 32 || L == 16 || L == 8; }

int attn_forward_tc(const bf16* qkv_hi, const bf16* qkv_lo, int n_samples, int L, bf16* out_hi, bf16* out_lo, cudaStream_t s) {
  if (L == 64) return launch_fwd_tc<64, 64>(qkv_hi, qkv_lo, n_samples, out_hi, out_lo, s);
  if (L == 32) return launch_fwd_tc<32, 32>(qkv_hi, qkv_lo, n_samples, out_hi, out_lo, s);
  if (L == 16) return launch_fwd_tc<16, 16>(qkv_hi, qkv_lo, n_samples, out_hi, out_lo, s);
  if (L == 8) return launch_fwd_tc<16, 8>(qkv_hi, qkv_lo, n_samples, out_hi, out_lo, s);
  RAMP_CHECK(false, "attn_forward_tc: L must be 64, 32, 16 or 8");
}
int attn_backward_tc(const bf16* qkv_hi, const bf16* qkv_lo, const bf16* do_hi, const bf16* do_lo, int n_samples, int L,
                     bf16* dq_hi, bf16* dq_lo, cudaStream_t s) {
  if (L == 64) return launch_bwd_tc<64, 64>(qkv_hi, qkv_lo, do_hi, do_lo, n_samples, dq_hi, dq_lo, s);
  if (L == 32) return launch_bwd_tc<32, 32>(qkv_hi, qkv_lo, do_hi, do_lo, n_samples, dq_hi, dq_lo, s);
  if (L == 16) return launch_bwd_tc<16, 16>(qkv_hi, qkv_lo, do_hi, do_lo, n_samples, dq_hi, dq_lo, s);
  if (L == 8) return launch_bwd_tc<16, 8>(qkv_hi, qkv_lo, do_hi, do_lo, n_samples, dq_hi, dq_lo, s);
  RAMP_CHECK(false, "attn_backward_tc: L must be 64, 32, 16 or 8");
}

}  // namespace ramp
