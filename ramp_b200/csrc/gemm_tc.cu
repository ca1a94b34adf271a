// tcgen05 / TMEM / TMA multi-tap split-precision GEMM for sm_100a — persistent, epilogue-overlapped.
//
// One kernel serves every dense contraction of the score network:
// linears, 1x1 / k5 / stride-2 / transposed 1-D convolutions and all of their input gradients.
//   D[128 rows x BN] (fp32, TMEM) = sum_{tap, kblock}  A_tap[128 x 64] * W_tap[BN x 64]^T
// * A rows are (sample, position) pairs of a channels-last activation; the tile for one tap is a 4-D
//   TMA box (64 channels, 1 parity, L_out positions, 128/L_out samples) whose position coordinate is
//   shifted by the tap, so conv padding is TMA out-of-bounds zero fill (no im2col, no halo code).
// * Operands are bf16 hi/lo planes (value = hi + lo).  precision==3 issues hi*hi + hi*lo + lo*hi
//   into the same fp32 TMEM accumulator (fp32-equivalent products, error ~2^-17); precision==1 issues
//   hi*hi only.
// * Persistent: one CTA per SM walks tiles (n fastest, so co-running CTAs share A tiles in L2).  The
//   accumulator is double buffered in TMEM (2 x BN columns), so the epilogue of tile i overlaps the
//   MMAs of tile i+1; the TMA producer runs ahead across tile boundaries through the smem ring.
// * Warp roles: warp 0 = TMA producer, warp 1 = TMEM owner + single-thread tcgen05.mma issuer,
//   warps 2..9 = two epilogue groups of four (TMEM lane quarter = warp % 4).  For BN >= 64 both groups drain
//   EVERY tile, each taking half of its columns ("cooperative": the tile's epilogue takes half as long and the
//   other accumulator buffer is being filled meanwhile, so neither the tensor pipe nor the epilogue warps wait on
//   each other as long as epilogue/2 >= MMA time or vice versa); for narrower tiles group g owns every other tile.
//   An epilogue thread owns one accumulator row; global loads / stores are transposed through a per-warp
//   XOR-swizzled smem staging tile so that every warp-wide access is row-contiguous.
// * Fused epilogues (GemmEpi): GEGLU forward / backward, LayerNorm forward / backward (Nout == BN == 256: the two
//   threads that own the halves of a row exchange partial statistics through shared memory and a 64-thread named
//   barrier; intermediate values are parked back in TMEM with tcgen05.st between the statistics pass and the
//   normalisation pass).
#include "common.cuh"
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>

namespace ramp {

static constexpr int BM = 128;
static constexpr int BK = 64;                 // bf16 elements = 128 B = one SWIZZLE_128B row
static constexpr int A_TILE_BYTES = BM * BK * 2;
static constexpr int EPI_WARPS = 8;            // two groups of four: group g drains accumulator buffer g
static constexpr int GEMM_THREADS = 64 + 32 * EPI_WARPS;
static constexpr int XCH_BYTES = 128 * 8;                 // static shared: LayerNorm partial-statistics exchange slots
static constexpr int STG_BYTES = EPI_WARPS * 32 * 32 * 4;   // per warp a 32 x 32 fp32 tile, XOR-swizzled 16-byte groups

struct GemmParams {
  int n_samples, L_out, nb;                   // nb = samples per 128-row tile
  int ntaps, kblocks, BN, nplanes, stages;
  int m_tiles, n_tiles;
  int tap_par[5], tap_joff[5], tap_wrow[5];   // tap_wrow = wtap * Nout
  const float* bias; const float* bias_tab; const int* row2scene; int Nout;
  const float* residual; int ldR;
  const float* residual2; int ldR2;
  float* out_f32; int ldO;
  bf16* out_hi; bf16* out_lo; int ldS;
  int o_L, o_stride, o_off;
  int epi, coop;                              // coop: both epilogue groups split the columns of every tile
  const float* ln_gamma; const float* ln_beta; float* ln_stats; const float* ln_x; const float* u;
  bf16* out2_hi; bf16* out2_lo; int ldS2;
};

// ---------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// K-major, SWIZZLE_128B operand tile: 8-row atoms of 1024 B (SBO), LBO unused, descriptor version 1.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
        "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]),
        "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]),
        "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void ld_acc(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  tmem_ld32(taddr, r);
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void st_acc(uint32_t taddr, const float (&v)[32]) {
  uint32_t r[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(v[i]);
  tmem_st32(taddr, r);
}

// ---------------------------------------------------------------------------------------------
// epilogue plumbing: a thread owns accumulator row (q*32 + lane) of the tile; for global traffic the
// warp re-maps to (row = (lane >> 3) + 4 i, 16-byte column group = lane & 7) through its staging tile.
// ---------------------------------------------------------------------------------------------
// staging tile: float4 group k of row r lives at float4 index r * 8 + (k ^ (r & 7))  (conflict-free both ways)
__device__ __forceinline__ int stg_idx(int r, int k) { return r * 8 + (k ^ (r & 7)); }
struct EpiCtx {
  float* stg;              // this warp's staging tile (32 x 32 fp32)
  int lane;
  bool my_ok; size_t my_row; int my_sample;
  unsigned ok8; uint32_t row8[8];
  float2* xch; int bar_id;                                   // LayerNorm statistics exchange slot shared with the other half's thread
};
// the two warps (same lane quarter, one per group) that share a tile's rows meet here
__device__ __forceinline__ void pair_sync(int bar_id) { asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory"); }

// v[32] (this thread's row, columns col0..col0+31) -> base[row * ld + col0 ..] for every live row of the warp
__device__ __forceinline__ void chunk_store_f32(const EpiCtx& e, const float (&v)[32], float* base, int ld, int col0) {
  float4* s4 = reinterpret_cast<float4*>(e.stg);
#pragma unroll
  for (int k = 0; k < 8; ++k) s4[stg_idx(e.lane, k)] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
  __syncwarp();
  const int rsub = e.lane >> 3, c4 = e.lane & 7;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float4 t = s4[stg_idx(rsub + 4 * i, c4)];
    if ((e.ok8 >> i) & 1u) *reinterpret_cast<float4*>(base + (size_t)e.row8[i] * ld + col0 + 4 * c4) = t;
  }
  __syncwarp();
}
// same, as bf16 hi / lo planes (ld, col0 in bf16 elements)
__device__ __forceinline__ void chunk_store_split(const EpiCtx& e, const float (&v)[32], bf16* hi, bf16* lo, int ld, int col0) {
  uint4* s4 = reinterpret_cast<uint4*>(e.stg);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    uint32_t h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) split2_bf16(v[8 * k + 2 * j], v[8 * k + 2 * j + 1], h[j], l[j]);
    s4[stg_idx(e.lane, k)] = make_uint4(h[0], h[1], h[2], h[3]);
    s4[stg_idx(e.lane, 4 + k)] = make_uint4(l[0], l[1], l[2], l[3]);
  }
  __syncwarp();
  const int rsub = e.lane >> 3, c4 = e.lane & 7;
  bf16* plane = (c4 < 4) ? hi : lo;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint4 t = s4[stg_idx(rsub + 4 * i, c4)];
    if ((e.ok8 >> i) & 1u) *reinterpret_cast<uint4*>(plane + (size_t)e.row8[i] * ld + col0 + 8 * (c4 & 3)) = t;
  }
  __syncwarp();
}
// base[row * ld + col0 ..] -> v[32] for this thread's row (zeros for dead rows)
__device__ __forceinline__ void chunk_load_f32(const EpiCtx& e, const float* __restrict__ base, int ld, int col0, float (&v)[32]) {
  const int rsub = e.lane >> 3, c4 = e.lane & 7;
  float4* s4 = reinterpret_cast<float4*>(e.stg);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    if ((e.ok8 >> i) & 1u) t = *reinterpret_cast<const float4*>(base + (size_t)e.row8[i] * ld + col0 + 4 * c4);
    s4[stg_idx(rsub + 4 * i, c4)] = t;
  }
  __syncwarp();
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float4 t = s4[stg_idx(e.lane, k)];
    v[4 * k] = t.x; v[4 * k + 1] = t.y; v[4 * k + 2] = t.z; v[4 * k + 3] = t.w;
  }
  __syncwarp();
}
// two-phase form: `chunk_ld_issue` puts the 8 row-contiguous 16-byte loads of the chunk in flight (registers),
// `chunk_ld_finish` transposes them through the staging tile into the thread's own row.  Issuing the next chunk's loads
// before the current chunk's arithmetic and stores keeps ~8 KB per warp in flight, which is what hides HBM latency here.
__device__ __forceinline__ void chunk_ld_issue(const EpiCtx& e, const float* __restrict__ base, int ld, int col0, float4 (&t)[8]) {
  const int c4 = e.lane & 7;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    t[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if ((e.ok8 >> i) & 1u) t[i] = __ldg(reinterpret_cast<const float4*>(base + (size_t)e.row8[i] * ld + col0 + 4 * c4));
  }
}
__device__ __forceinline__ void chunk_ld_finish(const EpiCtx& e, const float4 (&t)[8], float (&v)[32]) {
  const int rsub = e.lane >> 3, c4 = e.lane & 7;
  float4* s4 = reinterpret_cast<float4*>(e.stg);
#pragma unroll
  for (int i = 0; i < 8; ++i) s4[stg_idx(rsub + 4 * i, c4)] = t[i];
  __syncwarp();
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float4 x = s4[stg_idx(e.lane, k)];
    v[4 * k] = x.x; v[4 * k + 1] = x.y; v[4 * k + 2] = x.z; v[4 * k + 3] = x.w;
  }
  __syncwarp();
}
__device__ __forceinline__ void add_vec32(float (&v)[32], const float* __restrict__ b) {
#pragma unroll
  for (int i = 0; i < 32; i += 4) {
    const float4 t = *reinterpret_cast<const float4*>(b + i);
    v[i] += t.x; v[i + 1] += t.y; v[i + 2] += t.z; v[i + 3] += t.w;
  }
}
// acc + bias + per-scene bias + residual(s) for columns col .. col+31
__device__ __forceinline__ void epi_affine(const GemmParams& p, const EpiCtx& e, float (&v)[32], int col) {
  if (p.bias) add_vec32(v, p.bias + col);
  if (p.bias_tab && e.my_ok) add_vec32(v, p.bias_tab + (size_t)p.row2scene[e.my_sample] * p.Nout + col);
  if (p.residual) { float r[32]; chunk_load_f32(e, p.residual, p.ldR, col, r);
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] += r[i]; }
  if (p.residual2) { float r[32]; chunk_load_f32(e, p.residual2, p.ldR2, col, r);
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] += r[i]; }
}

// same with the residual chunk carried in flight: `t` holds the loads of columns `col` on entry and of `next_col` on exit
__device__ __forceinline__ void epi_affine_pf(const GemmParams& p, const EpiCtx& e, float (&v)[32], int col, float4 (&t)[8],
                                              int next_col, bool has_next) {
  if (p.bias) add_vec32(v, p.bias + col);
  if (p.bias_tab && e.my_ok) add_vec32(v, p.bias_tab + (size_t)p.row2scene[e.my_sample] * p.Nout + col);
  if (p.residual) {
    float r[32];
    chunk_ld_finish(e, t, r);
    if (has_next) chunk_ld_issue(e, p.residual, p.ldR, next_col, t);
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] += r[i];
  }
  if (p.residual2) { float r[32]; chunk_load_f32(e, p.residual2, p.ldR2, col, r);
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] += r[i]; }
}

// ---- one tile's epilogue, by mode.  acc = TMEM address of this warp's lane quarter, column 0 of the tile ----
__device__ __forceinline__ void epi_plain(const GemmParams& p, const EpiCtx& e, uint32_t acc, int colbase, int c0, int c1) {
  float4 t[8];
  if (p.residual) chunk_ld_issue(e, p.residual, p.ldR, colbase + c0, t);
#pragma unroll 1
  for (int c = c0; c < c1; c += 32) {
    float v[32];
    ld_acc(acc + c, v);
    epi_affine_pf(p, e, v, colbase + c, t, colbase + c + 32, c + 32 < c1);
    if (p.out_f32) chunk_store_f32(e, v, p.out_f32, p.ldO, colbase + c);
    if (p.out_hi) chunk_store_split(e, v, p.out_hi, p.out_lo, p.ldS, colbase + c);
  }
}
// BN == 16 tiles (Nout == 16): tiny outputs, direct per-row stores
__device__ __forceinline__ void epi_plain16(const GemmParams& p, const EpiCtx& e, uint32_t acc, int colbase) {
  uint32_t r[16];
  tmem_ld16(acc, r);
  tmem_ld_wait();
  if (!e.my_ok) return;
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
  if (p.bias)
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] += p.bias[colbase + i];
  if (p.bias_tab) {
    const float* bt = p.bias_tab + (size_t)p.row2scene[e.my_sample] * p.Nout + colbase;
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] += bt[i];
  }
  if (p.residual)
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] += p.residual[e.my_row * p.ldR + colbase + i];
  if (p.residual2)
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] += p.residual2[e.my_row * p.ldR2 + colbase + i];
  if (p.out_f32) {
    float* o = p.out_f32 + e.my_row * p.ldO + colbase;
#pragma unroll
    for (int i = 0; i < 16; i += 4) *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
  }
  if (p.out_hi) {
    __align__(16) bf16 h[16];
    __align__(16) bf16 l[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) split_bf16(v[i], h[i], l[i]);
    bf16* oh = p.out_hi + e.my_row * p.ldS + colbase;
    bf16* ol = p.out_lo + e.my_row * p.ldS + colbase;
    *reinterpret_cast<uint4*>(oh) = *reinterpret_cast<const uint4*>(h);
    *reinterpret_cast<uint4*>(oh + 8) = *reinterpret_cast<const uint4*>(h + 8);
    *reinterpret_cast<uint4*>(ol) = *reinterpret_cast<const uint4*>(l);
    *reinterpret_cast<uint4*>(ol + 8) = *reinterpret_cast<const uint4*>(l + 8);
  }
}
// GEGLU forward (layers_attention_mini.py:38-45): tile columns [0,128) = a, [128,256) = gate
__device__ __forceinline__ void epi_geglu_fwd(const GemmParams& p, const EpiCtx& e, uint32_t acc, int tile_n, int c0, int c1) {
  const int colbase = tile_n * 256;
  for (int c = c0 >> 1; c < (c1 >> 1); c += 32) {
    float a[32], g[32], s_[32];
    ld_acc(acc + c, a);
    ld_acc(acc + 128 + c, g);
    if (p.bias) { add_vec32(a, p.bias + colbase + c); add_vec32(g, p.bias + colbase + 128 + c); }
    // s = gelu(gate), q = a * gelu'(gate): everything the backward needs, so it is two multiplies per element there
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      float cdf, pdf;
      normal_cdf_pdf(g[i], cdf, pdf);
      const float sgl = g[i] * cdf;
      g[i] = a[i] * (cdf + g[i] * pdf);
      a[i] *= sgl;
      if (p.out_f32) s_[i] = sgl;
    }
    if (p.out_f32) {
      chunk_store_f32(e, s_, p.out_f32, p.ldO, colbase + c);
      chunk_store_f32(e, g, p.out_f32, p.ldO, colbase + 128 + c);
    }
    chunk_store_split(e, a, p.out2_hi, p.out2_lo, p.ldS2, tile_n * 128 + c);
  }
}
// GEGLU backward: acc = d(a * gelu(gate)) for natural columns colbase + c; u = (gelu(gate) | a gelu'(gate)) and du are
// tile-interleaved
__device__ __forceinline__ void epi_geglu_bwd(const GemmParams& p, const EpiCtx& e, uint32_t acc, int colbase, int c0, int c1) {
  auto u_col = [&](int c) { const int j = colbase + c; return (j >> 7) * 256 + (j & 127); };
  float4 t[8];                                   // the loads in flight: gelu(gate) of chunk c, then a gelu'(gate), then chunk c + 32 ...
  chunk_ld_issue(e, p.u, 2048, u_col(c0), t);
#pragma unroll 1
  for (int c = c0; c < c1; c += 32) {
    float d[32], x[32];
    const int pa = u_col(c), pg = pa + 128;
    ld_acc(acc + c, d);
    chunk_ld_finish(e, t, x);
    chunk_ld_issue(e, p.u, 2048, pg, t);
#pragma unroll
    for (int i = 0; i < 32; ++i) x[i] *= d[i];                      // u = (gelu(gate) | a * gelu'(gate))
    chunk_store_split(e, x, p.out_hi, p.out_lo, p.ldS, pa);
    chunk_ld_finish(e, t, x);
    if (c + 32 < c1) chunk_ld_issue(e, p.u, 2048, u_col(c + 32), t);
#pragma unroll
    for (int i = 0; i < 32; ++i) x[i] *= d[i];
    chunk_store_split(e, x, p.out_hi, p.out_lo, p.ldS, pg);
  }
}
// LayerNorm forward over the 256 channels of the row this thread owns (eps 1e-5, layers_attention_mini.py:138-140)
// This thread drains columns [c0, c1) of its row; with c1 - c0 == 128 the thread of the other half shares the statistics.
__device__ __forceinline__ void epi_ln_fwd(const GemmParams& p, const EpiCtx& e, uint32_t acc, int c0, int c1) {
  float shift = 0.f, s = 0.f, ss = 0.f;
  float4 t[8];
  if (p.residual) chunk_ld_issue(e, p.residual, p.ldR, c0, t);
#pragma unroll 1
  for (int c = c0; c < c1; c += 32) {
    float v[32];
    ld_acc(acc + c, v);
    epi_affine_pf(p, e, v, c, t, c + 32, c + 32 < c1);
    if (p.out_f32) chunk_store_f32(e, v, p.out_f32, p.ldO, c);
    if (p.out_hi) chunk_store_split(e, v, p.out_hi, p.out_lo, p.ldS, c);
    st_acc(acc + c, v);
    if (c == c0) shift = v[0];
#pragma unroll
    for (int i = 0; i < 32; ++i) { const float d = v[i] - shift; s += d; ss = fmaf(d, d, ss); }
  }
  tmem_st_wait();
  float mean, rstd;
  if (c1 - c0 == 256) {
    const float m0 = s * (1.f / 256.f);
    mean = shift + m0;
    rstd = rsqrtf(fmaxf(ss * (1.f / 256.f) - m0 * m0, 0.f) + 1e-5f);
  } else {
    // upper half sends (mean, M2) of its 128 values; lower half merges (Chan's update) and returns (mean, rstd)
    const float ml = s * (1.f / 128.f);
    const float2 mine = make_float2(shift + ml, fmaxf(ss - s * ml, 0.f));
    if (c0 != 0) *e.xch = mine;
    pair_sync(e.bar_id);
    if (c0 == 0) {
      const float2 peer = *e.xch;
      const float dm = peer.x - mine.x;
      mean = mine.x + 0.5f * dm;
      rstd = rsqrtf((mine.y + peer.y + dm * dm * 64.f) * (1.f / 256.f) + 1e-5f);
      *e.xch = make_float2(mean, rstd);
    }
    pair_sync(e.bar_id);
    if (c0 != 0) { const float2 t = *e.xch; mean = t.x; rstd = t.y; }
  }
  if (p.ln_stats && e.my_ok && c0 == 0) *reinterpret_cast<float2*>(p.ln_stats + e.my_row * 2) = make_float2(mean, rstd);
  for (int c = c0; c < c1; c += 32) {
    float v[32];
    ld_acc(acc + c, v);
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      const float4 g = *reinterpret_cast<const float4*>(p.ln_gamma + c + i);
      const float4 b = *reinterpret_cast<const float4*>(p.ln_beta + c + i);
      v[i] = (v[i] - mean) * rstd * g.x + b.x; v[i + 1] = (v[i + 1] - mean) * rstd * g.y + b.y;
      v[i + 2] = (v[i + 2] - mean) * rstd * g.z + b.z; v[i + 3] = (v[i + 3] - mean) * rstd * g.w + b.w;
    }
    chunk_store_split(e, v, p.out2_hi, p.out2_lo, p.ldS2, c);
  }
}
// out = residual + LayerNorm_backward(ln_x, acc): dx = rstd * (dyh - mean(dyh) - xh * mean(dyh * xh)), dyh = acc * gamma
__device__ __forceinline__ void epi_ln_bwd(const GemmParams& p, const EpiCtx& e, uint32_t acc, int c0, int c1) {
  float mean = 0.f, rstd = 0.f;
  if (e.my_ok) { const float2 st = *reinterpret_cast<const float2*>(p.ln_stats + e.my_row * 2); mean = st.x; rstd = st.y; }
  float s1 = 0.f, s2 = 0.f;
  float4 t[8];                                  // loads in flight (next ln_x chunk, or the residual chunk)
  chunk_ld_issue(e, p.ln_x, 256, c0, t);
#pragma unroll 1
  for (int c = c0; c < c1; c += 32) {
    float d[32], x[32];
    ld_acc(acc + c, d);
    chunk_ld_finish(e, t, x);
    chunk_ld_issue(e, p.ln_x, 256, c + 32 < c1 ? c + 32 : c0, t);      // the last one is the first chunk of pass 2
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      const float4 g = *reinterpret_cast<const float4*>(p.ln_gamma + c + i);
      d[i] *= g.x; d[i + 1] *= g.y; d[i + 2] *= g.z; d[i + 3] *= g.w;
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) { s1 += d[i]; s2 = fmaf(d[i], (x[i] - mean) * rstd, s2); }
    st_acc(acc + c, d);
  }
  tmem_st_wait();
  if (c1 - c0 != 256) {
    if (c0 != 0) *e.xch = make_float2(s1, s2);
    pair_sync(e.bar_id);
    if (c0 == 0) { const float2 peer = *e.xch; s1 += peer.x; s2 += peer.y; *e.xch = make_float2(s1, s2); }
    pair_sync(e.bar_id);
    if (c0 != 0) { const float2 t2 = *e.xch; s1 = t2.x; s2 = t2.y; }
  }
  s1 *= (1.f / 256.f); s2 *= (1.f / 256.f);
#pragma unroll 1
  for (int c = c0; c < c1; c += 32) {
    float d[32], x[32];
    const bool more = c + 32 < c1;
    ld_acc(acc + c, d);
    chunk_ld_finish(e, t, x);
    if (p.residual) chunk_ld_issue(e, p.residual, p.ldR, c, t);
    else if (more) chunk_ld_issue(e, p.ln_x, 256, c + 32, t);
#pragma unroll
    for (int i = 0; i < 32; ++i) d[i] = rstd * (d[i] - s1 - (x[i] - mean) * rstd * s2);
    if (p.residual) {
      chunk_ld_finish(e, t, x);
      if (more) chunk_ld_issue(e, p.ln_x, 256, c + 32, t);
#pragma unroll
      for (int i = 0; i < 32; ++i) d[i] += x[i];
    }
    if (p.out_f32) chunk_store_f32(e, d, p.out_f32, p.ldO, c);
    if (p.out_hi) chunk_store_split(e, d, p.out_hi, p.out_lo, p.ldS, c);
  }
}

// ---------------------------------------------------------------------------------------------
// kernel
// ---------------------------------------------------------------------------------------------
// One instantiation per epilogue mode: each gets its own register allocation (the modes' live sets differ a lot).
template <int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap mapAhi, const __grid_constant__ CUtensorMap mapAlo,
               const __grid_constant__ CUtensorMap mapBhi, const __grid_constant__ CUtensorMap mapBlo,
               const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b_tile_bytes = p.BN * BK * 2;
  const int stage_bytes = p.nplanes * (A_TILE_BYTES + b_tile_bytes);
  const uint32_t stg_base = smem_base + p.stages * stage_bytes;            // EPI_WARPS staging tiles
  const uint32_t bar_base = stg_base + STG_BYTES;
  const uint32_t full0 = bar_base, empty0 = bar_base + 8 * p.stages;
  const uint32_t tfull0 = bar_base + 16 * p.stages, tempty0 = tfull0 + 16;  // 2 + 2 accumulator barriers
  const uint32_t tmem_slot = tempty0 + 16;
  const int iters = p.ntaps * p.kblocks;
  const int total_tiles = p.m_tiles * p.n_tiles;
  uint32_t tmem_cols = 32;
  while ((int)tmem_cols < 2 * p.BN) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(tfull0 + 8 * b, 1); mbar_init(tempty0 + 8 * b, p.coop ? 8 : 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapAhi) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapBhi) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t tmem_acc;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_acc) : "r"(tmem_slot));

  if (warp == 0) {
    // ---------------- TMA producer ----------------
    if (lane == 0) {
      int it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int tile_m = tile / p.n_tiles, tile_n = tile - tile_m * p.n_tiles;
        const int c3 = tile_m * p.nb;
        for (int t = 0; t < p.ntaps; ++t) {
          for (int kb = 0; kb < p.kblocks; ++kb, ++it) {
            const int s = it % p.stages;
            const uint32_t ph = (it / p.stages) & 1;
            mbar_wait(empty0 + 8 * s, ph ^ 1);
            const uint32_t st = smem_base + s * stage_bytes;
            const uint32_t fb = full0 + 8 * s;
            mbar_expect_tx(fb, (uint32_t)(p.nplanes * (p.nb * p.L_out * BK * 2 + b_tile_bytes)));
            tma_load_4d(st, &mapAhi, fb, kb * BK, p.tap_par[t], p.tap_joff[t], c3);
            uint32_t off = A_TILE_BYTES;
            if (p.nplanes == 2) { tma_load_4d(st + off, &mapAlo, fb, kb * BK, p.tap_par[t], p.tap_joff[t], c3); off += A_TILE_BYTES; }
            tma_load_2d(st + off, &mapBhi, fb, kb * BK, p.tap_wrow[t] + tile_n * p.BN);
            off += b_tile_bytes;
            if (p.nplanes == 2) tma_load_2d(st + off, &mapBlo, fb, kb * BK, p.tap_wrow[t] + tile_n * p.BN);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer ----------------
    if (lane == 0) {
      // instruction descriptor: D=f32, A=B=bf16, both K-major, N = BN, M = 128
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      int it = 0, lt = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
        const int buf = lt & 1;
        mbar_wait(tempty0 + 8 * buf, ((lt >> 1) & 1) ^ 1);      // epilogue has drained this accumulator
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t dacc = tmem_acc + (uint32_t)(buf * p.BN);
        for (int i = 0; i < iters; ++i, ++it) {
          const int s = it % p.stages;
          const uint32_t ph = (it / p.stages) & 1;
          mbar_wait(full0 + 8 * s, ph);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t st = smem_base + s * stage_bytes;
          const uint32_t a_hi = st;
          const uint32_t a_lo = st + A_TILE_BYTES;
          const uint32_t b_hi = st + p.nplanes * A_TILE_BYTES;
          const uint32_t b_lo = b_hi + b_tile_bytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint32_t ko = k * 32;   // 16 bf16 = 32 B inside the 128 B swizzle row
            const uint32_t first = (i == 0 && k == 0) ? 0u : 1u;
            if (p.nplanes == 2) {
              umma_bf16(dacc, umma_desc_sw128(a_lo + ko), umma_desc_sw128(b_hi + ko), idesc, first);
              umma_bf16(dacc, umma_desc_sw128(a_hi + ko), umma_desc_sw128(b_lo + ko), idesc, 1u);
              umma_bf16(dacc, umma_desc_sw128(a_hi + ko), umma_desc_sw128(b_hi + ko), idesc, 1u);
            } else {
              umma_bf16(dacc, umma_desc_sw128(a_hi + ko), umma_desc_sw128(b_hi + ko), idesc, first);
            }
          }
          umma_commit(empty0 + 8 * s);           // frees the smem stage when these MMAs retire
        }
        umma_commit(tfull0 + 8 * buf);            // accumulator complete
      }
    }
  } else {
    // ---------------- epilogue: TMEM lane quarter = warp % 4, group = (warp - 2) / 4 ----------------
    const int q = warp & 3, group = (warp - 2) >> 2;
    __shared__ float2 xch[128];                             // XCH_BYTES static: one slot per tile row
    EpiCtx e;
    e.lane = lane;
    e.stg = reinterpret_cast<float*>(smem_gen + p.stages * stage_bytes) + (warp - 2) * 32 * 32;
    e.bar_id = 1 + q;
    e.xch = &xch[q * 32 + lane];
    // columns of a tile this group drains
    const int c0 = p.coop ? group * (p.BN >> 1) : 0, c1 = p.coop ? c0 + (p.BN >> 1) : p.BN;
    const int r = q * 32 + lane;
    const int n_local = r / p.L_out, l = r - n_local * p.L_out;
    auto my_row_of = [&](int tile_m, bool& ok) -> size_t {
      const int smp = tile_m * p.nb + n_local;
      ok = n_local < p.nb && smp < p.n_samples;
      return (size_t)smp * p.o_L + (size_t)l * p.o_stride + p.o_off;
    };
    const int step = p.coop ? (int)gridDim.x : 2 * (int)gridDim.x;
    int lt = p.coop ? 0 : group;
    int tile = blockIdx.x + lt * (int)gridDim.x;
    for (; tile < total_tiles; tile += step, lt += p.coop ? 1 : 2) {
      const int tile_m = tile / p.n_tiles, tile_n = tile - tile_m * p.n_tiles;
      const int buf = lt & 1;
      e.my_sample = tile_m * p.nb + n_local;
      e.my_row = my_row_of(tile_m, e.my_ok);
      e.ok8 = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int rr = q * 32 + (lane >> 3) + 4 * i;
        const int nl = rr / p.L_out, ll = rr - nl * p.L_out;
        const int smp = tile_m * p.nb + nl;
        if (nl < p.nb && smp < p.n_samples) e.ok8 |= 1u << i;
        e.row8[i] = (uint32_t)(smp * p.o_L + ll * p.o_stride + p.o_off);
      }
      mbar_wait(tfull0 + 8 * buf, (lt >> 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t acc = tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * p.BN);
      if (p.BN == 16) epi_plain16(p, e, acc, tile_n * 16);
      else if (EPI == EPI_PLAIN) epi_plain(p, e, acc, tile_n * p.BN, c0, c1);
      else if (EPI == EPI_GEGLU_FWD) epi_geglu_fwd(p, e, acc, tile_n, c0, c1);
      else if (EPI == EPI_GEGLU_BWD) epi_geglu_bwd(p, e, acc, tile_n * p.BN, c0, c1);
      else if (EPI == EPI_LN_FWD) epi_ln_fwd(p, e, acc, c0, c1);
      else epi_ln_bwd(p, e, acc, c0, c1);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty0 + 8 * buf);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_acc), "r"(tmem_cols) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// host side: tensor maps + launch
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static int g_max_smem = 0;
static int g_num_sms = 148;
// tuning knobs (environment, read once): RAMP_GEMM_COOP=0 disables the cooperative epilogue, RAMP_GEMM_L2PROMO=128|256
// picks the tensor-map L2 promotion
static int g_coop = 1, g_l2promo = 256;

int gemm_init() {
  if (g_encode) return 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  RAMP_CUDA_OK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  RAMP_CHECK(fn != nullptr && qres == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled not available");
  int dev = 0;
  RAMP_CUDA_OK(cudaGetDevice(&dev));
  RAMP_CUDA_OK(cudaDeviceGetAttribute(&g_max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  RAMP_CUDA_OK(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
  const int dyn = g_max_smem - XCH_BYTES;
  RAMP_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<EPI_PLAIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
  RAMP_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<EPI_GEGLU_FWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
  RAMP_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<EPI_GEGLU_BWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
  RAMP_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<EPI_LN_FWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
  RAMP_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<EPI_LN_BWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
  if (const char* v = getenv("RAMP_GEMM_COOP")) g_coop = atoi(v);
  if (const char* v = getenv("RAMP_GEMM_L2PROMO")) g_l2promo = atoi(v);
  g_encode = (EncodeTiledFn)fn;
  return 0;
}

struct MapKey {
  const void* ptr; uint64_t d[4]; uint64_t s[3]; uint32_t box[4]; int rank;
  bool operator==(const MapKey& o) const { return std::memcmp(this, &o, sizeof(MapKey)) == 0; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
    size_t h = 1469598103934665603ull;
    for (size_t i = 0; i < sizeof(MapKey) / 8; ++i) h = (h ^ w[i]) * 1099511628211ull;
    return h;
  }
};
static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;
static std::mutex g_maps_mu;

static int make_map(CUtensorMap* out, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box) {
  MapKey key;
  std::memset(&key, 0, sizeof(key));
  key.ptr = ptr; key.rank = rank;
  for (int i = 0; i < rank; ++i) { key.d[i] = dims[i]; key.box[i] = box[i]; }
  for (int i = 0; i < rank - 1; ++i) key.s[i] = strides_bytes[i];
  std::lock_guard<std::mutex> lk(g_maps_mu);
  auto it = g_maps.find(key);
  if (it != g_maps.end()) { *out = it->second; return 0; }
  cuuint64_t gd[4]; cuuint64_t gs[3]; cuuint32_t bx[4]; cuuint32_t es[4] = {1, 1, 1, 1};
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; }
  for (int i = 0; i < rank - 1; ++i) gs[i] = strides_bytes[i];
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), gd, gs, bx, es,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        g_l2promo == 128 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[256];
    snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled failed (%d): rank %d dims %llu %llu %llu %llu box %u %u %u %u",
             (int)r, rank, (unsigned long long)key.d[0], (unsigned long long)key.d[1], (unsigned long long)key.d[2],
             (unsigned long long)key.d[3], key.box[0], key.box[1], key.box[2], key.box[3]);
    set_error(buf);
    return -3;
  }
  if (g_maps.size() > 8192) g_maps.clear();
  g_maps[key] = *out;
  return 0;
}


static int pick_bn(int Nout, int m_tiles) {
  // Largest tile that divides Nout; drop to smaller tiles only while the grid would leave SMs idle.
  const int cands[5] = {256, 128, 64, 32, 16};
  int best = 0;
  for (int i = 0; i < 5; ++i) {
    int bn = cands[i];
    if (Nout % bn) continue;
    if (!best) best = bn;
    if ((long)m_tiles * (Nout / bn) >= g_num_sms || bn <= 64) return bn;
  }
  return best;
}

int gemm_launch(const GemmDesc& d, cudaStream_t stream) {
  if (gemm_init() != 0) return -1;
  RAMP_CHECK(d.L_out > 0 && d.L_out <= BM, "gemm: L_out must be in 1..128");
  RAMP_CHECK(d.ldA % 8 == 0 && d.Kw % 8 == 0, "gemm: leading dimensions must be multiples of 8 bf16");
  RAMP_CHECK(d.Nout % 16 == 0, "gemm: Nout must be a multiple of 16");
  RAMP_CHECK(d.ntaps >= 1 && d.ntaps <= 5, "gemm: 1..5 taps");
  RAMP_CHECK(d.precision == 1 || d.precision == 3, "gemm: precision must be 1 or 3");
  RAMP_CHECK(d.L_in % d.a_pstride == 0, "gemm: L_in must be a multiple of a_pstride");
  RAMP_CHECK(d.epi >= EPI_PLAIN && d.epi <= EPI_LN_BWD, "gemm: unknown epilogue");
  RAMP_CHECK((long)(d.n_samples + BM) * d.o_L + d.o_off < (1l << 31), "gemm: output row index must fit 31 bits");
  GemmParams p;
  std::memset(&p, 0, sizeof(p));
  p.n_samples = d.n_samples; p.L_out = d.L_out; p.nb = BM / d.L_out;
  p.ntaps = d.ntaps; p.kblocks = (d.K + BK - 1) / BK;
  p.m_tiles = (d.n_samples + p.nb - 1) / p.nb;
  p.BN = pick_bn(d.Nout, p.m_tiles);
  p.epi = d.epi;
  if (d.epi == EPI_GEGLU_FWD) {
    RAMP_CHECK(d.Nout % 256 == 0 && d.out2_hi && d.out2_lo, "gemm: GEGLU forward needs Nout % 256 == 0 and out2 planes");
    p.BN = 256;
  } else if (d.epi == EPI_GEGLU_BWD) {
    RAMP_CHECK(d.Nout % 128 == 0 && d.u && d.out_hi && d.out_lo, "gemm: GEGLU backward needs Nout % 128 == 0, u and out planes");
    if (p.BN < 32) p.BN = 128;
  } else if (d.epi == EPI_LN_FWD) {
    RAMP_CHECK(d.Nout == 256 && d.ln_gamma && d.ln_beta && d.out2_hi && d.out2_lo, "gemm: LN forward needs Nout == 256, gamma/beta, out2 planes");
    p.BN = 256;
  } else if (d.epi == EPI_LN_BWD) {
    RAMP_CHECK(d.Nout == 256 && d.ln_gamma && d.ln_stats && d.ln_x, "gemm: LN backward needs Nout == 256, gamma, stats, ln_x");
    p.BN = 256;
  }
  p.n_tiles = d.Nout / p.BN;
  p.coop = (p.BN >= 64 && g_coop) ? 1 : 0;
  p.nplanes = d.precision == 3 ? 2 : 1;
  const int stage_bytes = p.nplanes * (A_TILE_BYTES + p.BN * BK * 2);
  const int fixed = 1024 + STG_BYTES + 256;
  int stages = (g_max_smem - XCH_BYTES - fixed) / stage_bytes;
  if (stages > 6) stages = 6;
  const long total_iters = (long)p.ntaps * p.kblocks * ((long)p.m_tiles * p.n_tiles + g_num_sms - 1) / g_num_sms;
  if (stages > total_iters) stages = (int)total_iters;
  RAMP_CHECK(stages >= 1, "gemm: tile does not fit shared memory");
  p.stages = stages;
  for (int t = 0; t < d.ntaps; ++t) {
    p.tap_par[t] = d.taps[t].par; p.tap_joff[t] = d.taps[t].joff; p.tap_wrow[t] = d.taps[t].wtap * d.Nout;
  }
  p.bias = d.bias; p.bias_tab = d.bias_tab; p.row2scene = d.row2scene; p.Nout = d.Nout;
  p.residual = d.residual; p.ldR = d.ldR; p.residual2 = d.residual2; p.ldR2 = d.ldR2; p.out_f32 = d.out_f32; p.ldO = d.ldO;
  p.out_hi = d.out_hi; p.out_lo = d.out_lo; p.ldS = d.ldS;
  p.o_L = d.o_L; p.o_stride = d.o_stride; p.o_off = d.o_off;
  p.ln_gamma = d.ln_gamma; p.ln_beta = d.ln_beta; p.ln_stats = d.ln_stats; p.ln_x = d.ln_x; p.u = d.u;
  p.out2_hi = d.out2_hi; p.out2_lo = d.out2_lo; p.ldS2 = d.ldS2;
  RAMP_CHECK(p.BN == 16 || ((d.ldO % 4 == 0) && (d.ldS % 8 == 0) && (d.ldR % 4 == 0) && (d.ldR2 % 4 == 0) && (d.ldS2 % 8 == 0)),
             "gemm: output / residual leading dimensions must keep 16-byte alignment");

  CUtensorMap mAhi, mAlo, mBhi, mBlo;
  {
    uint64_t dims[4] = {(uint64_t)d.K, (uint64_t)d.a_pstride, (uint64_t)(d.L_in / d.a_pstride), (uint64_t)d.n_samples};
    uint64_t str[3] = {(uint64_t)d.ldA * 2, (uint64_t)d.a_pstride * d.ldA * 2, (uint64_t)d.L_in * d.ldA * 2};
    uint32_t box[4] = {(uint32_t)BK, 1u, (uint32_t)d.L_out, (uint32_t)p.nb};
    if (make_map(&mAhi, d.a_hi, 4, dims, str, box)) return -3;
    if (p.nplanes == 2) { if (make_map(&mAlo, d.a_lo, 4, dims, str, box)) return -3; } else mAlo = mAhi;
  }
  {
    uint64_t dims[2] = {(uint64_t)d.K, (uint64_t)d.n_wtaps * d.Nout};
    uint64_t str[1] = {(uint64_t)d.Kw * 2};
    uint32_t box[2] = {(uint32_t)BK, (uint32_t)p.BN};
    if (make_map(&mBhi, d.w_hi, 2, dims, str, box)) return -3;
    if (p.nplanes == 2) { if (make_map(&mBlo, d.w_lo, 2, dims, str, box)) return -3; } else mBlo = mBhi;
  }
  const size_t smem = (size_t)stages * stage_bytes + fixed;
  const long tiles = (long)p.m_tiles * p.n_tiles;
  dim3 grid((unsigned)(tiles < g_num_sms ? tiles : g_num_sms));
  switch (d.epi) {
    case EPI_PLAIN: gemm_tc_kernel<EPI_PLAIN><<<grid, GEMM_THREADS, smem, stream>>>(mAhi, mAlo, mBhi, mBlo, p); break;
    case EPI_GEGLU_FWD: gemm_tc_kernel<EPI_GEGLU_FWD><<<grid, GEMM_THREADS, smem, stream>>>(mAhi, mAlo, mBhi, mBlo, p); break;
    case EPI_GEGLU_BWD: gemm_tc_kernel<EPI_GEGLU_BWD><<<grid, GEMM_THREADS, smem, stream>>>(mAhi, mAlo, mBhi, mBlo, p); break;
    case EPI_LN_FWD: gemm_tc_kernel<EPI_LN_FWD><<<grid, GEMM_THREADS, smem, stream>>>(mAhi, mAlo, mBhi, mBlo, p); break;
    default: gemm_tc_kernel<EPI_LN_BWD><<<grid, GEMM_THREADS, smem, stream>>>(mAhi, mAlo, mBhi, mBlo, p); break;
  }
  RAMP_LAUNCH_OK();
  return 0;
}

}  // namespace ramp
