// tcgen05 / TMEM / TMA multi-tap split-precision GEMM for sm_100a — persistent, epilogue-overlapped.
//
// One kernel serves every dense contraction of the score network:
// linears, 1x1 / k5 / stride-2 / transposed 1-D convolutions and all of their input gradients.
//   D[128 rows x BN] (fp32, TMEM) = sum_{tap, kblock}  A_tap[128 x 64] * W_tap[BN x 64]^T
// * A rows are (sample, position) pairs of a channels-last activation; the tile for one tap is a 4-D
//   TMA box (64 channels, 1 parity, L_out positions, 128/L_out samples) whose position coordinate is
//   shifted by the tap, so conv padding is TMA out-of-bounds zero fill (no im2col, no halo code).
// * Operands are 16-bit hi/lo planes (value = hi + lo; fp16 + fp16 by default, bf16 + bf16 with
//   -DRAMP_PLANES_FP16=0, see common.cuh).  precision==3 issues hi*hi + hi*lo + lo*hi into the same fp32 TMEM
//   accumulator (fp32-equivalent products); precision==1 issues hi*hi only.
// * Persistent: one CTA per SM walks tiles (n fastest, so co-running CTAs share A tiles in L2).  The
//   accumulator is double buffered in TMEM (2 x BN columns), so the epilogue of tile i overlaps the
//   MMAs of tile i+1; the TMA producer runs ahead across tile boundaries through the smem ring.
// * Warp roles: warp 0 = TMA producer, warp 1 = TMEM owner + single-thread tcgen05.mma issuer,
//   warps 2..17 = four epilogue groups of four (TMEM lane quarter = warp % 4).  All groups drain EVERY tile, each a
//   quarter of its columns in chunks of 16 ("cooperative": the other accumulator buffer is being filled meanwhile, so
//   neither the tensor pipe nor the epilogue warps wait on each other).  The epilogues are latency-bound chains
//   (tcgen05.ld -> math -> smem transpose -> global), so four warps per SM sub-partition at <= 112 registers beat two
//   fatter ones.  An epilogue thread owns one accumulator row; global loads / stores are transposed through a
//   per-warp XOR-swizzled smem staging tile so that every warp-wide access covers whole row segments, and global
//   loads are issued one chunk ahead of their use.
// * Fused epilogues (GemmEpi), one kernel instantiation each: GEGLU forward / backward, LayerNorm forward /
//   backward (Nout == BN == 256: the four threads that own the quarters of a row exchange partial statistics
//   through the staging tiles and a 128-thread named barrier; intermediate values are parked back in TMEM with
//   tcgen05.st between the statistics pass and the normalisation pass).
#include "common.cuh"
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>

// RAMP_GEMM_DIAG (diagnostic builds only, results are WRONG): bit 0 drops the epilogues' global loads, bit 1 their global
// stores - timing the launches without one side shows which one binds an HBM-bound epilogue (scripts/gpu_r2_i.sh).
#ifndef RAMP_GEMM_DIAG
#define RAMP_GEMM_DIAG 0
#endif
#ifndef RAMP_GEMM_RELAXED_ARRIVE
#define RAMP_GEMM_RELAXED_ARRIVE 1      // 0: the .release form (MEMBAR.ALL.GPU + ERRBAR per tile), for same-box A/B runs
#endif

namespace ramp {

static constexpr int BM = 128;
static constexpr int BK = 64;                 // bf16 elements = 128 B = one SWIZZLE_128B row
static constexpr int A_TILE_BYTES = BM * BK * 2;
// Two epilogue configurations (template parameter CW = accumulator columns a thread handles per step):
//   CW = 32:  8 epilogue warps (2 groups), 128-byte row segments per global access - the memory-heavy epilogues
//   CW = 16: 16 epilogue warps (4 groups),  64-byte row segments, <= 96 registers  - the math-heavy GEGLU forward
// Both use 32 KB of staging tiles (one 32 x CW fp32 tile per warp).
__host__ __device__ constexpr int epi_warps(int cw) { return cw == 32 ? 8 : 16; }
// RAMP_GEMM_WG = 1: warp-group aligned roles.  Warps 0..3 = {TMA producer, MMA issuer, idle, idle}, epilogue warps start at
// warp 4, so every 4-warp group has ONE role and `setmaxnreg` can move registers from the producer / issuer group (40 each)
// to the epilogue groups (232 each at 8 epilogue warps instead of the 168 that 64 K / 384 threads allows): room for two
// chunk loads in flight per thread without spills.  0: round 1's layout (epilogue from warp 2, no reallocation).
#ifndef RAMP_GEMM_WG
#define RAMP_GEMM_WG 1
#endif
static constexpr int EPI_W0 = RAMP_GEMM_WG ? 4 : 2;          // first epilogue warp
__host__ __device__ constexpr int gemm_threads(int cw) { return 32 * (RAMP_GEMM_WG ? 4 : 2) + 32 * epi_warps(cw); }
__host__ __device__ constexpr int epi_regs(int cw) { return cw == 32 ? 232 : 104; }   // (65536 - 128 * 40) / epilogue threads, in 8s
static constexpr int STG_BYTES = 32 * 1024;

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
  int epi, ngroups;                           // ngroups = epilogue groups that drain a tile: min(BN / CW, epilogue warps / 4)
  const float* ln_gamma; const float* ln_beta; float* ln_stats; const float* ln_x; const float* u;
  bf16* out2_hi; bf16* out2_lo; int ldS2;
  int tma_store;                              // 1: dense output rows -> the epilogue stores through TMA (see EpiMaps)
  int stg_bufs;                               // staging tiles per epilogue warp (1..3; > 1 only with tma_store)
  int bk;                                     // k-block width in elements: 64 (128-byte rows, SWIZZLE_128B) or, for K <= 32,
                                              // 32 (64-byte rows, SWIZZLE_64B: half the shared-memory fill per stage, twice the stages)
};
// Tensor maps of the outputs (fp32 copy, hi / lo planes, second hi / lo planes) for the TMA store path.  Used when the
// tile's 128 accumulator rows are 128 consecutive output rows (o_L == L_out, o_stride == 1, o_off == 0: every launch of
// the transformer blocks and most of the convolutions); boxes are 32 rows x CW columns = one warp's staging tile.
struct EpiMaps { CUtensorMap f32, hi, lo, hi2, lo2; };

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
// Bounded wait: try_wait (which itself blocks for a hardware-defined slice), a 32 ns back-off after 2^16 failed tries, and
// after 2^26 (>= 2 s, four orders of magnitude beyond the longest legitimate wait of this kernel) a trap, so that a
// protocol bug surfaces as a launch failure through the C ABI (ramp_last_error) instead of a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
#pragma unroll 1
  for (uint32_t spins = 0;; ++spins) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) return;
    if (spins > (1u << 16)) __nanosleep(32);
    if (spins > (1u << 26)) __trap();
  }
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
// TMA store of one staging tile (box = 32 rows x CW columns) at (column c0, row c1) of an output tensor; completion is
// tracked by the issuing thread's bulk async-group
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(map), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read2() { asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// ---- CTA-pair (cta_group::2) variants: M = 256 tile over two SMs of a TPC; each CTA stages its own 128 A rows and HALF of
// the B rows, the leader CTA (cluster rank 0) issues the MMAs for both, commits are multicast to both CTAs' barriers ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// TMA loads of a CTA pair signal the LEADER's mbarrier: same smem offset, peer bit (24) of the cluster address cleared
static constexpr uint32_t PEER_BIT_MASK = 0xFEFFFFFFu;
__device__ __forceinline__ void tma_load_4d_pair(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                                 int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar & PEER_BIT_MASK), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar & PEER_BIT_MASK), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void umma_f16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {        // arrives on `bar` in BOTH CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"((uint16_t)3) : "memory");
}
// arrive on the barrier at the same smem offset in cluster CTA `rank`.  RELAXED on purpose: the arrival only tells the
// leader's MMA issuer that this warp's tcgen05.ld of the accumulator have completed (tcgen05.wait::ld + the tcgen05 fence
// before it); nothing this kernel's other threads read depends on the warp's global stores.  The default .release form
// compiles to MEMBAR.ALL.GPU + ERRBAR in front of the arrive, i.e. every epilogue warp of the pair's second CTA waited once
// per tile for ALL its outstanding global stores to be acknowledged (ncu r2 run K, GEGLU forward: 8 % of the stall samples
// on those three instructions, and the leader's MMAs wait for the slower CTA).
__device__ __forceinline__ void mbar_arrive_remote(uint32_t bar, uint32_t rank) {
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
#if RAMP_GEMM_RELAXED_ARRIVE
      "mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [ra];\n"
#else
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n"
#endif
      "}\n" ::"r"(bar), "r"(rank) : "memory");
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
// sw64: rows of 64 B (k-block of 32 elements), 8-row atoms of 512 B, layout type SWIZZLE_64B (4)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr, bool sw64 = false) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)((sw64 ? 512 : 1024) >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(sw64 ? 4 : 2) << 61;
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
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
        "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
// CW = epilogue chunk width: accumulator columns a thread handles per step (32 or 16)
template <int CW>
__device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t* r) {
  if constexpr (CW == 32) tmem_ld32(taddr, r); else tmem_ld16(taddr, r);
}
template <int CW>
__device__ __forceinline__ void ld_acc(uint32_t taddr, float (&v)[CW]) {
  uint32_t r[CW];
  tmem_ld<CW>(taddr, r);
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < CW; ++i) v[i] = __uint_as_float(r[i]);
}
// two accumulator chunks with one wait
template <int CW>
__device__ __forceinline__ void ld_acc2(uint32_t t0, float (&v0)[CW], uint32_t t1, float (&v1)[CW]) {
  uint32_t r0[CW], r1[CW];
  tmem_ld<CW>(t0, r0);
  tmem_ld<CW>(t1, r1);
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < CW; ++i) { v0[i] = __uint_as_float(r0[i]); v1[i] = __uint_as_float(r1[i]); }
}
template <int CW>
__device__ __forceinline__ void st_acc(uint32_t taddr, const float (&v)[CW]) {
  uint32_t r[CW];
#pragma unroll
  for (int i = 0; i < CW; ++i) r[i] = __float_as_uint(v[i]);
  if constexpr (CW == 32) tmem_st32(taddr, r); else tmem_st16(taddr, r);
}

// ---------------------------------------------------------------------------------------------
// epilogue plumbing: a thread owns accumulator row (q*32 + lane) of the tile and works on it in chunks of CW columns
// (G = CW/4 16-byte groups).  For global traffic the warp re-maps to (row = lane / G + (32/G) i, group = lane % G)
// through its staging tile, so every warp-wide access covers whole row segments of 4 CW bytes.
// ---------------------------------------------------------------------------------------------
// staging tile (32 rows x 4 CW bytes): group k of row r is XOR-swizzled so that both the row-owner accesses and the
// transposed accesses are bank-conflict free
template <int CW>
__device__ __forceinline__ int stg_idx(int r, int k) {
  if constexpr (CW == 32) return r * 8 + (k ^ (r & 7));
  else return r * 4 + (k ^ ((r >> 1) & 3));
}
template <int CW>
struct EpiCtx {
  static constexpr int G = CW / 4, RP = 32 / G, NP = G;   // groups per row, rows per pass, passes
  mutable float* stg;      // this warp's current staging tile
  float* stg0;             // its first staging tile (of nbuf, 32 x CW floats each)
  float* stg_q;            // first staging tile of group 0's warp for this lane quarter (+ 4 * nbuf * 32 * CW floats per group)
  int nbuf; mutable int cur;   // TMA store path: a bulk store leaves its tile busy until the TMA unit has read it, so the
                           // warp rotates over nbuf tiles and only waits until at most nbuf - 1 stores are still reading
  const EpiMaps* maps;     // TMA store path: output tensor maps, else nullptr
  int row0;                // first output row of this warp's 32 accumulator rows (TMA store path)
  int lane, bar_id;
  bool my_ok; size_t my_row; int my_sample;
  unsigned okn; uint32_t rown[NP];
};
// the NG warps (same TMEM lane quarter, one per group) that share a tile's rows meet here
template <int NG>
__device__ __forceinline__ void row_sync(int bar_id) { asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(32 * NG) : "memory"); }

// v[CW] (this thread's row, columns col0..col0+CW-1) -> base[row * ld + col0 ..] for every live row of the warp
// The staging tile may still be read by a TMA store of this warp: wait (elected lane, its bulk group) before writing it
template <int CW>
__device__ __forceinline__ void stg_acquire(const EpiCtx<CW>& e) {
  if (e.maps) {
    // the current tile's last store is nbuf stores back: all but the newest nbuf - 1 groups must have been read
    if (e.lane == 0) { if (e.nbuf == 1) tma_store_wait_read(); else if (e.nbuf == 2) tma_store_wait_read1(); else tma_store_wait_read2(); }
    __syncwarp();
  }
}
// after a bulk store was issued from the current tile: move on to the next one
template <int CW>
__device__ __forceinline__ void stg_rotate(const EpiCtx<CW>& e) {
  e.cur = e.cur + 1 == e.nbuf ? 0 : e.cur + 1;
  e.stg = e.stg0 + e.cur * (32 * CW);
}
template <int CW>
__device__ __forceinline__ void chunk_store_f32(const EpiCtx<CW>& e, const float (&v)[CW], float* base, int ld, int col0,
                                                const CUtensorMap* map = nullptr) {
  using E = EpiCtx<CW>;
  float4* s4 = reinterpret_cast<float4*>(e.stg);
  stg_acquire<CW>(e);
#pragma unroll
  for (int k = 0; k < E::G; ++k) s4[stg_idx<CW>(e.lane, k)] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
  if (map) {
    // row = lane, 16-byte group k at k ^ f(row): exactly the SWIZZLE_128B (CW = 32) / SWIZZLE_64B (CW = 16) image of a
    // [32 rows x CW fp32] box, so the tile goes out as ONE bulk tensor store (rows past the tensor's end are clipped)
    fence_async_smem();
    __syncwarp();
    if (e.lane == 0) {
      if (!(RAMP_GEMM_DIAG & 2)) tma_store_2d(map, smem_u32(e.stg), col0, e.row0);
      tma_store_commit();
    }
    stg_rotate<CW>(e);
    return;
  }
  __syncwarp();
  const int rsub = e.lane / E::G, c4 = e.lane % E::G;
  // all read-backs first, then the stores: with one register quad per pass the LDS -> STG pairs ran strictly one after
  // the other (ncu r2 run K: STG waiting on the short scoreboard = 11 % of the GEGLU launches' stall samples)
  float4 t[E::NP];
#pragma unroll
  for (int i = 0; i < E::NP; ++i) t[i] = s4[stg_idx<CW>(rsub + E::RP * i, c4)];
#pragma unroll
  for (int i = 0; i < E::NP; ++i)
    if (!(RAMP_GEMM_DIAG & 2) && ((e.okn >> i) & 1u)) *reinterpret_cast<float4*>(base + (size_t)e.rown[i] * ld + col0 + 4 * c4) = t[i];
  __syncwarp();
}
// same, as bf16 hi / lo planes (ld, col0 in bf16 elements): staging row = [hi 2 CW bytes | lo 2 CW bytes]
// 16-byte group k of row r in a [32 rows x CW 16-bit] plane tile (2 CW bytes per row): the SWIZZLE_64B (CW = 32) /
// SWIZZLE_32B (CW = 16) image of that box
template <int CW>
__device__ __forceinline__ int plane_idx(int r, int k) {
  if constexpr (CW == 32) return r * 4 + (k ^ ((r >> 1) & 3));
  else return r * 2 + (k ^ ((r >> 2) & 1));
}
template <int CW>
__device__ __forceinline__ void chunk_store_split(const EpiCtx<CW>& e, const float (&v)[CW], bf16* hi, bf16* lo, int ld, int col0,
                                                  const CUtensorMap* map_hi = nullptr, const CUtensorMap* map_lo = nullptr) {
  using E = EpiCtx<CW>;
  constexpr int H = E::G / 2;
  uint4* s4 = reinterpret_cast<uint4*>(e.stg);
  stg_acquire<CW>(e);
  if (map_hi) {
    // the staging tile holds the hi plane tile in its first half and the lo plane tile in its second: two bulk stores
    uint4* l4 = s4 + 32 * H;
#pragma unroll
    for (int k = 0; k < H; ++k) {
      uint32_t h[4], l[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) split2_bf16(v[8 * k + 2 * j], v[8 * k + 2 * j + 1], h[j], l[j]);
      s4[plane_idx<CW>(e.lane, k)] = make_uint4(h[0], h[1], h[2], h[3]);
      l4[plane_idx<CW>(e.lane, k)] = make_uint4(l[0], l[1], l[2], l[3]);
    }
    fence_async_smem();
    __syncwarp();
    if (e.lane == 0) {
      if (!(RAMP_GEMM_DIAG & 2)) {
        tma_store_2d(map_hi, smem_u32(s4), col0, e.row0);
        tma_store_2d(map_lo, smem_u32(l4), col0, e.row0);
      }
      tma_store_commit();
    }
    stg_rotate<CW>(e);
    return;
  }
#pragma unroll
  for (int k = 0; k < H; ++k) {
    uint32_t h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) split2_bf16(v[8 * k + 2 * j], v[8 * k + 2 * j + 1], h[j], l[j]);
    s4[stg_idx<CW>(e.lane, k)] = make_uint4(h[0], h[1], h[2], h[3]);
    s4[stg_idx<CW>(e.lane, H + k)] = make_uint4(l[0], l[1], l[2], l[3]);
  }
  __syncwarp();
  const int rsub = e.lane / E::G, c4 = e.lane % E::G;
  bf16* plane = (c4 < H) ? hi : lo;
  uint4 t[E::NP];
#pragma unroll
  for (int i = 0; i < E::NP; ++i) t[i] = s4[stg_idx<CW>(rsub + E::RP * i, c4)];
#pragma unroll
  for (int i = 0; i < E::NP; ++i)
    if (!(RAMP_GEMM_DIAG & 2) && ((e.okn >> i) & 1u)) *reinterpret_cast<uint4*>(plane + (size_t)e.rown[i] * ld + col0 + 8 * (c4 % H)) = t[i];
  __syncwarp();
}
// Loads come in two phases: `chunk_ld_issue` puts the row-contiguous 16-byte loads of a chunk in flight (registers),
// `chunk_ld_finish` transposes them through the staging tile into the thread's own row (zeros for dead rows).
// Issuing the next chunk's loads before the current chunk's arithmetic and stores is what hides HBM latency here.
template <int CW>
__device__ __forceinline__ void chunk_ld_issue(const EpiCtx<CW>& e, const float* __restrict__ base, int ld, int col0,
                                               float4 (&t)[CW / 4]) {
  using E = EpiCtx<CW>;
  const int c4 = e.lane % E::G;
#pragma unroll
  for (int i = 0; i < E::NP; ++i) {
    t[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!(RAMP_GEMM_DIAG & 1) && ((e.okn >> i) & 1u)) t[i] = __ldg(reinterpret_cast<const float4*>(base + (size_t)e.rown[i] * ld + col0 + 4 * c4));
  }
}
template <int CW>
__device__ __forceinline__ void chunk_ld_finish(const EpiCtx<CW>& e, const float4 (&t)[CW / 4], float (&v)[CW]) {
  using E = EpiCtx<CW>;
  const int rsub = e.lane / E::G, c4 = e.lane % E::G;
  float4* s4 = reinterpret_cast<float4*>(e.stg);
  stg_acquire<CW>(e);
#pragma unroll
  for (int i = 0; i < E::NP; ++i) s4[stg_idx<CW>(rsub + E::RP * i, c4)] = t[i];
  __syncwarp();
#pragma unroll
  for (int k = 0; k < E::G; ++k) {
    const float4 x = s4[stg_idx<CW>(e.lane, k)];
    v[4 * k] = x.x; v[4 * k + 1] = x.y; v[4 * k + 2] = x.z; v[4 * k + 3] = x.w;
  }
  __syncwarp();
}
template <int CW>
__device__ __forceinline__ void chunk_load_f32(const EpiCtx<CW>& e, const float* __restrict__ base, int ld, int col0, float (&v)[CW]) {
  float4 t[CW / 4];
  chunk_ld_issue<CW>(e, base, ld, col0, t);
  chunk_ld_finish<CW>(e, t, v);
}
template <int CW>
__device__ __forceinline__ void add_vec(float (&v)[CW], const float* __restrict__ b) {
#pragma unroll
  for (int i = 0; i < CW; i += 4) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(b + i));
    v[i] += t.x; v[i + 1] += t.y; v[i + 2] += t.z; v[i + 3] += t.w;
  }
}
// acc + bias + per-scene bias + residual(s) for columns col .. col+CW-1; the residual chunk is carried in flight: `t`
// holds the loads of columns `col` on entry and of `next_col` on exit
template <int CW>
__device__ __forceinline__ void epi_affine_pf(const GemmParams& p, const EpiCtx<CW>& e, float (&v)[CW], int col,
                                              float4 (&t)[CW / 4], int next_col, bool has_next) {
  if (p.bias) add_vec<CW>(v, p.bias + col);
  if (p.bias_tab && e.my_ok) add_vec<CW>(v, p.bias_tab + (size_t)p.row2scene[e.my_sample] * p.Nout + col);
  if (p.residual) {
    float r[CW];
    chunk_ld_finish<CW>(e, t, r);
    if (has_next) chunk_ld_issue<CW>(e, p.residual, p.ldR, next_col, t);
#pragma unroll
    for (int i = 0; i < CW; ++i) v[i] += r[i];
  }
  if (p.residual2) {
    float r[CW];
    chunk_load_f32<CW>(e, p.residual2, p.ldR2, col, r);
#pragma unroll
    for (int i = 0; i < CW; ++i) v[i] += r[i];
  }
}

// ---- one tile's epilogue, by mode.  acc = TMEM address of this warp's lane quarter, column 0 of the tile;
//      [c0, c1) = the tile columns this warp's group drains ----
template <int CW>
__device__ __forceinline__ void epi_plain(const GemmParams& p, const EpiCtx<CW>& e, uint32_t acc, int colbase, int c0, int c1) {
  float4 t[CW / 4];
  if (p.residual) chunk_ld_issue<CW>(e, p.residual, p.ldR, colbase + c0, t);
#pragma unroll 1
  for (int c = c0; c < c1; c += CW) {
    float v[CW];
    ld_acc<CW>(acc + c, v);
    epi_affine_pf<CW>(p, e, v, colbase + c, t, colbase + c + CW, c + CW < c1);
    if (p.out_f32) chunk_store_f32<CW>(e, v, p.out_f32, p.ldO, colbase + c, e.maps ? &e.maps->f32 : nullptr);
    if (p.out_hi) chunk_store_split<CW>(e, v, p.out_hi, p.out_lo, p.ldS, colbase + c, e.maps ? &e.maps->hi : nullptr, e.maps ? &e.maps->lo : nullptr);
  }
}
// GEGLU forward (layers_attention_mini.py:38-45): tile columns [0,128) = a, [128,256) = gate; a group drains the same
// sub-range of both
template <int CW>
__device__ __forceinline__ void epi_geglu_fwd(const GemmParams& p, const EpiCtx<CW>& e, uint32_t acc, int tile_n, int c0, int c1) {
  const int colbase = tile_n * 256;
#pragma unroll 1
  for (int c = c0 >> 1; c < (c1 >> 1); c += CW) {
    float a[CW], g[CW], s_[CW];
    ld_acc2<CW>(acc + c, a, acc + 128 + c, g);
    if (p.bias) { add_vec<CW>(a, p.bias + colbase + c); add_vec<CW>(g, p.bias + colbase + 128 + c); }
    // s = gelu(gate), q = a * gelu'(gate): everything the backward needs, so it is two multiplies per element there
#pragma unroll
    for (int i = 0; i < CW; ++i) {
      float cdf, pdf;
      normal_cdf_pdf(g[i], cdf, pdf);
      const float sgl = g[i] * cdf;
      g[i] = a[i] * (cdf + g[i] * pdf);
      a[i] *= sgl;
      s_[i] = sgl;
    }
    if (p.out_f32) {
      chunk_store_f32<CW>(e, s_, p.out_f32, p.ldO, colbase + c, e.maps ? &e.maps->f32 : nullptr);
      chunk_store_f32<CW>(e, g, p.out_f32, p.ldO, colbase + 128 + c, e.maps ? &e.maps->f32 : nullptr);
    }
    chunk_store_split<CW>(e, a, p.out2_hi, p.out2_lo, p.ldS2, tile_n * 128 + c, e.maps ? &e.maps->hi2 : nullptr, e.maps ? &e.maps->lo2 : nullptr);
  }
}
// GEGLU backward: acc = d(a * gelu(gate)) for natural columns colbase + c; u = (gelu(gate) | a gelu'(gate)) and du are
// tile-interleaved
template <int CW>
__device__ __forceinline__ void epi_geglu_bwd(const GemmParams& p, const EpiCtx<CW>& e, uint32_t acc, int colbase, int c0, int c1) {
  auto u_col = [&](int c) { const int j = colbase + c; return (j >> 7) * 256 + (j & 127); };
  // TWO chunk loads in flight per thread (gelu(gate) and a gelu'(gate) of the current chunk, then of the next one): this
  // launch is HBM-bound on the 8 KB / token of u it reads, and with one 128-byte load per thread in flight the 8 epilogue
  // warps covered only half of the bandwidth-delay product (ncu r1: long_scoreboard 38 %, 4.2 of 6.5 TB/s)
  float4 ts[CW / 4], tq[CW / 4];
  chunk_ld_issue<CW>(e, p.u, 2048, u_col(c0), ts);
  chunk_ld_issue<CW>(e, p.u, 2048, u_col(c0) + 128, tq);
#pragma unroll 1
  for (int c = c0; c < c1; c += CW) {
    float d[CW], x[CW];
    const int pa = u_col(c), pg = pa + 128;
    const bool more = c + CW < c1;
    ld_acc<CW>(acc + c, d);
    chunk_ld_finish<CW>(e, ts, x);
    if (more) chunk_ld_issue<CW>(e, p.u, 2048, u_col(c + CW), ts);
#pragma unroll
    for (int i = 0; i < CW; ++i) x[i] *= d[i];                      // u = (gelu(gate) | a * gelu'(gate))
    chunk_store_split<CW>(e, x, p.out_hi, p.out_lo, p.ldS, pa, e.maps ? &e.maps->hi : nullptr, e.maps ? &e.maps->lo : nullptr);
    chunk_ld_finish<CW>(e, tq, x);
    if (more) chunk_ld_issue<CW>(e, p.u, 2048, u_col(c + CW) + 128, tq);
#pragma unroll
    for (int i = 0; i < CW; ++i) x[i] *= d[i];
    chunk_store_split<CW>(e, x, p.out_hi, p.out_lo, p.ldS, pg, e.maps ? &e.maps->hi : nullptr, e.maps ? &e.maps->lo : nullptr);
  }
}
// The NG threads that own the parts of a 256-channel row publish a float2 each (through their warps' staging tiles,
// idle at that point) and read all of them back: everyone combines them in the same order, so they agree to the bit.
template <int CW, int NG>
__device__ __forceinline__ void row_exchange(const EpiCtx<CW>& e, float2 mine, float2 (&all)[NG]) {
  if (e.maps) {                                  // the exchange goes through tile 0 of every warp: nothing may still read it
    if (e.lane == 0) tma_store_wait_read();
    __syncwarp();
  }
  reinterpret_cast<float2*>(e.stg0)[e.lane] = mine;
  row_sync<NG>(e.bar_id);
#pragma unroll
  for (int g = 0; g < NG; ++g) all[g] = reinterpret_cast<const float2*>(e.stg_q + g * (4 * e.nbuf * 32 * CW))[e.lane];
  row_sync<NG>(e.bar_id);                        // nobody re-uses a staging tile before all have read it
}
// LayerNorm forward over the 256 channels of a row (eps 1e-5, layers_attention_mini.py:138-140); this thread drains
// columns [c0, c1) = 256 / NG of them
template <int CW, int NG>
__device__ __forceinline__ void epi_ln_fwd(const GemmParams& p, const EpiCtx<CW>& e, uint32_t acc, int c0, int c1) {
  float shift = 0.f, s = 0.f, ss = 0.f;
  float4 t[CW / 4];
  if (p.residual) chunk_ld_issue<CW>(e, p.residual, p.ldR, c0, t);
#pragma unroll 1
  for (int c = c0; c < c1; c += CW) {
    float v[CW];
    ld_acc<CW>(acc + c, v);
    epi_affine_pf<CW>(p, e, v, c, t, c + CW, c + CW < c1);
    if (p.out_f32) chunk_store_f32<CW>(e, v, p.out_f32, p.ldO, c, e.maps ? &e.maps->f32 : nullptr);
    if (p.out_hi) chunk_store_split<CW>(e, v, p.out_hi, p.out_lo, p.ldS, c, e.maps ? &e.maps->hi : nullptr, e.maps ? &e.maps->lo : nullptr);
    st_acc<CW>(acc + c, v);
    if (c == c0) shift = v[0];
#pragma unroll
    for (int i = 0; i < CW; ++i) { const float d = v[i] - shift; s += d; ss = fmaf(d, d, ss); }
  }
  tmem_st_wait();
  // (mean, M2) of this thread's 256 / NG values, merged pairwise with the other parts (Chan's update)
  constexpr float n = 256.f / NG;
  const float ml = s * (1.f / n);
  float2 part[NG];
  row_exchange<CW, NG>(e, make_float2(shift + ml, fmaxf(ss - s * ml, 0.f)), part);
  float mean, m2;
  if constexpr (NG == 2) {
    const float dm = part[1].x - part[0].x;
    mean = part[0].x + 0.5f * dm;
    m2 = part[0].y + part[1].y + dm * dm * (0.5f * n);
  } else {
    const float d01 = part[1].x - part[0].x, d23 = part[3].x - part[2].x;
    const float m01 = part[0].x + 0.5f * d01, m23 = part[2].x + 0.5f * d23;
    const float q01 = part[0].y + part[1].y + d01 * d01 * (0.5f * n), q23 = part[2].y + part[3].y + d23 * d23 * (0.5f * n);
    const float dm = m23 - m01;
    mean = m01 + 0.5f * dm;
    m2 = q01 + q23 + dm * dm * n;
  }
  const float rstd = rsqrtf(m2 * (1.f / 256.f) + 1e-5f);
  if (p.ln_stats && e.my_ok && c0 == 0) *reinterpret_cast<float2*>(p.ln_stats + e.my_row * 2) = make_float2(mean, rstd);
#pragma unroll 1
  for (int c = c0; c < c1; c += CW) {
    float v[CW];
    ld_acc<CW>(acc + c, v);
#pragma unroll
    for (int i = 0; i < CW; i += 4) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(p.ln_gamma + c + i));
      const float4 b = __ldg(reinterpret_cast<const float4*>(p.ln_beta + c + i));
      v[i] = (v[i] - mean) * rstd * g.x + b.x; v[i + 1] = (v[i + 1] - mean) * rstd * g.y + b.y;
      v[i + 2] = (v[i + 2] - mean) * rstd * g.z + b.z; v[i + 3] = (v[i + 3] - mean) * rstd * g.w + b.w;
    }
    chunk_store_split<CW>(e, v, p.out2_hi, p.out2_lo, p.ldS2, c, e.maps ? &e.maps->hi2 : nullptr, e.maps ? &e.maps->lo2 : nullptr);
  }
}
// out = residual + LayerNorm_backward(ln_x, acc): dx = rstd * (dyh - mean(dyh) - xh * mean(dyh * xh)), dyh = acc * gamma
template <int CW, int NG>
__device__ __forceinline__ void epi_ln_bwd(const GemmParams& p, const EpiCtx<CW>& e, uint32_t acc, int c0, int c1) {
  float mean = 0.f, rstd = 0.f;
  if (e.my_ok) { const float2 st = *reinterpret_cast<const float2*>(p.ln_stats + e.my_row * 2); mean = st.x; rstd = st.y; }
  float s1 = 0.f, s2 = 0.f;
  // two chunk loads in flight per thread throughout (ncu r1: long_scoreboard 63 % with one): pass 1 keeps ln_x of the
  // current and the next chunk in flight, pass 2 the ln_x and the residual chunk of the current column range
  float4 ta[CW / 4], tb[CW / 4];
  chunk_ld_issue<CW>(e, p.ln_x, 256, c0, ta);
  if (c0 + CW < c1) chunk_ld_issue<CW>(e, p.ln_x, 256, c0 + CW, tb);
  bool use_a = true;
#pragma unroll 1
  for (int c = c0; c < c1; c += CW) {
    float d[CW], x[CW];
    ld_acc<CW>(acc + c, d);
    // chunk c sits in ta (even iterations) or tb (odd); refill the buffer just consumed with chunk c + 2 CW
    if (use_a) { chunk_ld_finish<CW>(e, ta, x); if (c + 2 * CW < c1) chunk_ld_issue<CW>(e, p.ln_x, 256, c + 2 * CW, ta); }
    else { chunk_ld_finish<CW>(e, tb, x); if (c + 2 * CW < c1) chunk_ld_issue<CW>(e, p.ln_x, 256, c + 2 * CW, tb); }
    use_a = !use_a;
#pragma unroll
    for (int i = 0; i < CW; i += 4) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(p.ln_gamma + c + i));
      d[i] *= g.x; d[i + 1] *= g.y; d[i + 2] *= g.z; d[i + 3] *= g.w;
    }
#pragma unroll
    for (int i = 0; i < CW; ++i) { s1 += d[i]; s2 = fmaf(d[i], (x[i] - mean) * rstd, s2); }
    st_acc<CW>(acc + c, d);
  }
  // pass 2's first loads go out before the row statistics are exchanged
  chunk_ld_issue<CW>(e, p.ln_x, 256, c0, ta);
  if (p.residual) chunk_ld_issue<CW>(e, p.residual, p.ldR, c0, tb);
  tmem_st_wait();
  {
    float2 part[NG];
    row_exchange<CW, NG>(e, make_float2(s1, s2), part);
    if constexpr (NG == 2) { s1 = part[0].x + part[1].x; s2 = part[0].y + part[1].y; }
    else { s1 = (part[0].x + part[1].x) + (part[2].x + part[3].x); s2 = (part[0].y + part[1].y) + (part[2].y + part[3].y); }
    s1 *= (1.f / 256.f); s2 *= (1.f / 256.f);
  }
#pragma unroll 1
  for (int c = c0; c < c1; c += CW) {
    float d[CW], x[CW];
    const bool more = c + CW < c1;
    ld_acc<CW>(acc + c, d);
    chunk_ld_finish<CW>(e, ta, x);
    if (more) chunk_ld_issue<CW>(e, p.ln_x, 256, c + CW, ta);
#pragma unroll
    for (int i = 0; i < CW; ++i) d[i] = rstd * (d[i] - s1 - (x[i] - mean) * rstd * s2);
    if (p.residual) {
      chunk_ld_finish<CW>(e, tb, x);
      if (more) chunk_ld_issue<CW>(e, p.residual, p.ldR, c + CW, tb);
#pragma unroll
      for (int i = 0; i < CW; ++i) d[i] += x[i];
    }
    if (p.out_f32) chunk_store_f32<CW>(e, d, p.out_f32, p.ldO, c, e.maps ? &e.maps->f32 : nullptr);
    if (p.out_hi) chunk_store_split<CW>(e, d, p.out_hi, p.out_lo, p.ldS, c, e.maps ? &e.maps->hi : nullptr, e.maps ? &e.maps->lo : nullptr);
  }
}

// ---------------------------------------------------------------------------------------------
// kernel
// ---------------------------------------------------------------------------------------------
// One instantiation per epilogue mode: each gets its own register allocation (the modes' live sets differ a lot).
// CG = CTAs per tile: 1, or 2 = a CTA pair (cluster of two) on one 256 x BN tile (tcgen05 cta_group::2): every CTA stages its
// own 128 A rows but only BN / 2 rows of the weights, which cuts the L2 -> SM operand traffic per MMA by a third (the cap
// that bounds the weight-heavy launches: ~6300 B / clk chip-wide, B300_MICROARCH.md; FF1^T at L = 64 ran at 9.5 TB/s).
template <int EPI, int CW, int CG>
__global__ void __launch_bounds__(gemm_threads(CW), 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap mapAhi, const __grid_constant__ CUtensorMap mapAlo,
               const __grid_constant__ CUtensorMap mapBhi, const __grid_constant__ CUtensorMap mapBlo,
               const __grid_constant__ EpiMaps omaps, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cta_rank = CG == 2 ? (int)cluster_ctarank() : 0;
  const int a_tile_bytes = BM * p.bk * 2;
  const int b_tile_bytes = (p.BN / CG) * p.bk * 2;      // this CTA's share of the weight tile
  const int stage_bytes = p.nplanes * (a_tile_bytes + b_tile_bytes);
  const uint32_t stg_base = smem_base + p.stages * stage_bytes;            // the epilogue warps' staging tiles
  const uint32_t bar_base = stg_base + p.stg_bufs * STG_BYTES;
  const uint32_t full0 = bar_base, empty0 = bar_base + 8 * p.stages;
  const uint32_t tfull0 = bar_base + 16 * p.stages, tempty0 = tfull0 + 16;  // 2 + 2 accumulator barriers
  const uint32_t tmem_slot = tempty0 + 16;
  const int iters = p.ntaps * p.kblocks;
  // work units: (tile pair along M) x n-tile; unit u -> m-tile CG * (u / n_tiles) + cta_rank (an odd tile count leaves the
  // last pair's second CTA an out-of-range tile: its TMA boxes are zero-filled, its rows masked in the epilogue)
  const int total_tiles = ((p.m_tiles + CG - 1) / CG) * p.n_tiles;
  const int unit0 = blockIdx.x / CG, unit_step = gridDim.x / CG;
  uint32_t tmem_cols = 32;
  while ((int)tmem_cols < 2 * p.BN) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(tfull0 + 8 * b, 1); mbar_init(tempty0 + 8 * b, 4 * p.ngroups * CG); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapAhi) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapBhi) : "memory");
  }
  if (warp == 1) {
    if constexpr (CG == 2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(tmem_cols)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(tmem_cols)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();     // barriers and TMEM of BOTH CTAs exist from here on
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t tmem_acc;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_acc) : "r"(tmem_slot));
  // Programmatic dependent launch: everything above (barriers, TMEM, descriptor prefetch) touches no global data and may
  // run while the previous kernel of the stream drains; from here on its results are needed.  Both instructions are no-ops
  // when the launch carries no programmatic-serialization attribute.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");

#if RAMP_GEMM_WG
  if (warp < 4) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
#endif
  if (warp == 0) {
    // ---------------- TMA producer ----------------
    if (lane == 0) {
      int it = 0;
      for (int tile = unit0; tile < total_tiles; tile += unit_step) {
        const int pair_m = tile / p.n_tiles, tile_n = tile - pair_m * p.n_tiles;
        const int c3 = (pair_m * CG + cta_rank) * p.nb;
        for (int t = 0; t < p.ntaps; ++t) {
          for (int kb = 0; kb < p.kblocks; ++kb, ++it) {
            const int s = it % p.stages;
            const uint32_t ph = (it / p.stages) & 1;
            mbar_wait(empty0 + 8 * s, ph ^ 1);
            const uint32_t st = smem_base + s * stage_bytes;
            const uint32_t fb = full0 + 8 * s;
            const int brow = p.tap_wrow[t] + tile_n * p.BN + cta_rank * (p.BN / CG);
            if constexpr (CG == 2) {
              // the leader's barrier collects the bytes of both CTAs; only the leader arrives on it
              if (cta_rank == 0) mbar_expect_tx(fb, (uint32_t)(2 * p.nplanes * (a_tile_bytes + b_tile_bytes)));
              tma_load_4d_pair(st, &mapAhi, fb, kb * p.bk, p.tap_par[t], p.tap_joff[t], c3);
              uint32_t off = a_tile_bytes;
              if (p.nplanes == 2) { tma_load_4d_pair(st + off, &mapAlo, fb, kb * p.bk, p.tap_par[t], p.tap_joff[t], c3); off += a_tile_bytes; }
              tma_load_2d_pair(st + off, &mapBhi, fb, kb * p.bk, brow);
              off += b_tile_bytes;
              if (p.nplanes == 2) tma_load_2d_pair(st + off, &mapBlo, fb, kb * p.bk, brow);
            } else {
              mbar_expect_tx(fb, (uint32_t)(p.nplanes * (p.nb * p.L_out * p.bk * 2 + b_tile_bytes)));
              tma_load_4d(st, &mapAhi, fb, kb * p.bk, p.tap_par[t], p.tap_joff[t], c3);
              uint32_t off = a_tile_bytes;
              if (p.nplanes == 2) { tma_load_4d(st + off, &mapAlo, fb, kb * p.bk, p.tap_par[t], p.tap_joff[t], c3); off += a_tile_bytes; }
              tma_load_2d(st + off, &mapBhi, fb, kb * p.bk, brow);
              off += b_tile_bytes;
              if (p.nplanes == 2) tma_load_2d(st + off, &mapBlo, fb, kb * p.bk, brow);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer ----------------
    if (lane == 0 && cta_rank == 0) {       // a CTA pair's MMAs are issued by its leader alone
      // instruction descriptor: D = f32 (bit 4), A format bits [7,10), B format bits [10,13) (0 = f16, 1 = bf16), both
      // K-major, N = BN, M = 128 per CTA.  A and B of one MMA must share a format (a mixed f16 x bf16 descriptor is an
      // illegal instruction on B200).
      const uint32_t idesc0 = (1u << 4) | ((uint32_t)(p.BN >> 3) << 17) | ((uint32_t)((BM * CG) >> 4) << 24);
#if RAMP_PLANES_FP16
      const uint32_t idesc_lh = idesc0, idesc_hl = idesc0, idesc_hh = idesc0;                 // all planes fp16
#else
      const uint32_t idesc_lh = idesc0 | (1u << 7) | (1u << 10), idesc_hl = idesc_lh, idesc_hh = idesc_lh;   // all bf16
#endif
      const int ksteps = p.bk / 16;
      const bool sw64 = p.bk == 32;
      int it = 0, lt = 0;
      for (int tile = unit0; tile < total_tiles; tile += unit_step, ++lt) {
        const int buf = lt & 1;
        mbar_wait(tempty0 + 8 * buf, ((lt >> 1) & 1) ^ 1);      // epilogue (of both CTAs) has drained this accumulator
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t dacc = tmem_acc + (uint32_t)(buf * p.BN);
        for (int i = 0; i < iters; ++i, ++it) {
          const int s = it % p.stages;
          const uint32_t ph = (it / p.stages) & 1;
          mbar_wait(full0 + 8 * s, ph);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t st = smem_base + s * stage_bytes;
          const uint32_t a_hi = st;
          const uint32_t a_lo = st + a_tile_bytes;
          const uint32_t b_hi = st + p.nplanes * a_tile_bytes;
          const uint32_t b_lo = b_hi + b_tile_bytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            if (k >= ksteps) break;
            const uint32_t ko = k * 32;   // 16 bf16 = 32 B inside the 128 B swizzle row
            const uint32_t first = (i == 0 && k == 0) ? 0u : 1u;
            if constexpr (CG == 2) {
              if (p.nplanes == 2) {
                umma_f16_pair(dacc, umma_desc_sw128(a_lo + ko, sw64), umma_desc_sw128(b_hi + ko, sw64), idesc_lh, first);
                umma_f16_pair(dacc, umma_desc_sw128(a_hi + ko, sw64), umma_desc_sw128(b_lo + ko, sw64), idesc_hl, 1u);
                umma_f16_pair(dacc, umma_desc_sw128(a_hi + ko, sw64), umma_desc_sw128(b_hi + ko, sw64), idesc_hh, 1u);
              } else {
                umma_f16_pair(dacc, umma_desc_sw128(a_hi + ko, sw64), umma_desc_sw128(b_hi + ko, sw64), idesc_hh, first);
              }
            } else if (p.nplanes == 2) {
              umma_bf16(dacc, umma_desc_sw128(a_lo + ko, sw64), umma_desc_sw128(b_hi + ko, sw64), idesc_lh, first);
              umma_bf16(dacc, umma_desc_sw128(a_hi + ko, sw64), umma_desc_sw128(b_lo + ko, sw64), idesc_hl, 1u);
              umma_bf16(dacc, umma_desc_sw128(a_hi + ko, sw64), umma_desc_sw128(b_hi + ko, sw64), idesc_hh, 1u);
            } else {
              umma_bf16(dacc, umma_desc_sw128(a_hi + ko, sw64), umma_desc_sw128(b_hi + ko, sw64), idesc_hh, first);
            }
          }
          if constexpr (CG == 2) umma_commit_pair(empty0 + 8 * s); else umma_commit(empty0 + 8 * s);   // frees the smem stage (of both CTAs) when these MMAs retire
        }
        if constexpr (CG == 2) umma_commit_pair(tfull0 + 8 * buf); else umma_commit(tfull0 + 8 * buf);   // accumulator complete
      }
    }
  }
#if RAMP_GEMM_WG
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(epi_regs(CW)));
#else
  if (warp >= EPI_W0) {
#endif
    // ---------------- epilogue: TMEM lane quarter q = warp % 4, group = (warp - EPI_W0) / 4 ----------------
    constexpr int NG = epi_warps(CW) / 4;
    const int q = warp & 3, group = (warp - EPI_W0) >> 2;
    if (group < p.ngroups) {
      float* stg_all = reinterpret_cast<float*>(smem_gen + p.stages * stage_bytes);
      EpiCtx<CW> e;
      using E = EpiCtx<CW>;
      e.lane = lane;
      e.maps = p.tma_store ? &omaps : nullptr;
      e.bar_id = 1 + q;
      e.nbuf = p.stg_bufs; e.cur = 0;
      e.stg0 = stg_all + (warp - EPI_W0) * (p.stg_bufs * 32 * CW);
      e.stg = e.stg0;
      e.stg_q = stg_all + ((warp - EPI_W0) & 3) * (p.stg_bufs * 32 * CW);
      const int cols = p.BN / p.ngroups;                      // tile columns this group drains
      const int c0 = group * cols, c1 = c0 + cols;
      const int r = q * 32 + lane;
      const int n_local = r / p.L_out, l = r - n_local * p.L_out;
      int lt = 0;
      for (int tile = unit0; tile < total_tiles; tile += unit_step, ++lt) {
        const int pair_m = tile / p.n_tiles, tile_n = tile - pair_m * p.n_tiles;
        const int tile_m = pair_m * CG + cta_rank;
        const int buf = lt & 1;
        e.my_sample = tile_m * p.nb + n_local;
        e.my_ok = n_local < p.nb && e.my_sample < p.n_samples;
        e.my_row = (size_t)e.my_sample * p.o_L + (size_t)l * p.o_stride + p.o_off;
        e.row0 = tile_m * BM + q * 32;            // dense outputs: accumulator row r of tile m is output row m * 128 + r
        e.okn = 0;
#pragma unroll
        for (int i = 0; i < E::NP; ++i) {
          const int rr = q * 32 + lane / E::G + E::RP * i;
          const int nl = rr / p.L_out, ll = rr - nl * p.L_out;
          const int smp = tile_m * p.nb + nl;
          if (nl < p.nb && smp < p.n_samples) e.okn |= 1u << i;
          e.rown[i] = (uint32_t)(smp * p.o_L + ll * p.o_stride + p.o_off);
        }
        mbar_wait(tfull0 + 8 * buf, (lt >> 1) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t acc = tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * p.BN);
        if (EPI == EPI_PLAIN) epi_plain<CW>(p, e, acc, tile_n * p.BN, c0, c1);
        else if (EPI == EPI_GEGLU_FWD) epi_geglu_fwd<CW>(p, e, acc, tile_n, c0, c1);
        else if (EPI == EPI_GEGLU_BWD) epi_geglu_bwd<CW>(p, e, acc, tile_n * p.BN, c0, c1);
        else if (EPI == EPI_LN_FWD) epi_ln_fwd<CW, NG>(p, e, acc, c0, c1);
        else epi_ln_bwd<CW, NG>(p, e, acc, c0, c1);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
          if (CG == 2 && cta_rank != 0) mbar_arrive_remote(tempty0 + 8 * buf, 0);      // the leader's issuer waits for both CTAs
          else mbar_arrive(tempty0 + 8 * buf);
        }
      }
      if (p.tma_store && lane == 0) tma_store_wait_all();      // shared memory must outlive the last bulk stores
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if constexpr (CG == 2)
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_acc), "r"(tmem_cols) : "memory");
    else
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
// tuning knob (environment, read once): RAMP_GEMM_L2PROMO=128|256 picks the tensor-map L2 promotion
static int g_l2promo = 256;
// RAMP_GEMM_PAIR: 0 = every tile on one CTA; 1 = CTA pairs (cta_group::2) for launches with BN >= 128 and enough M tiles
#ifndef RAMP_GEMM_PAIR_DEFAULT
#define RAMP_GEMM_PAIR_DEFAULT 1      // measured on B200 (profiles/r2_cta_pair_ab.txt): +4.1 % on the headline, LN-backward launches -15 %
#endif
static int g_pair = RAMP_GEMM_PAIR_DEFAULT;
// RAMP_GEMM_TMA_STORE: 1 = launches whose tile rows are consecutive output rows store through TMA (bulk tensor stores of
// the epilogue warps' staging tiles); 0 = every launch uses the LDS -> STG read-back path (round 1 / A/B runs)
static int g_tma_store = 1;
// RAMP_GEMM_STG_BUFS = "e0,e1,e2,e3,e4": staging tiles per epilogue warp wanted by each epilogue mode (TMA store path; each
// extra tile costs 32 KB of the operand ring).  RAMP_GEMM_MIN_STAGES: operand stages that must remain.
static int g_stg_want[5] = {1, 1, 2, 2, 2};
static int g_min_stages = 2;
// RAMP_GEMM_BN_MODEL: 1 = pick the tile width by the wave model in pick_bn, 0 = round 1's rule (largest tile that divides
// Nout, smaller only while the grid would leave SMs idle)
static int g_bn_model = 1;
// RAMP_GEMM_BK32: 1 = contractions with K <= 32 (the 16- / 32-channel convolutions, proj_in at C = 32) use 32-wide k-blocks:
// 64-byte operand rows (SWIZZLE_64B), half the shared-memory fill per stage and up to ten stages in flight.  Those launches
// are bound by the TMA fill of five taps per tile, not by the MMAs: 32-channel k5 convolution at L = 32 76.6 -> 57.0 us,
// 16 -> 32 channels at L = 64 136 -> 98 us (profiles/r2_bk32_ab.txt)
static int g_bk32 = 1;
// RAMP_GEMM_E1_CW: epilogue configuration of the GEGLU forward (16 = 16 warps x 16-column chunks, 32 = 8 warps x 32)
static int g_e1_cw = 16;
// a launch gets its mode's extra staging tiles only while its k-loop is short (iterations per tile <= limit): long k-loops
// are tensor-bound and want the operand stages instead (profiles/r2_stg_bufs.txt)
static int g_stg_max_iters[5] = {0, 1 << 30, 1 << 30, 4, 12};

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
  const int dyn = g_max_smem;
#define RAMP_GEMM_ATTR(E, C)                                                                                         \
  RAMP_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<E, C, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));      \
  RAMP_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<E, C, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
  RAMP_GEMM_ATTR(EPI_PLAIN, 32) RAMP_GEMM_ATTR(EPI_PLAIN, 16) RAMP_GEMM_ATTR(EPI_GEGLU_FWD, 16) RAMP_GEMM_ATTR(EPI_GEGLU_FWD, 32)
  RAMP_GEMM_ATTR(EPI_GEGLU_BWD, 32) RAMP_GEMM_ATTR(EPI_LN_FWD, 32) RAMP_GEMM_ATTR(EPI_LN_BWD, 32)
#undef RAMP_GEMM_ATTR
  if (const char* v = getenv("RAMP_GEMM_L2PROMO")) g_l2promo = atoi(v);
  if (const char* v = getenv("RAMP_GEMM_PAIR")) g_pair = atoi(v);
  if (const char* v = getenv("RAMP_GEMM_TMA_STORE")) g_tma_store = atoi(v);
  if (const char* v = getenv("RAMP_GEMM_STG_BUFS")) {
    int w[5];
    if (sscanf(v, "%d,%d,%d,%d,%d", &w[0], &w[1], &w[2], &w[3], &w[4]) == 5)
      for (int i = 0; i < 5; ++i) g_stg_want[i] = w[i] < 1 ? 1 : (w[i] > 3 ? 3 : w[i]);
  }
  if (const char* v = getenv("RAMP_GEMM_BN_MODEL")) g_bn_model = atoi(v) != 0;
  if (const char* v = getenv("RAMP_GEMM_BK32")) g_bk32 = atoi(v) != 0;
  if (const char* v = getenv("RAMP_GEMM_E1_CW")) g_e1_cw = atoi(v) == 32 ? 32 : 16;
  if (const char* v = getenv("RAMP_GEMM_STG_ITERS")) {
    int w[5];
    if (sscanf(v, "%d,%d,%d,%d,%d", &w[0], &w[1], &w[2], &w[3], &w[4]) == 5) for (int i = 0; i < 5; ++i) g_stg_max_iters[i] = w[i];
  }
  if (const char* v = getenv("RAMP_GEMM_MIN_STAGES")) { g_min_stages = atoi(v); if (g_min_stages < 1) g_min_stages = 1; }
  g_encode = (EncodeTiledFn)fn;
  return 0;
}

struct MapKey {
  const void* ptr; uint64_t d[4]; uint64_t s[3]; uint32_t box[4]; int rank; int dtype; int swizzle; int pad_;
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
                    const uint32_t* box, CUtensorMapDataType dtype = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                    CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
  MapKey key;
  std::memset(&key, 0, sizeof(key));
  key.ptr = ptr; key.rank = rank; key.dtype = (int)dtype; key.swizzle = (int)swizzle;
  for (int i = 0; i < rank; ++i) { key.d[i] = dims[i]; key.box[i] = box[i]; }
  for (int i = 0; i < rank - 1; ++i) key.s[i] = strides_bytes[i];
  std::lock_guard<std::mutex> lk(g_maps_mu);
  auto it = g_maps.find(key);
  if (it != g_maps.end()) { *out = it->second; return 0; }
  cuuint64_t gd[4]; cuuint64_t gs[3]; cuuint32_t bx[4]; cuuint32_t es[4] = {1, 1, 1, 1};
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; }
  for (int i = 0; i < rank - 1; ++i) gs[i] = strides_bytes[i];
  CUresult r = g_encode(out, dtype, (cuuint32_t)rank, const_cast<void*>(ptr), gd, gs, bx, es,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
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
  const int cands[5] = {256, 128, 64, 32, 16};
  if (!g_bn_model) {
    int best = 0;
    for (int i = 0; i < 5; ++i) {
      int bn = cands[i];
      if (Nout % bn) continue;
      if (!best) best = bn;
      if ((long)m_tiles * (Nout / bn) >= g_num_sms || bn <= 64) return bn;
    }
    return best;
  }
  // A persistent CTA works through ceil(tiles / SMs) tiles, each costing its width (MMA and epilogue time are both
  // proportional to BN) plus a fixed part (pipeline fill, first-tile latency: about 32 columns' worth).  The old rule shrank
  // the tile until the grid covered the SMs and so turned 128 tiles of BN = 128 (one wave) into 256 tiles of BN = 64 (two
  // waves, the A tile fetched twice as often): the small launches of a replan tick (1024 rows: 64 row tiles at L = 8) are
  // mostly of that kind.  Wider tiles win ties.
  int best = 0; long best_cost = 0;
  for (int i = 0; i < 5; ++i) {
    const int bn = cands[i];
    if (Nout % bn) continue;
    const long tiles = (long)m_tiles * (Nout / bn);
    const long cost = ((tiles + g_num_sms - 1) / g_num_sms) * (bn + 32);
    if (!best || cost < best_cost) { best = bn; best_cost = cost; }
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
  p.bk = (g_bk32 && d.K <= 32) ? 32 : BK;
  p.ntaps = d.ntaps; p.kblocks = (d.K + p.bk - 1) / p.bk;
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
  const int cw = ((d.epi == EPI_GEGLU_FWD && g_e1_cw == 16) || p.BN == 16) ? 16 : 32;     // epilogue configuration (see epi_warps)
  p.ngroups = p.BN / cw < epi_warps(cw) / 4 ? p.BN / cw : epi_warps(cw) / 4;
  p.nplanes = d.precision == 3 ? 2 : 1;
  // CTA pairs: worth it where the weight tile is large (BN >= 128) and there are enough tile pairs to fill the pairs of SMs
  const int cg = (g_pair && p.BN >= 128 && (long)((p.m_tiles + 1) / 2) * p.n_tiles >= g_num_sms / 2) ? 2 : 1;
  const int stage_bytes = p.nplanes * (BM * p.bk * 2 + (p.BN / cg) * p.bk * 2);
  // TMA store path (see EpiMaps): only where a tile's accumulator rows are consecutive output rows
  p.tma_store = (g_tma_store && d.o_L == d.L_out && d.o_stride == 1 && d.o_off == 0 && BM % d.L_out == 0) ? 1 : 0;
  // staging tiles per epilogue warp: as many as the mode wants (g_stg_want) while at least g_min_stages operand stages fit
  int nbuf = (p.tma_store && p.ntaps * p.kblocks <= g_stg_max_iters[d.epi]) ? g_stg_want[d.epi] : 1;
  while (nbuf > 1 && (g_max_smem - (1024 + nbuf * STG_BYTES + 256)) / stage_bytes < g_min_stages) --nbuf;
  p.stg_bufs = nbuf;
  const int fixed = 1024 + nbuf * STG_BYTES + 256;
  int stages = (g_max_smem - fixed) / stage_bytes;
  const int max_stages = p.bk == 32 ? 10 : 6;
  if (stages > max_stages) stages = max_stages;
  const long total_iters = (long)p.ntaps * p.kblocks * ((long)((p.m_tiles + cg - 1) / cg) * p.n_tiles + g_num_sms / cg - 1) / (g_num_sms / cg);
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
  RAMP_CHECK((d.ldO % 4 == 0) && (d.ldS % 8 == 0) && (d.ldR % 4 == 0) && (d.ldR2 % 4 == 0) && (d.ldS2 % 8 == 0),
             "gemm: output / residual leading dimensions must keep 16-byte alignment");

  CUtensorMap mAhi, mAlo, mBhi, mBlo;
  const CUtensorMapSwizzle op_sw = p.bk == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B;
  {
    uint64_t dims[4] = {(uint64_t)d.K, (uint64_t)d.a_pstride, (uint64_t)(d.L_in / d.a_pstride), (uint64_t)d.n_samples};
    uint64_t str[3] = {(uint64_t)d.ldA * 2, (uint64_t)d.a_pstride * d.ldA * 2, (uint64_t)d.L_in * d.ldA * 2};
    uint32_t box[4] = {(uint32_t)p.bk, 1u, (uint32_t)d.L_out, (uint32_t)p.nb};
    if (make_map(&mAhi, d.a_hi, 4, dims, str, box, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, op_sw)) return -3;
    if (p.nplanes == 2) { if (make_map(&mAlo, d.a_lo, 4, dims, str, box, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, op_sw)) return -3; } else mAlo = mAhi;
  }
  {
    uint64_t dims[2] = {(uint64_t)d.K, (uint64_t)d.n_wtaps * d.Nout};
    uint64_t str[1] = {(uint64_t)d.Kw * 2};
    uint32_t box[2] = {(uint32_t)p.bk, (uint32_t)(p.BN / cg)};      // a CTA of a pair stages half of the tile's weight rows
    if (make_map(&mBhi, d.w_hi, 2, dims, str, box, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, op_sw)) return -3;
    if (p.nplanes == 2) { if (make_map(&mBlo, d.w_lo, 2, dims, str, box, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, op_sw)) return -3; } else mBlo = mBhi;
  }
  // output tensor maps of the TMA store path: [rows = n_samples * L_out][ld] arrays, boxes of 32 rows x cw columns
  EpiMaps om;
  std::memset(&om, 0, sizeof(om));
  if (p.tma_store) {
    const uint64_t rows = (uint64_t)d.n_samples * d.L_out;
    const uint32_t box[2] = {(uint32_t)cw, 32u};
    auto omap = [&](CUtensorMap* m, const void* ptr, int ld, bool f32) -> int {
      if (!ptr) return 0;
      uint64_t dims[2] = {(uint64_t)ld, rows};
      uint64_t str[1] = {(uint64_t)ld * (f32 ? 4 : 2)};
      const CUtensorMapSwizzle sw = f32 ? (cw == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B)
                                        : (cw == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
      return make_map(m, ptr, 2, dims, str, box, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, sw);
    };
    if (omap(&om.f32, d.out_f32, d.ldO, true) || omap(&om.hi, d.out_hi, d.ldS, false) || omap(&om.lo, d.out_lo, d.ldS, false) ||
        omap(&om.hi2, d.out2_hi, d.ldS2, false) || omap(&om.lo2, d.out2_lo, d.ldS2, false)) return -3;
  }
  const size_t smem = (size_t)stages * stage_bytes + fixed;
  const long tiles = (long)((p.m_tiles + cg - 1) / cg) * p.n_tiles * cg;      // CTAs that have work
  const long max_ctas = g_num_sms / cg * cg;
  dim3 grid((unsigned)(tiles < max_ctas ? tiles : max_ctas));
  // RAMP_PDL=1: launch with programmatic stream serialization, so this kernel's prologue overlaps the tail of the previous
  // kernel in the stream (it blocks at griddepcontrol.wait until that kernel has completed and flushed)
  const int use_pdl = pdl_mode() & 1;
  cudaLaunchConfig_t cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attrs[2];
  int na = 0;
  if (use_pdl) {
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (cg == 2) {
    attrs[na].id = cudaLaunchAttributeClusterDimension;
    attrs[na].val.clusterDim.x = 2; attrs[na].val.clusterDim.y = 1; attrs[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attrs; cfg.numAttrs = na;
#define RAMP_GEMM_LAUNCH(E, C)                                                                       \
  do {                                                                                               \
    cfg.blockDim = dim3(gemm_threads(C));                                                            \
    if (cg == 2) RAMP_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<E, C, 2>, mAhi, mAlo, mBhi, mBlo, om, p));  \
    else RAMP_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<E, C, 1>, mAhi, mAlo, mBhi, mBlo, om, p));          \
  } while (0)
  switch (d.epi) {
    case EPI_PLAIN: if (cw == 16) RAMP_GEMM_LAUNCH(EPI_PLAIN, 16); else RAMP_GEMM_LAUNCH(EPI_PLAIN, 32); break;
    case EPI_GEGLU_FWD: if (cw == 16) RAMP_GEMM_LAUNCH(EPI_GEGLU_FWD, 16); else RAMP_GEMM_LAUNCH(EPI_GEGLU_FWD, 32); break;
    case EPI_GEGLU_BWD: RAMP_GEMM_LAUNCH(EPI_GEGLU_BWD, 32); break;
    case EPI_LN_FWD: RAMP_GEMM_LAUNCH(EPI_LN_FWD, 32); break;
    default: RAMP_GEMM_LAUNCH(EPI_LN_BWD, 32); break;
  }
#undef RAMP_GEMM_LAUNCH
  RAMP_LAUNCH_OK();
  return 0;
}

}  // namespace ramp
