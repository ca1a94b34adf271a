// tcgen05 / TMEM / TMA multi-tap split-precision GEMM for sm_100a.
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
// * Warp roles: warp 0 = TMA producer, warp 1 = TMEM owner + single-thread tcgen05.mma issuer,
//   warps 2..5 = epilogue (tcgen05.ld -> bias / per-scene bias / residual -> fp32 and/or hi/lo stores).
#include "common.cuh"
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>

namespace ramp {

static constexpr int BM = 128;
static constexpr int BK = 64;                 // bf16 elements = 128 B = one SWIZZLE_128B row
static constexpr int A_TILE_BYTES = BM * BK * 2;
static constexpr int GEMM_THREADS = 192;

struct GemmParams {
  int n_samples, L_out, nb;                   // nb = samples per 128-row tile
  int ntaps, kblocks, BN, nplanes, stages;
  int tap_par[5], tap_joff[5], tap_wrow[5];   // tap_wrow = wtap * Nout
  const float* bias; const float* bias_tab; const int* row2scene; int Nout;
  const float* residual; int ldR;
  const float* residual2; int ldR2;
  float* out_f32; int ldO;
  bf16* out_hi; bf16* out_lo; int ldS;
  int o_L, o_stride, o_off;
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

// ---------------------------------------------------------------------------------------------
// kernel
// ---------------------------------------------------------------------------------------------
template <int W>  // W = accumulator columns handled per tcgen05.ld (32, or 16 for BN == 16)
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, const uint32_t* v, int col0, bool row_ok,
                                               size_t out_row, int sample) {
  if (!row_ok) return;
  float r[W];
#pragma unroll
  for (int i = 0; i < W; ++i) r[i] = __uint_as_float(v[i]);
  if (p.bias) {
#pragma unroll
    for (int i = 0; i < W; i += 4) {
      float4 b = *reinterpret_cast<const float4*>(p.bias + col0 + i);
      r[i] += b.x; r[i + 1] += b.y; r[i + 2] += b.z; r[i + 3] += b.w;
    }
  }
  if (p.bias_tab) {
    const float* bt = p.bias_tab + (size_t)p.row2scene[sample] * p.Nout + col0;
#pragma unroll
    for (int i = 0; i < W; i += 4) {
      float4 b = *reinterpret_cast<const float4*>(bt + i);
      r[i] += b.x; r[i + 1] += b.y; r[i + 2] += b.z; r[i + 3] += b.w;
    }
  }
  if (p.residual) {
    const float* rs = p.residual + out_row * p.ldR + col0;
#pragma unroll
    for (int i = 0; i < W; i += 4) {
      float4 b = *reinterpret_cast<const float4*>(rs + i);
      r[i] += b.x; r[i + 1] += b.y; r[i + 2] += b.z; r[i + 3] += b.w;
    }
  }
  if (p.residual2) {
    const float* rs = p.residual2 + out_row * p.ldR2 + col0;
#pragma unroll
    for (int i = 0; i < W; i += 4) {
      float4 b = *reinterpret_cast<const float4*>(rs + i);
      r[i] += b.x; r[i + 1] += b.y; r[i + 2] += b.z; r[i + 3] += b.w;
    }
  }
  if (p.out_f32) {
    float* o = p.out_f32 + out_row * p.ldO + col0;
#pragma unroll
    for (int i = 0; i < W; i += 4) *reinterpret_cast<float4*>(o + i) = make_float4(r[i], r[i + 1], r[i + 2], r[i + 3]);
  }
  if (p.out_hi) {
    bf16* oh = p.out_hi + out_row * p.ldS + col0;
    bf16* ol = p.out_lo + out_row * p.ldS + col0;
#pragma unroll
    for (int i = 0; i < W; i += 8) {
      __align__(16) bf16 h[8];
      __align__(16) bf16 l[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) split_bf16(r[i + j], h[j], l[j]);
      *reinterpret_cast<uint4*>(oh + i) = *reinterpret_cast<const uint4*>(h);
      *reinterpret_cast<uint4*>(ol + i) = *reinterpret_cast<const uint4*>(l);
    }
  }
}

__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap mapAhi, const __grid_constant__ CUtensorMap mapAlo,
               const __grid_constant__ CUtensorMap mapBhi, const __grid_constant__ CUtensorMap mapBlo,
               const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b_tile_bytes = p.BN * BK * 2;
  const int stage_bytes = p.nplanes * (A_TILE_BYTES + b_tile_bytes);
  const uint32_t bar_base = smem_base + p.stages * stage_bytes;   // full[s], empty[s], tmem_full, tmem_ptr
  const uint32_t full0 = bar_base, empty0 = bar_base + 8 * p.stages, tmem_full = bar_base + 16 * p.stages;
  const uint32_t tmem_slot = tmem_full + 8;
  const int iters = p.ntaps * p.kblocks;
  const int tile_m = blockIdx.x, tile_n = blockIdx.y;
  uint32_t tmem_cols = 32;
  while ((int)tmem_cols < p.BN) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
    mbar_init(tmem_full, 1);
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
      for (int t = 0; t < p.ntaps; ++t) {
        for (int kb = 0; kb < p.kblocks; ++kb, ++it) {
          const int s = it % p.stages;
          const uint32_t ph = (it / p.stages) & 1;
          mbar_wait(empty0 + 8 * s, ph ^ 1);
          const uint32_t st = smem_base + s * stage_bytes;
          const uint32_t fb = full0 + 8 * s;
          mbar_expect_tx(fb, (uint32_t)(p.nplanes * (p.nb * p.L_out * BK * 2 + b_tile_bytes)));
          const int c3 = tile_m * p.nb;
          tma_load_4d(st, &mapAhi, fb, kb * BK, p.tap_par[t], p.tap_joff[t], c3);
          uint32_t off = A_TILE_BYTES;
          if (p.nplanes == 2) { tma_load_4d(st + off, &mapAlo, fb, kb * BK, p.tap_par[t], p.tap_joff[t], c3); off += A_TILE_BYTES; }
          tma_load_2d(st + off, &mapBhi, fb, kb * BK, p.tap_wrow[t] + tile_n * p.BN);
          off += b_tile_bytes;
          if (p.nplanes == 2) tma_load_2d(st + off, &mapBlo, fb, kb * BK, p.tap_wrow[t] + tile_n * p.BN);
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer ----------------
    if (lane == 0) {
      // instruction descriptor: D=f32, A=B=bf16, both K-major, N = BN, M = 128
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      for (int it = 0; it < iters; ++it) {
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
          const uint32_t first = (it == 0 && k == 0) ? 0u : 1u;
          if (p.nplanes == 2) {
            umma_bf16(tmem_acc, umma_desc_sw128(a_lo + ko), umma_desc_sw128(b_hi + ko), idesc, first);
            umma_bf16(tmem_acc, umma_desc_sw128(a_hi + ko), umma_desc_sw128(b_lo + ko), idesc, 1u);
            umma_bf16(tmem_acc, umma_desc_sw128(a_hi + ko), umma_desc_sw128(b_hi + ko), idesc, 1u);
          } else {
            umma_bf16(tmem_acc, umma_desc_sw128(a_hi + ko), umma_desc_sw128(b_hi + ko), idesc, first);
          }
        }
        umma_commit(empty0 + 8 * s);           // frees the smem stage when these MMAs retire
      }
      umma_commit(tmem_full);                   // accumulator complete
    }
  } else {
    // ---------------- epilogue (warps 2..5 -> TMEM lane quarters 2,3,0,1) ----------------
    const int q = warp & 3;
    mbar_wait(tmem_full, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int r = q * 32 + lane;
    const int n_local = r / p.L_out, l = r - n_local * p.L_out;
    const int sample = tile_m * p.nb + n_local;
    const bool row_ok = n_local < p.nb && sample < p.n_samples;
    const size_t out_row = (size_t)sample * p.o_L + (size_t)l * p.o_stride + p.o_off;
    const uint32_t taddr = tmem_acc + ((uint32_t)(q * 32) << 16);
    const int colbase = tile_n * p.BN;
    if (p.BN >= 32) {
      for (int c = 0; c < p.BN; c += 32) {
        uint32_t v[32];
        tmem_ld32(taddr + c, v);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        epilogue_chunk<32>(p, v, colbase + c, row_ok, out_row, sample);
      }
    } else {
      uint32_t v[16];
      tmem_ld16(taddr, v);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      epilogue_chunk<16>(p, v, colbase, row_ok, out_row, sample);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
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

int gemm_init() {
  if (g_encode) return 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  RAMP_CUDA_OK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  RAMP_CHECK(fn != nullptr && qres == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled not available");
  int dev = 0;
  RAMP_CUDA_OK(cudaGetDevice(&dev));
  RAMP_CUDA_OK(cudaDeviceGetAttribute(&g_max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  RAMP_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, g_max_smem));
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
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
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
  // Largest tile that divides Nout; prefer >= 2 waves of 148 CTAs when the problem allows it.
  const int cands[5] = {256, 128, 64, 32, 16};
  int best = 0;
  for (int i = 0; i < 5; ++i) {
    int bn = cands[i];
    if (Nout % bn) continue;
    if (!best) best = bn;
    if ((long)m_tiles * (Nout / bn) >= 2 * 148 || bn <= 64) return bn;
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
  GemmParams p;
  std::memset(&p, 0, sizeof(p));
  p.n_samples = d.n_samples; p.L_out = d.L_out; p.nb = BM / d.L_out;
  p.ntaps = d.ntaps; p.kblocks = (d.K + BK - 1) / BK;
  const int m_tiles = (d.n_samples + p.nb - 1) / p.nb;
  p.BN = pick_bn(d.Nout, m_tiles);
  p.nplanes = d.precision == 3 ? 2 : 1;
  const int stage_bytes = p.nplanes * (A_TILE_BYTES + p.BN * BK * 2);
  int stages = (g_max_smem - 2048) / stage_bytes;
  if (stages > 4) stages = 4;
  if (stages > p.ntaps * p.kblocks) stages = p.ntaps * p.kblocks;
  RAMP_CHECK(stages >= 1, "gemm: tile does not fit shared memory");
  p.stages = stages;
  for (int t = 0; t < d.ntaps; ++t) {
    p.tap_par[t] = d.taps[t].par; p.tap_joff[t] = d.taps[t].joff; p.tap_wrow[t] = d.taps[t].wtap * d.Nout;
  }
  p.bias = d.bias; p.bias_tab = d.bias_tab; p.row2scene = d.row2scene; p.Nout = d.Nout;
  p.residual = d.residual; p.ldR = d.ldR; p.residual2 = d.residual2; p.ldR2 = d.ldR2; p.out_f32 = d.out_f32; p.ldO = d.ldO;
  p.out_hi = d.out_hi; p.out_lo = d.out_lo; p.ldS = d.ldS;
  p.o_L = d.o_L; p.o_stride = d.o_stride; p.o_off = d.o_off;

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
  const size_t smem = (size_t)stages * stage_bytes + 1024 + 256;
  dim3 grid(m_tiles, d.Nout / p.BN);
  gemm_tc_kernel<<<grid, GEMM_THREADS, smem, stream>>>(mAhi, mAlo, mBhi, mBlo, p);
  RAMP_LAUNCH_OK();
  return 0;
}

}  // namespace ramp
