// Point-cloud kernels: exact nearest-point query, artificial-potential-field
// pushes, trajectory cost terms and best-of-N selection.
//   nearest point : scipy cKDTree.query(k=1, distance_upper_bound) at APFhelper.py:17-35, APFhelper_dynamic.py:99-105
//   static APF    : APFhelper.py:37-104          dynamic APF : APFhelper_dynamic.py:107-142
//   cost terms    : cost.py:3-7,19-54            selection   : cost.py:56-88
// The reference's "potential field" is a k=1 nearest-neighbour push, so the hot loop is a min/argmin
// over the cloud per waypoint: FP32-pipe bound.  Cloud tiles are staged in shared memory by TMA bulk
// copies (double buffered, mbarrier completion) and re-laid there as point pairs; each thread keeps R waypoints in
// registers and evaluates |p|^2 - 2 q.p for both points of a pair with d packed FFMA2 plus one 3-input min; the argmin is
// recovered by re-scanning, in the direct form (q - p)^2, every 256-point chunk whose minimum key is within the key's
// rounding bound of the best one (normally one to three chunks; the whole cloud by the warp when four or more tie), so
// the index is the fp32-exact nearest point with the lowest index on ties, and the distance is re-derived exactly.
#include "../../include/ramp_b200.h"
#include "common.cuh"
#include <cstdio>
#include <cstdlib>
#include <cfloat>

namespace ramp {

static constexpr int NN_CHUNK = 256;         // argmin recovery granularity (points): the per-chunk bookkeeping is amortised over 256 keys
static constexpr int NN_TILE = 2048;         // cloud padding granularity (points); the kernel's shared-memory tiles divide it
static constexpr int NN_THREADS = 128;
// Per waypoint the scan tracks the three smallest chunk minima with their chunks, plus the VALUE of the fourth.  Only that
// value (the threshold a chunk minimum must beat to matter at all) lives in a register; the other six words per waypoint are
// touched once in a few hundred chunks and sit in shared memory ([word][waypoint slot][thread], conflict-free).  Round 2's
// first version kept all seven in registers: 168 registers at 10 waypoints per thread = 3 CTAs (12 warps) per SM, and ncu
// showed the FP32 pipe waiting on issue (smsp__issue_active 65 %, stall "wait" 30 %).  With the bookkeeping out of the
// register file the hot loop needs ~4 registers per waypoint and 5-6 CTAs fit.
static constexpr int NN_BK_WORDS = 6;        // best, second, third (float) + their chunks (int)

// cloud [P, d] -> packed float4 (x, y, z|0, |p|^2), padded to a multiple of NN_TILE with +inf norms
__global__ void nn_pack_kernel(const float* __restrict__ cloud, int P, int dim, int Ppad, float4* __restrict__ out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < Ppad; i += gridDim.x * blockDim.x) {
    float4 v = make_float4(0.f, 0.f, 0.f, FLT_MAX);
    if (i < P) {
      v.x = cloud[(size_t)i * dim]; v.y = cloud[(size_t)i * dim + 1]; v.z = dim == 3 ? cloud[(size_t)i * dim + 2] : 0.f;
      v.w = v.x * v.x + v.y * v.y + v.z * v.z;
    }
    out[i] = v;
  }
}

__device__ __forceinline__ float min3f(float a, float b, float c) {
  float d;
  asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
// bounded wait (see mbar_wait in gemm_tc.cu): back-off after 2^16 failed tries, trap after 2^26
__device__ __forceinline__ void nn_mbar_wait(uint32_t bar, uint32_t parity) {
#pragma unroll 1
  for (uint32_t spins = 0;; ++spins) {
    uint32_t done;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) return;
    if (spins > (1u << 16)) __nanosleep(32);
    if (spins > (1u << 26)) __trap();
  }
}

// Packed FP32 FMA (sm_100 FFMA2): two IEEE fused multiply-adds per issue slot, each half bit-identical to fmaf.  The scan is
// issue-bound with scalar FFMAs (d FFMA + 1/2 FMNMX3 per pair and one instruction per clock per scheduler: ncu r2 showed
// the FP32 pipe 49 % busy at 65 % issue utilisation), so the cloud tile is re-laid in shared memory as point PAIRS
// ({x0, x1, y0, y1}, {w0, w1} | {z0, z1, w0, w1}) and one FFMA2 produces the partial keys of both points for one waypoint.
__device__ __forceinline__ unsigned long long nn_fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ unsigned long long nn_pack2(float lo, float hi) {
  unsigned long long d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
  return d;
}
__device__ __forceinline__ void nn_unpack2(unsigned long long v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}

// shared memory per CTA: two staged cloud tiles (float4 per point, as packed in HBM) + the current tile as point pairs
// (16 B + 8 B per pair in 2-D, 16 B + 16 B in 3-D) + the bookkeeping words
__host__ __device__ constexpr int nn_pair_bytes(int DIM, int TILE) { return (TILE / 2) * (DIM == 2 ? 24 : 32); }
__host__ __device__ constexpr int nn_smem_bytes(int DIM, int R, int TILE) {
  return 2 * TILE * 16 + nn_pair_bytes(DIM, TILE) + NN_BK_WORDS * R * NN_THREADS * 4;
}
// CTAs per SM the launch bounds ask for: what shared memory allows (228 KB per SM, 1 KB reserved per CTA), at most what
// the hot loop's registers allow (~(DIM + 2) per waypoint + ~36 for the points in flight and addressing)
__host__ __device__ constexpr int nn_min_ctas(int DIM, int R, int TILE) {
  int by_smem = (228 * 1024) / (nn_smem_bytes(DIM, R, TILE) + 1024);
  int regs = (DIM + 2) * R + 48;          // -2 q per component (FFMA2 takes it as a broadcast scalar), chunk minimum, threshold
  regs = (regs + 7) / 8 * 8;
  int by_regs = 65536 / (NN_THREADS * regs);
  int n = by_smem < by_regs ? by_smem : by_regs;
  return n < 1 ? 1 : (n > 8 ? 8 : n);
}

// R = waypoints per thread (register blocking), TILE = cloud points per shared-memory buffer (two buffers per CTA)
template <int DIM, int R, int TILE>
__global__ void __launch_bounds__(NN_THREADS, nn_min_ctas(DIM, R, TILE))
nn_query_kernel(const float4* __restrict__ cloud4, const int* __restrict__ perm, int P, int Ppad,
                const float* __restrict__ q, int Q, int qstride, float thr, float* __restrict__ out_d,
                int* __restrict__ out_idx) {
  extern __shared__ __align__(128) unsigned char nn_smem[];
  float4* tile = reinterpret_cast<float4*>(nn_smem);                       // 2 x TILE, staged by TMA
  ulonglong2* pairA = reinterpret_cast<ulonglong2*>(nn_smem + 2 * TILE * 16);            // per point pair: {x0 x1 | y0 y1}
  unsigned long long* pairW = reinterpret_cast<unsigned long long*>(pairA + TILE / 2);   // 2-D: {w0 w1}
  ulonglong2* pairB = reinterpret_cast<ulonglong2*>(pairA + TILE / 2);                   // 3-D: {z0 z1 | w0 w1}
  float* bkf = reinterpret_cast<float*>(nn_smem + 2 * TILE * 16 + nn_pair_bytes(DIM, TILE));   // [3][R][threads]: best, second, third
  int* bki = reinterpret_cast<int*>(bkf + 3 * R * NN_THREADS);              // [3][R][threads]: their chunks
  __shared__ __align__(8) unsigned long long bars[2];
  const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&bars[0]);
  const uint32_t tile0 = (uint32_t)__cvta_generic_to_shared(tile);
  const int tid = threadIdx.x;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int ntiles = Ppad / TILE;
  auto issue = [&](int t) {
    const uint32_t bar = bar0 + 8 * (t & 1);
    const uint32_t dst = tile0 + (t & 1) * TILE * 16;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(TILE * 16) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(cloud4 + (size_t)t * TILE), "r"(TILE * 16), "r"(bar) : "memory");
  };
  if (tid == 0) issue(0);

  // word w of waypoint slot r of this thread (only this thread ever touches it: no synchronisation needed)
#define NN_BKF(w, r) bkf[((w) * R + (r)) * NN_THREADS + tid]
#define NN_BKI(w, r) bki[((w) * R + (r)) * NN_THREADS + tid]
  unsigned long long mx[R], my[R], mz[R];      // (-2 q, -2 q): both halves of an FFMA2 operand
  float fourth[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int qi = (blockIdx.x * R + r) * NN_THREADS + tid;
    float ax = 0.f, ay = 0.f, az = 0.f;
    if (qi < Q) {
      ax = -2.f * q[(size_t)qi * qstride];
      ay = -2.f * q[(size_t)qi * qstride + 1];
      if (DIM == 3) az = -2.f * q[(size_t)qi * qstride + 2];
    }
    mx[r] = nn_pack2(ax, ax); my[r] = nn_pack2(ay, ay); mz[r] = nn_pack2(az, az);
    fourth[r] = FLT_MAX;
    NN_BKF(0, r) = FLT_MAX; NN_BKF(1, r) = FLT_MAX; NN_BKF(2, r) = FLT_MAX;
    NN_BKI(0, r) = 0; NN_BKI(1, r) = 0; NN_BKI(2, r) = 0;
  }
  for (int t = 0; t < ntiles; ++t) {
    if (tid == 0 && t + 1 < ntiles) issue(t + 1);
    nn_mbar_wait(bar0 + 8 * (t & 1), (t >> 1) & 1);
    const float4* tp = tile + (t & 1) * TILE;
    // staged tile -> point pairs (the previous tile's pairs were last read before the barrier that ended its scan)
    for (int i = tid; i < TILE / 2; i += NN_THREADS) {
      const float4 pa = tp[2 * i], pb = tp[2 * i + 1];
      pairA[i] = make_ulonglong2(nn_pack2(pa.x, pb.x), nn_pack2(pa.y, pb.y));
      if (DIM == 2) pairW[i] = nn_pack2(pa.w, pb.w);
      else pairB[i] = make_ulonglong2(nn_pack2(pa.z, pb.z), nn_pack2(pa.w, pb.w));
    }
    __syncthreads();
#pragma unroll 1
    for (int c = 0; c < TILE / NN_CHUNK; ++c) {
      float cm[R];
#pragma unroll
      for (int r = 0; r < R; ++r) cm[r] = FLT_MAX;
#pragma unroll 8
      for (int j = 0; j < NN_CHUNK / 2; ++j) {
        // keys k = |p|^2 - 2 q.p of both points of the pair, accumulated in the order y, (z,) x - the same three roundings
        // per key as the scalar form the exactness bound below is derived for
        const ulonglong2 a = pairA[c * (NN_CHUNK / 2) + j];
        if (DIM == 2) {
          const unsigned long long w = pairW[c * (NN_CHUNK / 2) + j];
#pragma unroll
          for (int r = 0; r < R; ++r) {
            float k0, k1;
            nn_unpack2(nn_fma2(mx[r], a.x, nn_fma2(my[r], a.y, w)), k0, k1);
            cm[r] = min3f(cm[r], k0, k1);
          }
        } else {
          const ulonglong2 b = pairB[c * (NN_CHUNK / 2) + j];
#pragma unroll
          for (int r = 0; r < R; ++r) {
            float k0, k1;
            nn_unpack2(nn_fma2(mx[r], a.x, nn_fma2(mz[r], b.x, nn_fma2(my[r], a.y, b.y))), k0, k1);
            cm[r] = min3f(cm[r], k0, k1);
          }
        }
      }
      const int gc = t * (TILE / NN_CHUNK) + c;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        if (cm[r] < fourth[r]) {      // rare after the first few chunks: most chunks are nowhere near this waypoint
          const float v = cm[r], b = NN_BKF(0, r), s2 = NN_BKF(1, r), t3 = NN_BKF(2, r);
          if (v < t3) {
            fourth[r] = t3;
            if (v < s2) {
              NN_BKF(2, r) = s2; NN_BKI(2, r) = NN_BKI(1, r);
              if (v < b) { NN_BKF(1, r) = b; NN_BKI(1, r) = NN_BKI(0, r); NN_BKF(0, r) = v; NN_BKI(0, r) = gc; }
              else { NN_BKF(1, r) = v; NN_BKI(1, r) = gc; }
            } else { NN_BKF(2, r) = v; NN_BKI(2, r) = gc; }
          } else {
            fourth[r] = v;
          }
        }
      }
    }
    __syncthreads();
  }
  // Exact decision.  The scan key k = |p|^2 - 2 q.p is three roundings of |p|^2 (pack) and two fused multiply-adds away
  // from d^2 - |q|^2: |error| <= 2^-24 (4 |p|^2 + 2 |q| |p| + |k|), with |p| <= |q| + d.  Every chunk whose minimum key lies
  // within twice that bound (+ margin) of the best key may hold the true nearest point and is re-scanned with the direct
  // form (q - p)^2, lowest index winning ties: the winning chunk, the two runner-up chunks, and - when a FOURTH chunk is
  // that close too (dense clouds: the 1 M-point sweep has a median nearest d^2 of 1e-6, the size of the bound itself) - the
  // whole cloud, by the warp together (the lanes stride over the points, then an argmin reduction).  That fallback reads the
  // whole cloud from L2 per query, so large clouds are packed in MORTON ORDER (ramp_b200/ops.py, `perm` = sorted position ->
  // caller's index): near-tie points are spatial neighbours of each other, hence of the same or the adjacent chunk, and a
  // third tying chunk becomes rare (measured: 2-D sweep shape 131 ms unsorted -> see profiles/r2_*).
  const int lane = tid & 31;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int qi = (blockIdx.x * R + r) * NN_THREADS + tid;
    const bool live = qi < Q;
    const size_t qo = (size_t)(live ? qi : 0) * qstride;
    const float qx = q[qo], qy = q[qo + 1];
    const float qz = DIM == 3 ? q[qo + 2] : 0.f;
    const float q2 = qx * qx + qy * qy + qz * qz;
    const float best = NN_BKF(0, r), second = NN_BKF(1, r), third = NN_BKF(2, r);
    const float qn = sqrtf(q2), pn = qn + sqrtf(fmaxf(best + q2, 0.f));
    const float tol = 1.5e-7f * (4.f * pn * pn + 2.f * qn * pn + fabsf(best)) + 1e-30f;    // 2.5 x 2^-24 x (...)
    float bd2 = FLT_MAX; int gi = P;
    const bool crowded = live && fourth[r] <= best + tol;
    const int ncand = second <= best + tol ? (third <= best + tol ? 3 : 2) : 1;
    const unsigned need = __ballot_sync(0xffffffffu, crowded);
    if (live && !crowded) {
      for (int cnd = 0; cnd < ncand; ++cnd) {
        const int ch = NN_BKI(cnd, r);
        const float4* cp = cloud4 + (size_t)ch * NN_CHUNK;
        for (int j = 0; j < NN_CHUNK; ++j) {
          const float4 p = cp[j];
          if (p.w == FLT_MAX) continue;               // padding
          const float dx = qx - p.x, dy = qy - p.y;
          float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
          if (DIM == 3) { const float dz = qz - p.z; d2 = __fadd_rn(d2, __fmul_rn(dz, dz)); }
          const int g = perm ? perm[ch * NN_CHUNK + j] : ch * NN_CHUNK + j;      // index in the caller's order
          if (d2 < bd2 || (d2 == bd2 && g < gi)) { bd2 = d2; gi = g; }
        }
      }
    }
    for (unsigned m = need; m; m &= m - 1) {         // warp-uniform loop: one crowded query at a time
      const int src = __ffs(m) - 1;
      const float sx = __shfl_sync(0xffffffffu, qx, src), sy = __shfl_sync(0xffffffffu, qy, src);
      const float sz = __shfl_sync(0xffffffffu, qz, src);
      float wd = FLT_MAX; int wi = P;
      for (int g = lane; g < P; g += 32) {
        const float4 p = cloud4[g];
        const float dx = sx - p.x, dy = sy - p.y;
        float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
        if (DIM == 3) { const float dz = sz - p.z; d2 = __fadd_rn(d2, __fmul_rn(dz, dz)); }
        if (d2 <= wd) {                                 // ties: lowest index in the caller's order
          const int og = perm ? perm[g] : g;
          if (d2 < wd || og < wi) { wd = d2; wi = og; }
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float od = __shfl_xor_sync(0xffffffffu, wd, o);
        const int oi = __shfl_xor_sync(0xffffffffu, wi, o);
        if (od < wd || (od == wd && oi < wi)) { wd = od; wi = oi; }
      }
      if (lane == src) { bd2 = wd; gi = wi; }
    }
    if (live) {
      const float d = sqrtf(bd2);
      const bool hit = gi < P && d < thr;
      out_d[qi] = hit ? d : INFINITY;
      out_idx[qi] = hit ? gi : P;
    }
  }
#undef NN_BKF
#undef NN_BKI
}

// ------------------------------------------------------------------------------------------------
// static APF (APFhelper.py:37-104): one CTA per trajectory
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) apf_static_kernel(const float* __restrict__ x, int H, int D, int dim,
                                                         const float* __restrict__ dist, const int* __restrict__ idx,
                                                         const float4* __restrict__ cloud4, int P, float thr,
                                                         float strength, int window, float* __restrict__ out) {
  __shared__ float fx[128], fy[128], fz[128];
  const int b = blockIdx.x, t = threadIdx.x;
  const float* xb = x + (size_t)b * H * D;
  float px = 0.f, py = 0.f, pz = 0.f;
  if (t < H) {
    px = xb[t * D]; py = xb[t * D + 1]; if (dim == 3) pz = xb[t * D + 2];
    const float d = dist[(size_t)b * H + t];
    const int id = idx[(size_t)b * H + t];
    float ax = 0.f, ay = 0.f, az = 0.f;
    if (d < thr && id < P) {
      const float4 p = cloud4[id];
      float dx = px - p.x, dy = py - p.y, dz = dim == 3 ? pz - p.z : 0.f;
      const float n = sqrtf(dx * dx + dy * dy + dz * dz) + 1e-8f;
      const float mag = strength * expf(-d / thr);
      ax = mag * (dx / n); ay = mag * (dy / n); az = mag * (dz / n);
    }
    fx[t] = ax; fy[t] = ay; fz[t] = az;
  }
  __syncthreads();
  if (t < H) {
    float sx = 0.f, sy = 0.f, sz = 0.f;
    const float inv = (window * 0.5f) * (window * 0.5f);
    const int lo = max(0, t - window), hi = min(H - 1, t + window);
    for (int s = lo; s <= hi; ++s) {
      const int k = t - s;
      const float w = expf(-0.5f * (float)(k * k) / inv);
      sx += fx[s] * w; sy += fy[s] * w; sz += fz[s] * w;
    }
    float* ob = out + (size_t)b * H * D + (size_t)t * D;
    ob[0] = px + sx; ob[1] = py + sy;
    if (dim == 3) ob[2] = pz + sz;
    for (int c = dim; c < D; ++c) ob[c] = xb[t * D + c];
  }
}

// ------------------------------------------------------------------------------------------------
// dynamic APF (APFhelper_dynamic.py:107-142): per trajectory, in place on the positions
//   mode 0: push waypoints [c - w, min(H-1, c + w)) around the closest-approach index c
//   mode 1: push waypoints [0, affected) away from the pursuer disc, mixed 0.9 / 0.1 with the goal direction
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) apf_dynamic_kernel(float* __restrict__ x, int H, int D,
                                                          const float* __restrict__ dist, const int* __restrict__ idx,
                                                          const float4* __restrict__ cloud4, int P, int mode, int window,
                                                          int affected, float strength, float thr_force,
                                                          const float* __restrict__ goal,
                                                          const unsigned char* __restrict__ active,
                                                          const float* __restrict__ gate_x, int gate_row,
                                                          const float* __restrict__ gate_pos, float gate_thr) {
  __shared__ float sd[128];
  __shared__ int si[128];
  const int b = blockIdx.x, t = threadIdx.x;
  if (active && !active[b]) return;
  if (gate_x) {   // pursuer gate: |x[b, gate_row, :2] - pursuer| < threshold (diffusion_model_dynamic.py:408-425)
    const float gx = gate_x[((size_t)b * H + gate_row) * D] - gate_pos[0];
    const float gy = gate_x[((size_t)b * H + gate_row) * D + 1] - gate_pos[1];
    if (!(sqrtf(__fadd_rn(__fmul_rn(gx, gx), __fmul_rn(gy, gy))) < gate_thr)) return;
  }
  float* xb = x + (size_t)b * H * D;
  const int n = mode == 1 ? min(affected, H) : H;
  sd[t] = (t < n) ? dist[(size_t)b * H + t] : INFINITY;
  si[t] = t;
  __syncthreads();
  int lo = 0, hi = n;
  if (mode == 0) {
    for (int o = 64; o > 0; o >>= 1) {          // first index of the minimum (np.argmin)
      if (t < o) {
        const float a = sd[t], c = sd[t + o];
        if (c < a || (c == a && si[t + o] < si[t])) { sd[t] = c; si[t] = si[t + o]; }
      }
      __syncthreads();
    }
    const int c = si[0];
    lo = max(0, c - window); hi = min(H - 1, c + window);
  }
  if (t >= lo && t < hi) {
    const int id = idx[(size_t)b * H + t];
    if (id < P) {
      const float d = dist[(size_t)b * H + t];
      const float4 p = cloud4[id];
      const float qx = xb[t * D], qy = xb[t * D + 1];
      float ax = qx - p.x, ay = qy - p.y;
      const float an = sqrtf(ax * ax + ay * ay) + 1e-8f;
      ax /= an; ay /= an;
      if (goal) {
        float gx = goal[0] - qx, gy = goal[1] - qy;
        const float gn = sqrtf(gx * gx + gy * gy) + 1e-8f;
        gx /= gn; gy /= gn;
        ax = 0.9f * ax + 0.1f * gx; ay = 0.9f * ay + 0.1f * gy;
        const float cn = sqrtf(ax * ax + ay * ay) + 1e-8f;
        ax /= cn; ay /= cn;
      }
      const float force = strength * expf(-d / thr_force);
      xb[t * D] = qx + force * ax;
      xb[t * D + 1] = qy + force * ay;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// cost terms (cost.py:3-7,19-54): one warp per trajectory
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) cost_terms_kernel(const float* __restrict__ x, int B, int H, int D, int dim,
                                                         const float* __restrict__ dist, float thr,
                                                         float* __restrict__ out /* [B,3]: free, length, smoothness */) {
  const int b = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (b >= B) return;
  const float* xb = x + (size_t)b * H * D;
  float len = 0.f, sm = 0.f;
  int hit = 0;
  for (int t = lane; t < H; t += 32) {
    if (dist[(size_t)b * H + t] < thr) hit = 1;
    if (t + 1 < H) {
      float a = 0.f, v = 0.f;
      for (int c = 0; c < dim; ++c) { const float d = xb[(t + 1) * D + c] - xb[t * D + c]; a += d * d; }
      for (int c = dim; c < D; ++c) { const float d = xb[(t + 1) * D + c] - xb[t * D + c]; v += d * d; }
      len += sqrtf(a); sm += sqrtf(v);
    }
  }
  len = warp_sum(len); sm = warp_sum(sm);
  hit = __any_sync(0xffffffffu, hit);
  if (lane == 0) { out[b * 3] = hit ? 0.f : 1.f; out[b * 3 + 1] = len; out[b * 3 + 2] = sm; }
}

// selection (cost.py:56-88) over n gathered trajectories: min-max normalise over survivors, weighted sum, argmin
__global__ void __launch_bounds__(1024) select_best_kernel(const float* __restrict__ terms, int n, float w_smooth,
                                                           float w_len, int normalize, float* __restrict__ out /* [8] */,
                                                           float* __restrict__ costs_out) {
  __shared__ float s0[1024], s1[1024], s2[1024], s3[1024];
  __shared__ int si[1024];
  const int t = threadIdx.x;
  float lmin = INFINITY, lmax = -INFINITY, smin = INFINITY, smax = -INFINITY;
  int cnt = 0;
  for (int i = t; i < n; i += 1024)
    if (terms[i * 3] != 0.f) {
      const float l = terms[i * 3 + 1], s = terms[i * 3 + 2];
      lmin = fminf(lmin, l); lmax = fmaxf(lmax, l); smin = fminf(smin, s); smax = fmaxf(smax, s); ++cnt;
    }
  s0[t] = lmin; s1[t] = lmax; s2[t] = smin; s3[t] = smax; si[t] = cnt;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (t < o) {
      s0[t] = fminf(s0[t], s0[t + o]); s1[t] = fmaxf(s1[t], s1[t + o]);
      s2[t] = fminf(s2[t], s2[t + o]); s3[t] = fmaxf(s3[t], s3[t + o]); si[t] += si[t + o];
    }
    __syncthreads();
  }
  lmin = s0[0]; lmax = s1[0]; smin = s2[0]; smax = s3[0];
  const int n_free = si[0];
  __syncthreads();
  // cost = w_s * (s - smin)/(smax - smin) + w_l * (l - lmin)/(lmax - lmin); 0/0 = NaN as in the reference.
  float best = INFINITY; int bi = -1; bool bnan = false;
  for (int i = t; i < n; i += 1024) {
    if (terms[i * 3] != 0.f) {
      const float sv = terms[i * 3 + 2], lv = terms[i * 3 + 1];
      const float c = normalize ? w_smooth * ((sv - smin) / (smax - smin)) + w_len * ((lv - lmin) / (lmax - lmin))
                                : w_smooth * sv + w_len * lv;
      const bool cn = c != c;
      if (bi < 0 || (cn && !bnan) || (!bnan && c < best)) { best = c; bi = i; bnan = cn; }
      if (costs_out) costs_out[i] = c;
    } else if (costs_out) {
      costs_out[i] = INFINITY;
    }
  }
  s0[t] = best; si[t] = bi; s1[t] = bnan ? 1.f : 0.f;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (t < o) {
      const int ia = si[t], ib = si[t + o];
      const bool na = s1[t] != 0.f, nb = s1[t + o] != 0.f;
      bool take = false;
      if (ib >= 0) {
        if (ia < 0) take = true;
        else if (na != nb) take = nb;                       // NaN sorts first (torch.argmin propagates NaN)
        else if (na) take = ib < ia;
        else take = s0[t + o] < s0[t] || (s0[t + o] == s0[t] && ib < ia);
      }
      if (take) { s0[t] = s0[t + o]; si[t] = ib; s1[t] = s1[t + o]; }
    }
    __syncthreads();
  }
  const int best_i = si[0];
  const float best_c = s0[0];
  __syncthreads();
  // rank of the winner among the collision-free trajectories (= cost.py's best_index into the filtered batch)
  int below = 0;
  for (int i = t; i < best_i; i += 1024) below += terms[i * 3] != 0.f ? 1 : 0;
  si[t] = below;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (t < o) si[t] += si[t + o];
    __syncthreads();
  }
  if (t == 0) {
    out[0] = (float)best_i; out[1] = best_c; out[2] = (float)n_free; out[3] = lmin; out[4] = lmax; out[5] = smin; out[6] = smax;
    out[7] = best_i >= 0 ? (float)si[0] : -1.f;
  }
}

// FP32 FMA micro-benchmark: 8 independent FFMA chains per thread, 256-thread CTAs, 8 CTAs per SM.
__global__ void __launch_bounds__(256) fma_probe_kernel(float* __restrict__ sink, int iters, float b, float c) {
  float a[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) a[j] = (float)(threadIdx.x + j) * 1e-3f;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = fmaf(a[j], b, c);
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) s += a[j];
  if (s == 123456.789f) sink[0] = s;      // keeps the chains alive
}

}  // namespace ramp

using namespace ramp;

template <int DIM, int R, int TILE>
static int nn_launch(const float4* c4, const int* perm, int P, int Ppad, const float* q, int Q, int qstride, float thr,
                     float* od, int* oi, cudaStream_t s) {
  static bool attr = false;
  const int smem = nn_smem_bytes(DIM, R, TILE);
  if (!attr) { RAMP_CUDA_OK(cudaFuncSetAttribute(nn_query_kernel<DIM, R, TILE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); attr = true; }
  const int per = NN_THREADS * R;
  nn_query_kernel<DIM, R, TILE><<<(Q + per - 1) / per, NN_THREADS, smem, s>>>(c4, perm, P, Ppad, q, Q, qstride, thr, od, oi);
  RAMP_LAUNCH_OK();
  return 0;
}
// dispatch over the instantiated (R, TILE) pairs
template <int DIM>
static int nn_dispatch(int R, int TILE, const float4* c4, const int* perm, int P, int Ppad, const float* q, int Q, int qstride,
                       float thr, float* od, int* oi, cudaStream_t s) {
#define RAMP_NN_CASE(r) \
  case r: return nn_launch<DIM, r, 512>(c4, perm, P, Ppad, q, Q, qstride, thr, od, oi, s);
  switch (R) {
    RAMP_NN_CASE(1) RAMP_NN_CASE(2) RAMP_NN_CASE(3) RAMP_NN_CASE(4) RAMP_NN_CASE(5) RAMP_NN_CASE(6) RAMP_NN_CASE(7) RAMP_NN_CASE(8)
    RAMP_NN_CASE(9) RAMP_NN_CASE(10) RAMP_NN_CASE(12) RAMP_NN_CASE(14) RAMP_NN_CASE(16)
  }
#undef RAMP_NN_CASE
  set_error("ramp_nn_query: unsupported register blocking");
  return -2;
}
// The grid is one CTA per 128 R waypoints, each scanning the whole cloud, so an SM works through
// ceil(CTAs / SMs) * R blocks of 128 waypoints.  The cost of one block depends little on R since the chunk bookkeeping left
// the register file (more waypoints per thread amortise the shared-memory loads of a point pair over more FFMA2s, fewer
// leave more CTAs resident), so the choice is mostly about filling whole waves: the launch takes the R with the smallest
// ceil(CTAs / SMs) * R * cost[R].  cost[R] = ms per block measured on B200 at 8192 x 64 waypoints x 1 M points
// (profiles/r2_nn_sweep_ffma2.txt, second sweep: all 13 instantiated R).  A model that charged a lone last CTA per SM as two
// (the FP32 pipe needs several warps) was tried and did not predict the measurements - R = 7 (four co-resident CTAs per SM,
// no tail) runs the 2-D sweep shape in 61.6 ms, R = 4 (six resident + a seventh) in 60.9 ms - so it is not in the formula.
template <int DIM>
static int nn_choose(int Q, int* R_out, int* TILE_out) {
  static const int Rs[13] = {1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 14, 16};
  static const double cost2[13] = {3.62, 2.30, 2.05, 2.175, 2.19, 2.165, 2.20, 2.125, 2.15, 2.16, 2.11, 2.25, 2.20};
  static const double cost3[13] = {4.63, 2.81, 2.56, 2.50, 2.49, 2.47, 2.456, 2.47, 2.46, 2.47, 2.48, 2.52, 2.50};
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms < 1) sms = 148;
  }
  *TILE_out = 512;                                    // the only instantiated tile: 16 KB staged + 6-8 KB as pairs per CTA
  if (const char* v = getenv("RAMP_NN_FORCE")) {      // "R" - tuning knob
    int r = 0, t = 0;
    if (sscanf(v, "%d,%d", &r, &t) >= 1) { *R_out = r; return 0; }
  }
  double best = 1e300;
  for (int ri = 0; ri < 13; ++ri) {
    const long ctas = ((long)Q + NN_THREADS * Rs[ri] - 1) / (NN_THREADS * Rs[ri]);
    const double t = (double)((ctas + sms - 1) / sms) * Rs[ri] * (DIM == 2 ? cost2[ri] : cost3[ri]);
    if (t < best) { best = t; *R_out = Rs[ri]; }
  }
  return 0;
}

extern "C" {

int ramp_fp32_fma_probe(float* tflops_host, void* stream) {
  RAMP_CHECK(tflops_host, "ramp_fp32_fma_probe: null argument");
  cudaStream_t s = (cudaStream_t)stream;
  int dev = 0, sms = 148;
  RAMP_CUDA_OK(cudaGetDevice(&dev));
  RAMP_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  static float* sink = nullptr;
  if (!sink) RAMP_CUDA_OK(cudaMalloc(&sink, 256));
  cudaEvent_t e0, e1;
  RAMP_CUDA_OK(cudaEventCreate(&e0));
  RAMP_CUDA_OK(cudaEventCreate(&e1));
  const int iters = 1 << 18, ctas = sms * 8;
  fma_probe_kernel<<<ctas, 256, 0, s>>>(sink, 1 << 10, 0.999f, 1e-3f);          // warm-up
  RAMP_LAUNCH_OK();
  RAMP_CUDA_OK(cudaEventRecord(e0, s));
  fma_probe_kernel<<<ctas, 256, 0, s>>>(sink, iters, 0.999f, 1e-3f);
  RAMP_LAUNCH_OK();
  RAMP_CUDA_OK(cudaEventRecord(e1, s));
  RAMP_CUDA_OK(cudaEventSynchronize(e1));
  float ms = 0.f;
  RAMP_CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *tflops_host = (float)((double)ctas * 256.0 * iters * 8.0 * 2.0 / (ms * 1e-3) / 1e12);
  return 0;
}

size_t ramp_cloud_packed_points(int P) { return (size_t)((P + NN_TILE - 1) / NN_TILE) * NN_TILE; }

int ramp_cloud_pack(const float* cloud, int P, int dim, void* packed, void* stream) {
  RAMP_CHECK(cloud && packed, "ramp_cloud_pack: null argument");
  RAMP_CHECK(dim == 2 || dim == 3, "ramp_cloud_pack: dim must be 2 or 3");
  RAMP_CHECK(P >= 1, "ramp_cloud_pack: empty cloud");
  const int Ppad = (int)ramp_cloud_packed_points(P);
  nn_pack_kernel<<<(Ppad + 255) / 256, 256, 0, (cudaStream_t)stream>>>(cloud, P, dim, Ppad, (float4*)packed);
  RAMP_LAUNCH_OK();
  return 0;
}

int ramp_nn_query(const void* packed, const int32_t* perm, int P, int dim, const float* q, int Q, int qstride,
                  float upper_bound, float* out_dist, int32_t* out_idx, void* stream) {
  RAMP_CHECK(dim == 2 || dim == 3, "ramp_nn_query: dim must be 2 or 3");
  if (Q == 0) return 0;
  RAMP_CHECK(packed && q && out_dist && out_idx, "ramp_nn_query: null argument");
  const int Ppad = (int)ramp_cloud_packed_points(P);
  const float4* c4 = (const float4*)packed;
  cudaStream_t s = (cudaStream_t)stream;
  int R = 1, TILE = 2048;
  if (dim == 2) {
    if (nn_choose<2>(Q, &R, &TILE)) return -1;
    return nn_dispatch<2>(R, TILE, c4, perm, P, Ppad, q, Q, qstride, upper_bound, out_dist, out_idx, s);
  }
  if (nn_choose<3>(Q, &R, &TILE)) return -1;
  return nn_dispatch<3>(R, TILE, c4, perm, P, Ppad, q, Q, qstride, upper_bound, out_dist, out_idx, s);
}

int ramp_apf_static(const float* x, int B, int H, int D, int dim, const float* dist, const int32_t* idx,
                    const void* packed, int P, float thr, float strength, int window, float* out, void* stream) {
  RAMP_CHECK(x && dist && idx && packed && out, "ramp_apf_static: null argument");
  RAMP_CHECK(H <= 128, "ramp_apf_static: horizon must be <= 128");
  if (B == 0) return 0;
  apf_static_kernel<<<B, 128, 0, (cudaStream_t)stream>>>(x, H, D, dim, dist, idx, (const float4*)packed, P, thr, strength,
                                                         window, out);
  RAMP_LAUNCH_OK();
  return 0;
}

int ramp_apf_dynamic(float* x, int B, int H, int D, const float* dist, const int32_t* idx, const void* packed, int P,
                     int mode, int window, int affected, float strength, float thr_force, const float* goal,
                     const unsigned char* active, const float* gate_x, int gate_row, const float* gate_pos,
                     float gate_thr, void* stream) {
  RAMP_CHECK(x && dist && idx && packed, "ramp_apf_dynamic: null argument");
  RAMP_CHECK(H <= 128, "ramp_apf_dynamic: horizon must be <= 128");
  RAMP_CHECK(mode == 0 || mode == 1, "ramp_apf_dynamic: mode must be 0 (static) or 1 (pursuer)");
  RAMP_CHECK(!gate_x || (gate_pos && gate_row >= 0 && gate_row < H), "ramp_apf_dynamic: bad gate");
  if (B == 0) return 0;
  apf_dynamic_kernel<<<B, 128, 0, (cudaStream_t)stream>>>(x, H, D, dist, idx, (const float4*)packed, P, mode, window,
                                                          affected, strength, thr_force, goal, active, gate_x, gate_row,
                                                          gate_pos, gate_thr);
  RAMP_LAUNCH_OK();
  return 0;
}

int ramp_cost_terms(const float* x, int B, int H, int D, int dim, const float* dist, float thr, float* terms,
                    void* stream) {
  RAMP_CHECK(x && dist && terms, "ramp_cost_terms: null argument");
  if (B == 0) return 0;
  cost_terms_kernel<<<(B + 7) / 8, 256, 0, (cudaStream_t)stream>>>(x, B, H, D, dim, dist, thr, terms);
  RAMP_LAUNCH_OK();
  return 0;
}

int ramp_select_best(const float* terms, int n, float w_smooth, float w_len, int normalize, float* out8,
                     float* costs_out, void* stream) {
  RAMP_CHECK(terms && out8, "ramp_select_best: null argument");
  select_best_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(terms, n, w_smooth, w_len, normalize, out8, costs_out);
  RAMP_LAUNCH_OK();
  return 0;
}

}  // extern "C"
