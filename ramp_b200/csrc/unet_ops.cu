// See unet_ops.cuh.  Reference sites: mpd/models/layers/layers.py:225-297,327-361 and
// mpd/models/layers/layers_attention_mini.py:38-65,83-149 (forward math); backward formulas are the
// analytic input-gradients that torch.autograd.grad evaluates at UnetInference.py:23-27.
#include "unet_ops.cuh"

namespace ramp {

static __device__ __forceinline__ void store_split(bf16* hi, bf16* lo, size_t idx, float v) {
  bf16 h, l;
  split_bf16(v, h, l);
  hi[idx] = h;
  lo[idx] = l;
}

// ------------------------------------------------------------------------------------------------
__global__ void prep_input_kernel(const float* __restrict__ x, int n_samples, int n_pad, int L, int D, int rep, bf16* hi,
                                  bf16* lo) {
  const size_t total = (size_t)n_pad * L * 16;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i & 15);
    const size_t row = i >> 4;
    const size_t n = row / L;
    const size_t l = row - n * L;
    float v = (c < D && n < (size_t)n_samples) ? x[((n / rep) * L + l) * D + c] : 0.f;
    store_split(hi, lo, i, v);
  }
}
int prep_input(const float* x, int n_samples, int n_pad, int L, int D, int x_repeat, bf16* hi, bf16* lo, cudaStream_t s) {
  const size_t total = (size_t)n_pad * L * 16;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  prep_input_kernel<<<blocks, 256, 0, s>>>(x, n_samples, n_pad, L, D, x_repeat < 1 ? 1 : x_repeat, hi, lo);
  RAMP_LAUNCH_OK();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// GroupNorm(8 groups) over one sample's [L, C] tile (L*C <= 2048 floats): one 256-thread CTA per sample.
// The tile is read ONCE with coalesced float4 loads and stays in registers (<= 2 float4 per thread); group
// statistics are reduced by segmented warp shuffles (the lanes of one group are adjacent) + a fixed-order pass.
// Mish(x) = x * tanh(softplus(x)) is evaluated as x * n / (n + 2), n = e^x (e^x + 2)  (exact identity).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float mish_fast(float x) {
  if (x > 20.f) return x;
  const float w = __expf(x), n = w * (w + 2.f);
  return x * __fdividef(n, n + 2.f);
}
__device__ __forceinline__ float mish_grad_fast(float x) {
  if (x > 20.f) return 1.f;
  const float w = __expf(x), n = w * (w + 2.f), d = __fdividef(1.f, n + 2.f);
  return n * d + x * (4.f * w * (w + 1.f)) * d * d;
}
// sum of v over the lanes that share a group: groups are `seg` adjacent lanes (seg = 1, 2, 4 or 8)
__device__ __forceinline__ float seg_sum(float v, int seg) {
  for (int o = 1; o < seg; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
struct GnTile {           // the (<= 2) float4 of the sample tile owned by a thread
  float4 v[2]; int row[2], col[2]; bool ok[2];
};
__device__ __forceinline__ void gn_load(GnTile& t, const float* __restrict__ x, int ldX, size_t row0, int L, int C) {
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int idx = (threadIdx.x + i * 256) * 4;
    t.ok[i] = idx < L * C;
    t.row[i] = t.ok[i] ? idx / C : 0;
    t.col[i] = t.ok[i] ? idx - t.row[i] * C : 0;
    t.v[i] = t.ok[i] ? *reinterpret_cast<const float4*>(x + (row0 + t.row[i]) * ldX + t.col[i]) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}
// red[g] = sum over group g of (a[0], a[1]) (one value per owned float4), in a FIXED order (no atomics, so the
// sampler is bit-reproducible run to run): segment sums are parked in `part` (one slot per float4 of the tile), then
// warp g adds the <= 64 segment leaders of group g lane-strided and finishes with a butterfly.  Ends with a CTA barrier.
__device__ __forceinline__ void gn_group_reduce(const GnTile& t, const float (&a)[2], int L, int C, float* part, float* red) {
  const int cg = C / 8, seg = cg >= 32 ? 8 : cg / 4, nl = cg >= 32 ? cg / 32 : 1;
#pragma unroll
  for (int i = 0; i < 2; ++i) part[threadIdx.x + i * 256] = seg_sum(a[i], seg);
  __syncthreads();
  const int g = threadIdx.x >> 5, lane = threadIdx.x & 31, n = L * nl;
  float s = 0.f;
  for (int k = lane; k < n; k += 32) {
    const int r = k / nl, j = k - r * nl;
    s += part[(r * C + g * cg) / 4 + j * seg];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) red[g] = s;
  __syncthreads();
}
// two quantities in one pass (part holds 2 x 512 slots)
__device__ __forceinline__ void gn_group_reduce2(const GnTile& t, const float (&a)[2], const float (&b)[2], int L, int C,
                                                 float* part, float* redA, float* redB) {
  const int cg = C / 8, seg = cg >= 32 ? 8 : cg / 4, nl = cg >= 32 ? cg / 32 : 1;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    part[threadIdx.x + i * 256] = seg_sum(a[i], seg);
    part[512 + threadIdx.x + i * 256] = seg_sum(b[i], seg);
  }
  __syncthreads();
  const int g = threadIdx.x >> 5, lane = threadIdx.x & 31, n = L * nl;
  float s = 0.f, u = 0.f;
  for (int k = lane; k < n; k += 32) {
    const int r = k / nl, j = k - r * nl, slot = (r * C + g * cg) / 4 + j * seg;
    s += part[slot]; u += part[512 + slot];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); u += __shfl_xor_sync(0xffffffffu, u, o); }
  if (lane == 0) { redA[g] = s; redB[g] = u; }
  __syncthreads();
}
__device__ __forceinline__ void gn_store(float4 y, size_t row, int c, float* out_f32, int ldO, bf16* hi, bf16* lo, int ldS) {
  if (out_f32) *reinterpret_cast<float4*>(out_f32 + row * ldO + c) = y;
  if (hi) {
    uint32_t h0, l0, h1, l1;
    split2_bf16(y.x, y.y, h0, l0);
    split2_bf16(y.z, y.w, h1, l1);
    *reinterpret_cast<uint2*>(hi + row * ldS + c) = make_uint2(h0, h1);
    *reinterpret_cast<uint2*>(lo + row * ldS + c) = make_uint2(l0, l1);
  }
}
// mean / rstd of every group into stat[0..7] / stat[8..15]
__device__ __forceinline__ void gn_stats(const GnTile& t, int L, int C, float eps, float* part, float* red, float* stat) {
  const int cg = C / 8;
  const float inv_m = 1.f / (float)(L * cg);
  float a[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) a[i] = (t.v[i].x + t.v[i].y) + (t.v[i].z + t.v[i].w);
  gn_group_reduce(t, a, L, C, part, red);
  if (threadIdx.x < 8) stat[threadIdx.x] = red[threadIdx.x] * inv_m;          // read after the next reduction's barriers
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const float m = red[t.col[i] / cg] * inv_m;
    const float dx = t.v[i].x - m, dy = t.v[i].y - m, dz = t.v[i].z - m, dw = t.v[i].w - m;
    a[i] = t.ok[i] ? (dx * dx + dy * dy) + (dz * dz + dw * dw) : 0.f;
  }
  gn_group_reduce(t, a, L, C, part, red + 8);
  if (threadIdx.x < 8) stat[8 + threadIdx.x] = rsqrtf(red[8 + threadIdx.x] * inv_m + eps);
  __syncthreads();
}

// CT = the channel count when it is one of the U-Net's (32 / 64 / 128 / 256: every index computation folds to shifts;
// the generic CT = 0 instantiation divides by runtime values and spends most of its issue slots on that)
template <int CT>
__global__ void __launch_bounds__(256) gn_forward_kernel(GnArgs a) {
  __shared__ float red[16], stat[16], part[512];
  pdl_wait();
  const size_t row0 = (size_t)blockIdx.x * a.L;
  const int C = CT ? CT : a.C;
  const int cg = C / 8;
  GnTile t;
  gn_load(t, a.x, a.ldX, row0, a.L, C);
  gn_stats(t, a.L, C, a.eps, part, red, stat);
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    if (!t.ok[i]) continue;
    const int c = t.col[i], g = c / cg;
    const size_t row = row0 + t.row[i];
    const float mean = stat[g], rstd = stat[8 + g];
    const float4 gm = *reinterpret_cast<const float4*>(a.gamma + c), bt = *reinterpret_cast<const float4*>(a.beta + c);
    float y[4] = {(t.v[i].x - mean) * rstd * gm.x + bt.x, (t.v[i].y - mean) * rstd * gm.y + bt.y,
                  (t.v[i].z - mean) * rstd * gm.z + bt.z, (t.v[i].w - mean) * rstd * gm.w + bt.w};
    if (a.mish)
#pragma unroll
      for (int j = 0; j < 4; ++j) y[j] = mish_fast(y[j]);
    if (a.cond) {
      const float4 cd = *reinterpret_cast<const float4*>(a.cond + c);
      y[0] += cd.x; y[1] += cd.y; y[2] += cd.z; y[3] += cd.w;
    }
    if (a.add) {
      const float4 ad = *reinterpret_cast<const float4*>(a.add + row * a.ldAdd + c);
      y[0] += ad.x; y[1] += ad.y; y[2] += ad.z; y[3] += ad.w;
    }
    gn_store(make_float4(y[0], y[1], y[2], y[3]), row, c, a.out_f32, a.ldO, a.out_hi, a.out_lo, a.ldS);
  }
}
int gn_forward(const GnArgs& a, cudaStream_t s) {
  RAMP_CHECK(a.C % 32 == 0 && a.L * a.C <= 2048, "gn: C must be a multiple of 32 and L*C <= 2048");
  RAMP_CHECK(a.ldX % 4 == 0 && (!a.add || a.ldAdd % 4 == 0) && (!a.out_f32 || a.ldO % 4 == 0) && (!a.out_hi || a.ldS % 4 == 0),
             "gn: leading dimensions must be multiples of 4");
  switch (a.C) {
    case 32: RAMP_CUDA_OK(launch_pdl(2, gn_forward_kernel<32>, dim3(a.n_samples), dim3(256), 0, s, a)); break;
    case 64: RAMP_CUDA_OK(launch_pdl(2, gn_forward_kernel<64>, dim3(a.n_samples), dim3(256), 0, s, a)); break;
    case 128: RAMP_CUDA_OK(launch_pdl(2, gn_forward_kernel<128>, dim3(a.n_samples), dim3(256), 0, s, a)); break;
    case 256: RAMP_CUDA_OK(launch_pdl(2, gn_forward_kernel<256>, dim3(a.n_samples), dim3(256), 0, s, a)); break;
    default: RAMP_CUDA_OK(launch_pdl(2, gn_forward_kernel<0>, dim3(a.n_samples), dim3(256), 0, s, a)); break;
  }
  RAMP_LAUNCH_OK();
  return 0;
}

template <int CT>
__global__ void __launch_bounds__(256) gn_backward_kernel(GnBwdArgs a) {
  __shared__ float red[16], stat[16], red2[16], part[1024];
  pdl_wait();
  const size_t row0 = (size_t)blockIdx.x * a.L;
  const int C = CT ? CT : a.C;
  const int cg = C / 8;
  const float inv_m = 1.f / (float)(a.L * cg);
  GnTile t;
  gn_load(t, a.x, a.ldX, row0, a.L, C);
  gn_stats(t, a.L, C, a.eps, part, red, stat);
  float xh[2][4], dyh[2][4], p1[2], p2[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int c = t.col[i], g = c / cg;
    const float mean = stat[g], rstd = stat[8 + g];
    const float4 gm = *reinterpret_cast<const float4*>(a.gamma + c), bt = *reinterpret_cast<const float4*>(a.beta + c);
    float4 dy = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t.ok[i]) dy = *reinterpret_cast<const float4*>(a.dy + (row0 + t.row[i]) * a.ldDy + c);
    const float xv[4] = {t.v[i].x, t.v[i].y, t.v[i].z, t.v[i].w}, dv[4] = {dy.x, dy.y, dy.z, dy.w};
    const float gv[4] = {gm.x, gm.y, gm.z, gm.w}, bv[4] = {bt.x, bt.y, bt.z, bt.w};
    p1[i] = 0.f; p2[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      xh[i][j] = (xv[j] - mean) * rstd;
      float d = dv[j] * gv[j];
      if (a.mish) d *= mish_grad_fast(xh[i][j] * gv[j] + bv[j]);
      dyh[i][j] = t.ok[i] ? d : 0.f;
      p1[i] += dyh[i][j];
      p2[i] += dyh[i][j] * xh[i][j];
    }
  }
  gn_group_reduce2(t, p1, p2, a.L, C, part, red2, red2 + 8);
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    if (!t.ok[i]) continue;
    const int c = t.col[i], g = c / cg;
    const size_t row = row0 + t.row[i];
    const float rstd = stat[8 + g], s1 = red2[g] * inv_m, s2 = red2[8 + g] * inv_m;
    float dx[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) dx[j] = rstd * (dyh[i][j] - s1 - xh[i][j] * s2);
    if (a.add) {
      const float4 ad = *reinterpret_cast<const float4*>(a.add + row * a.ldAdd + c);
      dx[0] += ad.x; dx[1] += ad.y; dx[2] += ad.z; dx[3] += ad.w;
    }
    gn_store(make_float4(dx[0], dx[1], dx[2], dx[3]), row, c, a.out_f32, a.ldO, a.out_hi, a.out_lo, a.ldS);
  }
}
int gn_backward(const GnBwdArgs& a, cudaStream_t s) {
  RAMP_CHECK(a.C % 32 == 0 && a.L * a.C <= 2048, "gn: C must be a multiple of 32 and L*C <= 2048");
  RAMP_CHECK(a.ldX % 4 == 0 && a.ldDy % 4 == 0 && (!a.add || a.ldAdd % 4 == 0) && (!a.out_f32 || a.ldO % 4 == 0) &&
             (!a.out_hi || a.ldS % 4 == 0), "gn: leading dimensions must be multiples of 4");
  switch (a.C) {
    case 32: RAMP_CUDA_OK(launch_pdl(2, gn_backward_kernel<32>, dim3(a.n_samples), dim3(256), 0, s, a)); break;
    case 64: RAMP_CUDA_OK(launch_pdl(2, gn_backward_kernel<64>, dim3(a.n_samples), dim3(256), 0, s, a)); break;
    case 128: RAMP_CUDA_OK(launch_pdl(2, gn_backward_kernel<128>, dim3(a.n_samples), dim3(256), 0, s, a)); break;
    case 256: RAMP_CUDA_OK(launch_pdl(2, gn_backward_kernel<256>, dim3(a.n_samples), dim3(256), 0, s, a)); break;
    default: RAMP_CUDA_OK(launch_pdl(2, gn_backward_kernel<0>, dim3(a.n_samples), dim3(256), 0, s, a)); break;
  }
  RAMP_LAUNCH_OK();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// LayerNorm over 256 channels: one warp per token, 8 channels per lane (two float4).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ln_forward_kernel(const float* __restrict__ x, int rows,
                                                         const float* __restrict__ gamma,
                                                         const float* __restrict__ beta, bf16* hi, bf16* lo) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * 256);
  float4 a = xr[lane], b = xr[32 + lane];
  float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += v[i];
  const float mean = warp_sum(s) * (1.f / 256.f);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) { v[i] -= mean; q += v[i] * v[i]; }
  const float rstd = rsqrtf(warp_sum(q) * (1.f / 256.f) + 1e-5f);
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int c0 = h * 128 + lane * 4;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = c0 + i;
      store_split(hi, lo, (size_t)row * 256 + c, v[h * 4 + i] * rstd * gamma[c] + beta[c]);
    }
  }
}
int ln_forward(const float* x, int rows, int C, const float* gamma, const float* beta, bf16* out_hi, bf16* out_lo,
               cudaStream_t s) {
  RAMP_CHECK(C == 256, "ln: C must be 256");
  ln_forward_kernel<<<(rows + 7) / 8, 256, 0, s>>>(x, rows, gamma, beta, out_hi, out_lo);
  RAMP_LAUNCH_OK();
  return 0;
}

__global__ void __launch_bounds__(256) ln_backward_kernel(const float* __restrict__ x, int rows,
                                                          const float* __restrict__ gamma,
                                                          const float* __restrict__ dy, const float* __restrict__ add,
                                                          float* out, bf16* hi, bf16* lo) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * 256);
  const float4* dr = reinterpret_cast<const float4*>(dy + (size_t)row * 256);
  float4 a = xr[lane], b = xr[32 + lane], da = dr[lane], db = dr[32 + lane];
  float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  float d[8] = {da.x, da.y, da.z, da.w, db.x, db.y, db.z, db.w};
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += v[i];
  const float mean = warp_sum(s) * (1.f / 256.f);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) { v[i] -= mean; q += v[i] * v[i]; }
  const float rstd = rsqrtf(warp_sum(q) * (1.f / 256.f) + 1e-5f);
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int h = 0; h < 2; ++h)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = h * 4 + i;
      v[k] *= rstd;                               // xhat
      d[k] *= gamma[h * 128 + lane * 4 + i];      // dyhat
      s1 += d[k];
      s2 += d[k] * v[k];
    }
  s1 = warp_sum(s1) * (1.f / 256.f);
  s2 = warp_sum(s2) * (1.f / 256.f);
#pragma unroll
  for (int h = 0; h < 2; ++h)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = h * 4 + i;
      const size_t idx = (size_t)row * 256 + h * 128 + lane * 4 + i;
      float dx = rstd * (d[k] - s1 - v[k] * s2);
      if (add) dx += add[idx];
      if (out) out[idx] = dx;
      if (hi) store_split(hi, lo, idx, dx);
    }
}
int ln_backward(const float* x, int rows, int C, const float* gamma, const float* dy, const float* add, float* out_f32,
                bf16* out_hi, bf16* out_lo, cudaStream_t s) {
  RAMP_CHECK(C == 256, "ln: C must be 256");
  ln_backward_kernel<<<(rows + 7) / 8, 256, 0, s>>>(x, rows, gamma, dy, add, out_f32, out_hi, out_lo);
  RAMP_LAUNCH_OK();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Self-attention core, 4 heads x 64, L <= 64 tokens: one CTA per (sample, head), operands in smem.
// ------------------------------------------------------------------------------------------------
static constexpr int AH = 64;       // head dim
static constexpr int AP = AH + 1;   // padded smem row

__device__ __forceinline__ void attn_load_head(const float* __restrict__ src, int ld, int L, float* dst) {
  for (int i = threadIdx.x; i < L * AH; i += blockDim.x) dst[(i / AH) * AP + (i % AH)] = src[(size_t)(i / AH) * ld + (i % AH)];
}
// S = softmax(scale * Q K^T) into P [L][L+1]
__device__ __forceinline__ void attn_probs(const float* Q, const float* K, float* P, int L) {
  const int LP = L + 1;
  for (int idx = threadIdx.x; idx < L * L; idx += blockDim.x) {
    const int i = idx / L, j = idx - i * L;
    float acc = 0.f;
#pragma unroll 16
    for (int d = 0; d < AH; ++d) acc = fmaf(Q[i * AP + d], K[j * AP + d], acc);
    P[i * LP + j] = acc * 0.125f;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int i = warp; i < L; i += nw) {
    float m = -INFINITY;
    for (int j = lane; j < L; j += 32) m = fmaxf(m, P[i * LP + j]);
    m = warp_max(m);
    float s = 0.f;
    for (int j = lane; j < L; j += 32) { float e = expf(P[i * LP + j] - m); P[i * LP + j] = e; s += e; }
    s = 1.f / warp_sum(s);
    for (int j = lane; j < L; j += 32) P[i * LP + j] *= s;
  }
  __syncthreads();
}

__global__ void __launch_bounds__(128) attn_forward_kernel(const float* __restrict__ qkv, int L, bf16* hi, bf16* lo) {
  extern __shared__ float sm[];
  const int n = blockIdx.x >> 2, h = blockIdx.x & 3;
  float* Q = sm; float* K = Q + L * AP; float* V = K + L * AP; float* P = V + L * AP;
  const float* base = qkv + (size_t)n * L * 768 + h * AH;
  attn_load_head(base, 768, L, Q);
  attn_load_head(base + 256, 768, L, K);
  attn_load_head(base + 512, 768, L, V);
  __syncthreads();
  attn_probs(Q, K, P, L);
  const int LP = L + 1;
  for (int idx = threadIdx.x; idx < L * AH; idx += blockDim.x) {
    const int i = idx / AH, d = idx - i * AH;
    float acc = 0.f;
    for (int j = 0; j < L; ++j) acc = fmaf(P[i * LP + j], V[j * AP + d], acc);
    store_split(hi, lo, ((size_t)n * L + i) * 256 + h * AH + d, acc);
  }
}
static size_t attn_smem(int L, bool bwd) { return sizeof(float) * ((bwd ? 4 : 3) * L * AP + (bwd ? 2 : 1) * L * (L + 1)); }
int attn_forward_generic(const float* qkv, int n_samples, int L, bf16* out_hi, bf16* out_lo, cudaStream_t s) {
  RAMP_CHECK(L <= 64, "attention: L must be <= 64");
  static bool attr = false;
  if (!attr) { RAMP_CUDA_OK(cudaFuncSetAttribute(attn_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_smem(64, false))); attr = true; }
  attn_forward_kernel<<<n_samples * 4, 128, attn_smem(L, false), s>>>(qkv, L, out_hi, out_lo);
  RAMP_LAUNCH_OK();
  return 0;
}

__global__ void __launch_bounds__(128) attn_backward_kernel(const float* __restrict__ qkv,
                                                            const float* __restrict__ d_out, int L, bf16* hi, bf16* lo) {
  extern __shared__ float sm[];
  const int n = blockIdx.x >> 2, h = blockIdx.x & 3;
  const int LP = L + 1;
  float* Q = sm; float* K = Q + L * AP; float* V = K + L * AP; float* dO = V + L * AP;
  float* P = dO + L * AP; float* dS = P + L * LP;
  const float* base = qkv + (size_t)n * L * 768 + h * AH;
  attn_load_head(base, 768, L, Q);
  attn_load_head(base + 256, 768, L, K);
  attn_load_head(base + 512, 768, L, V);
  attn_load_head(d_out + (size_t)n * L * 256 + h * AH, 256, L, dO);
  __syncthreads();
  attn_probs(Q, K, P, L);
  // dP = dO V^T
  for (int idx = threadIdx.x; idx < L * L; idx += blockDim.x) {
    const int i = idx / L, j = idx - i * L;
    float acc = 0.f;
#pragma unroll 16
    for (int d = 0; d < AH; ++d) acc = fmaf(dO[i * AP + d], V[j * AP + d], acc);
    dS[i * LP + j] = acc;
  }
  __syncthreads();
  // dS = P * (dP - rowsum(P * dP)) * scale
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int i = warp; i < L; i += nw) {
    float s = 0.f;
    for (int j = lane; j < L; j += 32) s += P[i * LP + j] * dS[i * LP + j];
    s = warp_sum(s);
    for (int j = lane; j < L; j += 32) dS[i * LP + j] = P[i * LP + j] * (dS[i * LP + j] - s) * 0.125f;
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < L * AH; idx += blockDim.x) {
    const int i = idx / AH, d = idx - i * AH;   // i doubles as the key index j for dK / dV
    float dq = 0.f, dk = 0.f, dv = 0.f;
    for (int j = 0; j < L; ++j) {
      dq = fmaf(dS[i * LP + j], K[j * AP + d], dq);
      dk = fmaf(dS[j * LP + i], Q[j * AP + d], dk);
      dv = fmaf(P[j * LP + i], dO[j * AP + d], dv);
    }
    const size_t o = ((size_t)n * L + i) * 768 + h * AH + d;
    store_split(hi, lo, o, dq);
    store_split(hi, lo, o + 256, dk);
    store_split(hi, lo, o + 512, dv);
  }
}
int attn_backward_generic(const float* qkv, const float* d_out, int n_samples, int L, bf16* dq_hi, bf16* dq_lo, cudaStream_t s) {
  RAMP_CHECK(L <= 64, "attention: L must be <= 64");
  static bool attr = false;
  if (!attr) { RAMP_CUDA_OK(cudaFuncSetAttribute(attn_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_smem(64, true))); attr = true; }
  attn_backward_kernel<<<n_samples * 4, 128, attn_smem(L, true), s>>>(qkv, d_out, L, dq_hi, dq_lo);
  RAMP_LAUNCH_OK();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// GEGLU (layers_attention_mini.py:38-45)
// ------------------------------------------------------------------------------------------------
__global__ void geglu_forward_kernel(const float* __restrict__ u, size_t rows, int inner, bf16* hi, bf16* lo) {
  const size_t total = rows * inner;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / inner; const int c = (int)(i - r * inner);
    const float a = u[r * 2 * inner + c], g = u[r * 2 * inner + inner + c];
    store_split(hi, lo, i, a * gelu_f(g));
  }
}
__global__ void geglu_backward_kernel(const float* __restrict__ u, const float* __restrict__ dg, size_t rows, int inner,
                                      bf16* hi, bf16* lo) {
  const size_t total = rows * inner;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / inner; const int c = (int)(i - r * inner);
    const float a = u[r * 2 * inner + c], g = u[r * 2 * inner + inner + c], d = dg[i];
    store_split(hi, lo, r * 2 * inner + c, d * gelu_f(g));
    store_split(hi, lo, r * 2 * inner + inner + c, d * a * gelu_grad_f(g));
  }
}
static int ew_blocks(size_t total) {
  size_t b = (total + 255) / 256;
  return (int)(b > 148 * 16 ? 148 * 16 : b);
}
int geglu_forward(const float* u, int rows, int inner, bf16* out_hi, bf16* out_lo, cudaStream_t s) {
  geglu_forward_kernel<<<ew_blocks((size_t)rows * inner), 256, 0, s>>>(u, rows, inner, out_hi, out_lo);
  RAMP_LAUNCH_OK();
  return 0;
}
int geglu_backward(const float* u, const float* dg, int rows, int inner, bf16* du_hi, bf16* du_lo, cudaStream_t s) {
  geglu_backward_kernel<<<ew_blocks((size_t)rows * inner), 256, 0, s>>>(u, dg, rows, inner, du_hi, du_lo);
  RAMP_LAUNCH_OK();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Time embedding (layers.py:233-259) + cond_mlp of every residual block (layers.py:338-346)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) time_cond_kernel(const float* __restrict__ tw, float t, const CondBlock* blocks) {
  __shared__ float pe[32], hid[128], emb[32];
  const int tid = threadIdx.x;
  if (tid < 16) {
    const float f = expf((float)tid * -(logf(10000.f) / 15.f));
    pe[tid] = sinf(t * f);
    pe[tid + 16] = cosf(t * f);
  }
  __syncthreads();
  {
    const float* w = tw + tid * 32;
    float acc = tw[128 * 32 + tid];
    for (int k = 0; k < 32; ++k) acc = fmaf(w[k], pe[k], acc);
    hid[tid] = mish_f(acc);
  }
  __syncthreads();
  if (tid < 32) {
    const float* w = tw + 128 * 32 + 128 + tid * 128;
    float acc = tw[128 * 32 + 128 + 32 * 128 + tid];
    for (int k = 0; k < 128; ++k) acc = fmaf(w[k], hid[k], acc);
    emb[tid] = acc / (1.f + expf(-acc));   // SiLU feeds every cond_mlp
  }
  __syncthreads();
  const CondBlock b = blocks[blockIdx.x];
  for (int c = tid; c < b.cout; c += blockDim.x) {
    float acc = b.b[c];
    for (int k = 0; k < 32; ++k) acc = fmaf(b.w[c * 32 + k], emb[k], acc);
    b.out[c] = acc;
  }
}
int time_cond(const float* tw, float t, const CondBlock* blocks_dev, int n_blocks, cudaStream_t s) {
  time_cond_kernel<<<n_blocks, 128, 0, s>>>(tw, t, blocks_dev);
  RAMP_LAUNCH_OK();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// attn2 (single context token) per-scene constants
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) scene_const_kernel(const SceneBlock* blocks, const float* __restrict__ latents,
                                                          int n_scenes, int ctx_dim) {
  __shared__ float v[256];
  const SceneBlock b = blocks[blockIdx.x];
  const int sidx = blockIdx.y, tid = threadIdx.x;
  const float* lat = latents + (size_t)sidx * ctx_dim;
  float acc = 0.f;
  for (int k = 0; k < ctx_dim; ++k) acc = fmaf(b.wv[(size_t)tid * ctx_dim + k], lat[k], acc);
  v[tid] = acc;
  __syncthreads();
  acc = b.bo2[tid] + b.bo1[tid];
  for (int k = 0; k < 256; ++k) acc = fmaf(b.wo[tid * 256 + k], v[k], acc);
  b.tab[(size_t)sidx * 256 + tid] = acc;
}
int scene_constants(const SceneBlock* blocks_dev, int n_blocks, const float* latents, int n_scenes, int ctx_dim,
                    cudaStream_t s) {
  scene_const_kernel<<<dim3(n_blocks, n_scenes), 256, 0, s>>>(blocks_dev, latents, n_scenes, ctx_dim);
  RAMP_LAUNCH_OK();
  return 0;
}

}  // namespace ramp
