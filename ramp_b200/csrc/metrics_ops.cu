// Evaluation metrics of the sampled trajectories on the device (the step right after the loop, and the
// definition of the "success" verdict): scripts/inference/core/metrics.py:8-126.
//   * collision intensity against axis-aligned boxes (metrics.py:48-82),
//   * path length / smoothness sums (metrics.py:21-45),
//   * waypoint variance: for every horizon step, the variance over ALL n x n entries of triu(cdist(p, p), 1) + eps
//     (metrics.py:9-19; the zeros of the lower triangle and the diagonal are part of the population), summed over the
//     horizon.  O(H n^2) pair distances, computed directly (no |a|^2 + |b|^2 - 2ab cancellation) and reduced in a
//     fixed order with fp64 partials, so the result is bit-reproducible,
//   * mean / unbiased std of a vector (path-length statistics over the collision-free trajectories).
#include "common.cuh"

namespace ramp {

__device__ __forceinline__ float mt_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double mt_warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// one warp per trajectory: fraction of waypoints inside any box (bounds = centre -+ size / 2, both inclusive)
__global__ void __launch_bounds__(256) collision_intensity_kernel(const float* __restrict__ x, int B, int H, int D,
                                                                  const float* __restrict__ centers,
                                                                  const float* __restrict__ sizes, int size_stride,
                                                                  int n_boxes, float* __restrict__ out) {
  extern __shared__ float sbox[];                  // [n_boxes][4] = lo_x, lo_y, hi_x, hi_y
  for (int i = threadIdx.x; i < n_boxes; i += blockDim.x) {
    const float cx = centers[2 * i], cy = centers[2 * i + 1];
    const float sx = sizes[i * size_stride], sy = sizes[i * size_stride + (size_stride > 1 ? 1 : 0)];
    sbox[4 * i] = cx - sx / 2.f; sbox[4 * i + 1] = cy - sy / 2.f;
    sbox[4 * i + 2] = cx + sx / 2.f; sbox[4 * i + 3] = cy + sy / 2.f;
  }
  __syncthreads();
  const int b = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (b >= B) return;
  float cnt = 0.f;
  for (int h = lane; h < H; h += 32) {
    const float px = x[((size_t)b * H + h) * D], py = x[((size_t)b * H + h) * D + 1];
    bool hit = false;
    for (int i = 0; i < n_boxes; ++i)
      hit = hit || (px >= sbox[4 * i] && px <= sbox[4 * i + 2] && py >= sbox[4 * i + 1] && py <= sbox[4 * i + 3]);
    cnt += hit ? 1.f : 0.f;
  }
  cnt = mt_warp_sum(cnt);
  if (lane == 0) out[b] = cnt / (float)H;
}

// one warp per trajectory: sum_t || x[t+1, c0:c1] - x[t, c0:c1] ||
__global__ void __launch_bounds__(256) path_sums_kernel(const float* __restrict__ x, int B, int H, int D, int c0, int c1,
                                                        float* __restrict__ out) {
  const int b = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (b >= B) return;
  float s = 0.f;
  for (int t = lane; t + 1 < H; t += 32) {
    const float* p = x + ((size_t)b * H + t) * D;
    float q = 0.f;
    for (int c = c0; c < c1; ++c) { const float d = p[D + c] - p[c]; q = fmaf(d, d, q); }
    s += sqrtf(q);
  }
  s = mt_warp_sum(s);
  if (lane == 0) out[b] = s;
}

// pair sums of one (horizon step, 128 x 128 tile pair): part[(h * npairs + pair) * 2 + {0,1}] = (sum d, sum d^2), i < j
static constexpr int WV_TILE = 128;
__global__ void __launch_bounds__(256) waypoint_pairs_kernel(const float* __restrict__ x, int n, int H, int D, int ntiles,
                                                             double* __restrict__ part) {
  __shared__ float2 pi[WV_TILE], pj[WV_TILE];
  __shared__ double red[2][8];
  const int npairs = ntiles * (ntiles + 1) / 2;
  const int h = blockIdx.x / npairs;
  int pair = blockIdx.x - h * npairs, ti = 0;
  while (pair >= ntiles - ti) { pair -= ntiles - ti; ++ti; }      // row ti of the upper-triangular tile matrix
  const int tj = ti + pair;
  if (threadIdx.x < WV_TILE) {
    const int i = ti * WV_TILE + threadIdx.x;
    pi[threadIdx.x] = i < n ? make_float2(x[((size_t)i * H + h) * D], x[((size_t)i * H + h) * D + 1]) : make_float2(0.f, 0.f);
  } else {
    const int j = tj * WV_TILE + threadIdx.x - WV_TILE;
    pj[threadIdx.x - WV_TILE] = j < n ? make_float2(x[((size_t)j * H + h) * D], x[((size_t)j * H + h) * D + 1]) : make_float2(0.f, 0.f);
  }
  __syncthreads();
  const int il = threadIdx.x & (WV_TILE - 1), jh = threadIdx.x >> 7;
  const int i = ti * WV_TILE + il;
  const float2 a = pi[il];
  float s1 = 0.f, s2 = 0.f;
#pragma unroll 8
  for (int jj = 0; jj < WV_TILE / 2; ++jj) {
    const int jl = jh * (WV_TILE / 2) + jj, j = tj * WV_TILE + jl;
    const float2 b = pj[jl];
    const float dx = a.x - b.x, dy = a.y - b.y;
    const float d2 = fmaf(dx, dx, dy * dy);
    if (i < j && j < n) { s1 += sqrtf(d2); s2 += d2; }
  }
  double d1 = mt_warp_sum((double)s1), d2 = mt_warp_sum((double)s2);
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = d1; red[1][threadIdx.x >> 5] = d2; }
  __syncthreads();
  if (threadIdx.x < 2) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += red[threadIdx.x][w];
    part[((size_t)h * npairs + (blockIdx.x - h * npairs)) * 2 + threadIdx.x] = t;
  }
}
// out = sum_h var_h,  var_h = (S2 - S1^2 / N) / (N - 1) over the N = n^2 entries (adding eps to every entry does not
// change a variance)
__global__ void __launch_bounds__(128) waypoint_var_final_kernel(const double* __restrict__ part, int n, int H, int npairs,
                                                                 float* __restrict__ out) {
  __shared__ double red[4];
  double acc = 0.0;
  const double N = (double)n * (double)n;
  for (int h = threadIdx.x; h < H; h += blockDim.x) {
    double s1 = 0.0, s2 = 0.0;
    for (int p = 0; p < npairs; ++p) { s1 += part[((size_t)h * npairs + p) * 2]; s2 += part[((size_t)h * npairs + p) * 2 + 1]; }
    acc += (s2 - s1 * s1 / N) / (N - 1.0);
  }
  acc = mt_warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) out[0] = (float)((red[0] + red[1]) + (red[2] + red[3]));
}

// out3 = (mean, unbiased std, sum) of v[0..n)
__global__ void __launch_bounds__(256) vector_stats_kernel(const float* __restrict__ v, int n, float* __restrict__ out3) {
  __shared__ double red[8];
  __shared__ double s_mean;
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) s += (double)v[i];
  s = mt_warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += red[w];
    s_mean = t / (double)n;
    out3[0] = (float)s_mean; out3[2] = (float)t;
  }
  __syncthreads();
  const double m = s_mean;
  double q = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) { const double d = (double)v[i] - m; q += d * d; }
  q = mt_warp_sum(q);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = q;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += red[w];
    out3[1] = (float)sqrt(t / (double)(n - 1));        // n == 1: 0 / 0 = NaN, as torch.std
  }
}

// ---- no-allocation variant of the waypoint variance (no reference counterpart) ----
// The partial sums of waypoint_pairs_kernel live in THIS file-scope device array instead of caller scratch: every GPU that
// loads the library holds WV_STATIC_BYTES (1 MiB) until exit.  In host code `g_wv_scratch` is only a host-side stand-in for
// the symbol, not a device address: the address is fetched once per device with cudaGetSymbolAddress (wv_static_ptr).
static constexpr size_t WV_STATIC_BYTES = 1u << 20;
static __device__ __align__(256) double g_wv_scratch[WV_STATIC_BYTES / sizeof(double)];
static constexpr int WV_MAX_DEVICES = 64;
static double* g_wv_ptr[WV_MAX_DEVICES] = {nullptr};

// set-up, once per device, with that device current: device address of the static scratch (no allocation)
static int wv_static_ptr(double** out) {
  int dev = 0;
  RAMP_CUDA_OK(cudaGetDevice(&dev));
  RAMP_CHECK(dev >= 0 && dev < WV_MAX_DEVICES, "ramp_waypoint_variance_static: device index out of range");
  if (!g_wv_ptr[dev]) {
    double* p = nullptr;
    RAMP_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_wv_scratch));
    RAMP_CHECK(p != nullptr, "ramp_waypoint_variance_static: cudaGetSymbolAddress returned null");
    g_wv_ptr[dev] = p;
  }
  *out = g_wv_ptr[dev];
  return 0;
}

}  // namespace ramp

using namespace ramp;

extern "C" {

size_t ramp_waypoint_variance_static_max_bytes(void) { return WV_STATIC_BYTES; }

int ramp_waypoint_variance_static(const float* x, int n, int H, int D, float* out1, void* stream) {
  RAMP_CHECK(x && out1, "ramp_waypoint_variance_static: null argument");
  RAMP_CHECK(n >= 2 && H >= 1 && D >= 2, "ramp_waypoint_variance_static: need at least 2 trajectories");
  const int nt = (n + WV_TILE - 1) / WV_TILE, npairs = nt * (nt + 1) / 2;
  const size_t need = (size_t)H * npairs * 2 * sizeof(double);
  if (need > WV_STATIC_BYTES) {
    char buf[224];
    snprintf(buf, sizeof(buf), "ramp_waypoint_variance_static: n = %d trajectories x H = %d needs %zu bytes of scratch, the "
             "static array holds %zu (use ramp_waypoint_variance with caller scratch)", n, H, need, (size_t)WV_STATIC_BYTES);
    set_error(buf);
    return -2;
  }
  double* ws = nullptr;
  if (wv_static_ptr(&ws)) return -1;
  // every partial the final kernel reads is written by the pairs kernel of THIS call: nothing has to start at zero
  waypoint_pairs_kernel<<<H * npairs, 256, 0, (cudaStream_t)stream>>>(x, n, H, D, nt, ws);
  RAMP_LAUNCH_OK();
  waypoint_var_final_kernel<<<1, 128, 0, (cudaStream_t)stream>>>(ws, n, H, npairs, out1);
  RAMP_LAUNCH_OK();
  return 0;
}

int ramp_collision_intensity(const float* x, int B, int H, int D, const float* box_centers, const float* box_sizes,
                             int size_stride, int n_boxes, float* out, void* stream) {
  if (B == 0) return 0;
  RAMP_CHECK(x && out && (n_boxes == 0 || (box_centers && box_sizes)), "ramp_collision_intensity: null argument");
  RAMP_CHECK(H >= 1 && D >= 2 && n_boxes >= 0 && n_boxes <= 2048 && (size_stride == 1 || size_stride == 2),
             "ramp_collision_intensity: need H >= 1, D >= 2, at most 2048 boxes, sizes [n] or [n, 2]");
  collision_intensity_kernel<<<(B + 7) / 8, 256, (size_t)(n_boxes > 0 ? n_boxes : 1) * 16, (cudaStream_t)stream>>>(
      x, B, H, D, box_centers, box_sizes, size_stride, n_boxes, out);
  RAMP_LAUNCH_OK();
  return 0;
}

int ramp_path_sums(const float* x, int B, int H, int D, int c0, int c1, float* out, void* stream) {
  if (B == 0) return 0;
  RAMP_CHECK(x && out, "ramp_path_sums: null argument");
  RAMP_CHECK(H >= 1 && 0 <= c0 && c0 < c1 && c1 <= D, "ramp_path_sums: need 0 <= c0 < c1 <= D");
  path_sums_kernel<<<(B + 7) / 8, 256, 0, (cudaStream_t)stream>>>(x, B, H, D, c0, c1, out);
  RAMP_LAUNCH_OK();
  return 0;
}

size_t ramp_waypoint_variance_workspace_bytes(int n, int H) {
  if (n <= 0 || H <= 0) return 0;
  const size_t nt = ((size_t)n + WV_TILE - 1) / WV_TILE;
  return (size_t)H * (nt * (nt + 1) / 2) * 2 * sizeof(double);
}

int ramp_waypoint_variance(const float* x, int n, int H, int D, void* workspace, float* out1, void* stream) {
  RAMP_CHECK(x && workspace && out1, "ramp_waypoint_variance: null argument");
  RAMP_CHECK(n >= 2 && H >= 1 && D >= 2, "ramp_waypoint_variance: need at least 2 trajectories");
  const int nt = (n + WV_TILE - 1) / WV_TILE, npairs = nt * (nt + 1) / 2;
  RAMP_CHECK((long)H * npairs < (1l << 31), "ramp_waypoint_variance: too many tiles");
  waypoint_pairs_kernel<<<H * npairs, 256, 0, (cudaStream_t)stream>>>(x, n, H, D, nt, (double*)workspace);
  RAMP_LAUNCH_OK();
  waypoint_var_final_kernel<<<1, 128, 0, (cudaStream_t)stream>>>((const double*)workspace, n, H, npairs, out1);
  RAMP_LAUNCH_OK();
  return 0;
}

int ramp_vector_stats(const float* v, int n, float* out3, void* stream) {
  RAMP_CHECK(v && out3 && n >= 1, "ramp_vector_stats: need a non-empty vector");
  vector_stats_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(v, n, out3);
  RAMP_LAUNCH_OK();
  return 0;
}

}  // extern "C"
