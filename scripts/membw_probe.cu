// HBM bandwidth probe for the GEMM epilogue's store path (round 2, run J): how fast can one B200 READ, WRITE and COPY, and what
// do the epilogue's access shapes cost?  The per-launch diagnostic (profiles/r2_gemm_diag.txt) showed that every write-heavy
// launch of the score network moves its output at ~3.2 TB/s; this program tells a hardware write ceiling apart from an
// access-pattern loss.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/_bin/membw_probe scripts/membw_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__global__ void read_k(const float4* __restrict__ a, size_t n4, float* sink) {
  float s = 0.f;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const float4 v = __ldg(a + i);
    s += v.x + v.y + v.z + v.w;
  }
  if (s == 123.456f) *sink = s;
}
// fully coalesced: a warp writes 512 contiguous bytes per instruction
template <int HINT>
__global__ void write_k(float4* __restrict__ a, size_t n4) {
  const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    if (HINT == 0) a[i] = v;
    else if (HINT == 1) __stcs(a + i, v);
    else __stwt(a + i, v);
  }
}
__global__ void copy_k(const float4* __restrict__ a, float4* __restrict__ b, size_t n4) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) b[i] = __ldg(a + i);
}
// The epilogue's shape: a warp owns 32 consecutive rows of a [rows][row_bytes] array and walks the row in segments of SEG
// bytes; per segment it issues 32 * SEG / 512 instructions, each covering 512 / SEG rows x SEG bytes (16 B per lane).
// WARPS_PER_ROWBLOCK warps split the row between them (the epilogue groups).  READ = same shape as loads.
template <int SEG, bool READ>
__global__ void seg_k(float4* __restrict__ a, size_t rows, int row_bytes, int col_split, float* sink) {
  constexpr int LPR = SEG / 16;            // lanes per row segment
  constexpr int RPI = 32 / LPR;            // rows per instruction
  const int lane = threadIdx.x & 31;
  const size_t gw = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5, nw = ((size_t)gridDim.x * blockDim.x) >> 5;
  const size_t units = (rows / 32) * col_split;
  const int span = row_bytes / col_split;   // bytes of a row this warp covers
  float s = 0.f;
  const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
  for (size_t u = gw; u < units; u += nw) {
    const size_t rb = u / col_split;
    const int cs = (int)(u % col_split);
    char* base = reinterpret_cast<char*>(a) + rb * 32 * (size_t)row_bytes + (size_t)cs * span;
    for (int c = 0; c < span; c += SEG) {
#pragma unroll
      for (int i = 0; i < 32 / RPI; ++i) {
        const int r = lane / LPR + RPI * i;
        float4* p = reinterpret_cast<float4*>(base + (size_t)r * row_bytes + c + 16 * (lane % LPR));
        if (READ) { const float4 t = __ldg(p); s += t.x + t.w; } else *p = v;
      }
    }
  }
  if (READ && s == 123.456f) *sink = s;
}

template <typename F>
static void timeit(const char* name, double bytes, F f) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  f(); f();
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  printf("%-58s %8.3f ms  %7.1f GB/s\n", name, best, bytes / best / 1e6);
  fflush(stdout);
}

int main() {
  const size_t bytes = (size_t)4 << 30;            // 4 GiB per buffer: 32x the L2
  float4 *a, *b; float* sink;
  CK(cudaMalloc(&a, bytes)); CK(cudaMalloc(&b, bytes)); CK(cudaMalloc(&sink, 256));
  CK(cudaMemset(a, 0, bytes)); CK(cudaMemset(b, 0, bytes));
  int sms = 148; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const size_t n4 = bytes / 16;
  const int G = sms * 8, T = 256;
  timeit("read   LDG.128 coalesced", (double)bytes, [&] { read_k<<<G, T>>>(a, n4, sink); });
  timeit("write  STG.128 coalesced", (double)bytes, [&] { write_k<0><<<G, T>>>(a, n4); });
  timeit("write  STG.128 coalesced .cs", (double)bytes, [&] { write_k<1><<<G, T>>>(a, n4); });
  timeit("write  STG.128 coalesced .wt", (double)bytes, [&] { write_k<2><<<G, T>>>(a, n4); });
  timeit("write  cudaMemsetAsync", (double)bytes, [&] { CK(cudaMemsetAsync(a, 1, bytes)); });
  timeit("copy   LDG.128 -> STG.128 (read + write bytes)", 2.0 * bytes, [&] { copy_k<<<G, T>>>(a, b, n4); });
  timeit("copy   cudaMemcpyAsync D2D (read + write bytes)", 2.0 * bytes, [&] { CK(cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice)); });
  // epilogue shapes: rows of 8 KB (u: 2048 fp32), 4 KB (du planes: 2048 x 2 B), 1 KB (256 fp32), 512 B (256 x 2 B planes)
  const int rbs[4] = {8192, 4096, 1024, 512};
  for (int k = 0; k < 4; ++k) {
    const int rb = rbs[k];
    const size_t rows = bytes / rb;
    char nm[128];
    // one warp per 32-row block walks the whole row (col_split 1) or eight warps split it (8)
    for (int split = 1; split <= 8; split *= 8) {
      if (rb / split < 128) continue;
      snprintf(nm, sizeof nm, "write  rows of %4d B, 128 B segments, %d warp(s)/row block", rb, split);
      timeit(nm, (double)bytes, [&] { seg_k<128, false><<<G, T>>>(a, rows, rb, split, sink); });
      snprintf(nm, sizeof nm, "write  rows of %4d B,  64 B segments, %d warp(s)/row block", rb, split);
      timeit(nm, (double)bytes, [&] { seg_k<64, false><<<G, T>>>(a, rows, rb, split, sink); });
      snprintf(nm, sizeof nm, "write  rows of %4d B,  32 B segments, %d warp(s)/row block", rb, split);
      timeit(nm, (double)bytes, [&] { seg_k<32, false><<<G, T>>>(a, rows, rb, split, sink); });
      snprintf(nm, sizeof nm, "read   rows of %4d B, 128 B segments, %d warp(s)/row block", rb, split);
      timeit(nm, (double)bytes, [&] { seg_k<128, true><<<G, T>>>(a, rows, rb, split, sink); });
    }
  }
  // low occupancy like the epilogue: 8 warps per SM only
  timeit("write  STG.128 coalesced, 148 CTAs x 256 threads", (double)bytes, [&] { write_k<0><<<sms, 256>>>(a, n4); });
  timeit("write  rows of 8192 B, 128 B segments, 148 CTAs x 256 thr", (double)bytes, [&] { seg_k<128, false><<<sms, 256>>>(a, bytes / 8192, 8192, 8, sink); });
  timeit("read   LDG.128 coalesced, 148 CTAs x 256 threads", (double)bytes, [&] { read_k<<<sms, 256>>>(a, n4, sink); });
  return 0;
}
