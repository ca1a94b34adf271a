// Scene encoders (point cloud -> conditioning latent) as CUDA kernels, FP32 on the CUDA cores.
//
// 2-D: ObstacleEncoderSet (mpd/models/diffusion_models/obstacle_encoder.py:94-152): per-point embedding
// (Linear 2->64, LayerNorm, GELU) ++ sinusoidal code of the obstacle centre ++ code of the max-normalised offset
// -> Linear 192->64, LayerNorm, GELU -> 384 tokens x 64; three independent stacks of three pre-norm transformer
// blocks (4 heads x 16, MLP 64->256->64, exact GELU); mean over tokens; Linear-GELU-Linear heads to 64 / 96 / 160.
// The whole encoder is 8 launches: embed, 3 x (LayerNorm + QKV | attention + projection + LayerNorm + MLP), pool.
// The three stacks and all scenes ride in the grid of every launch (S * 3 * 12 CTAs of 32 tokens), activations stay
// in a small L2-resident workspace.  Weights are read from ONE flat fp32 buffer in the reference's state-dict
// layout ([out, in] row-major); each CTA stages what it needs into shared memory transposed with a padded stride,
// so that a warp (lanes = output channels) reads them conflict-free.
#include "common.cuh"
#include <cstring>
#include <string>
#include <vector>

namespace ramp {

static constexpr int SE_HID = 64, SE_TOK = 384, SE_OBS = 6, SE_PTS = 64, SE_STACKS = 3, SE_BLOCKS = 3, SE_MLP = 256;
static constexpr int SE_TG = 32;                         // tokens per CTA
static constexpr int SE_GROUPS = SE_TOK / SE_TG;         // 12
static constexpr int SE_OUT = 320;

// ---- parameter table: names are the reference state-dict keys under "scene_encoder." -----------------------------
struct SeParam { std::string name; size_t numel; size_t offset; };
struct SeTable {
  std::vector<SeParam> p;
  size_t total = 0;
  void add(const std::string& n, size_t numel) { p.push_back({n, numel, total}); total += numel; }
  long find(const char* n) const {
    for (size_t i = 0; i < p.size(); ++i) if (p[i].name == n) return (long)i;
    return -1;
  }
};
struct BlockOff { size_t n1w, n1b, qkv, pw, pb, n2w, n2b, w1, b1, w2, b2; };
struct Se2dLayout {
  SeTable t;
  size_t div, pe_w, pe_b, pe_lnw, pe_lnb, ce_w, ce_b, ce_lnw, ce_lnb;
  BlockOff blk[SE_STACKS][SE_BLOCKS];
  size_t pool_w0[SE_STACKS], pool_b0[SE_STACKS], pool_w2[SE_STACKS], pool_b2[SE_STACKS];
  Se2dLayout() {
    const int dims[SE_STACKS] = {64, 96, 160};
    auto A = [&](const std::string& n, size_t k) { size_t o = t.total; t.add(n, k); return o; };
    div = A("pos_encoder.div_term", SE_HID / 2);
    pe_w = A("point_embedding.0.weight", SE_HID * 2); pe_b = A("point_embedding.0.bias", SE_HID);
    pe_lnw = A("point_embedding.1.weight", SE_HID); pe_lnb = A("point_embedding.1.bias", SE_HID);
    ce_w = A("combined_encoder.0.weight", SE_HID * 3 * SE_HID); ce_b = A("combined_encoder.0.bias", SE_HID);
    ce_lnw = A("combined_encoder.1.weight", SE_HID); ce_lnb = A("combined_encoder.1.bias", SE_HID);
    for (int s = 0; s < SE_STACKS; ++s)
      for (int b = 0; b < SE_BLOCKS; ++b) {
        const std::string q = "set_transformers." + std::to_string(s) + "." + std::to_string(b) + ".";
        BlockOff& o = blk[s][b];
        o.n1w = A(q + "norm1.weight", SE_HID); o.n1b = A(q + "norm1.bias", SE_HID);
        o.qkv = A(q + "attn.qkv.weight", 3 * SE_HID * SE_HID);
        o.pw = A(q + "attn.proj.weight", SE_HID * SE_HID); o.pb = A(q + "attn.proj.bias", SE_HID);
        o.n2w = A(q + "norm2.weight", SE_HID); o.n2b = A(q + "norm2.bias", SE_HID);
        o.w1 = A(q + "mlp.0.weight", SE_MLP * SE_HID); o.b1 = A(q + "mlp.0.bias", SE_MLP);
        o.w2 = A(q + "mlp.3.weight", SE_HID * SE_MLP); o.b2 = A(q + "mlp.3.bias", SE_HID);
      }
    for (int s = 0; s < SE_STACKS; ++s) {
      const std::string q = "poolings." + std::to_string(s) + ".";
      pool_w0[s] = A(q + "0.weight", (size_t)dims[s] * SE_HID); pool_b0[s] = A(q + "0.bias", dims[s]);
      pool_w2[s] = A(q + "2.weight", (size_t)dims[s] * dims[s]); pool_b2[s] = A(q + "2.bias", dims[s]);
    }
  }
};
static const Se2dLayout& se2d() { static Se2dLayout l; return l; }

// ---- device helpers ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float se_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float se_gelu(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752f)); }
// LayerNorm over 64 channels held two per lane (c = lane, lane + 32), eps 1e-5 (torch default, biased variance)
__device__ __forceinline__ void se_ln64(float& a, float& b, const float* __restrict__ w, const float* __restrict__ bias, int lane) {
  const float mean = se_warp_sum(a + b) * (1.f / 64.f);
  const float da = a - mean, db = b - mean;
  const float var = se_warp_sum(da * da + db * db) * (1.f / 64.f);
  const float r = rsqrtf(var + 1e-5f);
  a = da * r * w[lane] + bias[lane];
  b = db * r * w[lane + 32] + bias[lane + 32];
}
// stage W [N][K] (row-major, global) into shared memory as sW[k * (N + 1) + n]
__device__ __forceinline__ void se_stage_t(float* sW, const float* __restrict__ W, int N, int K) {
  for (int i = threadIdx.x; i < N * K; i += blockDim.x) {
    const int n = i / K, k = i - n * K;
    sW[k * (N + 1) + n] = __ldg(W + i);
  }
}

// ---- kernel E: points -> tokens (written once per stack) ---------------------------------------------------------
struct SeEmbedArgs { const float* prm; const float* pts; float* z; size_t div, pe_w, pe_b, pe_lnw, pe_lnb, ce_w, ce_b, ce_lnw, ce_lnb; };
__global__ void __launch_bounds__(256) se2d_embed_kernel(SeEmbedArgs a) {
  extern __shared__ float sm[];
  float* sW = sm;                                  // [192][65]
  float* sP = sW + 192 * 65;                       // [64][2] points, then offsets
  float* sObs = sP + 128;                          // [64] centre code
  float* sIn = sObs + 64;                          // [8 warps][192]
  __shared__ float sC[2], sPeak;
  const int s = blockIdx.x / SE_OBS, o = blockIdx.x - s * SE_OBS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float* P = a.prm;
  se_stage_t(sW, P + a.ce_w, SE_HID, 3 * SE_HID);
  if (threadIdx.x < 128) sP[threadIdx.x] = a.pts[((size_t)s * SE_OBS + o) * 128 + threadIdx.x];
  __syncthreads();
  if (warp == 0) {                                 // centre = mean over the 64 points
    const float cx = se_warp_sum(sP[2 * lane] + sP[2 * (lane + 32)]) * (1.f / 64.f);
    const float cy = se_warp_sum(sP[2 * lane + 1] + sP[2 * (lane + 32) + 1]) * (1.f / 64.f);
    float m = 0.f;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int p = lane + 32 * j;
      m = fmaxf(m, fmaxf(fabsf(sP[2 * p] - cx), fabsf(sP[2 * p + 1] - cy)));
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
    if (lane == 0) { sC[0] = cx; sC[1] = cy; sPeak = m; }
  }
  __syncthreads();
  const float cx = sC[0], cy = sC[1], inv = 1.f / (sPeak + 1e-8f);
  if (threadIdx.x < 64) {                          // code of the centre: even slots sin sum, odd slots cos sum
    const float d = P[a.div + (threadIdx.x >> 1)];
    sObs[threadIdx.x] = (threadIdx.x & 1) ? cosf(cx * d) + cosf(cy * d) : sinf(cx * d) + sinf(cy * d);
  }
  __syncthreads();
  float* in = sIn + warp * 192;
  for (int pp = 0; pp < 8; ++pp) {
    const int p = warp * 8 + pp;
    const float x = sP[2 * p], y = sP[2 * p + 1];
    // point embedding: Linear(2 -> 64), LayerNorm, GELU
    float e0 = P[a.pe_w + 2 * lane] * x + P[a.pe_w + 2 * lane + 1] * y + P[a.pe_b + lane];
    float e1 = P[a.pe_w + 2 * (lane + 32)] * x + P[a.pe_w + 2 * (lane + 32) + 1] * y + P[a.pe_b + lane + 32];
    se_ln64(e0, e1, P + a.pe_lnw, P + a.pe_lnb, lane);
    in[lane] = se_gelu(e0); in[lane + 32] = se_gelu(e1);
    in[64 + lane] = sObs[lane]; in[96 + lane] = sObs[lane + 32];
    const float rx = (x - cx) * inv, ry = (y - cy) * inv;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int c = lane + 32 * j;
      const float d = P[a.div + (c >> 1)];
      in[128 + c] = (c & 1) ? cosf(rx * d) + cosf(ry * d) : sinf(rx * d) + sinf(ry * d);
    }
    __syncwarp();
    float t0 = P[a.ce_b + lane], t1 = P[a.ce_b + lane + 32];
#pragma unroll 8
    for (int k = 0; k < 192; ++k) {
      const float v = in[k];
      t0 = fmaf(v, sW[k * 65 + lane], t0);
      t1 = fmaf(v, sW[k * 65 + lane + 32], t1);
    }
    se_ln64(t0, t1, P + a.ce_lnw, P + a.ce_lnb, lane);
    t0 = se_gelu(t0); t1 = se_gelu(t1);
    const int tok = o * SE_PTS + p;
#pragma unroll
    for (int st = 0; st < SE_STACKS; ++st) {
      float* z = a.z + (((size_t)s * SE_STACKS + st) * SE_TOK + tok) * SE_HID;
      z[lane] = t0; z[lane + 32] = t1;
    }
    __syncwarp();
  }
}

// ---- kernel A: LayerNorm(norm1) + QKV projection -----------------------------------------------------------------
struct SeQkvArgs { const float* prm; const float* z; float* qkv; size_t n1w[SE_STACKS], n1b[SE_STACKS], w[SE_STACKS]; };
__global__ void __launch_bounds__(256) se2d_qkv_kernel(SeQkvArgs a) {
  extern __shared__ float sm[];
  float* sW = sm;                                  // [64][193]
  float* sA = sW + 64 * 193;                       // [32 tokens][64]
  const int g = blockIdx.x % SE_GROUPS, ss = blockIdx.x / SE_GROUPS, st = ss % SE_STACKS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float* P = a.prm;
  se_stage_t(sW, P + a.w[st], 3 * SE_HID, SE_HID);
  const size_t tok0 = (size_t)ss * SE_TOK + g * SE_TG;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const int tl = warp * 4 + t;
    const float* z = a.z + (tok0 + tl) * SE_HID;
    float v0 = z[lane], v1 = z[lane + 32];
    se_ln64(v0, v1, P + a.n1w[st], P + a.n1b[st], lane);
    sA[tl * 64 + lane] = v0; sA[tl * 64 + lane + 32] = v1;
  }
  __syncthreads();
  float acc[4][6];
#pragma unroll
  for (int t = 0; t < 4; ++t)
#pragma unroll
    for (int j = 0; j < 6; ++j) acc[t][j] = 0.f;
  const float* A0 = sA + warp * 4 * 64;
#pragma unroll 4
  for (int k = 0; k < 64; ++k) {
    float w[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) w[j] = sW[k * 193 + lane + 32 * j];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float v = A0[t * 64 + k];
#pragma unroll
      for (int j = 0; j < 6; ++j) acc[t][j] = fmaf(v, w[j], acc[t][j]);
    }
  }
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    float* o = a.qkv + (tok0 + warp * 4 + t) * (3 * SE_HID);
#pragma unroll
    for (int j = 0; j < 6; ++j) o[lane + 32 * j] = acc[t][j];
  }
}

// ---- kernel B: attention + projection + residual + LayerNorm(norm2) + MLP + residual -----------------------------
struct SeBlockArgs {
  const float* prm; const float* qkv; float* z;
  size_t pw[SE_STACKS], pb[SE_STACKS], n2w[SE_STACKS], n2b[SE_STACKS], w1[SE_STACKS], b1[SE_STACKS], w2[SE_STACKS], b2[SE_STACKS];
};
static constexpr int SEB_R1 = 256 * 65;                  // region 1 floats: K/V block + merge buffer | W1^T | W2^T
static constexpr int SEB_SMEM_FLOATS = SEB_R1 + 64 * 65 + 3 * 32 * 65 + 32 * 257;
__global__ void __launch_bounds__(256) se2d_block_kernel(SeBlockArgs a) {
  extern __shared__ float sm[];
  float* R1 = sm;
  float* sWp = R1 + SEB_R1;                        // [64][65]
  float* sO = sWp + 64 * 65;                       // [32][65] attention output, later LayerNorm(norm2) output
  float* sZ = sO + 32 * 65;                        // [32][65] z + attention branch
  float* sT = sZ + 32 * 65;                        // [32][65] scratch
  float* sH = sT + 32 * 65;                        // [32][257] MLP hidden
  const int g = blockIdx.x % SE_GROUPS, ss = blockIdx.x / SE_GROUPS, st = ss % SE_STACKS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float* P = a.prm;
  const size_t tok0 = (size_t)ss * SE_TOK + g * SE_TG;
  se_stage_t(sWp, P + a.pw[st], SE_HID, SE_HID);

  // ---- attention: warp = (head, key half), lane = query; online softmax over 6 blocks of 64 keys ----
  {
    const int h = warp & 3, half = warp >> 2;
    float q[16], acc[16];
    const float* qp = a.qkv + (tok0 + lane) * 192 + h * 16;
#pragma unroll
    for (int d = 0; d < 16; ++d) { q[d] = qp[d] * 0.25f; acc[d] = 0.f; }      // scale = 16^-1/2
    float m = -INFINITY, l = 0.f;
    float* sKV = R1;                               // [64 keys][128] = K (64) | V (64)
    for (int kb = 0; kb < SE_TOK / 64; ++kb) {
      __syncthreads();
      const float* src = a.qkv + ((size_t)ss * SE_TOK + kb * 64) * 192 + 64;
      for (int i = threadIdx.x; i < 64 * 32; i += 256) {                         // float4 units: 32 per key
        const int key = i >> 5, c4 = i & 31;
        reinterpret_cast<float4*>(sKV)[i] = __ldg(reinterpret_cast<const float4*>(src + (size_t)key * 192) + c4);
      }
      __syncthreads();
      for (int j = half * 32; j < half * 32 + 32; ++j) {
        const float4* kr = reinterpret_cast<const float4*>(sKV + j * 128 + h * 16);
        float s = 0.f;
#pragma unroll
        for (int d4 = 0; d4 < 4; ++d4) {
          const float4 kk = kr[d4];
          s = fmaf(q[4 * d4], kk.x, s); s = fmaf(q[4 * d4 + 1], kk.y, s);
          s = fmaf(q[4 * d4 + 2], kk.z, s); s = fmaf(q[4 * d4 + 3], kk.w, s);
        }
        const float mn = fmaxf(m, s);
        const float sc = __expf(m - mn), p = __expf(s - mn);
        m = mn; l = l * sc + p;
        const float4* vr = reinterpret_cast<const float4*>(sKV + j * 128 + 64 + h * 16);
#pragma unroll
        for (int d4 = 0; d4 < 4; ++d4) {
          const float4 vv = vr[d4];
          acc[4 * d4] = fmaf(p, vv.x, acc[4 * d4] * sc); acc[4 * d4 + 1] = fmaf(p, vv.y, acc[4 * d4 + 1] * sc);
          acc[4 * d4 + 2] = fmaf(p, vv.z, acc[4 * d4 + 2] * sc); acc[4 * d4 + 3] = fmaf(p, vv.w, acc[4 * d4 + 3] * sc);
        }
      }
    }
    __syncthreads();
    float* mg = R1 + 64 * 128 + (h * 32 + lane) * 18;     // merge buffer behind the K/V block
    if (half == 1) {
      mg[0] = m; mg[1] = l;
#pragma unroll
      for (int d = 0; d < 16; ++d) mg[2 + d] = acc[d];
    }
    __syncthreads();
    if (half == 0) {
      const float m1 = mg[0], l1 = mg[1];
      const float mn = fmaxf(m, m1), s0 = __expf(m - mn), s1 = __expf(m1 - mn);
      const float invl = 1.f / (l * s0 + l1 * s1);
#pragma unroll
      for (int d = 0; d < 16; ++d) sO[lane * 65 + h * 16 + d] = (acc[d] * s0 + mg[2 + d] * s1) * invl;
    }
    __syncthreads();
  }
  // region 1 is free: start staging W1^T while the projection runs
  se_stage_t(R1, P + a.w1[st], SE_MLP, SE_HID);          // [64][257]

  // ---- projection + residual, LayerNorm(norm2): warp = 4 tokens, lane = channels (lane, lane + 32) ----
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const int tl = warp * 4 + t;
    float o0 = P[a.pb[st] + lane], o1 = P[a.pb[st] + lane + 32];
#pragma unroll 8
    for (int k = 0; k < 64; ++k) {
      const float v = sO[tl * 65 + k];
      o0 = fmaf(v, sWp[k * 65 + lane], o0);
      o1 = fmaf(v, sWp[k * 65 + lane + 32], o1);
    }
    const float* z = a.z + (tok0 + tl) * SE_HID;
    o0 += z[lane]; o1 += z[lane + 32];
    sZ[tl * 65 + lane] = o0; sZ[tl * 65 + lane + 32] = o1;
    se_ln64(o0, o1, P + a.n2w[st], P + a.n2b[st], lane);
    sT[tl * 65 + lane] = o0; sT[tl * 65 + lane + 32] = o1;
  }
  __syncthreads();
  // ---- MLP 64 -> 256 (GELU): warp = 4 tokens, lane = hidden units lane + 32 j ----
  {
    float acc[4][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float b = P[a.b1[st] + lane + 32 * j];
#pragma unroll
      for (int t = 0; t < 4; ++t) acc[t][j] = b;
    }
#pragma unroll 2
    for (int k = 0; k < 64; ++k) {
      float w[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = R1[k * 257 + lane + 32 * j];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float v = sT[(warp * 4 + t) * 65 + k];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[t][j] = fmaf(v, w[j], acc[t][j]);
      }
    }
#pragma unroll
    for (int t = 0; t < 4; ++t)
#pragma unroll
      for (int j = 0; j < 8; ++j) sH[(warp * 4 + t) * 257 + lane + 32 * j] = se_gelu(acc[t][j]);
  }
  __syncthreads();
  se_stage_t(R1, P + a.w2[st], SE_HID, SE_MLP);          // [256][65]
  __syncthreads();
  // ---- MLP 256 -> 64 + residual ----
  {
    float acc[4][2];
#pragma unroll
    for (int t = 0; t < 4; ++t) { acc[t][0] = P[a.b2[st] + lane]; acc[t][1] = P[a.b2[st] + lane + 32]; }
#pragma unroll 4
    for (int k = 0; k < 256; ++k) {
      const float w0 = R1[k * 65 + lane], w1 = R1[k * 65 + lane + 32];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float v = sH[(warp * 4 + t) * 257 + k];
        acc[t][0] = fmaf(v, w0, acc[t][0]);
        acc[t][1] = fmaf(v, w1, acc[t][1]);
      }
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int tl = warp * 4 + t;
      float* z = a.z + (tok0 + tl) * SE_HID;
      z[lane] = sZ[tl * 65 + lane] + acc[t][0];
      z[lane + 32] = sZ[tl * 65 + lane + 32] + acc[t][1];
    }
  }
}

// ---- kernel P: mean over tokens, Linear - GELU - Linear ------------------------------------------------------------
struct SePoolArgs { const float* prm; const float* z; float* out; size_t w0[SE_STACKS], b0[SE_STACKS], w2[SE_STACKS], b2[SE_STACKS]; };
__global__ void __launch_bounds__(256) se2d_pool_kernel(SePoolArgs a) {
  __shared__ float part[4][64], pooled[64], hid[160];
  const int s = blockIdx.x / SE_STACKS, st = blockIdx.x - s * SE_STACKS;
  const int dims[SE_STACKS] = {64, 96, 160}, offs[SE_STACKS] = {0, 64, 160};
  const int d = dims[st];
  const float* z = a.z + (size_t)blockIdx.x * SE_TOK * SE_HID;
  const int c = threadIdx.x & 63, pq = threadIdx.x >> 6;
  float acc = 0.f;
  for (int t = pq * 96; t < pq * 96 + 96; ++t) acc += z[t * 64 + c];
  part[pq][c] = acc;
  __syncthreads();
  if (threadIdx.x < 64) pooled[c] = ((part[0][c] + part[1][c]) + (part[2][c] + part[3][c])) * (1.f / SE_TOK);
  __syncthreads();
  const float* P = a.prm;
  if (threadIdx.x < d) {
    float v = P[a.b0[st] + threadIdx.x];
    const float* w = P + a.w0[st] + (size_t)threadIdx.x * 64;
    for (int k = 0; k < 64; ++k) v = fmaf(pooled[k], w[k], v);
    hid[threadIdx.x] = se_gelu(v);
  }
  __syncthreads();
  if (threadIdx.x < d) {
    float v = P[a.b2[st] + threadIdx.x];
    const float* w = P + a.w2[st] + (size_t)threadIdx.x * d;
    for (int k = 0; k < d; ++k) v = fmaf(hid[k], w[k], v);
    a.out[(size_t)s * SE_OUT + offs[st] + threadIdx.x] = v;
  }
}


// =====================================================================================================================
// 3-D: ObstacleEncoder (mpd/models/diffusion_models/obstacle_encoder3d.py:5-94): per-obstacle PointNet (1x1 convs
// 3 -> 64 -> 256, eval-mode BatchNorm, SELU, max over points), two nn.MultiheadAttention blocks over the obstacles
// (E = 256, 4 heads x 64, FFN 256 -> 512 -> 256 with SELU), Linear, max over obstacles, Linear - SELU - Linear.
// Two launches: one CTA per (scene, obstacle) for the PointNet, one CTA per scene for everything after it (<= 32
// obstacle tokens live in shared memory; every Linear streams its weights through a transposed 64-output tile).
// =====================================================================================================================
static constexpr int S3_E = 256, S3_FF = 512, S3_MAXT = 32;
struct S3Block { size_t n1w, n1b, inw, inb, ow, ob, n2w, n2b, f1w, f1b, f2w, f2b; };
struct Se3dLayout {
  SeTable t;
  size_t c1w, c1b, c2w, c2b, bn1w, bn1b, bn1m, bn1v, bn2w, bn2b, bn2m, bn2v, opw, opb, g0w, g0b, g2w, g2b;
  S3Block blk[2];
  Se3dLayout() {
    auto A = [&](const std::string& n, size_t k) { size_t o = t.total; t.add(n, k); return o; };
    c1w = A("point_processor.conv1.weight", 64 * 3); c1b = A("point_processor.conv1.bias", 64);
    c2w = A("point_processor.conv2.weight", S3_E * 64); c2b = A("point_processor.conv2.bias", S3_E);
    bn1w = A("point_processor.bn1.weight", 64); bn1b = A("point_processor.bn1.bias", 64);
    bn1m = A("point_processor.bn1.running_mean", 64); bn1v = A("point_processor.bn1.running_var", 64);
    bn2w = A("point_processor.bn2.weight", S3_E); bn2b = A("point_processor.bn2.bias", S3_E);
    bn2m = A("point_processor.bn2.running_mean", S3_E); bn2v = A("point_processor.bn2.running_var", S3_E);
    for (int b = 0; b < 2; ++b) {
      const std::string q = "set_transformer_blocks." + std::to_string(b) + ".";
      S3Block& o = blk[b];
      o.n1w = A(q + "norm1.weight", S3_E); o.n1b = A(q + "norm1.bias", S3_E);
      o.inw = A(q + "mha.in_proj_weight", 3 * S3_E * S3_E); o.inb = A(q + "mha.in_proj_bias", 3 * S3_E);
      o.ow = A(q + "mha.out_proj.weight", S3_E * S3_E); o.ob = A(q + "mha.out_proj.bias", S3_E);
      o.n2w = A(q + "norm2.weight", S3_E); o.n2b = A(q + "norm2.bias", S3_E);
      o.f1w = A(q + "ffn.0.weight", S3_FF * S3_E); o.f1b = A(q + "ffn.0.bias", S3_FF);
      o.f2w = A(q + "ffn.3.weight", S3_E * S3_FF); o.f2b = A(q + "ffn.3.bias", S3_E);
    }
    opw = A("output_proj.weight", S3_E * S3_E); opb = A("output_proj.bias", S3_E);
    g0w = A("global_pooling.0.weight", S3_E * S3_E); g0b = A("global_pooling.0.bias", S3_E);
    g2w = A("global_pooling.2.weight", S3_E * S3_E); g2b = A("global_pooling.2.bias", S3_E);
  }
};
static const Se3dLayout& se3d() { static Se3dLayout l; return l; }

__device__ __forceinline__ float se_selu(float x) {
  return 1.0507009873554805f * (x > 0.f ? x : 1.6732632423543772f * expm1f(x));
}

// ---- PointNet: one CTA per (scene, obstacle) -----------------------------------------------------------------------
struct S3PointArgs { const float* prm; const float* pts; float* x0; int n_pts; size_t c1w, c1b, c2w, c2b, bn1w, bn1b, bn1m, bn1v, bn2w, bn2b, bn2m, bn2v; };
__global__ void __launch_bounds__(256) se3d_pointnet_kernel(S3PointArgs a) {
  extern __shared__ float sm[];
  float* sW = sm;                                  // conv2 weight transposed [64][257]
  float* sH = sW + 64 * 257;                       // [8 warps][4 points][64]
  float* sMax = sH + 8 * 4 * 64;                   // [8 warps][256]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float* P = a.prm;
  se_stage_t(sW, P + a.c2w, S3_E, 64);
  __syncthreads();
  const float* pts = a.pts + (size_t)blockIdx.x * a.n_pts * 3;
  // eval-mode BatchNorm folded into scale / shift per channel
  float s1[2], t1[2], s2[8], t2[8], cb2[8], mx[8];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int c = lane + 32 * j;
    s1[j] = P[a.bn1w + c] * rsqrtf(P[a.bn1v + c] + 1e-5f);
    t1[j] = P[a.bn1b + c] - P[a.bn1m + c] * s1[j];
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = lane + 32 * j;
    s2[j] = P[a.bn2w + c] * rsqrtf(P[a.bn2v + c] + 1e-5f);
    t2[j] = P[a.bn2b + c] - P[a.bn2m + c] * s2[j];
    cb2[j] = P[a.c2b + c];
    mx[j] = -INFINITY;
  }
  float* h = sH + warp * 4 * 64;
  for (int p0 = warp * 4; p0 < a.n_pts; p0 += 32) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int p = min(p0 + t, a.n_pts - 1);      // a repeated point does not change the max
      const float x = pts[3 * p], y = pts[3 * p + 1], z = pts[3 * p + 2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int c = lane + 32 * j;
        const float v = P[a.c1w + 3 * c] * x + P[a.c1w + 3 * c + 1] * y + P[a.c1w + 3 * c + 2] * z + P[a.c1b + c];
        h[t * 64 + c] = se_selu(v * s1[j] + t1[j]);
      }
    }
    __syncwarp();
    float acc[4][8];
#pragma unroll
    for (int t = 0; t < 4; ++t)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[t][j] = cb2[j];
#pragma unroll 2
    for (int k = 0; k < 64; ++k) {
      float w[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = sW[k * 257 + lane + 32 * j];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float v = h[t * 64 + k];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[t][j] = fmaf(v, w[j], acc[t][j]);
      }
    }
#pragma unroll
    for (int t = 0; t < 4; ++t)
#pragma unroll
      for (int j = 0; j < 8; ++j) mx[j] = fmaxf(mx[j], se_selu(acc[t][j] * s2[j] + t2[j]));
    __syncwarp();
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) sMax[warp * 256 + lane + 32 * j] = mx[j];
  __syncthreads();
  {
    float m = sMax[threadIdx.x];
#pragma unroll
    for (int w = 1; w < 8; ++w) m = fmaxf(m, sMax[w * 256 + threadIdx.x]);
    a.x0[(size_t)blockIdx.x * S3_E + threadIdx.x] = m;
  }
}

// ---- everything after the PointNet: one CTA per scene --------------------------------------------------------------
struct S3SceneArgs { const float* prm; const float* x0; float* out; int n_obs; S3Block blk[2]; size_t opw, opb, g0w, g0b, g2w, g2b; };
// acc[t][j] += sum_k in[4 warp + t][k] * W[n0 + lane + 32 j][k], k < K (chunks of 256 staged transposed in sW)
__device__ __forceinline__ void s3_linear_tile(const float* sIn, int ldIn, const float* __restrict__ W, int K, int n0, float* sW,
                                               float (&acc)[4][2]) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int kc = 0; kc < K; kc += 256) {
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * 256; i += 256) {
      const int n = i >> 8, k = i & 255;
      sW[k * 65 + n] = __ldg(W + (size_t)(n0 + n) * K + kc + k);
    }
    __syncthreads();
#pragma unroll 4
    for (int k = 0; k < 256; ++k) {
      const float w0 = sW[k * 65 + lane], w1 = sW[k * 65 + lane + 32];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float v = sIn[(warp * 4 + t) * ldIn + kc + k];
        acc[t][0] = fmaf(v, w0, acc[t][0]);
        acc[t][1] = fmaf(v, w1, acc[t][1]);
      }
    }
  }
}
// LayerNorm over 256 channels of token `tok` (one warp), eps 1e-5
__device__ __forceinline__ void s3_ln_token(const float* x, float* y, const float* __restrict__ w, const float* __restrict__ b, int lane) {
  float v[8], s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) { v[j] = x[lane + 32 * j]; s += v[j]; }
  const float mean = se_warp_sum(s) * (1.f / 256.f);
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) { v[j] -= mean; q = fmaf(v[j], v[j], q); }
  const float r = rsqrtf(se_warp_sum(q) * (1.f / 256.f) + 1e-5f);
#pragma unroll
  for (int j = 0; j < 8; ++j) y[lane + 32 * j] = v[j] * r * w[lane + 32 * j] + b[lane + 32 * j];
}
__global__ void __launch_bounds__(256) se3d_scene_kernel(S3SceneArgs a) {
  extern __shared__ float sm[];
  float* sX = sm;                                  // [32][256] residual stream
  float* sA = sX + 32 * 256;                       // [32][256] LayerNorm output / attention output
  float* sB = sA + 32 * 256;                       // [32][768] qkv, or [32][512] FFN hidden, or [32][256] projected tokens
  float* sW = sB + 32 * 768;                       // [256][65] weight tile
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int T = a.n_obs;
  const float* P = a.prm;
  for (int i = threadIdx.x; i < 32 * 256; i += 256) {
    const int t = i >> 8;
    sX[i] = t < T ? a.x0[((size_t)blockIdx.x * T + t) * S3_E + (i & 255)] : 0.f;
    sA[i] = 0.f;
  }
  __syncthreads();
  for (int b = 0; b < 2; ++b) {
    const S3Block& o = a.blk[b];
    for (int t = warp; t < T; t += 8) s3_ln_token(sX + t * 256, sA + t * 256, P + o.n1w, P + o.n1b, lane);
    // qkv = in_proj(a)
    for (int n0 = 0; n0 < 3 * S3_E; n0 += 64) {
      float acc[4][2];
#pragma unroll
      for (int t = 0; t < 4; ++t) { acc[t][0] = P[o.inb + n0 + lane]; acc[t][1] = P[o.inb + n0 + lane + 32]; }
      s3_linear_tile(sA, 256, P + o.inw, S3_E, n0, sW, acc);
#pragma unroll
      for (int t = 0; t < 4; ++t) { sB[(warp * 4 + t) * 768 + n0 + lane] = acc[t][0]; sB[(warp * 4 + t) * 768 + n0 + lane + 32] = acc[t][1]; }
    }
    __syncthreads();
    // attention: thread = (head, query), 4 x 32 threads; keys / values broadcast from shared memory
    if (threadIdx.x < 128) {
      const int h = threadIdx.x >> 5, tq = lane;
      float q[64], p[S3_MAXT], out[64];
#pragma unroll
      for (int d = 0; d < 64; ++d) q[d] = sB[tq * 768 + h * 64 + d] * 0.125f;
      float m = -INFINITY;
#pragma unroll
      for (int j = 0; j < S3_MAXT; ++j) {
        float s = -INFINITY;
        if (j < T) {
          s = 0.f;
          const float4* kr = reinterpret_cast<const float4*>(sB + j * 768 + 256 + h * 64);
#pragma unroll
          for (int d4 = 0; d4 < 16; ++d4) {
            const float4 kk = kr[d4];
            s = fmaf(q[4 * d4], kk.x, s); s = fmaf(q[4 * d4 + 1], kk.y, s);
            s = fmaf(q[4 * d4 + 2], kk.z, s); s = fmaf(q[4 * d4 + 3], kk.w, s);
          }
        }
        p[j] = s; m = fmaxf(m, s);
      }
      float l = 0.f;
#pragma unroll
      for (int j = 0; j < S3_MAXT; ++j) { p[j] = j < T ? expf(p[j] - m) : 0.f; l += p[j]; }
      const float invl = 1.f / l;
#pragma unroll
      for (int d = 0; d < 64; ++d) out[d] = 0.f;
#pragma unroll
      for (int j = 0; j < S3_MAXT; ++j) {
        if (j < T) {
          const float4* vr = reinterpret_cast<const float4*>(sB + j * 768 + 512 + h * 64);
          const float pj = p[j] * invl;
#pragma unroll
          for (int d4 = 0; d4 < 16; ++d4) {
            const float4 vv = vr[d4];
            out[4 * d4] = fmaf(pj, vv.x, out[4 * d4]); out[4 * d4 + 1] = fmaf(pj, vv.y, out[4 * d4 + 1]);
            out[4 * d4 + 2] = fmaf(pj, vv.z, out[4 * d4 + 2]); out[4 * d4 + 3] = fmaf(pj, vv.w, out[4 * d4 + 3]);
          }
        }
      }
      if (tq < T)
#pragma unroll
        for (int d = 0; d < 64; ++d) sA[tq * 256 + h * 64 + d] = out[d];
    }
    __syncthreads();
    // x += out_proj(attention)
    for (int n0 = 0; n0 < S3_E; n0 += 64) {
      float acc[4][2];
#pragma unroll
      for (int t = 0; t < 4; ++t) { acc[t][0] = P[o.ob + n0 + lane]; acc[t][1] = P[o.ob + n0 + lane + 32]; }
      s3_linear_tile(sA, 256, P + o.ow, S3_E, n0, sW, acc);
#pragma unroll
      for (int t = 0; t < 4; ++t) { sX[(warp * 4 + t) * 256 + n0 + lane] += acc[t][0]; sX[(warp * 4 + t) * 256 + n0 + lane + 32] += acc[t][1]; }
    }
    __syncthreads();
    for (int t = warp; t < T; t += 8) s3_ln_token(sX + t * 256, sA + t * 256, P + o.n2w, P + o.n2b, lane);
    // hidden = SELU(ffn.0(a))
    for (int n0 = 0; n0 < S3_FF; n0 += 64) {
      float acc[4][2];
#pragma unroll
      for (int t = 0; t < 4; ++t) { acc[t][0] = P[o.f1b + n0 + lane]; acc[t][1] = P[o.f1b + n0 + lane + 32]; }
      s3_linear_tile(sA, 256, P + o.f1w, S3_E, n0, sW, acc);
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        sB[(warp * 4 + t) * 768 + n0 + lane] = se_selu(acc[t][0]);
        sB[(warp * 4 + t) * 768 + n0 + lane + 32] = se_selu(acc[t][1]);
      }
    }
    // x += ffn.3(hidden)
    for (int n0 = 0; n0 < S3_E; n0 += 64) {
      float acc[4][2];
#pragma unroll
      for (int t = 0; t < 4; ++t) { acc[t][0] = P[o.f2b + n0 + lane]; acc[t][1] = P[o.f2b + n0 + lane + 32]; }
      s3_linear_tile(sB, 768, P + o.f2w, S3_FF, n0, sW, acc);
#pragma unroll
      for (int t = 0; t < 4; ++t) { sX[(warp * 4 + t) * 256 + n0 + lane] += acc[t][0]; sX[(warp * 4 + t) * 256 + n0 + lane + 32] += acc[t][1]; }
    }
    __syncthreads();
  }
  // output_proj, then max over the obstacles
  for (int n0 = 0; n0 < S3_E; n0 += 64) {
    float acc[4][2];
#pragma unroll
    for (int t = 0; t < 4; ++t) { acc[t][0] = P[a.opb + n0 + lane]; acc[t][1] = P[a.opb + n0 + lane + 32]; }
    s3_linear_tile(sX, 256, P + a.opw, S3_E, n0, sW, acc);
#pragma unroll
    for (int t = 0; t < 4; ++t) { sB[(warp * 4 + t) * 768 + n0 + lane] = acc[t][0]; sB[(warp * 4 + t) * 768 + n0 + lane + 32] = acc[t][1]; }
  }
  __syncthreads();
  float* vec = sA;                                 // [256] pooled, [256..511] hidden
  {
    float m = -INFINITY;
    for (int t = 0; t < T; ++t) m = fmaxf(m, sB[t * 768 + threadIdx.x]);
    vec[threadIdx.x] = m;
  }
  __syncthreads();
  // global_pooling: Linear - SELU - Linear; warp-cooperative mat-vec (lanes over k)
  for (int stage = 0; stage < 2; ++stage) {
    const float* W = P + (stage == 0 ? a.g0w : a.g2w);
    const float* B = P + (stage == 0 ? a.g0b : a.g2b);
    const float* in = vec + stage * 256;
    for (int n = warp; n < S3_E; n += 8) {
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) s = fmaf(in[lane + 32 * j], __ldg(W + (size_t)n * S3_E + lane + 32 * j), s);
      s = se_warp_sum(s) + B[n];
      if (lane == 0) {
        if (stage == 0) vec[256 + n] = se_selu(s);
        else a.out[(size_t)blockIdx.x * S3_E + n] = s;
      }
    }
    __syncthreads();
  }
}

}  // namespace ramp

// ---- C ABI ---------------------------------------------------------------------------------------------------------
using namespace ramp;

extern "C" {

int ramp_scene2d_param_count(void) { return (int)se2d().t.p.size(); }
const char* ramp_scene2d_param_name(int i) {
  const auto& p = se2d().t.p;
  return (i >= 0 && i < (int)p.size()) ? p[i].name.c_str() : nullptr;
}
size_t ramp_scene2d_param_numel(int i) {
  const auto& p = se2d().t.p;
  return (i >= 0 && i < (int)p.size()) ? p[i].numel : 0;
}
size_t ramp_scene2d_param_floats(void) { return se2d().t.total; }
size_t ramp_scene2d_workspace_floats(int n_scenes) {
  return (size_t)(n_scenes > 0 ? n_scenes : 0) * SE_STACKS * SE_TOK * (SE_HID + 3 * SE_HID);
}

int ramp_scene2d_encode(const float* params, const float* pts, int n_scenes, float* workspace, float* out, void* stream) {
  if (n_scenes == 0) return 0;
  RAMP_CHECK(params && pts && workspace && out, "ramp_scene2d_encode: null argument");
  RAMP_CHECK(n_scenes > 0, "ramp_scene2d_encode: negative scene count");
  const Se2dLayout& L = se2d();
  cudaStream_t s = (cudaStream_t)stream;
  float* z = workspace;
  float* qkv = workspace + (size_t)n_scenes * SE_STACKS * SE_TOK * SE_HID;
  static bool attr = false;
  const int smem_e = (192 * 65 + 128 + 64 + 8 * 192) * 4, smem_a = (64 * 193 + 32 * 64) * 4, smem_b = SEB_SMEM_FLOATS * 4;
  if (!attr) {
    RAMP_CUDA_OK(cudaFuncSetAttribute(se2d_embed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_e));
    RAMP_CUDA_OK(cudaFuncSetAttribute(se2d_qkv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_a));
    RAMP_CUDA_OK(cudaFuncSetAttribute(se2d_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_b));
    attr = true;
  }
  {
    SeEmbedArgs a{params, pts, z, L.div, L.pe_w, L.pe_b, L.pe_lnw, L.pe_lnb, L.ce_w, L.ce_b, L.ce_lnw, L.ce_lnb};
    se2d_embed_kernel<<<n_scenes * SE_OBS, 256, smem_e, s>>>(a);
    RAMP_LAUNCH_OK();
  }
  const int grid = n_scenes * SE_STACKS * SE_GROUPS;
  for (int b = 0; b < SE_BLOCKS; ++b) {
    SeQkvArgs qa{}; qa.prm = params; qa.z = z; qa.qkv = qkv;
    SeBlockArgs ba{}; ba.prm = params; ba.qkv = qkv; ba.z = z;
    for (int st = 0; st < SE_STACKS; ++st) {
      const BlockOff& o = L.blk[st][b];
      qa.n1w[st] = o.n1w; qa.n1b[st] = o.n1b; qa.w[st] = o.qkv;
      ba.pw[st] = o.pw; ba.pb[st] = o.pb; ba.n2w[st] = o.n2w; ba.n2b[st] = o.n2b;
      ba.w1[st] = o.w1; ba.b1[st] = o.b1; ba.w2[st] = o.w2; ba.b2[st] = o.b2;
    }
    se2d_qkv_kernel<<<grid, 256, smem_a, s>>>(qa);
    RAMP_LAUNCH_OK();
    se2d_block_kernel<<<grid, 256, smem_b, s>>>(ba);
    RAMP_LAUNCH_OK();
  }
  {
    SePoolArgs pa{}; pa.prm = params; pa.z = z; pa.out = out;
    for (int st = 0; st < SE_STACKS; ++st) { pa.w0[st] = L.pool_w0[st]; pa.b0[st] = L.pool_b0[st]; pa.w2[st] = L.pool_w2[st]; pa.b2[st] = L.pool_b2[st]; }
    se2d_pool_kernel<<<n_scenes * SE_STACKS, 256, 0, s>>>(pa);
    RAMP_LAUNCH_OK();
  }
  return 0;
}

int ramp_scene3d_param_count(void) { return (int)se3d().t.p.size(); }
const char* ramp_scene3d_param_name(int i) {
  const auto& p = se3d().t.p;
  return (i >= 0 && i < (int)p.size()) ? p[i].name.c_str() : nullptr;
}
size_t ramp_scene3d_param_numel(int i) {
  const auto& p = se3d().t.p;
  return (i >= 0 && i < (int)p.size()) ? p[i].numel : 0;
}
size_t ramp_scene3d_param_floats(void) { return se3d().t.total; }
size_t ramp_scene3d_workspace_floats(int n_scenes, int n_obs) {
  return (size_t)(n_scenes > 0 ? n_scenes : 0) * (n_obs > 0 ? n_obs : 0) * S3_E;
}

int ramp_scene3d_encode(const float* params, const float* pts, int n_scenes, int n_obs, int n_pts, float* workspace, float* out,
                        void* stream) {
  if (n_scenes == 0) return 0;
  RAMP_CHECK(params && pts && workspace && out, "ramp_scene3d_encode: null argument");
  RAMP_CHECK(n_scenes > 0 && n_obs >= 1 && n_obs <= S3_MAXT && n_pts >= 1, "ramp_scene3d_encode: need 1..32 obstacles with >= 1 point");
  const Se3dLayout& L = se3d();
  cudaStream_t s = (cudaStream_t)stream;
  static bool attr = false;
  const int smem_p = (64 * 257 + 8 * 4 * 64 + 8 * 256) * 4, smem_s = (2 * 32 * 256 + 32 * 768 + 256 * 65) * 4;
  if (!attr) {
    RAMP_CUDA_OK(cudaFuncSetAttribute(se3d_pointnet_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_p));
    RAMP_CUDA_OK(cudaFuncSetAttribute(se3d_scene_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_s));
    attr = true;
  }
  S3PointArgs pa{params, pts, workspace, n_pts, L.c1w, L.c1b, L.c2w, L.c2b, L.bn1w, L.bn1b, L.bn1m, L.bn1v, L.bn2w, L.bn2b, L.bn2m, L.bn2v};
  se3d_pointnet_kernel<<<n_scenes * n_obs, 256, smem_p, s>>>(pa);
  RAMP_LAUNCH_OK();
  S3SceneArgs sa{}; sa.prm = params; sa.x0 = workspace; sa.out = out; sa.n_obs = n_obs;
  sa.blk[0] = L.blk[0]; sa.blk[1] = L.blk[1];
  sa.opw = L.opw; sa.opb = L.opb; sa.g0w = L.g0w; sa.g0b = L.g0b; sa.g2w = L.g2w; sa.g2b = L.g2b;
  se3d_scene_kernel<<<n_scenes, 256, smem_s, s>>>(sa);
  RAMP_LAUNCH_OK();
  return 0;
}

}  // extern "C"
