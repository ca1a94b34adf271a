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

}  // extern "C"
