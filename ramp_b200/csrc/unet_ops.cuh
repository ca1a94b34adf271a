// CUDA-core ops of the score network (norms, activations, attention core, embeddings) — fwd and
// input-gradient.  All tensors are channels-last fp32 [rows, C] with an explicit leading dimension;
// GEMM operands are produced directly as bf16 hi/lo planes.
#pragma once
#include "common.cuh"

namespace ramp {

// x [ceil(n_samples / x_repeat), L, D] fp32 (row n reads source row n / x_repeat) -> hi/lo planes [n_samples_pad, L, 16] (channels >= D and samples >= n zero)
int prep_input(const float* x, int n_samples, int n_pad, int L, int D, int x_repeat, bf16* hi, bf16* lo, cudaStream_t s);

// GroupNorm(8 groups) [+ Mish] [+ cond[C]] [+ add]  (layers.py:280-297, layers_attention_mini.py:77-80)
struct GnArgs {
  const float* x; int ldX;          // [n_samples*L, C]
  int n_samples, L, C;
  const float* gamma; const float* beta; float eps;
  int mish;                          // apply Mish after the affine
  const float* cond;                 // [C] added after the activation, or null
  const float* add; int ldAdd;       // optional tensor added last
  float* out_f32; int ldO;
  bf16* out_hi; bf16* out_lo; int ldS;
};
int gn_forward(const GnArgs& a, cudaStream_t s);
// dx = GN/Mish backward of dy (recomputes statistics from x) [+ add]
struct GnBwdArgs {
  const float* x; int ldX; int n_samples, L, C;
  const float* gamma; const float* beta; float eps; int mish;
  const float* dy; int ldDy;
  const float* add; int ldAdd;
  float* out_f32; int ldO;
  bf16* out_hi; bf16* out_lo; int ldS;
};
int gn_backward(const GnBwdArgs& a, cudaStream_t s);

// LayerNorm over C=256 channels per token (eps 1e-5)
int ln_forward(const float* x, int rows, int C, const float* gamma, const float* beta, bf16* out_hi, bf16* out_lo,
               cudaStream_t s);
// out = add + LN_backward(x, dy)
int ln_backward(const float* x, int rows, int C, const float* gamma, const float* dy, const float* add, float* out_f32,
                bf16* out_hi, bf16* out_lo, cudaStream_t s);

// softmax(q k^T / sqrt(64)) v for 4 heads x 64; qkv [n_samples*L, 768] fp32 (q | k | v), out hi/lo [rows, 256]
int attn_forward(const float* qkv, int n_samples, int L, bf16* out_hi, bf16* out_lo, cudaStream_t s);
// dqkv hi/lo [rows, 768] from d_out fp32 [rows, 256]
int attn_backward(const float* qkv, const float* d_out, int n_samples, int L, bf16* dq_hi, bf16* dq_lo, cudaStream_t s);

// tensor-core variants (attn_tc.cu, L = 64 / 32): operands and gradients are bf16 hi/lo planes end to end
bool attn_tc_supported(int L);
int attn_forward_tc(const bf16* qkv_hi, const bf16* qkv_lo, int n_samples, int L, bf16* out_hi, bf16* out_lo, cudaStream_t s);
int attn_backward_tc(const bf16* qkv_hi, const bf16* qkv_lo, const bf16* do_hi, const bf16* do_lo, int n_samples, int L,
                     bf16* dq_hi, bf16* dq_lo, cudaStream_t s);
// any L <= 64 (one CTA per (sample, head), scalar smem loops) — used when L is not 8/16/32/64
int attn_forward_generic(const float* qkv, int n_samples, int L, bf16* out_hi, bf16* out_lo, cudaStream_t s);
int attn_backward_generic(const float* qkv, const float* d_out, int n_samples, int L, bf16* dq_hi, bf16* dq_lo, cudaStream_t s);

// GEGLU: u [rows, 2*inner] (a | gate) -> a * gelu(gate)
int geglu_forward(const float* u, int rows, int inner, bf16* out_hi, bf16* out_lo, cudaStream_t s);
int geglu_backward(const float* u, const float* dg, int rows, int inner, bf16* du_hi, bf16* du_lo, cudaStream_t s);

// Time embedding + every residual block's conditioning vector for one scalar timestep.
// tw: packed [W1(128x32) | b1(128) | W2(32x128) | b2(32)]; per block r: cw[r] -> [Cout x 32 | Cout], out[r] -> [Cout]
struct CondBlock { const float* w; const float* b; float* out; int cout; };
int time_cond(const float* tw, float t, const CondBlock* blocks_dev, int n_blocks, cudaStream_t s);

// Cross-attention with a single context token collapses to a per-scene constant:
// tab[b][s] = bo1[b] + Wo2[b] (Wv2[b] latent[s]) + bo2[b]
struct SceneBlock { const float* wv; const float* wo; const float* bo2; const float* bo1; float* tab; };
int scene_constants(const SceneBlock* blocks_dev, int n_blocks, const float* latents, int n_scenes, int ctx_dim,
                    cudaStream_t s);

}  // namespace ramp
