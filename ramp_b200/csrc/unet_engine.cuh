// Score-network engine: TemporalUnet forward + analytic input-gradient (energy gradient) on the GPU.
// Mirrors TemporalUnetInference.forward / forward_no_energy / EnergyGradFunction
// (mpd/models/diffusion_models/UnetInference.py:19-37,157-224).
#pragma once
#include "common.cuh"
#include <string>
#include <unordered_map>
#include <vector>

namespace ramp {

struct UnetConfig {
  int state_dim;     // D (4 in 2-D, 6 in 3-D)
  int horizon;       // H, multiple of 8, <= 64
  int ctx_dim;       // 320 (2-D scene encoder) / 256 (3-D)
  int precision;     // 1 or 3 (see gemm_tc.cu)
  int max_rows;      // largest number of network rows per call (workspace is sized for this)
  int max_scenes;    // largest number of distinct conditioning latents
};

struct PackedW {     // [ntaps][Nout][Kw] bf16 hi/lo planes
  bf16* hi = nullptr; bf16* lo = nullptr; int ntaps = 0, Nout = 0, Kw = 0, K = 0;
};
struct ConvW {       // a conv / linear layer: forward and input-gradient packings
  PackedW f, b; const float* bias = nullptr; int cin = 0, cout = 0, kw = 0;
};
struct ResW {
  ConvW c1, c2, res; bool has_res = false;
  const float *g1, *b1, *g2, *b2; float* cond = nullptr; int cin = 0, cout = 0;
};
struct TBlockW {
  PackedW qkv_f, qkv_b, o_f, o_b, ff1_f, ff1_b, ff2_f, ff2_b;
  const float *ln1g, *ln1b, *ln3g, *ln3b, *ff1_bias, *ff2_bias;   // ff1_bias: tile-interleaved copy (geglu_perm)
  float* bias_tab = nullptr;   // [max_scenes, 256]: attn1.to_out bias + collapsed attn2 output per scene
};
struct STW {
  const float *gn_g, *gn_b; ConvW pin, pout; TBlockW blk[2]; int C = 0;
};

// forward activations kept for the backward sweep
struct ResSave { float* c1; float* c2; };
struct TBlockSave { float* zin; float* qkv; Split qkv_s; float* z1; float* u; float* stats1; float* stats3; };   // qkv_s: planes (tensor-core attention)   // stats: (mean, rstd) per token
struct GemmFuse {      // fused-epilogue request for UnetEngine::gemm (see GemmEpi in common.cuh)
  int epi = EPI_PLAIN;
  const float* ln_gamma = nullptr; const float* ln_beta = nullptr; float* ln_stats = nullptr; const float* ln_x = nullptr;
  const float* u = nullptr; bf16* out2_hi = nullptr; bf16* out2_lo = nullptr; int ldS2 = 0;
};
struct STSave { const float* x; int ldx; TBlockSave blk[2]; };

class UnetEngine {
 public:
  explicit UnetEngine(const UnetConfig& cfg) : cfg_(cfg) {}
  ~UnetEngine();
  int set_param(const std::string& name, const float* data, size_t numel);
  int finalize(cudaStream_t s);
  int set_scenes(const float* latents_dev, int n_scenes, cudaStream_t s);
  // eps_out / f_out: [n_rows, H, D] (either may be null)
  int energy_grad(const float* x, int n_rows, int x_repeat, float t, const int* row2scene, float* eps_out,
                  float* f_out, cudaStream_t s);
  size_t workspace_bytes() const { return arena_bytes_; }
  int profile(int enable, float* ms4, int* launches4);
  int profile_dump(const char* path);   // per-launch CSV (class, tag, ms) of the events recorded since the last fetch
  const UnetConfig& cfg() const { return cfg_; }

 private:
  struct Param { float* ptr; size_t numel; };
  UnetConfig cfg_;
  std::unordered_map<std::string, Param> params_;
  std::vector<void*> owned_;
  bool finalized_ = false;
  int n_scenes_ = 0;

  // weights
  ResW down_rb_[4][2], mid_rb_[2], up_rb_[3][2];
  STW down_st_[4], mid_st_, up_st_[3];
  ConvW down_conv_[3], up_conv_[3], final_c5_, final_c1_;
  const float *final_g_, *final_b_;
  float* time_w_ = nullptr;          // packed time MLP
  void* cond_blocks_dev_ = nullptr;  // CondBlock[16]
  void* scene_blocks_dev_ = nullptr; // SceneBlock[16]
  std::vector<STW*> all_st_;

  // workspace arena (bump allocator; `dry_` = size-measuring pass)
  char* arena_ = nullptr; size_t arena_bytes_ = 0, arena_off_ = 0; bool dry_ = false;
  float* allocf(size_t n);
  Split allocs(size_t n);
  void* alloc_bytes(size_t n);

  // helpers
  const float* P(const std::string& name, size_t numel);
  int pack(PackedW& w, const std::string& name, int T, int Nd, int Kd, int Ns, int Ks, long st, long sn, long sk,
           int noff = 0, int koff = 0, bool alloc = true, long src_off = 0, size_t full_numel = 0, float scale = 1.f);
  int load_conv(ConvW& c, const std::string& prefix, int cin, int cout, int kw, bool transposed = false, float bscale = 1.f);
  int load_res(ResW& r, const std::string& prefix, int cin, int cout, int cin_true);
  int load_st(STW& st, const std::string& prefix, int C);
  void* dev_alloc(size_t bytes);

  struct Act {  // channels-last activation view
    const float* f = nullptr; int ldf = 0;
    const bf16* hi = nullptr; const bf16* lo = nullptr; int lds = 0;
    int C = 0;
  };
  struct OutSpec {  // where a layer should write its result
    bool want_f = true;
    bf16* hi = nullptr; bf16* lo = nullptr; int lds = 0;   // external split destination (e.g. concat buffer), or
    bool want_s = true;                                       // allocate a private one
  };
  int run(const float* x, int n_rows, int x_repeat, float t, const int* row2scene, float* eps_out, float* f_out,
          cudaStream_t s);
  int gemm(const Act& a, int K, const PackedW& w, const GemmTap* taps, int ntaps, int L_in, int L_out, int pstride,
           const float* bias, const float* bias_tab, const float* res, int ldr, const float* res2, int ldr2,
           float* of, int ldo, bf16* oh, bf16* ol, int lds, int o_L, int o_stride, int o_off,
           const GemmFuse* fuse = nullptr);
  int res_fwd(const ResW& w, const Act& x, int L, const OutSpec& o, Act& y, ResSave& sv);
  int res_bwd(const ResW& w, const ResSave& sv, const Act& dy, int L, const float* extra, int ldextra, Act& dx,
              bool want_split);
  int st_fwd(const STW& w, const Act& x, int L, const OutSpec& o, Act& y, STSave& sv);
  int st_bwd(const STW& w, const STSave& sv, const Act& dy, int L, Act& dx);

  int n_ = 0, npad_ = 0; const int* r2s_ = nullptr; cudaStream_t s_ = nullptr;

  // Row-split execution: a call over >= split_min_rows_ rows runs as `split_` independent parts on side streams, forked
  // from and joined back into the caller's stream with events.  Each launch of a dependent chain owns every SM with one
  // persistent CTA, so its last wave and its prologue leave SMs idle; the other part's launches fill them.
  int split_ = 2, split_min_rows_ = 2048; bool skip_time_cond_ = false;      // measured +1.4 % on the headline (profiles/r2_*)
  // RAMP_UNET_SPLIT_ALIGN = 1: parts are whole WAVES of the deepest level's tiles (SMs x 128 / (H / 8) rows each, rounded down
  // to whole trajectories) plus one small remainder part, instead of equal parts: 7168 rows are 448 tiles at L = 8 = 3 waves
  // of 148 + 4 tiles, and those 4 cost every launch of the level a fourth wave (see plan_parts)
  int split_align_ = 0, sms_ = 148;
  int plan_parts(int n_rows, int unit, int* rows) const;
  cudaStream_t side_[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t fork_ev_ = nullptr, join_ev_[4] = {nullptr, nullptr, nullptr, nullptr};
  int ensure_side_streams();

  // optional per-class timing
  bool prof_ = false;
  struct ProfEv { cudaEvent_t a, b; int cat; char tag[56]; };
  std::vector<ProfEv> prof_events_; size_t prof_used_ = 0;
 public:
  struct ProfScope {
    UnetEngine* e; int idx;
    ProfScope(UnetEngine* eng, int cat);
    void tag(const char* fmt, ...);
    ~ProfScope();
  };
};

}  // namespace ramp
