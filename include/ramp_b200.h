/* ramp_b200 — C ABI of the B200-native RAMP reverse-diffusion hot path.
 *
 * Plain pointers and sizes only.  Every pointer is a DEVICE pointer unless its name ends in
 * `_host`.  `stream` is a cudaStream_t passed as void*.  Every function returns 0 on success and
 * a negative status otherwise; ramp_last_error() then returns a description (thread-local).
 * Nothing here allocates after *_create; callers own all buffers they pass.
 *
 * The reference (wondmgezahu/RAMP) has no FFI: its seam is Python class lookup
 * (mpd/trainer/train_loaders.py:11-25).  Each entry point below cites the reference code whose
 * arithmetic it replaces; the Python classes in ramp_b200/ that mirror the reference API bind
 * these symbols with ctypes.
 */
#ifndef RAMP_B200_H
#define RAMP_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* ramp_last_error(void);
int ramp_version(void);
/* layout of the 16-bit operand planes (value = hi + lo) this build reads and writes: 2 = fp16 hi + fp16 residual (default),
 * 0 = bf16 hi + bf16 lo (-DRAMP_PLANES_FP16=0, `python -m ramp_b200.build --planes-bf16`) */
int ramp_plane_format(void);
/* number of CUDA kernels this library has launched in this process */
unsigned long long ramp_launch_count(void);
/* a CUDA-graph replay launches the kernels that were counted once, while the graph was captured: the host-side owner of
 * the graph (ramp_b200/graphs.py) adds that number per replay so that the counter keeps meaning "kernels launched" */
void ramp_launch_count_add(unsigned long long n);

/* ---- low-level dense contraction (building block of the score network; exposed for tests) ----
 * D[(n,l), :] = sum_t A[n, par_t + a_pstride*(l + joff_t), 0:K] . W[wtap_t, :, 0:K]^T (+bias +bias_tab[row2scene[n]] +residual)
 * A / W are 16-bit hi/lo planes (value = hi + lo, format = ramp_plane_format()); precision 3 = hi*hi + hi*lo + lo*hi, 1 = hi*hi.
 * taps_host: ntaps triples (par, joff, wtap).  Output row of (n,l) = n*o_L + l*o_stride + o_off.
 * Replaces torch F.linear / F.conv1d / F.conv_transpose1d calls of mpd/models/layers/layers.py:262-297
 * and mpd/models/layers/layers_attention_mini.py:41,58-62,92-99 and their autograd input-gradients. */
int ramp_gemm(const void* a_hi, const void* a_lo, int n_samples, int L_in, int ldA, int K, int a_pstride, int L_out,
              int ntaps, const int* taps_host, const void* w_hi, const void* w_lo, int Nout, int Kw, int n_wtaps,
              const float* bias, const float* bias_tab, const int* row2scene, const float* residual, int ldR,
              float* out_f32, int ldO, void* out_hi, void* out_lo, int ldS, int o_L, int o_stride, int o_off,
              int precision, void* stream);
/* fp32 -> bf16 hi/lo planes */
int ramp_split_bf16(const float* x, size_t n, void* hi, void* lo, void* stream);

/* ---- score network: TemporalUnetInference energy gradient ------------------------------------
 * Replaces TemporalUnetInference.forward / forward_no_energy / EnergyGradFunction
 * (mpd/models/diffusion_models/UnetInference.py:19-37,157-224): eps = d/dx 0.5*sum f(x)^2, with the
 * backward sweep written out analytically (no autograd). */
typedef struct ramp_unet_config {
  int state_dim;   /* D: 4 (2-D models) or 6 (3-D) */
  int horizon;     /* H: multiple of 8, 8..64 */
  int ctx_dim;     /* 320 (ObstacleEncoderSet) or 256 (ObstacleEncoder) */
  int precision;   /* 3 = split-bf16 x3 (fp32-equivalent, the parity mode); 1 = plain bf16 */
  int max_rows;    /* most network rows per call */
  int max_scenes;  /* most distinct conditioning latents */
} ramp_unet_config;
int ramp_unet_create(const ramp_unet_config* cfg, void** handle);
int ramp_unet_destroy(void* handle);
/* name = reference state-dict key without the leading "model." (e.g. "downs.0.0.blocks.0.block.0.weight");
 * data = fp32, host or device, copied. */
int ramp_unet_set_param(void* handle, const char* name, const float* data, size_t numel);
int ramp_unet_finalize(void* handle, void* stream);      /* packs weights, allocates the workspace */
size_t ramp_unet_workspace_bytes(void* handle);
/* latents [n_scenes, ctx_dim]: cached scene-encoder outputs (UnetInference.py:146-156); an all-zero row is
 * the unconditional branch (UnetInference.py:190-197). */
int ramp_unet_set_scenes(void* handle, const float* latents, int n_scenes, void* stream);
/* Network row n reads trajectory x[n / x_repeat] (x_repeat = G+1 evaluates every trajectory under each
 * conditioning without materialising repeat_interleave, diffusion_model_static.py:142-145,192-206);
 * x [ceil(n_rows/x_repeat), H, D]; t = the (batch-uniform) diffusion timestep; row2scene [n_rows] int32;
 * eps_out / f_out [n_rows, H, D], either may be NULL (f_out = forward_no_energy output). */
int ramp_unet_energy_grad(void* handle, const float* x, int n_rows, int x_repeat, float t, const int32_t* row2scene,
                          float* eps_out, float* f_out, void* stream);

/* Per-kernel-class device timing (CUDA events on the launch stream around every launch of the engine).
 * enable != 0 starts recording; a call with ms4_host != NULL synchronises and returns the accumulated
 * milliseconds and launch counts of {tensor-core GEMM, norms, attention core, elementwise} and resets them. */
int ramp_unet_profile(void* handle, int enable, float* ms4_host, int* launches4_host);
/* writes one CSV line (class, shape tag, ms) per launch recorded since the last fetch; call BEFORE the fetch */
int ramp_unet_profile_dump(void* handle, const char* path);

/* ---- fused sampler step (one launch per reverse-diffusion step) --------------------------------
 * eps rows come from ramp_unet_energy_grad; row of (trajectory b, group g) = b*eps_bstride + g*eps_gstride,
 * g == G is the unconditional row.  eps = eps_u + sum_g w_g (eps_g - eps_u)   [G == 1: (1+w) eps_c - w eps_u].
 * mode 0 (DDPM, sample_functions.py:19-48 + diffusion_model_static.py:149-172):
 *     x0 = clamp(sqrt_recip*x - sqrt_recipm1*eps, +-1); out = coef1*x0 + coef2*x + sigma*noise*noise_scale
 * mode 1 (DDIM eta=0, diffusion_model_static.py:259-333): out = sa_prev*x0 + sdir*((x - sa_t*x0)/sb_t)
 * mode 2: only x0_out / eps_out are written (the APF runs between the two halves of a DDIM step).
 * mode 3: out = x + sigma*noise*noise_scale (+ inpainting) only: the second half of a DDPM step whose posterior mean
 *         was pushed by the potential field first (diffusion_model_static.py:172-181); eps is not read.
 * Inpainting (sample_functions.py:5-10): waypoint h of every trajectory is overwritten by
 * pin_val[b*pin_bstride + h*D ...] when pin_mask[h] != 0.  out may alias x. */
typedef struct ramp_step_args {
  const float* x; const float* eps; long eps_bstride; long eps_gstride;
  int B, H, D, G; float w[8];
  float sqrt_recip, sqrt_recipm1;
  float coef1, coef2, sigma, noise_scale; const float* noise;
  float sa_t, sb_t, sa_prev, sdir;
  const unsigned char* pin_mask; const float* pin_val; long pin_bstride;
  float* out; float* x0_out; float* eps_out; int mode;
  int predict_epsilon;   /* 1: network output is noise (scripts); 0: it is x0 itself (diffusion_model_static.py:109-118) */
} ramp_step_args;
int ramp_step(const ramp_step_args* args, void* stream);
int ramp_ddim_update(const float* x, const float* x0, int B, int H, int D, float sa_t, float sb_t, float sa_prev,
                     float sdir, const unsigned char* pin_mask, const float* pin_val, long pin_bstride, float* out,
                     void* stream);
int ramp_inpaint(float* x, int B, int H, int D, const unsigned char* pin_mask, const float* pin_val, long pin_bstride,
                 void* stream);
/* q_sample (diffusion_model_dynamic.py:671-680); x0_bstride 0 broadcasts one trajectory over the batch */
int ramp_q_sample(const float* x0, long x0_bstride, const float* noise, int B, int H, int D, float sqrt_ac,
                  float sqrt_1mac, const unsigned char* pin_mask, const float* pin_val, long pin_bstride, float* out,
                  void* stream);

/* sm (diffusion_model_dynamic.py:192-214), in place: waypoints start+1 .. start+n become a constant-velocity segment
 * from waypoint `start` toward waypoint start+n (speed capped at max_vel, time step dt); D must be 4 */
int ramp_smooth_segment(float* x, int B, int H, int D, int start, int n, float dt, float max_vel, void* stream);

/* ---- point cloud: nearest point, APF, cost, selection ------------------------------------------
 * ramp_cloud_pack: cloud [P, dim] fp32 -> packed buffer of ramp_cloud_packed_points(P) float4
 * ramp_nn_query  : cKDTree.query(k=1, distance_upper_bound) (APFhelper.py:17-35): for q[i*qstride + 0..dim-1],
 *                  out_dist = distance (inf if >= upper_bound), out_idx = index (P if none), exact in fp32 with the lowest
 *                  index on ties.  perm (may be NULL): the packed cloud is a permutation of the caller's cloud (a spatially
 *                  sorted copy, which keeps the exact-tie fallback rare on dense clouds); perm[i] = caller's index of
 *                  packed point i, and out_idx / the tie rule refer to the caller's order. */
size_t ramp_cloud_packed_points(int P);
/* FP32 FMA micro-benchmark (measurement aid): fills every SM with independent FFMA chains on `stream`, waits,
 * and writes the achieved TFLOP/s (2 flops per FFMA) to *tflops_host: the denominator of the nearest-point kernel's roofline. */
int ramp_fp32_fma_probe(float* tflops_host, void* stream);
int ramp_cloud_pack(const float* cloud, int P, int dim, void* packed, void* stream);
int ramp_nn_query(const void* packed, const int32_t* perm, int P, int dim, const float* q, int Q, int qstride, float upper_bound,
                  float* out_dist, int32_t* out_idx, void* stream);
/* avoidance (APFhelper.py:37-104): windowed Gaussian push away from the nearest point; out may alias x */
int ramp_apf_static(const float* x, int B, int H, int D, int dim, const float* dist, const int32_t* idx,
                    const void* packed, int P, float thr, float strength, int window, float* out, void* stream);
/* avoidance (APFhelper_dynamic.py:107-142), in place; mode 0 = static boxes, 1 = pursuer disc (goal mixing).
 * Optional gate (diffusion_model_dynamic.py:408-425): trajectory b is pushed only if
 * |gate_x[b, gate_row, 0:2] - gate_pos[0:2]| < gate_thr (gate_x NULL = no gate); `active` is a second, explicit mask. */
int ramp_apf_dynamic(float* x, int B, int H, int D, const float* dist, const int32_t* idx, const void* packed, int P,
                     int mode, int window, int affected, float strength, float thr_force, const float* goal,
                     const unsigned char* active, const float* gate_x, int gate_row, const float* gate_pos,
                     float gate_thr, void* stream);
/* cost.py:3-54: terms[b] = (collision_free, path_length, smoothness); dist from ramp_nn_query(upper_bound = thr) */
int ramp_cost_terms(const float* x, int B, int H, int D, int dim, const float* dist, float thr, float* terms,
                    void* stream);
/* cost.py:56-88 over n (all-gathered) trajectories: out8 = (best index or -1, best cost, n_free, len min/max,
 * smooth min/max, rank of the best among the collision-free ones = cost.py's best_index).  normalize != 0 applies the
 * min-max normalisation (0/0 = NaN as in the reference).  costs_out (nullable) [n]: total cost per trajectory,
 * +inf where it collides. */
int ramp_select_best(const float* terms, int n, float w_smooth, float w_len, int normalize, float* out8,
                     float* costs_out, void* stream);

/* ---- scene encoder, 2-D: ObstacleEncoderSet.forward -----------------------------------------------
 * Replaces mpd/models/diffusion_models/obstacle_encoder.py:52-152 (ObstaclePositionalEncoding, SetAttention,
 * SetTransformerBlock, ObstacleEncoderSet): pts [n_scenes, 6, 64, 2] -> out [n_scenes, 320], fp32, eval mode.
 * params = ONE flat fp32 device buffer holding the module's state-dict tensors ([out, in] row-major, as stored)
 * concatenated in the order ramp_scene2d_param_name(0 .. count-1) (names are the keys under "scene_encoder.",
 * e.g. "set_transformers.1.2.attn.qkv.weight"; "pos_encoder.div_term" is the registered buffer).
 * workspace: ramp_scene2d_workspace_floats(n_scenes) floats of caller scratch. */
int ramp_scene2d_param_count(void);
const char* ramp_scene2d_param_name(int i);
size_t ramp_scene2d_param_numel(int i);
size_t ramp_scene2d_param_floats(void);
size_t ramp_scene2d_workspace_floats(int n_scenes);
int ramp_scene2d_encode(const float* params, const float* pts, int n_scenes, float* workspace, float* out, void* stream);

/* ---- scene encoder, 3-D: ObstacleEncoder.forward (eval mode) -----------------------------------------
 * Replaces mpd/models/diffusion_models/obstacle_encoder3d.py:5-94 (PointNetEncoder with running-stat BatchNorm,
 * SetTransformerBlock on nn.MultiheadAttention, ObstacleEncoder): pts [n_scenes, n_obs, n_pts, 3] -> out
 * [n_scenes, 256]; 1 <= n_obs <= 32, n_pts >= 1.  params / names / workspace as for the 2-D encoder. */
int ramp_scene3d_param_count(void);
const char* ramp_scene3d_param_name(int i);
size_t ramp_scene3d_param_numel(int i);
size_t ramp_scene3d_param_floats(void);
size_t ramp_scene3d_workspace_floats(int n_scenes, int n_obs);
int ramp_scene3d_encode(const float* params, const float* pts, int n_scenes, int n_obs, int n_pts, float* workspace,
                        float* out, void* stream);

/* ---- evaluation metrics of the sampled batch (scripts/inference/core/metrics.py) ------------------------
 * ramp_collision_intensity: Metrics.compute_collision_intensity (metrics.py:48-82): out[b] = fraction of the H
 *     waypoints of trajectory b inside any axis-aligned box [centre - size/2, centre + size/2] (inclusive);
 *     box_sizes is [n_boxes] (size_stride 1, square boxes) or [n_boxes, 2] (size_stride 2).
 * ramp_path_sums: out[b] = sum_t |x[b,t+1,c0:c1] - x[b,t,c0:c1]|  (compute_path_length with columns 0:2,
 *     compute_smoothness with the velocity columns; metrics.py:21-45).
 * ramp_waypoint_variance: Metrics.compute_variance_waypoints (metrics.py:9-19) over n >= 2 trajectories:
 *     out1[0] = sum_h var(triu(cdist(p_h, p_h), 1).view(-1) + eps), unbiased, population = all n*n entries.
 *     workspace: ramp_waypoint_variance_workspace_bytes(n, H) bytes, 8-byte aligned.
 * ramp_vector_stats: out3 = (mean, unbiased std, sum) of v[0..n)  (path-length mean / std, metrics.py:112-114). */
int ramp_collision_intensity(const float* x, int B, int H, int D, const float* box_centers, const float* box_sizes,
                             int size_stride, int n_boxes, float* out, void* stream);
int ramp_path_sums(const float* x, int B, int H, int D, int c0, int c1, float* out, void* stream);
size_t ramp_waypoint_variance_workspace_bytes(int n, int H);
int ramp_waypoint_variance(const float* x, int n, int H, int D, void* workspace, float* out1, void* stream);
int ramp_vector_stats(const float* v, int n, float* out3, void* stream);
/* ramp_waypoint_variance_static: NO COUNTERPART IN THE REFERENCE — added for callers
 * that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same result
 * as ramp_waypoint_variance, but the scratch is a file-scope `static __device__` array of
 * ramp_waypoint_variance_static_max_bytes() = 1 MiB that every GPU loading the library holds until exit; the call allocates
 * nothing (no cudaMalloc / cudaFree / allocating library routine).  Limit: H * T (T + 1) / 2 * 16 bytes <= 1 MiB with
 * T = ceil(n / 128) tiles, e.g. n <= 4096 trajectories at H = 64 (n <= 5632 is the last size that fits); larger sizes are
 * rejected with status -2 and a message in ramp_last_error().  All calls on a GPU share the one array, which keeps the last
 * call's contents (every entry read is rewritten by the same call, so nothing needs clearing): calls that overlap on the GPU
 * (different streams, or graph branches running concurrently) RACE — serialise them.  The first call on each device fetches
 * the array's device address (cudaGetSymbolAddress, with that device current); make it once outside a graph capture. */
size_t ramp_waypoint_variance_static_max_bytes(void);
int ramp_waypoint_variance_static(const float* x, int n, int H, int D, float* out1, void* stream);

/* ---- caller-side state of the dynamic planner on the device --------------------------------
 * ramp_pursuer_tick: one launch = MultiSphereFieldDynamics.update_centers (deps/torch_robotics/.../primitives.py:
 *     100-107) with the pursuit law of DynamicsGenerator.create_pursuit_dynamics (scripts/inference/core/utils.py:
 *     84-137), followed by generate_sphere_points (mpd/models/diffusion_models/APFhelper_dynamic.py:18-39).
 *     center [2] fp32 device, in-out (the reference keeps a float32 tensor between ticks); robot [>=2] fp32 device
 *     (x, y of the state the pursuer chases); time t = t0 + *tick (tick: device int, may be NULL = 0; `advance`
 *     increments it after use so graph replays walk forward); do_step = 0 only regenerates the points.
 *     points [n_rim + n_in, 2] fp32 (NULL = motion only): n_rim golden-angle rim points, then n_in interior points
 *     from row (*tick mod tape_ticks) of tape [tape_ticks, 2*n_in] fp64 = n_in draws of U(0,1) (radius) then n_in
 *     draws of U(0, 2pi) (angle), i.e. numpy's draws in the reference's order.  center_out [2] (may be NULL or alias
 *     another buffer) receives the new centre.  The reference's n_rim = int(num_points * 0.9).
 * ramp_box_points: generate_box_points (APFhelper_dynamic.py:41-68) for n_boxes boxes of float32 centres / sizes
 *     [n_boxes, 2]: out [n_boxes, n_per, 2]; n_edge [n_boxes] = the reference's randint draw per box; u_edge
 *     [n_boxes, n_per] (first n_edge[b] used) and u_in [n_boxes, n_per, 2] (first n_per - n_edge[b] pairs) = its
 *     np.random.rand draws, fp64. */
int ramp_pursuer_tick(float* center, const float* robot, int* tick, double t0, int do_step, int advance, float vel_x,
                      float vel_y, double pursuit_strength, double random_strength, double dt, double radius, int n_rim,
                      int n_in, const double* tape, int tape_ticks, float* points, float* center_out, void* stream);
int ramp_box_points(const float* centers, const float* sizes, int n_boxes, int n_per, const int* n_edge,
                    const double* u_edge, const double* u_in, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif
