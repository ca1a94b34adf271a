// Caller-side state of the dynamic planner, moved onto the device: the pursuer's motion law
// (scripts/inference/core/utils.py:84-137 + MultiSphereFieldDynamics.update_centers, primitives.py:100-107), the
// pursuer disc sampler (APFhelper_dynamic.py:18-39) and the static box sampler (APFhelper_dynamic.py:41-68).
// The reference does these in numpy on the host every tick; here one launch moves the pursuer and rewrites its point
// cloud in HBM, so a graph-replayed replan tick needs no host arithmetic.  Random draws are INPUTS (a tape of numpy
// draws uploaded once), so a seeded run produces the reference's clouds.  Arithmetic follows the reference's dtype flow
// (float32 centres / sizes / velocity from torch tensors, float64 wherever numpy promotes) with explicitly rounded
// operations (no FMA contraction).
#include "../../include/ramp_b200.h"
#include "common.cuh"

namespace ramp {

struct PursuerLaw {
  float vx, vy;          // per-axis speed (float32 tensor in the reference)
  double pursuit, wobble;  // pursuit_strength, random_strength
  double dt;
};

// new centre (float32, as `to_torch(new_centers)` stores it) from the previous one
__device__ __forceinline__ void pursuer_law(float cx, float cy, float rx, float ry, double t, const PursuerLaw& w,
                                            float& ox, float& oy) {
  float dx = __fsub_rn(rx, cx), dy = __fsub_rn(ry, cy);
  const float dist = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
  if (dist > 0.f) {
    dx = __fdiv_rn(dx, dist);
    dy = __fdiv_rn(dy, dist);
  }
  const double ang = __dmul_rn(6.283185307179586, t);   // 2*np.pi*t
  const double hx = __dadd_rn((double)__fmul_rn((float)w.pursuit, dx), __dmul_rn(w.wobble, sin(ang)));
  const double hy = __dadd_rn((double)__fmul_rn((float)w.pursuit, dy), __dmul_rn(w.wobble, cos(ang)));
  double nx = __dadd_rn((double)cx, __dmul_rn(__dmul_rn(hx, (double)w.vx), w.dt));
  double ny = __dadd_rn((double)cy, __dmul_rn(__dmul_rn(hy, (double)w.vy), w.dt));
  nx = fmin(fmax(nx, -1.0), 1.0);
  ny = fmin(fmax(ny, -1.0), 1.0);
  ox = (float)nx;
  oy = (float)ny;
}

// One CTA: thread 0 advances the pursuer (when `do_step`), then all threads rewrite the disc cloud around the new centre.
// tick counter: t = t0 + (tick ? *tick : 0); `advance` increments *tick afterwards (graph replays walk the tape).
__global__ void __launch_bounds__(256) pursuer_tick_kernel(float* center, const float* robot, int* tick, double t0,
                                                           int do_step, int advance, PursuerLaw law, double radius,
                                                           int n_rim, int n_in, const double* tape, int tape_ticks,
                                                           float* points, float* center_out) {
  __shared__ float s_c[2];
  __shared__ int s_tick;
  if (threadIdx.x == 0) {
    const int k = tick ? *tick : 0;
    s_tick = k;
    float cx = center[0], cy = center[1];
    if (do_step) {
      pursuer_law(cx, cy, robot[0], robot[1], t0 + (double)k, law, cx, cy);
      center[0] = cx;
      center[1] = cy;
    }
    if (center_out) {
      center_out[0] = cx;
      center_out[1] = cy;
    }
    s_c[0] = cx;
    s_c[1] = cy;
    if (tick && advance) *tick = k + 1;
  }
  __syncthreads();
  if (!points) return;
  const double cx = (double)s_c[0], cy = (double)s_c[1];
  const double golden = __dmul_rn(3.141592653589793, __dsub_rn(3.0, sqrt(5.0)));
  const double* u = (tape && n_in > 0) ? tape + (size_t)(tape_ticks > 0 ? s_tick % tape_ticks : 0) * 2 * n_in : nullptr;
  for (int i = threadIdx.x; i < n_rim + n_in; i += blockDim.x) {
    double r, th;
    if (i < n_rim) {
      r = radius;
      th = __dmul_rn(golden, (double)i);
    } else {
      r = __dmul_rn(radius, sqrt(u[i - n_rim]));
      th = u[n_in + (i - n_rim)];
    }
    points[2 * i] = (float)__dadd_rn(__dmul_rn(r, cos(th)), cx);
    points[2 * i + 1] = (float)__dadd_rn(__dmul_rn(r, sin(th)), cy);
  }
}

// generate_box_points for float32 centres / sizes: one CTA per box.  u_edge [n_boxes, n_per] (only the first n_edge[b]
// of a row are used), u_in [n_boxes, n_per, 2] (first n_per - n_edge[b] pairs).
__global__ void __launch_bounds__(128) box_points_kernel(const float* centers, const float* sizes, int n_per,
                                                         const int* n_edge, const double* u_edge, const double* u_in,
                                                         float* out) {
  const int b = blockIdx.x;
  const float cx = centers[2 * b], cy = centers[2 * b + 1], w = sizes[2 * b], h = sizes[2 * b + 1];
  const float hw = __fdiv_rn(w, 2.f), hh = __fdiv_rn(h, 2.f);
  const float left = __fsub_rn(cx, hw), right = __fadd_rn(cx, hw), top = __fadd_rn(cy, hh), bottom = __fsub_rn(cy, hh);
  // the reference walks an 8-entry length table [w, w, h, h, w, w, h, h]: float32 cumulative sums (sequential) and a
  // float32 total (numpy's 8-accumulator pairwise order)
  float len[8] = {w, w, h, h, w, w, h, h}, cum[8];
  float run = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    run = i == 0 ? len[0] : __fadd_rn(run, len[i]);
    cum[i] = run;
  }
  const float half = __fadd_rn(__fadd_rn(w, w), __fadd_rn(h, h));
  const double total = (double)__fadd_rn(half, half);
  const float ex[4] = {left, right, right, left}, ey[4] = {top, top, bottom, bottom};
  const int ne = n_edge[b];
  for (int j = threadIdx.x; j < n_per; j += blockDim.x) {
    float* o = out + ((size_t)b * n_per + j) * 2;
    if (j < ne) {
      const double s = __dmul_rn(u_edge[(size_t)b * n_per + j], total);
      int seg = 0;
      while (seg < 7 && (double)cum[seg] < s) ++seg;          // np.searchsorted(..., side='left')
      const double before = seg == 0 ? 0.0 : (double)cum[seg - 1];
      const double frac = __ddiv_rn(__dsub_rn(s, before), (double)len[seg]);
      const int a = seg & 3, e = (seg + 1) & 3;
      o[0] = (float)__dadd_rn((double)ex[a], __dmul_rn(frac, (double)__fsub_rn(ex[e], ex[a])));
      o[1] = (float)__dadd_rn((double)ey[a], __dmul_rn(frac, (double)__fsub_rn(ey[e], ey[a])));
    } else {
      const double* u = u_in + ((size_t)b * n_per + (j - ne)) * 2;
      o[0] = (float)__dadd_rn(__dmul_rn(u[0], (double)w), (double)left);
      o[1] = (float)__dadd_rn(__dmul_rn(u[1], (double)h), (double)bottom);
    }
  }
}

}  // namespace ramp

using namespace ramp;

extern "C" {

int ramp_pursuer_tick(float* center, const float* robot, int* tick, double t0, int do_step, int advance, float vel_x,
                      float vel_y, double pursuit_strength, double random_strength, double dt, double radius, int n_rim,
                      int n_in, const double* tape, int tape_ticks, float* points, float* center_out, void* stream) {
  RAMP_CHECK(center, "ramp_pursuer_tick: null centre");
  RAMP_CHECK(!do_step || robot, "ramp_pursuer_tick: the motion law needs the robot position");
  RAMP_CHECK(n_rim >= 0 && n_in >= 0, "ramp_pursuer_tick: negative point count");
  RAMP_CHECK(!points || n_in == 0 || tape, "ramp_pursuer_tick: interior points need a tape of uniform draws");
  PursuerLaw law{vel_x, vel_y, pursuit_strength, random_strength, dt};
  pursuer_tick_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(center, robot, tick, t0, do_step, advance, law, radius, n_rim,
                                                           n_in, tape, tape_ticks, points, center_out);
  RAMP_LAUNCH_OK();
  return 0;
}

int ramp_box_points(const float* centers, const float* sizes, int n_boxes, int n_per, const int* n_edge,
                    const double* u_edge, const double* u_in, float* out, void* stream) {
  RAMP_CHECK(centers && sizes && n_edge && u_edge && u_in && out, "ramp_box_points: null argument");
  RAMP_CHECK(n_per >= 1, "ramp_box_points: at least one point per box");
  if (n_boxes <= 0) return 0;
  box_points_kernel<<<n_boxes, 128, 0, (cudaStream_t)stream>>>(centers, sizes, n_per, n_edge, u_edge, u_in, out);
  RAMP_LAUNCH_OK();
  return 0;
}

}  // extern "C"
