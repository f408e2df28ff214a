// C ABI of the B200-native sea-current planning path (include/seacurrent_b200.h).
//
// Host side of the drop-in boundary: owns the CUDA stream, the device scratch that the fused
// kernels need between stages (arc-length tables, controllable sets, Thomas scratch) and, for
// the host-buffer entry point, the device mirrors of the caller's arrays.  No planning
// arithmetic happens on the host: every stage is a kernel launch (path_lane_kernel.cu / path_kernel.cu,
// arc_kernel.cu, toppra_fast_kernel.cu / toppra_kernel.cu, param_kernel.cu).  There is no CPU fallback — without an sm_100 device
// scb_create fails with SCB_E_NO_DEVICE.
//
// Reference call sequence replaced (examples/json_test.cpp:28-58):
//   from_path -> arclength -> gen_vel_prof<1> -> resample -> angular_velocity.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <thread>
#include <vector>

#include "scb_common.cuh"

namespace {

struct Slot { void* p = nullptr; size_t bytes = 0; };
constexpr int kAutoHostChunk = 8192;  // scb_plan_batch: problems per chunk when the caller set none

enum SlotId {
  // scratch used when the caller did not ask for the corresponding output
  SL_L, SL_TABLE, SL_CTRL_X, SL_CTRL_Y, SL_K_LO, SL_K_HI, SL_X, SL_T, SL_XB, SL_CP, SL_KAPPA, SL_STATUS, SL_FAIL, SL_NS,
  // device mirrors of host inputs
  SL_IN_WX, SL_IN_WY, SL_IN_SA, SL_IN_VLO, SL_IN_VHI, SL_IN_VLG, SL_IN_VHG, SL_IN_ALO, SL_IN_AHI, SL_IN_VS,
  SL_IN_VE, SL_IN_PL, SL_IN_EDGES, SL_IN_CX, SL_IN_CY, SL_IN_TAB,
  // device mirrors of host outputs
  SL_O_STATUS, SL_O_FAIL, SL_O_CTRL_X, SL_O_CTRL_Y, SL_O_L, SL_O_TABLE, SL_O_T_TOTAL, SL_O_COST, SL_O_K_LO,
  SL_O_K_HI, SL_O_X, SL_O_T, SL_O_PX, SL_O_PY, SL_O_V, SL_O_A, SL_O_OMEGA, SL_O_NS, SL_O_STIME, SL_O_SPOS,
  SL_O_SVEL, SL_O_SACC, SL_O_SPX, SL_O_SPY, SL_O_SOMEGA,
  // FMT* (scb_fmt_batch): one float block of inputs, one of outputs
  SL_FMT_IN, SL_FMT_OUT,
  // obstacle-distance speed limit: curve points when the caller did not ask for them, capped per-gridpoint boxes
  SL_PX, SL_PY, SL_CAP_LO, SL_CAP_HI,
  SL_PP_IN, SL_PP_OUT,  // scb_toppra_ppoly: one block of inputs, one of outputs
  // scb_plan_batch_summary: every output stays resident in HBM until the next such call; scb_fetch gathers from it
  SL_R_STATUS, SL_R_FAIL, SL_R_CTRL_X, SL_R_CTRL_Y, SL_R_L, SL_R_TABLE, SL_R_T_TOTAL, SL_R_COST, SL_R_K_LO, SL_R_K_HI,
  SL_R_X, SL_R_T, SL_R_PX, SL_R_PY, SL_R_V, SL_R_A, SL_R_OMEGA, SL_R_SUMMARY, SL_F_IDS, SL_F_STAGE,
  SL_COUNT
};

}  // namespace

struct scb_ctx {
  scb_config cfg;
  cudaStream_t stream = nullptr;        // the ctx's own stream
  cudaStream_t run = nullptr;           // stream the current call launches on (= stream, or the caller's: scb_plan_batch_async)
  cudaStream_t last = nullptr;          // stream of the previous call; ev_last marks its end
  cudaEvent_t ev_last = nullptr;
  std::vector<void*> retired;           // scratch buffers replaced by larger ones: freed when the ctx is idle
  // results retained by scb_plan_batch_summary
  scb_batch_out kept = {};
  int kept_P = 0, kept_W = 0, kept_G = 0, kept_rec32 = 0;
  cudaStream_t hi = nullptr;            // highest-priority stream: the latency-bound stages (toppra, parametrizer) run here
  cudaEvent_t ev_hi[2] = {};            // hand-over stream -> hi, hi -> stream
  cudaStream_t copy_stream = nullptr;   // D2H of chunk c while chunk c+1 computes (scb_plan_batch)
  cudaEvent_t ev[9] = {}, ev_done[2] = {}, ev_copied[2] = {};
  Slot slot[2 * SL_COUNT];              // second half: mirror set 1 of the chunked host call
  int host_chunk = 0;                   // scb_set_host_chunk
  bool piped = false;
  scb_timing timing = {};
  bool timing_pending = false, host_call = false, split_spline = false;
  char err[256] = {0};
};

namespace {

using namespace scb;

int fail_cuda(scb_ctx* c, cudaError_t e, const char* where) {
  snprintf(c->err, sizeof(c->err), "%s: %s", where, cudaGetErrorString(e));
  return SCB_E_CUDA;
}
#define SCB_CUDA(call)                                         \
  do {                                                         \
    cudaError_t e_ = (call);                                   \
    if (e_ != cudaSuccess) return fail_cuda(ctx, e_, #call);   \
  } while (0)

// grow-only device buffer.  A buffer that has to grow may still be read by kernels in flight (an asynchronous call, or
// the other half of the chunk pipeline): it is retired, not freed - cudaFree would have to wait for the device.
int ensure(scb_ctx* ctx, int id, size_t bytes, void** out) {
  Slot& s = ctx->slot[id];
  if (bytes == 0) bytes = 16;
  if (s.bytes < bytes) {
    if (s.p) { ctx->retired.push_back(s.p); s.p = nullptr; s.bytes = 0; }
    cudaError_t e = cudaMalloc(&s.p, bytes);
    if (e != cudaSuccess) { fail_cuda(ctx, e, "cudaMalloc"); return e == cudaErrorMemoryAllocation ? SCB_E_NOMEM : SCB_E_CUDA; }
    s.bytes = bytes;
  }
  *out = s.p;
  return SCB_OK;
}

// free the scratch buffers that were replaced by larger ones; called only when the ctx's streams are idle
void release_retired(scb_ctx* ctx) {
  for (void* p : ctx->retired) cudaFree(p);
  ctx->retired.clear();
}

size_t real_size(const scb_ctx* c) { return c->cfg.precision == SCB_FP64 ? 8 : 4; }
int degree_of(const scb_ctx* c) { return c->cfg.spline_mode == SCB_REF_CUBIC ? 3 : 5; }

int validate(scb_ctx* ctx, const scb_batch_in* in, const scb_batch_out* out, int& stages) {
  if (!ctx || !in || !out) return SCB_E_INVALID;
  stages = in->stages ? in->stages : (SCB_STAGE_SPLINE | SCB_STAGE_TIMING | SCB_STAGE_GEOMETRY);
  if (in->P < 0 || in->P > ctx->cfg.max_problems) return SCB_E_INVALID;
  if (in->lp_mode < SCB_LP_DEFAULT || in->lp_mode > SCB_LP_CLOSED_FORM) return SCB_E_INVALID;
  if (in->param_mode != SCB_PARAM_SPLINE && in->param_mode != SCB_PARAM_CONST_ACCEL) return SCB_E_INVALID;
  if ((stages & SCB_STAGE_SPLINE) && (in->W < 2 || in->W > ctx->cfg.max_waypoints || !in->wx || !in->wy)) return SCB_E_INVALID;
  if ((stages & SCB_STAGE_TIMING) && (in->G < 2 || in->G > ctx->cfg.max_gridpoints)) return SCB_E_INVALID;
  if ((stages & SCB_STAGE_TIMING) && (!in->acc_lo || !in->acc_hi)) return SCB_E_INVALID;
  if ((stages & SCB_STAGE_TIMING) && !((in->vel_lo && in->vel_hi) || (in->vel_lo_grid && in->vel_hi_grid))) return SCB_E_INVALID;
  if ((stages & SCB_STAGE_TIMING) && !(stages & SCB_STAGE_SPLINE) && !in->path_len) return SCB_E_INVALID;
  if ((stages & SCB_STAGE_GEOMETRY) && !(stages & SCB_STAGE_SPLINE)) {
    // geometry of an EXISTING spline: control points and tables of an earlier SCB_STAGE_SPLINE call are inputs
    if ((stages & SCB_STAGE_TIMING) || !out->ctrl_x || !out->ctrl_y || !out->arclen_table) return SCB_E_INVALID;
    if (in->W < 2 || in->W > ctx->cfg.max_waypoints) return SCB_E_INVALID;
  }
  if (!(in->omega_max >= 0) || !(in->acc_lat_max >= 0)) return SCB_E_INVALID;
  if (in->n_obstacle_edges < 0 || (in->n_obstacle_edges > 0 && !in->obstacle_edges)) return SCB_E_INVALID;
  if (in->collision_chords < 0 || in->collision_chords > 4096) return SCB_E_INVALID;
  if (in->n_obstacle_edges > 0 && (stages & SCB_STAGE_SPLINE) && ctx->cfg.spline_mode != SCB_REF_CUBIC && in->collision_chords == 0)
    return SCB_E_INVALID;  // the tangent clipping is the cubic's; the quintic takes edges for the collision term only
  if (in->n_obstacle_edges > 0 && (stages & SCB_STAGE_SPLINE) && ctx->cfg.spline_mode != SCB_REF_CUBIC && in->W > 512)
    return SCB_E_INVALID;  // the collision term lives in the lane kernels' optimiser loop (W <= 512)
  if ((in->omega_max > 0 || in->acc_lat_max > 0) &&
      (stages & (SCB_STAGE_SPLINE | SCB_STAGE_TIMING | SCB_STAGE_GEOMETRY)) != (SCB_STAGE_SPLINE | SCB_STAGE_TIMING | SCB_STAGE_GEOMETRY))
    return SCB_E_INVALID;  // the caps need the curvature at the gridpoints
  if (!(in->a_brake >= 0) || !(in->t_react >= 0)) return SCB_E_INVALID;
  if (in->a_brake > 0 && in->n_obstacle_edges > 0 &&
      (stages & (SCB_STAGE_SPLINE | SCB_STAGE_TIMING | SCB_STAGE_GEOMETRY)) != (SCB_STAGE_SPLINE | SCB_STAGE_TIMING | SCB_STAGE_GEOMETRY))
    return SCB_E_INVALID;  // the limit needs the curve points of the gridpoints
  if (out->S_cap < 0 || out->S_cap > ctx->cfg.max_samples) return SCB_E_INVALID;
  if ((stages & SCB_STAGE_GEOMETRY) && !(stages & SCB_STAGE_TIMING) && (out->S_cap <= 0 || !out->n_samples || !out->s_pos)) return SCB_E_INVALID;
  if (!out->status) return SCB_E_INVALID;
  // the path kernel keeps one path's state in shared memory
  if (stages & SCB_STAGE_SPLINE) {
    int pb, th; bool tab; size_t sm;
    if (!path_launch_shape(in->W, ctx->cfg.spline_mode, (int)real_size(ctx), pb, th, tab, sm)) return SCB_E_INVALID;
  }
  return SCB_OK;
}

// Enqueue every stage on the ctx stream.  All pointers are device pointers.
template <class R>
int run_device(scb_ctx* ctx, const scb_batch_in* in, const scb_batch_out* out, int stages) {
  const int P = in->P, W = in->W, G = in->G, deg = degree_of(ctx);
  const size_t rs = sizeof(R);
  cudaStream_t st = ctx->run ? ctx->run : ctx->stream;
  int launches = 0;
  if (P == 0) return SCB_OK;
  // scratch is shared by every call on the ctx: a call on another stream than the previous one waits for it
  if (ctx->last && ctx->last != st) SCB_CUDA(cudaStreamWaitEvent(st, ctx->ev_last, 0));
  // float records (scb_batch_out.record_f32): only a conversion in the stores of the FP64 kernels
  const bool rec32 = out->record_f32 != 0 && sizeof(R) == 8;
  const bool do_spline = stages & SCB_STAGE_SPLINE, do_timing = stages & SCB_STAGE_TIMING,
             do_geo = stages & SCB_STAGE_GEOMETRY;
  int rc;
  void *dL = out->arclen_total, *dTab = out->arclen_table, *dCx = out->ctrl_x, *dCy = out->ctrl_y;
  int32_t* dFail = out->fail_stage;
  if (!dFail && (rc = ensure(ctx, SL_FAIL, (size_t)P * 4, (void**)&dFail))) return rc;
  SCB_CUDA(cudaEventRecord(ctx->ev[1], st));
  const bool want_stream = out->S_cap > 0 && (out->s_time || out->s_pos || out->s_vel || out->s_acc || out->s_px ||
                                              out->s_py || out->s_omega || out->n_samples);
  void* dKappa = nullptr;
  const bool curv = in->omega_max > 0 || in->acc_lat_max > 0;  // validate(): needs all three stages
  const bool obst = in->a_brake > 0 && in->n_obstacle_edges > 0 && in->obstacle_edges;  // ditto
  void *dPx = rec32 ? nullptr : out->px, *dPy = rec32 ? nullptr : out->py;  // R-typed curve points (the obstacle kernel's input)
  if (do_spline) {
    const size_t nseg = W - 1;
    // the control points and the arc-length tables are the hand-over between the path kernel and the
    // arc kernel (and the uniform-dt sampler): scratch when the caller did not ask for them
    if (!dL && (rc = ensure(ctx, SL_L, P * rs, &dL))) return rc;
    if (!dTab && (rc = ensure(ctx, SL_TABLE, (size_t)P * nseg * kTable * rs, &dTab))) return rc;
    if (!dCx && (rc = ensure(ctx, SL_CTRL_X, (size_t)P * nseg * (deg + 1) * rs, &dCx))) return rc;
    if (!dCy && (rc = ensure(ctx, SL_CTRL_Y, (size_t)P * nseg * (deg + 1) * rs, &dCy))) return rc;
    PathArgs a{};
    a.P = P; a.W = W; a.G = G; a.wx = in->wx; a.wy = in->wy; a.start_angle = in->start_angle;
    a.vel_start = in->vel_start; a.vel_end = in->vel_end;
    const bool clip = ctx->cfg.spline_mode == SCB_REF_CUBIC && in->obstacle_edges;  // shrink_tangent: cubic only
    a.edges = clip ? in->obstacle_edges : nullptr; a.n_edges = clip ? in->n_obstacle_edges : 0;
    if (ctx->cfg.spline_mode == SCB_LAU_QUINTIC && in->collision_chords > 0 && in->obstacle_edges && in->n_obstacle_edges > 0) {
      // Lau09 IV-B inside the optimiser loop: colliding trial iterates are rejected (path_lane_kernel.cu)
      a.coll_edges = in->obstacle_edges; a.n_coll_edges = in->n_obstacle_edges; a.coll_chords = in->collision_chords;
    }
    a.ctrl_x = dCx; a.ctrl_y = dCy; a.table = dTab; a.L = dL; a.cost = out->cost; a.status = out->status;
    if (do_geo && do_timing) {
      if (obst) {
        if (!dPx && (rc = ensure(ctx, SL_PX, (size_t)G * P * rs, &dPx))) return rc;
        if (!dPy && (rc = ensure(ctx, SL_PY, (size_t)G * P * rs, &dPy))) return rc;
      }
      a.px = dPx; a.py = dPy;
      if (rec32) { a.px32 = (float*)out->px; a.py32 = (float*)out->py; }
      if (out->omega || curv) { if ((rc = ensure(ctx, SL_KAPPA, (size_t)G * P * rs, &dKappa))) return rc; a.kappa = dKappa; }
    }
    a.opt = in->opt;
    launch_path<R>(a, ctx->cfg.spline_mode, st, ctx->ev[7]);
    launches += 2;  // path kernel + arc kernel
    if (in->collision_chords > 0 && in->obstacle_edges && in->n_obstacle_edges > 0) {
      CollideArgs ca{};
      ca.P = P; ca.W = W; ca.E = in->n_obstacle_edges; ca.M = in->collision_chords;
      ca.ctrl_x = dCx; ca.ctrl_y = dCy; ca.edges = in->obstacle_edges; ca.status = out->status; ca.cost = out->cost;
      launch_collide<R>(ca, deg, st);
      ++launches;
    }
    ctx->split_spline = true;
  } else {
    SCB_CUDA(cudaMemsetAsync(out->status, 0, (size_t)P * 4, st));
    dL = (void*)in->path_len;
    ctx->split_spline = false;
  }
  SCB_CUDA(cudaEventRecord(ctx->ev[2], st));
  // internal state of the recurrences is FP64 in both precision modes: in FP64 mode the caller's
  // K_lo / K_hi / t arrays double as that state, in FP32 mode it lives in scratch
  const bool f64 = sizeof(R) == 8, t_is_state = f64 && !rec32;  // the caller's FP64 t array doubles as the knot-time state
  void *dT = t_is_state ? out->t : nullptr, *dXb = nullptr, *dCp = nullptr, *dXsq = nullptr;
  if (do_timing) {
    void *dKlo = f64 ? out->K_lo : nullptr, *dKhi = f64 ? out->K_hi : nullptr, *dX = out->x;
    const size_t gp = (size_t)G * P * rs, gp8 = (size_t)G * P * 8;
    if (!dKlo && (rc = ensure(ctx, SL_K_LO, gp8, &dKlo))) return rc;
    if (!dKhi && (rc = ensure(ctx, SL_K_HI, gp8, &dKhi))) return rc;
    if (!dX && (rc = ensure(ctx, SL_X, gp, &dX))) return rc;
    dXsq = dX;
    if (!dT && (rc = ensure(ctx, SL_T, gp8, &dT))) return rc;
    if (in->param_mode == SCB_PARAM_SPLINE) {  // Thomas scratch; the const-accel parametrizer solves nothing
      if ((rc = ensure(ctx, SL_XB, gp8, &dXb))) return rc;
      if ((rc = ensure(ctx, SL_CP, gp8, &dCp))) return rc;
    }
    // the latency-bound stages go to the ctx's high-priority stream (own stream only: a caller's stream stays single)
    const bool use_hi = ctx->hi && st == ctx->stream;
    cudaStream_t st0 = st;
    if (use_hi) {
      SCB_CUDA(cudaEventRecord(ctx->ev_hi[0], st));
      SCB_CUDA(cudaStreamWaitEvent(ctx->hi, ctx->ev_hi[0], 0));
      st = ctx->hi;
    }
    ToppraArgs t{};
    t.P = P; t.G = G; t.L = dL; t.vel_lo = in->vel_lo; t.vel_hi = in->vel_hi;
    t.vel_lo_grid = in->vel_lo_grid; t.vel_hi_grid = in->vel_hi_grid;
    t.acc_lo = in->acc_lo; t.acc_hi = in->acc_hi; t.vel_start = in->vel_start; t.vel_end = in->vel_end;
    t.x_cap = in->x_cap > 0 ? in->x_cap : 100.0;
    if (curv) { t.kappa = dKappa; t.omega_max = in->omega_max; t.acc_lat_max = in->acc_lat_max; }
    if (obst) {  // Lau09 III-B eq. 2: the per-gridpoint boxes under the obstacle-distance limit replace the caller's
      void *dLo, *dHi;
      if ((rc = ensure(ctx, SL_CAP_LO, gp, &dLo)) || (rc = ensure(ctx, SL_CAP_HI, gp, &dHi))) return rc;
      ObstArgs oa{};
      oa.P = P; oa.G = G; oa.E = in->n_obstacle_edges; oa.px = dPx; oa.py = dPy; oa.edges = in->obstacle_edges;
      oa.vel_lo = in->vel_lo; oa.vel_hi = in->vel_hi; oa.vel_lo_grid = in->vel_lo_grid; oa.vel_hi_grid = in->vel_hi_grid;
      oa.a_brake = in->a_brake; oa.t_react = in->t_react; oa.out_lo = dLo; oa.out_hi = dHi;
      launch_obst_cap<R>(oa, st);
      ++launches;
      t.vel_lo_grid = dLo; t.vel_hi_grid = dHi;
    }
    t.K_lo = dKlo; t.K_hi = dKhi; t.x = dX; t.status = out->status; t.fail_stage = dFail;
    if (!f64) { t.K_lo_out = out->K_lo; t.K_hi_out = out->K_hi; }
    if (in->lp_mode == SCB_LP_SEIDEL) launch_toppra<R>(t, st);
    else launch_toppra_fast<R>(t, st);
    ++launches;
    SCB_CUDA(cudaEventRecord(ctx->ev[3], st));
    ParamArgs q{};
    q.P = P; q.G = G; q.L = dL; q.vel_start = in->vel_start; q.vel_end = in->vel_end;
    q.x = dX; q.kappa = dKappa;
    q.t = t_is_state ? nullptr : out->t; q.v = out->v; q.a = out->a; q.omega = out->omega;
    q.rec_f32 = rec32 ? 1 : 0;
    q.t_buf = dT; q.X_buf = dXb; q.cp_buf = dCp; q.T_total = out->T_total; q.status = out->status;
    q.keep_X = want_stream ? 1 : 0;
    q.param_mode = in->param_mode;
    launch_param<R>(q, st);
    ++launches;
    if (use_hi) {
      SCB_CUDA(cudaEventRecord(ctx->ev_hi[1], st));
      st = st0;
      SCB_CUDA(cudaStreamWaitEvent(st, ctx->ev_hi[1], 0));
    }
  } else {
    SCB_CUDA(cudaMemsetAsync(dFail, 0xff, (size_t)P * 4, st));
    SCB_CUDA(cudaEventRecord(ctx->ev[3], st));
  }
  SCB_CUDA(cudaEventRecord(ctx->ev[4], st));
  if (want_stream && (do_timing || do_geo)) {
    int32_t* dNs = out->n_samples;
    if (!dNs && (rc = ensure(ctx, SL_NS, (size_t)P * 4, (void**)&dNs))) return rc;
    SampleArgs s{};
    s.P = P; s.G = G; s.W = W; s.deg = deg; s.S_cap = out->S_cap;
    s.dt = in->dt > 0 ? in->dt : (double)0.02f;
    s.L = dL; s.vel_start = in->vel_start; s.vel_end = in->vel_end; s.t_buf = dT; s.X_buf = dXb;
    s.x = dXsq; s.param_mode = in->param_mode;
    s.ctrl_x = dCx; s.ctrl_y = dCy; s.table = dTab; s.n_samples = dNs;
    s.s_time = out->s_time; s.s_pos = out->s_pos; s.s_vel = out->s_vel; s.s_acc = out->s_acc;
    s.s_px = out->s_px; s.s_py = out->s_py; s.s_omega = out->s_omega; s.status = out->status;
    s.timing = do_timing ? 1 : 0; s.geometry = do_geo ? 1 : 0; s.nudge = in->nudge_positions ? 1 : 0;
    s.rec_f32 = rec32 ? 1 : 0;
    launch_sample<R>(s, st);
    ++launches;
  }
  SCB_CUDA(cudaEventRecord(ctx->ev[5], st));
  SCB_CUDA(cudaEventRecord(ctx->ev_last, st));
  ctx->last = st;
  SCB_CUDA(cudaGetLastError());
  ctx->timing.launches = launches;
  ctx->timing_pending = true;
  return SCB_OK;
}

template <class R>
__global__ void pack_summary_kernel(int P, const int32_t* status, const int32_t* fail, const R* T, const R* L, const R* J,
                                    scb_summary* dst) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  scb_summary s;
  s.status = status[p]; s.fail_stage = fail ? fail[p] : -1;
  s.T_total = T ? (double)T[p] : 0.0; s.arclen_total = L ? (double)L[p] : 0.0; s.cost = J ? (double)J[p] : 0.0;
  dst[p] = s;
}

int run_device_any(scb_ctx* ctx, const scb_batch_in* in, const scb_batch_out* out, int stages) {
  return ctx->cfg.precision == SCB_FP64 ? run_device<double>(ctx, in, out, stages) : run_device<float>(ctx, in, out, stages);
}

void collect_timing(scb_ctx* ctx) {
  if (!ctx->timing_pending) return;
  scb_timing& t = ctx->timing;
  auto el = [&](int a, int b) { float ms = 0; cudaEventElapsedTime(&ms, ctx->ev[a], ctx->ev[b]); return ms; };
  t.spline_ms = el(1, 2); t.toppra_ms = el(2, 3); t.param_ms = el(3, 4); t.sample_ms = el(4, 5);
  t.path_ms = ctx->split_spline ? el(1, 7) : 0; t.arc_ms = ctx->split_spline ? el(7, 2) : 0;
  if (ctx->host_call && ctx->piped) { t.h2d_ms = 0; t.d2h_ms = el(8, 6); t.total_ms = el(0, 6); }  // stages: last chunk
  else if (ctx->host_call) { t.h2d_ms = el(0, 1); t.d2h_ms = el(5, 6); t.total_ms = el(0, 6); }
  else { t.h2d_ms = 0; t.d2h_ms = 0; t.total_ms = el(1, 5); }
  ctx->timing_pending = false;
}

}  // namespace

extern "C" {

int scb_version(void) { return SCB_VERSION; }

const char* scb_strerror(int code) {
  switch (code) {
    case SCB_OK: return "ok";
    case SCB_E_INVALID: return "invalid argument";
    case SCB_E_NO_DEVICE: return "no usable sm_100 CUDA device (there is no CPU fallback)";
    case SCB_E_CUDA: return "CUDA runtime error (see scb_last_cuda_error)";
    case SCB_E_NOMEM: return "out of device memory";
    case SCB_E_TOO_LARGE: return "G * P exceeds the static scratch of the allocation-free entry (SCB_STATIC_MAX_GP)";
    default: return "unknown error";
  }
}

void scb_default_opt(scb_opt* o) {
  if (!o) return;
  o->iters = 32; o->move_waypoints = 0; o->lambda = 0.05;
  o->e_init = 1.0; o->e_min = 0.25; o->e_max = 4.0;
  o->step_e = 0.1; o->step_w = 0.05; o->w_radius = 0.25;
}

int64_t scb_algorithmic_bytes(int32_t precision, int32_t spline_mode, int32_t P, int32_t W, int32_t G) {
  // b*[(2W+8) + 2(d+1)(W-1) + 9G + 3] + 4 bytes per trajectory
  const int64_t b = precision == SCB_FP64 ? 8 : 4, d = spline_mode == SCB_REF_CUBIC ? 3 : 5;
  const int64_t per = b * ((2 * (int64_t)W + 8) + 2 * (d + 1) * ((int64_t)W - 1) + 9 * (int64_t)G + 3) + 4;
  return per * P;
}

int scb_create(scb_ctx** out, const scb_config* cfg) {
  if (!out || !cfg) return SCB_E_INVALID;
  *out = nullptr;
  if (cfg->precision != SCB_FP64 && cfg->precision != SCB_FP32) return SCB_E_INVALID;
  if (cfg->spline_mode != SCB_REF_CUBIC && cfg->spline_mode != SCB_LAU_QUINTIC) return SCB_E_INVALID;
  if (cfg->max_problems < 1 || cfg->max_waypoints < 2 || cfg->max_gridpoints < 2 || cfg->max_samples < 0) return SCB_E_INVALID;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || cfg->device < 0 || cfg->device >= n) { cudaGetLastError(); return SCB_E_NO_DEVICE; }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, cfg->device) != cudaSuccess) return SCB_E_NO_DEVICE;
  if (prop.major != 10) return SCB_E_NO_DEVICE;  // kernels are built for sm_100a only
  scb_ctx* ctx = new (std::nothrow) scb_ctx();
  if (!ctx) return SCB_E_NOMEM;
  ctx->cfg = *cfg;
  if (cudaSetDevice(cfg->device) != cudaSuccess || cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete ctx; return SCB_E_CUDA;
  }
  for (auto& e : ctx->ev) if (cudaEventCreate(&e) != cudaSuccess) { scb_destroy(ctx); return SCB_E_CUDA; }
  if (cudaEventCreateWithFlags(&ctx->ev_last, cudaEventDisableTiming) != cudaSuccess) { scb_destroy(ctx); return SCB_E_CUDA; }
  {
    // Priority hand-over (SCB_PRIO=0 turns it off): the block scheduler offers a freed SM slot to the highest-priority
    // pending work, so with several batches in flight the toppra / parametrizer CTAs of one batch start beside the
    // FP64-bound path / arc kernels of another instead of behind their whole grid.
    const char* e = getenv("SCB_PRIO");
    if (!e || atoi(e) != 0) {
      int least = 0, greatest = 0;
      if (cudaDeviceGetStreamPriorityRange(&least, &greatest) == cudaSuccess && greatest < least &&
          cudaStreamCreateWithPriority(&ctx->hi, cudaStreamNonBlocking, greatest) == cudaSuccess) {
        for (auto& ev : ctx->ev_hi)
          if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) { scb_destroy(ctx); return SCB_E_CUDA; }
      } else {
        cudaGetLastError();
        ctx->hi = nullptr;
      }
    }
  }
  *out = ctx;
  return SCB_OK;
}

int scb_destroy(scb_ctx* ctx) {
  if (!ctx) return SCB_E_INVALID;
  cudaSetDevice(ctx->cfg.device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  if (ctx->last && ctx->last != ctx->stream && ctx->ev_last) cudaEventSynchronize(ctx->ev_last);  // a caller's stream
  release_retired(ctx);
  if (ctx->ev_last) cudaEventDestroy(ctx->ev_last);
  for (auto& s : ctx->slot) if (s.p) cudaFree(s.p);
  for (auto& e : ctx->ev) if (e) cudaEventDestroy(e);
  for (auto& e : ctx->ev_done) if (e) cudaEventDestroy(e);
  for (auto& e : ctx->ev_copied) if (e) cudaEventDestroy(e);
  if (ctx->copy_stream) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamDestroy(ctx->copy_stream); }
  if (ctx->hi) { cudaStreamSynchronize(ctx->hi); cudaStreamDestroy(ctx->hi); }
  for (auto& e : ctx->ev_hi) if (e) cudaEventDestroy(e);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return SCB_OK;
}

int scb_plan_batch_device(scb_ctx* ctx, const scb_batch_in* in, const scb_batch_out* out, int async) {
  int stages;
  int rc = validate(ctx, in, out, stages);
  if (rc) return rc;
  SCB_CUDA(cudaSetDevice(ctx->cfg.device));
  ctx->host_call = false;
  ctx->run = nullptr;
  rc = run_device_any(ctx, in, out, stages);
  if (rc) return rc;
  if (!async) { SCB_CUDA(cudaStreamSynchronize(ctx->stream)); collect_timing(ctx); release_retired(ctx); }
  return SCB_OK;
}

int scb_plan_batch_async(scb_ctx* ctx, const scb_batch_in* in, const scb_batch_out* out, void* stream) {
  int stages;
  int rc = validate(ctx, in, out, stages);
  if (rc) return rc;
  SCB_CUDA(cudaSetDevice(ctx->cfg.device));
  ctx->host_call = false;
  ctx->run = stream ? (cudaStream_t)stream : ctx->stream;
  rc = run_device_any(ctx, in, out, stages);
  ctx->run = nullptr;
  return rc;
}

int scb_synchronize(scb_ctx* ctx) {
  if (!ctx) return SCB_E_INVALID;
  SCB_CUDA(cudaSetDevice(ctx->cfg.device));
  SCB_CUDA(cudaStreamSynchronize(ctx->stream));
  if (ctx->last && ctx->last != ctx->stream) SCB_CUDA(cudaEventSynchronize(ctx->ev_last));  // a caller's stream
  collect_timing(ctx);
  release_retired(ctx);
  return SCB_OK;
}

// Host buffers in, host buffers out, for problems [base, base + Pn) of the caller's arrays (leading dimension
// in->P): scb_plan_batch is the whole range, scb_multi_plan_batch gives every device its block.  A large range is
// planned in chunks of `chunk` problems: the kernels of chunk c+1 run on the ctx stream while the trajectory records
// of chunk c travel to the host on a second stream (two sets of device mirrors, ping-pong), so the call costs the
// D2H time plus ONE chunk of kernels instead of their sum.  Chunks are compact on the device ([G][Pc]); the
// caller's [G][P] arrays are filled by pitched copies.  Problems are independent, so the results do not depend on
// the chunking (tests/test_gpu_parity.py::test_host_chunks_bit_identical).
}  // extern "C"

namespace {

// H2D of the inputs of problems [p0, p0 + Pc) (host arrays of leading dimension P) into the ctx's input mirrors,
// on stream st; din receives the device pointers.  The obstacle edge list is shared: uploaded when `edges_too`.
int upload_inputs(scb_ctx* ctx, const scb_batch_in* in, int stages, int p0, int Pc, cudaStream_t st, bool edges_too,
                  const void** dEdges, scb_batch_in& din) {
  const int P = in->P, W = in->W, G = in->G;
  const size_t rs = real_size(ctx);
  const bool do_spline = stages & SCB_STAGE_SPLINE, do_timing = stages & SCB_STAGE_TIMING;
  int rc;
  din = *in;
  din.P = Pc;
  auto up = [&](int id, const void* src, size_t rows, size_t elem, const void** dst) -> int {
    if (!src) { *dst = nullptr; return SCB_OK; }
    void* d;
    int r = ensure(ctx, id, rows * Pc * elem, &d);
    if (r) return r;
    const size_t w = (size_t)Pc * elem, hp = (size_t)P * elem;
    const char* h = (const char*)src + (size_t)p0 * elem;
    cudaError_t e = (rows == 1 || Pc == P) ? cudaMemcpyAsync(d, h, w * rows, cudaMemcpyHostToDevice, st)
                                           : cudaMemcpy2DAsync(d, w, h, hp, w, rows, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return fail_cuda(ctx, e, "cudaMemcpyAsync H2D");
    *dst = d;
    return SCB_OK;
  };
  if (do_spline) {
    if ((rc = up(SL_IN_WX, in->wx, 1, W * rs, &din.wx))) return rc;
    if ((rc = up(SL_IN_WY, in->wy, 1, W * rs, &din.wy))) return rc;
    if ((rc = up(SL_IN_SA, in->start_angle, 1, rs, &din.start_angle))) return rc;
    if (in->n_obstacle_edges > 0) {  // shared by every problem: uploaded once
      if (edges_too) {
        void* d;
        if ((rc = ensure(ctx, SL_IN_EDGES, (size_t)in->n_obstacle_edges * 4 * rs, &d))) return rc;
        SCB_CUDA(cudaMemcpyAsync(d, in->obstacle_edges, (size_t)in->n_obstacle_edges * 4 * rs, cudaMemcpyHostToDevice, st));
        *dEdges = d;
      }
      din.obstacle_edges = *dEdges;
    }
  }
  if (do_timing) {
    if ((rc = up(SL_IN_VLO, in->vel_lo, 1, rs, &din.vel_lo))) return rc;
    if ((rc = up(SL_IN_VHI, in->vel_hi, 1, rs, &din.vel_hi))) return rc;
    if ((rc = up(SL_IN_VLG, in->vel_lo_grid, G, rs, &din.vel_lo_grid))) return rc;
    if ((rc = up(SL_IN_VHG, in->vel_hi_grid, G, rs, &din.vel_hi_grid))) return rc;
    if ((rc = up(SL_IN_ALO, in->acc_lo, 1, rs, &din.acc_lo))) return rc;
    if ((rc = up(SL_IN_AHI, in->acc_hi, 1, rs, &din.acc_hi))) return rc;
    if (!do_spline && (rc = up(SL_IN_PL, in->path_len, 1, rs, &din.path_len))) return rc;
  }
  // the end slopes of q(s) are read by the arc kernel (gridpoint geometry) as well as by the timing stage:
  // mirrored whenever the caller passes them, whatever the stages
  if ((rc = up(SL_IN_VS, in->vel_start, 1, rs, &din.vel_start))) return rc;
  if ((rc = up(SL_IN_VE, in->vel_end, 1, rs, &din.vel_end))) return rc;
  return SCB_OK;
}

int plan_host_range(scb_ctx* ctx, const scb_batch_in* in, const scb_batch_out* out, int stages, int base, int Pn) {
  SCB_CUDA(cudaSetDevice(ctx->cfg.device));
  const int P = in->P, W = in->W, G = in->G, deg = degree_of(ctx), S = out->S_cap;
  if (Pn == 0) return SCB_OK;
  const size_t rs = real_size(ctx);
  const size_t rr = out->record_f32 ? 4 : rs;  // element size of the record arrays
  ctx->run = nullptr;
  cudaStream_t st = ctx->stream;
  int rc;
  int chunk = Pn;
  if (ctx->host_chunk > 0) chunk = ctx->host_chunk < Pn ? ctx->host_chunk : Pn;
  else if (ctx->host_chunk == 0 && Pn >= 2 * kAutoHostChunk) chunk = kAutoHostChunk;
  const int nchunks = (Pn + chunk - 1) / chunk;
  const bool piped = nchunks > 1;
  if (piped && !ctx->copy_stream) {
    SCB_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (auto& e : ctx->ev_done) SCB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto& e : ctx->ev_copied) SCB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  cudaStream_t cs = piped ? ctx->copy_stream : st;
  const bool do_spline = stages & SCB_STAGE_SPLINE, do_timing = stages & SCB_STAGE_TIMING;
  const size_t nc = (size_t)(W > 1 ? W - 1 : 0) * (deg + 1), nt = (size_t)(W > 1 ? W - 1 : 0) * kTable;
  SCB_CUDA(cudaEventRecord(ctx->ev[0], st));
  const void* dEdges = nullptr;
  for (int c = 0; c < nchunks; ++c) {
    const int p0 = base + c * chunk, Pc = (Pn - c * chunk < chunk) ? Pn - c * chunk : chunk, b = c & 1;
    if (c >= 2) SCB_CUDA(cudaStreamWaitEvent(st, ctx->ev_copied[b], 0));  // mirror set b is free again
    scb_batch_in din;
    scb_batch_out dout = *out;
    // rows x [P] host array <-> rows x [Pc] device array (rows = 1 and a wider element for the [P][n] arrays)
    auto copy = [&](void* dst, const void* src, size_t rows, size_t elem, bool to_host, cudaStream_t s) -> cudaError_t {
      const size_t w = (size_t)Pc * elem, hp = (size_t)P * elem;
      const cudaMemcpyKind k = to_host ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice;
      if (rows == 1 || Pc == P) return cudaMemcpyAsync(dst, src, w * rows, k, s);
      return to_host ? cudaMemcpy2DAsync(dst, hp, src, w, w, rows, k, s) : cudaMemcpy2DAsync(dst, w, src, hp, w, rows, k, s);
    };
    // ---- H2D of this chunk's inputs, on the compute stream (full speed when the caller's memory is pinned)
    if ((rc = upload_inputs(ctx, in, stages, p0, Pc, st, c == 0, &dEdges, din))) return rc;
    auto up = [&](int id, const void* src, size_t rows, size_t elem, const void** dst) -> int {
      if (!src) { *dst = nullptr; return SCB_OK; }
      void* d;
      int r = ensure(ctx, id, rows * Pc * elem, &d);
      if (r) return r;
      cudaError_t e = copy(d, (const char*)src + (size_t)p0 * elem, rows, elem, false, st);
      if (e != cudaSuccess) return fail_cuda(ctx, e, "cudaMemcpyAsync H2D");
      *dst = d;
      return SCB_OK;
    };
    // ---- device mirrors (set b) of the requested outputs
    struct Down { void* host; void* dev; size_t rows, elem; };
    Down downs[32];
    int nd = 0;
    auto mirror = [&](int id, void* host, size_t rows, size_t elem, void** dev) -> int {
      if (!host) { *dev = nullptr; return SCB_OK; }
      void* d;
      int r = ensure(ctx, id + b * SL_COUNT, rows * Pc * elem, &d);
      if (r) return r;
      *dev = d;
      downs[nd++] = {(char*)host + (size_t)p0 * elem, d, rows, elem};
      return SCB_OK;
    };
#define MIR(id, field, rows, elem) if ((rc = mirror(id, out->field, rows, elem, (void**)&dout.field))) return rc
    MIR(SL_O_STATUS, status, 1, 4); MIR(SL_O_FAIL, fail_stage, 1, 4);
    if (do_spline) {
      MIR(SL_O_CTRL_X, ctrl_x, 1, nc * rs); MIR(SL_O_CTRL_Y, ctrl_y, 1, nc * rs);
      MIR(SL_O_L, arclen_total, 1, rs); MIR(SL_O_TABLE, arclen_table, 1, nt * rs); MIR(SL_O_COST, cost, 1, rs);
    } else if (stages & SCB_STAGE_GEOMETRY) {  // the spline is an input
      dout.arclen_total = dout.cost = nullptr;
      if ((rc = up(SL_IN_CX, out->ctrl_x, 1, nc * rs, (const void**)&dout.ctrl_x))) return rc;
      if ((rc = up(SL_IN_CY, out->ctrl_y, 1, nc * rs, (const void**)&dout.ctrl_y))) return rc;
      if ((rc = up(SL_IN_TAB, out->arclen_table, 1, nt * rs, (const void**)&dout.arclen_table))) return rc;
    } else { dout.ctrl_x = dout.ctrl_y = dout.arclen_total = dout.arclen_table = dout.cost = nullptr; }
    if (do_timing) {
      MIR(SL_O_T_TOTAL, T_total, 1, rs);
      MIR(SL_O_K_LO, K_lo, G, rs); MIR(SL_O_K_HI, K_hi, G, rs); MIR(SL_O_X, x, G, rs); MIR(SL_O_T, t, G, rr);
      MIR(SL_O_V, v, G, rr); MIR(SL_O_A, a, G, rr);
      if (stages & SCB_STAGE_GEOMETRY) { MIR(SL_O_PX, px, G, rr); MIR(SL_O_PY, py, G, rr); MIR(SL_O_OMEGA, omega, G, rr); }
      else { dout.px = dout.py = dout.omega = nullptr; }
    } else { dout.T_total = dout.K_lo = dout.K_hi = dout.x = dout.t = dout.v = dout.a = dout.px = dout.py = dout.omega = nullptr; }
    if (S > 0) {
      MIR(SL_O_NS, n_samples, 1, 4);
      MIR(SL_O_STIME, s_time, S, rr); MIR(SL_O_SPOS, s_pos, S, rr); MIR(SL_O_SVEL, s_vel, S, rr); MIR(SL_O_SACC, s_acc, S, rr);
      MIR(SL_O_SPX, s_px, S, rr); MIR(SL_O_SPY, s_py, S, rr); MIR(SL_O_SOMEGA, s_omega, S, rr);
      if (!do_timing) {  // positions (and velocities) to resample at are inputs
        SCB_CUDA(copy(dout.n_samples, (const char*)out->n_samples + (size_t)p0 * 4, 1, 4, false, st));
        SCB_CUDA(copy(dout.s_pos, (const char*)out->s_pos + (size_t)p0 * rr, S, rr, false, st));
        if (out->s_vel) SCB_CUDA(copy(dout.s_vel, (const char*)out->s_vel + (size_t)p0 * rr, S, rr, false, st));
      }
    }
#undef MIR
    ctx->host_call = true;
    rc = run_device_any(ctx, &din, &dout, stages);
    if (rc) return rc;
    if (piped) {
      SCB_CUDA(cudaEventRecord(ctx->ev_done[b], st));
      SCB_CUDA(cudaStreamWaitEvent(cs, ctx->ev_done[b], 0));
      if (c == 0) SCB_CUDA(cudaEventRecord(ctx->ev[8], cs));
    }
    for (int i = 0; i < nd; ++i) SCB_CUDA(copy(downs[i].host, downs[i].dev, downs[i].rows, downs[i].elem, true, cs));
    if (piped) SCB_CUDA(cudaEventRecord(ctx->ev_copied[b], cs));
  }
  if (piped) {
    SCB_CUDA(cudaStreamWaitEvent(st, ctx->ev_copied[0], 0));
    SCB_CUDA(cudaStreamWaitEvent(st, ctx->ev_copied[1], 0));
  }
  ctx->piped = piped;
  ctx->timing.launches *= nchunks;
  SCB_CUDA(cudaEventRecord(ctx->ev[6], st));
  SCB_CUDA(cudaEventRecord(ctx->ev_last, st));
  SCB_CUDA(cudaStreamSynchronize(st));
  collect_timing(ctx);
  release_retired(ctx);
  return SCB_OK;
}

// ---- rank, then fetch -------------------------------------------------------------------------------------------
struct GatherJob { const void* src; void* dst; int rows, per, words; };  // rows > 0: [rows][P] columns; else [P][per] rows
struct GatherArgs { GatherJob job[20]; int njobs, P, n; const int32_t* ids; };

// dst[i][j] = src[i][ids[j]]  (column gather of a [rows][P] array)  or  dst[j][:] = src[ids[j]][:]  (row gather of a
// [P][per] array); elements are `words` 32-bit words.  blockIdx.y = job.
__global__ void gather_kernel(GatherArgs a) {
  const GatherJob jb = a.job[blockIdx.y];
  const uint32_t* __restrict__ src = (const uint32_t*)jb.src;
  uint32_t* __restrict__ dst = (uint32_t*)jb.dst;
  const size_t w = (size_t)jb.words;
  if (jb.rows > 0) {
    const size_t total = (size_t)jb.rows * a.n * w;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
      const size_t k = e % w, ij = e / w, j = ij % a.n, i = ij / a.n;
      dst[e] = src[((size_t)i * a.P + a.ids[j]) * w + k];
    }
  } else {
    const size_t row = (size_t)jb.per * w, total = row * a.n;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
      const size_t j = e / row, k = e % row;
      dst[e] = src[(size_t)a.ids[j] * row + k];
    }
  }
}

// Plan problems [base, base + Pn) of the caller's host inputs, keep every output in the ctx, copy the summaries back.
int plan_summary_range(scb_ctx* ctx, const scb_batch_in* in, int rec32, int base, int Pn, scb_summary* summaries) {
  SCB_CUDA(cudaSetDevice(ctx->cfg.device));
  ctx->kept_P = 0;
  if (Pn == 0) return SCB_OK;
  const int W = in->W, G = in->G, deg = degree_of(ctx);
  const size_t rs = real_size(ctx), rr = rec32 ? 4 : rs;
  const int stages = SCB_STAGE_SPLINE | SCB_STAGE_TIMING | SCB_STAGE_GEOMETRY;
  ctx->run = nullptr;
  cudaStream_t st = ctx->stream;
  int rc;
  SCB_CUDA(cudaEventRecord(ctx->ev[0], st));
  scb_batch_in din;
  const void* dEdges = nullptr;
  if ((rc = upload_inputs(ctx, in, stages, base, Pn, st, true, &dEdges, din))) return rc;
  scb_batch_out k = {};
  const size_t nc = (size_t)(W - 1) * (deg + 1), nt = (size_t)(W - 1) * kTable, gp = (size_t)G * Pn;
#define KEEP(id, field, bytes) if ((rc = ensure(ctx, id, (bytes), (void**)&k.field))) return rc
  KEEP(SL_R_STATUS, status, (size_t)Pn * 4); KEEP(SL_R_FAIL, fail_stage, (size_t)Pn * 4);
  KEEP(SL_R_CTRL_X, ctrl_x, Pn * nc * rs); KEEP(SL_R_CTRL_Y, ctrl_y, Pn * nc * rs);
  KEEP(SL_R_L, arclen_total, Pn * rs); KEEP(SL_R_TABLE, arclen_table, Pn * nt * rs);
  KEEP(SL_R_T_TOTAL, T_total, Pn * rs); KEEP(SL_R_COST, cost, Pn * rs);
  KEEP(SL_R_K_LO, K_lo, gp * rs); KEEP(SL_R_K_HI, K_hi, gp * rs); KEEP(SL_R_X, x, gp * rs);
  KEEP(SL_R_T, t, gp * rr); KEEP(SL_R_PX, px, gp * rr); KEEP(SL_R_PY, py, gp * rr);
  KEEP(SL_R_V, v, gp * rr); KEEP(SL_R_A, a, gp * rr); KEEP(SL_R_OMEGA, omega, gp * rr);
#undef KEEP
  k.record_f32 = rec32 ? 1 : 0;
  ctx->host_call = true;
  if ((rc = run_device_any(ctx, &din, &k, stages))) return rc;
  scb_summary* dsum;
  if ((rc = ensure(ctx, SL_R_SUMMARY, (size_t)Pn * sizeof(scb_summary), (void**)&dsum))) return rc;
  if ((rc = scb_pack_summary_device(ctx, Pn, k.status, k.fail_stage, k.T_total, k.arclen_total, k.cost, dsum))) return rc;
  SCB_CUDA(cudaMemcpyAsync(summaries, dsum, (size_t)Pn * sizeof(scb_summary), cudaMemcpyDeviceToHost, st));
  ctx->piped = false;
  ctx->timing.launches += 1;
  SCB_CUDA(cudaEventRecord(ctx->ev[6], st));
  SCB_CUDA(cudaEventRecord(ctx->ev_last, st));
  SCB_CUDA(cudaStreamSynchronize(st));
  collect_timing(ctx);
  release_retired(ctx);
  ctx->kept = k; ctx->kept_P = Pn; ctx->kept_W = W; ctx->kept_G = G; ctx->kept_rec32 = rec32 ? 1 : 0;
  return SCB_OK;
}

// Gather problems ids[0..n) of the retained results into host arrays of leading dimension ld, columns [col0, col0 + n).
int fetch_cols(scb_ctx* ctx, int n, const int32_t* ids, const scb_batch_out* out, int col0, int ld) {
  if (n == 0) return SCB_OK;
  SCB_CUDA(cudaSetDevice(ctx->cfg.device));
  const int P = ctx->kept_P, W = ctx->kept_W, G = ctx->kept_G, deg = degree_of(ctx);
  if (P <= 0) return SCB_E_INVALID;
  for (int j = 0; j < n; ++j) if (ids[j] < 0 || ids[j] >= P) return SCB_E_INVALID;
  const size_t rs = real_size(ctx), rr = ctx->kept_rec32 ? 4 : rs;
  const size_t nc = (size_t)(W - 1) * (deg + 1), nt = (size_t)(W - 1) * kTable;
  cudaStream_t st = ctx->stream;
  int rc;
  int32_t* dIds;
  if ((rc = ensure(ctx, SL_F_IDS, (size_t)n * 4, (void**)&dIds))) return rc;
  SCB_CUDA(cudaMemcpyAsync(dIds, ids, (size_t)n * 4, cudaMemcpyHostToDevice, st));
  const scb_batch_out& k = ctx->kept;
  struct Item { void* host; const void* dev; int rows, per; size_t elem; };
  const Item items[] = {
      {out->status, k.status, 0, 1, 4}, {out->fail_stage, k.fail_stage, 0, 1, 4},
      {out->ctrl_x, k.ctrl_x, 0, (int)nc, rs}, {out->ctrl_y, k.ctrl_y, 0, (int)nc, rs},
      {out->arclen_total, k.arclen_total, 0, 1, rs}, {out->arclen_table, k.arclen_table, 0, (int)nt, rs},
      {out->T_total, k.T_total, 0, 1, rs}, {out->cost, k.cost, 0, 1, rs},
      {out->K_lo, k.K_lo, G, 1, rs}, {out->K_hi, k.K_hi, G, 1, rs}, {out->x, k.x, G, 1, rs},
      {out->t, k.t, G, 1, rr}, {out->px, k.px, G, 1, rr}, {out->py, k.py, G, 1, rr},
      {out->v, k.v, G, 1, rr}, {out->a, k.a, G, 1, rr}, {out->omega, k.omega, G, 1, rr}};
  size_t total = 0;
  for (const Item& it : items) if (it.host) total += ((size_t)(it.rows > 0 ? it.rows : it.per) * n * it.elem + 15) / 16 * 16;
  char* stage;
  if ((rc = ensure(ctx, SL_F_STAGE, total, (void**)&stage))) return rc;
  GatherArgs ga{};
  ga.P = P; ga.n = n; ga.ids = dIds;
  size_t off = 0, offs[20];
  size_t most = 0;
  for (const Item& it : items) {
    if (!it.host) continue;
    const size_t bytes = (size_t)(it.rows > 0 ? it.rows : it.per) * n * it.elem;
    offs[ga.njobs] = off;
    ga.job[ga.njobs++] = {it.dev, stage + off, it.rows, it.per, (int)(it.elem / 4)};
    off += (bytes + 15) / 16 * 16;
    most = bytes / 4 > most ? bytes / 4 : most;
  }
  if (ga.njobs == 0) return SCB_OK;
  SCB_CUDA(cudaEventRecord(ctx->ev[0], st));
  const int gx = (int)((most + 255) / 256 < 296 ? (most + 255) / 256 : 296);
  gather_kernel<<<dim3(gx ? gx : 1, ga.njobs), 256, 0, st>>>(ga);
  SCB_CUDA(cudaGetLastError());
  int j = 0;
  for (const Item& it : items) {
    if (!it.host) continue;
    const char* src = stage + offs[j++];
    if (it.rows > 0) {  // [rows][n] -> columns [col0, col0 + n) of [rows][ld]
      const size_t w = (size_t)n * it.elem;
      if (ld == n) SCB_CUDA(cudaMemcpyAsync(it.host, src, w * it.rows, cudaMemcpyDeviceToHost, st));
      else SCB_CUDA(cudaMemcpy2DAsync((char*)it.host + (size_t)col0 * it.elem, (size_t)ld * it.elem, src, w, w, it.rows, cudaMemcpyDeviceToHost, st));
    } else {
      const size_t row = (size_t)it.per * it.elem;
      SCB_CUDA(cudaMemcpyAsync((char*)it.host + (size_t)col0 * row, src, row * n, cudaMemcpyDeviceToHost, st));
    }
  }
  SCB_CUDA(cudaEventRecord(ctx->ev[6], st));
  SCB_CUDA(cudaStreamSynchronize(st));
  scb_timing& t = ctx->timing;
  float ms = 0;
  cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[6]);
  t.d2h_ms = ms;  // the plan's stage times stay; d2h_ms = this fetch
  return SCB_OK;
}

int validate_summary(scb_ctx* ctx, const scb_batch_in* in) {
  if (!ctx || !in) return SCB_E_INVALID;
  int32_t dummy = 0;
  scb_batch_out o{};
  o.status = &dummy;
  int stages;
  if (in->stages != 0 && in->stages != (SCB_STAGE_SPLINE | SCB_STAGE_TIMING | SCB_STAGE_GEOMETRY)) return SCB_E_INVALID;
  return validate(ctx, in, &o, stages);
}

}  // namespace

struct scb_multi {
  std::vector<scb_ctx*> ctx;
  int kept_P = 0;  // problems of the last scb_multi_plan_summary
};

namespace {
// one host thread per device, as one scb_ctx per host thread is the ABI's rule
template <class F> int on_every_device(scb_multi* m, F&& body) {
  const int R = (int)m->ctx.size();
  std::vector<int> rcs((size_t)R, SCB_OK);
  std::vector<std::thread> th;
  for (int r = 1; r < R; ++r) th.emplace_back([&, r] { rcs[(size_t)r] = body(r); });
  rcs[0] = body(0);
  for (auto& t : th) t.join();
  for (int rc : rcs) if (rc) return rc;
  return SCB_OK;
}

}  // namespace

extern "C" {

int scb_plan_batch(scb_ctx* ctx, const scb_batch_in* in, const scb_batch_out* out) {
  int stages;
  int rc = validate(ctx, in, out, stages);
  if (rc) return rc;
  return plan_host_range(ctx, in, out, stages, 0, in->P);
}

int scb_plan_batch_summary(scb_ctx* ctx, const scb_batch_in* in, int32_t record_f32, scb_summary* summaries) {
  int rc = validate_summary(ctx, in);
  if (rc) return rc;
  if (!summaries) return SCB_E_INVALID;
  return plan_summary_range(ctx, in, record_f32, 0, in->P, summaries);
}

int scb_fetch(scb_ctx* ctx, int32_t n, const int32_t* problem_ids, const scb_batch_out* out) {
  if (!ctx || n < 0 || (n > 0 && !problem_ids) || !out) return SCB_E_INVALID;
  return fetch_cols(ctx, n, problem_ids, out, 0, n);
}

void scb_shard_range(int32_t P, int32_t r, int32_t R, int32_t* lo, int32_t* hi) {
  const int64_t per = R > 0 ? ((int64_t)P + R - 1) / R : P;
  int64_t a = (int64_t)r * per, b;
  if (a > P) a = P;
  b = a + per;
  if (b > P) b = P;
  if (lo) *lo = (int32_t)a;
  if (hi) *hi = (int32_t)b;
}

int scb_multi_create(scb_multi** out, const scb_config* cfg, int32_t n_devices, const int32_t* devices) {
  if (!out || !cfg || n_devices < 1) return SCB_E_INVALID;
  *out = nullptr;
  scb_multi* m = new (std::nothrow) scb_multi();
  if (!m) return SCB_E_NOMEM;
  for (int r = 0; r < n_devices; ++r) {
    scb_config c = *cfg;
    c.device = devices ? devices[r] : r;
    scb_ctx* x = nullptr;
    const int rc = scb_create(&x, &c);
    if (rc) { scb_multi_destroy(m); return rc; }
    m->ctx.push_back(x);
  }
  *out = m;
  return SCB_OK;
}

int scb_multi_destroy(scb_multi* m) {
  if (!m) return SCB_E_INVALID;
  for (scb_ctx* x : m->ctx) scb_destroy(x);
  delete m;
  return SCB_OK;
}

int32_t scb_multi_devices(const scb_multi* m) { return m ? (int32_t)m->ctx.size() : 0; }

int scb_multi_get_timing(scb_multi* m, int32_t r, scb_timing* out) {
  if (!m || r < 0 || r >= (int32_t)m->ctx.size()) return SCB_E_INVALID;
  return scb_get_timing(m->ctx[(size_t)r], out);
}

int scb_multi_plan_batch(scb_multi* m, const scb_batch_in* in, const scb_batch_out* out) {
  if (!m || m->ctx.empty() || !in || !out) return SCB_E_INVALID;
  const int R = (int)m->ctx.size();
  int stages = 0;
  for (int r = 0; r < R; ++r) {  // every block must fit its device's capacity
    int32_t lo, hi;
    scb_shard_range(in->P, r, R, &lo, &hi);
    scb_batch_in probe = *in;
    probe.P = hi - lo;
    const int rc = validate(m->ctx[(size_t)r], &probe, out, stages);
    if (rc) return rc;
  }
  return on_every_device(m, [&](int r) {
    int32_t lo, hi;
    scb_shard_range(in->P, r, R, &lo, &hi);
    return plan_host_range(m->ctx[(size_t)r], in, out, stages, lo, hi - lo);
  });
}

int scb_multi_plan_summary(scb_multi* m, const scb_batch_in* in, int32_t record_f32, scb_summary* summaries) {
  if (!m || m->ctx.empty() || !in || !summaries) return SCB_E_INVALID;
  const int R = (int)m->ctx.size();
  m->kept_P = 0;
  for (int r = 0; r < R; ++r) {
    int32_t lo, hi;
    scb_shard_range(in->P, r, R, &lo, &hi);
    scb_batch_in probe = *in;
    probe.P = hi - lo;
    const int rc = validate_summary(m->ctx[(size_t)r], &probe);
    if (rc) return rc;
  }
  const int rc = on_every_device(m, [&](int r) {
    int32_t lo, hi;
    scb_shard_range(in->P, r, R, &lo, &hi);
    return plan_summary_range(m->ctx[(size_t)r], in, record_f32, lo, hi - lo, summaries + lo);
  });
  if (rc == SCB_OK) m->kept_P = in->P;
  return rc;
}

int scb_multi_fetch(scb_multi* m, int32_t n, const int32_t* problem_ids, const scb_batch_out* out) {
  if (!m || m->ctx.empty() || n < 0 || (n > 0 && !problem_ids) || !out || m->kept_P <= 0) return SCB_E_INVALID;
  const int R = (int)m->ctx.size(), P = m->kept_P;
  // ids of one device are fetched into consecutive columns: the caller's order is restored by asking for them in
  // runs; an MPC loop asks for a handful of ids, so every run of consecutive same-owner ids is one small fetch
  for (int j = 0; j < n; ++j) if (problem_ids[j] < 0 || problem_ids[j] >= P) return SCB_E_INVALID;
  int j0 = 0;
  while (j0 < n) {
    int32_t lo = 0, hi = 0;
    int owner = 0;
    for (owner = 0; owner < R; ++owner) { scb_shard_range(P, owner, R, &lo, &hi); if (problem_ids[j0] < hi) break; }
    int j1 = j0;
    std::vector<int32_t> local;
    while (j1 < n && problem_ids[j1] >= lo && problem_ids[j1] < hi) local.push_back(problem_ids[j1++] - lo);
    const int rc = fetch_cols(m->ctx[(size_t)owner], j1 - j0, local.data(), out, j0, n);
    if (rc) return rc;
    j0 = j1;
  }
  return SCB_OK;
}

int scb_set_host_chunk(scb_ctx* ctx, int32_t problems) {
  if (!ctx) return SCB_E_INVALID;
  ctx->host_chunk = problems;
  return SCB_OK;
}

int scb_pack_summary_device(scb_ctx* ctx, int32_t P, const int32_t* status, const int32_t* fail_stage, const void* T_total,
                            const void* arclen_total, const void* cost, scb_summary* dst) {
  if (!ctx || P < 0 || !status || !dst) return SCB_E_INVALID;
  if (P == 0) return SCB_OK;
  SCB_CUDA(cudaSetDevice(ctx->cfg.device));
  const int block = 256, grid = (P + block - 1) / block;
  if (ctx->cfg.precision == SCB_FP64)
    pack_summary_kernel<double><<<grid, block, 0, ctx->stream>>>(P, status, fail_stage, (const double*)T_total,
                                                                 (const double*)arclen_total, (const double*)cost, dst);
  else
    pack_summary_kernel<float><<<grid, block, 0, ctx->stream>>>(P, status, fail_stage, (const float*)T_total,
                                                                (const float*)arclen_total, (const float*)cost, dst);
  SCB_CUDA(cudaGetLastError());
  return SCB_OK;
}

// FMT* over a shared node set, one CTA per query (fmt_kernel.cu).  Inputs are packed into one device block
// (nodes, edges, queries), outputs into another; both are a few hundred KB, so the call is one H2D, one launch,
// one D2H per output array.  total_ms of scb_timing = the call, spline_ms = the kernel.
int scb_fmt_batch(scb_ctx* ctx, const scb_fmt_in* in, const scb_fmt_out* out) {
  if (!ctx || !in || !out) return SCB_E_INVALID;
  const int Q = in->Q, n = in->n_nodes, E = in->n_edges, M = in->max_path;
  if (Q < 0 || n < 0 || n > 8192 || E < 0 || M < 0) return SCB_E_INVALID;
  if (n > 0 && (!in->node_x || !in->node_y)) return SCB_E_INVALID;
  if (E > 0 && !in->edges) return SCB_E_INVALID;
  if (!(in->rn >= 0)) return SCB_E_INVALID;
  if (Q == 0) return SCB_OK;
  if (!in->init_x || !in->init_y || !in->goal_x || !in->goal_y || !out->status || !out->path_len || !out->cost) return SCB_E_INVALID;
  if ((out->path_x == nullptr) != (out->path_y == nullptr) || (out->path_x && M < 1)) return SCB_E_INVALID;
  SCB_CUDA(cudaSetDevice(ctx->cfg.device));
  cudaStream_t st = ctx->stream;
  const size_t n_in = 2 * (size_t)n + 4 * (size_t)E + 4 * (size_t)Q;
  const size_t n_out = 4 * (size_t)Q + (out->path_x ? 2 * (size_t)Q * M : 0);
  float *din, *dout;
  int rc;
  if ((rc = ensure(ctx, SL_FMT_IN, n_in * 4, (void**)&din))) return rc;
  if ((rc = ensure(ctx, SL_FMT_OUT, n_out * 4, (void**)&dout))) return rc;
  scb::FmtArgs a{};
  a.Q = Q; a.n = n; a.E = E; a.max_path = M; a.rn = in->rn;
  float* p = din;
  auto up = [&](const float* src, size_t cnt, const float** dst) -> cudaError_t {
    *dst = p;
    cudaError_t e = cnt ? cudaMemcpyAsync(p, src, cnt * 4, cudaMemcpyHostToDevice, st) : cudaSuccess;
    p += cnt;
    return e;
  };
  SCB_CUDA(cudaEventRecord(ctx->ev[0], st));
  SCB_CUDA(up(in->node_x, n, &a.nx)); SCB_CUDA(up(in->node_y, n, &a.ny)); SCB_CUDA(up(in->edges, 4 * (size_t)E, &a.edges));
  SCB_CUDA(up(in->init_x, Q, &a.ix)); SCB_CUDA(up(in->init_y, Q, &a.iy));
  SCB_CUDA(up(in->goal_x, Q, &a.gx)); SCB_CUDA(up(in->goal_y, Q, &a.gy));
  a.status = (int32_t*)dout; a.path_len = a.status + Q; a.closed = a.path_len + Q; a.cost = dout + 3 * (size_t)Q;
  if (out->path_x) { a.path_x = dout + 4 * (size_t)Q; a.path_y = a.path_x + (size_t)Q * M; }
  SCB_CUDA(cudaEventRecord(ctx->ev[1], st));
  SCB_CUDA(scb::launch_fmt(a, st));
  SCB_CUDA(cudaEventRecord(ctx->ev[2], st));
  SCB_CUDA(cudaMemcpyAsync(out->status, a.status, (size_t)Q * 4, cudaMemcpyDeviceToHost, st));
  SCB_CUDA(cudaMemcpyAsync(out->path_len, a.path_len, (size_t)Q * 4, cudaMemcpyDeviceToHost, st));
  if (out->closed) SCB_CUDA(cudaMemcpyAsync(out->closed, a.closed, (size_t)Q * 4, cudaMemcpyDeviceToHost, st));
  SCB_CUDA(cudaMemcpyAsync(out->cost, a.cost, (size_t)Q * 4, cudaMemcpyDeviceToHost, st));
  if (out->path_x) {
    SCB_CUDA(cudaMemcpyAsync(out->path_x, a.path_x, (size_t)Q * M * 4, cudaMemcpyDeviceToHost, st));
    SCB_CUDA(cudaMemcpyAsync(out->path_y, a.path_y, (size_t)Q * M * 4, cudaMemcpyDeviceToHost, st));
  }
  SCB_CUDA(cudaEventRecord(ctx->ev[6], st));
  SCB_CUDA(cudaStreamSynchronize(st));
  scb_timing& t = ctx->timing;
  t = scb_timing{};
  cudaEventElapsedTime(&t.h2d_ms, ctx->ev[0], ctx->ev[1]);
  cudaEventElapsedTime(&t.spline_ms, ctx->ev[1], ctx->ev[2]);
  cudaEventElapsedTime(&t.d2h_ms, ctx->ev[2], ctx->ev[6]);
  cudaEventElapsedTime(&t.total_ms, ctx->ev[0], ctx->ev[6]);
  t.launches = 1;
  ctx->timing_pending = false;
  return SCB_OK;
}

// toppra on a caller-supplied piecewise-cubic path (toppra_kernel.cu: toppra_ppoly_kernel).  One H2D block, one launch,
// one D2H block.
int scb_toppra_ppoly(scb_ctx* ctx, const scb_ppoly_in* in, const scb_ppoly_out* out) {
  if (!ctx || !in || !out) return SCB_E_INVALID;
  const int P = in->P, ns = in->n_seg, m = in->dof, G = in->G;
  if (P < 0 || ns < 1 || ns > 64 || G < 2 || m < 1 || m > SCB_MAX_DOF) return SCB_E_INVALID;
  if (P == 0) return SCB_OK;
  const bool vgrid = in->vel_lo_grid && in->vel_hi_grid;
  if (!in->coeffs || !in->breaks || !(vgrid || (in->vel_lo && in->vel_hi)) || !in->acc_lo || !in->acc_hi || !out->status || !out->K_lo || !out->K_hi)
    return SCB_E_INVALID;
  for (int k = 0; k < ns; ++k) if (!(in->breaks[k] < in->breaks[k + 1])) return SCB_E_INVALID;
  SCB_CUDA(cudaSetDevice(ctx->cfg.device));
  cudaStream_t st = ctx->stream;
  const size_t nco = (size_t)P * ns * 4 * m, nl = (size_t)P * m, gp = (size_t)G * P, ng = vgrid ? gp * m : 0;
  const size_t n_in = nco + (ns + 1) + 4 * nl + 2 * ng;
  double* din; char* dout;
  int rc;
  if ((rc = ensure(ctx, SL_PP_IN, n_in * 8, (void**)&din))) return rc;
  if ((rc = ensure(ctx, SL_PP_OUT, 5 * gp * 8 + 3 * (size_t)P * 4 + 64, (void**)&dout))) return rc;
  scb::PpolyArgs a{};
  a.P = P; a.n_seg = ns; a.dof = m; a.G = G; a.interpolation = in->interpolation;
  a.sd_start = in->sd_start; a.sd_end = in->sd_end; a.x_cap = in->x_cap > 0 ? in->x_cap : 100.0;
  double* q = din;
  auto up = [&](const double* src, size_t n, const double** dst) -> cudaError_t { *dst = q; cudaError_t e = cudaMemcpyAsync(q, src, n * 8, cudaMemcpyHostToDevice, st); q += n; return e; };
  SCB_CUDA(up(in->coeffs, nco, &a.coeffs)); SCB_CUDA(up(in->breaks, ns + 1, &a.breaks));
  if (vgrid) { SCB_CUDA(up(in->vel_lo_grid, ng, &a.vel_lo_grid)); SCB_CUDA(up(in->vel_hi_grid, ng, &a.vel_hi_grid)); }
  else { SCB_CUDA(up(in->vel_lo, nl, &a.vel_lo)); SCB_CUDA(up(in->vel_hi, nl, &a.vel_hi)); }
  SCB_CUDA(up(in->acc_lo, nl, &a.acc_lo)); SCB_CUDA(up(in->acc_hi, nl, &a.acc_hi));
  double* o = (double*)dout;
  a.K_lo = o; a.K_hi = o + gp; a.x = o + 2 * gp;
  a.feas_lo = (out->feas_lo || out->feas_hi) ? o + 3 * gp : nullptr; a.feas_hi = a.feas_lo ? o + 4 * gp : nullptr;
  int32_t* oi = (int32_t*)(o + 5 * gp);
  a.status = oi; a.fail_stage = oi + P; a.feas_status = oi + 2 * (size_t)P;
  scb::launch_toppra_ppoly(a, st);
  SCB_CUDA(cudaGetLastError());
  auto down = [&](void* dst, const void* src, size_t bytes) -> cudaError_t { return dst ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st) : cudaSuccess; };
  SCB_CUDA(down(out->status, a.status, (size_t)P * 4)); SCB_CUDA(down(out->fail_stage, a.fail_stage, (size_t)P * 4));
  SCB_CUDA(down(out->feas_status, a.feas_status, (size_t)P * 4));
  SCB_CUDA(down(out->K_lo, a.K_lo, gp * 8)); SCB_CUDA(down(out->K_hi, a.K_hi, gp * 8)); SCB_CUDA(down(out->x, a.x, gp * 8));
  if (a.feas_lo) { SCB_CUDA(down(out->feas_lo, a.feas_lo, gp * 8)); SCB_CUDA(down(out->feas_hi, a.feas_hi, gp * 8)); }
  SCB_CUDA(cudaStreamSynchronize(st));
  release_retired(ctx);
  return SCB_OK;
}

// gen_vel_prof<N> (sc:1191-1265) for a batch: Hermite coefficients on the host (tp:geometric_path/
// piecewise_poly_path.cpp:364-375, the reference's order of operations), toppra_ppoly_kernel<N, interpolation> for the
// LP passes, then the timing stage's parametrizer and sampler once per degree of freedom on the shared sdot^2: degree
// of freedom k is the 1-DOF Hermite path 0 -> pos_end[k] - pos_start[k] with its own end slopes (the offset
// pos_start[k] is added to the record on the way out).
int scb_gen_vel_prof(scb_ctx* ctx, const scb_ndof_in* in, const scb_ndof_out* out) {
  if (!ctx || !in || !out) return SCB_E_INVALID;
  const int P = in->P, m = in->dof, G = in->G, S = out->S_cap;
  if (P < 0 || P > ctx->cfg.max_problems || G < 2 || G > ctx->cfg.max_gridpoints || m < 1 || m > SCB_MAX_DOF) return SCB_E_INVALID;
  if (S < 0 || S > ctx->cfg.max_samples) return SCB_E_INVALID;
  if (in->param_mode != SCB_PARAM_SPLINE && in->param_mode != SCB_PARAM_CONST_ACCEL) return SCB_E_INVALID;
  if (P == 0) return SCB_OK;
  const bool vgrid = in->vel_lo_grid && in->vel_hi_grid;
  if (!in->pos_start || !in->pos_end || !(vgrid || (in->vel_lo && in->vel_hi)) || !in->acc_lo || !in->acc_hi || !out->status)
    return SCB_E_INVALID;
  SCB_CUDA(cudaSetDevice(ctx->cfg.device));
  cudaStream_t st = ctx->stream;
  const size_t nl = (size_t)P * m, gp = (size_t)G * P, ng = vgrid ? gp * m : 0, sp = (size_t)S * P;
  // ---- host staging: coefficients [P][1][4][m], breaks, and the per-dof arrays of the parametrizer [m][P]
  std::vector<double> stage(4 * nl + 2 + 3 * nl);
  double* co = stage.data(); double* br = co + 4 * nl; double* Lk = br + 2; double* vsk = Lk + nl; double* vek = vsk + nl;
  br[0] = 0.0; br[1] = 1.0;
  for (int p = 0; p < P; ++p)
    for (int k = 0; k < m; ++k) {
      const double p0 = in->pos_start[(size_t)p * m + k], p1 = in->pos_end[(size_t)p * m + k];
      const double v0 = in->vel_start ? in->vel_start[(size_t)p * m + k] : 0.0, v1 = in->vel_end ? in->vel_end[(size_t)p * m + k] : 0.0;
      const double dt = 1.0;
      double* c = co + (size_t)p * 4 * m + k;  // rows 0..3 at stride m
      c[3 * m] = p0; c[2 * m] = v0;
      c[0] = (v1 * dt - 2 * p1 + c[2 * m] * dt + 2 * c[3 * m]) / (dt * dt * dt);
      c[m] = (v1 - c[2 * m] - 3 * c[0] * dt * dt) / (2 * dt);
      Lk[(size_t)k * P + p] = p1 - p0; vsk[(size_t)k * P + p] = v0; vek[(size_t)k * P + p] = v1;
    }
  const size_t n_in = stage.size() + 4 * nl + 2 * ng;
  double* din; char* dout;
  int rc;
  if ((rc = ensure(ctx, SL_PP_IN, n_in * 8, (void**)&din))) return rc;
  const bool spline = in->param_mode == SCB_PARAM_SPLINE;
  const size_t n_out = 3 * gp + (spline ? 3 : 1) * gp + (size_t)P + sp + 3 * sp * m;   // K_lo K_hi x | t X c' | T | s_time | s_pos s_vel s_acc
  if ((rc = ensure(ctx, SL_PP_OUT, n_out * 8 + 3 * (size_t)P * 4 + 64, (void**)&dout))) return rc;
  scb::PpolyArgs a{};
  a.P = P; a.n_seg = 1; a.dof = m; a.G = G; a.interpolation = in->interpolation;
  a.sd_start = 0.0; a.sd_end = 0.0; a.x_cap = in->x_cap > 0 ? in->x_cap : 100.0;
  SCB_CUDA(cudaMemcpyAsync(din, stage.data(), stage.size() * 8, cudaMemcpyHostToDevice, st));
  a.coeffs = din; a.breaks = din + 4 * nl;
  const double *dL = din + 4 * nl + 2, *dVs = dL + nl, *dVe = dVs + nl;
  double* q = din + stage.size();
  auto up = [&](const double* src, size_t n, const double** dst) -> cudaError_t { *dst = q; cudaError_t e = cudaMemcpyAsync(q, src, n * 8, cudaMemcpyHostToDevice, st); q += n; return e; };
  if (vgrid) { SCB_CUDA(up(in->vel_lo_grid, ng, &a.vel_lo_grid)); SCB_CUDA(up(in->vel_hi_grid, ng, &a.vel_hi_grid)); }
  else { SCB_CUDA(up(in->vel_lo, nl, &a.vel_lo)); SCB_CUDA(up(in->vel_hi, nl, &a.vel_hi)); }
  SCB_CUDA(up(in->acc_lo, nl, &a.acc_lo)); SCB_CUDA(up(in->acc_hi, nl, &a.acc_hi));
  double* o = (double*)dout;
  a.K_lo = o; a.K_hi = o + gp; a.x = o + 2 * gp;
  double* dT = o + 3 * gp; double* dXb = spline ? dT + gp : nullptr; double* dCp = spline ? dT + 2 * gp : nullptr;
  double* dTt = dT + (spline ? 3 : 1) * gp; double* dSt = dTt + P; double* dSp = dSt + sp; double* dSv = dSp + sp * m; double* dSa = dSv + sp * m;
  int32_t* oi = (int32_t*)(dSa + sp * m);
  a.status = oi; a.fail_stage = oi + P;
  int32_t* dNs = oi + 2 * (size_t)P;
  scb::launch_toppra_ppoly(a, st);
  for (int k = 0; k < m; ++k) {
    ParamArgs pa{};
    pa.P = P; pa.G = G; pa.L = dL + (size_t)k * P; pa.vel_start = dVs + (size_t)k * P; pa.vel_end = dVe + (size_t)k * P;
    pa.x = a.x; pa.t_buf = dT; pa.X_buf = dXb; pa.cp_buf = dCp; pa.T_total = dTt; pa.status = a.status;
    pa.keep_X = 1; pa.param_mode = in->param_mode;
    launch_param<double>(pa, st);
    if (S > 0) {
      SampleArgs sa{};
      sa.P = P; sa.G = G; sa.W = 2; sa.deg = 3; sa.S_cap = S; sa.dt = in->dt > 0 ? in->dt : (double)0.02f;
      sa.L = pa.L; sa.vel_start = pa.vel_start; sa.vel_end = pa.vel_end; sa.t_buf = dT; sa.X_buf = dXb; sa.x = a.x;
      sa.param_mode = in->param_mode; sa.n_samples = dNs; sa.s_time = dSt;
      sa.s_pos = dSp + (size_t)k * sp; sa.s_vel = dSv + (size_t)k * sp; sa.s_acc = dSa + (size_t)k * sp;
      sa.status = a.status; sa.timing = 1;
      launch_sample<double>(sa, st);
    }
  }
  SCB_CUDA(cudaGetLastError());
  auto down = [&](void* dst, const void* src, size_t bytes) -> cudaError_t { return dst ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st) : cudaSuccess; };
  SCB_CUDA(down(out->status, a.status, (size_t)P * 4)); SCB_CUDA(down(out->fail_stage, a.fail_stage, (size_t)P * 4));
  SCB_CUDA(down(out->K_lo, a.K_lo, gp * 8)); SCB_CUDA(down(out->K_hi, a.K_hi, gp * 8)); SCB_CUDA(down(out->x, a.x, gp * 8));
  SCB_CUDA(down(out->t, dT, gp * 8)); SCB_CUDA(down(out->T_total, dTt, (size_t)P * 8));
  std::vector<int32_t> ns_host(S > 0 ? P : 0);
  if (S > 0) {
    SCB_CUDA(cudaMemcpyAsync(ns_host.data(), dNs, (size_t)P * 4, cudaMemcpyDeviceToHost, st)); SCB_CUDA(down(out->s_time, dSt, sp * 8));
    SCB_CUDA(down(out->s_pos, dSp, sp * m * 8)); SCB_CUDA(down(out->s_vel, dSv, sp * m * 8)); SCB_CUDA(down(out->s_acc, dSa, sp * m * 8));
  }
  SCB_CUDA(cudaStreamSynchronize(st));
  // positions are recorded relative to pos_start (the 1-DOF kernels' Hermite path starts at 0): add the offset to the
  // rows that were written (rows beyond n_samples stay 0, failed problems have none)
  if (S > 0 && out->n_samples) std::memcpy(out->n_samples, ns_host.data(), (size_t)P * 4);
  if (S > 0 && out->s_pos) {
    for (int k = 0; k < m; ++k)
      for (int p = 0; p < P; ++p) {
        const double p0 = in->pos_start[(size_t)p * m + k];
        if (p0 == 0.0 || out->status[p] != SCB_ST_OK) continue;
        const int n = ns_host[p] < S ? ns_host[p] : S;
        double* col = out->s_pos + (size_t)k * sp + p;
        for (int i = 0; i < n; ++i) col[(size_t)i * P] += p0;
      }
  }
  release_retired(ctx);
  return SCB_OK;
}

void* scb_alloc_pinned(size_t bytes) {
  void* p = nullptr;
  if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}
void scb_free_pinned(void* p) { if (p) cudaFreeHost(p); }

int scb_get_timing(scb_ctx* ctx, scb_timing* out) {
  if (!ctx || !out) return SCB_E_INVALID;
  *out = ctx->timing;
  return SCB_OK;
}

const char* scb_last_cuda_error(scb_ctx* ctx) { return ctx ? ctx->err : ""; }
void* scb_stream(scb_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

}  // extern "C"
