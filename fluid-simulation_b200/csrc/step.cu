// step.cu -- the time step (loop body of main(), fluid.cc:1378-1490) and the stage-level C ABI.
#include "common.cuh"
#include <string.h>

namespace {

struct PhaseTimer {
  flipb200_ctx* c;
  explicit PhaseTimer(flipb200_ctx* ctx) : c(ctx) {}
  int mark(int e) { return cudaEventRecord(c->ev[e], c->stream) == cudaSuccess ? 0 : -1; }
  float ms(int a, int b) { float t = 0.f; cudaEventElapsedTime(&t, c->ev[a], c->ev[b]); return t; }
};

// do { setRHS; setDiver; setA; setA2; cg.compute; cg.solve; velUpdate(dt/10); setRHS; setDiver;
//      setOnlyB; error } while (error > 0.1)                                   fluid.cc:1457-1484
// A depends only on the flags and dt, both fixed within a step, so nothing is re-assembled; b of
// pass k+1 equals b2 of pass k (same velocities, same dt), so it is not recomputed either.
int relax_loop(flipb200_ctx* c, double dt, flipb200_stats* st, float* ms_pcg) {
  int rc = fb_rhs_div(c, dt, 0);
  if (rc) return rc;
  double error = 0.0;
  int passes = 0;
  *ms_pcg = 0.f;
  do {
    int iters = 0; double rel = 0.0;
    cudaEventRecord(c->ev[4], c->stream);
    c->pass_idx = passes;
    rc = fb_pcg_solve(c, dt, &iters, &rel);
    cudaEventRecord(c->ev[5], c->stream);
    if (rc) return rc;
    rc = fb_vel_update(c, dt / c->P.relax_div);                         // fluid.cc:1475
    if (rc) return rc;
    rc = fb_rhs_div(c, dt, 1);                                          // :1477-1481 (syncs)
    if (rc) return rc;
    float t = 0.f; cudaEventElapsedTime(&t, c->ev[4], c->ev[5]); *ms_pcg += t;
    error = sqrt(c->hs->err_num) / sqrt(c->hs->err_den);                // :1483  (x/0 = inf loops, 0/0 = NaN exits)
    ++passes;
    st->pcg_iters_total += iters; st->pcg_iters_last = iters; st->pcg_rel_resid = rel;
    if (passes >= c->P.max_passes) break;
  } while (error > c->P.relax_tol);                                     // :1484
  st->passes = passes;
  st->last_error = error;
  c->last_step_passes = passes;          // warm start (pcg.cu): which solutions x / xprev hold for the next step
  return 0;
}

int step_device(flipb200_ctx* c, double* dt_inout, flipb200_stats* st_out) {
  flipb200_stats st; memset(&st, 0, sizeof(st));
  const int64_t l0 = c->launches;
  PhaseTimer T(c);
  const double dt = *dt_inout;
  int rc;
  T.mark(0);
  if ((rc = fb_sort_particles(c))) return rc;
  T.mark(1);
  if ((rc = fb_p2g(c))) return rc;                                      // fluid.cc:1378,1384
  T.mark(2);
  if ((rc = fb_classify(c))) return rc;                                 // :1388-1455
  st.n_fluid = c->n_fluid;
  float ms_pcg = 0.f;
  if ((rc = relax_loop(c, dt, &st, &ms_pcg))) return rc;                // :1457-1484
  if ((rc = fb_cell_delta(c))) return rc;
  T.mark(3);
  if ((rc = fb_g2p_advect(c))) return rc;                               // :1490
  T.mark(6);
  if ((rc = fb_fetch_scalars(c))) return rc;
  st.dt_out = c->hs->dt_out;
  unsigned long long mb = c->hs->max_speed_bits;
  memcpy(&st.max_speed, &mb, sizeof(double));
  st.ms_sort = T.ms(0, 1); st.ms_p2g = T.ms(1, 2); st.ms_pcg = ms_pcg;
  st.ms_grid = T.ms(2, 3) - ms_pcg; st.ms_g2p = T.ms(3, 6); st.ms_total = T.ms(0, 6);
  st.launches = c->launches - l0;
  *dt_inout = st.dt_out;
  if (st_out) *st_out = st;
  return FLIPB200_OK;
}

}  // namespace

extern "C" {

int flipb200_step(flipb200_ctx* c, double* dt_inout, flipb200_stats* st) {
  if (!c || !dt_inout) return FLIPB200_ERR_ARG;
  CK(cudaSetDevice(c->device));
  return step_device(c, dt_inout, st);
}

static int step_host_impl(flipb200_ctx* c, double* const arr[6], int64_t n, int64_t cap, int64_t* n_out,
                          double* dt_inout, float* output_grid, flipb200_stats* st) {
  if (!c || !dt_inout || n < 0 || cap < n) return FLIPB200_ERR_ARG;
  CK(cudaSetDevice(c->device));
  if (n >= (int64_t)2147483647) return FLIPB200_ERR_ARG;
  c->pt.n = 0;
  int rc = fb_reserve_particles(c, n);
  if (rc) return rc;
  for (int q = 0; q < 6; ++q)
    if (n) CK(cudaMemcpyAsync(c->pt.a[q], arr[q], n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  c->pt.n = n; c->sorted = false; c->sparse_ok = false;
  if (fb_init_ids(c, 0, n, 0)) return FLIPB200_ERR_CUDA;
  rc = step_device(c, dt_inout, st);
  if (rc) return rc;
  // With params.track_ids the arrays come back in the caller's order (the reference never
  // reorders its PointList); otherwise cell-sorted, i.e. a permutation of the input.  On a slab the
  // rank's particle set itself changes (migration): *n_out particles come back, cell-sorted.
  if (n_out) *n_out = c->pt.n;
  if (c->pt.n > cap) return FLIPB200_ERR_ARG;
  if (fb_download_particles(c, arr)) return FLIPB200_ERR_CUDA;
  if (output_grid) {
    const Grid& g = c->g;
    CK(cudaMemcpy2DAsync(output_grid, g.nx * sizeof(float), c->wf + (size_t)g.k0 * g.plane, g.px * sizeof(float),
                         g.nx * sizeof(float), (size_t)g.ny * (g.k1 - g.k0), cudaMemcpyDeviceToHost, c->stream));
  }
  CK(cudaStreamSynchronize(c->stream));
  return FLIPB200_OK;
}

int flipb200_step_host(flipb200_ctx* c, double* px, double* py, double* pz, double* vx, double* vy, double* vz,
                       int64_t n, double* dt_inout, float* output_grid, flipb200_stats* st) {
  if (c && c->world > 1) return FLIPB200_ERR_STATE;   // slab contexts: flipb200_step_host_slab
  double* arr[6] = {px, py, pz, vx, vy, vz};
  return step_host_impl(c, arr, n, n, nullptr, dt_inout, output_grid, st);
}

int flipb200_step_host_slab(flipb200_ctx* c, double* px, double* py, double* pz, double* vx, double* vy, double* vz,
                            int64_t n, int64_t cap, int64_t* n_out, double* dt_inout, float* output_grid,
                            flipb200_stats* st) {
  if (!n_out) return FLIPB200_ERR_ARG;
  double* arr[6] = {px, py, pz, vx, vy, vz};
  return step_host_impl(c, arr, n, cap, n_out, dt_inout, output_grid, st);
}

int flipb200_seed_box(flipb200_ctx* c, const int32_t lo[3], const int32_t hi[3], uint64_t seed) {
  if (!c || !lo || !hi) return FLIPB200_ERR_ARG;
  CK(cudaSetDevice(c->device));
  return fb_seed_box(c, lo, hi, seed);
}

int flipb200_emit_box(flipb200_ctx* c, const int32_t lo[3], const int32_t hi[3], uint64_t seed) {
  if (!c || !lo || !hi) return FLIPB200_ERR_ARG;
  CK(cudaSetDevice(c->device));
  return fb_emit_box(c, lo, hi, seed);
}

int flipb200_p2g(flipb200_ctx* c) {
  if (!c) return FLIPB200_ERR_ARG;
  CK(cudaSetDevice(c->device));
  int rc = fb_sort_particles(c);
  if (rc) return rc;
  rc = fb_p2g(c);
  if (rc) return rc;
  CK(cudaStreamSynchronize(c->stream));
  c->stage = 1;
  return FLIPB200_OK;
}

int flipb200_classify(flipb200_ctx* c, int32_t* n_fluid) {
  if (!c) return FLIPB200_ERR_ARG;
  CK(cudaSetDevice(c->device));
  int rc = fb_classify(c);
  if (rc) return rc;
  if (n_fluid) *n_fluid = c->n_fluid;
  c->stage = 2;
  return FLIPB200_OK;
}

int flipb200_rhs_div(flipb200_ctx* c, double dt, double* b_norm2) {
  if (!c) return FLIPB200_ERR_ARG;
  if (c->stage < 2) return FLIPB200_ERR_STATE;
  CK(cudaSetDevice(c->device));
  int rc = fb_rhs_div(c, dt, 0);
  if (rc) return rc;
  if (b_norm2) *b_norm2 = c->hs->bnorm2;
  c->stage = 3;
  return FLIPB200_OK;
}

int flipb200_pcg_solve(flipb200_ctx* c, double dt, int32_t* iters, double* rel_resid) {
  if (!c) return FLIPB200_ERR_ARG;
  if (c->stage < 3) return FLIPB200_ERR_STATE;
  CK(cudaSetDevice(c->device));
  int it = 0; double rel = 0.0;
  c->pass_idx = 0;
  c->last_step_passes = 1;               // stage-level call: the guess is whatever the last solve left in x
  int rc = fb_pcg_solve(c, dt, &it, &rel);
  if (iters) *iters = it;
  if (rel_resid) *rel_resid = rel;
  if (rc) return rc;
  CK(cudaStreamSynchronize(c->stream));
  return FLIPB200_OK;
}

int flipb200_vel_update(flipb200_ctx* c, double dt_sub) {
  if (!c) return FLIPB200_ERR_ARG;
  if (c->stage < 2) return FLIPB200_ERR_STATE;
  CK(cudaSetDevice(c->device));
  int rc = fb_vel_update(c, dt_sub);
  if (rc) return rc;
  CK(cudaStreamSynchronize(c->stream));
  return FLIPB200_OK;
}

int flipb200_relax_error(flipb200_ctx* c, double dt, double* error) {
  if (!c) return FLIPB200_ERR_ARG;
  if (c->stage < 3) return FLIPB200_ERR_STATE;
  CK(cudaSetDevice(c->device));
  int rc = fb_rhs_div(c, dt, 1);
  if (rc) return rc;
  if (error) *error = sqrt(c->hs->err_num) / sqrt(c->hs->err_den);
  return FLIPB200_OK;
}

int flipb200_g2p_advect(flipb200_ctx* c, double* dt_out) {
  if (!c) return FLIPB200_ERR_ARG;
  if (c->stage < 2) return FLIPB200_ERR_STATE;
  CK(cudaSetDevice(c->device));
  int rc = fb_cell_delta(c);
  if (rc) return rc;
  rc = fb_g2p_advect(c);
  if (rc) return rc;
  rc = fb_fetch_scalars(c);
  if (rc) return rc;
  if (dt_out) *dt_out = c->hs->dt_out;
  c->stage = 0;
  return FLIPB200_OK;
}

int flipb200_time_kernel(flipb200_ctx* c, int which, int32_t iters, float* ms_per_launch, int64_t* units) {
  if (!c || !ms_per_launch || !units || iters < 1) return FLIPB200_ERR_ARG;
  CK(cudaSetDevice(c->device));
  switch (which) {
    case FLIPB200_KERNEL_SPMV:
    case FLIPB200_KERNEL_PCG_ITER:
      return fb_pcg_time(c, which, iters, ms_per_launch, units);
    case FLIPB200_KERNEL_SORT: {
      for (int w = 0; w < 2; ++w) { int rc = fb_sort_particles(c); if (rc) return rc; }
      CK(cudaEventRecord(c->ev[6], c->stream));
      for (int i = 0; i < iters; ++i) { int rc = fb_sort_particles(c); if (rc) return rc; }
      CK(cudaEventRecord(c->ev[7], c->stream));
      CK(cudaStreamSynchronize(c->stream));
      float e; CK(cudaEventElapsedTime(&e, c->ev[6], c->ev[7]));
      *ms_per_launch = e / iters; *units = c->pt.n;
      return FLIPB200_OK;
    }
    case FLIPB200_KERNEL_P2G: {
      if (!c->sorted) { int rc = fb_sort_particles(c); if (rc) return rc; }
      for (int w = 0; w < 2; ++w) { int rc = fb_p2g(c); if (rc) return rc; }
      CK(cudaEventRecord(c->ev[6], c->stream));
      for (int i = 0; i < iters; ++i) { int rc = fb_p2g(c); if (rc) return rc; }
      CK(cudaEventRecord(c->ev[7], c->stream));
      CK(cudaStreamSynchronize(c->stream));
      float e; CK(cudaEventElapsedTime(&e, c->ev[6], c->ev[7]));
      *ms_per_launch = e / iters; *units = c->pt.n;
      c->stage = 0;   // accumulators hold un-normalised sums
      return FLIPB200_OK;
    }
    default:
      return FLIPB200_ERR_ARG;
  }
}

int flipb200_profile_iteration(flipb200_ctx* c, int32_t iters, int32_t cap, char* names, float* us, int32_t* n_out) {
  if (!c || !names || !us || !n_out || iters < 1 || cap < 1) return FLIPB200_ERR_ARG;
  CK(cudaSetDevice(c->device));
  if (!c->prof) c->prof = (IterProf*)calloc(1, sizeof(IterProf));
  if (!c->prof) return FLIPB200_ERR_ARG;
  c->prof->request = true;
  c->prof->out_n = 0;
  float ms = 0.f; int64_t units = 0;
  int rc = fb_pcg_time(c, FLIPB200_KERNEL_PCG_ITER, iters, &ms, &units);
  c->prof->request = false;
  if (rc) return rc;
  const int n = c->prof->out_n < cap ? c->prof->out_n : cap;
  for (int m = 0; m < n; ++m) {
    us[m] = c->prof->out_us[m];
    snprintf(names + (size_t)m * FLIPB200_PROFILE_NAME, FLIPB200_PROFILE_NAME, "%s", c->prof->out_name[m]);
  }
  *n_out = n;
  return FLIPB200_OK;
}

}  // extern "C"
