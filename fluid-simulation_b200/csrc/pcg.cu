// pcg.cu -- matrix-free 7-point Jacobi-preconditioned conjugate gradient on the fluid cells.
//   replaces: setA fluid.cc:304-412, setA2 :481-541 (never materialised), cg.compute(A) :1473,
//             p = cg.solve(b) :1474 (Eigen ConjugateGradient.h:26-90 loop, same stop rule).
// Matrix (from setA): for a fluid cell c with k non-solid neighbours, A_cc = Adiag = k
// fp32-accumulated additions of scale = dt/(rho dx^2) (fluid.cc:333-370), A_cn = float(-scale) for
// every fluid neighbour n (:335); air neighbours are Dirichlet 0, solid neighbours Neumann.
// All vectors hold 0 on non-fluid nodes, so A.p = diag[k]*p_c + off * (sum of the 6 neighbours).
//
// One iteration = 3 kernels (algorithmic bytes per swept node, fp64: 17 + 49 + 25 = 91):
//   K1 k_spmv      Ap = A p,  pAp += p.Ap                       read p 8, flag 1; write Ap 8
//   K2 k_update_xr x += a p, r -= a Ap, rr += r.r, rz += r.M^-1 r   read x,p,r,Ap 32, flag 1; write x,r 16
//   K3 k_update_p  p = M^-1 r + b p  (or stop: rr < tol^2 |b|^2)    read r,p 16, flag 1; write p 8
// Scalars never visit the host inside the loop: dot products land in a 4-slot ring in device
// memory (slot = iteration & 3), alpha/beta are formed by the consuming kernel, convergence sets
// ds->pcg_done and the remaining launches of a batch return immediately.
#include "common.cuh"

struct PcgCoef {
  double diag[7];   // Adiag for k = 0..6 non-solid neighbours (fp32-accumulated, widened)
  double inv[7];    // Jacobi: 1/diag
  double off;       // (double)(float)(-scale)
};

#define PCG_TX 64
#define PCG_TY 8
#define PCG_ZC 16

struct Tiling { int tiles_x, tiles_y, chunks_z, nitems; };

__device__ __forceinline__ void item_to_tile(const Tiling& t, int item, int& tx, int& ty, int& tz) {
  tz = item / (t.tiles_x * t.tiles_y);
  int rem = item - tz * t.tiles_x * t.tiles_y;
  ty = rem / t.tiles_x;
  tx = rem - ty * t.tiles_x;
}

__global__ void k_pcg_reset(DevScalars* ds) {
  for (int s = 0; s < 4; ++s) { ds->pAp[s] = 0.0; ds->rz[s] = 0.0; ds->rr[s] = 0.0; }
  ds->pcg_done = 0; ds->pcg_iters = 0; ds->pcg_status = 0; ds->rr_final = 0.0;
}

// x = 0, r = b, p = M^-1 r, rz[3] = r.p   (Eigen ConjugateGradient.h:43,63,66 with x0 = 0)
__global__ void __launch_bounds__(PCG_TX * PCG_TY)
k_pcg_init(Grid g, Box box, Tiling t, PcgCoef cf, const uint8_t* __restrict__ fl, const float* __restrict__ b,
           double* __restrict__ x, double* __restrict__ r, double* __restrict__ p, DevScalars* __restrict__ ds,
           int mg) {
  __shared__ double sm[32];
  double acc[1] = {0.0};
  for (int item = blockIdx.x; item < t.nitems; item += gridDim.x) {
    int tx, ty, tz; item_to_tile(t, item, tx, ty, tz);
    int i = box.x0 + tx * PCG_TX + threadIdx.x, j = box.y0 + ty * PCG_TY + threadIdx.y;
    if (i >= box.x1 || j >= box.y1) continue;
    int zs = box.z0 + tz * PCG_ZC, ze = min(zs + PCG_ZC, box.z1);
    int q = gidx(g, i, j, zs);
    for (int z = zs; z < ze; ++z, q += g.plane) {
      uint8_t f = fl[q];
      double rv = 0.0, pv = 0.0;
      if (f & FB_FLUID) {
        int k = 6 - __popc(f & FB_SMASK);
        rv = (double)b[q];
        pv = rv * cf.inv[k];
        acc[0] += rv * pv;
      }
      x[q] = 0.0; r[q] = rv; p[q] = mg ? 0.0 : pv;
    }
  }
  if (mg) return;   // multigrid: z = M^-1 r and r.z come from the V-cycle
  block_sum<1>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0) atomicAdd(&ds->rz[3], acc[0]);
}

// K1: Ap = A p and p.Ap.  The block marches along z keeping p(z-1), p(z), p(z+1) in registers, so
// each p value is fetched from L2/HBM once per column; the four in-plane neighbours hit L1.
__global__ void __launch_bounds__(PCG_TX * PCG_TY)
k_spmv(Grid g, Box box, Tiling t, PcgCoef cf, const uint8_t* __restrict__ fl, const double* __restrict__ p,
       double* __restrict__ Ap, DevScalars* __restrict__ ds, int it, int noexit) {
  __shared__ double sm[32];
  if (!noexit && ds->pcg_done) return;
  double acc[1] = {0.0};
  for (int item = blockIdx.x; item < t.nitems; item += gridDim.x) {
    int tx, ty, tz; item_to_tile(t, item, tx, ty, tz);
    int i = box.x0 + tx * PCG_TX + threadIdx.x, j = box.y0 + ty * PCG_TY + threadIdx.y;
    if (i >= box.x1 || j >= box.y1) continue;
    int zs = box.z0 + tz * PCG_ZC, ze = min(zs + PCG_ZC, box.z1);
    int q = gidx(g, i, j, zs);
    double pm = zs > 0 ? p[q - g.plane] : 0.0;
    double pc = p[q];
    for (int z = zs; z < ze; ++z, q += g.plane) {
      double pn = (z + 1 < g.nz) ? p[q + g.plane] : 0.0;
      uint8_t f = fl[q];
      if (f & FB_FLUID) {   // fluid cells lie inside the shell: all six neighbours are in the box
        int k = 6 - __popc(f & FB_SMASK);
        double s = (pm + pn) + (p[q - 1] + p[q + 1]) + (p[q - g.px] + p[q + g.px]);
        double ap = cf.diag[k] * pc + cf.off * s;
        Ap[q] = ap;
        acc[0] += pc * ap;
      }
      pm = pc; pc = pn;
    }
  }
  block_sum<1>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0) atomicAdd(&ds->pAp[it & 3], acc[0]);
}

// K2: alpha = rz_old / pAp;  x += alpha p;  r -= alpha Ap;  rr = |r|^2;  rz = r.M^-1 r
template <bool JAC>
__global__ void __launch_bounds__(PCG_TX * PCG_TY)
k_update_xr(Grid g, Box box, Tiling t, PcgCoef cf, const uint8_t* __restrict__ fl, double* __restrict__ x,
            double* __restrict__ r, const double* __restrict__ p, const double* __restrict__ Ap,
            DevScalars* __restrict__ ds, int it, int noexit) {
  __shared__ double sm[64];
  if (!noexit && ds->pcg_done) return;
  const double pAp = ds->pAp[it & 3];
  const double rz_old = ds->rz[(it + 3) & 3];
  if (!noexit && !(pAp > 0.0)) {   // breakdown (also catches NaN)
    if (blockIdx.x == 0 && threadIdx.x == 0 && threadIdx.y == 0) {
      ds->pcg_status = FLIPB200_ERR_BREAKDOWN; ds->pcg_done = 1; ds->pcg_iters = it;
      ds->rr_final = ds->rr[(it + 3) & 3];
    }
    return;
  }
  const double alpha = rz_old / pAp;                   // ConjugateGradient.h:72
  double acc[2] = {0.0, 0.0};
  for (int item = blockIdx.x; item < t.nitems; item += gridDim.x) {
    int tx, ty, tz; item_to_tile(t, item, tx, ty, tz);
    int i = box.x0 + tx * PCG_TX + threadIdx.x, j = box.y0 + ty * PCG_TY + threadIdx.y;
    if (i >= box.x1 || j >= box.y1) continue;
    int zs = box.z0 + tz * PCG_ZC, ze = min(zs + PCG_ZC, box.z1);
    int q = gidx(g, i, j, zs);
    for (int z = zs; z < ze; ++z, q += g.plane) {
      uint8_t f = fl[q];
      if (f & FB_FLUID) {
        int k = 6 - __popc(f & FB_SMASK);
        double xn = x[q] + alpha * p[q];               // :73
        double rn = r[q] - alpha * Ap[q];              // :74
        x[q] = xn; r[q] = rn;
        acc[0] += rn * rn;                             // :76
        if (JAC) acc[1] += rn * (rn * cf.inv[k]);      // :80,83
      }
    }
  }
  block_sum<2>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0) {
    atomicAdd(&ds->rr[it & 3], acc[0]);
    if (JAC) atomicAdd(&ds->rz[it & 3], acc[1]);
  }
}

// K3: stop test (ConjugateGradient.h:77-78), beta = rz/rz_old (:84), p = z + beta p (:85)
__global__ void __launch_bounds__(PCG_TX * PCG_TY)
k_update_p(Grid g, Box box, Tiling t, PcgCoef cf, const uint8_t* __restrict__ fl, const double* __restrict__ r,
           double* __restrict__ p, DevScalars* __restrict__ ds, int it, int max_iter, double threshold, int noexit) {
  if (!noexit && ds->pcg_done) return;
  const double rr = ds->rr[it & 3];
  const bool lead = blockIdx.x == 0 && threadIdx.x == 0 && threadIdx.y == 0;
  if (!noexit && rr < threshold) {
    if (lead) { ds->pcg_done = 1; ds->pcg_iters = it; ds->rr_final = rr; }
    return;
  }
  const double beta = ds->rz[it & 3] / ds->rz[(it + 3) & 3];
  if (lead) {
    int s = (it + 1) & 3;
    ds->pAp[s] = 0.0; ds->rz[s] = 0.0; ds->rr[s] = 0.0;
    if (!noexit && it + 1 >= max_iter) {
      ds->pcg_done = 1; ds->pcg_iters = max_iter; ds->rr_final = rr; ds->pcg_status = FLIPB200_ERR_NOCONV;
    }
  }
  for (int item = blockIdx.x; item < t.nitems; item += gridDim.x) {
    int tx, ty, tz; item_to_tile(t, item, tx, ty, tz);
    int i = box.x0 + tx * PCG_TX + threadIdx.x, j = box.y0 + ty * PCG_TY + threadIdx.y;
    if (i >= box.x1 || j >= box.y1) continue;
    int zs = box.z0 + tz * PCG_ZC, ze = min(zs + PCG_ZC, box.z1);
    int q = gidx(g, i, j, zs);
    for (int z = zs; z < ze; ++z, q += g.plane) {
      uint8_t f = fl[q];
      if (f & FB_FLUID) {
        int k = 6 - __popc(f & FB_SMASK);
        p[q] = r[q] * cf.inv[k] + beta * p[q];
      }
    }
  }
}

// ---- multigrid-preconditioned variant: z = M^-1 r comes from fb_mg_vcycle (mg.cu) ---------------
// stop test of ConjugateGradient.h:77-78, evaluated before the preconditioner is applied
__global__ void k_pcg_check(DevScalars* __restrict__ ds, int it, double threshold) {
  if (ds->pcg_done) return;
  const double rr = ds->rr[it & 3];
  if (rr < threshold) { ds->pcg_done = 1; ds->pcg_iters = it; ds->rr_final = rr; }
}

// p = z + beta p, beta = rz/rz_old (first: p = z).  it = -1 for the initial direction.
__global__ void __launch_bounds__(PCG_TX * PCG_TY)
k_update_p_mg(Grid g, Box box, Tiling t, const uint8_t* __restrict__ fl, const double* __restrict__ z,
              double* __restrict__ p, DevScalars* __restrict__ ds, int it, int max_iter) {
  if (ds->pcg_done) return;
  const bool first = it < 0;
  const bool lead = blockIdx.x == 0 && threadIdx.x == 0 && threadIdx.y == 0;
  const double beta = first ? 0.0 : ds->rz[it & 3] / ds->rz[(it + 3) & 3];
  if (lead) {
    int s = (it + 1) & 3;
    ds->pAp[s] = 0.0; ds->rz[s] = 0.0; ds->rr[s] = 0.0;
    if (!first && it + 1 >= max_iter) {
      ds->pcg_done = 1; ds->pcg_iters = max_iter; ds->rr_final = ds->rr[it & 3]; ds->pcg_status = FLIPB200_ERR_NOCONV;
    }
  }
  for (int item = blockIdx.x; item < t.nitems; item += gridDim.x) {
    int tx, ty, tz; item_to_tile(t, item, tx, ty, tz);
    int i = box.x0 + tx * PCG_TX + threadIdx.x, j = box.y0 + ty * PCG_TY + threadIdx.y;
    if (i >= box.x1 || j >= box.y1) continue;
    int zs = box.z0 + tz * PCG_ZC, ze = min(zs + PCG_ZC, box.z1);
    int q = gidx(g, i, j, zs);
    for (int zz = zs; zz < ze; ++zz, q += g.plane) {
      if (fl[q] & FB_FLUID) p[q] = first ? z[q] : z[q] + beta * p[q];
    }
  }
}

static void make_coef(const flipb200_ctx* c, double dt, PcgCoef* cf) {
  const double scale = dt / (c->P.rho * c->P.dx * c->P.dx);       // fluid.cc:306
  float acc = 0.0f;
  cf->diag[0] = 0.0; cf->inv[0] = 0.0;
  for (int k = 1; k <= 6; ++k) {
    acc = (float)((double)acc + scale);                           // daccessor.setValue(c, getValue(c) + scale), :333
    cf->diag[k] = (double)acc;
    cf->inv[k] = 1.0 / (double)acc;
  }
  cf->off = (double)(float)(-1 * scale);                          // :335
}

static Tiling make_tiling(const Box& b) {
  Tiling t;
  t.tiles_x = cdiv(b.x1 - b.x0, PCG_TX);
  t.tiles_y = cdiv(b.y1 - b.y0, PCG_TY);
  t.chunks_z = cdiv(b.z1 - b.z0, PCG_ZC);
  t.nitems = t.tiles_x * t.tiles_y * t.chunks_z;
  return t;
}

int fb_pcg_solve(flipb200_ctx* c, double dt, int* iters_out, double* rel_out) {
  const Grid& g = c->g;
  const Box& b = c->box;
  if (iters_out) *iters_out = 0;
  if (rel_out) *rel_out = 0.0;
  if (c->n_fluid == 0 || !(b.x1 > b.x0)) return 0;
  PcgCoef cf; make_coef(c, dt, &cf);
  Tiling t = make_tiling(b);
  const int grid = min(t.nitems, c->num_sms * 2);
  const dim3 blk(PCG_TX, PCG_TY, 1);
  const double bnorm2 = c->hs->bnorm2;             // fetched by fb_rhs_div
  const float* bvec = c->bf[c->bcur];
  const bool use_mg = c->mg_levels >= 2;
  k_pcg_reset<<<1, 1, 0, c->stream>>>(c->ds);
  k_pcg_init<<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, bvec, c->x, c->r, c->pd, c->ds, use_mg ? 1 : 0);
  CK_LAUNCH(); c->launches += 2;
  if (bnorm2 == 0.0) return 0;                     // rhsNorm2 == 0 => x = 0, 0 iterations (ConjugateGradient.h:46-52)
  if (!(bnorm2 == bnorm2)) return FLIPB200_ERR_BREAKDOWN;
  const double tol = c->P.pcg_tol;
  const double threshold = tol * tol * bnorm2;     // :53
  const int max_iter = c->P.pcg_max_iter < 0 ? 2 * c->n_fluid : c->P.pcg_max_iter;   // IterativeSolverBase.h:283
  const int batch = use_mg ? 8 : 32;
  int it = 0;
  bool done = max_iter == 0;
  if (use_mg && !done) {
    int rc = fb_mg_setup(c, dt);                   // hierarchy of this step (flags are per step; cheap)
    if (rc) return rc;
    rc = fb_mg_vcycle(c, -1);                      // z0 = M^-1 r0, rz[3] = r0.z0
    if (rc) return rc;
    k_update_p_mg<<<grid, blk, 0, c->stream>>>(g, b, t, c->flags, c->Ap, c->pd, c->ds, -1, max_iter);
    c->launches++;
  }
  while (!done && use_mg) {
    int stop = min(it + batch, max_iter);
    for (; it < stop; ++it) {
      k_spmv<<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, c->pd, c->Ap, c->ds, it, 0);
      k_update_xr<false><<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, c->x, c->r, c->pd, c->Ap, c->ds, it, 0);
      k_pcg_check<<<1, 1, 0, c->stream>>>(c->ds, it, threshold);
      c->launches += 3;
      int rc = fb_mg_vcycle(c, it);
      if (rc) return rc;
      k_update_p_mg<<<grid, blk, 0, c->stream>>>(g, b, t, c->flags, c->Ap, c->pd, c->ds, it, max_iter);
      c->launches++;
    }
    CK_LAUNCH();
    int rc = fb_fetch_scalars(c);
    if (rc) return rc;
    done = c->hs->pcg_done != 0 || it >= max_iter;
  }
  while (!done) {
    int stop = min(it + batch, max_iter);
    for (; it < stop; ++it) {
      k_spmv<<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, c->pd, c->Ap, c->ds, it, 0);
      k_update_xr<true><<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, c->x, c->r, c->pd, c->Ap, c->ds, it, 0);
      k_update_p<<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, c->r, c->pd, c->ds, it, max_iter, threshold, 0);
      c->launches += 3;
    }
    CK_LAUNCH();
    int rc = fb_fetch_scalars(c);
    if (rc) return rc;
    done = c->hs->pcg_done != 0 || it >= max_iter;
  }
  if (iters_out) *iters_out = c->hs->pcg_iters;
  if (rel_out) *rel_out = sqrt(c->hs->rr_final / bnorm2);          // cg.error(), ConjugateGradient.h:88
  if (c->hs->pcg_status == FLIPB200_ERR_BREAKDOWN) return FLIPB200_ERR_BREAKDOWN;
  return 0;   // hitting max_iter is reported through iters == max_iter, as Eigen does
}

// Timing hook: `iters` launches of K1 alone (which == SPMV) or of the whole 3-kernel iteration
// (which == PCG_ITER) on the current system, no convergence exit.  units = swept nodes per launch.
int fb_pcg_time(flipb200_ctx* c, int which, int iters, float* ms, int64_t* units) {
  const Grid& g = c->g;
  const Box& b = c->box;
  if (!(b.x1 > b.x0)) return FLIPB200_ERR_STATE;
  PcgCoef cf; make_coef(c, c->P.dt_max, &cf);
  Tiling t = make_tiling(b);
  const int grid = min(t.nitems, c->num_sms * 2);
  const dim3 blk(PCG_TX, PCG_TY, 1);
  k_pcg_reset<<<1, 1, 0, c->stream>>>(c->ds);
  k_pcg_init<<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, c->bf[c->bcur], c->x, c->r, c->pd, c->ds, 0);
  for (int w = 0; w < 3; ++w) {
    k_spmv<<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, c->pd, c->Ap, c->ds, w, 1);
    if (which == FLIPB200_KERNEL_PCG_ITER) {
      k_update_xr<true><<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, c->x, c->r, c->pd, c->Ap, c->ds, w, 1);
      k_update_p<<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, c->r, c->pd, c->ds, w, 1 << 30, 0.0, 1);
    }
  }
  CK(cudaEventRecord(c->ev[6], c->stream));
  for (int it = 3; it < 3 + iters; ++it) {
    k_spmv<<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, c->pd, c->Ap, c->ds, it, 1);
    if (which == FLIPB200_KERNEL_PCG_ITER) {
      k_update_xr<true><<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, c->x, c->r, c->pd, c->Ap, c->ds, it, 1);
      k_update_p<<<grid, blk, 0, c->stream>>>(g, b, t, cf, c->flags, c->r, c->pd, c->ds, it, 1 << 30, 0.0, 1);
    }
  }
  CK(cudaEventRecord(c->ev[7], c->stream));
  CK(cudaStreamSynchronize(c->stream));
  CK_LAUNCH();
  float e = 0.f;
  CK(cudaEventElapsedTime(&e, c->ev[6], c->ev[7]));
  *ms = e / iters;
  *units = (int64_t)(b.x1 - b.x0) * (b.y1 - b.y0) * (b.z1 - b.z0);
  c->launches += (which == FLIPB200_KERNEL_PCG_ITER ? 3 : 1) * (iters + 3) + 2;
  return 0;
}
