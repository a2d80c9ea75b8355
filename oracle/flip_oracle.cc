/*
 * flip_oracle.cc -- CPU ORACLE (test infrastructure only; see flip_oracle.h).
 *
 * Dense-array restatement of the FLIP substep of /root/reference/fluid.cc.  Every function
 * cites the reference lines it follows.  Build flags mirror run.sh:3 (-std=c++11 -O3 -DNDEBUG);
 * no -ffast-math, no FMA contraction (x86-64 baseline), so the arithmetic order is the
 * reference's.  Parity status: "unpinned by the reference" except for the Eigen solve and the
 * analytic KATs listed in flip_oracle.h.
 */
#include "flip_oracle.h"
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <cstdio>
#include <vector>
#include <chrono>
#include <algorithm>
#ifdef _OPENMP
#include <omp.h>
#endif

/* eigen_solver.cc (reference's vendored Eigen); absent when built without ORACLE_WITH_EIGEN */
#ifdef ORACLE_WITH_EIGEN
extern "C" {
void* oracle_eig_create(void);
void  oracle_eig_destroy(void*);
int   oracle_eig_compute_solve(void*, int n, int64_t nnz, const int32_t* rows, const int32_t* cols,
                               const double* vals, const double* b, double* x, int max_iter,
                               int* iters, double* err);
int   oracle_eig_dense(int n, const double* A, const double* b, double* x);
}
#endif

namespace {
typedef std::chrono::steady_clock Clock;
inline double secs(Clock::time_point a, Clock::time_point b) {
  return std::chrono::duration<double>(b - a).count();
}
}  // namespace

struct oracle_ctx {
  oracle_params P;
  int64_t N;
  std::vector<uint8_t> solid;
  std::vector<float> fg[9];
  std::vector<double> dg[7];
  std::vector<int32_t> indices;
  int numActive;
  std::vector<double> px, py, pz, vx, vy, vz;
  std::vector<double> b, b2, p;
  std::vector<int32_t> trow, tcol;
  std::vector<double> tval;
  std::vector<double> hist_err;
  std::vector<int32_t> hist_it;
  void* eig;
  double last_max_speed;

  inline int64_t id(int i, int j, int k) const { return ((int64_t)k * P.ny + j) * P.nx + i; }
  inline bool inBox(int i, int j, int k) const {
    return i >= 0 && j >= 0 && k >= 0 && i < P.nx && j < P.ny && k < P.nz;
  }
  /* isSolid, fluid.cc:46-57: out-of-box reads return the VDB background 0 => not solid */
  inline bool isSolid(int i, int j, int k) const { return inBox(i, j, k) && solid[id(i, j, k)] == 1; }
  /* isWithinBounds(c, bound-2) / (c, 58), fluid.cc:38-45, in index space */
  inline bool inside(int i, int j, int k) const {
    const int w = P.wall;
    return i >= w && j >= w && k >= w && i <= P.nx - 1 - w && j <= P.ny - 1 - w && k <= P.nz - 1 - w;
  }
  inline float fget(int g, int i, int j, int k) const { return inBox(i, j, k) ? fg[g][id(i, j, k)] : 0.0f; }
  inline double dget(int g, int i, int j, int k) const { return inBox(i, j, k) ? dg[g][id(i, j, k)] : 0.0; }
};

extern "C" {

void oracle_default_params(oracle_params* p) {
  memset(p, 0, sizeof(*p));
  p->nx = p->ny = p->nz = 121;
  p->ox = p->oy = p->oz = 60;
  p->wall = 2;
  p->relax_div = 10;
  p->dx = 1.0; p->rho = 1.0;
  p->gx = 0; p->gy = -10.0; p->gz = 0;
  p->dt_max = 0.1;
  p->relax_tol = 0.1;
  p->restitution = 0.0;
  p->solver = 0;
  p->max_passes = 200;
  p->threads = 1;
  p->pcg_max_iter = -1;
  p->flip_ratio = 1.0;
  p->extrapolate = 0;
}

int oracle_has_eigen(void) {
#ifdef ORACLE_WITH_EIGEN
  return 1;
#else
  return 0;
#endif
}

/* fluid.cc:22-37 */
double oracle_spline(double x) {
  if (x < 0) x *= -1.0;
  if (x < 0.5) return 1.5 * (4.0 * x * x * x - 4.0 * x * x + 2.0 / 3.0);
  if (x < 1.0) return 1.5 * ((-8.0 * (x * x * x) / 6.0) + 4.0 * x * x - 4.0 * x + 4.0 / 3.0);
  return 0;
}

oracle_ctx* oracle_create(const oracle_params* p, const uint8_t* solid_in) {
  oracle_ctx* c = new oracle_ctx();
  c->P = *p;
  c->N = (int64_t)p->nx * p->ny * p->nz;
  c->solid.assign(c->N, 0);
  for (int g = 0; g < 9; ++g) c->fg[g].assign(c->N, 0.0f);
  for (int g = 0; g < 7; ++g) c->dg[g].assign(c->N, 0.0);
  c->indices.assign(c->N, -1);
  c->numActive = 0;
  c->eig = 0;
  c->last_max_speed = 0;
  /* solid shell: fluid.cc:1256-1266 (|c| > 58 <=> outside the inner box) */
  for (int k = 0; k < p->nz; ++k)
    for (int j = 0; j < p->ny; ++j)
      for (int i = 0; i < p->nx; ++i) {
        int64_t q = c->id(i, j, k);
        if (!c->inside(i, j, k)) c->solid[q] = 1;
        else if (solid_in && solid_in[q]) c->solid[q] = 1;
      }
#ifdef ORACLE_WITH_EIGEN
  c->eig = oracle_eig_create();  /* one cg object for the whole run, fluid.cc:1352 */
#endif
  return c;
}

void oracle_destroy(oracle_ctx* c) {
  if (!c) return;
#ifdef ORACLE_WITH_EIGEN
  if (c->eig) oracle_eig_destroy(c->eig);
#endif
  delete c;
}

void oracle_set_particles(oracle_ctx* c, const double* px, const double* py, const double* pz,
                          const double* vx, const double* vy, const double* vz, int64_t n) {
  c->px.assign(px, px + n); c->py.assign(py, py + n); c->pz.assign(pz, pz + n);
  c->vx.assign(vx, vx + n); c->vy.assign(vy, vy + n); c->vz.assign(vz, vz + n);
}
int64_t oracle_num_particles(oracle_ctx* c) { return (int64_t)c->px.size(); }
void oracle_get_particles(oracle_ctx* c, double* px, double* py, double* pz, double* vx, double* vy,
                          double* vz) {
  size_t n = c->px.size() * sizeof(double);
  memcpy(px, c->px.data(), n); memcpy(py, c->py.data(), n); memcpy(pz, c->pz.data(), n);
  memcpy(vx, c->vx.data(), n); memcpy(vy, c->vy.data(), n); memcpy(vz, c->vz.data(), n);
}

/* fluid.cc:1378  vels->fill(0) */
void oracle_clear_vel(oracle_ctx* c) {
  for (int g = ORACLE_U; g <= ORACLE_WV; ++g) std::fill(c->dg[g].begin(), c->dg[g].end(), 0.0);
}

/* window of p2gCatmullRom / CatmullRomFLIP / interpolate: fluid.cc:267-276, 212-221, 848-857.
 * round() is C99 half-away-from-zero, evaluated in reference coordinates. */
static inline void window(const oracle_ctx* c, double x, double y, double z, int* lo, int* hi) {
  int fcx = (int)round(x), fcy = (int)round(y), fcz = (int)round(z);
  int fi = fcx + c->P.ox, fj = fcy + c->P.oy, fk = fcz + c->P.oz;
  lo[0] = fi - 1 > 0 ? fi - 1 : 0;           hi[0] = fi + 1 < c->P.nx - 1 ? fi + 1 : c->P.nx - 1;
  lo[1] = fj - 1 > 0 ? fj - 1 : 0;           hi[1] = fj + 1 < c->P.ny - 1 ? fj + 1 : c->P.ny - 1;
  lo[2] = fk - 1 > 0 ? fk - 1 : 0;           hi[2] = fk + 1 < c->P.nz - 1 ? fk + 1 : c->P.nz - 1;
}

/* PointList::P2Gtransfer fluid.cc:1106-1148 with p2gCatmullRom :265-299.
 * Serial particle order = one legal schedule of the reference's tbb::parallel_for. */
void oracle_p2g(oracle_ctx* c) {
  std::vector<float>& W = c->fg[ORACLE_W];
  std::vector<double>&U = c->dg[ORACLE_U], &V = c->dg[ORACLE_V], &Wv = c->dg[ORACLE_WV];
  std::fill(W.begin(), W.end(), 0.0f);                        /* :1108 */
  const int64_t np = (int64_t)c->px.size();
  const int ox = c->P.ox, oy = c->P.oy, oz = c->P.oz;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) if (c->P.threads > 1) num_threads(c->P.threads > 1 ? c->P.threads : 1)
#endif
  for (int64_t n = 0; n < np; ++n) {
    const double x = c->px[n], y = c->py[n], z = c->pz[n];
    const double ux = c->vx[n], uy = c->vy[n], uz = c->vz[n];
    int lo[3], hi[3];
    window(c, x, y, z, lo, hi);
    for (int i = lo[0]; i <= hi[0]; ++i)
      for (int j = lo[1]; j <= hi[1]; ++j)
        for (int k = lo[2]; k <= hi[2]; ++k) {
          if (!c->isSolid(i, j, k) && c->inside(i, j, k)) {   /* :288 */
            double cw = oracle_spline(x - (i - ox)) * oracle_spline(y - (j - oy)) * oracle_spline(z - (k - oz));
            int64_t q = c->id(i, j, k);
            if (c->P.threads > 1) {
              /* timing mode: atomics stand in for the per-node std::mutex (:290-294) */
#ifdef _OPENMP
#pragma omp atomic
#endif
              W[q] += cw;
#ifdef _OPENMP
#pragma omp atomic
#endif
              U[q] += cw * ux;
#ifdef _OPENMP
#pragma omp atomic
#endif
              V[q] += cw * uy;
#ifdef _OPENMP
#pragma omp atomic
#endif
              Wv[q] += cw * uz;
            } else {
              W[q] = (float)((double)W[q] + cw);              /* :292 fp32 store */
              U[q] = U[q] + cw * ux;                          /* :293 */
              V[q] = V[q] + cw * uy;
              Wv[q] = Wv[q] + cw * uz;
            }
          }
        }
  }
  /* :1131-1146 normalise where w > 0 */
  for (int64_t q = 0; q < c->N; ++q) {
    double w = W[q];
    if (w > 0) { U[q] /= w; V[q] /= w; Wv[q] /= w; }
  }
  if (c->P.extrapolate) oracle_extrapolate(c);               /* :1147 (commented out in the reference) */
}

/* extrapolate fluid.cc:705-802 with the `defined` grid P2Gtransfer builds (:1109-1125, :1141): defined =
 * outside 58 | solid | w > 0.  Wave 0 (:714-748): every defined, non-solid, in-bounds node (= w > 0) adds
 * its velocity to each undefined node of its 3x3x3 neighbourhood (clamped to the box, :720-725) and
 * counts it; the touched nodes are divided by their count and become defined (:749-753).  Then the
 * frontier repeats the same from the nodes defined by the previous wave only (:754-800) until nothing
 * new is reached.  The sums start from the node's own value (:739), 0 for an air node. */
void oracle_extrapolate(oracle_ctx* c) {
  const int nx = c->P.nx, ny = c->P.ny, nz = c->P.nz;
  std::vector<float>& W = c->fg[ORACLE_W];
  std::vector<double>& U = c->dg[ORACLE_U];
  std::vector<double>& V = c->dg[ORACLE_V];
  std::vector<double>& Wv = c->dg[ORACLE_WV];
  std::vector<uint8_t> defined(c->N, 0);
  std::vector<int32_t> count(c->N, 0);
  std::vector<int64_t> frontier, next;
  for (int i = 0; i < nx; ++i) for (int j = 0; j < ny; ++j) for (int k = 0; k < nz; ++k) {
    const int64_t q = c->id(i, j, k);
    defined[q] = (!c->inside(i, j, k) || c->isSolid(i, j, k) || W[q] > 0) ? 1 : 0;
  }
  auto spread = [&](int i, int j, int k, std::vector<int64_t>& out) {
    const int64_t q = c->id(i, j, k);
    for (int a = (i == 0 ? i : i - 1); a <= (i == nx - 1 ? i : i + 1); ++a)
      for (int b = (j == 0 ? j : j - 1); b <= (j == ny - 1 ? j : j + 1); ++b)
        for (int d = (k == 0 ? k : k - 1); d <= (k == nz - 1 ? k : k + 1); ++d) {
          const int64_t n = c->id(a, b, d);
          if (defined[n]) continue;
          if (count[n] == 0) out.push_back(n);
          count[n] += 1;
          U[n] = U[q] + U[n]; V[n] = V[q] + V[n]; Wv[n] = Wv[q] + Wv[n];
        }
  };
  auto settle = [&](const std::vector<int64_t>& v) {
    for (size_t t = 0; t < v.size(); ++t) {
      const int64_t n = v[t];
      U[n] = U[n] / count[n]; V[n] = V[n] / count[n]; Wv[n] = Wv[n] / count[n];
      defined[n] = 1;
    }
  };
  for (int i = 0; i < nx; ++i) for (int j = 0; j < ny; ++j) for (int k = 0; k < nz; ++k) {   /* :714-748, x outer */
    const int64_t q = c->id(i, j, k);
    if (defined[q] && !c->isSolid(i, j, k) && c->inside(i, j, k)) spread(i, j, k, frontier);
  }
  settle(frontier);
  while (!frontier.empty()) {                                  /* :754-800 */
    next.clear();
    for (size_t t = 0; t < frontier.size(); ++t) {
      const int64_t q = frontier[t];
      const int i = (int)(q % nx), j = (int)((q / nx) % ny), k = (int)(q / ((int64_t)nx * ny));
      spread(i, j, k, next);
    }
    settle(next);
    frontier.swap(next);
  }
}

/* main fluid.cc:1388-1448: indices=-1, container=0, PointList::interpolate (:843-882),
 * fluid-cell numbering in x-outer/y/z-inner order, outputGrid. */
int oracle_classify(oracle_ctx* c) {
  std::vector<float>& C = c->fg[ORACLE_CONTAINER];
  std::fill(c->indices.begin(), c->indices.end(), -1);       /* :1388 */
  std::fill(C.begin(), C.end(), 0.0f);                       /* :1396-1409 */
  const int64_t np = (int64_t)c->px.size();
  const int ox = c->P.ox, oy = c->P.oy, oz = c->P.oz;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) if (c->P.threads > 1) num_threads(c->P.threads > 1 ? c->P.threads : 1)
#endif
  for (int64_t n = 0; n < np; ++n) {                         /* :845-881 */
    const double x = c->px[n], y = c->py[n], z = c->pz[n];
    int lo[3], hi[3];
    window(c, x, y, z, lo, hi);
    for (int i = lo[0]; i <= hi[0]; ++i)
      for (int j = lo[1]; j <= hi[1]; ++j)
        for (int k = lo[2]; k <= hi[2]; ++k) {
          double cw = oracle_spline(x - (i - ox)) * oracle_spline(y - (j - oy)) * oracle_spline(z - (k - oz));
          if (!c->isSolid(i, j, k) && cw > 0) {                /* :870 */
            int64_t q = c->id(i, j, k);
            if (c->P.threads > 1) {
#ifdef _OPENMP
#pragma omp atomic
#endif
              C[q] += cw;
            } else {
              C[q] = (float)((double)C[q] + cw);              /* :873 */
            }
          }
        }
  }
  int numActive = 0;                                           /* :1416-1433 */
  for (int i = 0; i < c->P.nx; ++i)
    for (int j = 0; j < c->P.ny; ++j)
      for (int k = 0; k < c->P.nz; ++k)
        if (!c->isSolid(i, j, k) && c->inside(i, j, k)) {
          int64_t q = c->id(i, j, k);
          if (C[q] > 0) c->indices[q] = numActive++;
        }
  std::vector<float>& O = c->fg[ORACLE_OUTPUT];                /* :1434-1448 */
  for (int64_t q = 0; q < c->N; ++q)
    if (c->solid[q] != 1) O[q] = C[q];
  c->numActive = numActive;
  c->b.assign(numActive, 0.0);   /* reference leaves these uninitialised (:1449-1453) */
  c->b2.assign(numActive, 0.0);
  c->p.assign(numActive, 0.0);
  return numActive;
}

/* fluid.cc:1455 velBeforeUpdate = vels->deepCopy() */
void oracle_save_vel_before(oracle_ctx* c) {
  c->dg[ORACLE_UB] = c->dg[ORACLE_U];
  c->dg[ORACLE_VB] = c->dg[ORACLE_V];
  c->dg[ORACLE_WB] = c->dg[ORACLE_WV];
}

/* rhs.fill(0), diver.fill(0), setRHS (fluid.cc:414-479), setDiver (:566-610) */
void oracle_rhs_div(oracle_ctx* c, double dt) {
  std::vector<float>&R = c->fg[ORACLE_RHS], &D = c->fg[ORACLE_DIVER], &C = c->fg[ORACLE_CONTAINER];
  std::fill(R.begin(), R.end(), 0.0f);
  std::fill(D.begin(), D.end(), 0.0f);
  const double dx = c->P.dx;
  const double scale = 1.0 / dx;                               /* :416 */
  const double gx = c->P.gx * dt, gy = c->P.gy * dt, gz = c->P.gz * dt;   /* :420 */
  const int nx = c->P.nx, ny = c->P.ny, nz = c->P.nz;
  for (int i = 0; i < nx; ++i)
    for (int j = 0; j < ny; ++j)
      for (int k = 0; k < nz; ++k) {
        int64_t q = c->id(i, j, k);
        double val = C[q];
        if (val > 0 && !c->isSolid(i, j, k)) {                 /* :445 */
          double vx = c->dg[ORACLE_U][q], vy = c->dg[ORACLE_V][q], vz = c->dg[ORACLE_WV][q];
          double vix = c->dget(ORACLE_U, i + 1, j, k);
          double vjy = c->dget(ORACLE_V, i, j + 1, k);
          double vkz = c->dget(ORACLE_WV, i, j, k + 1);
          /* isWithinBounds(nbr, 60) == nbr in box; isSolid() already false out of box */
          if (c->isSolid(i - 1, j, k)) R[q] = (float)((double)R[q] - (scale * (vx + gx)));    /* :449 */
          if (c->isSolid(i + 1, j, k)) R[q] = (float)((double)R[q] + (scale * (vix + gx)));   /* :453 */
          if (c->isSolid(i, j - 1, k)) R[q] = (float)((double)R[q] - (scale * (vy + gy)));    /* :458 */
          if (c->isSolid(i, j + 1, k)) R[q] = (float)((double)R[q] + (scale * (vjy + gy)));   /* :462 */
          if (c->isSolid(i, j, k - 1)) R[q] = (float)((double)R[q] - (scale * (vz + gz)));    /* :467 */
          if (c->isSolid(i, j, k + 1)) R[q] = (float)((double)R[q] + (scale * (vkz + gz)));   /* :471 */
        }
      }
  for (int i = 0; i < nx; ++i)
    for (int j = 0; j < ny; ++j)
      for (int k = 0; k < nz; ++k) {
        int64_t q = c->id(i, j, k);
        if (C[q] > 0 && !c->isSolid(i, j, k)) {                /* :579 */
          double vx = c->dg[ORACLE_U][q], vy = c->dg[ORACLE_V][q], vz = c->dg[ORACLE_WV][q];
          double vix = c->dget(ORACLE_U, i + 1, j, k);
          double vjy = c->dget(ORACLE_V, i, j + 1, k);
          double vkz = c->dget(ORACLE_WV, i, j, k + 1);
          double u = 0, vd = 0, w = 0;
          if (!c->isSolid(i + 1, j, k)) u = (vix - vx) / dx;   /* :592-595 */
          if (!c->isSolid(i, j + 1, k)) vd = (vjy - vy) / dx;
          if (!c->isSolid(i, j, k + 1)) w = (vkz - vz) / dx;
          D[q] = (float)((double)R[q] - u - vd - w);           /* :605 */
        }
      }
}

/* Adiag/Aplus fills (:1461-1464), setA (fluid.cc:304-412), setA2 (:481-541) -> triplets, b */
void oracle_assemble(oracle_ctx* c, double dt) {
  std::vector<float>&AD = c->fg[ORACLE_ADIAG], &AI = c->fg[ORACLE_APLUSI], &AJ = c->fg[ORACLE_APLUSJ],
                    &AK = c->fg[ORACLE_APLUSK], &C = c->fg[ORACLE_CONTAINER], &D = c->fg[ORACLE_DIVER];
  std::fill(AD.begin(), AD.end(), 0.0f); std::fill(AI.begin(), AI.end(), 0.0f);
  std::fill(AJ.begin(), AJ.end(), 0.0f); std::fill(AK.begin(), AK.end(), 0.0f);
  const double scale = dt / (c->P.rho * c->P.dx * c->P.dx);   /* :306 */
  const int nx = c->P.nx, ny = c->P.ny, nz = c->P.nz;
#define ADD_DIAG(Q) AD[Q] = (float)((double)AD[Q] + scale)
  for (int i = 0; i < nx; ++i)
    for (int j = 0; j < ny; ++j)
      for (int k = 0; k < nz; ++k) {
        int64_t q = c->id(i, j, k);
        double val = C[q];
        const int di[3] = {1, 0, 0}, dj[3] = {0, 1, 0}, dk[3] = {0, 0, 1};
        std::vector<float>* AP[3] = {&AI, &AJ, &AK};
        if (!c->isSolid(i, j, k) && val > 0) {                 /* :326 */
          for (int a = 0; a < 3; ++a) {
            int ii = i + di[a], jj = j + dj[a], kk = k + dk[a];
            if (!c->isSolid(ii, jj, kk)) {                     /* :328 */
              double val2 = c->fget(ORACLE_CONTAINER, ii, jj, kk);
              if (val2 > 0) {
                ADD_DIAG(q);                                   /* :333 */
                if (c->inBox(ii, jj, kk)) { int64_t q2 = c->id(ii, jj, kk); ADD_DIAG(q2); }  /* :334 */
                (*AP[a])[q] = (float)(-1 * scale);             /* :335 */
              } else {
                ADD_DIAG(q);                                   /* :339 */
              }
            }
          }
        } else if (!c->isSolid(i, j, k) && c->inside(i, j, k)) {   /* :376 */
          for (int a = 0; a < 3; ++a) {
            int ii = i + di[a], jj = j + dj[a], kk = k + dk[a];
            if (!c->isSolid(ii, jj, kk)) {
              double val2 = c->fget(ORACLE_CONTAINER, ii, jj, kk);
              if (val2 > 0 && c->inBox(ii, jj, kk)) { int64_t q2 = c->id(ii, jj, kk); ADD_DIAG(q2); }  /* :383 */
            }
          }
        }
      }
#undef ADD_DIAG
  /* setA2 :481-541 -- triplet order kept exactly */
  c->trow.clear(); c->tcol.clear(); c->tval.clear();
  auto push = [&](int r, int cc, double v) { c->trow.push_back(r); c->tcol.push_back(cc); c->tval.push_back(v); };
  for (int i = 0; i < nx; ++i)
    for (int j = 0; j < ny; ++j)
      for (int k = 0; k < nz; ++k) {
        int64_t q = c->id(i, j, k);
        if (AD[q] != 0) {                                      /* :498 */
          int index = c->indices[q];
          if (AI[q] != 0) push(index, c->indices[c->id(i + 1, j, k)], AI[q]);
          if (AJ[q] != 0) push(index, c->indices[c->id(i, j + 1, k)], AJ[q]);
          if (AK[q] != 0) push(index, c->indices[c->id(i, j, k + 1)], AK[q]);
          if (c->fget(ORACLE_APLUSI, i - 1, j, k) != 0) push(index, c->indices[c->id(i - 1, j, k)], AI[c->id(i - 1, j, k)]);
          if (c->fget(ORACLE_APLUSJ, i, j - 1, k) != 0) push(index, c->indices[c->id(i, j - 1, k)], AJ[c->id(i, j - 1, k)]);
          if (c->fget(ORACLE_APLUSK, i, j, k - 1) != 0) push(index, c->indices[c->id(i, j, k - 1)], AK[c->id(i, j, k - 1)]);
          push(index, index, AD[q]);                           /* :534 */
          c->b[index] = D[q];                                  /* :535 */
        }
      }
}

/* own Jacobi-PCG (solver=1): same loop as Eigen's conjugate_gradient
 * (Eigen/src/IterativeLinearSolvers/ConjugateGradient.h:26-90) with a diagonal preconditioner.
 * Used only where the IC factorisation is too slow/large (>= 128^3) -- a "port", not the reference. */
static int jacobi_pcg(oracle_ctx* c, int max_iter, double* err) {
  const int n = c->numActive;
  std::vector<int64_t> rp(n + 1, 0);
  const size_t nt = c->trow.size();
  for (size_t t = 0; t < nt; ++t) rp[c->trow[t] + 1]++;
  for (int r = 0; r < n; ++r) rp[r + 1] += rp[r];
  std::vector<int32_t> ci(nt); std::vector<double> va(nt); std::vector<int64_t> cur(rp.begin(), rp.end() - 1);
  std::vector<double> diag(n, 0.0);
  for (size_t t = 0; t < nt; ++t) {
    int r = c->trow[t]; int64_t s = cur[r]++;
    ci[s] = c->tcol[t]; va[s] = c->tval[t];
    if (c->tcol[t] == r) diag[r] += c->tval[t];
  }
  std::vector<double> x(n, 0.0), r(c->b), z(n), pv(n), tmp(n);
  double rhs2 = 0; for (int i = 0; i < n; ++i) rhs2 += c->b[i] * c->b[i];
  if (rhs2 == 0) { std::fill(c->p.begin(), c->p.end(), 0.0); *err = 0; return 0; }
  const double tol = 2.220446049250313e-16;
  const double threshold = tol * tol * rhs2;
  double rn2 = rhs2;
  if (rn2 < threshold) { *err = std::sqrt(rn2 / rhs2); return 0; }
  for (int i = 0; i < n; ++i) pv[i] = r[i] / diag[i];
  double absNew = 0; for (int i = 0; i < n; ++i) absNew += r[i] * pv[i];
  int it = 0;
  while (it < max_iter) {
    for (int i = 0; i < n; ++i) { double s = 0; for (int64_t e = rp[i]; e < rp[i + 1]; ++e) s += va[e] * pv[ci[e]]; tmp[i] = s; }
    double pAp = 0; for (int i = 0; i < n; ++i) pAp += pv[i] * tmp[i];
    double alpha = absNew / pAp;
    for (int i = 0; i < n; ++i) { x[i] += alpha * pv[i]; r[i] -= alpha * tmp[i]; }
    rn2 = 0; for (int i = 0; i < n; ++i) rn2 += r[i] * r[i];
    if (rn2 < threshold) break;
    for (int i = 0; i < n; ++i) z[i] = r[i] / diag[i];
    double absOld = absNew; absNew = 0; for (int i = 0; i < n; ++i) absNew += r[i] * z[i];
    double beta = absNew / absOld;
    for (int i = 0; i < n; ++i) pv[i] = z[i] + beta * pv[i];
    ++it;
  }
  c->p = x;
  *err = std::sqrt(rn2 / rhs2);
  return it;
}

/* fluid.cc:1473-1474: cg.compute(A); p = cg.solve(b) */
int oracle_solve(oracle_ctx* c, double* rel_err) {
  int n = c->numActive;
  int iters = 0; double err = 0;
  if (n == 0) { if (rel_err) *rel_err = 0; return 0; }
  int max_iter = c->P.pcg_max_iter < 0 ? 2 * n : c->P.pcg_max_iter;   /* IterativeSolverBase.h:283 */
  if (c->P.solver == 0) {
#ifdef ORACLE_WITH_EIGEN
    int rc = oracle_eig_compute_solve(c->eig, n, (int64_t)c->trow.size(), c->trow.data(), c->tcol.data(),
                                      c->tval.data(), c->b.data(), c->p.data(), c->P.pcg_max_iter, &iters, &err);
    (void)rc;
#else
    fprintf(stderr, "flip_oracle: built without the reference's Eigen; solver=0 unavailable\n");
    abort();
#endif
  } else {
    iters = jacobi_pcg(c, max_iter, &err);
  }
  /* scatter p to the grid view for convenience (not in the reference) */
  std::vector<double>& PG = c->dg[ORACLE_PGRID];
  std::fill(PG.begin(), PG.end(), 0.0);
  for (int64_t q = 0; q < c->N; ++q) if (c->indices[q] >= 0) PG[q] = c->p[c->indices[q]];
  if (rel_err) *rel_err = err;
  return iters;
}

/* velUpdate fluid.cc:612-703 (sequential read-modify-write, reference scan order) */
void oracle_vel_update(oracle_ctx* c, double dt) {
  const double scale = dt / (c->P.rho * c->P.dx);              /* :614 */
  std::vector<float>& C = c->fg[ORACLE_CONTAINER];
  std::vector<double>&U = c->dg[ORACLE_U], &V = c->dg[ORACLE_V], &W = c->dg[ORACLE_WV];
  const int nx = c->P.nx, ny = c->P.ny, nz = c->P.nz;
  const double gx = c->P.gx * dt, gy = c->P.gy * dt, gz = c->P.gz * dt;   /* :638 */
  for (int i = 0; i < nx; ++i)
    for (int j = 0; j < ny; ++j)
      for (int k = 0; k < nz; ++k) {
        int64_t q = c->id(i, j, k);
        if (!c->isSolid(i, j, k) && C[q] > 0) {                /* :633-635 */
          double pre = c->p[c->indices[q]];                    /* :637 */
          U[q] = U[q] - scale * pre + gx;                      /* :639-642 */
          V[q] = V[q] - scale * pre + gy;
          W[q] = W[q] - scale * pre + gz;
          if (i + 1 < nx) { int64_t q2 = c->id(i + 1, j, k); U[q2] = U[q2] + scale * pre; }   /* :644-649 */
          if (j + 1 < ny) { int64_t q2 = c->id(i, j + 1, k); V[q2] = V[q2] + scale * pre; }   /* :651-656 */
          if (k + 1 < nz) { int64_t q2 = c->id(i, j, k + 1); W[q2] = W[q2] + scale * pre; }   /* :658-663 */
        }
      }
  for (int i = 0; i < nx; ++i)                                 /* :670-702 */
    for (int j = 0; j < ny; ++j)
      for (int k = 0; k < nz; ++k)
        if (c->isSolid(i, j, k)) {
          int64_t q = c->id(i, j, k);
          U[q] = 0; V[q] = 0; W[q] = 0;
          if (i + 1 < nx) U[c->id(i + 1, j, k)] = 0;
          if (j + 1 < ny) V[c->id(i, j + 1, k)] = 0;
          if (k + 1 < nz) W[c->id(i, j, k + 1)] = 0;
        }
}

/* fluid.cc:1477-1483: diver=0, rhs=0, setRHS, setDiver, setOnlyB(b2) (:543-564), error */
double oracle_relax_error(oracle_ctx* c, double dt) {
  oracle_rhs_div(c, dt);
  std::vector<float>&AD = c->fg[ORACLE_ADIAG], &D = c->fg[ORACLE_DIVER];
  for (int64_t q = 0; q < c->N; ++q)
    if (AD[q] != 0) c->b2[c->indices[q]] = D[q];
  /* Eigen (b - b2).norm() / b.norm(): sqrt of sum of squares in double */
  double num = 0, den = 0;
  for (int i = 0; i < c->numActive; ++i) {
    double d = c->b[i] - c->b2[i];
    num += d * d; den += c->b[i] * c->b[i];
  }
  return std::sqrt(num) / std::sqrt(den);
}

/* getVelocity fluid.cc:59-70 */
static inline void get_velocity(const oracle_ctx* c, int gu, int i, int j, int k, double* out) {
  out[0] = (c->dget(gu, i, j, k) + c->dget(gu, i + 1, j, k)) / 2.0;
  out[1] = (c->dget(gu + 1, i, j, k) + c->dget(gu + 1, i, j + 1, k)) / 2.0;
  out[2] = (c->dget(gu + 2, i, j, k) + c->dget(gu + 2, i, j, k + 1)) / 2.0;
}

/* PointList::FLIPadvect fluid.cc:972-1038 with CatmullRomFLIP :210-263.
 * flip_ratio r < 1 (an extension: main() only ever runs r = 1, the PIC call at :1488 is commented
 * out) blends in the PIC velocity of PointList::advect :904-971, i.e. clampedCatmullRom :125-207
 * (same window, same weights, velocity = getVelocity(node, vels), 0 when the weight sum is 0 :185):
 *   v_p = r * (v_p + delta) + (1 - r) * v_pic */
void oracle_flip_advect(oracle_ctx* c, double* dt_inout) {
  const int64_t np = (int64_t)c->px.size();
  const int ox = c->P.ox, oy = c->P.oy, oz = c->P.oz;
  const double e = c->P.restitution;                           /* :974 */
  double maxSpeed = 0.0;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) reduction(max : maxSpeed) if (c->P.threads > 1) num_threads(c->P.threads > 1 ? c->P.threads : 1)
#endif
  for (int64_t n = 0; n < np; ++n) {                           /* :978-991 */
    const double x = c->px[n], y = c->py[n], z = c->pz[n];
    int lo[3], hi[3];
    window(c, x, y, z, lo, hi);
    double weight = 0, dxs = 0, dys = 0, dzs = 0, pu = 0, pv = 0, pw = 0;
    for (int i = lo[0]; i <= hi[0]; ++i)
      for (int j = lo[1]; j <= hi[1]; ++j)
        for (int k = lo[2]; k <= hi[2]; ++k)
          if (c->inside(i, j, k)) {                            /* :237 (no solid test) */
            double velc[3], velp[3];
            get_velocity(c, ORACLE_U, i, j, k, velc);
            get_velocity(c, ORACLE_UB, i, j, k, velp);
            double cw = oracle_spline(x - (i - ox)) * oracle_spline(y - (j - oy)) * oracle_spline(z - (k - oz));
            weight += cw;
            dxs += (velc[0] - velp[0]) * cw;                   /* :252 */
            dys += (velc[1] - velp[1]) * cw;
            dzs += (velc[2] - velp[2]) * cw;
            pu += velc[0] * cw; pv += velc[1] * cw; pw += velc[2] * cw;   /* clampedCatmullRom :171-176 */
          }
    if (weight != 0) {                                         /* :258-262 */
      c->vx[n] += dxs / weight; c->vy[n] += dys / weight; c->vz[n] += dzs / weight;   /* :981 */
    }
    if (c->P.flip_ratio != 1.0) {
      const double r = c->P.flip_ratio;
      const double qu = weight != 0 ? pu / weight : 0.0, qv = weight != 0 ? pv / weight : 0.0, qw = weight != 0 ? pw / weight : 0.0;   /* :178-186 */
      c->vx[n] = r * c->vx[n] + (1.0 - r) * qu; c->vy[n] = r * c->vy[n] + (1.0 - r) * qv; c->vz[n] = r * c->vz[n] + (1.0 - r) * qw;
    }
    double len = std::sqrt(c->vx[n] * c->vx[n] + c->vy[n] * c->vy[n] + c->vz[n] * c->vz[n]);
    if (maxSpeed < len) maxSpeed = len;
  }
  double timestep;
  if (maxSpeed != 0) timestep = c->P.dt_max < c->P.dx / maxSpeed ? c->P.dt_max : c->P.dx / maxSpeed;   /* :992-999 */
  else timestep = c->P.dt_max;
  *dt_inout = timestep;
  c->last_max_speed = maxSpeed;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) if (c->P.threads > 1) num_threads(c->P.threads > 1 ? c->P.threads : 1)
#endif
  for (int64_t n = 0; n < np; ++n) {                           /* :1000-1036 */
    double x = c->px[n], y = c->py[n], z = c->pz[n];
    double nxp = x + timestep * c->vx[n], nyp = y + timestep * c->vy[n], nzp = z + timestep * c->vz[n];
    int rx = (int)round(nxp), ry = (int)round(nyp), rz = (int)round(nzp);
    if (c->isSolid(rx + ox, ry + oy, rz + oz)) {               /* :1011 */
      double sx = c->vx[n] * timestep, sy = c->vy[n] * timestep, sz = c->vz[n] * timestep;
      /* Coord(double,double,double) truncates toward zero (openvdb/math/Coord.h:61) */
      if (c->isSolid((int)round(x + sx) + ox, (int)y + oy, (int)z + oz)) c->vx[n] *= -1.0 * e;   /* :1017 */
      if (c->isSolid((int)x + ox, (int)round(y + sy) + oy, (int)z + oz)) c->vy[n] *= -1.0 * e;   /* :1021 */
      if (c->isSolid((int)x + ox, (int)y + oy, (int)round(z + sz) + oz)) c->vz[n] *= -1.0 * e;   /* :1025 */
      c->px[n] = x + c->vx[n] * timestep;                      /* :1029 */
      c->py[n] = y + c->vy[n] * timestep;
      c->pz[n] = z + c->vz[n] * timestep;
    } else {
      c->px[n] = nxp; c->py[n] = nyp; c->pz[n] = nzp;          /* :1033 */
    }
  }
}

/* PointList::interpFromGrid fluid.cc:883-895 for the particles [first, first + count) (the ones whose
 * initializedVelocities flag would be false): v = clampedCatmullRom(p, vels) :125-207 -- the window of
 * :127-136, nodes within 58 (:162), velocity getVelocity(node, vels) (:164), weights spline^3 (:178),
 * u /= weight, (0,0,0) when the weight sum is 0 (:185-197).  The min/max of :150-176 are computed by
 * the reference but never applied. */
void oracle_interp_from_grid(oracle_ctx* c, int64_t first, int64_t count) {
  const int ox = c->P.ox, oy = c->P.oy, oz = c->P.oz;
  const int64_t np = (int64_t)c->px.size();
  for (int64_t n = first; n < first + count && n < np; ++n) {
    const double x = c->px[n], y = c->py[n], z = c->pz[n];
    int lo[3], hi[3];
    window(c, x, y, z, lo, hi);
    double weight = 0, u = 0, v = 0, w = 0;
    for (int i = lo[0]; i <= hi[0]; ++i)
      for (int j = lo[1]; j <= hi[1]; ++j)
        for (int k = lo[2]; k <= hi[2]; ++k)
          if (c->inside(i, j, k)) {
            double velc[3];
            get_velocity(c, ORACLE_U, i, j, k, velc);
            double cw = oracle_spline(x - (i - ox)) * oracle_spline(y - (j - oy)) * oracle_spline(z - (k - oz));
            weight += cw;
            u += velc[0] * cw; v += velc[1] * cw; w += velc[2] * cw;
          }
    if (weight != 0) { c->vx[n] = u / weight; c->vy[n] = v / weight; c->vz[n] = w / weight; }
    else { c->vx[n] = 0; c->vy[n] = 0; c->vz[n] = 0; }
  }
}

/* the step body, fluid.cc:1378-1490 */
int oracle_step(oracle_ctx* c, double* dt_inout, oracle_stats* st) {
  oracle_stats s; memset(&s, 0, sizeof(s));
  double dt = *dt_inout;
  Clock::time_point t0 = Clock::now();
  oracle_clear_vel(c);
  oracle_p2g(c);
  Clock::time_point t1 = Clock::now();
  s.n_fluid = oracle_classify(c);
  oracle_save_vel_before(c);
  Clock::time_point t2 = Clock::now();
  s.t_p2g = secs(t0, t1) ;
  s.t_grid = secs(t1, t2);
  c->hist_err.clear(); c->hist_it.clear();
  double error = 0;
  do {
    Clock::time_point a = Clock::now();
    oracle_rhs_div(c, dt);
    Clock::time_point b = Clock::now();
    oracle_assemble(c, dt);
    Clock::time_point d = Clock::now();
    double rel = 0;
    int it = oracle_solve(c, &rel);
    Clock::time_point e = Clock::now();
    oracle_vel_update(c, dt / c->P.relax_div);
    error = oracle_relax_error(c, dt);
    Clock::time_point f = Clock::now();
    s.t_grid += secs(a, b) + secs(e, f);
    s.t_assemble += secs(b, d);
    s.t_solve += secs(d, e);
    s.passes++; s.pcg_iters_total += it; s.pcg_iters_last = it;
    c->hist_err.push_back(error); c->hist_it.push_back(it);
    if (s.passes >= c->P.max_passes) break;
  } while (error > c->P.relax_tol);                            /* :1484; inf loops, NaN exits */
  s.last_error = error;
  Clock::time_point t3 = Clock::now();
  oracle_flip_advect(c, &dt);
  Clock::time_point t4 = Clock::now();
  s.t_g2p = secs(t3, t4);
  s.dt_out = dt;
  s.max_speed = c->last_max_speed;
  *dt_inout = dt;
  if (st) *st = s;
  return 0;
}

int oracle_pass_history(oracle_ctx* c, double* errors, int32_t* iters, int32_t cap) {
  int n = (int)c->hist_err.size();
  for (int i = 0; i < n && i < cap; ++i) { errors[i] = c->hist_err[i]; iters[i] = c->hist_it[i]; }
  return n;
}

const float* oracle_fgrid(oracle_ctx* c, int which) { return c->fg[which].data(); }
double* oracle_dgrid(oracle_ctx* c, int which) { return c->dg[which].data(); }
const int32_t* oracle_indices(oracle_ctx* c) { return c->indices.data(); }
const uint8_t* oracle_solid(oracle_ctx* c) { return c->solid.data(); }
int oracle_num_active(oracle_ctx* c) { return c->numActive; }
const double* oracle_b(oracle_ctx* c) { return c->b.data(); }
const double* oracle_p(oracle_ctx* c) { return c->p.data(); }
const double* oracle_b2(oracle_ctx* c) { return c->b2.data(); }
int64_t oracle_triplets(oracle_ctx* c, const int32_t** rows, const int32_t** cols, const double** vals) {
  *rows = c->trow.data(); *cols = c->tcol.data(); *vals = c->tval.data();
  return (int64_t)c->trow.size();
}

int oracle_kat_solve_dense(int n, const double* A, const double* b, double* x) {
#ifdef ORACLE_WITH_EIGEN
  return oracle_eig_dense(n, A, b, x);
#else
  (void)n; (void)A; (void)b; (void)x;
  return -1;
#endif
}

}  // extern "C"
