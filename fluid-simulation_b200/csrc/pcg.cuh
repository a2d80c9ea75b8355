// pcg.cuh -- coefficient tables shared by pcg.cu and mg.cu
#pragma once
#include "common.cuh"

struct PcgCoef {
  double diag[7];   // Adiag for k = 0..6 non-solid neighbours (fp32-accumulated as in setA, widened)
  double inv[7];    // Jacobi: 1/diag
  double off;       // (double)(float)(-scale)
  float winv[7];    // multigrid level 0: omega / (k * h2inv)  (first damped-Jacobi sweep from x = 0)
};

// damped-Jacobi weights of the V-cycle: pre-smoothing uses (W1, W2), post-smoothing the reversed
// order (W2, W1), which keeps M^-1 symmetric.  Chebyshev-like pair found by search on the tank and
// column scenes (tools/mg_proto5.py): PCG convergence rate 0.167 -> 0.124 against (0.8, 0.8).
#define MG_OMEGA1 0.6f
#define MG_OMEGA2 1.7f

// per-level multigrid coefficients (fp32)
struct MgCoef {
  float h2inv;      // operator scale of the level
  float winv[7];    // omega / (k * h2inv)
};
static inline MgCoef make_mg_coef(double h2inv, double omega) {
  MgCoef m; m.h2inv = (float)h2inv; m.winv[0] = 0.f;
  for (int k = 1; k <= 6; ++k) m.winv[k] = (float)(omega / (k * h2inv));
  return m;
}

// kernels take the iteration number as an argument, or IT_DEV to read it from ds->pcg_it (graph replay)
#define IT_DEV (-1000)
__device__ __forceinline__ int pcg_iter(const DevScalars* ds, int it_arg) { return it_arg == IT_DEV ? ds->pcg_it : it_arg; }

// params.mixed bit 1 (default ON since round 2: measured -26 us per iteration at 256^3, profiles/r2a_variants_ab.txt): the level-0 right-hand side b0 and the
// level-0 residual handed to the restriction are STORED as fp16 (arithmetic stays fp32).  Prototype
// (tools/mg_proto7.py, full tank 128^3): 21 CG iterations either way, with r.z taken from the stored b0
// as the fused dot does; fp16 storage of the iterates x costs 2 iterations, bf16 of anything more.
// Saves 14 of the ~88 B/node the level-0 part of an iteration moves.  fp16 needs O(1) entries: the
// scaled residual (|b0| = 1, entries ~ 1/sqrt(N)) is multiplied by bscale = 2^floor(log2 sqrt(N)) / 4; the
// V-cycle is linear, k_update_p4 divides z by it again and the fused r.z by its square.
// Four-node sweeps only (rows >= 128 nodes); the direction update runs as k_update_p4.  z-slab runs: the fp16 residual's
// ghost planes travel as fp16 (halo_push4h).
static inline bool mg_h16_active(const flipb200_ctx* c) {
  if (!(c->P.mixed & 2) || c->mg_levels < 2) return false;
  if (!c->opt_st4) return false;
  const int x0 = c->box.x0 & ~3, xe = min(c->g.px, (c->box.x1 + 3) & ~3);      // make_sweep(g, box, 4)
  return xe - x0 >= 128;
}
static inline double mg_h16_bscale(const flipb200_ctx* c) {
  // the GLOBAL sweep box: every rank of a z-slab run must scale the one global vector by the same power of two
  const double n = (double)(c->gbox.x1 - c->gbox.x0) * (c->gbox.y1 - c->gbox.y0) * (c->gbox.z1 - c->gbox.z0);
  int e = 0;
  while ((double)(1ll << (2 * (e + 1))) <= n && e < 14) ++e;     // 2^e <= sqrt(n)
  // a quarter of that: typical entries ~0.25, and |r| may grow by an order of magnitude against the
  // |r_prev| it is scaled with (CG residuals are not monotone) before a lone entry nears fp16's 65504
  return (double)(1 << (e > 2 ? e - 2 : 0));
}
// mg_h16.cu: host wrappers of the fp16-storage copies of the level-0 sweeps
struct Sweep;
struct PeerHalo;
struct GatherPush;
void fb_h16_smooth4(flipb200_ctx* c, int dot, dim3 g, const Sweep& S, const MgCoef& m, const uint8_t* fl, const float* b0,
                    const float* xi, float* xo, int it, float rzs, int dist, const PeerHalo& H);
void fb_h16_resid4(flipb200_ctx* c, dim3 g, const Sweep& S, const MgCoef& m, const uint8_t* fl, const uint8_t* nrm, const float* b0,
                   const float* x, float* rn, const PeerHalo& H);
void fb_h16_restrict4(flipb200_ctx* c, dim3 g, dim3 blk, const Grid& gf, const Box& fbox, const Grid& gc, const Box& cbox, const MgCoef& mcc,
                      const uint8_t* fc, const float* rn, float* bc, float* x1c, int zc, const PeerHalo& H, const GatherPush& G);

// params.mixed bit 0 (default ON since round 2: measured -32 us per iteration, same iteration counts; pcg_p32.cu): the CG search direction stored as fp32.
// Same conditions as the fp16 variant: multigrid on, four-node sweeps; on z-slabs the ghost planes of p travel as fp32.
static inline bool pcg_p32_active(const flipb200_ctx* c) {
  if (!(c->P.mixed & 1) || c->mg_levels < 2) return false;
  if (!c->opt_st4) return false;
  const int x0 = c->box.x0 & ~3, xe = min(c->g.px, (c->box.x1 + 3) & ~3);      // make_sweep(g, box, 4)
  return xe - x0 >= 128;
}
struct PcgCoef;
void fb_p32_spmv(flipb200_ctx* c, const PcgCoef& cf, int it_arg, int noexit, int dist, const PeerHalo& H);
void fb_p32_update_p(flipb200_ctx* c, const float* zf, int it_arg, int noexit, double inv_bscale, bool down, const PeerHalo& H);
void fb_p32_finish_x(flipb200_ctx* c);
