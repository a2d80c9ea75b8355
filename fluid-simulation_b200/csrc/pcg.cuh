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
