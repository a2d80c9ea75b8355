/*
 * flip_oracle.h -- CPU ORACLE for the FLIP substep of the reference's fluid.cc.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product library
 * (libflipb200.so) never links, loads or calls anything in oracle/.
 *
 * What it is: a dense-array restatement of the hot-path functions of
 * /root/reference/fluid.cc (spline :22-37, getVelocity :59-70, CatmullRomFLIP :210-263,
 * p2gCatmullRom :265-299, setA :304-412, setRHS :414-479, setA2 :481-541, setOnlyB :543-564,
 * setDiver :566-610, velUpdate :612-703, PointList::interpolate :843-882,
 * PointList::FLIPadvect :972-1038, PointList::P2Gtransfer :1106-1148, main step body
 * :1378-1490).  The pressure solve is NOT restated when built with ORACLE_WITH_EIGEN: it
 * calls the reference's own vendored Eigen 3.3.4
 * ConjugateGradient<SparseMatrix<double>, Lower|Upper, IncompleteCholesky<double>>
 * exactly as fluid.cc:1352,1473-1474 does (see eigen_solver.cc).
 *
 * Parity pinning: PINNED BY THE REFERENCE'S OWN CODE.  The reference has no tests or golden vectors
 * for fluid.cc, so the pin is the reference itself: /root/reference/fluid.cc compiles UNMODIFIED (main
 * renamed) over the dense-array OpenVDB/TBB stand-in of oracle/shim/ with the reference's vendored Eigen
 * (oracle/ref_bridge.cc -> oracle/_ref/libfluid_ref.so), and tests/test_ref_pin.py checks BIT FOR BIT that
 * (a) every function of the step body (P2Gtransfer, interpolate, setRHS, setDiver, setA, setA2 + the matrix
 * handed to Eigen, cg.solve, velUpdate, setOnlyB, FLIPadvect, extrapolate, advect, interpFromGrid) returns
 * what this restatement returns on three adversarial 121^3 scenes (walls, interior obstacle, strays outside
 * the box, round() ties), and (b) the reference's own main() run for three steps of its own loop on its own
 * scene ends in the same particles, grid velocities, cell numbering, pass counts and dt as oracle_step x 3.
 * Also: the known-answer tests the reference tree holds next to the path
 * (openvdb/unittest/TestConjGradient.cc:57-101, TestPoissonSolver.cc:215-290) and the hand analysis of
 * SURVEY.md 3.2 (tests/test_oracle.py).  Only the scatterer's voxel visiting order (VDB tree order) is not
 * reproduced: initial positions are distribution-equivalent (irrelevant to parity, both sides get the same arrays).
 *
 * Layout: grid arrays are flat, idx = (k*ny + j)*nx + i with i = x + ox etc. (x fastest);
 * reference coordinate c = (i-ox, j-oy, k-oz).  Loops that carry sequential semantics
 * (setA, velUpdate, cell numbering) run in the reference's scan order x-outer, y, z-inner.
 * FloatGrid-typed grids are float arrays, so every store rounds to fp32 like the reference.
 */
#ifndef FLIP_ORACLE_H
#define FLIP_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int32_t nx, ny, nz;     /* nodes per axis (121 in fluid.cc)                         */
  int32_t ox, oy, oz;     /* index of reference coordinate 0 (60 in fluid.cc)        */
  int32_t wall;           /* solid shell thickness: 2 <=> |c| > 58  (fluid.cc:1264)   */
  int32_t relax_div;      /* 10  (fluid.cc:1475, dt/10)                               */
  double dx, rho;         /* 1, 1 (fluid.cc:1358,1471)                                */
  double gx, gy, gz;      /* (0,-10,0) (fluid.cc:1357)                                */
  double dt_max;          /* 0.1 (fluid.cc:1490)                                      */
  double relax_tol;       /* 0.1 (fluid.cc:1484)                                      */
  double restitution;     /* e = 0 (fluid.cc:974)                                     */
  int32_t solver;         /* 0 = reference Eigen IC-PCG; 1 = oracle's own Jacobi-PCG  */
  int32_t max_passes;     /* safety cap on the do..while loop (reference has none)    */
  int32_t threads;        /* 1 = serial particle loops (parity); >1 OpenMP (timing)   */
  int32_t pcg_max_iter;   /* <0 => 2n (IterativeSolverBase.h:283)                     */
  double  flip_ratio;     /* 1 = pure FLIP (fluid.cc:981, what main() runs); 0 = PIC of advect (:904-971) */
  int32_t extrapolate;    /* 0 = what main() runs (the call at fluid.cc:1147 is commented out); 1 = run extrapolate :705-802 at the end of P2Gtransfer */
  int32_t reserved0;
} oracle_params;

typedef struct {
  int32_t passes, n_fluid, pcg_iters_total, pcg_iters_last;
  double last_error, dt_out, max_speed;
  double t_p2g, t_grid, t_assemble, t_solve, t_g2p; /* seconds, steady_clock */
} oracle_stats;

typedef struct oracle_ctx oracle_ctx;

void oracle_default_params(oracle_params* p);
int  oracle_has_eigen(void);
oracle_ctx* oracle_create(const oracle_params* p, const uint8_t* solid /* nx*ny*nz or NULL => shell only */);
void oracle_destroy(oracle_ctx*);
void oracle_set_particles(oracle_ctx*, const double* px, const double* py, const double* pz,
                          const double* vx, const double* vy, const double* vz, int64_t n);
int64_t oracle_num_particles(oracle_ctx*);
void oracle_get_particles(oracle_ctx*, double* px, double* py, double* pz,
                          double* vx, double* vy, double* vz);

/* scalar kernels */
double oracle_spline(double x);                                   /* fluid.cc:22-37 */

/* stages (mirror the step-loop call sequence, fluid.cc:1378-1490) */
void oracle_clear_vel(oracle_ctx*);                               /* :1378 */
void oracle_p2g(oracle_ctx*);                                     /* :1384 -> :1106-1148 */
void oracle_extrapolate(oracle_ctx*);                             /* :1147 (commented out) -> :705-802 */
int  oracle_classify(oracle_ctx*);                                /* :1388-1448 (container, indices, output) -> numActive */
void oracle_save_vel_before(oracle_ctx*);                         /* :1455 */
void oracle_rhs_div(oracle_ctx*, double dt);                      /* :1465-1470 */
void oracle_assemble(oracle_ctx*, double dt);                     /* :1461-1464,1471-1472 (setA+setA2 -> A, b) */
int  oracle_solve(oracle_ctx*, double* rel_err);                  /* :1473-1474 -> iterations */
void oracle_vel_update(oracle_ctx*, double dt_sub);               /* :1475 -> :612-703 */
double oracle_relax_error(oracle_ctx*, double dt);                /* :1477-1483 */
void oracle_flip_advect(oracle_ctx*, double* dt_inout);           /* :1490 -> :972-1038 */
void oracle_interp_from_grid(oracle_ctx*, int64_t first, int64_t count);   /* :883-895 (re-seeding hook :1495-1497) */
int  oracle_step(oracle_ctx*, double* dt_inout, oracle_stats*);   /* whole body :1378-1490 */

/* per-pass history of the last oracle_step (errors, iterations); returns #passes */
int  oracle_pass_history(oracle_ctx*, double* errors, int32_t* iters, int32_t cap);

/* raw views (owned by ctx, valid until destroy); N = nx*ny*nz */
enum { ORACLE_W = 0, ORACLE_CONTAINER, ORACLE_RHS, ORACLE_DIVER, ORACLE_ADIAG,
       ORACLE_APLUSI, ORACLE_APLUSJ, ORACLE_APLUSK, ORACLE_OUTPUT };        /* float grids  */
enum { ORACLE_U = 0, ORACLE_V, ORACLE_WV, ORACLE_UB, ORACLE_VB, ORACLE_WB, ORACLE_PGRID }; /* double grids */
const float*   oracle_fgrid(oracle_ctx*, int which);
double*        oracle_dgrid(oracle_ctx*, int which);     /* mutable: tests may inject fields */
const int32_t* oracle_indices(oracle_ctx*);
const uint8_t* oracle_solid(oracle_ctx*);
int            oracle_num_active(oracle_ctx*);
const double*  oracle_b(oracle_ctx*);       /* numActive */
const double*  oracle_p(oracle_ctx*);       /* numActive */
const double*  oracle_b2(oracle_ctx*);      /* numActive: setOnlyB's vector (fluid.cc:1481) */
/* the triplet list setA2 hands to setFromTriplets (fluid.cc:498-540), in push order; returns its length */
int64_t oracle_triplets(oracle_ctx*, const int32_t** rows, const int32_t** cols, const double** vals);

/* Known-answer helper: solve a small dense SPD system with the reference's solver type
 * (used with openvdb/unittest/TestConjGradient.cc:66-92).  Returns iterations, <0 if no Eigen. */
int oracle_kat_solve_dense(int n, const double* A_rowmajor, const double* b, double* x);

#ifdef __cplusplus
}
#endif
#endif
