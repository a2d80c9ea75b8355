/*
 * flipb200.h -- C ABI of the B200-native FLIP substep (drop-in for the step body of the
 * reference's fluid.cc).  Plain C types only: no exceptions, STL or torch types cross this
 * boundary.  Every entry point cites the reference interface it replaces (file:line relative to
 * the reference repository).  The reference has no FFI today; the seam is the call sequence of
 * its step loop, fluid.cc:1378-1490 (see INTEGRATION.md for the binding a maintainer would add).
 *
 * Conventions
 *  - all calls return FLIPB200_OK (0) or a negative status; flipb200_strerror() names it.
 *  - caller owns host buffers, the library owns device buffers.
 *  - host grid arrays are dense, idx = (k*ny + j)*nx + i (x fastest), i = x + ox, where
 *    (x,y,z) is the reference's integer coordinate (fluid.cc uses [-60,60]^3, ox = 60).
 *  - particle positions/velocities are SoA doubles in REFERENCE coordinates.
 *  - one ctx is not thread-safe; independent ctxs are.  Calls are synchronous for the caller.
 *  - there is NO CPU fallback: without a CUDA device every compute call fails with
 *    FLIPB200_ERR_CUDA.
 */
#ifndef FLIPB200_H
#define FLIPB200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define FLIPB200_VERSION 100

enum {
  FLIPB200_OK = 0,
  FLIPB200_ERR_ARG = -1,        /* bad argument                                   */
  FLIPB200_ERR_CUDA = -2,       /* CUDA runtime error (no device, OOM, launch)    */
  FLIPB200_ERR_BREAKDOWN = -3,  /* PCG breakdown: p.Ap <= 0 or non-finite scalar  */
  FLIPB200_ERR_NOCONV = -4,     /* PCG hit max iterations (solution still stored) */
  FLIPB200_ERR_STATE = -5,      /* stage called out of order                      */
  FLIPB200_ERR_COMM = -6,       /* NCCL / multi-GPU exchange error                */
  FLIPB200_ERR_LIMIT = -7       /* size beyond a compiled-in limit (flipb200_p2g_noalloc) */
};

/* POD parameters; defaults (flipb200_default_params) equal the literals of fluid.cc. */
typedef struct {
  int32_t nx, ny, nz;        /* nodes per axis; 121 (box [-60,60], fluid.cc:1159)            */
  int32_t ox, oy, oz;        /* index of reference coordinate 0; 60                           */
  int32_t wall;              /* solid shell thickness; 2 <=> |c| > 58 (fluid.cc:1264), >= 1   */
  int32_t relax_div;         /* 10: velUpdate(dt/10) (fluid.cc:1475)                          */
  double  dx, rho;           /* 1, 1 (fluid.cc:1358,1471)                                     */
  double  gx, gy, gz;        /* gravity (0,-10,0) (fluid.cc:1357)                             */
  double  dt_max;            /* 0.1 (fluid.cc:1490)                                           */
  double  relax_tol;         /* 0.1 (fluid.cc:1484)                                           */
  double  restitution;       /* e = 0 (fluid.cc:974)                                          */
  double  flip_ratio;        /* 1.0: pure FLIP, what main() runs (fluid.cc:981); r < 1 blends in
                                (1-r) x the PIC velocity of advect/clampedCatmullRom (:904-971, :125-207) */
  double  pcg_tol;           /* 2.220446049250313e-16 (Eigen IterativeSolverBase.h:363)       */
  int32_t pcg_max_iter;      /* <0 => 2*n_fluid (IterativeSolverBase.h:283)                   */
  int32_t max_passes;        /* safety cap on the do..while loop (reference: unbounded); 200  */
  int32_t precond;           /* 1 = multigrid V-cycle in fp32 (default); 0 = Jacobi           */
  int32_t device;            /* CUDA device ordinal; <0 => current device                     */
  int32_t track_ids;         /* 1: carry particle ids through the cell sort so that get_particles /
                                step_host return the caller's original order (the reference never
                                reorders its PointList); 0: arrays come back cell-sorted          */
  int32_t extrapolate;       /* 0: what main() runs (the call at fluid.cc:1147 is commented out); 1: run
                                extrapolate (fluid.cc:705-802) on the P2G velocities -- layer by layer every
                                air node gets the average of the nodes defined one layer earlier (single GPU) */
  int64_t max_particles;     /* capacity to reserve; 0 => grow on set_particles               */
  /* ---- solver options without a counterpart in the reference (round 2) ---- */
  int32_t warm_start;        /* 1 (default): CG starts from the previous pressure -- the last pass's / step's solution on the
                                cells that are still fluid, extrapolated over the relaxation passes (b_{k+1} = (1 - 1/relax_div) b_k + G
                                within a step) and scaled by the residual-minimising factor; same stop rule |r|^2 < tol^2 |b|^2.
                                0: x0 = 0, what cg.solve() does (Eigen ConjugateGradient.h:233-237)                               */
  int32_t mixed;             /* storage precision INSIDE the fp64 solve (arithmetic stays fp64 / fp32): bit 0 = CG search direction
                                stored as fp32, bit 1 = level-0 multigrid right-hand side and residual stored as fp16.
                                default 3; 0 = everything at full width.  Rows >= 128 nodes (also on z-slabs); ignored otherwise.       */
  int32_t precision;         /* 64 (default): the reference's arithmetic -- fp64 CG solved to pcg_tol = DBL_EPSILON.
                                32: "fp32 mode" (BASELINE north_star): the pressure solve stops at FLT_EPSILON relative residual
                                (pcg_tol is set to 1.1920928955078125e-07 unless the caller changed it from the default); the
                                state, the transfers and the CG recurrences stay fp64, so results differ from precision = 64 only
                                by the solve's tolerance (tests/test_gpu_parity.py::test_fp32_mode states the bound).           */
  int32_t reserved[5];
} flipb200_params;

/* per-step statistics (the reference prints none of these except dt and error) */
typedef struct {
  int32_t passes;            /* relaxation passes K of the do..while (fluid.cc:1457-1484)     */
  int32_t n_fluid;           /* numActive (fluid.cc:1395-1433)                                */
  int32_t pcg_iters_total;   /* sum of cg.iterations() over the passes                        */
  int32_t pcg_iters_last;
  double  last_error;        /* error printed at fluid.cc:1499                                */
  double  dt_out;            /* dt returned by FLIPadvect (fluid.cc:992-999)                  */
  double  max_speed;         /* maxSpeed (fluid.cc:976-991)                                   */
  double  pcg_rel_resid;     /* sqrt(|r|^2/|b|^2) of the last solve (cg.error())              */
  /* device-timed phases, milliseconds (CUDA events on the ctx stream) */
  float   ms_sort, ms_p2g, ms_grid, ms_pcg, ms_g2p, ms_total;
  int64_t launches;          /* kernels launched by this call                                 */
} flipb200_stats;

typedef struct flipb200_ctx flipb200_ctx;

void        flipb200_default_params(flipb200_params*);
const char* flipb200_strerror(int status);
int         flipb200_version(void);

/* Scene: replaces the grid construction of main(), fluid.cc:1157-1246, and the solid shell loop
 * :1256-1331.  solid_host: nx*ny*nz bytes (non-zero = solid obstacle) or NULL; the shell of
 * thickness `wall` is always made solid, as the reference does. */
int  flipb200_create (const flipb200_params*, const uint8_t* solid_host, flipb200_ctx**);
void flipb200_destroy(flipb200_ctx*);

/* ---- multi-GPU: z-slab decomposition (no counterpart in the reference, which is single-process;
 * its only parallelism is tbb::parallel_for over particles, fluid.cc:845-1126) -------------------
 * One context per GPU / process.  `params` and `solid_host` describe the WHOLE box on every rank;
 * rank r owns the planes [z0, z1) reported by flipb200_slab_range plus ghost planes.  After
 * create_slab the ranks must be connected with ONE of:
 *   flipb200_comm_init_nccl  -- one process per GPU: rank 0 calls flipb200_comm_unique_id and ships
 *                               the 128 bytes to the others (torch.distributed / MPI / a file);
 *   flipb200_comm_init_local -- all ranks are host threads of one process (tests, also on one GPU).
 * Every per-step call (step, p2g, classify, ...) is then COLLECTIVE: all ranks call it, in the same
 * order.  Particles: set_particles takes the whole scene on every rank and keeps what the rank
 * owns; seed_box seeds the rank's part of a global box; get_particles / get_grid return the
 * rank's particles / owned planes.  step_host is not available on slab contexts. */
int  flipb200_create_slab(const flipb200_params*, const uint8_t* solid_host, int32_t rank, int32_t world, flipb200_ctx**);
/* create_slab with slab boundaries chosen by the caller (static load balancing along z): zsplit[0..world], 0 = zsplit[0] < ... <
 * zsplit[world] = nz, interior boundaries multiples of 2^a (a >= 1), every slab >= 2 * 2^a planes; same array on every rank */
int  flipb200_create_slab_split(const flipb200_params*, const uint8_t* solid_host, int32_t rank, int32_t world,
                                const int32_t* zsplit, flipb200_ctx**);
int  flipb200_slab_range(flipb200_ctx*, int32_t* z0, int32_t* z1);
int  flipb200_comm_unique_id(void* id128);                         /* ncclGetUniqueId, 128 bytes */
int  flipb200_comm_init_nccl(flipb200_ctx*, const void* id128);    /* ncclCommInitRank(world, id, rank) */
int  flipb200_comm_init_local(flipb200_ctx*, int32_t group_id);
int  flipb200_get_particle_ids(flipb200_ctx*, uint32_t* ids, int64_t cap);   /* needs params.track_ids */
/* how the PCG ghost planes travel, known after the first solve: 0 = separate copies (NCCL send/recv),
 * 1 = pushed by the producing kernels + host barrier (in-process ranks), 2 = pushed over NVLink peer
 * mappings with device-side flags (one process per GPU; csrc/halo.cuh) */
int  flipb200_halo_mode(flipb200_ctx*);
/* diagnostics: out8 = { halo mode (+10: dot products exchanged over peer mailboxes, +20: first replicated
 * multigrid level gathered by peer stores), iteration graph active, multigrid levels, distributed levels,
 * kernels per PCG iteration, ghost-pushing kernels per iteration, rank, world } */
int  flipb200_debug_info(flipb200_ctx*, int32_t* out8);
/* flipb200_step_host for a slab (collective): the rank uploads ITS n particles (every base plane
 * must be owned by the rank or an adjacent one), steps, and gets back the *n_out particles it owns
 * after migration (host arrays hold `cap` entries) and, optionally, its owned planes of outputGrid. */
int  flipb200_step_host_slab(flipb200_ctx*, double* px, double* py, double* pz, double* vx, double* vy, double* vz,
                             int64_t n, int64_t cap, int64_t* n_out, double* dt_inout, float* output_grid,
                             flipb200_stats*);

/* PointList state (fluid.cc:804-813, add :841).  Host SoA arrays are copied. */
int  flipb200_set_particles(flipb200_ctx*, const double* px, const double* py, const double* pz,
                            const double* vx, const double* vy, const double* vz, int64_t n);
int  flipb200_get_particles(flipb200_ctx*, double* px, double* py, double* pz,
                            double* vx, double* vy, double* vz, int64_t cap, int64_t* n);
int  flipb200_num_particles(flipb200_ctx*, int64_t* n);
/* Device-side seeding (the role of UniformPointScatter at fluid.cc:1348-1350, stratified):
 * exactly 8 particles per voxel of the index box [lo,hi] (inclusive), zero velocity. */
int  flipb200_seed_box(flipb200_ctx*, const int32_t lo[3], const int32_t hi[3], uint64_t seed);
/* Re-seeding hook (commented out in the reference driver, fluid.cc:1374-1382 and :1495-1497): appends
 * 8 particles per voxel of the box like flipb200_seed_box and gives each new particle the PIC
 * velocity of PointList::interpFromGrid (fluid.cc:883-895, clampedCatmullRom :125-207) gathered from
 * the current grid velocities (those of the last step / p2g + classify).  Collective on z-slab runs. */
int  flipb200_emit_box(flipb200_ctx*, const int32_t lo[3], const int32_t hi[3], uint64_t seed);

/* One time step == the loop body fluid.cc:1378-1490 with device-resident particles.
 * dt_inout: in = dt used by the solve (fluid.cc:1469-1475), out = FLIPadvect's dt (:1490). */
int  flipb200_step(flipb200_ctx*, double* dt_inout, flipb200_stats*);
/* Same step through HOST buffers (what the reference's main() holds): uploads the six
 * particle arrays, steps, downloads them and (optionally) the fp32 output grid
 * (outputGrid, fluid.cc:1434-1451; nx*ny*nz floats or NULL). */
int  flipb200_step_host(flipb200_ctx*, double* px, double* py, double* pz,
                        double* vx, double* vy, double* vz, int64_t n,
                        double* dt_inout, float* output_grid, flipb200_stats*);

/* Stage-level entry points (one per seam of the reference's step loop; used by parity tests). */
int  flipb200_p2g(flipb200_ctx*);                                /* vels.fill(0) :1378 + PointList::P2Gtransfer :1106-1148 (+ interpolate :843-882 fused) */
int  flipb200_classify(flipb200_ctx*, int32_t* n_fluid);         /* main :1388-1448 (+ velBeforeUpdate copy :1455) */
int  flipb200_rhs_div(flipb200_ctx*, double dt, double* b_norm2);/* setRHS :414-479 + setDiver :566-610 (+ b of setA2 :535) */
int  flipb200_pcg_solve(flipb200_ctx*, double dt, int32_t* iters, double* rel_resid); /* setA :304-412, setA2 :481-541, cg.compute/solve :1473-1474 */
int  flipb200_vel_update(flipb200_ctx*, double dt_sub);          /* velUpdate :612-703 */
int  flipb200_relax_error(flipb200_ctx*, double dt, double* error); /* :1477-1483 (setRHS, setDiver, setOnlyB, norm ratio) */
int  flipb200_g2p_advect(flipb200_ctx*, double* dt_out);         /* PointList::FLIPadvect :972-1038 */

/* Grid access for tests / output.  Double grids: 8 bytes per node; W/B float; FLAGS uint8. */
enum {
  FLIPB200_GRID_U = 0, FLIPB200_GRID_V = 1, FLIPB200_GRID_W = 2,     /* vels (Vec3dGrid, fluid.cc:1240), double */
  FLIPB200_GRID_UB = 3, FLIPB200_GRID_VB = 4, FLIPB200_GRID_WB = 5,  /* velBeforeUpdate (:1455), double          */
  FLIPB200_GRID_P = 6,                                               /* pressure scattered to the grid, double   */
  FLIPB200_GRID_WEIGHT = 7,                                          /* weights == containerGrid == outputGrid off solids, float (:809,1157,1161) */
  FLIPB200_GRID_B = 8,                                               /* diver (= b), float (:1231)               */
  FLIPB200_GRID_FLAGS = 9                                            /* uint8: bit0..5 solid nbr (-x,+x,-y,+y,-z,+z), bit6 fluid row, bit7 solid */
};
int  flipb200_get_grid(flipb200_ctx*, int which, void* host_out, size_t bytes);
int  flipb200_set_grid(flipb200_ctx*, int which, const void* host_in, size_t bytes);

/* Micro-benchmark hooks (bench.py roofline lines): run `iters` PCG iterations / kernels on the
 * current system without convergence exit and return the average device time per launch. */
enum { FLIPB200_KERNEL_SPMV = 0, FLIPB200_KERNEL_PCG_ITER = 1, FLIPB200_KERNEL_P2G = 2,
       FLIPB200_KERNEL_G2P = 3, FLIPB200_KERNEL_SORT = 4 };
int  flipb200_time_kernel(flipb200_ctx*, int which, int32_t iters, float* ms_per_launch, int64_t* units);

/* Stage-by-stage device times of ONE PCG iteration (plain launches with a CUDA event between the stages,
 * averaged over `iters` iterations of the current system, no convergence exit): names[m] (FLIPB200_PROFILE_NAME
 * bytes each, e.g. "k_spmv", "L0 halo+smooth", "tail (small levels + coarsest)") and us[m], m < *n <= cap.
 * A stage name can occur more than once (levels >= 2 share one name).  bench.py's roofline.by_kernel. */
#define FLIPB200_PROFILE_NAME 40
int  flipb200_profile_iteration(flipb200_ctx*, int32_t iters, int32_t cap, char* names, float* us, int32_t* n);

/* NO COUNTERPART IN THE REFERENCE.
 * The particle -> grid transfer of flipb200_p2g + the normalisation of flipb200_classify as ONE call that allocates
 * nothing: no cudaMalloc / cudaFree or allocating library routine, for callers that cannot allocate (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  Its only scratch, the fp64 weight accumulator, is
 * a static __device__ array of FLIPB200_NOALLOC_MAX_NODES doubles (16 MiB) that every GPU loading the library holds
 * until exit; grids with nx*ny*nz above that limit are rejected with FLIPB200_ERR_LIMIT (message on stderr).
 * All pointers are DEVICE pointers owned by the caller, dense [z][y][x] grids of nx*ny*nz elements:
 *   in : solid_dev (bytes, non-zero = obstacle; NULL = none; the shell of thickness `wall` is always solid),
 *        px..vz_dev (n particles, SoA, any order)
 *   out: u_dev, v_dev, w_dev  P2G velocities, divided by the float-rounded weight where it is > 0 (= GRID_U..W after
 *        flipb200_classify), weight_dev the FloatGrid `weights` (= GRID_WEIGHT); the call clears them first.
 * Work is queued on `stream` (a cudaStream_t; NULL = the default stream) of the CURRENT device; the call does not
 * synchronise.  Every call shares the one scratch array: calls that overlap on the GPU (different streams without a
 * dependency) race on it -- serialise them.  The first call on a device looks up the array's address
 * (cudaGetSymbolAddress); make it before starting a graph capture. */
#define FLIPB200_NOALLOC_MAX_NODES 2097152   /* 128^3 */
int  flipb200_p2g_noalloc(int32_t nx, int32_t ny, int32_t nz, int32_t ox, int32_t oy, int32_t oz, int32_t wall,
                          const uint8_t* solid_dev,
                          const double* px_dev, const double* py_dev, const double* pz_dev,
                          const double* vx_dev, const double* vy_dev, const double* vz_dev, int64_t n,
                          double* u_dev, double* v_dev, double* w_dev, float* weight_dev, void* stream);

/* Pinned host memory helpers for the e2e path (cudaHostAlloc / cudaFreeHost). */
int  flipb200_host_alloc(void** ptr, size_t bytes);
int  flipb200_host_free(void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* FLIPB200_H */
