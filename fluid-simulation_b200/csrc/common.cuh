// common.cuh -- shared definitions of the B200-native FLIP substep library (libflipb200.so).
// Product code: never includes or links anything under oracle/.
#pragma once
#include <stdlib.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>
#include "flipb200.h"

// ---------------------------------------------------------------------------------------------
// Grid layout in HBM.  All node arrays share one geometry: x fastest with a padded row pitch
// `px` (multiple of 32 elements => every row of an 8-byte array starts 256-byte aligned, rows of
// the 1-byte flag array 32-byte aligned), then y, then z:  idx = (k*ny + j)*px + i.
// z is the slowest index so that z-slabs and ghost planes are contiguous (multi-GPU split).
//
// Multi-GPU: the box is cut into z-slabs.  A rank's arrays hold its owned planes
// [k0, k1) plus SLAB_GHOST ghost planes on either side (also past the ends of the global box,
// where they are inert), so every kernel indexes a slab exactly like a whole grid; `kg0` is the
// global z index of local plane 0 and `gnz` the global plane count.  Single GPU: kg0 = 0,
// gnz = nz, k0 = 0, k1 = nz.
// ---------------------------------------------------------------------------------------------
#define SLAB_GHOST 2
struct Grid {
  int nx, ny, nz;   // nodes per axis held in the local arrays (nz includes ghost planes)
  int px;           // row pitch in elements
  int ox, oy, oz;   // local index of reference coordinate 0
  int wall;         // solid shell thickness
  int plane;        // px*ny
  int total;        // px*ny*nz  (< 2^31, checked at create)
  int kg0;          // global z index of local plane 0 (negative on rank 0 of a slab split)
  int gnz;          // global number of z planes
  int k0, k1;       // owned local planes [k0, k1)
};

__host__ __device__ __forceinline__ int gidx(const Grid& g, int i, int j, int k) {
  return (k * g.ny + j) * g.px + i;
}
// inside the GLOBAL box and inside the local arrays
__host__ __device__ __forceinline__ bool in_box(const Grid& g, int i, int j, int k) {
  return i >= 0 && j >= 0 && k >= 0 && i < g.nx && j < g.ny && k < g.nz && k + g.kg0 >= 0 && k + g.kg0 < g.gnz;
}
// isWithinBounds(c, 58) of fluid.cc:38-45 in index space: inside the solid shell
__host__ __device__ __forceinline__ bool inside_shell(const Grid& g, int i, int j, int k) {
  return i >= g.wall && j >= g.wall && k + g.kg0 >= g.wall && i <= g.nx - 1 - g.wall &&
         j <= g.ny - 1 - g.wall && k + g.kg0 <= g.gnz - 1 - g.wall;
}

// flag byte per node
enum : uint8_t {
  FB_SXM = 1, FB_SXP = 2, FB_SYM = 4, FB_SYP = 8, FB_SZM = 16, FB_SZP = 32,  // neighbour is solid
  FB_FLUID = 64,   // fluid cell == row of the pressure system (fluid.cc:1423-1428)
  FB_SOLID = 128   // cell itself is solid (fluid.cc:46-57)
};
#define FB_SMASK 63

// A sub-box of the grid that kernels sweep (bounding box of the fluid cells +-1, x-aligned)
struct Box { int x0, y0, z0, x1, y1, z1; };   // half-open

// Scalars that live in device memory (one struct per ctx) -- reductions, PCG control, dt.
struct DevScalars {
  double pAp[4], rz[4], rr[4];       // PCG dot products, ring indexed by iteration & 3
  double bnorm2;                     // |b|^2 of the current right-hand side
  double err_num, err_den;           // |b-b2|^2, |b|^2 of the relaxation test (fluid.cc:1483)
  double threshold;                  // tol^2 * |b|^2 (Eigen ConjugateGradient.h:53)
  double rr_final;                   // |r|^2 at exit
  unsigned long long max_speed_bits; // max |v_p| as ordered bits (fluid.cc:976-991)
  double dt_out;                     // FLIPadvect's dt (fluid.cc:992-999)
  int n_fluid;
  int bbox[6];                       // min x,y,z / max x,y,z of fluid cells
  int pcg_done, pcg_iters, pcg_status;
  int pcg_max_iter;                  // iteration limit of the current solve
  int pcg_it;                        // current iteration (kernels launched from a CUDA graph read it here)
  int n_invalid;                     // particles whose support lies outside the box
  int x_pending;                     // the stop test fired before k_update_p<1> applied x += alpha p (k_pcg_finish_x)
  double warm_gamma;                 // warm start: x0 = gamma * guess (residual-minimising scale)
  double warm_rr0;                   // |b - A x0|^2 as accumulated by the warm init (exact)
  // z-slab runs: kernels accumulate this rank's share of the current dot products here; after the
  // all-reduce k_pcg_publish moves the sum into the ring slot (single GPU: straight into the ring)
  double part[4];                    // [0] p.Ap  [1] r.z  [2] r.r
  int halo_base;                     // sequence number base of the fused ghost-plane pushes (halo.cuh)
  int pad2;
  unsigned blk_cnt[4];               // fused all-reduce (halo.cuh): finished blocks of the current dot-product kernel
  unsigned red_seq[4];               //   and the sequence number of the last collected sum, per PART_* id
  int n_fluid_part[64];              // k_classify's per-block counts, spread over 64 words (k_classify_sum adds them up)
};
#define PART_PAP 0
#define PART_RZ 1
#define PART_RR 2

struct Particles {
  double* a[6];      // px,py,pz,vx,vy,vz  (current, cell-sorted after sort)
  double* b[6];      // alternate buffers (sort destination)
  uint32_t* key;     // cell key of each particle (base cell of its 2x2x2 support)
  uint32_t* rank;    // arrival rank inside the cell
  uint32_t* id;      // original index of each particle (only with params.track_ids)
  uint32_t* id_b;
  int64_t n, cap;
};

// One level of the geometric multigrid hierarchy.  The V-cycle runs in fp32 (it only preconditions
// the fp64 CG); level 0 shares the fine flags.
struct MgLevel {
  Grid g;             // z-slab runs: a slab with ghost planes on the distributed levels, the whole grid on the replicated ones
  bool dist;          // level is slab-distributed (needs halo fills); false: every rank holds the whole level
  int zs[17];         // first global plane of every rank at this level (+ end)
  Box gbox;           // sweep box over the whole domain (z in global planes of the level)
  Box boxe;           // `box` extended into the ghost planes the global box covers (input range of the restriction)
  Box box;            // sweep box of this step at this level (local planes)
  uint8_t* flags;     // same encoding as the fine flags (FB_*)
  uint8_t* nrm;       // 64 x (sum of the prolongation weights of the non-solid coarse donors), 27..64
  float *xa, *xb, *b; // solution ping-pong and right-hand side
  double h2inv;       // operator scale: A x = h2inv * (k x_c - sum of fluid neighbours)
};

#define MG_MAX_LEVELS 12
// FLIPB200_PROFILE_ITER=1: stage-by-stage CUDA-event timing of the PCG iteration inside
// flipb200_time_kernel(PCG_ITER) (plain launches, one stream sync per iteration); printed to stderr.
#define PROF_MAX 96
struct IterProf {
  bool on;
  int n;                       // marks of the current iteration
  cudaEvent_t ev[PROF_MAX];
  const char* name[PROF_MAX];
  double acc[PROF_MAX];
  int iters;
  bool request;                // flipb200_profile_iteration: profile the next fb_pcg_time(PCG_ITER) and keep the result below
  int out_n;                   // stages of the last profiled run
  float out_us[PROF_MAX];
  const char* out_name[PROF_MAX];
};
struct Comm;   // comm.cu: NCCL (one process per GPU) or in-process threaded ranks (tests)

// host state of the ghost-plane pushes fused into the PCG kernels (halo.cuh, comm.cu)
enum { HA_XA = 0, HA_XB = 1, HA_PD = 2 };
#define HALO_DBG_MAX 64
struct HaloState {
  unsigned long long* dbg;   // FLIPB200_HALO_DEBUG=1: 8 timestamps per fb_halo_make call of an iteration
  int dbg_n;
  int mode;                          // 0 off (halo copies), 1 push + host barrier (in-process ranks), 2 push + device flags
  int acq_sys;                       // consumer acquires at system scope (default 1; FLIPB200_HALO_ACQ=gpu -> 0), halo.cuh
  unsigned* flags;                   // device: [0..1] flags written by the lower / upper neighbour, [2..3] block counters
  unsigned* peer_flags[2];           // the neighbours' flag blocks (peer mappings): [0] lower, [1] upper
  void* peer[2][MG_MAX_LEVELS][3];   // the neighbours' arrays: [dir][level][HA_*]
  long long abs_seq;                 // producer kernels enqueued so far
  long long seg_base;                // host mirror of DevScalars::halo_base for the segment being enqueued
  long long last_prod[MG_MAX_LEVELS][3];
  // all-reduce fused into the dot-product kernels (mode 2 only): every rank's mailbox, peer-mapped
  struct RedMail* mail;              // own
  struct RedMail* peer_mail[16];     // [rank] (own included); null when the fused all-reduce is off
  int red_on;
  // all-gather of the first replicated level by peer stores (halo.cuh GatherPush)
  int gather_on;
  float* gather_b[16];               // [rank] that rank's b / xa array of level mg_nd
  float* gather_xa[16];
  long long gather_seq;              // sequence number of the latest gather
};

struct flipb200_ctx {
  flipb200_params P;
  Grid g;                  // local grid (a z-slab with ghost planes when world > 1)
  int rank, world;         // z-slab decomposition; world == 1: the whole box
  int gz0, gz1;            // owned global planes [gz0, gz1)
  int zsplit[17];          // global first plane of every rank (+ end), level 0
  Comm* comm;
  int slab_align;          // log2 of the alignment of the slab boundaries (mg.cu: distributed levels)
  double* halo_tmp;        // 4 planes of doubles: receive buffer of the P2G ghost sum
  HaloState halo;
  IterProf* prof;
  bool halo_ready;         // fb_halo_setup has run (lazily, at the first collective solve)
  int device;
  int num_sms;
  cudaStream_t stream;
  // node arrays
  uint8_t* solid;          // input solid mask incl. shell (1 byte)
  uint8_t* flags;          // FB_* (static solid part + per-step fluid bit)
  uint8_t* flags_static;   // solid part only
  int32_t* ext_layer;      // params.extrapolate: wave that defined each node (+ a counter behind it), lazily allocated
  double *u[3];            // vels; during P2G: momentum accumulators
  double *ub[3];           // velBeforeUpdate; after the relaxation loop: cell-centred FLIP delta
  double *wacc;            // weight accumulator (fp64), rounded once to fp32 into wf
  float  *wf;              // weights == containerGrid
  float  *bf[2];           // diver (= b) ping-pong: [cur] is b of this pass, [cur^1] receives b2
  int     bcur;
  double *x, *r, *pd, *Ap; // PCG vectors (x = pressure on the grid)
  double *xprev;           // warm start: the solution of the pass before the last one (params.warm_start)
  double *xfirst;          // warm start: the first pass's solution of the previous step (guess of the next step's first pass)
  double *xG;              // warm start: A^-1 G of the previous step (x_2 - th x_1), guess of the second pass: x_2 = th x_1 + xG
  bool have_first, have_G; // xfirst / xG hold data
  int last_step_passes;    // passes of the last finished step (0: unknown): 1 => x is that step's first solution, 2 => (xprev, x) = (x_1, x_2)
  bool x_valid;            // c->x holds a converged solution (0 off the current fluid cells: k_classify keeps that invariant)
  int pass_idx;            // relaxation pass of the solve about to run (0 = first of the step; set by the step loop)
  int xprev_pass;          // pass whose solution xprev holds (-1: none in this step)
  double last_rate;        // decimal digits of |r|^2 gained per iteration in the previous solve (sizes the first launch batch)
  MgLevel mg[MG_MAX_LEVELS];
  int mg_levels;           // 0 => hierarchy not allocated (precond != 1)
  int mg_nd;               // levels [0, mg_nd) are slab-distributed, [mg_nd, mg_levels) replicated (0 when world == 1)
  int mg_tail0;            // levels [mg_tail0, mg_levels) run inside the one-block tail kernel (mg.cu)
  // cell-sort arrays
  int32_t* cell_count;     // total+2 entries (last bucket: invalid particles)
  int32_t* cell_start;
  int32_t* scan_tmp;
  Particles pt;
  int64_t n_sorted;        // particles [0, n_sorted) are in cell order; [n_sorted, n) arrived by migration this step
  bool sorted;             // particle arrays are in cell order and cell_start is valid
  // experiment switches, all read from the environment ONCE, in fb_create (a live context never re-reads them, so the
  // captured iteration graph and the solve prologue / epilogue cannot disagree about a storage format)
  bool opt_sparse;         // FLIPB200_SPARSE != 0: per-step grid passes only sweep the region the fluid can reach (fb_active_region)
  bool opt_st4;            // FLIPB200_ST4 != 0: four nodes per lane on rows >= 128 nodes
  bool opt_rz64;           // FLIPB200_RZ64 = 1
  bool opt_no_graph;       // FLIPB200_NO_GRAPH = 1
  bool opt_warm_debug;     // FLIPB200_WARM_DEBUG
  int opt_cheb_steps;      // FLIPB200_CHEB = steps,lmin (0: the compiled-in polynomial)
  float opt_cheb_lmin;
  char opt_p2g, opt_g2p;   // kernel variants (FLIPB200_P2G / FLIPB200_G2P, read once at create): see particles.cu
  DevScalars* ds;          // device
  DevScalars* hs;          // pinned host mirror
  // Region outside which every node array is in its EMPTY state (wf = u = ub = x = 0, flags = static flags): the bounding box
  // `wet` (global node indices, min x,y,z / max x,y,z) of the nodes with weight > 0 found by the last fb_classify.  Valid
  // (sparse_ok) until particles or grids are replaced from outside (set_particles, seed_box, set_grid, step_host).
  bool sparse_ok;
  int wet[6];
  Box box;                 // sweep box of this step (local planes, clipped to the owned range)
  Box gbox;                // the same box over the whole domain (z in GLOBAL planes); == box when world == 1
  int n_fluid;
  int last_iters;          // iterations of the previous solve (sizes the first launch batch)
  bool mg_dirty;           // cell classes changed since the hierarchy was built (set by classify)
  // cached CUDA graph of one PCG iteration; valid while box / dt / mode are unchanged
  cudaGraph_t it_graph; cudaGraphExec_t it_exec; int64_t it_kernels; int it_nprod;
  Box it_box, it_gbox; double it_dt; int it_mode;
  int stage;               // bookkeeping for stage-level calls
  cudaEvent_t ev[8];
  int64_t launches;
  char errmsg[256];
};

#define CK(call)                                                                          \
  do {                                                                                    \
    cudaError_t e__ = (call);                                                             \
    if (e__ != cudaSuccess) {                                                             \
      snprintf(c->errmsg, sizeof(c->errmsg), "%s:%d %s", __FILE__, __LINE__,              \
               cudaGetErrorString(e__));                                                  \
      fprintf(stderr, "flipb200: CUDA error %s\n", c->errmsg);                            \
      return FLIPB200_ERR_CUDA;                                                           \
    }                                                                                     \
  } while (0)

#define CK_LAUNCH() CK(cudaGetLastError())

// Minimum resident blocks per SM asked of the sweep kernels (second argument of __launch_bounds__; 0 = not specified --
// 1 is NOT neutral, ptxas then spends registers freely: k_spmv4f 62 -> 72 us).  Without a hint the register allocator
// settles on 40 / 48 / 64 registers, one or two blocks per SM short of the next occupancy step.  Every value was chosen
// by A/B builds on B200 (tools/gpu_ab_lib.sh; profiles/r2_launch_bounds_ab.txt), per kernel and per
// instantiation -- the effect is not monotonic (k_mg_smooth4h: 12 costs 4 us at equal registers, 16 gains 2-3 us):
//   k_spmv4f       one GPU 9 (64 -> 56 regs, 62.0 -> 60.2 us); PEER none (9 spills there: 64 -> 67.6 us)
//   k_mg_smooth4h  one GPU 16 (40 -> 32 regs, no spill: 41.5 / 40.5 / 46.7 -> 39.5 / 39.0 / 43.3 us); PEER 12 (48 -> 40 regs, 53 -> 49.5 us)
//   k_mg_resid4h   11 (one GPU 45.0 -> 43.8 us, PEER 61 -> 53 us; 14+ spills)
//   k_mg_prolong4  5 (41.0 -> 38.8 us, PEER 47 -> 44 us)      k_update_p4f  8 (40 -> 32 regs: 90.5 -> 78.5 us)
//   k_mg_restrict4h 5: slower (32.4 -> 36.9 us, spills), left alone
#ifndef FB_MINB_SMOOTH4H_DOT
#define FB_MINB_SMOOTH4H_DOT 16
#endif
#ifndef FB_MINB_RESTRICT4H
#define FB_MINB_RESTRICT4H 0
#endif
#ifndef FB_MINB_UPDATE_P4F
#define FB_MINB_UPDATE_P4F 8      /* 40 -> 32 registers (16 bytes spilled), full occupancy: 90.5 -> 78.5 us */
#endif
#ifndef FB_MINB_RHS_DIV
#define FB_MINB_RHS_DIV 6        /* 48 -> 40 registers: grid phase of a 2-pass step 1.718 -> 1.684 ms */
#endif
#ifndef FB_MINB_CLASSIFY
#define FB_MINB_CLASSIFY 6       /* 48 -> 40 registers (68 bytes spilled): 1.718 -> 1.667 ms */
#endif
#ifndef FB_MINB_UPDATE_XR4
#define FB_MINB_UPDATE_XR4 0
#endif
#define FB_MINB_SPMV4F(PEER) ((PEER) ? 0 : 9)
#ifndef FB_MINB_SMOOTH4H_PRE
#define FB_MINB_SMOOTH4H_PRE 16
#endif
#ifndef FB_MINB_SMOOTH4H_PEER
#define FB_MINB_SMOOTH4H_PEER 16
#endif
#define FB_MINB_SMOOTH4H(DOT, PEER) ((PEER) ? FB_MINB_SMOOTH4H_PEER : ((DOT) ? FB_MINB_SMOOTH4H_DOT : FB_MINB_SMOOTH4H_PRE))
#define FB_MINB_RESID4H 11
#define FB_MINB_PROLONG4 5

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
// spline() of fluid.cc:22-37.  Value-identical to the reference away from |x| -> 1; within 1e-4
// of the support edge the polynomial is pure cancellation noise (~1e-16) whose SIGN decides
// `w > 0`, so there the reference's exact operation sequence is reproduced (no FMA, true division).
__device__ __forceinline__ double spline_ref(double x) {
  x = fabs(x);
  if (x < 0.5) return 1.5 * (4.0 * x * x * x - 4.0 * x * x + 2.0 / 3.0);
  if (x < 1.0) {
    if (x > 1.0 - 1e-4) {
      double x3 = __dmul_rn(__dmul_rn(x, x), x);
      double t1 = __ddiv_rn(__dmul_rn(-8.0, x3), 6.0);
      double t2 = __dmul_rn(__dmul_rn(4.0, x), x);
      double t3 = __dmul_rn(4.0, x);
      double s = __dadd_rn(__dsub_rn(__dadd_rn(t1, t2), t3), 4.0 / 3.0);
      return __dmul_rn(1.5, s);
    }
    return 1.5 * ((-4.0 / 3.0) * (x * x * x) + 4.0 * x * x - 4.0 * x + 4.0 / 3.0);
  }
  return 0.0;
}

// The two non-zero 1-D weights of a particle along one axis: w0 = spline(x - n0), w1 = spline(x - (n0 + 1)) with n0 the
// base node (x - n0 in [0, 1)).  |x - n0| + |x - n0 - 1| = 1, so one argument takes the inner polynomial of
// fluid.cc:22-37 and the other the outer one: evaluating spline_ref twice makes a warp run BOTH branches for both
// calls (lanes disagree), 28 fp64 operations and 6 branches per axis; here each polynomial is evaluated once on
// the argument it belongs to, and selects put the results back in node order.  Same expressions as spline_ref;
// the rare cases (support edge, ties at 0.5, arguments outside [0, 1]) fall back to spline_ref itself.
__device__ __forceinline__ void spline_pair(double x, double n0, double& w0, double& w1) {
  const double d0 = x - n0, d1 = x - (n0 + 1.0);
  const double a0 = fabs(d0), a1 = fabs(d1);
  const double lo = fmin(a0, a1), hi = fmax(a0, a1);
  if (lo < 0.5 && hi >= 0.5 && hi <= 1.0 - 1e-4) {
    const double wl = 1.5 * (4.0 * lo * lo * lo - 4.0 * lo * lo + 2.0 / 3.0);
    const double wh = 1.5 * ((-4.0 / 3.0) * (hi * hi * hi) + 4.0 * hi * hi - 4.0 * hi + 4.0 / 3.0);
    const bool first_near = a0 <= a1;
    w0 = first_near ? wl : wh;
    w1 = first_near ? wh : wl;
  } else {
    w0 = spline_ref(d0);
    w1 = spline_ref(d1);
  }
}

// Base cell (lower corner of the 2x2x2 non-zero support) of a coordinate along one axis.
// The reference visits round(p)-1..round(p)+1 (fluid.cc:267-276); spline(|d| >= 1) == 0 leaves
// round(p) and the neighbour on the side of the fractional part.  Returns index-space base.
__device__ __forceinline__ int base_cell(double p, int o) {
  double f = round(p);            // C99 half-away-from-zero, in reference coordinates
  int fi = (int)f + o;
  return (p - f) < 0.0 ? fi - 1 : fi;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide sum of up to NV doubles per thread; result valid in thread 0.  blockDim <= 1024.
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* smem /* >= 32*NV */) {
  const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  const int nthreads = blockDim.x * blockDim.y * blockDim.z;
  const int lane = tid & 31, wid = tid >> 5, nw = (nthreads + 31) >> 5;
#pragma unroll
  for (int q = 0; q < NV; ++q) v[q] = warp_sum(v[q]);
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int q = 0; q < NV; ++q) smem[q * 32 + wid] = v[q];
  }
  __syncthreads();
  if (wid == 0) {
#pragma unroll
    for (int q = 0; q < NV; ++q) {
      double t = lane < nw ? smem[q * 32 + lane] : 0.0;
      v[q] = warp_sum(t);
    }
  }
}

// internal API between translation units
int fb_sort_particles(flipb200_ctx* c);
int fb_p2g(flipb200_ctx* c);
int fb_classify(flipb200_ctx* c);
// The region (local node indices, half-open, z clipped to the owned planes; z1 == z0: nothing of it on this rank) that
// this step's per-node passes -- P2G clears and cell sweep, k_classify, k_cell_delta -- have to visit: the whole grid,
// or, while ctx::sparse_ok holds, the last step's wet bounding box grown by SPARSE_MARGIN nodes.  Particles move at
// most one cell per step (dt <= dx / maxSpeed, fluid.cc:992-999), so their 2x2x2 supports stay inside wet +- 2.
#define SPARSE_MARGIN 3
Box fb_active_region(const flipb200_ctx* c, bool* full);
int fb_rhs_div(flipb200_ctx* c, double dt, int mode);
int fb_pcg_solve(flipb200_ctx* c, double dt, int* iters, double* rel);
int fb_vel_update(flipb200_ctx* c, double dt_sub);
int fb_cell_delta(flipb200_ctx* c);
int fb_cell_delta_r(flipb200_ctx* c, double ratio);
int fb_emit_box(flipb200_ctx* c, const int32_t lo[3], const int32_t hi[3], uint64_t seed);
int fb_g2p_advect(flipb200_ctx* c);
int fb_fetch_scalars(flipb200_ctx* c);
int fb_pcg_time(flipb200_ctx* c, int which, int iters, float* ms, int64_t* units);
void fb_graph_invalidate(flipb200_ctx* c);
int fb_mg_alloc(flipb200_ctx* c);
void fb_mg_free(flipb200_ctx* c);
int fb_mg_setup(flipb200_ctx* c, double dt);                    // per step: coarse flags + boxes
int fb_mg_vcycle(flipb200_ctx* c, int it);                      // z (= mg[0].xb, fp32, scaled) = M^-1 r; r.z -> ds->rz[it & 3]
void fb_pcg_publish(flipb200_ctx* c, int mask, int it_arg);     // ds->part[] -> ring slots after an all-reduce
int fb_pcg_reduce_dot(flipb200_ctx* c, int id, int it_arg);     // all-reduce of ds->part[id] into the ring (fused mailboxes or NCCL)
int fb_seed_box(flipb200_ctx* c, const int32_t lo[3], const int32_t hi[3], uint64_t seed);
int fb_reserve_particles(flipb200_ctx* c, int64_t n);
int fb_init_ids(flipb200_ctx* c, int64_t first, int64_t n, int64_t id0);
int fb_sort_particles_ex(flipb200_ctx* c, int filter);
int fb_download_particles(flipb200_ctx* c, double* const dst[6]);
int fb_create(const flipb200_params* p, const uint8_t* solid_host, int rank, int world, flipb200_ctx** out, const int32_t* zsplit_in);

// ---- z-slab exchanges (comm.cu); every call is a no-op returning 0 when world == 1 -------------
enum { FB_F64 = 0, FB_I32 = 1, FB_U64 = 2 };
enum { FB_SUM = 0, FB_MIN = 1, FB_MAX = 2 };
// fill my ghost planes from the neighbours' owned planes: nlo planes below k0, nhi planes above k1
int fb_halo_fill(flipb200_ctx* c, const Grid& g, void* base, size_t elem, int nlo, int nhi);
// P2G ghost sum: my upper ghost plane k1 of each array is added into the upper neighbour's plane k0
int fb_halo_sum_up(flipb200_ctx* c, double* const arr[4]);
int fb_allreduce(flipb200_ctx* c, void* dev, int count, int type, int op);
// every rank ends with all owned planes of a REPLICATED (global) grid; zs[r] = first plane of rank r
int fb_allgather_planes(flipb200_ctx* c, const Grid& gglobal, void* base, size_t elem, const int* zs);
int fb_allgather_planes2(flipb200_ctx* c, const Grid& g, void* base0, void* base1, size_t elem, const int* zs);
// exchange the particles of the trailing down / up buckets with the neighbours (particles.cu)
int fb_comm_exchange_counts(flipb200_ctx* c, const int64_t send[2], int64_t recv[2]);
int fb_comm_exchange_particles(flipb200_ctx* c, int narr, void* const src[], void* const dst[], const size_t elem[],
                               const int64_t send_off[2], const int64_t send_n[2],
                               const int64_t recv_off[2], const int64_t recv_n[2]);
void fb_comm_free(flipb200_ctx* c);
int fb_comm_barrier(flipb200_ctx* c);
bool fb_comm_capturable(flipb200_ctx* c);
int fb_halo_setup(flipb200_ctx* c);
int fb_halo_point(flipb200_ctx* c);          // a place where a halo copy used to be (host barrier of the test transport)

static inline void fb_prof(flipb200_ctx* c, const char* name) {
  IterProf* p = c->prof;
  if (!p || !p->on || p->n >= PROF_MAX) return;
  if (!p->ev[p->n]) cudaEventCreate(&p->ev[p->n]);
  p->name[p->n] = name;
  cudaEventRecord(p->ev[p->n], c->stream);
  p->n++;
}
static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
