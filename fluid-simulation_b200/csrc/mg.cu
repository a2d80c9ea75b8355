// mg.cu -- geometric multigrid V-cycle used as the (IC-free) preconditioner of the pressure CG.
//
// The reference preconditions with Eigen's IncompleteCholesky (fluid.cc:1352,1473); the north_star
// asks for a "Jacobi/IC-free" preconditioner.  Jacobi-PCG needs ~10 n iterations on an n^3 tank
// (2600 at 256^3); this V-cycle brings that to ~25-30, independent of n (tools/mg_proto*.py).
//
// Design (after McAdams, Sifakis, Teran 2010, adapted to the reference's cell classes):
//  * level l+1 cell (I,J,K) covers level-l cells (2I..2I+1, 2J..2J+1, 2K..2K+1); children outside
//    the grid count as solid.  Coarse type: AIR if any child is air (Dirichlet wins), else FLUID if
//    any child is fluid, else SOLID.
//  * operator on every level: rediscretised 7-point, A x = h2inv*(k x_c - sum fluid nbrs), k = number
//    of non-solid neighbours (air = Dirichlet 0, solid = Neumann); h2inv_l = scale / 4^l.
//  * smoother: damped Jacobi, 2 pre sweeps with weights (0.6, 1.7) + 2 post sweeps in the reversed order (symmetric).
//  * prolongation P: cell-centred trilinear (3/4,1/4 per axis) with the weights of SOLID coarse
//    cells dropped and the rest renormalised (constant extrapolation at Neumann walls) -- the
//    normaliser of every fine cell is an integer n64/64 kept in a byte array per level;
//    restriction R = P^T / 8.  => M^-1 is symmetric positive definite, as CG requires.
//  * levels of <= 4096 cells run inside ONE thread block (k_mg_tail); the coarsest (<= 512 cells) is
//    solved by 40 Chebyshev steps in shared memory.
//  * precision: the whole cycle runs in fp32 on a residual scaled to O(1) (the outer CG stays fp64
//    and converges to the same 2.2e-16; prototype: 23 vs 22 iterations, tools/mg_proto3.py), which
//    halves the preconditioner's HBM traffic.  r.z is accumulated in fp64 against the fp64 residual.
// Kernel shapes: sweep.cuh.  Finest-level bytes per node and V-cycle: smooth 13, residual 14,
// restrict 4.5, prolong 9.5, smooth 13, smooth+dot 21  = 75 B (fp32) vs 91 B for ONE Jacobi-PCG
// iteration in fp64.  Levels whose rows are >= 128 nodes wide use the four-nodes-per-lane variants
// (k_mg_smooth4 / k_mg_resid4 / k_mg_restrict4 / k_mg_prolong4): the two-node kernels are
// issue-bound there (73-80 % of the issue slots at <= 50 % DRAM, profiles/r1c_summary.txt).
#include "sweep.cuh"
#include "pcg.cuh"
#include "halo.cuh"

#define MG_TAIL_CELLS 4096       // levels up to this size run inside the one-block tail kernel
#define MG_COARSEST_CELLS 512    // coarsening stops here (Chebyshev solve)
#define MG_TAIL_MAX 4

#define MG_REPLICATED_CELLS (1 << 19)   // z-slab runs: levels with at most this many cells are solved on every rank

// dimensions of the whole level below `f` (f: whole-domain dimensions of the finer level)
static void mg_level_dims(const Grid& f, Grid* c) {
  c->nx = (f.nx + 1) / 2; c->ny = (f.ny + 1) / 2; c->nz = (f.gnz + 1) / 2;
  c->px = (c->nx + 31) & ~31;
  c->ox = c->oy = c->oz = 0; c->wall = 0;
  c->plane = c->px * c->ny;
  c->total = c->plane * c->nz;
  c->kg0 = 0; c->gnz = c->nz; c->k0 = 0; c->k1 = c->nz;
}

static int mg_alloc_level(flipb200_ctx* c, MgLevel& L, bool own_flags) {
  size_t T = (size_t)L.g.total;
  if (own_flags) {
    CK(cudaMalloc((void**)&L.flags, T));
    CK(cudaMemsetAsync(L.flags, FB_SOLID | FB_SMASK, T, c->stream));
  }
  CK(cudaMalloc((void**)&L.nrm, T));
  CK(cudaMalloc((void**)&L.xa, T * sizeof(float)));
  CK(cudaMalloc((void**)&L.xb, T * sizeof(float)));
  CK(cudaMalloc((void**)&L.b, T * sizeof(float)));
  CK(cudaMemsetAsync(L.nrm, 64, T, c->stream));
  CK(cudaMemsetAsync(L.xa, 0, T * sizeof(float), c->stream));
  CK(cudaMemsetAsync(L.xb, 0, T * sizeof(float), c->stream));
  CK(cudaMemsetAsync(L.b, 0, T * sizeof(float), c->stream));
  return 0;
}

struct MgTail;
struct DevScalars;
typedef void (*TailKernel)(const MgTail, const DevScalars*);
static TailKernel tail_kernel(int n);

int fb_mg_alloc(flipb200_ctx* c) {
  c->mg_levels = 0; c->mg_nd = 0;
  if (c->P.precond != 1) return 0;
  // count levels first (whole-domain dimensions): a grid that is already "coarsest" keeps the Jacobi preconditioner
  Grid f = c->g; f.nz = f.gnz;
  int nl = 1;
  Grid dims[MG_MAX_LEVELS]; dims[0] = f;
  const char* te = getenv("FLIPB200_MG_TAIL");      // =0: stop coarsening at MG_TAIL_CELLS (tail = coarsest level only)
  const int64_t stop_cells = (te && te[0] == '0') ? MG_TAIL_CELLS : MG_COARSEST_CELLS;
  while (nl < MG_MAX_LEVELS) {
    const Grid& p = dims[nl - 1];
    if ((int64_t)p.nx * p.ny * p.nz <= stop_cells || min(p.nx, min(p.ny, p.nz)) < 4) break;
    mg_level_dims(p, &dims[nl]);
    ++nl;
  }
  if (nl < 2) return 0;
  // z-slab runs: how many levels stay distributed.  Slab boundaries are multiples of 2^slab_align
  // and level nd is written plane-aligned too, so nd <= slab_align; the first level small enough
  // to replicate ends the distributed part.
  int nd = 0;
  if (c->world > 1) {
    nd = 1;
    while (nd < c->slab_align && nd < nl - 1 && (int64_t)dims[nd].nx * dims[nd].ny * dims[nd].nz > MG_REPLICATED_CELLS) ++nd;
    const char* e = getenv("FLIPB200_MG_ND");
    if (e && e[0]) { int v = atoi(e); nd = max(1, min(v, min(c->slab_align, nl - 1))); }
    if (c->slab_align < 1) return FLIPB200_ERR_ARG;
  }
  for (int l = 0; l < nl; ++l) {
    MgLevel& L = c->mg[l];
    L.dist = l < nd;
    for (int r = 0; r < c->world; ++r) L.zs[r] = c->zsplit[r] >> l;
    L.zs[c->world] = dims[l].nz;
    if (l == 0) { L.g = c->g; L.flags = c->flags; }
    else if (L.dist) {
      Grid g = dims[l];
      const int own = L.zs[c->rank + 1] - L.zs[c->rank];
      g.gnz = dims[l].nz; g.nz = own + 2 * SLAB_GHOST; g.kg0 = L.zs[c->rank] - SLAB_GHOST;
      g.k0 = SLAB_GHOST; g.k1 = SLAB_GHOST + own; g.total = g.plane * g.nz;
      L.g = g;
    } else L.g = dims[l];
    if (mg_alloc_level(c, L, l > 0)) return FLIPB200_ERR_CUDA;
  }
  for (int n = 1; n <= MG_TAIL_MAX; ++n)
    CK(cudaFuncSetAttribute((const void*)tail_kernel(n), cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
  c->mg_levels = nl; c->mg_nd = nd;
  // first level of the one-block tail: replicated, not level 0, small enough, at most MG_TAIL_MAX levels
  int t0 = nl - 1;
  while (t0 - 1 >= max(1, nd) && nl - (t0 - 1) <= MG_TAIL_MAX &&
         (int64_t)dims[t0 - 1].nx * dims[t0 - 1].ny * dims[t0 - 1].nz <= MG_TAIL_CELLS) --t0;
  c->mg_tail0 = t0;
  return 0;
}

void fb_mg_free(flipb200_ctx* c) {
  for (int l = 0; l < c->mg_levels; ++l) {
    MgLevel& L = c->mg[l];
    if (l > 0) cudaFree(L.flags);
    cudaFree(L.nrm); cudaFree(L.xa); cudaFree(L.xb); cudaFree(L.b);
  }
  c->mg_levels = 0;
}

// ---- per-step hierarchy: coarse cell types, solid-neighbour masks, prolongation normalisers ----
// Sweeps the coarse planes [K0, K0 + gridDim.z) (local indices of gc).  Slab runs: the children of an
// owned coarse plane are owned fine planes (slab boundaries are aligned, ctx.cu slab_split).
__global__ void k_mg_coarsen(Grid gf, Grid gc, int K0, const uint8_t* __restrict__ ff, uint8_t* __restrict__ fc) {
  int I = blockIdx.x * blockDim.x + threadIdx.x;
  int J = blockIdx.y * blockDim.y + threadIdx.y;
  int K = K0 + blockIdx.z;
  if (I >= gc.px || J >= gc.ny) return;
  int q = gidx(gc, I, J, K);
  if (I >= gc.nx) { fc[q] = FB_SOLID; return; }
  bool any_air = false, any_fluid = false;
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    int i = 2 * I + (m & 1), j = 2 * J + ((m >> 1) & 1), k = 2 * (K + gc.kg0) + (m >> 2) - gf.kg0;
    if (i < gf.nx && j < gf.ny && k + gf.kg0 < gf.gnz) {
      uint8_t f = ff[gidx(gf, i, j, k)];
      if (f & FB_FLUID) any_fluid = true;
      else if (!(f & FB_SOLID)) any_air = true;
    }
  }
  fc[q] = any_air ? 0 : (any_fluid ? FB_FLUID : FB_SOLID);
}

__global__ void k_mg_nbrmask(Grid g, int K0, uint8_t* __restrict__ fl) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int j = blockIdx.y * blockDim.y + threadIdx.y;
  int k = K0 + blockIdx.z;
  if (i >= g.nx || j >= g.ny) return;
  int q = gidx(g, i, j, k);
  auto sol = [&](int a, int b, int cc) -> bool {
    return !in_box(g, a, b, cc) || (fl[gidx(g, a, b, cc)] & FB_SOLID);   // outside the coarse grid: Neumann
  };
  uint8_t f = fl[q] & (FB_FLUID | FB_SOLID);
  uint8_t m = 0;
  if (sol(i - 1, j, k)) m |= FB_SXM;
  if (sol(i + 1, j, k)) m |= FB_SXP;
  if (sol(i, j - 1, k)) m |= FB_SYM;
  if (sol(i, j + 1, k)) m |= FB_SYP;
  if (sol(i, j, k - 1)) m |= FB_SZM;
  if (sol(i, j, k + 1)) m |= FB_SZP;
  // only bit 7 of the neighbours is read and it never changes here: the in-place update is safe
  if ((f & FB_FLUID) && m == FB_SMASK) f &= ~FB_FLUID;   // no non-solid neighbour: empty row
  fl[q] = f | m;
}

// n64 = 64 * sum of the trilinear weights of the non-solid, in-grid coarse donors of a fine cell
__global__ void k_mg_norm(Grid gf, Box box, Grid gc, const uint8_t* __restrict__ ff, const uint8_t* __restrict__ fc,
                          uint8_t* __restrict__ nrm) {
  int i = box.x0 + blockIdx.x * blockDim.x + threadIdx.x;
  int j = box.y0 + blockIdx.y * blockDim.y + threadIdx.y;
  int k = box.z0 + blockIdx.z;
  if (i >= box.x1 || j >= box.y1) return;
  int q = gidx(gf, i, j, k);
  int n64 = 64;
  if (ff[q] & FB_FLUID) {
    n64 = 0;
    const int kg = k + gf.kg0;                       // global fine plane
    int I = i >> 1, J = j >> 1, K = (kg >> 1) - gc.kg0;
    int sx = (i & 1) ? 1 : -1, sy = (j & 1) ? 1 : -1, sz = (kg & 1) ? 1 : -1;
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      int ci = I + ((m & 1) ? sx : 0), cj = J + ((m & 2) ? sy : 0), ck = K + ((m & 4) ? sz : 0);
      if (!in_box(gc, ci, cj, ck)) continue;
      if (fc[gidx(gc, ci, cj, ck)] & FB_SOLID) continue;
      n64 += ((m & 1) ? 1 : 3) * ((m & 2) ? 1 : 3) * ((m & 4) ? 1 : 3);
    }
    if (n64 == 0) n64 = 64;   // cannot happen: the parent of a fluid cell is never solid
  }
  nrm[q] = (uint8_t)n64;
}

// local sweep boxes of a level from its whole-domain box
static void mg_local_boxes(const flipb200_ctx* c, MgLevel& L) {
  const Box& gb = L.gbox;
  L.box = gb; L.boxe = gb;
  if (!L.dist || !(gb.x1 > gb.x0)) return;
  const int z0 = L.zs[c->rank], z1 = L.zs[c->rank + 1], kg0 = L.g.kg0;
  L.box.z0 = max(gb.z0, z0) - kg0; L.box.z1 = min(gb.z1, z1) - kg0;
  L.boxe.z0 = max(gb.z0, z0 - 1) - kg0; L.boxe.z1 = min(gb.z1, z1 + 1) - kg0;
  if (L.box.z1 <= L.box.z0) {          // the box does not reach this slab: one inert plane (uniform launches)
    L.box.z0 = L.g.k0; L.box.z1 = L.g.k0 + 1;
    L.boxe = L.box;
  }
}

int fb_mg_setup(flipb200_ctx* c, double dt) {
  if (c->mg_levels == 0) return 0;
  const double scale = dt / (c->P.rho * c->P.dx * c->P.dx);   // fluid.cc:306
  c->mg[0].gbox = c->gbox;
  c->mg[0].box = c->box;
  mg_local_boxes(c, c->mg[0]);
  c->mg[0].h2inv = scale;
  const dim3 blk(32, 8, 1);
  int rc;
  for (int l = 1; l < c->mg_levels; ++l) {
    MgLevel& F = c->mg[l - 1];
    MgLevel& L = c->mg[l];
    L.h2inv = F.h2inv * 0.25;
    if (!c->mg_dirty) continue;                     // same cell classes as the previous solve of this step
    // cell types: this rank coarsens the planes whose children it owns ...
    const int K0 = (F.dist ? L.zs[c->rank] : 0) - L.g.kg0;
    const int nK = F.dist ? L.zs[c->rank + 1] - L.zs[c->rank] : L.g.gnz;
    k_mg_coarsen<<<dim3(cdiv(L.g.px, 32), cdiv(L.g.ny, 8), nK), blk, 0, c->stream>>>(F.g, L.g, K0, F.flags, L.flags);
    c->launches++;
    // ... and gets the rest: ghost planes from the z neighbours, or every plane of a replicated level
    if (L.dist) { if ((rc = fb_halo_fill(c, L.g, L.flags, 1, 1, 1))) return rc; }
    else if (F.dist) { if ((rc = fb_allgather_planes(c, L.g, L.flags, 1, L.zs))) return rc; }
    const int M0 = L.dist ? L.g.k0 : 0, nM = L.dist ? L.g.k1 - L.g.k0 : L.g.nz;
    k_mg_nbrmask<<<dim3(cdiv(L.g.px, 32), cdiv(L.g.ny, 8), nM), blk, 0, c->stream>>>(L.g, M0, L.flags);
    c->launches++;
    // coarse box: the parents of the fine box plus a 1-cell margin (prolongation donors)
    Box b;
    const Box& fg = F.gbox;
    b.x0 = max(0, fg.x0 / 2 - 1) & ~1; b.y0 = max(0, fg.y0 / 2 - 1); b.z0 = max(0, fg.z0 / 2 - 1);
    b.x1 = min(L.g.nx, (fg.x1 + 1) / 2 + 1); b.y1 = min(L.g.ny, (fg.y1 + 1) / 2 + 1);
    b.z1 = min(L.g.gnz, (fg.z1 + 1) / 2 + 1);
    if (!(fg.x1 > fg.x0)) b = Box{0, 0, 0, 0, 0, 0};
    L.gbox = b;
    mg_local_boxes(c, L);
    const Box& fb = F.box;
    if (fb.x1 > fb.x0 && fb.y1 > fb.y0 && fb.z1 > fb.z0) {
      dim3 g2(cdiv(fb.x1 - fb.x0, 32), cdiv(fb.y1 - fb.y0, 8), fb.z1 - fb.z0);
      k_mg_norm<<<g2, blk, 0, c->stream>>>(F.g, fb, L.g, F.flags, L.flags, F.nrm);
      c->launches++;
    }
  }
  c->mg_dirty = false;
  CK_LAUNCH();
  return 0;
}

// ---- V-cycle kernels (fp32) --------------------------------------------------------------------
// damped-Jacobi sweep  xo = xi + w D^-1 (b - A xi)   (STENCIL shape).
// DOT (last sweep of level 0): also the CG's r.z, accumulated in fp64:
//   DOT 2 (default): rz[it&3] += sum b * xo / sc^2 -- r taken as b / sc, the fp32 right-hand side this
//          sweep reads anyway (z = M b exactly, so b.z is the inner product M defines; the 6e-8
//          rounding of r only perturbs alpha and beta, x and r stay consistent: same iteration counts);
//   DOT 1 (FLIPB200_RZ64=1): rz[it&3] += sum r64 * xo / sc, 8 more bytes per node.
template <int DOT>
__global__ void __launch_bounds__(32 * ST_BY)
k_mg_smooth(const __grid_constant__ Sweep S, const __grid_constant__ MgCoef mc, const uint8_t* __restrict__ fl,
            const float* __restrict__ b, const float* __restrict__ xi, float* __restrict__ xo,
            const double* __restrict__ r64, DevScalars* __restrict__ ds, int it_arg, int dist,
            const __grid_constant__ PeerHalo H) {
  __shared__ double sm[32];
  if (ds->pcg_done) return;
  const int it = pcg_iter(ds, it_arg);
  const int kz0 = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * S.zc, kz1 = min(kz0 + S.zc, S.box.z1);
  const bool bdn = kz0 <= H.kdn && H.kdn < kz1, bup = kz0 <= H.kup && H.kup < kz1;
  halo_wait(H, ds, bdn, bup);
  double acc[1] = {0.0};
  StencilMarch<float> m;
  if (m.begin(S, xi)) {
    const int kend = min(m.k + S.zc, S.box.z1);
#pragma unroll
    for (int z = 0; z < ST_ZP; ++z) {
      if (m.k >= kend) break;
      m.load(S, xi);
      uchar2 f = make_uchar2(0, 0);
      float2 bv = make_float2(0.f, 0.f);
      double2 rv = make_double2(0.0, 0.0);
      if (m.act) {
        f = *reinterpret_cast<const uchar2*>(fl + m.q);
        bv = ld2<float>(b + m.q);
        if (DOT == 1) rv = ld2<double>(r64 + m.q);
      }
      float o0 = 0.f, o1 = 0.f;
      if (f.x & FB_FLUID) { int kk = 6 - __popc(f.x & FB_SMASK); o0 = m.c0() + mc.winv[kk] * (bv.x - mc.h2inv * ((float)kk * m.c0() - m.sum0())); }
      if (f.y & FB_FLUID) { int kk = 6 - __popc(f.y & FB_SMASK); o1 = m.c1() + mc.winv[kk] * (bv.y - mc.h2inv * ((float)kk * m.c1() - m.sum1())); }
      if (m.act) { st2<float>(xo + m.q, o0, o1); halo_push2<float>(H, m.q, m.k, o0, o1); }
      if (DOT == 1) acc[0] += rv.x * (double)o0 + rv.y * (double)o1;
      if (DOT == 2) acc[0] += (double)bv.x * (double)o0 + (double)bv.y * (double)o1;
      m.advance(S);
    }
  }
  halo_signal(H, ds, bdn, bup);
  if (DOT) {
    block_sum<1>(acc, sm);
    if (threadIdx.x == 0 && threadIdx.y == 0 && acc[0] != 0.0) {
      const double rrp = ds->rr[(it + 3) & 3];
      atomicAdd(dist ? &ds->part[PART_RZ] : &ds->rz[it & 3], acc[0] * (rrp > 0.0 ? (DOT == 2 ? rrp : sqrt(rrp)) : 1.0));
    }
  }
}

// normalised residual  rn = (b - A x) * 64 / n64   (input of the restriction)
__global__ void __launch_bounds__(32 * ST_BY)
k_mg_resid(const __grid_constant__ Sweep S, const __grid_constant__ MgCoef mc, const uint8_t* __restrict__ fl,
           const uint8_t* __restrict__ nrm, const float* __restrict__ b, const float* __restrict__ x,
           float* __restrict__ rn, DevScalars* __restrict__ ds, const __grid_constant__ PeerHalo H) {
  if (ds->pcg_done) return;
  const int kz0 = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * S.zc, kz1 = min(kz0 + S.zc, S.box.z1);
  const bool bdn = kz0 <= H.kdn && H.kdn < kz1, bup = kz0 <= H.kup && H.kup < kz1;
  halo_wait(H, ds, bdn, bup);
  StencilMarch<float> m;
  if (m.begin(S, x)) {
    const int kend = min(m.k + S.zc, S.box.z1);
#pragma unroll
    for (int z = 0; z < ST_ZP; ++z) {
      if (m.k >= kend) break;
      m.load(S, x);
      uchar2 f = make_uchar2(0, 0), nn = make_uchar2(64, 64);
      float2 bv = make_float2(0.f, 0.f);
      if (m.act) {
        f = *reinterpret_cast<const uchar2*>(fl + m.q);
        nn = *reinterpret_cast<const uchar2*>(nrm + m.q);
        bv = ld2<float>(b + m.q);
      }
      float o0 = 0.f, o1 = 0.f;
      if (f.x & FB_FLUID) { int kk = 6 - __popc(f.x & FB_SMASK); o0 = (bv.x - mc.h2inv * ((float)kk * m.c0() - m.sum0())) * (64.f / (float)nn.x); }
      if (f.y & FB_FLUID) { int kk = 6 - __popc(f.y & FB_SMASK); o1 = (bv.y - mc.h2inv * ((float)kk * m.c1() - m.sum1())) * (64.f / (float)nn.y); }
      if (m.act) { st2<float>(rn + m.q, o0, o1); halo_push2<float>(H, m.q, m.k, o0, o1); }
      m.advance(S);
    }
  }
  halo_signal(H, ds, bdn, bup);
}

// ---- STENCIL4 variants of the two sweeps (levels whose rows are >= 128 nodes wide) ----------------
__device__ __forceinline__ float smooth_node(const MgCoef& mc, uint32_t f, float c, float b, float sum) {
  if (!(f & FB_FLUID)) return 0.f;
  const int kk = 6 - __popc(f & FB_SMASK);
  return c + mc.winv[kk] * (b - mc.h2inv * ((float)kk * c - sum));
}
__device__ __forceinline__ float resid_node(const MgCoef& mc, uint32_t f, uint32_t nn, float c, float b, float sum) {
  if (!(f & FB_FLUID)) return 0.f;
  const int kk = 6 - __popc(f & FB_SMASK);
  return (b - mc.h2inv * ((float)kk * c - sum)) * (64.f / (float)nn);
}
// PEER = false (single GPU): the ghost-plane wait / push / signal code is not compiled in (its
// presence alone costs ~5 us per level-0 sweep, tools/kbench.cu T9).
template <int DOT, bool PEER>
__global__ void __launch_bounds__(32 * ST_BY)
k_mg_smooth4(const __grid_constant__ Sweep S, const __grid_constant__ MgCoef mc, const uint8_t* __restrict__ fl,
             const float* __restrict__ b, const float* __restrict__ xi, float* __restrict__ xo,
             const double* __restrict__ r64, DevScalars* __restrict__ ds, int it_arg, int dist,
             const __grid_constant__ PeerHalo H) {
  __shared__ double sm[32];
  if (ds->pcg_done) return;
  const int it = pcg_iter(ds, it_arg);
  const int kz0 = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * S.zc, kz1 = min(kz0 + S.zc, S.box.z1);
  const bool bdn = kz0 <= H.kdn && H.kdn < kz1, bup = kz0 <= H.kup && H.kup < kz1;
  if (PEER) halo_wait(H, ds, bdn, bup);
  double acc[1] = {0.0};
  StencilMarch4 m;
  if (m.begin(S, xi)) {
    const int kend = min(m.k + S.zc, S.box.z1);
#pragma unroll
    for (int z = 0; z < ST_ZP; ++z) {
      if (m.k >= kend) break;
      m.load(S, xi);
      uint32_t f = 0;
      float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
      double2 ra = make_double2(0.0, 0.0), rb = ra;
      if (m.act) {
        f = *reinterpret_cast<const uint32_t*>(fl + m.q);
        bv = *reinterpret_cast<const float4*>(b + m.q);
        if (DOT == 1) { ra = ld2<double>(r64 + m.q); rb = ld2<double>(r64 + m.q + 2); }
      }
      float4 o;
      o.x = smooth_node(mc, f & 0xffu, m.pc.x, bv.x, m.sum0());
      o.y = smooth_node(mc, (f >> 8) & 0xffu, m.pc.y, bv.y, m.sum1());
      o.z = smooth_node(mc, (f >> 16) & 0xffu, m.pc.z, bv.z, m.sum2());
      o.w = smooth_node(mc, f >> 24, m.pc.w, bv.w, m.sum3());
      if (m.act) { *reinterpret_cast<float4*>(xo + m.q) = o; if (PEER) halo_push4(H, m.q, m.k, o); }
      if (DOT == 1) acc[0] += (ra.x * (double)o.x + ra.y * (double)o.y) + (rb.x * (double)o.z + rb.y * (double)o.w);
      if (DOT == 2) acc[0] += ((double)bv.x * (double)o.x + (double)bv.y * (double)o.y) + ((double)bv.z * (double)o.z + (double)bv.w * (double)o.w);
      m.advance(S);
    }
  }
  if (PEER) halo_signal(H, ds, bdn, bup);
  if (DOT) {
    block_sum<1>(acc, sm);
    if (threadIdx.x == 0 && threadIdx.y == 0 && acc[0] != 0.0) {
      const double rrp = ds->rr[(it + 3) & 3];
      atomicAdd(dist ? &ds->part[PART_RZ] : &ds->rz[it & 3], acc[0] * (rrp > 0.0 ? (DOT == 2 ? rrp : sqrt(rrp)) : 1.0));
    }
  }
}
template <bool PEER>
__global__ void __launch_bounds__(32 * ST_BY)
k_mg_resid4(const __grid_constant__ Sweep S, const __grid_constant__ MgCoef mc, const uint8_t* __restrict__ fl,
            const uint8_t* __restrict__ nrm, const float* __restrict__ b, const float* __restrict__ x,
            float* __restrict__ rn, DevScalars* __restrict__ ds, const __grid_constant__ PeerHalo H) {
  if (ds->pcg_done) return;
  const int kz0 = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * S.zc, kz1 = min(kz0 + S.zc, S.box.z1);
  const bool bdn = kz0 <= H.kdn && H.kdn < kz1, bup = kz0 <= H.kup && H.kup < kz1;
  if (PEER) halo_wait(H, ds, bdn, bup);
  StencilMarch4 m;
  if (m.begin(S, x)) {
    const int kend = min(m.k + S.zc, S.box.z1);
#pragma unroll
    for (int z = 0; z < ST_ZP; ++z) {
      if (m.k >= kend) break;
      m.load(S, x);
      uint32_t f = 0, nn = 0x40404040u;
      float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m.act) {
        f = *reinterpret_cast<const uint32_t*>(fl + m.q);
        nn = *reinterpret_cast<const uint32_t*>(nrm + m.q);
        bv = *reinterpret_cast<const float4*>(b + m.q);
      }
      float4 o;
      o.x = resid_node(mc, f & 0xffu, nn & 0xffu, m.pc.x, bv.x, m.sum0());
      o.y = resid_node(mc, (f >> 8) & 0xffu, (nn >> 8) & 0xffu, m.pc.y, bv.y, m.sum1());
      o.z = resid_node(mc, (f >> 16) & 0xffu, (nn >> 16) & 0xffu, m.pc.z, bv.z, m.sum2());
      o.w = resid_node(mc, f >> 24, nn >> 24, m.pc.w, bv.w, m.sum3());
      if (m.act) { *reinterpret_cast<float4*>(rn + m.q) = o; if (PEER) halo_push4(H, m.q, m.k, o); }
      m.advance(S);
    }
  }
  if (PEER) halo_signal(H, ds, bdn, bup);
}

// restriction b_c[I] = (1/8) sum_i P[i,I] res_i = (1/512) sum over the 4x4x4 fine cells of
// {1,3,3,1}^3 weights x rn_i, fused with the coarse level's first Jacobi sweep x1 = w D^-1 b_c.
// One lane per coarse cell: the lane loads its two children of each of the 16 fine rows as one
// float2 (coalesced across the warp) and gets the two outer taps from its neighbours by shuffle.
// The block marches RS_ZC coarse planes: the y/x-reduced sum of a fine plane is shared by the two
// coarse planes it feeds, so a coarse cell costs ~9 fine-row loads instead of 16 (RS_ZC = 8; small
// levels are latency-bound and keep one plane per block, mg_zchunk).
__global__ void __launch_bounds__(256)
k_mg_restrict(Grid gf, Box fbox, Grid gc, Box cbox, MgCoef mcc, const uint8_t* __restrict__ fc,
              const float* __restrict__ rn, float* __restrict__ bc, float* __restrict__ x1c,
              DevScalars* __restrict__ ds, int RS_ZC, const __grid_constant__ PeerHalo H,
              const __grid_constant__ GatherPush G) {
  if (ds->pcg_done) return;
  const int lane = threadIdx.x;
  const int I = cbox.x0 + blockIdx.x * 32 + lane;
  const int J = cbox.y0 + blockIdx.y * blockDim.y + threadIdx.y;
  const int K0 = cbox.z0 + zperm(blockIdx.z, gridDim.z, H.mode ? ZP_EDGE : ZP_UP) * RS_ZC, K1 = min(K0 + RS_ZC, cbox.z1);
  const bool bdn = K0 <= H.kdn && H.kdn < K1, bup = K0 <= H.kup && H.kup < K1;
  halo_wait(H, ds, bdn, bup);                       // boundary coarse planes read the fine ghost planes
  if (J < cbox.y1) {                                // whole warp
    const bool in = I < cbox.x1;
    // the fine sweeps write pairs over [fx0, fxe): everything in that x-range is fresh (0 off the fluid)
    const int fx0 = fbox.x0 & ~1, fxe = min(gf.px, (fbox.x1 + 1) & ~1);
    const int xf = 2 * I;                             // children xf, xf+1
    const bool cv = in && xf >= fx0 && xf < fxe;
    const bool eL = lane == 0 && in && xf - 1 >= fx0 && xf - 1 < fxe, eR = lane == 31 && in && xf + 2 >= fx0 && xf + 2 < fxe;
    const int xc = cv ? xf : fx0;                     // clamped (value discarded when !cv)
    // rows 2J-1 .. 2J+2 with weights 1 3 3 1 (0 outside the fine box)
    int jr[4]; float wy[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int jf = 2 * J + t - 1;
      wy[t] = (jf >= fbox.y0 && jf < fbox.y1) ? ((t == 1 || t == 2) ? 3.f : 1.f) : 0.f;
      jr[t] = min(max(jf, fbox.y0), fbox.y1 - 1);
    }
    // y/x-reduced sum of fine plane kf for this coarse column: sum_rows wy * (a + 3 vx + 3 vy + d)
    auto plane_sum = [&](int kf) -> float {
      const bool pin = kf >= fbox.z0 && kf < fbox.z1;               // block-uniform
      const int kc = min(max(kf, fbox.z0), fbox.z1 - 1);
      float2 v[4]; float ea[4], ed[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float* row = rn + gidx(gf, 0, jr[t], kc);
        v[t] = *reinterpret_cast<const float2*>(row + xc);
        ea[t] = eL ? row[xf - 1] : 0.f;
        ed[t] = eR ? row[xf + 2] : 0.f;
      }
      float sum = 0.f;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float vx = cv ? v[t].x : 0.f, vy = cv ? v[t].y : 0.f;
        float a = __shfl_up_sync(0xffffffffu, vy, 1), d = __shfl_down_sync(0xffffffffu, vx, 1);
        if (lane == 0) a = ea[t];
        if (lane == 31) d = ed[t];
        sum += wy[t] * ((a + d) + 3.f * (vx + vy));
      }
      return pin ? sum : 0.f;
    };
    const int kb = 2 * (K0 + gc.kg0) - gf.kg0;       // local fine plane of the first child of K0
    float sm1 = plane_sum(kb - 1), s0 = plane_sum(kb);
    for (int K = K0; K < K1; ++K) {
      const int kf = kb + 2 * (K - K0);
      const float s1 = plane_sum(kf + 1), s2 = plane_sum(kf + 2);
      const float sum = (sm1 + s2) + 3.f * (s0 + s1);
      if (in) {
        const int qc = gidx(gc, I, J, K);
        const uint8_t f = fc[qc];
        const float r = (f & FB_FLUID) ? sum * (1.f / 512.f) : 0.f;
        const float x1 = (f & FB_FLUID) ? mcc.winv[6 - __popc(f & FB_SMASK)] * r : 0.f;
        bc[qc] = r;
        x1c[qc] = x1;
        halo_push1<float>(H, qc, K, x1);
        gather_push1(G, qc, r, x1);
      }
      sm1 = s1; s0 = s2;
    }
  }
  halo_signal(H, ds, bdn, bup);
  gather_signal(G, ds);
}

// Two coarse cells (four fine children per row, one float4) per lane; used when the coarse rows are
// at least 64 cells wide.  Same association of the sums as k_mg_restrict.
__global__ void __launch_bounds__(256)
k_mg_restrict4(Grid gf, Box fbox, Grid gc, Box cbox, MgCoef mcc, const uint8_t* __restrict__ fc,
               const float* __restrict__ rn, float* __restrict__ bc, float* __restrict__ x1c,
               DevScalars* __restrict__ ds, int RS_ZC, const __grid_constant__ PeerHalo H,
               const __grid_constant__ GatherPush G) {
  if (ds->pcg_done) return;
  const int lane = threadIdx.x;
  const int I = (cbox.x0 & ~1) + blockIdx.x * 64 + 2 * lane;         // cells I, I + 1
  const int J = cbox.y0 + blockIdx.y * blockDim.y + threadIdx.y;
  const int K0 = cbox.z0 + zperm(blockIdx.z, gridDim.z, H.mode ? ZP_EDGE : ZP_UP) * RS_ZC, K1 = min(K0 + RS_ZC, cbox.z1);
  const bool bdn = K0 <= H.kdn && H.kdn < K1, bup = K0 <= H.kup && H.kup < K1;
  halo_wait(H, ds, bdn, bup);
  if (J < cbox.y1) {                                // whole warp
    const bool in = I < cbox.x1;
    const int fx0 = fbox.x0 & ~1, fxe = min(gf.px, (fbox.x1 + 1) & ~1);   // fresh x-range of the fine sweeps (pairs)
    const int xf = 2 * I;                             // children xf .. xf+3
    const bool cv0 = in && xf >= fx0 && xf < fxe, cv1 = in && xf + 2 >= fx0 && xf + 2 < fxe;
    const bool eL = lane == 0 && in && xf - 1 >= fx0 && xf - 1 < fxe, eR = lane == 31 && in && xf + 4 >= fx0 && xf + 4 < fxe;
    const int xc = in ? xf : 0;                       // in-bounds address (values masked by cv0 / cv1)
    int jr[4]; float wy[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int jf = 2 * J + t - 1;
      wy[t] = (jf >= fbox.y0 && jf < fbox.y1) ? ((t == 1 || t == 2) ? 3.f : 1.f) : 0.f;
      jr[t] = min(max(jf, fbox.y0), fbox.y1 - 1);
    }
    auto plane_sum = [&](int kf) -> float2 {
      const bool pin = kf >= fbox.z0 && kf < fbox.z1;               // block-uniform
      const int kc = min(max(kf, fbox.z0), fbox.z1 - 1);
      float4 v[4]; float ea[4], ed[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float* row = rn + gidx(gf, 0, jr[t], kc);
        v[t] = *reinterpret_cast<const float4*>(row + xc);
        ea[t] = eL ? row[xf - 1] : 0.f;
        ed[t] = eR ? row[xf + 4] : 0.f;
      }
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float vx = cv0 ? v[t].x : 0.f, vy = cv0 ? v[t].y : 0.f, vz = cv1 ? v[t].z : 0.f, vw = cv1 ? v[t].w : 0.f;
        float a = __shfl_up_sync(0xffffffffu, vw, 1), d = __shfl_down_sync(0xffffffffu, vx, 1);
        if (lane == 0) a = ea[t];
        if (lane == 31) d = ed[t];
        s0 += wy[t] * ((a + vz) + 3.f * (vx + vy));
        s1 += wy[t] * ((vy + d) + 3.f * (vz + vw));
      }
      return pin ? make_float2(s0, s1) : make_float2(0.f, 0.f);
    };
    const int kb = 2 * (K0 + gc.kg0) - gf.kg0;       // local fine plane of the first child of K0
    float2 sm1 = plane_sum(kb - 1), s0 = plane_sum(kb);
    for (int K = K0; K < K1; ++K) {
      const int kf = kb + 2 * (K - K0);
      const float2 s1 = plane_sum(kf + 1), s2 = plane_sum(kf + 2);
      if (in) {
        const int qc = gidx(gc, I, J, K);
        const uchar2 f = *reinterpret_cast<const uchar2*>(fc + qc);
        const bool in1 = I + 1 < cbox.x1;
        float r0 = 0.f, r1 = 0.f, y0 = 0.f, y1 = 0.f;
        if (f.x & FB_FLUID) { r0 = ((sm1.x + s2.x) + 3.f * (s0.x + s1.x)) * (1.f / 512.f); y0 = mcc.winv[6 - __popc(f.x & FB_SMASK)] * r0; }
        if (in1 && (f.y & FB_FLUID)) { r1 = ((sm1.y + s2.y) + 3.f * (s0.y + s1.y)) * (1.f / 512.f); y1 = mcc.winv[6 - __popc(f.y & FB_SMASK)] * r1; }
        st2<float>(bc + qc, r0, r1);
        st2<float>(x1c + qc, y0, y1);
        halo_push2<float>(H, qc, K, y0, y1);
        gather_push2(G, qc, r0, r1, y0, y1);
      }
      sm1 = s1; s0 = s2;
    }
  }
  halo_signal(H, ds, bdn, bup);
  gather_signal(G, ds);
}

// a rank with nothing to restrict into the first replicated level still has to announce that
__global__ void k_mg_gather_signal(DevScalars* __restrict__ ds, const __grid_constant__ GatherPush G) {
  if (ds->pcg_done) return;
  gather_signal(G, ds);
}

// prolongation + correction  x_f[i] += (sum_I w64(i,I) e_c[I]) / n64_i   (e_c is 0 off the fluid cells).
// One lane per coarse column I = two fine nodes (2I, 2I+1); the x-neighbours I-1 / I+1 come by
// shuffle.  The block marches PR_ZC fine planes: the y-combined coarse value Y(K) of a coarse plane
// feeds the four fine planes 2K-1 .. 2K+2, so it stays in a three-plane register window and a fine
// row costs ~1.5 coarse-row loads instead of 4.
__global__ void __launch_bounds__(256)
k_mg_prolong(Grid gf, Box fbox, Grid gc, const uint8_t* __restrict__ ff, const uint8_t* __restrict__ nrm,
             const float* __restrict__ ec, float* __restrict__ x, DevScalars* __restrict__ ds,
             int PR_ZC, const __grid_constant__ PeerHalo H) {
  if (ds->pcg_done) return;
  const int lane = threadIdx.x;
  const int i = (fbox.x0 & ~1) + blockIdx.x * 64 + 2 * lane;
  const int j = fbox.y0 + blockIdx.y * blockDim.y + threadIdx.y;
  const int k0c = fbox.z0 + zperm(blockIdx.z, gridDim.z, H.mode ? ZP_EDGE : ZP_UP) * PR_ZC, k1c = min(k0c + PR_ZC, fbox.z1);
  const bool bdn = k0c <= H.kdn && H.kdn < k1c, bup = k0c <= H.kup && H.kup < k1c;
  halo_wait(H, ds, bdn, bup);                       // boundary fine planes read the coarse ghost planes
  if (j < fbox.y1) {                                // whole warp
    const int I = i >> 1, J = j >> 1;
    int J2 = J + ((j & 1) ? 1 : -1);
    const bool yin = J2 >= 0 && J2 < gc.ny;
    J2 = yin ? J2 : J;
    const float wy = yin ? 1.f : 0.f;
    const bool cin = I < gc.nx;
    const int Ie = lane == 0 ? I - 1 : I + 1;         // the column the edge lanes cannot get by shuffle
    const bool ein = (lane == 0 || lane == 31) && Ie >= 0 && Ie < gc.nx;
    // y-combined coarse value (weights 3|1) of global coarse plane Kg for column I (and Ie on the edge lanes)
    auto ycomb = [&](int Kg, float& ye) -> float {
      ye = 0.f;
      if (Kg < 0 || Kg >= gc.gnz) return 0.f;         // block-uniform
      const float* r0 = ec + gidx(gc, 0, J, Kg - gc.kg0);
      const float* r1 = ec + gidx(gc, 0, J2, Kg - gc.kg0);
      if (ein) ye = 3.f * r0[Ie] + wy * r1[Ie];
      return cin ? 3.f * r0[I] + wy * r1[I] : 0.f;
    };
    const int xe = min(gf.px, (fbox.x1 + 1) & ~1);
    const bool act = i < xe;
    int Kc = (k0c + gf.kg0) >> 1;
    float em, e0, ep;
    float Ym = ycomb(Kc - 1, em), Y0 = ycomb(Kc, e0), Yp = ycomb(Kc + 1, ep);
    // the fine-plane operands are fetched one plane ahead (the march is latency-bound otherwise)
    int q = gidx(gf, act ? i : 0, j, k0c);
    uchar2 f = make_uchar2(0, 0), nn = make_uchar2(1, 1), f1 = f, nn1 = nn;
    float2 xv = make_float2(0.f, 0.f), xv1 = xv;
    if (act) {
      f = *reinterpret_cast<const uchar2*>(ff + q);
      nn = *reinterpret_cast<const uchar2*>(nrm + q);
      xv = *reinterpret_cast<const float2*>(x + q);
    }
    for (int k = k0c; k < k1c; ++k, q += gf.plane) {
      if (act && k + 1 < k1c) {
        const int qn = q + gf.plane;
        f1 = *reinterpret_cast<const uchar2*>(ff + qn);
        nn1 = *reinterpret_cast<const uchar2*>(nrm + qn);
        xv1 = *reinterpret_cast<const float2*>(x + qn);
      }
      const int kg = k + gf.kg0;                       // global fine plane; its donors are kg>>1 (3) and kg>>1 -+ 1 (1)
      if ((kg >> 1) != Kc) {
        Kc = kg >> 1;
        Ym = Y0; em = e0; Y0 = Yp; e0 = ep;
        Yp = ycomb(Kc + 1, ep);
      }
      const bool up = kg & 1;
      const float E = 3.f * Y0 + (up ? Yp : Ym);
      const float Ee = 3.f * e0 + (up ? ep : em);
      float EL = __shfl_up_sync(0xffffffffu, E, 1), ER = __shfl_down_sync(0xffffffffu, E, 1);
      if (lane == 0) EL = Ee;
      if (lane == 31) ER = Ee;
      if ((f.x | f.y) & FB_FLUID) {                    // never true on inactive lanes (f = 0)
        if (f.x & FB_FLUID) xv.x += __fdividef(3.f * E + EL, (float)nn.x);     // node 2I   : donors I (3) and I-1 (1)
        if (f.y & FB_FLUID) xv.y += __fdividef(3.f * E + ER, (float)nn.y);     // node 2I+1 : donors I (3) and I+1 (1)
        *reinterpret_cast<float2*>(x + q) = xv;
        halo_push2<float>(H, q, k, xv.x, xv.y);
      }
      f = f1; nn = nn1; xv = xv1;
    }
  }
  halo_signal(H, ds, bdn, bup);
}

// Four fine nodes (two coarse columns) per lane: the issue-bound 2-node shape spends most of its
// instructions on per-lane overhead.  Used on the levels whose rows fill a 128-node warp.
__global__ void __launch_bounds__(256, FB_MINB_PROLONG4)
k_mg_prolong4(Grid gf, Box fbox, Grid gc, const uint8_t* __restrict__ ff, const uint8_t* __restrict__ nrm,
              const float* __restrict__ ec, float* __restrict__ x, DevScalars* __restrict__ ds,
              int PR_ZC, const __grid_constant__ PeerHalo H) {
  if (ds->pcg_done) return;
  const int lane = threadIdx.x;
  const int i = (fbox.x0 & ~3) + blockIdx.x * 128 + 4 * lane;
  const int j = fbox.y0 + blockIdx.y * blockDim.y + threadIdx.y;
  const int k0c = fbox.z0 + zperm(blockIdx.z, gridDim.z, H.mode ? ZP_EDGE : ZP_UP) * PR_ZC, k1c = min(k0c + PR_ZC, fbox.z1);
  const bool bdn = k0c <= H.kdn && H.kdn < k1c, bup = k0c <= H.kup && H.kup < k1c;
  halo_wait(H, ds, bdn, bup);
  if (j < fbox.y1) {                                // whole warp
    const int I = i >> 1, J = j >> 1;               // coarse columns I, I + 1 (I even)
    int J2 = J + ((j & 1) ? 1 : -1);
    const bool yin = J2 >= 0 && J2 < gc.ny;
    J2 = yin ? J2 : J;
    const float wy = yin ? 1.f : 0.f;
    const bool c0in = I < gc.nx, c1in = I + 1 < gc.nx;
    const int Ie = lane == 0 ? I - 1 : I + 2;         // the column the edge lanes cannot get by shuffle
    const bool ein = (lane == 0 || lane == 31) && Ie >= 0 && Ie < gc.nx;
    auto ycomb = [&](int Kg, float& ye) -> float2 {
      ye = 0.f;
      if (Kg < 0 || Kg >= gc.gnz) return make_float2(0.f, 0.f);     // block-uniform
      const float* r0 = ec + gidx(gc, 0, J, Kg - gc.kg0);
      const float* r1 = ec + gidx(gc, 0, J2, Kg - gc.kg0);
      if (ein) ye = 3.f * r0[Ie] + wy * r1[Ie];
      float2 a = make_float2(0.f, 0.f), b = a;
      if (c0in) { a = *reinterpret_cast<const float2*>(r0 + I); b = *reinterpret_cast<const float2*>(r1 + I); }   // px is even: I + 1 is addressable
      return make_float2(3.f * a.x + wy * b.x, c1in ? 3.f * a.y + wy * b.y : 0.f);
    };
    const int xe = min(gf.px, (fbox.x1 + 3) & ~3);
    const bool act = i < xe;
    int Kc = (k0c + gf.kg0) >> 1;
    float em, e0, ep;
    float2 Ym = ycomb(Kc - 1, em), Y0 = ycomb(Kc, e0), Yp = ycomb(Kc + 1, ep);
    int q = gidx(gf, act ? i : 0, j, k0c);
    uint32_t f = 0, nn = 0x01010101u, f1 = 0, nn1 = 0x01010101u;
    float4 xv = make_float4(0.f, 0.f, 0.f, 0.f), xv1 = xv;
    if (act) {
      f = *reinterpret_cast<const uint32_t*>(ff + q);
      nn = *reinterpret_cast<const uint32_t*>(nrm + q);
      xv = *reinterpret_cast<const float4*>(x + q);
    }
    for (int k = k0c; k < k1c; ++k, q += gf.plane) {
      if (act && k + 1 < k1c) {
        const int qn = q + gf.plane;
        f1 = *reinterpret_cast<const uint32_t*>(ff + qn);
        nn1 = *reinterpret_cast<const uint32_t*>(nrm + qn);
        xv1 = *reinterpret_cast<const float4*>(x + qn);
      }
      const int kg = k + gf.kg0;
      if ((kg >> 1) != Kc) {
        Kc = kg >> 1;
        Ym = Y0; em = e0; Y0 = Yp; e0 = ep;
        Yp = ycomb(Kc + 1, ep);
      }
      const bool up = kg & 1;
      const float E0 = 3.f * Y0.x + (up ? Yp.x : Ym.x);          // column I
      const float E1 = 3.f * Y0.y + (up ? Yp.y : Ym.y);          // column I + 1
      const float Ee = 3.f * e0 + (up ? ep : em);
      float EL = __shfl_up_sync(0xffffffffu, E1, 1), ER = __shfl_down_sync(0xffffffffu, E0, 1);
      if (lane == 0) EL = Ee;
      if (lane == 31) ER = Ee;
      if (f & ((uint32_t)FB_FLUID * 0x01010101u)) {               // any of the four fluid (inactive lanes: f = 0)
        if (f & FB_FLUID) xv.x += __fdividef(3.f * E0 + EL, (float)(nn & 0xffu));                    // node 2I     : I (3), I-1 (1)
        if ((f >> 8) & FB_FLUID) xv.y += __fdividef(3.f * E0 + E1, (float)((nn >> 8) & 0xffu));      // node 2I + 1 : I (3), I+1 (1)
        if ((f >> 16) & FB_FLUID) xv.z += __fdividef(3.f * E1 + E0, (float)((nn >> 16) & 0xffu));    // node 2I + 2 : I+1 (3), I (1)
        if ((f >> 24) & FB_FLUID) xv.w += __fdividef(3.f * E1 + ER, (float)(nn >> 24));              // node 2I + 3 : I+1 (3), I+2 (1)
        *reinterpret_cast<float4*>(x + q) = xv;
        halo_push4(H, q, k, xv);
      }
      f = f1; nn = nn1; xv = xv1;
    }
  }
  halo_signal(H, ds, bdn, bup);
}

// ---- the tail of the V-cycle in ONE block ---------------------------------------------------------
// Levels with <= MG_TAIL_CELLS cells (16^3 and below at 256^3) cost more in launch latency than in
// work, so one 1024-thread block runs that whole part of the cycle: the level loops are plain
// per-cell restatements of k_mg_smooth / k_mg_resid / k_mg_restrict / k_mg_prolong over the level's
// (L1-resident) global arrays with a __syncthreads between the sweeps, and the coarsest level
// (<= MG_COARSEST_CELLS) is solved by MG_CHEB_STEPS steps of the Chebyshev semi-iteration on the
// Jacobi-scaled operator D^-1 A (spectrum inside [MG_CHEB_LMIN, 2]) -- a fixed polynomial, hence
// still a symmetric linear preconditioner, and close to an exact coarse solve (22 vs 27 outer
// iterations against 41 plain Jacobi sweeps, tools/mg_proto4.py).  Its search direction d lives in
// shared memory with a zero halo (ping-pong); x, r of the (<= 4) cells of a thread in registers.
// The arrays are rewritten inside the kernel: no __restrict__ / read-only loads here.
#define MG_CHEB_STEPS 40
#define MG_CHEB_LMIN 0.005f
struct TailLevel {
  Grid g; Box box;
  MgCoef m1, m2;            // weights W1 / W2
  const uint8_t* fl; const uint8_t* nrm;
  float *xa, *xb, *b;
};
struct MgTail { int n, cheb_steps; float cheb_lmin; TailLevel L[MG_TAIL_MAX]; };   // L[0] finest tail level ... L[n-1] coarsest

// the (<= 4) cells of a thread on one tail level: computed once, reused by every sweep of the level
struct TailCells {
  int q[4];                 // array index, -1: no cell
  int ijk[4];               // i | j << 10 | k << 20
  uint8_t f[4];             // flag byte (0 when !fluid: the sweeps only need fluid cells)
  uint8_t in[4];            // bit d set: neighbour d (x-,x+,y-,y+,z-,z+) is inside the grid
};
__device__ __forceinline__ void tail_cells(const TailLevel& L, TailCells& c) {
  const int bx = L.box.x1 - L.box.x0, by = L.box.y1 - L.box.y0, bz = L.box.z1 - L.box.z0;
  const int n = bx * by * bz;
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int e = threadIdx.x + u * 1024;
    c.q[u] = -1; c.f[u] = 0; c.in[u] = 0; c.ijk[u] = 0;
    if (e < n) {
      const int i = L.box.x0 + e % bx, j = L.box.y0 + (e / bx) % by, k = L.box.z0 + e / (bx * by);
      c.q[u] = gidx(L.g, i, j, k);
      c.ijk[u] = i | (j << 10) | (k << 20);
      c.f[u] = L.fl[c.q[u]];
      c.in[u] = (i > 0 ? 1 : 0) | (i < L.g.nx - 1 ? 2 : 0) | (j > 0 ? 4 : 0) | (j < L.g.ny - 1 ? 8 : 0) |
                (k > 0 ? 16 : 0) | (k < L.g.nz - 1 ? 32 : 0);
    }
  }
}
// sum of the six neighbours (0 outside the grid; non-fluid cells hold 0)
__device__ __forceinline__ float tail_nsum(const Grid& g, int q, uint8_t in, const float* x) {
  const float xm = (in & 1) ? x[q - 1] : 0.f, xp = (in & 2) ? x[q + 1] : 0.f;
  const float ym = (in & 4) ? x[q - g.px] : 0.f, yp = (in & 8) ? x[q + g.px] : 0.f;
  const float zm = (in & 16) ? x[q - g.plane] : 0.f, zp = (in & 32) ? x[q + g.plane] : 0.f;
  return (zm + zp) + (ym + yp) + (xm + xp);
}
__device__ __forceinline__ void tail_smooth(const TailLevel& L, const TailCells& c, const MgCoef& mc, const float* xi, float* xo) {
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    if (c.q[u] < 0) continue;
    float o = 0.f;
    if (c.f[u] & FB_FLUID) {
      const int kk = 6 - __popc(c.f[u] & FB_SMASK);
      const float xc = xi[c.q[u]];
      o = xc + mc.winv[kk] * (L.b[c.q[u]] - mc.h2inv * ((float)kk * xc - tail_nsum(L.g, c.q[u], c.in[u], xi)));
    }
    xo[c.q[u]] = o;
  }
  __syncthreads();
}
__device__ __forceinline__ void tail_resid(const TailLevel& L, const TailCells& c, const float* x, float* rn) {
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    if (c.q[u] < 0) continue;
    float o = 0.f;
    if (c.f[u] & FB_FLUID) {
      const int kk = 6 - __popc(c.f[u] & FB_SMASK);
      o = (L.b[c.q[u]] - L.m1.h2inv * ((float)kk * x[c.q[u]] - tail_nsum(L.g, c.q[u], c.in[u], x))) *
          (64.f / (float)L.nrm[c.q[u]]);
    }
    rn[c.q[u]] = o;
  }
  __syncthreads();
}
// C.b = R rn, C.xa = W1 D^-1 C.b  (k_mg_restrict); cc: the thread's cells on the COARSE level
__device__ __forceinline__ void tail_restrict(const TailLevel& F, const TailLevel& C, const TailCells& cc, const float* rn) {
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    if (cc.q[u] < 0) continue;
    float r = 0.f, x1 = 0.f;
    if (cc.f[u] & FB_FLUID) {
      const int ci = cc.ijk[u] & 1023, cj = (cc.ijk[u] >> 10) & 1023, ck = cc.ijk[u] >> 20;
      float sum = 0.f;
#pragma unroll 1
      for (int tz = 0; tz < 4; ++tz) {
        const int kf = 2 * ck + tz - 1;
        if (kf < F.box.z0 || kf >= F.box.z1) continue;
        const float wz = (tz == 1 || tz == 2) ? 3.f : 1.f;
#pragma unroll
        for (int ty = 0; ty < 4; ++ty) {
          const int jf = 2 * cj + ty - 1;
          if (jf < F.box.y0 || jf >= F.box.y1) continue;
          const float wzy = wz * ((ty == 1 || ty == 2) ? 3.f : 1.f);
          const float* row = rn + gidx(F.g, 0, jf, kf);
          float rs = 0.f;
#pragma unroll
          for (int tx = 0; tx < 4; ++tx) {
            const int xf = 2 * ci + tx - 1;
            if (xf >= F.box.x0 && xf < F.box.x1) rs += ((tx == 1 || tx == 2) ? 3.f : 1.f) * row[xf];
          }
          sum += wzy * rs;
        }
      }
      r = sum * (1.f / 512.f);
      x1 = C.m1.winv[6 - __popc(cc.f[u] & FB_SMASK)] * r;
    }
    C.b[cc.q[u]] = r;
    C.xa[cc.q[u]] = x1;
  }
  __syncthreads();
}
// F.xb += P C.xb  (k_mg_prolong); c: the thread's cells on the FINE level
__device__ __forceinline__ void tail_prolong(const TailLevel& F, const TailLevel& C, const TailCells& c) {
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    if (c.q[u] < 0 || !(c.f[u] & FB_FLUID)) continue;
    const int i = c.ijk[u] & 1023, j = (c.ijk[u] >> 10) & 1023, k = c.ijk[u] >> 20;
    const int I = i >> 1, J = j >> 1, K = k >> 1;
    const int I2 = I + ((i & 1) ? 1 : -1), J2 = J + ((j & 1) ? 1 : -1), K2 = K + ((k & 1) ? 1 : -1);
    float sum = 0.f;
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      const int ci = (m & 1) ? I2 : I, cj = (m & 2) ? J2 : J, ck = (m & 4) ? K2 : K;
      if (ci < 0 || ci >= C.g.nx || cj < 0 || cj >= C.g.ny || ck < 0 || ck >= C.g.nz) continue;
      sum += (float)(((m & 1) ? 1 : 3) * ((m & 2) ? 1 : 3) * ((m & 4) ? 1 : 3)) * C.xb[gidx(C.g, ci, cj, ck)];
    }
    F.xb[c.q[u]] += __fdividef(sum, (float)F.nrm[c.q[u]]);
  }
  __syncthreads();
}
// coarsest level: b -> xb
__device__ void tail_cheb(const TailLevel& L, float* smx, int steps, float lmin) {
  const int bx = L.box.x1 - L.box.x0, by = L.box.y1 - L.box.y0, bz = L.box.z1 - L.box.z0;
  const int n = bx * by * bz;
  const int sx = bx + 2, sxy = sx * (by + 2), stot = sxy * (bz + 2);
  float* A = smx; float* B = smx + stot;
  for (int e = threadIdx.x; e < 2 * stot; e += blockDim.x) smx[e] = 0.f;
  const float lmax = 2.f;
  const float theta = 0.5f * (lmax + lmin), delta = 0.5f * (lmax - lmin), sigma = theta / delta;
  float xr[4], rr[4], kinv[4]; int sidx[4], gq[4];
  __syncthreads();
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    int e = threadIdx.x + u * 1024;
    xr[u] = 0.f; rr[u] = 0.f; kinv[u] = 0.f; sidx[u] = -1; gq[u] = 0;
    if (e < n) {
      int li = e % bx, lj = (e / bx) % by, lk = e / (bx * by);
      int q = gidx(L.g, L.box.x0 + li, L.box.y0 + lj, L.box.z0 + lk);
      uint8_t f = L.fl[q];
      sidx[u] = (lk + 1) * sxy + (lj + 1) * sx + li + 1; gq[u] = q;
      if (f & FB_FLUID) {
        int kk = 6 - __popc(f & FB_SMASK);
        kinv[u] = 1.f / (float)kk;
        rr[u] = L.b[q] * kinv[u] / L.m1.h2inv;      // r0 = D^-1 b
        A[sidx[u]] = rr[u] / theta;                 // d0
      }
    }
  }
  __syncthreads();
  float rho = 1.f / sigma;
  for (int s = 0; s < steps; ++s) {
    const float* di = (s & 1) ? B : A;
    float* dn = (s & 1) ? A : B;
    const float rho_new = 1.f / (2.f * sigma - rho);
    const float c1 = rho_new * rho, c2 = 2.f * rho_new / delta;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (sidx[u] < 0 || kinv[u] == 0.f) continue;
      const int q = sidx[u];
      const float dc = di[q];
      const float sn = (di[q - 1] + di[q + 1]) + (di[q - sx] + di[q + sx]) + (di[q - sxy] + di[q + sxy]);
      xr[u] += dc;
      rr[u] -= dc - sn * kinv[u];                   // r -= D^-1 A d
      dn[q] = c1 * dc + c2 * rr[u];
    }
    rho = rho_new;
    __syncthreads();
  }
#pragma unroll
  for (int u = 0; u < 4; ++u)
    if (sidx[u] >= 0) L.xb[gq[u]] = xr[u];
  __syncthreads();
}

template <int N>                                     // N = number of tail levels (T.n)
__global__ void __launch_bounds__(1024)
k_mg_tail(const __grid_constant__ MgTail T, const DevScalars* __restrict__ ds) {
  extern __shared__ float smx[];
  if (ds->pcg_done) return;
  TailCells c[N];                                    // every loop below unrolls: registers
#pragma unroll
  for (int l = 0; l < N; ++l) tail_cells(T.L[l], c[l]);
#pragma unroll
  for (int l = 0; l < N - 1; ++l) {                  // down: b, xa (first pre-sweep) come from the restriction above
    const TailLevel& L = T.L[l];
    tail_smooth(L, c[l], L.m2, L.xa, L.xb);
    tail_resid(L, c[l], L.xb, L.xa);
    tail_restrict(L, T.L[l + 1], c[l + 1], L.xa);
  }
  tail_cheb(T.L[N - 1], smx, T.cheb_steps, T.cheb_lmin);
#pragma unroll
  for (int l = N - 2; l >= 0; --l) {                 // up
    const TailLevel& L = T.L[l];
    tail_prolong(L, T.L[l + 1], c[l]);
    tail_smooth(L, c[l], L.m2, L.xb, L.xa);
    tail_smooth(L, c[l], L.m1, L.xa, L.xb);
  }
}
static TailKernel tail_kernel(int n) {
  switch (n) { case 1: return k_mg_tail<1>; case 2: return k_mg_tail<2>; case 3: return k_mg_tail<3>; default: return k_mg_tail<4>; }
}

// ---- host orchestration ----------------------------------------------------------------------
// z = M^-1 r.  Input: level-0 b (= float(r*sc)) and xa (= first sweep) written by the CG's K2/init;
// output: level-0 xb (fp32, still scaled by sc); r.z is added to ds->rz[it & 3] in fp64.
// z-slab runs: on the distributed levels every sweep is preceded by a one-plane halo
// fill of its input; the first replicated level is assembled with an all-gather of the restricted
// residual and solved redundantly by every rank; r.z is all-reduced at the end.
// Ghost planes of `arr` before a sweep reads them: a copy (NCCL / in-process pulls), or -- with the
// pushes fused into the producing kernels (halo.cuh) -- nothing but the test transport's host barrier.
// planes a restrict / prolong block marches: the largest chunk that still leaves `min_blocks` blocks
// (8 on the 256^3 level, 1 on the latency-bound coarse levels).
static int mg_zchunk(int blocks_xy, int nz, int min_blocks) {
  for (int zc = 8; zc > 1; zc >>= 1)
    if ((int64_t)blocks_xy * cdiv(nz, zc) >= min_blocks) return zc;
  return 1;
}
// the four-node sweeps without the peer code when the run has no z neighbours
template <int DOT, typename... A>
static void launch_smooth4(dim3 g, cudaStream_t st, const PeerHalo& H, A... a) {
  if (H.mode) k_mg_smooth4<DOT, true><<<g, STENCIL_BLOCK, 0, st>>>(a..., H);
  else k_mg_smooth4<DOT, false><<<g, STENCIL_BLOCK, 0, st>>>(a..., H);
}
template <typename... A>
static void launch_resid4(dim3 g, cudaStream_t st, const PeerHalo& H, A... a) {
  if (H.mode) k_mg_resid4<true><<<g, STENCIL_BLOCK, 0, st>>>(a..., H);
  else k_mg_resid4<false><<<g, STENCIL_BLOCK, 0, st>>>(a..., H);
}
#define HALO(L, arr) do { if ((L).dist) { int rc__ = fused ? fb_halo_point(c) : fb_halo_fill(c, (L).g, (arr), sizeof(float), 1, 1); if (rc__) return rc__; } } while (0)
int fb_mg_vcycle(flipb200_ctx* c, int it) {
  const int nl = c->mg_levels;
  const int dist = c->world > 1 ? 1 : 0;
  const bool fused = c->halo.mode != 0;
  const dim3 pblk(32, 8, 1);
  Sweep S[MG_MAX_LEVELS];
  MgCoef mc[MG_MAX_LEVELS], mc2[MG_MAX_LEVELS];   // mc: weight W1 (first pre-sweep, last post-sweep), mc2: W2
  // STENCIL4 (four nodes per lane) on the levels whose rows fill a 128-node warp; FLIPB200_ST4=0: never
  Sweep S4[MG_MAX_LEVELS];
  bool w4[MG_MAX_LEVELS];
  const bool st4_on = c->opt_st4, rz64 = c->opt_rz64;
  const bool h16 = mg_h16_active(c);                // level 0: b0 and the residual stored as fp16 (experimental, pcg.cuh, mg_h16.cu)
  const float rzs = h16 ? (float)(1.0 / (mg_h16_bscale(c) * mg_h16_bscale(c))) : 1.f;
  for (int l = 0; l < nl; ++l) {
    S[l] = make_sweep(c->mg[l].g, c->mg[l].box);
    S4[l] = make_sweep(c->mg[l].g, c->mg[l].box, 4);
    w4[l] = st4_on && (S4[l].xe - S4[l].box.x0) >= 128;
    if (!w4[l] && !c->mg[l].dist) S[l].zc = small_level_zc(S[l]);    // k_mg_smooth / k_mg_resid of the small levels
    if (w4[l] && l >= 1 && !c->mg[l].dist) S4[l].zc = FB_L1_ZC;
    mc[l] = make_mg_coef(c->mg[l].h2inv, MG_OMEGA1); mc2[l] = make_mg_coef(c->mg[l].h2inv, MG_OMEGA2);
  }
  if (sweep_empty(S[0])) return 0;
  // Sweep directions on a single GPU (sweep.cuh ZP_DOWN): k_spmv up, k_update_xr down, pre-smooth up,
  // residual down, restrict up | prolong up, post-smooth 1 down, post-smooth 2 up, k_update_p down.
  const size_t F = sizeof(float);
  const int t0 = c->mg_tail0;
  bool gathered = false;                            // the level below was assembled by peer stores: its first reader waits for them
  for (int l = 0; l < t0; ++l) {
    MgLevel& L = c->mg[l];
    MgLevel& C = c->mg[l + 1];
    const dim3 tg = w4[l] ? stencil_grid(S4[l], 4) : stencil_grid(S[l]);
    const int nbs = tg.x * tg.y, k0 = L.g.k0, k1 = L.g.k1 - 1;
    HALO(L, L.xa);
    {
      PeerHalo Hs = fb_halo_make(c, l, HA_XA, l, HA_XB, F, k0, k1, nbs);
      if (gathered) { fb_gather_wait(c, &Hs); gathered = false; }
      if (l == 0 && h16) fb_h16_smooth4(c, 0, tg, S4[l], mc2[l], L.flags, L.b, L.xa, L.xb, it, rzs, dist, Hs);
      else if (w4[l]) launch_smooth4<0>(tg, c->stream, Hs, S4[l], mc2[l], L.flags, L.b, L.xa, L.xb, (const double*)nullptr, c->ds, it, dist);
      else k_mg_smooth<0><<<tg, STENCIL_BLOCK, 0, c->stream>>>(S[l], mc2[l], L.flags, L.b, L.xa, L.xb, nullptr, c->ds, it, dist, Hs);
    }
    fb_prof(c, l == 0 ? "L0 halo+smooth" : (l == 1 ? "L1 halo+smooth" : "L2+ halo+smooth"));
    HALO(L, L.xb);
    {
      // the level-0 residual is stored (and pushed to the neighbours' ghost planes) as fp16 with params.mixed bit 1
      const PeerHalo Hr = fb_halo_make(c, l, HA_XB, l, HA_XA, (l == 0 && h16) ? sizeof(__half) : F, k0, k1, nbs);
      if (l == 0 && h16) fb_h16_resid4(c, tg, sweep_down(S4[l]), mc[l], L.flags, L.nrm, L.b, L.xb, L.xa, Hr);
      else if (w4[l]) launch_resid4(tg, c->stream, Hr, sweep_down(S4[l]), mc[l], L.flags, L.nrm, L.b, L.xb, L.xa, c->ds);
      else k_mg_resid<<<tg, STENCIL_BLOCK, 0, c->stream>>>(sweep_down(S[l]), mc[l], L.flags, L.nrm, L.b, L.xb, L.xa, c->ds, Hr);
    }
    fb_prof(c, l == 0 ? "L0 halo+resid" : (l == 1 ? "L1 halo+resid" : "L2+ halo+resid"));
    if (l == 0 && h16) {   // the residual in L.xa is an array of halves: its planes are half as wide in bytes
      if (L.dist) { int rc__ = fused ? fb_halo_point(c) : fb_halo_fill(c, L.g, L.xa, sizeof(__half), 1, 1); if (rc__) return rc__; }
    } else
    HALO(L, L.xa);
    Box cb = C.box;
    int ck0 = C.g.k0, ck1 = C.g.k1 - 1;          // boundary planes in the coarse index space
    if (L.dist && !C.dist) {            // this rank restricts into the planes whose children it owns
      cb.z0 = max(C.gbox.z0, C.zs[c->rank]); cb.z1 = min(C.gbox.z1, C.zs[c->rank + 1]);
      ck0 = C.zs[c->rank]; ck1 = C.zs[c->rank + 1] - 1;
    }
    {
      // the sequence number is taken even when this rank has nothing to restrict
      const bool r4 = w4[l] && cb.x1 - (cb.x0 & ~1) >= 64;
      dim3 cg(max(1, r4 ? cdiv(cb.x1 - (cb.x0 & ~1), 64) : cdiv(cb.x1 - cb.x0, 32)), max(1, cdiv(cb.y1 - cb.y0, 8)), 1);
      const int zc = mg_zchunk(cg.x * cg.y, cb.z1 - cb.z0, r4 ? 512 : 1024);
      cg.z = max(1, cdiv(cb.z1 - cb.z0, zc));
      PeerHalo H = fb_halo_make(c, l, HA_XA, C.dist ? l + 1 : -1, C.dist ? HA_XA : -1, F, ck0, ck1, cg.x * cg.y);
      const bool some = cb.z1 > cb.z0 && cb.x1 > cb.x0 && cb.y1 > cb.y0;
      // the restriction into the first replicated level pushes its planes to every rank (halo.cuh GatherPush)
      gathered = L.dist && !C.dist && c->halo.gather_on;
      const GatherPush G = gathered ? fb_gather_make(c, some ? cg.x * cg.y * cg.z : 1) : gather_off();
      if (some) {
        if (l == 0 && h16 && !r4) return FLIPB200_ERR_STATE;       // the fp16 residual has no two-node reader
        if (l == 0 && h16) fb_h16_restrict4(c, cg, pblk, L.g, L.boxe, C.g, cb, mc[l + 1], C.flags, L.xa, C.b, C.xa, zc, H, G);
        else if (r4) k_mg_restrict4<<<cg, pblk, 0, c->stream>>>(L.g, L.boxe, C.g, cb, mc[l + 1], C.flags, L.xa, C.b, C.xa, c->ds, zc, H, G);
        else k_mg_restrict<<<cg, pblk, 0, c->stream>>>(L.g, L.boxe, C.g, cb, mc[l + 1], C.flags, L.xa, C.b, C.xa, c->ds, zc, H, G);
      } else if (gathered && G.mode == 2) k_mg_gather_signal<<<1, 32, 0, c->stream>>>(c->ds, G);
    }
    fb_prof(c, l == 0 ? "L0 halo+restrict" : (l == 1 ? "L1 halo+restrict" : "L2+ halo+restrict"));
    if (L.dist && !C.dist) {
      int rc = gathered ? fb_halo_point(c) : fb_allgather_planes2(c, C.g, C.b, C.xa, sizeof(float), C.zs);
      if (rc) return rc;
      fb_prof(c, gathered ? "gather b, xa (peer stores)" : "allgather b, xa");
    }
    c->launches += 3;
  }
  {
    MgTail T;
    memset(&T, 0, sizeof(T));
    T.n = nl - t0;
    for (int l = t0; l < nl; ++l) {
      MgLevel& L = c->mg[l];
      TailLevel& t = T.L[l - t0];
      t.g = L.g; t.box = L.box; t.m1 = mc[l]; t.m2 = mc2[l];
      t.fl = L.flags; t.nrm = L.nrm; t.xa = L.xa; t.xb = L.xb; t.b = L.b;
    }
    const Box& b = c->mg[nl - 1].box;
    // 40 steps on [0.005, 2] also on the 8^3 coarsest level: shorter polynomials on a narrower interval
    // (30 on [0.01, 2], 24 on [0.01, 2] ...) saved 5 us per cycle but cost 1-2 CG iterations per step.
    T.cheb_steps = MG_CHEB_STEPS;
    T.cheb_lmin = MG_CHEB_LMIN;
    if (c->opt_cheb_steps > 0) { T.cheb_steps = c->opt_cheb_steps; T.cheb_lmin = c->opt_cheb_lmin; }   // FLIPB200_CHEB (experiments)
    size_t smem = 2 * sizeof(float) * (size_t)(b.x1 - b.x0 + 2) * (b.y1 - b.y0 + 2) * (b.z1 - b.z0 + 2);
    if (smem > 160 * 1024 || (int64_t)(b.x1 - b.x0) * (b.y1 - b.y0) * (b.z1 - b.z0) > 4096) return FLIPB200_ERR_ARG;
    tail_kernel(T.n)<<<1, 1024, smem, c->stream>>>(T, c->ds);
    c->launches++;
    fb_prof(c, "tail (small levels + coarsest)");
  }
  for (int l = t0 - 1; l >= 0; --l) {
    MgLevel& L = c->mg[l];
    MgLevel& C = c->mg[l + 1];
    const Box& fb = L.box;
    dim3 pg(w4[l] ? cdiv(((fb.x1 + 3) & ~3) - (fb.x0 & ~3), 128) : cdiv(((fb.x1 + 1) & ~1) - (fb.x0 & ~1), 64), cdiv(fb.y1 - fb.y0, 8), 1);
    const int pzc = mg_zchunk(pg.x * pg.y, fb.z1 - fb.z0, w4[l] ? 512 : 1024);
    pg.z = cdiv(fb.z1 - fb.z0, pzc);
    const dim3 tg = w4[l] ? stencil_grid(S4[l], 4) : stencil_grid(S[l]);
    const int nbs = tg.x * tg.y, k0 = L.g.k0, k1 = L.g.k1 - 1;
    HALO(C, C.xb);
    {
      PeerHalo PH = fb_halo_make(c, C.dist ? l + 1 : -1, C.dist ? HA_XB : -1, l, HA_XB, F, k0, k1, pg.x * pg.y);
      if (w4[l]) k_mg_prolong4<<<pg, pblk, 0, c->stream>>>(L.g, fb, C.g, L.flags, L.nrm, C.xb, L.xb, c->ds, pzc, PH);
      else k_mg_prolong<<<pg, pblk, 0, c->stream>>>(L.g, fb, C.g, L.flags, L.nrm, C.xb, L.xb, c->ds, pzc, PH);
    }
    fb_prof(c, l == 0 ? "L0 halo+prolong" : (l == 1 ? "L1 halo+prolong" : "L2+ halo+prolong"));
    HALO(L, L.xb);
    {
      const PeerHalo H1 = fb_halo_make(c, l, HA_XB, l, HA_XA, F, k0, k1, nbs);
      if (l == 0 && h16) fb_h16_smooth4(c, 0, tg, sweep_down(S4[l]), mc2[l], L.flags, L.b, L.xb, L.xa, it, rzs, dist, H1);
      else if (w4[l]) launch_smooth4<0>(tg, c->stream, H1, sweep_down(S4[l]), mc2[l], L.flags, L.b, L.xb, L.xa, (const double*)nullptr, c->ds, it, dist);
      else k_mg_smooth<0><<<tg, STENCIL_BLOCK, 0, c->stream>>>(sweep_down(S[l]), mc2[l], L.flags, L.b, L.xb, L.xa, nullptr, c->ds, it, dist, H1);
    }
    fb_prof(c, l == 0 ? "L0 halo+smooth (post 1)" : (l == 1 ? "L1 halo+smooth (post 1)" : "L2+ halo+smooth (post 1)"));
    HALO(L, L.xa);
    // the last sweep's output is read through ghosts only as the coarse correction of the level above
    if (l == 0) {
      const PeerHalo Hd = fb_halo_make(c, l, HA_XA, -1, -1, F, k0, k1, nbs);
      if (h16) fb_h16_smooth4(c, 2, tg, S4[l], mc[l], L.flags, L.b, L.xa, L.xb, it, rzs, dist, Hd);
      else if (w4[l] && rz64) launch_smooth4<1>(tg, c->stream, Hd, S4[l], mc[l], L.flags, L.b, L.xa, L.xb, (const double*)c->r, c->ds, it, dist);
      else if (w4[l]) launch_smooth4<2>(tg, c->stream, Hd, S4[l], mc[l], L.flags, L.b, L.xa, L.xb, (const double*)c->r, c->ds, it, dist);
      else if (rz64) k_mg_smooth<1><<<tg, STENCIL_BLOCK, 0, c->stream>>>(S[l], mc[l], L.flags, L.b, L.xa, L.xb, c->r, c->ds, it, dist, Hd);
      else k_mg_smooth<2><<<tg, STENCIL_BLOCK, 0, c->stream>>>(S[l], mc[l], L.flags, L.b, L.xa, L.xb, c->r, c->ds, it, dist, Hd);
    }
    else {
      const PeerHalo H2 = fb_halo_make(c, l, HA_XA, l, HA_XB, F, k0, k1, nbs);
      if (w4[l]) launch_smooth4<0>(tg, c->stream, H2, S4[l], mc[l], L.flags, L.b, L.xa, L.xb, (const double*)nullptr, c->ds, it, dist);
      else k_mg_smooth<0><<<tg, STENCIL_BLOCK, 0, c->stream>>>(S[l], mc[l], L.flags, L.b, L.xa, L.xb, nullptr, c->ds, it, dist, H2);
    }
    fb_prof(c, l == 0 ? "L0 halo+smooth (post 2)" : (l == 1 ? "L1 halo+smooth (post 2)" : "L2+ halo+smooth (post 2)"));
    c->launches += 3;
  }
  CK_LAUNCH();
  if (dist) {
    int rc = fb_pcg_reduce_dot(c, PART_RZ, it);
    if (rc) return rc;
  }
  fb_prof(c, "allreduce rz");
  return 0;
}
