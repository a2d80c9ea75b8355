// mg.cu -- geometric multigrid V-cycle used as the (IC-free) preconditioner of the pressure CG.
//
// The reference preconditions with Eigen's IncompleteCholesky (fluid.cc:1352,1473); the north_star
// asks for a "Jacobi/IC-free" preconditioner.  Jacobi-PCG needs ~10 n iterations on an n^3 tank
// (2600 at 256^3); this V-cycle brings that to ~25, independent of n (tools/mg_proto*.py).
//
// Design (after McAdams, Sifakis, Teran 2010, adapted to the reference's cell classes):
//  * level l+1 cell (I,J,K) covers level-l cells (2I..2I+1, 2J..2J+1, 2K..2K+1); children outside
//    the grid count as solid.  Coarse type: AIR if any child is air (Dirichlet wins), else FLUID if
//    any child is fluid, else SOLID.
//  * operator on every level: rediscretised 7-point, A x = h2inv*(k x_c - sum fluid nbrs), k = number
//    of non-solid neighbours (air = Dirichlet 0, solid = Neumann); h2inv_l = scale / 4^l.
//  * smoother: damped Jacobi (omega = 0.8), 2 pre + 2 post sweeps (symmetric).
//  * prolongation P: cell-centred trilinear (3/4,1/4 per axis) with the weights of SOLID coarse
//    cells dropped and the rest renormalised (constant extrapolation at Neumann walls);
//    restriction R = P^T / 8.  => M^-1 is symmetric positive definite, as CG requires.
//  * coarsest level (<= 4096 cells): 2x20 Jacobi sweeps in one thread block.
#include "common.cuh"

#define MG_OMEGA 0.8
#define MG_COARSE_SWEEPS 40
#define MG_COARSEST_CELLS 4096

static void mg_level_dims(const Grid& f, Grid* c) {
  c->nx = (f.nx + 1) / 2; c->ny = (f.ny + 1) / 2; c->nz = (f.nz + 1) / 2;
  c->px = (c->nx + 31) & ~31;
  c->ox = c->oy = c->oz = 0; c->wall = 0;
  c->plane = c->px * c->ny;
  c->total = c->plane * c->nz;
}

int fb_mg_alloc(flipb200_ctx* c) {
  c->mg_levels = 0;
  if (c->P.precond != 1) return 0;
  MgLevel& L0 = c->mg[0];
  L0.g = c->g; L0.flags = c->flags; L0.b = c->r; L0.t = c->Ap;
  CK(cudaMalloc((void**)&c->mgx, (size_t)c->g.total * sizeof(double)));
  CK(cudaMemsetAsync(c->mgx, 0, (size_t)c->g.total * sizeof(double), c->stream));
  L0.x = c->mgx;
  int nl = 1;
  while (nl < MG_MAX_LEVELS) {
    const Grid& f = c->mg[nl - 1].g;
    if ((int64_t)f.nx * f.ny * f.nz <= MG_COARSEST_CELLS || min(f.nx, min(f.ny, f.nz)) < 4) break;
    MgLevel& L = c->mg[nl];
    mg_level_dims(f, &L.g);
    size_t T = (size_t)L.g.total;
    CK(cudaMalloc((void**)&L.flags, T));
    CK(cudaMalloc((void**)&L.x, T * sizeof(double)));
    CK(cudaMalloc((void**)&L.t, T * sizeof(double)));
    CK(cudaMalloc((void**)&L.b, T * sizeof(double)));
    CK(cudaMemsetAsync(L.flags, FB_SOLID | FB_SMASK, T, c->stream));
    CK(cudaMemsetAsync(L.x, 0, T * sizeof(double), c->stream));
    CK(cudaMemsetAsync(L.t, 0, T * sizeof(double), c->stream));
    CK(cudaMemsetAsync(L.b, 0, T * sizeof(double), c->stream));
    ++nl;
  }
  if (nl < 2) { cudaFree(c->mgx); c->mgx = nullptr; nl = 0; }   // grid too small: Jacobi preconditioner
  c->mg_levels = nl;
  return 0;
}

void fb_mg_free(flipb200_ctx* c) {
  if (c->mg_levels == 0) return;
  cudaFree(c->mgx);
  for (int l = 1; l < c->mg_levels; ++l) {
    cudaFree(c->mg[l].flags); cudaFree(c->mg[l].x); cudaFree(c->mg[l].t); cudaFree(c->mg[l].b);
  }
  c->mg_levels = 0;
}

// ---- per-step hierarchy: coarse cell types, then solid-neighbour masks -------------------------
__global__ void k_mg_coarsen(Grid gf, Grid gc, const uint8_t* __restrict__ ff, uint8_t* __restrict__ fc) {
  int I = blockIdx.x * blockDim.x + threadIdx.x;
  int J = blockIdx.y * blockDim.y + threadIdx.y;
  int K = blockIdx.z;
  if (I >= gc.px || J >= gc.ny) return;
  int q = gidx(gc, I, J, K);
  if (I >= gc.nx) { fc[q] = FB_SOLID; return; }
  bool any_air = false, any_fluid = false;
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    int i = 2 * I + (m & 1), j = 2 * J + ((m >> 1) & 1), k = 2 * K + (m >> 2);
    if (i < gf.nx && j < gf.ny && k < gf.nz) {
      uint8_t f = ff[gidx(gf, i, j, k)];
      if (f & FB_FLUID) any_fluid = true;
      else if (!(f & FB_SOLID)) any_air = true;
    }
  }
  fc[q] = any_air ? 0 : (any_fluid ? FB_FLUID : FB_SOLID);
}

__global__ void k_mg_nbrmask(Grid g, uint8_t* __restrict__ fl) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int j = blockIdx.y * blockDim.y + threadIdx.y;
  int k = blockIdx.z;
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
  // note: only bits 6/7 of the neighbours are read above and only bits 0..5 (+ bit 6) of this
  // cell change, so the in-place update is race-free.
  if ((f & FB_FLUID) && m == FB_SMASK) f &= ~FB_FLUID;   // no non-solid neighbour: empty row
  fl[q] = f | m;
}

int fb_mg_setup(flipb200_ctx* c, double dt) {
  if (c->mg_levels == 0) return 0;
  const double scale = dt / (c->P.rho * c->P.dx * c->P.dx);   // fluid.cc:306
  c->mg[0].box = c->box;
  c->mg[0].h2inv = scale;
  c->mg[0].g = c->g;
  for (int l = 1; l < c->mg_levels; ++l) {
    MgLevel& F = c->mg[l - 1];
    MgLevel& L = c->mg[l];
    dim3 blk(32, 8, 1), grd(cdiv(L.g.px, 32), cdiv(L.g.ny, 8), L.g.nz);
    k_mg_coarsen<<<grd, blk, 0, c->stream>>>(F.g, L.g, F.flags, L.flags);
    k_mg_nbrmask<<<grd, blk, 0, c->stream>>>(L.g, L.flags);
    c->launches += 2;
    L.h2inv = F.h2inv * 0.25;
    // coarse box: covers the parents of the fine box, +1 cell margin for the prolongation reads
    Box b;
    b.x0 = max(0, F.box.x0 / 2 - 1) & ~1; b.y0 = max(0, F.box.y0 / 2 - 1); b.z0 = max(0, F.box.z0 / 2 - 1);
    b.x1 = min(L.g.nx, (F.box.x1 + 1) / 2 + 1); b.y1 = min(L.g.ny, (F.box.y1 + 1) / 2 + 1);
    b.z1 = min(L.g.nz, (F.box.z1 + 1) / 2 + 1);
    if (!(F.box.x1 > F.box.x0)) b = Box{0, 0, 0, 0, 0, 0};
    L.box = b;
  }
  CK_LAUNCH();
  return 0;
}

// ---- V-cycle kernels -----------------------------------------------------------------------------
// All sweep the level's box with one thread per node; non-fluid nodes inside the box are written
// as 0 so that the 6-neighbour sum needs no masks (same convention as pcg.cu).
struct MgArgs {
  Grid g; Box box; double h2inv;
};

#define MG_NODE_PROLOGUE                                             \
  int i = a.box.x0 + blockIdx.x * blockDim.x + threadIdx.x;          \
  int j = a.box.y0 + blockIdx.y * blockDim.y + threadIdx.y;          \
  int k = a.box.z0 + blockIdx.z;                                     \
  bool inb = i < a.box.x1 && j < a.box.y1;

// sum of the six neighbours; solid / out-of-grid neighbours (mask bit set) are not loaded: coarse
// levels have no solid shell, so a fluid cell may sit on the grid boundary.
__device__ __forceinline__ double mg_nbr_sum(const double* __restrict__ x, int q, const Grid& g, uint8_t f) {
  double s = 0.0;
  if (!(f & FB_SXM)) s += x[q - 1];
  if (!(f & FB_SXP)) s += x[q + 1];
  if (!(f & FB_SYM)) s += x[q - g.px];
  if (!(f & FB_SYP)) s += x[q + g.px];
  if (!(f & FB_SZM)) s += x[q - g.plane];
  if (!(f & FB_SZP)) s += x[q + g.plane];
  return s;
}

// first Jacobi sweep from x = 0:  x = omega D^-1 b
__global__ void __launch_bounds__(256)
k_mg_smooth_first(MgArgs a, const uint8_t* __restrict__ fl, const double* __restrict__ b, double* __restrict__ xo,
                  const DevScalars* __restrict__ ds) {
  if (ds->pcg_done) return;
  MG_NODE_PROLOGUE
  if (!inb) return;
  int q = gidx(a.g, i, j, k);
  uint8_t f = fl[q];
  double v = 0.0;
  if (f & FB_FLUID) {
    int kk = 6 - __popc(f & FB_SMASK);
    v = MG_OMEGA * b[q] / ((double)kk * a.h2inv);
  }
  xo[q] = v;
}

// Jacobi sweep: xo = xi + omega D^-1 (b - A xi);  DOT: also accumulate sum(b * xo) into ds->rz[slot]
template <bool DOT>
__global__ void __launch_bounds__(256)
k_mg_smooth(MgArgs a, const uint8_t* __restrict__ fl, const double* __restrict__ b, const double* __restrict__ xi,
            double* __restrict__ xo, DevScalars* __restrict__ ds, int slot) {
  __shared__ double sm[32];
  if (ds->pcg_done) return;
  MG_NODE_PROLOGUE
  double acc[1] = {0.0};
  if (inb) {
    int q = gidx(a.g, i, j, k);
    uint8_t f = fl[q];
    double v = 0.0;
    if (f & FB_FLUID) {
      int kk = 6 - __popc(f & FB_SMASK);
      double xc = xi[q];
      double s = mg_nbr_sum(xi, q, a.g, f);
      double bq = b[q];
      double Ax = a.h2inv * ((double)kk * xc - s);
      v = xc + MG_OMEGA * (bq - Ax) / ((double)kk * a.h2inv);
      if (DOT) acc[0] = bq * v;
    }
    xo[q] = v;
  }
  if (DOT) {
    block_sum<1>(acc, sm);
    if (threadIdx.x == 0 && threadIdx.y == 0 && acc[0] != 0.0) atomicAdd(&ds->rz[slot], acc[0]);
  }
}

// res = b - A x
__global__ void __launch_bounds__(256)
k_mg_residual(MgArgs a, const uint8_t* __restrict__ fl, const double* __restrict__ b, const double* __restrict__ x,
              double* __restrict__ res, const DevScalars* __restrict__ ds) {
  if (ds->pcg_done) return;
  MG_NODE_PROLOGUE
  if (!inb) return;
  int q = gidx(a.g, i, j, k);
  uint8_t f = fl[q];
  double v = 0.0;
  if (f & FB_FLUID) {
    int kk = 6 - __popc(f & FB_SMASK);
    double s = mg_nbr_sum(x, q, a.g, f);
    v = b[q] - a.h2inv * ((double)kk * x[q] - s);
  }
  res[q] = v;
}

// per-axis geometry of the transfer stencil: fine cell at offset o in {-1,0,1,2} from 2I has
// parent I+pr (weight 3/4) and second donor I+sr (weight 1/4)
__device__ __forceinline__ void mg_axis(int o, int& pr, int& sr) {
  pr = (o < 0) ? -1 : (o > 1 ? 1 : 0);
  sr = (o == -1 || o == 2) ? 0 : (o == 0 ? -1 : 1);
}

// restriction  b_c[I] = (1/8) sum_i P[i,I] res_f[i]  with P's row normalisation (see header)
__global__ void __launch_bounds__(256)
k_mg_restrict(MgArgs fa, MgArgs ca, const uint8_t* __restrict__ fc, const double* __restrict__ res,
              double* __restrict__ bc, const DevScalars* __restrict__ ds) {
  if (ds->pcg_done) return;
  int I = ca.box.x0 + blockIdx.x * blockDim.x + threadIdx.x;
  int J = ca.box.y0 + blockIdx.y * blockDim.y + threadIdx.y;
  int K = ca.box.z0 + blockIdx.z;
  if (I >= ca.box.x1 || J >= ca.box.y1) return;
  int qc = gidx(ca.g, I, J, K);
  uint8_t f = fc[qc];
  if (!(f & FB_FLUID)) { bc[qc] = 0.0; return; }
  // non-solid mask of the 3x3x3 coarse neighbourhood
  unsigned m27 = 0;
#pragma unroll
  for (int dz = -1; dz <= 1; ++dz)
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
      for (int dx = -1; dx <= 1; ++dx) {
        int a = I + dx, b = J + dy, cc = K + dz;
        bool ns = in_box(ca.g, a, b, cc) && !(fc[gidx(ca.g, a, b, cc)] & FB_SOLID);
        if (ns) m27 |= 1u << (((dz + 1) * 3 + (dy + 1)) * 3 + (dx + 1));
      }
  const bool all = m27 == 0x7FFFFFFu;
  double sum = 0.0;
  for (int oz = -1; oz <= 2; ++oz) {
    int kf = 2 * K + oz;
    if (kf < fa.box.z0 || kf >= fa.box.z1) continue;
    int pz, sz; mg_axis(oz, pz, sz);
    double wz = (oz == 0 || oz == 1) ? 0.75 : 0.25;
    for (int oy = -1; oy <= 2; ++oy) {
      int jf = 2 * J + oy;
      if (jf < fa.box.y0 || jf >= fa.box.y1) continue;
      int py, sy; mg_axis(oy, py, sy);
      double wy = (oy == 0 || oy == 1) ? 0.75 : 0.25;
#pragma unroll
      for (int ox = -1; ox <= 2; ++ox) {
        int xf = 2 * I + ox;
        if (xf < fa.box.x0 || xf >= fa.box.x1) continue;
        double r = res[gidx(fa.g, xf, jf, kf)];
        if (r == 0.0) continue;
        double wx = (ox == 0 || ox == 1) ? 0.75 : 0.25;
        double w = wx * wy * wz;
        if (!all) {
          int px_, sx; mg_axis(ox, px_, sx);
          double norm = 0.0;
#pragma unroll
          for (int t = 0; t < 8; ++t) {
            int rx = (t & 1) ? sx : px_, ry = (t & 2) ? sy : py, rz = (t & 4) ? sz : pz;
            double ww = ((t & 1) ? 0.25 : 0.75) * ((t & 2) ? 0.25 : 0.75) * ((t & 4) ? 0.25 : 0.75);
            if (m27 & (1u << (((rz + 1) * 3 + (ry + 1)) * 3 + (rx + 1)))) norm += ww;
          }
          w /= norm;   // norm >= 27/64: the parent of a fluid cell is never solid
        }
        sum += w * r;
      }
    }
  }
  bc[qc] = 0.125 * sum;
}

// prolongation + correction:  t_f[i] += sum_I P[i,I] e_c[I]
__global__ void __launch_bounds__(256)
k_mg_prolong(MgArgs a, MgArgs ca, const uint8_t* __restrict__ ff, const uint8_t* __restrict__ fc,
             const double* __restrict__ ec, double* __restrict__ t, const DevScalars* __restrict__ ds) {
  if (ds->pcg_done) return;
  MG_NODE_PROLOGUE
  if (!inb) return;
  int q = gidx(a.g, i, j, k);
  if (!(ff[q] & FB_FLUID)) return;
  int I = i >> 1, J = j >> 1, K = k >> 1;
  int sx = (i & 1) ? 1 : -1, sy = (j & 1) ? 1 : -1, sz = (k & 1) ? 1 : -1;
  double norm = 0.0, sum = 0.0;
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    int ci = I + ((m & 1) ? sx : 0), cj = J + ((m & 2) ? sy : 0), ck = K + ((m & 4) ? sz : 0);
    if (!in_box(ca.g, ci, cj, ck)) continue;
    int qc = gidx(ca.g, ci, cj, ck);
    uint8_t f = fc[qc];
    if (f & FB_SOLID) continue;
    double w = ((m & 1) ? 0.25 : 0.75) * ((m & 2) ? 0.25 : 0.75) * ((m & 4) ? 0.25 : 0.75);
    norm += w;
    if (f & FB_FLUID) sum += w * ec[qc];
  }
  t[q] += sum / norm;
}

// coarsest level: MG_COARSE_SWEEPS+1 damped-Jacobi sweeps in one block; result in L.t
__global__ void __launch_bounds__(1024)
k_mg_coarsest(MgArgs a, const uint8_t* __restrict__ fl, const double* __restrict__ b, double* __restrict__ x,
              double* __restrict__ t, const DevScalars* __restrict__ ds) {
  if (ds->pcg_done) return;
  const int bx = a.box.x1 - a.box.x0, by = a.box.y1 - a.box.y0, bz = a.box.z1 - a.box.z0;
  const int n = bx * by * bz;
  for (int s = 0; s <= MG_COARSE_SWEEPS; ++s) {
    const double* xi = (s & 1) ? t : x;
    double* xo = (s & 1) ? x : t;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
      int i = a.box.x0 + e % bx, j = a.box.y0 + (e / bx) % by, k = a.box.z0 + e / (bx * by);
      int q = gidx(a.g, i, j, k);
      uint8_t f = fl[q];
      double v = 0.0;
      if (f & FB_FLUID) {
        int kk = 6 - __popc(f & FB_SMASK);
        double d = (double)kk * a.h2inv;
        if (s == 0) v = MG_OMEGA * b[q] / d;
        else {
          double xc = xi[q];
          double sn = mg_nbr_sum(xi, q, a.g, f);
          v = xc + MG_OMEGA * (b[q] - a.h2inv * ((double)kk * xc - sn)) / d;
        }
      }
      xo[q] = v;
    }
    __syncthreads();
  }
}

// ---- host orchestration ----------------------------------------------------------------------
static inline MgArgs mg_args(const MgLevel& L) { MgArgs a; a.g = L.g; a.box = L.box; a.h2inv = L.h2inv; return a; }
static inline bool mg_grid(const MgLevel& L, dim3* grd) {
  const Box& b = L.box;
  if (!(b.x1 > b.x0 && b.y1 > b.y0 && b.z1 > b.z0)) return false;
  *grd = dim3(cdiv(b.x1 - b.x0, 32), cdiv(b.y1 - b.y0, 8), b.z1 - b.z0);
  return true;
}

// z = M^-1 r : input r = mg[0].b (the CG residual), output mg[0].t; r.z is added to ds->rz[it & 3].
int fb_mg_vcycle(flipb200_ctx* c, int it) {
  const int nl = c->mg_levels;
  const dim3 blk(32, 8, 1);
  dim3 grd;
  for (int l = 0; l < nl - 1; ++l) {
    MgLevel& L = c->mg[l];
    MgLevel& C = c->mg[l + 1];
    if (!mg_grid(L, &grd)) return 0;
    MgArgs a = mg_args(L), ca = mg_args(C);
    k_mg_smooth_first<<<grd, blk, 0, c->stream>>>(a, L.flags, L.b, L.x, c->ds);
    k_mg_smooth<false><<<grd, blk, 0, c->stream>>>(a, L.flags, L.b, L.x, L.t, c->ds, 0);
    k_mg_residual<<<grd, blk, 0, c->stream>>>(a, L.flags, L.b, L.t, L.x, c->ds);
    dim3 cg;
    if (!mg_grid(C, &cg)) return 0;
    k_mg_restrict<<<cg, blk, 0, c->stream>>>(a, ca, C.flags, L.x, C.b, c->ds);
    c->launches += 4;
  }
  {
    MgLevel& L = c->mg[nl - 1];
    k_mg_coarsest<<<1, 1024, 0, c->stream>>>(mg_args(L), L.flags, L.b, L.x, L.t, c->ds);
    c->launches++;
  }
  for (int l = nl - 2; l >= 0; --l) {
    MgLevel& L = c->mg[l];
    MgLevel& C = c->mg[l + 1];
    mg_grid(L, &grd);
    MgArgs a = mg_args(L), ca = mg_args(C);
    k_mg_prolong<<<grd, blk, 0, c->stream>>>(a, ca, L.flags, C.flags, C.t, L.t, c->ds);
    k_mg_smooth<false><<<grd, blk, 0, c->stream>>>(a, L.flags, L.b, L.t, L.x, c->ds, 0);
    if (l == 0) k_mg_smooth<true><<<grd, blk, 0, c->stream>>>(a, L.flags, L.b, L.x, L.t, c->ds, it & 3);
    else k_mg_smooth<false><<<grd, blk, 0, c->stream>>>(a, L.flags, L.b, L.x, L.t, c->ds, 0);
    c->launches += 3;
  }
  CK_LAUNCH();
  return 0;
}
