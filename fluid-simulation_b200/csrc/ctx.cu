// ctx.cu -- context lifetime, particle/grid transfer, error strings (C ABI of include/flipb200.h).
#include "common.cuh"
#include <string.h>
#include <stdlib.h>
#include <vector>

extern "C" {

void flipb200_default_params(flipb200_params* p) {
  memset(p, 0, sizeof(*p));
  p->nx = p->ny = p->nz = 121;   // fluid.cc:1159
  p->ox = p->oy = p->oz = 60;
  p->wall = 2;                   // fluid.cc:1264
  p->relax_div = 10;             // fluid.cc:1475
  p->dx = 1.0; p->rho = 1.0;     // fluid.cc:1358,1471
  p->gx = 0.0; p->gy = -10.0; p->gz = 0.0;   // fluid.cc:1357
  p->dt_max = 0.1;               // fluid.cc:1490
  p->relax_tol = 0.1;            // fluid.cc:1484
  p->restitution = 0.0;          // fluid.cc:974
  p->flip_ratio = 1.0;           // fluid.cc:981
  p->pcg_tol = 2.220446049250313e-16;   // NumTraits<double>::epsilon(), IterativeSolverBase.h:363
  p->pcg_max_iter = -1;
  p->max_passes = 200;
  p->precond = 1;                // multigrid V-cycle (IC-free); 0 = Jacobi
  p->device = -1;
  p->track_ids = 0;
  p->extrapolate = 0;            // fluid.cc:1147: the call is commented out
  p->max_particles = 0;
  p->warm_start = 1;
  p->mixed = 3;
  p->precision = 64;
}

const char* flipb200_strerror(int s) {
  switch (s) {
    case FLIPB200_OK: return "ok";
    case FLIPB200_ERR_ARG: return "bad argument";
    case FLIPB200_ERR_CUDA: return "CUDA error (no device, out of memory or launch failure)";
    case FLIPB200_ERR_BREAKDOWN: return "PCG breakdown (p.Ap <= 0 or non-finite)";
    case FLIPB200_ERR_NOCONV: return "PCG reached the iteration limit";
    case FLIPB200_ERR_STATE: return "stage called out of order";
    case FLIPB200_ERR_LIMIT: return "size exceeds a compiled-in limit (FLIPB200_NOALLOC_MAX_NODES)";
    case FLIPB200_ERR_COMM: return "multi-GPU communication error";
    default: return "unknown status";
  }
}

int flipb200_version(void) { return FLIPB200_VERSION; }

}  // extern "C"

// static solid-neighbour mask (the solid grid never changes after scene setup, fluid.cc:1256-1331)
__global__ void k_static_flags(Grid g, const uint8_t* __restrict__ solid, uint8_t* __restrict__ fl) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int j = blockIdx.y * blockDim.y + threadIdx.y;
  int k = blockIdx.z;
  if (i >= g.px || j >= g.ny) return;
  int q = gidx(g, i, j, k);
  if (i >= g.nx || k + g.kg0 < 0 || k + g.kg0 >= g.gnz) { fl[q] = FB_SOLID | FB_SMASK; return; }   // pitch padding / ghost planes past the box: inert
  auto sol = [&](int a, int b, int cc) -> bool {
    return in_box(g, a, b, cc) && solid[gidx(g, a, b, cc)] != 0;   // out of box: VDB background => not solid
  };
  uint8_t f = 0;
  if (sol(i, j, k)) f |= FB_SOLID;
  if (sol(i - 1, j, k)) f |= FB_SXM;
  if (sol(i + 1, j, k)) f |= FB_SXP;
  if (sol(i, j - 1, k)) f |= FB_SYM;
  if (sol(i, j + 1, k)) f |= FB_SYP;
  if (sol(i, j, k - 1)) f |= FB_SZM;
  if (sol(i, j, k + 1)) f |= FB_SZP;
  fl[q] = f;
}

template <typename T>
static int alloc_n(flipb200_ctx* c, T** p, size_t n) {
  CK(cudaMalloc((void**)p, n * sizeof(T)));
  CK(cudaMemsetAsync(*p, 0, n * sizeof(T), c->stream));
  return 0;
}

int fb_reserve_particles(flipb200_ctx* c, int64_t n) {
  Particles& pt = c->pt;
  if (n <= pt.cap) return 0;
  // grow: keep contents of the current buffers
  int64_t ncap = n + n / 16 + 1024;
  ncap = (ncap + 63) & ~(int64_t)63;
  double* na[6]; double* nb[6];
  for (int q = 0; q < 6; ++q) {
    CK(cudaMalloc((void**)&na[q], ncap * sizeof(double)));
    CK(cudaMalloc((void**)&nb[q], ncap * sizeof(double)));
    if (pt.n > 0) CK(cudaMemcpyAsync(na[q], pt.a[q], pt.n * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  }
  uint32_t *nk, *nr;
  CK(cudaMalloc((void**)&nk, ncap * sizeof(uint32_t)));
  CK(cudaMalloc((void**)&nr, ncap * sizeof(uint32_t)));
  uint32_t *ni = nullptr, *nib = nullptr;
  if (c->P.track_ids) {
    CK(cudaMalloc((void**)&ni, ncap * sizeof(uint32_t)));
    CK(cudaMalloc((void**)&nib, ncap * sizeof(uint32_t)));
    if (pt.n > 0) CK(cudaMemcpyAsync(ni, pt.id, pt.n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, c->stream));
  }
  CK(cudaStreamSynchronize(c->stream));
  for (int q = 0; q < 6; ++q) { cudaFree(pt.a[q]); cudaFree(pt.b[q]); pt.a[q] = na[q]; pt.b[q] = nb[q]; }
  cudaFree(pt.key); cudaFree(pt.rank); cudaFree(pt.id); cudaFree(pt.id_b);
  pt.key = nk; pt.rank = nr; pt.id = ni; pt.id_b = nib; pt.cap = ncap;
    // key / rank arrays were reallocated
  return 0;
}

__global__ void k_iota(uint32_t* __restrict__ id, int64_t first, int64_t n, int64_t id0) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) id[first + i] = (uint32_t)(id0 + i);
}
// ids of the particles [first, first+n) := id0, id0+1, ...
int fb_init_ids(flipb200_ctx* c, int64_t first, int64_t n, int64_t id0) {
  if (!c->P.track_ids || n <= 0) return 0;
  k_iota<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->pt.id, first, n, id0);
  CK_LAUNCH(); c->launches++;
  return 0;
}

// scatter back to the caller's order: dst[id[i]] = src[i]
__global__ void k_unpermute(int64_t n, const uint32_t* __restrict__ id,
                            const double* __restrict__ a0, const double* __restrict__ a1, const double* __restrict__ a2,
                            const double* __restrict__ a3, const double* __restrict__ a4, const double* __restrict__ a5,
                            double* __restrict__ b0, double* __restrict__ b1, double* __restrict__ b2,
                            double* __restrict__ b3, double* __restrict__ b4, double* __restrict__ b5) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t d = id[i];
  b0[d] = a0[i]; b1[d] = a1[i]; b2[d] = a2[i]; b3[d] = a3[i]; b4[d] = a4[i]; b5[d] = a5[i];
}

// device -> host copy of the six particle arrays (original order when ids are tracked)
int fb_download_particles(flipb200_ctx* c, double* const dst[6]) {
  Particles& pt = c->pt;
  if (pt.n == 0) return 0;
  double* const* src = pt.a;
  if (c->P.track_ids && c->world == 1) {   // slab contexts return local order; ids via flipb200_get_particle_ids
    k_unpermute<<<(unsigned)((pt.n + 255) / 256), 256, 0, c->stream>>>(pt.n, pt.id, pt.a[0], pt.a[1], pt.a[2], pt.a[3],
                                                                       pt.a[4], pt.a[5], pt.b[0], pt.b[1], pt.b[2],
                                                                       pt.b[3], pt.b[4], pt.b[5]);
    CK_LAUNCH(); c->launches++;
    src = pt.b;
  }
  for (int q = 0; q < 6; ++q)
    if (dst[q]) CK(cudaMemcpyAsync(dst[q], src[q], pt.n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  return 0;
}

// z-slab split: first owned plane of every rank, aligned to 2^a so that `a` multigrid levels
// coarsen without crossing a slab boundary (mg.cu); every slab keeps >= 2 planes on level a.
static int slab_split(int nz, int world, int* zs, int* align_log2) {
  if (world == 1) { zs[0] = 0; zs[1] = nz; *align_log2 = 0; return 0; }
  const int base = nz / world;
  if (base < 4) return -1;
  int a = 0;
  while (a < 6 && (4 << a) <= base) ++a;     // 2^(a+1) <= base
  const int A = 1 << a;
  for (int r = 0; r < world; ++r) zs[r] = (int)((int64_t)r * nz / world) & ~(A - 1);
  zs[world] = nz;
  for (int r = 0; r < world; ++r) if (zs[r + 1] - zs[r] < 2 * A) return -1;
  *align_log2 = a;
  return 0;
}

// caller-given slab boundaries (flipb200_create_slab_split): zs[0] = 0 < zs[1] < ... < zs[world] = nz; the alignment is the
// largest 2^a (a <= 5) that divides every interior boundary and leaves every slab >= 2 * 2^a planes; a >= 1 is required
static int slab_split_given(int nz, int world, const int32_t* in, int* zs, int* align_log2) {
  if (in[0] != 0 || in[world] != nz) return -1;
  for (int r = 0; r < world; ++r) if (in[r + 1] <= in[r]) return -1;
  int a = 5;
  for (int r = 1; r < world; ++r) { int t = 0; while (t < 5 && !((in[r] >> t) & 1)) ++t; if (t < a) a = t; }
  for (;;) {
    bool ok = true;
    for (int r = 0; r < world; ++r) if (in[r + 1] - in[r] < (2 << a)) ok = false;
    if (ok || a == 0) break;
    --a;
  }
  if (a < 1) return -1;
  for (int r = 0; r < world; ++r) if (in[r + 1] - in[r] < (2 << a) || in[r + 1] - in[r] < 4) return -1;
  for (int r = 0; r <= world; ++r) zs[r] = in[r];
  *align_log2 = a;
  return 0;
}

int fb_create(const flipb200_params* p, const uint8_t* solid_host, int rank, int world, flipb200_ctx** out, const int32_t* zsplit_in) {
  if (!p || !out) return FLIPB200_ERR_ARG;
  *out = nullptr;
  if (p->nx < 2 * p->wall + 2 || p->ny < 2 * p->wall + 2 || p->nz < 2 * p->wall + 2) return FLIPB200_ERR_ARG;
  if (p->wall < 1 || p->relax_div < 1 || !(p->dx > 0) || !(p->rho > 0)) return FLIPB200_ERR_ARG;
  if (!(p->flip_ratio >= 0.0 && p->flip_ratio <= 1.0)) return FLIPB200_ERR_ARG;   // 1 = the reference's pure FLIP (fluid.cc:981)
  if (p->extrapolate != 0 && (p->extrapolate != 1 || world > 1)) return FLIPB200_ERR_ARG;   // single GPU only (grid.cu fb_extrapolate)
  if (p->precision != 0 && p->precision != 32 && p->precision != 64) return FLIPB200_ERR_ARG;
  flipb200_ctx* c = (flipb200_ctx*)calloc(1, sizeof(flipb200_ctx));
  if (!c) return FLIPB200_ERR_ARG;
  c->P = *p;
  if (c->P.precision == 0) c->P.precision = 64;                  // zero-filled params of an older caller
  // fp32 mode: solve to single-precision round-off unless the caller set a tolerance of their own
  if (c->P.precision == 32 && c->P.pcg_tol == 2.220446049250313e-16) c->P.pcg_tol = 1.1920928955078125e-07;
  // experiment switches are read ONCE, here; after that only c->P / c->opt decide (two contexts of one process can differ)
  if (const char* e = getenv("FLIPB200_WARM")) c->P.warm_start = atoi(e);
  if (const char* e = getenv("FLIPB200_MIXED")) c->P.mixed = atoi(e);
  c->opt_p2g = 'c'; c->opt_g2p = 't';
  { const char* e = getenv("FLIPB200_SPARSE"); c->opt_sparse = !(e && e[0] == '0'); }
  { const char* e = getenv("FLIPB200_ST4"); c->opt_st4 = !(e && e[0] == '0'); }
  { const char* e = getenv("FLIPB200_RZ64"); c->opt_rz64 = e && e[0] == '1'; }
  { const char* e = getenv("FLIPB200_NO_GRAPH"); c->opt_no_graph = e && e[0] == '1'; }
  c->opt_warm_debug = getenv("FLIPB200_WARM_DEBUG") != nullptr;
  c->opt_cheb_steps = 0; c->opt_cheb_lmin = 0.f;
  if (const char* e = getenv("FLIPB200_CHEB")) {   // "steps,lmin"
    int st = 0; float lm = 0.f;
    if (sscanf(e, "%d,%f", &st, &lm) == 2 && st > 0 && st <= 200 && lm > 0.f && lm < 1.f) { c->opt_cheb_steps = st; c->opt_cheb_lmin = lm; }
  }
  if (const char* e = getenv("FLIPB200_P2G")) if (e[0] == 'c' || e[0] == 'p' || e[0] == 't') c->opt_p2g = e[0];
  if (const char* e = getenv("FLIPB200_G2P")) if (e[0] == 'c' || e[0] == 'p' || e[0] == 't') c->opt_g2p = e[0];
  c->rank = rank; c->world = world;
  if (zsplit_in && world > 1 ? slab_split_given(p->nz, world, zsplit_in, c->zsplit, &c->slab_align)
                             : slab_split(p->nz, world, c->zsplit, &c->slab_align)) { free(c); return FLIPB200_ERR_ARG; }
  c->gz0 = c->zsplit[rank]; c->gz1 = c->zsplit[rank + 1];
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    fprintf(stderr, "flipb200: no CUDA device -- this library has no CPU fallback\n");
    free(c);
    return FLIPB200_ERR_CUDA;
  }
  int dev = p->device;
  if (dev < 0) { if (cudaGetDevice(&dev) != cudaSuccess) { free(c); return FLIPB200_ERR_CUDA; } }
  c->device = dev;
  CK(cudaSetDevice(dev));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, dev));
  c->num_sms = prop.multiProcessorCount;
  CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  for (int e = 0; e < 8; ++e) CK(cudaEventCreate(&c->ev[e]));

  Grid& g = c->g;
  const int gh = world > 1 ? SLAB_GHOST : 0;
  g.nx = p->nx; g.ny = p->ny; g.nz = (c->gz1 - c->gz0) + 2 * gh;
  g.px = (p->nx + 31) & ~31;
  g.kg0 = c->gz0 - gh; g.gnz = p->nz; g.k0 = gh; g.k1 = g.nz - gh;
  g.ox = p->ox; g.oy = p->oy; g.oz = p->oz - g.kg0;
  g.wall = p->wall;
  int64_t plane = (int64_t)g.px * g.ny, total = plane * g.nz;
  if (total >= (int64_t)2147483647 - 8) { flipb200_destroy(c); return FLIPB200_ERR_ARG; }
  g.plane = (int)plane; g.total = (int)total;
  const size_t T = (size_t)total;

  if (alloc_n(c, &c->solid, T) || alloc_n(c, &c->flags, T) || alloc_n(c, &c->flags_static, T)) goto fail;
  for (int a = 0; a < 3; ++a)
    if (alloc_n(c, &c->u[a], T) || alloc_n(c, &c->ub[a], T)) goto fail;
  if (alloc_n(c, &c->wacc, T) || alloc_n(c, &c->wf, T) || alloc_n(c, &c->bf[0], T) || alloc_n(c, &c->bf[1], T)) goto fail;
  if (alloc_n(c, &c->x, T) || alloc_n(c, &c->r, T) || alloc_n(c, &c->pd, T) || alloc_n(c, &c->Ap, T)) goto fail;
  if (c->P.warm_start && (alloc_n(c, &c->xprev, T) || alloc_n(c, &c->xfirst, T) || alloc_n(c, &c->xG, T))) goto fail;
  if (alloc_n(c, &c->cell_count, T + 8) || alloc_n(c, &c->cell_start, T + 8) || alloc_n(c, &c->scan_tmp, T / 2048 + 64)) goto fail;
  if (alloc_n(c, &c->ds, 1)) goto fail;
  if (world > 1 && alloc_n(c, &c->halo_tmp, 4 * (size_t)plane)) goto fail;
  if (fb_mg_alloc(c)) goto fail;
  if (cudaHostAlloc((void**)&c->hs, sizeof(DevScalars), cudaHostAllocDefault) != cudaSuccess) goto fail;
  memset(c->hs, 0, sizeof(DevScalars));

  {
    // solid mask with the shell forced solid (fluid.cc:1264-1266), laid out with the row pitch;
    // ghost planes carry the neighbours' values (the mask is static, every rank has the whole box)
    std::vector<uint8_t> h(T, 0);
    for (int k = 0; k < g.nz; ++k) {
      const int kk = k + g.kg0;
      if (kk < 0 || kk >= g.gnz) continue;
      for (int j = 0; j < g.ny; ++j)
        for (int i = 0; i < g.nx; ++i) {
          bool s = !inside_shell(g, i, j, k);
          if (!s && solid_host) s = solid_host[((size_t)kk * g.ny + j) * g.nx + i] != 0;
          h[gidx(g, i, j, k)] = s ? 1 : 0;
        }
    }
    if (cudaMemcpyAsync(c->solid, h.data(), T, cudaMemcpyHostToDevice, c->stream) != cudaSuccess) goto fail;
    dim3 blk(32, 8, 1), grd(cdiv(g.px, 32), cdiv(g.ny, 8), g.nz);
    k_static_flags<<<grd, blk, 0, c->stream>>>(g, c->solid, c->flags_static);
    if (cudaMemcpyAsync(c->flags, c->flags_static, T, cudaMemcpyDeviceToDevice, c->stream) != cudaSuccess) goto fail;
    if (cudaStreamSynchronize(c->stream) != cudaSuccess) goto fail;
  }
  c->box = Box{0, 0, g.k0, g.nx, g.ny, g.k1};
  if (p->max_particles > 0 && fb_reserve_particles(c, p->max_particles)) goto fail;
  *out = c;
  return FLIPB200_OK;
fail:
  fprintf(stderr, "flipb200_create failed: %s\n", cudaGetErrorString(cudaGetLastError()));
  flipb200_destroy(c);
  return FLIPB200_ERR_CUDA;
}

extern "C" {

int flipb200_create(const flipb200_params* p, const uint8_t* solid_host, flipb200_ctx** out) {
  return fb_create(p, solid_host, 0, 1, out, nullptr);
}

// One z-slab of the box: rank `rank` of `world` owns the global planes reported by
// flipb200_slab_range; params and solid_host describe the WHOLE box on every rank.  The context is
// usable after flipb200_comm_init_nccl / _local.  New design: the reference is single-process.
int flipb200_create_slab(const flipb200_params* p, const uint8_t* solid_host, int32_t rank, int32_t world,
                         flipb200_ctx** out) {
  if (world < 1 || world > 16 || rank < 0 || rank >= world) return FLIPB200_ERR_ARG;
  return fb_create(p, solid_host, rank, world, out, nullptr);
}

// The same with slab boundaries chosen by the caller (static load balancing: the fluid of a scene rarely fills the
// box evenly along z, and the pressure solve only sweeps the planes that hold fluid).  zsplit: world + 1 global
// plane indices, 0 = zsplit[0] < ... < zsplit[world] = nz, interior boundaries multiples of 2^a with a >= 1 and every
// slab >= 2 * 2^a planes (a decides how many multigrid levels stay distributed).  The same array on every rank.
int flipb200_create_slab_split(const flipb200_params* p, const uint8_t* solid_host, int32_t rank, int32_t world,
                               const int32_t* zsplit, flipb200_ctx** out) {
  if (world < 1 || world > 16 || rank < 0 || rank >= world || !zsplit) return FLIPB200_ERR_ARG;
  return fb_create(p, solid_host, rank, world, out, zsplit);
}

int flipb200_slab_range(flipb200_ctx* c, int32_t* z0, int32_t* z1) {
  if (!c) return FLIPB200_ERR_ARG;
  if (z0) *z0 = c->gz0;
  if (z1) *z1 = c->gz1;
  return FLIPB200_OK;
}

void flipb200_destroy(flipb200_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  fb_graph_invalidate(c);
  fb_mg_free(c);
  fb_comm_free(c);
  cudaFree(c->halo_tmp); cudaFree(c->halo.flags); cudaFree(c->halo.mail);
  cudaFree(c->solid); cudaFree(c->flags); cudaFree(c->flags_static); cudaFree(c->ext_layer);
  for (int a = 0; a < 3; ++a) { cudaFree(c->u[a]); cudaFree(c->ub[a]); }
  cudaFree(c->wacc); cudaFree(c->wf); cudaFree(c->bf[0]); cudaFree(c->bf[1]);
  cudaFree(c->x); cudaFree(c->r); cudaFree(c->pd); cudaFree(c->Ap); cudaFree(c->xprev); cudaFree(c->xfirst); cudaFree(c->xG);
  cudaFree(c->cell_count); cudaFree(c->cell_start); cudaFree(c->scan_tmp);
  for (int q = 0; q < 6; ++q) { cudaFree(c->pt.a[q]); cudaFree(c->pt.b[q]); }
  cudaFree(c->pt.key); cudaFree(c->pt.rank); cudaFree(c->pt.id); cudaFree(c->pt.id_b);
  cudaFree(c->ds);
  if (c->hs) cudaFreeHost(c->hs);
  for (int e = 0; e < 8; ++e) if (c->ev[e]) cudaEventDestroy(c->ev[e]);
  if (c->stream) cudaStreamDestroy(c->stream);
  free(c);
}

int flipb200_set_particles(flipb200_ctx* c, const double* px, const double* py, const double* pz,
                           const double* vx, const double* vy, const double* vz, int64_t n) {
  if (!c || n < 0 || (n > 0 && (!px || !py || !pz || !vx || !vy || !vz))) return FLIPB200_ERR_ARG;
  if (n >= (int64_t)2147483647) return FLIPB200_ERR_ARG;
  CK(cudaSetDevice(c->device));
  c->pt.n = 0;
  if (fb_reserve_particles(c, n)) return FLIPB200_ERR_CUDA;
  const double* src[6] = {px, py, pz, vx, vy, vz};
  for (int q = 0; q < 6; ++q)
    if (n) CK(cudaMemcpyAsync(c->pt.a[q], src[q], n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  c->pt.n = n;
  if (fb_init_ids(c, 0, n, 0)) return FLIPB200_ERR_CUDA;
  CK(cudaStreamSynchronize(c->stream));
  c->sorted = false;
  c->sparse_ok = false;            // new particles: the next step sweeps the whole grid (grid.cu fb_active_region)
 
  if (c->world > 1) {
    // every rank is handed the whole scene and keeps the particles whose base plane it owns
    int rc = fb_sort_particles_ex(c, 1);
    if (rc) return rc;
    CK(cudaStreamSynchronize(c->stream));
  }
  return FLIPB200_OK;
}

/* ids of the particles in the order flipb200_get_particles returns them on a slab context (index
 * into the arrays given to flipb200_set_particles; needs params.track_ids) */
int flipb200_get_particle_ids(flipb200_ctx* c, uint32_t* ids, int64_t cap) {
  if (!c || !ids || cap < c->pt.n) return FLIPB200_ERR_ARG;
  if (!c->P.track_ids) return FLIPB200_ERR_STATE;
  CK(cudaSetDevice(c->device));
  if (c->pt.n) CK(cudaMemcpyAsync(ids, c->pt.id, c->pt.n * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return FLIPB200_OK;
}

int flipb200_get_particles(flipb200_ctx* c, double* px, double* py, double* pz, double* vx, double* vy,
                           double* vz, int64_t cap, int64_t* n) {
  if (!c) return FLIPB200_ERR_ARG;
  if (n) *n = c->pt.n;
  if (cap < c->pt.n) return FLIPB200_ERR_ARG;
  CK(cudaSetDevice(c->device));
  double* dst[6] = {px, py, pz, vx, vy, vz};
  if (fb_download_particles(c, dst)) return FLIPB200_ERR_CUDA;
  CK(cudaStreamSynchronize(c->stream));
  return FLIPB200_OK;
}

int flipb200_num_particles(flipb200_ctx* c, int64_t* n) {
  if (!c || !n) return FLIPB200_ERR_ARG;
  *n = c->pt.n;
  return FLIPB200_OK;
}

static int grid_ptr(flipb200_ctx* c, int which, void** p, size_t* esz) {
  switch (which) {
    case FLIPB200_GRID_U: case FLIPB200_GRID_V: case FLIPB200_GRID_W: *p = c->u[which]; *esz = 8; return 0;
    case FLIPB200_GRID_UB: case FLIPB200_GRID_VB: case FLIPB200_GRID_WB: *p = c->ub[which - 3]; *esz = 8; return 0;
    case FLIPB200_GRID_P: *p = c->x; *esz = 8; return 0;
    case FLIPB200_GRID_WEIGHT: *p = c->wf; *esz = 4; return 0;
    case FLIPB200_GRID_B: *p = c->bf[c->bcur]; *esz = 4; return 0;
    case FLIPB200_GRID_FLAGS: *p = c->flags; *esz = 1; return 0;
    default: return -1;
  }
}

int flipb200_get_grid(flipb200_ctx* c, int which, void* host_out, size_t bytes) {
  if (!c || !host_out) return FLIPB200_ERR_ARG;
  void* d; size_t esz;
  if (grid_ptr(c, which, &d, &esz)) return FLIPB200_ERR_ARG;
  const Grid& g = c->g;   // a slab returns its owned planes [gz0, gz1)
  if (bytes != (size_t)g.nx * g.ny * (g.k1 - g.k0) * esz) return FLIPB200_ERR_ARG;
  CK(cudaSetDevice(c->device));
  CK(cudaMemcpy2DAsync(host_out, g.nx * esz, (char*)d + (size_t)g.k0 * g.plane * esz, g.px * esz, g.nx * esz,
                       (size_t)g.ny * (g.k1 - g.k0), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return FLIPB200_OK;
}

int flipb200_set_grid(flipb200_ctx* c, int which, const void* host_in, size_t bytes) {
  if (!c || !host_in) return FLIPB200_ERR_ARG;
  void* d; size_t esz;
  if (grid_ptr(c, which, &d, &esz)) return FLIPB200_ERR_ARG;
  const Grid& g = c->g;
  if (bytes != (size_t)g.nx * g.ny * (g.k1 - g.k0) * esz) return FLIPB200_ERR_ARG;
  CK(cudaSetDevice(c->device));
  CK(cudaMemcpy2DAsync((char*)d + (size_t)g.k0 * g.plane * esz, g.px * esz, host_in, g.nx * esz, g.nx * esz,
                       (size_t)g.ny * (g.k1 - g.k0), cudaMemcpyHostToDevice, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  c->sparse_ok = false;            // a grid replaced from outside: nothing is known about where it is empty
  return FLIPB200_OK;
}

int flipb200_host_alloc(void** ptr, size_t bytes) {
  if (!ptr) return FLIPB200_ERR_ARG;
  return cudaHostAlloc(ptr, bytes, cudaHostAllocDefault) == cudaSuccess ? FLIPB200_OK : FLIPB200_ERR_CUDA;
}
int flipb200_host_free(void* ptr) {
  return cudaFreeHost(ptr) == cudaSuccess ? FLIPB200_OK : FLIPB200_ERR_CUDA;
}

}  // extern "C"

int fb_fetch_scalars(flipb200_ctx* c) {
  CK(cudaMemcpyAsync(c->hs, c->ds, sizeof(DevScalars), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}
