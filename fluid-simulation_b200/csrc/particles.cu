// particles.cu -- particle side of the FLIP substep: cell sort, P2G splat, G2P FLIP update,
// advection with solid collision, device seeding.
//   reference: PointList::P2Gtransfer fluid.cc:1106-1148, p2gCatmullRom :265-299,
//              PointList::interpolate :843-882, PointList::FLIPadvect :972-1038, CatmullRomFLIP :210-263
#include "common.cuh"
#include <stdlib.h>

// ---------------------------------------------------------------------------------------------
// cell sort: key = linear index of the base cell of the particle's 2x2x2 support; particles
// whose support leaves the box (they touch no writable node because the shell is solid) go to
// the trailing bucket `total`.  Counting sort: histogram (atomic, returns the in-cell rank),
// exclusive scan, scatter.
// ---------------------------------------------------------------------------------------------
// z-slabs: a particle belongs to the rank that owns its base plane; base planes below / above the
// owned range go to the buckets total+1 (lower neighbour) / total+2 (upper neighbour).
__device__ __forceinline__ uint32_t particle_key(const Grid& g, double x, double y, double z) {
  // guard against NaN / huge coordinates before the int conversion
  if (!(fabs(x) < 1.0e9) || !(fabs(y) < 1.0e9) || !(fabs(z) < 1.0e9)) return (uint32_t)g.total;
  int bi = base_cell(x, g.ox), bj = base_cell(y, g.oy), bk = base_cell(z, g.oz);
  if (bi < 0 || bj < 0 || bk + g.kg0 < 0 || bi > g.nx - 2 || bj > g.ny - 2 || bk + g.kg0 > g.gnz - 2) return (uint32_t)g.total;
  if (bk < g.k0) return (uint32_t)g.total + 1u;
  if (bk >= g.k1) return (uint32_t)g.total + 2u;
  return (uint32_t)gidx(g, bi, bj, bk);
}

__global__ void k_key_count(Grid g, int64_t n, const double* __restrict__ px, const double* __restrict__ py,
                            const double* __restrict__ pz, uint32_t* __restrict__ key,
                            uint32_t* __restrict__ rank, int32_t* __restrict__ count, int drop_outside) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t kx = particle_key(g, px[i], py[i], pz[i]);
  if (drop_outside && kx == (uint32_t)g.total) kx = (uint32_t)g.total + 1u;   // slab filter: rank 0 keeps the strays
  key[i] = kx;
  rank[i] = (uint32_t)atomicAdd(&count[kx], 1);
}

// exclusive scan of int32[n]: 2048 items per block (1024 threads x 2)
#define SCAN_ITEMS 2048
__global__ void k_scan_block(const int32_t* __restrict__ in, int32_t* __restrict__ out,
                             int32_t* __restrict__ block_sum, int n) {
  __shared__ int32_t wsum[32];
  int base = blockIdx.x * SCAN_ITEMS;
  int t = threadIdx.x;
  int i0 = base + 2 * t, i1 = i0 + 1;
  int a = i0 < n ? in[i0] : 0, b = i1 < n ? in[i1] : 0;
  int s = a + b;
  int lane = t & 31, w = t >> 5;
  int v = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += y; }
  if (lane == 31) wsum[w] = v;
  __syncthreads();
  if (w == 0) {
    int x = wsum[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    wsum[lane] = x;
  }
  __syncthreads();
  int excl = v - s + (w > 0 ? wsum[w - 1] : 0);
  if (i0 < n) out[i0] = excl;
  if (i1 < n) out[i1] = excl + a;
  if (t == 1023) block_sum[blockIdx.x] = wsum[31];
}
__global__ void k_scan_sums(int32_t* __restrict__ sums, int nb) {
  // single block: sequential over chunks of 1024
  __shared__ int32_t wsum[32];
  __shared__ int32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  int t = threadIdx.x, lane = t & 31, w = t >> 5;
  for (int base = 0; base < nb; base += 1024) {
    int i = base + t;
    int s = i < nb ? sums[i] : 0;
    int v = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += y; }
    if (lane == 31) wsum[w] = v;
    __syncthreads();
    if (w == 0) {
      int x = wsum[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
      wsum[lane] = x;
    }
    __syncthreads();
    int excl = v - s + (w > 0 ? wsum[w - 1] : 0) + carry;
    if (i < nb) sums[i] = excl;
    __syncthreads();
    if (t == 1023) carry = excl + s;
    __syncthreads();
  }
}
__global__ void k_scan_add(int32_t* __restrict__ out, const int32_t* __restrict__ sums, int n) {
  int i = blockIdx.x * SCAN_ITEMS + threadIdx.x;
  int add = sums[blockIdx.x];
  if (i < n) out[i] += add;
  if (i + 1024 < n) out[i + 1024] += add;
}

__global__ void k_reorder(int64_t n, const uint32_t* __restrict__ key, const uint32_t* __restrict__ rank,
                          const int32_t* __restrict__ start,
                          const double* __restrict__ a0, const double* __restrict__ a1, const double* __restrict__ a2,
                          const double* __restrict__ a3, const double* __restrict__ a4, const double* __restrict__ a5,
                          double* __restrict__ b0, double* __restrict__ b1, double* __restrict__ b2,
                          double* __restrict__ b3, double* __restrict__ b4, double* __restrict__ b5,
                          const uint32_t* __restrict__ id, uint32_t* __restrict__ id_b) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int64_t d = (int64_t)start[key[i]] + rank[i];
  b0[d] = a0[i]; b1[d] = a1[i]; b2[d] = a2[i];
  b3[d] = a3[i]; b4[d] = a4[i]; b5[d] = a5[i];
  if (id) id_b[d] = id[i];
}

// filter != 0 (slab contexts, after set_particles with the whole scene): keep what this rank owns,
// drop the rest without any exchange.
int fb_sort_particles_ex(flipb200_ctx* c, int filter) {
  const Grid& g = c->g;
  Particles& pt = c->pt;
  const int nb = g.total + 3;                 // cells + "outside the box" + "to lower rank" + "to upper rank"
  CK(cudaMemsetAsync(c->cell_count, 0, (size_t)(nb + 1) * sizeof(int32_t), c->stream));
  if (pt.n > 0) {
    int blocks = (int)((pt.n + 255) / 256);
    k_key_count<<<blocks, 256, 0, c->stream>>>(g, pt.n, pt.a[0], pt.a[1], pt.a[2], pt.key, pt.rank, c->cell_count,
                                               (filter && c->rank != 0) ? 1 : 0);
    CK_LAUNCH(); c->launches++;
  }
  // scan nb+1 entries so that cell_start[nb] == n
  int ns = nb + 1;
  int sb = cdiv(ns, SCAN_ITEMS);
  k_scan_block<<<sb, 1024, 0, c->stream>>>(c->cell_count, c->cell_start, c->scan_tmp, ns);
  k_scan_sums<<<1, 1024, 0, c->stream>>>(c->scan_tmp, sb);
  k_scan_add<<<sb, 1024, 0, c->stream>>>(c->cell_start, c->scan_tmp, ns);
  CK_LAUNCH(); c->launches += 3;
  if (pt.n > 0) {
    int blocks = (int)((pt.n + 255) / 256);
    k_reorder<<<blocks, 256, 0, c->stream>>>(pt.n, pt.key, pt.rank, c->cell_start, pt.a[0], pt.a[1], pt.a[2],
                                             pt.a[3], pt.a[4], pt.a[5], pt.b[0], pt.b[1], pt.b[2], pt.b[3],
                                             pt.b[4], pt.b[5], pt.id, pt.id_b);
    CK_LAUNCH(); c->launches++;
    for (int q = 0; q < 6; ++q) { double* t = pt.a[q]; pt.a[q] = pt.b[q]; pt.b[q] = t; }
    { uint32_t* t = pt.id; pt.id = pt.id_b; pt.id_b = t; }
  }
  c->n_sorted = pt.n;
  c->sorted = true;
  if (c->world == 1) return 0;
  // ---- migration: the two trailing buckets leave, the neighbours' arrive as an
  // unsorted tail [n_sorted, n) that this step's P2G / G2P handle particle by particle.
  int32_t s3[3];
  CK(cudaMemcpyAsync(s3, c->cell_start + g.total + 1, sizeof(s3), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  const int64_t n_keep = s3[0];
  const int64_t send_n[2] = {s3[1] - s3[0], s3[2] - s3[1]};
  c->n_sorted = n_keep;
  if (filter) { pt.n = n_keep; return 0; }
  int64_t recv_n[2];
  int rc = fb_comm_exchange_counts(c, send_n, recv_n);
  if (rc) return rc;
  const int64_t nnew = n_keep + recv_n[0] + recv_n[1];
  if (nnew >= (int64_t)2147483647) return FLIPB200_ERR_ARG;
  if ((rc = fb_reserve_particles(c, nnew > pt.n ? nnew : pt.n))) return rc;
  void* src[8]; void* dst[8]; size_t elem[8]; int na = 0;
  for (int q = 0; q < 6; ++q) { src[na] = pt.a[q]; dst[na] = pt.b[q]; elem[na] = sizeof(double); ++na; }
  if (c->P.track_ids) { src[na] = pt.id; dst[na] = pt.id_b; elem[na] = sizeof(uint32_t); ++na; }
  const int64_t send_off[2] = {n_keep, n_keep + send_n[0]};
  const int64_t recv_off[2] = {0, recv_n[0]};
  if ((rc = fb_comm_exchange_particles(c, na, src, dst, elem, send_off, send_n, recv_off, recv_n))) return rc;
  const int64_t nr = recv_n[0] + recv_n[1];
  if (nr > 0)
    for (int a = 0; a < na; ++a)
      CK(cudaMemcpyAsync((char*)src[a] + n_keep * elem[a], dst[a], nr * elem[a], cudaMemcpyDeviceToDevice, c->stream));
  pt.n = nnew;
  return 0;
}
int fb_sort_particles(flipb200_ctx* c) { return fb_sort_particles_ex(c, 0); }

// ---------------------------------------------------------------------------------------------
// P2G.  One thread per base cell: every particle of the cell touches the same 2x2x2 nodes, so
// the 8 x (w, w*vx, w*vy, w*vz) partial sums stay in registers and are flushed once per cell with
// native fp64 RED atomics (REDG.E.ADD.F64) -- 32 atomics per cell instead of 32 per particle.
// Node predicate of the reference (fluid.cc:288): !solid && inside the shell; the shell is always
// solid here, so the flag byte alone decides.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
k_p2g_cell(Grid g, Box box, const int32_t* __restrict__ start, const uint8_t* __restrict__ fl,
           const double* __restrict__ px, const double* __restrict__ py, const double* __restrict__ pz,
           const double* __restrict__ vx, const double* __restrict__ vy, const double* __restrict__ vz,
           double* __restrict__ wacc, double* __restrict__ ux, double* __restrict__ uy, double* __restrict__ uz) {
  int i = box.x0 + blockIdx.x * blockDim.x + threadIdx.x;
  int j = box.y0 + blockIdx.y * blockDim.y + threadIdx.y;
  int k = box.z0 + blockIdx.z;
  if (i >= box.x1 || j >= box.y1 || i > g.nx - 2 || j > g.ny - 2 || k + g.kg0 > g.gnz - 2) return;
  int q = gidx(g, i, j, k);
  int s = start[q], e = start[q + 1];
  if (s == e) return;
  double aw[8], ax[8], ay[8], az[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) { aw[m] = 0; ax[m] = 0; ay[m] = 0; az[m] = 0; }
  const double nx0 = (double)(i - g.ox), ny0 = (double)(j - g.oy), nz0 = (double)(k - g.oz);
  // (measured and dropped in round 2, profiles/r2_particles.md: issuing the loads of particle n + 1 before the arithmetic
  //  of particle n, and handing the x+1 partial sums to the next lane so that a cell sends 16 REDs instead of 32 --
  //  2.10 and 2.24 ms against 2.06 ms: the kernel is bound by neither the loads in flight nor the RED rate)
  for (int n = s; n < e; ++n) {
    double x = px[n], y = py[n], z = pz[n];
    double a = vx[n], b = vy[n], cc = vz[n];
    double wx0, wx1, wy0, wy1, wz0, wz1;
    spline_pair(x, nx0, wx0, wx1); spline_pair(y, ny0, wy0, wy1); spline_pair(z, nz0, wz0, wz1);
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      double cw = ((m & 1) ? wx1 : wx0) * ((m & 2) ? wy1 : wy0) * ((m & 4) ? wz1 : wz0);   // (sx*sy)*sz, fluid.cc:291
      aw[m] += cw; ax[m] += cw * a; ay[m] += cw * b; az[m] += cw * cc;
    }
  }
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    int qq = q + (m & 1) + ((m & 2) ? g.px : 0) + ((m & 4) ? g.plane : 0);
    if (fl[qq] & FB_SOLID) continue;            // fluid.cc:288
    atomicAdd(&wacc[qq], aw[m]);
    atomicAdd(&ux[qq], ax[m]);
    atomicAdd(&uy[qq], ay[m]);
    atomicAdd(&uz[qq], az[m]);
  }
}

// Reference-shaped variant: one thread per particle, 32 atomics each.  Kept as the cross-check of
// the cell kernel (FLIPB200_P2G=particle) and as the path for unsorted input.
__global__ void k_p2g_particle(Grid g, int64_t n, const uint8_t* __restrict__ fl,
                               const double* __restrict__ px, const double* __restrict__ py, const double* __restrict__ pz,
                               const double* __restrict__ vx, const double* __restrict__ vy, const double* __restrict__ vz,
                               double* __restrict__ wacc, double* __restrict__ ux, double* __restrict__ uy, double* __restrict__ uz) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  double x = px[t], y = py[t], z = pz[t];
  uint32_t key = particle_key(g, x, y, z);
  if (key >= (uint32_t)g.total) return;
  int q = (int)key;
  int k = q / g.plane, rem = q - k * g.plane, j = rem / g.px, i = rem - j * g.px;
  double a = vx[t], b = vy[t], cc = vz[t];
  const double nx0 = (double)(i - g.ox), ny0 = (double)(j - g.oy), nz0 = (double)(k - g.oz);
  double wx0, wx1, wy0, wy1, wz0, wz1;
  spline_pair(x, nx0, wx0, wx1); spline_pair(y, ny0, wy0, wy1); spline_pair(z, nz0, wz0, wz1);
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    int qq = q + (m & 1) + ((m & 2) ? g.px : 0) + ((m & 4) ? g.plane : 0);
    if (fl[qq] & FB_SOLID) continue;
    double cw = ((m & 1) ? wx1 : wx0) * ((m & 2) ? wy1 : wy0) * ((m & 4) ? wz1 : wz0);
    atomicAdd(&wacc[qq], cw);
    atomicAdd(&ux[qq], cw * a);
    atomicAdd(&uy[qq], cw * b);
    atomicAdd(&uz[qq], cw * cc);
  }
}

int fb_p2g_rows(flipb200_ctx* c);                          // rows.cu
// bounding box (in base cells) of the particles, from the sorted cell histogram is expensive;
// P2G sweeps every base cell of the box instead -- a cell with no particles costs two int loads.
// c->opt_p2g (FLIPB200_P2G, read once at create; the parity tests cross-check the variants):
//   'c' one thread per cell (default: 2.06 ms at 256^3 / 127 M particles), 't' rows staged by TMA (rows.cu, 2.16 ms),
//   'p' one thread per particle with 32 REDs each (reference-shaped; also the path of unsorted input)
int fb_p2g(flipb200_ctx* c) {
  const Grid& g = c->g;
  // vels.fill(0) (fluid.cc:1378), weights.fill(0) (:1108) -- on the planes of the active region (grid.cu fb_active_region:
  // everything outside is zero already) plus the plane above it that the 2x2x2 splat reaches, and, on z-slabs, the ghost
  // planes: plane k1 is shipped whole to the upper neighbour (fb_halo_sum_up) whatever the region is
  bool full = true;
  const Box AR = fb_active_region(c, &full);
  double* acc4[4] = {c->wacc, c->u[0], c->u[1], c->u[2]};
  {
    int z0 = full ? 0 : AR.z0, z1 = full ? g.nz : min(AR.z1 + 1, g.nz);
    if (!full && c->world > 1) {
      // ghost ranges [0, k0) and [k1, nz) always; merged with the region's planes when they touch
      if (g.k0 > 0) { if (z1 > z0 && z0 <= g.k0) z0 = 0; else for (int a = 0; a < 4; ++a) CK(cudaMemsetAsync(acc4[a], 0, (size_t)g.k0 * g.plane * sizeof(double), c->stream)); }
      if (g.k1 < g.nz) { if (z1 > z0 && z1 >= g.k1) z1 = g.nz; else for (int a = 0; a < 4; ++a) CK(cudaMemsetAsync(acc4[a] + (size_t)g.k1 * g.plane, 0, (size_t)(g.nz - g.k1) * g.plane * sizeof(double), c->stream)); }
    }
    if (z1 > z0)
      for (int a = 0; a < 4; ++a) CK(cudaMemsetAsync(acc4[a] + (size_t)z0 * g.plane, 0, (size_t)(z1 - z0) * g.plane * sizeof(double), c->stream));
  }
  Particles& pt = c->pt;
  const char mode = c->opt_p2g;
  if (mode == 'p' && c->world == 1) {
    if (pt.n == 0) return 0;
    int blocks = (int)((pt.n + 127) / 128);
    k_p2g_particle<<<blocks, 128, 0, c->stream>>>(g, pt.n, c->flags_static, pt.a[0], pt.a[1], pt.a[2], pt.a[3],
                                                  pt.a[4], pt.a[5], c->wacc, c->u[0], c->u[1], c->u[2]);
    CK_LAUNCH(); c->launches++;
    return 0;
  }
  if (!c->sorted) { int rc = fb_sort_particles(c); if (rc) return rc; }
  if (c->n_sorted > 0) {
    if (mode == 't') {
      int rc = fb_p2g_rows(c);
      if (rc) return rc;
    } else {
      Box b{AR.x0, AR.y0, AR.z0, min(AR.x1, g.nx - 1), min(AR.y1, g.ny - 1), AR.z1};   // base cells of the owned planes of the active region
      dim3 blk(32, 4, 1), grd(cdiv(b.x1 - b.x0, 32), cdiv(b.y1 - b.y0, 4), b.z1 - b.z0);
      if (b.z1 > b.z0 && b.x1 > b.x0 && b.y1 > b.y0)
      k_p2g_cell<<<grd, blk, 0, c->stream>>>(g, b, c->cell_start, c->flags_static, pt.a[0], pt.a[1], pt.a[2], pt.a[3],
                                             pt.a[4], pt.a[5], c->wacc, c->u[0], c->u[1], c->u[2]);
      CK_LAUNCH(); c->launches++;
    }
  }
  if (pt.n > c->n_sorted) {   // particles that migrated in this step (unsorted tail)
    const int64_t o = c->n_sorted, m = pt.n - o;
    k_p2g_particle<<<(unsigned)((m + 127) / 128), 128, 0, c->stream>>>(g, m, c->flags_static, pt.a[0] + o, pt.a[1] + o,
                                                                      pt.a[2] + o, pt.a[3] + o, pt.a[4] + o, pt.a[5] + o,
                                                                      c->wacc, c->u[0], c->u[1], c->u[2]);
    CK_LAUNCH(); c->launches++;
  }
  // nodes of plane k1 belong to the upper neighbour: add my contributions there
  double* arr[4] = {c->wacc, c->u[0], c->u[1], c->u[2]};
  return fb_halo_sum_up(c, arr);
}

// ---------------------------------------------------------------------------------------------
// G2P (FLIP) + advection.  `dc` holds the cell-centred FLIP delta
//   dc_a[c] = getVelocity(c, vels)_a - getVelocity(c, velBeforeUpdate)_a      (fluid.cc:59-70,239-240,252)
// precomputed once per step for every node inside the shell (zero elsewhere, fluid.cc:237).
// Phase 1 (fluid.cc:978-991): v_p += sum(cw*dc)/sum(cw), maxSpeed.  Phase 2 (:1000-1036) needs the
// global max => second kernel.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_g2p(Grid g, int64_t n, const uint8_t* __restrict__ fl,
      const double* __restrict__ px, const double* __restrict__ py, const double* __restrict__ pz,
      double* __restrict__ vx, double* __restrict__ vy, double* __restrict__ vz,
      const double* __restrict__ dcx, const double* __restrict__ dcy, const double* __restrict__ dcz,
      DevScalars* __restrict__ ds, double ratio) {
  __shared__ double red[32];
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double speed = 0.0;
  if (t < n) {
    double x = px[t], y = py[t], z = pz[t];
    double a = vx[t], b = vy[t], cc = vz[t];
    uint32_t key = particle_key(g, x, y, z);
    if (key < (uint32_t)g.total) {
      int q = (int)key;
      int k = q / g.plane, rem = q - k * g.plane, j = rem / g.px, i = rem - j * g.px;
      const double nx0 = (double)(i - g.ox), ny0 = (double)(j - g.oy), nz0 = (double)(k - g.oz);
      double wx0, wx1, wy0, wy1, wz0, wz1;
      spline_pair(x, nx0, wx0, wx1); spline_pair(y, ny0, wy0, wy1); spline_pair(z, nz0, wz0, wz1);
      double wsum = 0, dx = 0, dy = 0, dz = 0;
#pragma unroll
      for (int m = 0; m < 8; ++m) {
        int ii = i + (m & 1), jj = j + ((m >> 1) & 1), kk = k + ((m >> 2) & 1);
        if (!inside_shell(g, ii, jj, kk)) continue;          // fluid.cc:237 (no solid test)
        int qq = q + (m & 1) + ((m & 2) ? g.px : 0) + ((m & 4) ? g.plane : 0);
        double cw = ((m & 1) ? wx1 : wx0) * ((m & 2) ? wy1 : wy0) * ((m & 4) ? wz1 : wz0);
        wsum += cw;
        dx += dcx[qq] * cw; dy += dcy[qq] * cw; dz += dcz[qq] * cw;
      }
      if (ratio != 1.0) { a *= ratio; b *= ratio; cc *= ratio; }            // PIC/FLIP blend (see k_cell_delta)
      if (wsum != 0.0) { a += dx / wsum; b += dy / wsum; cc += dz / wsum; }   // fluid.cc:258-262, 981
      vx[t] = a; vy[t] = b; vz[t] = cc;
    } else if (ratio != 1.0) {
      a *= ratio; b *= ratio; cc *= ratio;
      vx[t] = a; vy[t] = b; vz[t] = cc;
    }
    speed = sqrt(a * a + b * b + cc * cc);                  // Vec3::length, fluid.cc:982
  }
  speed = warp_max(speed);
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) red[w] = speed;
  __syncthreads();
  if (w == 0) {
    double s = lane < (blockDim.x >> 5) ? red[lane] : 0.0;
    s = warp_max(s);
    if (lane == 0 && s > 0.0) atomicMax(&ds->max_speed_bits, (unsigned long long)__double_as_longlong(s));
  }
}

// Cell-ordered G2P: one thread per base cell.  The 8 x 3 cell-centred deltas (and which of the 8
// nodes lie inside the shell) are loaded once per cell and reused by all of its particles, so the
// gather costs 24 loads per cell instead of 24 per particle.  Requires cell-sorted particles.
__global__ void __launch_bounds__(128)
k_g2p_cell(Grid g, const int32_t* __restrict__ start,
           const double* __restrict__ px, const double* __restrict__ py, const double* __restrict__ pz,
           double* __restrict__ vx, double* __restrict__ vy, double* __restrict__ vz,
           const double* __restrict__ dcx, const double* __restrict__ dcy, const double* __restrict__ dcz,
           DevScalars* __restrict__ ds, double ratio) {
  __shared__ double red[4];
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int j = blockIdx.y * blockDim.y + threadIdx.y;
  int k = g.k0 + blockIdx.z;
  double speed = 0.0;
  int s = 0, e = 0, q = 0;
  if (i <= g.nx - 2 && j <= g.ny - 2 && k + g.kg0 <= g.gnz - 2) { q = gidx(g, i, j, k); s = start[q]; e = start[q + 1]; }
  if (s < e) {
    double ddx[8], ddy[8], ddz[8];
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      int ii = i + (m & 1), jj = j + ((m >> 1) & 1), kk = k + ((m >> 2) & 1);
      int qq = q + (m & 1) + ((m & 2) ? g.px : 0) + ((m & 4) ? g.plane : 0);
      bool in = inside_shell(g, ii, jj, kk);                     // fluid.cc:237 (no solid test)
      ddx[m] = in ? dcx[qq] : 0.0; ddy[m] = in ? dcy[qq] : 0.0; ddz[m] = in ? dcz[qq] : 0.0;
    }
    unsigned inmask = 0;
#pragma unroll
    for (int m = 0; m < 8; ++m)
      if (inside_shell(g, i + (m & 1), j + ((m >> 1) & 1), k + ((m >> 2) & 1))) inmask |= 1u << m;
    const double nx0 = (double)(i - g.ox), ny0 = (double)(j - g.oy), nz0 = (double)(k - g.oz);
    for (int n = s; n < e; ++n) {
      double x = px[n], y = py[n], z = pz[n];
      double a = vx[n], b = vy[n], cc = vz[n];
      double wx0, wx1, wy0, wy1, wz0, wz1;
      spline_pair(x, nx0, wx0, wx1); spline_pair(y, ny0, wy0, wy1); spline_pair(z, nz0, wz0, wz1);
      double wsum = 0, dx = 0, dy = 0, dz = 0;
#pragma unroll
      for (int m = 0; m < 8; ++m) {
        double cw = ((m & 1) ? wx1 : wx0) * ((m & 2) ? wy1 : wy0) * ((m & 4) ? wz1 : wz0);
        if (inmask & (1u << m)) wsum += cw;
        dx += ddx[m] * cw; dy += ddy[m] * cw; dz += ddz[m] * cw;
      }
      if (ratio != 1.0) { a *= ratio; b *= ratio; cc *= ratio; }          // PIC/FLIP blend (see k_cell_delta)
      // fluid.cc:258-262, 981.  One reciprocal instead of three fp64 divisions (<= 1 ulp apart; the
      // divisions were a quarter of the kernel's fp64 instructions).
      if (wsum != 0.0) { const double iw = 1.0 / wsum; a += dx * iw; b += dy * iw; cc += dz * iw; }
      if (wsum != 0.0 || ratio != 1.0) { vx[n] = a; vy[n] = b; vz[n] = cc; }
      speed = fmax(speed, a * a + b * b + cc * cc);            // squared: sqrt is monotone, taken once below
    }
  }
  speed = sqrt(warp_max(speed));                               // fluid.cc:982 (Vec3::length)
  int tid = threadIdx.x + blockDim.x * threadIdx.y, lane = tid & 31, w = tid >> 5;
  if (lane == 0) red[w] = speed;
  __syncthreads();
  if (tid == 0) {
    double m = fmax(fmax(red[0], red[1]), fmax(red[2], red[3]));
    if (m > 0.0) atomicMax(&ds->max_speed_bits, (unsigned long long)__double_as_longlong(m));
  }
}

// particles of the trailing "outside" bucket: no grid contact, but their speed counts (fluid.cc:982)
__global__ void k_speed_tail(const int32_t* __restrict__ start, int total, int64_t n /* = n_sorted */, double* __restrict__ vx,
                             double* __restrict__ vy, double* __restrict__ vz, DevScalars* __restrict__ ds, double ratio) {
  double speed = 0.0;
  for (int64_t t = (int64_t)start[total] + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n;
       t += (int64_t)gridDim.x * blockDim.x) {
    double a = vx[t], b = vy[t], cc = vz[t];
    if (ratio != 1.0) {   // no node in reach: weight sum 0 => the PIC part is 0 (clampedCatmullRom :185)
      a *= ratio; b *= ratio; cc *= ratio;
      vx[t] = a; vy[t] = b; vz[t] = cc;
    }
    speed = fmax(speed, sqrt(a * a + b * b + cc * cc));
  }
  speed = warp_max(speed);
  if ((threadIdx.x & 31) == 0 && speed > 0.0) atomicMax(&ds->max_speed_bits, (unsigned long long)__double_as_longlong(speed));
}

__device__ __forceinline__ bool solid_at(const Grid& g, const uint8_t* __restrict__ fl, int i, int j, int k) {
  return in_box(g, i, j, k) && (fl[gidx(g, i, j, k)] & FB_SOLID);
}

__global__ void __launch_bounds__(256)
k_advect(Grid g, int64_t n, const uint8_t* __restrict__ fl, double dt_max, double dxc, double e,
         double* __restrict__ px, double* __restrict__ py, double* __restrict__ pz,
         double* __restrict__ vx, double* __restrict__ vy, double* __restrict__ vz,
         DevScalars* __restrict__ ds) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double maxSpeed = __longlong_as_double((long long)ds->max_speed_bits);
  double timestep = dt_max;                                   // fluid.cc:992-999
  if (maxSpeed != 0.0) timestep = dt_max < dxc / maxSpeed ? dt_max : dxc / maxSpeed;
  if (t == 0) ds->dt_out = timestep;
  if (t >= n) return;
  double x = px[t], y = py[t], z = pz[t];
  double a = vx[t], b = vy[t], cc = vz[t];
  // position += timestep * velocity without FMA contraction (fluid.cc:1003)
  double nxp = __dadd_rn(x, __dmul_rn(timestep, a));
  double nyp = __dadd_rn(y, __dmul_rn(timestep, b));
  double nzp = __dadd_rn(z, __dmul_rn(timestep, cc));
  bool bad = !(fabs(nxp) < 1.0e9) || !(fabs(nyp) < 1.0e9) || !(fabs(nzp) < 1.0e9);
  int rx = bad ? 0 : (int)round(nxp), ry = bad ? 0 : (int)round(nyp), rz = bad ? 0 : (int)round(nzp);
  if (!bad && solid_at(g, fl, rx + g.ox, ry + g.oy, rz + g.oz)) {   // fluid.cc:1011
    double sx = __dmul_rn(a, timestep), sy = __dmul_rn(b, timestep), sz = __dmul_rn(cc, timestep);
    // openvdb::Coord(double,double,double) truncates toward zero (openvdb/math/Coord.h:61)
    int tx = (int)x, ty = (int)y, tz = (int)z;
    bool changed = false;
    if (solid_at(g, fl, (int)round(__dadd_rn(x, sx)) + g.ox, ty + g.oy, tz + g.oz)) { a *= -1.0 * e; changed = true; }   // :1017-1020
    if (solid_at(g, fl, tx + g.ox, (int)round(__dadd_rn(y, sy)) + g.oy, tz + g.oz)) { b *= -1.0 * e; changed = true; }   // :1021-1024
    if (solid_at(g, fl, tx + g.ox, ty + g.oy, (int)round(__dadd_rn(z, sz)) + g.oz)) { cc *= -1.0 * e; changed = true; }  // :1025-1028
    px[t] = __dadd_rn(x, __dmul_rn(a, timestep));             // fluid.cc:1029
    py[t] = __dadd_rn(y, __dmul_rn(b, timestep));
    pz[t] = __dadd_rn(z, __dmul_rn(cc, timestep));
    if (changed) { vx[t] = a; vy[t] = b; vz[t] = cc; }
  } else {
    px[t] = nxp; py[t] = nyp; pz[t] = nzp;                    // fluid.cc:1033
  }
}

__global__ void k_reset_max(DevScalars* ds) { ds->max_speed_bits = 0ull; }

// one particle of FLIPadvect phase 2 (fluid.cc:1000-1036), exactly as k_advect does it
__device__ __forceinline__ void advect_one(const Grid& g, const uint8_t* __restrict__ fl, double timestep, double e,
                                           double x, double y, double z, double& a, double& b, double& cc,
                                           double& ox, double& oy, double& oz) {
  const double nxp = __dadd_rn(x, __dmul_rn(timestep, a));
  const double nyp = __dadd_rn(y, __dmul_rn(timestep, b));
  const double nzp = __dadd_rn(z, __dmul_rn(timestep, cc));
  const bool bad = !(fabs(nxp) < 1.0e9) || !(fabs(nyp) < 1.0e9) || !(fabs(nzp) < 1.0e9);
  const int rx = bad ? 0 : (int)round(nxp), ry = bad ? 0 : (int)round(nyp), rz = bad ? 0 : (int)round(nzp);
  if (!bad && solid_at(g, fl, rx + g.ox, ry + g.oy, rz + g.oz)) {   // fluid.cc:1011
    const double sx = __dmul_rn(a, timestep), sy = __dmul_rn(b, timestep), sz = __dmul_rn(cc, timestep);
    const int tx = (int)x, ty = (int)y, tz = (int)z;               // Coord(double...) truncates (openvdb/math/Coord.h:61)
    if (solid_at(g, fl, (int)round(__dadd_rn(x, sx)) + g.ox, ty + g.oy, tz + g.oz)) a *= -1.0 * e;    // :1017-1020
    if (solid_at(g, fl, tx + g.ox, (int)round(__dadd_rn(y, sy)) + g.oy, tz + g.oz)) b *= -1.0 * e;    // :1021-1024
    if (solid_at(g, fl, tx + g.ox, ty + g.oy, (int)round(__dadd_rn(z, sz)) + g.oz)) cc *= -1.0 * e;   // :1025-1028
    ox = __dadd_rn(x, __dmul_rn(a, timestep));                     // :1029
    oy = __dadd_rn(y, __dmul_rn(b, timestep));
    oz = __dadd_rn(z, __dmul_rn(cc, timestep));
  } else {
    ox = nxp; oy = nyp; oz = nzp;                                  // :1033
  }
}

// the trailing "outside the box" bucket of the fused path: no grid contact (k_speed_tail), then advection
__global__ void k_g2p_advect_tail(Grid g, const int32_t* __restrict__ start, int64_t n /* = n_sorted */, const uint8_t* __restrict__ fl,
                                  const double* __restrict__ px, const double* __restrict__ py, const double* __restrict__ pz,
                                  const double* __restrict__ vx, const double* __restrict__ vy, const double* __restrict__ vz,
                                  double* __restrict__ opx, double* __restrict__ opy, double* __restrict__ opz,
                                  double* __restrict__ ovx, double* __restrict__ ovy, double* __restrict__ ovz,
                                  DevScalars* __restrict__ ds, double ratio, double timestep, double e_rest) {
  double speed = 0.0;
  for (int64_t t = (int64_t)start[g.total] + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n;
       t += (int64_t)gridDim.x * blockDim.x) {
    double x = px[t], y = py[t], z = pz[t];
    double a = vx[t], b = vy[t], cc = vz[t];
    if (ratio != 1.0) { a *= ratio; b *= ratio; cc *= ratio; }
    speed = fmax(speed, sqrt(a * a + b * b + cc * cc));
    advect_one(g, fl, timestep, e_rest, x, y, z, a, b, cc, x, y, z);
    opx[t] = x; opy[t] = y; opz[t] = z; ovx[t] = a; ovy[t] = b; ovz[t] = cc;
  }
  speed = warp_max(speed);
  if ((threadIdx.x & 31) == 0 && speed > 0.0) atomicMax(&ds->max_speed_bits, (unsigned long long)__double_as_longlong(speed));
}

// dt of fluid.cc:992-999 from the collected max speed (the two-kernel path does this inside k_advect)
__global__ void k_set_dt(DevScalars* ds, double dt_max, double dxc) {
  const double maxSpeed = __longlong_as_double((long long)ds->max_speed_bits);
  double timestep = dt_max;
  if (maxSpeed != 0.0) timestep = dt_max < dxc / maxSpeed ? dt_max : dxc / maxSpeed;
  ds->dt_out = timestep;
}

int fb_g2p_advect_rows(flipb200_ctx* c, double timestep);   // rows.cu

// c->opt_g2p (FLIPB200_G2P, read once at create):
//   't' (default) ONE pass over the particles: G2P + advection with the rows staged by TMA (rows.cu), betting on
//       dt == dt_max (true whenever maxSpeed <= dx / dt_max; tried when the previous step ended that way); the new state
//       goes to the alternate buffers, so a lost bet costs one wasted pass and the two-kernel path below runs on the
//       untouched inputs.  96 B / particle instead of 72 + 72: 3.0 ms against 4.5 ms at 256^3 / 127 M particles.
//   'c' two kernels: k_g2p_cell (one thread per cell) then k_advect;  'p' k_g2p (one thread per particle) then k_advect
int fb_g2p_advect(flipb200_ctx* c) {
  const Grid& g = c->g;
  Particles& pt = c->pt;
  k_reset_max<<<1, 1, 0, c->stream>>>(c->ds);
  c->launches++;
  int blocks = (int)((pt.n + 255) / 256);
  if (blocks == 0) blocks = 1;
  const char mode = c->opt_g2p;
  int rc;
  if (mode == 't' && c->sorted && pt.n == c->n_sorted && pt.n > 0 && (c->hs->dt_out == c->P.dt_max || c->hs->dt_out == 0.0)) {
    const double dt_spec = c->P.dt_max;
    if ((rc = fb_g2p_advect_rows(c, dt_spec))) return rc;
    // the trailing bucket (key == total) holds the particles outside the box: no grid contact, free flight
    k_g2p_advect_tail<<<64, 256, 0, c->stream>>>(g, c->cell_start, c->n_sorted, c->flags, pt.a[0], pt.a[1], pt.a[2], pt.a[3], pt.a[4], pt.a[5],
                                                 pt.b[0], pt.b[1], pt.b[2], pt.b[3], pt.b[4], pt.b[5], c->ds, c->P.flip_ratio, dt_spec,
                                                 c->P.restitution);
    CK_LAUNCH(); c->launches++;
    if ((rc = fb_allreduce(c, &c->ds->max_speed_bits, 1, FB_U64, FB_MAX))) return rc;   // maxSpeed is global (fluid.cc:976-991)
    k_set_dt<<<1, 1, 0, c->stream>>>(c->ds, c->P.dt_max, c->P.dx);
    CK_LAUNCH(); c->launches++;
    if ((rc = fb_fetch_scalars(c))) return rc;
    if (c->hs->dt_out == dt_spec) {                  // bet won: the alternate buffers hold the new state
      for (int q = 0; q < 6; ++q) { double* t = pt.a[q]; pt.a[q] = pt.b[q]; pt.b[q] = t; }
      c->sorted = false;                             // positions moved: cell order is stale (ids did not move: same index)
      return 0;
    }
    k_reset_max<<<1, 1, 0, c->stream>>>(c->ds);     // bet lost: pt.a is untouched
    c->launches++;
  }
  if (c->sorted && !(mode == 'p' && c->world == 1)) {
    if (c->n_sorted > 0) {
      dim3 blk(32, 4, 1), grd(cdiv(g.nx - 1, 32), cdiv(g.ny - 1, 4), g.k1 - g.k0);
      k_g2p_cell<<<grd, blk, 0, c->stream>>>(g, c->cell_start, pt.a[0], pt.a[1], pt.a[2], pt.a[3], pt.a[4], pt.a[5],
                                             c->r, c->pd, c->Ap, c->ds, c->P.flip_ratio);   // dc arrays written by fb_cell_delta
      c->launches++;
      k_speed_tail<<<64, 256, 0, c->stream>>>(c->cell_start, g.total, c->n_sorted, pt.a[3], pt.a[4], pt.a[5], c->ds, c->P.flip_ratio);
      c->launches++;
    }
    if (pt.n > c->n_sorted) {   // migrated-in tail
      const int64_t o = c->n_sorted, m = pt.n - o;
      k_g2p<<<(unsigned)((m + 255) / 256), 256, 0, c->stream>>>(g, m, c->flags, pt.a[0] + o, pt.a[1] + o, pt.a[2] + o,
                                                               pt.a[3] + o, pt.a[4] + o, pt.a[5] + o, c->r, c->pd, c->Ap, c->ds, c->P.flip_ratio);
      c->launches++;
    }
  } else {
    k_g2p<<<blocks, 256, 0, c->stream>>>(g, pt.n, c->flags, pt.a[0], pt.a[1], pt.a[2], pt.a[3], pt.a[4], pt.a[5],
                                         c->r, c->pd, c->Ap, c->ds, c->P.flip_ratio);   // dc arrays written by fb_cell_delta
    c->launches++;
  }
  CK_LAUNCH();
  // maxSpeed is a global max (fluid.cc:976-991): every rank advects with the same dt
  if ((rc = fb_allreduce(c, &c->ds->max_speed_bits, 1, FB_U64, FB_MAX))) return rc;
  k_advect<<<blocks, 256, 0, c->stream>>>(g, pt.n, c->flags, c->P.dt_max, c->P.dx, c->P.restitution, pt.a[0],
                                          pt.a[1], pt.a[2], pt.a[3], pt.a[4], pt.a[5], c->ds);
  CK_LAUNCH(); c->launches++;
  c->sorted = false;   // positions moved: cell order is stale
  return 0;
}

// ---------------------------------------------------------------------------------------------
// device seeding: 8 particles per voxel, one per 2x2x2 sub-cell, jittered with a counter hash
// (role of UniformPointScatter at fluid.cc:1348-1350; stratified like SURVEY 8d config 2/3)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__device__ __forceinline__ double u01(uint64_t h) {   // (0,1) open
  return ((double)(h >> 11) + 0.5) * (1.0 / 9007199254740992.0);
}
// t0: index of the first (voxel, sub-cell) pair this launch seeds within the WHOLE box, so that a
// z-slab rank draws exactly the jitters the single-GPU run draws for the same voxels.
__global__ void k_seed_box(Grid g, int lx, int ly, int lz, int dxn, int dyn, int64_t t0, int64_t count, uint64_t seed,
                           int64_t base, double* __restrict__ px, double* __restrict__ py, double* __restrict__ pz,
                           double* __restrict__ vx, double* __restrict__ vy, double* __restrict__ vz) {
  int64_t tl = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tl >= count) return;
  const int64_t t = t0 + tl;
  int64_t v = t >> 3; int sub = (int)(t & 7);
  int i = (int)(v % dxn), j = (int)((v / dxn) % dyn), k = (int)(v / ((int64_t)dxn * dyn));
  uint64_t h = splitmix64(seed ^ (uint64_t)t * 0xD1B54A32D192ED03ull);
  double jx = u01(h); h = splitmix64(h); double jy = u01(h); h = splitmix64(h); double jz = u01(h);
  int64_t o = base + tl;
  px[o] = (double)(lx + i - g.ox) - 0.5 + ((sub & 1) + jx) * 0.5;
  py[o] = (double)(ly + j - g.oy) - 0.5 + (((sub >> 1) & 1) + jy) * 0.5;
  pz[o] = (double)(lz + k - g.oz) - 0.5 + (((sub >> 2) & 1) + jz) * 0.5;   // lz, g.oz: local plane indices
  vx[o] = 0.0; vy[o] = 0.0; vz[o] = 0.0;
}

// lo/hi are GLOBAL voxel indices; a slab seeds the voxels of its owned planes (the first sort moves
// the particles whose base plane belongs to a neighbour).
int fb_seed_box(flipb200_ctx* c, const int32_t lo[3], const int32_t hi[3], uint64_t seed) {
  const Grid& g = c->g;
  int dxn = hi[0] - lo[0] + 1, dyn = hi[1] - lo[1] + 1, dzn = hi[2] - lo[2] + 1;
  if (dxn <= 0 || dyn <= 0 || dzn <= 0) return FLIPB200_ERR_ARG;
  const int zl = lo[2] > c->gz0 ? lo[2] : c->gz0, zh = hi[2] < c->gz1 - 1 ? hi[2] : c->gz1 - 1;
  if (zh < zl) return 0;
  const int64_t per_plane = (int64_t)dxn * dyn * 8;
  const int64_t t0 = (int64_t)(zl - lo[2]) * per_plane, add = (int64_t)(zh - zl + 1) * per_plane;
  int64_t base = c->pt.n;
  if (base + add >= (int64_t)2147483647) return FLIPB200_ERR_ARG;
  int rc = fb_reserve_particles(c, base + add);
  if (rc) return rc;
  int64_t blocks = (add + 255) / 256;
  k_seed_box<<<(unsigned)blocks, 256, 0, c->stream>>>(g, lo[0], lo[1], lo[2] - g.kg0, dxn, dyn, t0, add, seed, base,
                                                      c->pt.a[0], c->pt.a[1], c->pt.a[2], c->pt.a[3], c->pt.a[4], c->pt.a[5]);
  CK_LAUNCH(); c->launches++;
  if (fb_init_ids(c, base, add, c->world > 1 ? t0 : base)) return FLIPB200_ERR_CUDA;
  CK(cudaStreamSynchronize(c->stream));
  c->pt.n = base + add;
  c->sorted = false;
  c->sparse_ok = false;            // particles may have appeared anywhere: the next step sweeps the whole grid
  return 0;
}

// Re-seeding hook of the reference driver (commented out at fluid.cc:1374-1382 and :1495-1497):
// scatter new particles into a source box, then PointList::interpFromGrid (:883-895) gives every new
// particle the PIC velocity clampedCatmullRom(p, vels) (:125-207): sum(cw * getVelocity(node)) / sum(cw)
// over the in-shell nodes of its window, 0 when the weight sum is 0.  Here: fb_seed_box appends the
// particles, k_cell_delta with ratio 0 leaves getVelocity(c, vels) in the idle PCG vectors and the
// per-particle gather kernel with ratio 0 turns it into the particle velocity (v = 0*v + interp).
// Collective on z-slab runs (ghost fill of the cell-centred velocities).
int fb_emit_box(flipb200_ctx* c, const int32_t lo[3], const int32_t hi[3], uint64_t seed) {
  const int64_t base = c->pt.n;
  int rc = fb_seed_box(c, lo, hi, seed);
  if (rc) return rc;
  const int64_t add = c->pt.n - base;
  if ((rc = fb_cell_delta_r(c, 0.0))) return rc;
  if (add > 0) {
    Particles& pt = c->pt;
    k_g2p<<<(unsigned)((add + 255) / 256), 256, 0, c->stream>>>(c->g, add, c->flags, pt.a[0] + base, pt.a[1] + base, pt.a[2] + base,
                                                                pt.a[3] + base, pt.a[4] + base, pt.a[5] + base, c->r, c->pd, c->Ap, c->ds, 0.0);
    CK_LAUNCH(); c->launches++;
  }
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}
