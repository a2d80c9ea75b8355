// rows.cu -- particle <-> grid transfers with the particles of a cell ROW staged in shared memory by TMA.
//   P2G:  PointList::P2Gtransfer fluid.cc:1106-1148 (p2gCatmullRom :265-299)
//   G2P + advection in one pass: PointList::FLIPadvect :972-1038 (CatmullRomFLIP :210-263)
//
// Why (ncu, profiles/r1d_summary.txt): the one-thread-per-cell kernels of particles.cu keep the 8 x 4 partial
// sums of a cell in registers (no cross-lane traffic, 32 REDs per cell), but every lane reads ITS cell's
// particles with 8-byte loads that sit 64 B apart: 32 sectors per warp request, 45 % L1 hits, 16 resident
// warps with six loads in flight each -- 37 % of the DRAM bandwidth.  Particle-parallel kernels load
// perfectly but then have to move 32 contributions per particle across lanes (shuffle / shared-memory
// volume 10x the input), and a particle-parallel G2P is bound by its 24 gathered loads (measured 4.4 ms
// against 3.1 ms, profiles/r2e).  So the mapping stays one thread per cell, and the LOADS change:
// cell-sorted particles of 64 consecutive cells of an x-row are ONE contiguous range of the SoA arrays; one
// elected thread streams it into shared memory with 1-D bulk copies (cp.async.bulk, completion on an
// mbarrier: no registers, no LSU instructions, full 128-byte lines) two stages ahead of the arithmetic,
// and every lane walks its own cell's particles out of shared memory.  Lanes start their walk rotated by
// lane/2 so that the 64-bit shared loads of a half-warp fall into 16 different bank pairs at ~8 particles
// per cell.  G2P + advection write the new state back into the stage and a bulk store returns it to HBM in
// full lines -- the per-cell kernels could not write positions without a 64-byte-stride scatter.
#include "common.cuh"
#include <stdlib.h>

#define RS_CH 512                      // particles per stage (even)
#define RS_NST 2                       // stages
#define RS_T 64                        // threads per block of the P2G kernel = cells per group
#define RS_CELLS 64                    // cells per group
#define RS_GT 128                      // threads per block of the G2P kernel: TWO threads per cell (even / odd particles)
#define RS_LD (RS_CH + 8)              // doubles per attribute and stage (chunk start is aligned down to 16 bytes)
#define RS_STAGE_BYTES (6 * RS_LD * 8)
#define RS_SMEM (RS_NST * RS_STAGE_BYTES + 64)

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int cnt) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(cnt)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
// global -> shared, 16-byte aligned on both sides, size a multiple of 16
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// shared -> global
__device__ __forceinline__ void bulk_store(void* dst, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// the row's work items: groups of RS_T cells, each group's particle range cut into chunks of RS_CH
struct RowIter {
  int gi, ng, q0, cells;
  int glo, ghi, lo, hi;
  __device__ __forceinline__ void init(int q0_, int cells_) { q0 = q0_; cells = cells_; ng = (cells_ + RS_CELLS - 1) / RS_CELLS; gi = -1; glo = ghi = lo = hi = 0; }
  __device__ __forceinline__ bool next(const int32_t* __restrict__ start) {
    lo = hi;
    if (lo >= ghi) {
      do {
        if (++gi >= ng) return false;
        const int c0 = gi * RS_CELLS, c1 = min(c0 + RS_CELLS, cells);
        glo = start[q0 + c0]; ghi = start[q0 + c1];
      } while (ghi <= glo);
      lo = glo;
    }
    hi = min(lo + RS_CH, ghi);
    return true;
  }
};

template <int NATTR>
__device__ __forceinline__ void issue_load(double* stage, uint64_t* bar, const RowIter& it, const double* const* __restrict__ src) {
  const int alo = it.lo & ~1;
  const uint32_t n = (uint32_t)((it.hi - alo + 1) & ~1);
  mbar_expect(bar, NATTR * n * 8u);
#pragma unroll
  for (int a = 0; a < NATTR; ++a) bulk_load(stage + a * RS_LD, src[a] + alo, n * 8u, bar);
}

struct Ptrs6 { const double* p[6]; };
struct OPtrs6 { double* p[6]; };

// ---------------------------------------------------------------------------------------------
// P2G: one block per x-row of base cells (j, k); thread t <-> cell 64 gi + t of the current group.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(RS_GT, 4)
k_p2g_rows(Grid g, const int32_t* __restrict__ start, const uint8_t* __restrict__ fl, Ptrs6 in,
           double* __restrict__ wacc, double* __restrict__ ux, double* __restrict__ uy, double* __restrict__ uz) {
  extern __shared__ __align__(128) unsigned char smraw[];
  double* st = reinterpret_cast<double*>(smraw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smraw + RS_NST * RS_STAGE_BYTES);
  const int j = blockIdx.x, k = g.k0 + blockIdx.y;
  const int tid = threadIdx.x;
  const int q0 = gidx(g, 0, j, k);
  const int cells = g.nx - 1;
  if (start[q0] == start[q0 + cells]) return;                  // empty row (block-uniform)
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < RS_NST; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  RowIter prod, cons;
  prod.init(q0, cells); cons.init(q0, cells);
  bool pmore = true;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < RS_NST; ++s) { pmore = prod.next(start); if (pmore) issue_load<6>(st + s * 6 * RS_LD, &full[s], prod, in.p); else break; }
  }
  double aw[8], ax[8], ay[8], az[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) { aw[m] = 0; ax[m] = 0; ay[m] = 0; az[m] = 0; }
  int cur_g = -1, s_c = 0, e_c = 0, qc = 0;
  bool any = false;
  double nx0 = 0, ny0 = (double)(j - g.oy), nz0 = (double)(k - g.oz);
  // two threads per cell (even / odd particles, each with its own 8 x 4 partial sums): twice the warps on the same stage
  const int part = tid & 1;
  const int rot = 2 * (((tid & 31) >> 2) & 3);
  // called by every thread of the block at the same points: the pair of a cell adds its two partial sums with one
  // shuffle per value and the even lane sends the 32 REDs of the cell (as many atomics as one thread per cell)
  auto flush = [&]() {
    const int other = __shfl_xor_sync(0xffffffffu, (int)any, 1);      // unconditional: every lane of the warp takes part
    const bool anyc = any || other != 0;
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      aw[m] += __shfl_xor_sync(0xffffffffu, aw[m], 1); ax[m] += __shfl_xor_sync(0xffffffffu, ax[m], 1);
      ay[m] += __shfl_xor_sync(0xffffffffu, ay[m], 1); az[m] += __shfl_xor_sync(0xffffffffu, az[m], 1);
    }
    // the x+1 nodes of cell c are the x nodes of cell c + 1 (two lanes to the right): hand them over by shuffle, so a
    // cell sends 16 REDs instead of 32 (the kernel is bound by the fp64 RED rate); the last cell of the warp keeps its own
    const int lane = tid & 31;
    const bool left = __shfl_up_sync(0xffffffffu, (int)anyc, 2) != 0 && lane >= 2;
#pragma unroll
    for (int m = 0; m < 8; m += 2) {
      const double tw = __shfl_up_sync(0xffffffffu, aw[m + 1], 2), tx = __shfl_up_sync(0xffffffffu, ax[m + 1], 2);
      const double ty = __shfl_up_sync(0xffffffffu, ay[m + 1], 2), tz = __shfl_up_sync(0xffffffffu, az[m + 1], 2);
      if (left) { aw[m] += tw; ax[m] += tx; ay[m] += ty; az[m] += tz; }
    }
    if (part == 0 && (anyc || left)) {
      const int qn = q0 + cur_g * RS_CELLS + (tid >> 1);            // this lane's cell even when it lies past the row's last cell
#pragma unroll
      for (int m = 0; m < 8; ++m) {
        if ((m & 1) ? !(lane == 30 && anyc) : false) continue;
        const int qq = qn + (m & 1) + ((m & 2) ? g.px : 0) + ((m & 4) ? g.plane : 0);
        if (!(fl[qq] & FB_SOLID)) {                               // fluid.cc:288
          atomicAdd(&wacc[qq], aw[m]); atomicAdd(&ux[qq], ax[m]); atomicAdd(&uy[qq], ay[m]); atomicAdd(&uz[qq], az[m]);
        }
      }
    }
#pragma unroll
    for (int m = 0; m < 8; ++m) { aw[m] = 0; ax[m] = 0; ay[m] = 0; az[m] = 0; }
    any = false;
  };
  for (int item = 0; cons.next(start); ++item) {
    const int s = item % RS_NST;
    if (cons.gi != cur_g) {                                      // new group: flush the old cell, take the new one
      flush();
      cur_g = cons.gi;
      const int c = cur_g * RS_CELLS + (tid >> 1);
      if (c < cells) { qc = q0 + c; s_c = start[qc]; e_c = start[qc + 1]; } else { s_c = e_c = 0; }
      nx0 = (double)(c - g.ox);
    }
    mbar_wait(&full[s], (uint32_t)((item / RS_NST) & 1));
    const double* S = st + s * 6 * RS_LD;
    const int alo = cons.lo & ~1;
    const int lo_t = max(s_c, cons.lo), hi_t = min(e_c, cons.hi), cnt = hi_t - lo_t;
    const int mine = cnt > 0 ? (cnt - part + 1) >> 1 : 0;
    if (mine > 0) {
      any = true;
      int n = lo_t + part + (rot % (2 * mine));
      for (int t = 0; t < mine; ++t) {
        const int u = n - alo;
        const double x = S[u], y = S[RS_LD + u], z = S[2 * RS_LD + u];
        const double a = S[3 * RS_LD + u], b = S[4 * RS_LD + u], cc = S[5 * RS_LD + u];
        double wx0, wx1, wy0, wy1, wz0, wz1;
        spline_pair(x, nx0, wx0, wx1); spline_pair(y, ny0, wy0, wy1); spline_pair(z, nz0, wz0, wz1);
#pragma unroll
        for (int m = 0; m < 8; ++m) {
          const double cw = ((m & 1) ? wx1 : wx0) * ((m & 2) ? wy1 : wy0) * ((m & 4) ? wz1 : wz0);   // (sx*sy)*sz, fluid.cc:291
          aw[m] += cw; ax[m] += cw * a; ay[m] += cw * b; az[m] += cw * cc;
        }
        n += 2;
        if (n >= hi_t) n = lo_t + part;
      }
    }
    __syncthreads();                                             // everyone is done with stage s
    if (tid == 0 && pmore) { pmore = prod.next(start); if (pmore) issue_load<6>(st + s * 6 * RS_LD, &full[s], prod, in.p); }
  }
  flush();
}

// one particle of FLIPadvect phase 2 (fluid.cc:1000-1036), the same sequence as k_advect of particles.cu
__device__ __forceinline__ bool solid_at_r(const Grid& g, const uint8_t* __restrict__ fl, int i, int j, int k) {
  return in_box(g, i, j, k) && (fl[gidx(g, i, j, k)] & FB_SOLID);
}
__device__ __forceinline__ void advect_one_r(const Grid& g, const uint8_t* __restrict__ fl, double timestep, double e,
                                             double x, double y, double z, double& a, double& b, double& cc,
                                             double& ox, double& oy, double& oz) {
  const double nxp = __dadd_rn(x, __dmul_rn(timestep, a));
  const double nyp = __dadd_rn(y, __dmul_rn(timestep, b));
  const double nzp = __dadd_rn(z, __dmul_rn(timestep, cc));
  const bool bad = !(fabs(nxp) < 1.0e9) || !(fabs(nyp) < 1.0e9) || !(fabs(nzp) < 1.0e9);
  const int rx = bad ? 0 : (int)round(nxp), ry = bad ? 0 : (int)round(nyp), rz = bad ? 0 : (int)round(nzp);
  if (!bad && solid_at_r(g, fl, rx + g.ox, ry + g.oy, rz + g.oz)) {   // fluid.cc:1011
    const double sx = __dmul_rn(a, timestep), sy = __dmul_rn(b, timestep), sz = __dmul_rn(cc, timestep);
    const int tx = (int)x, ty = (int)y, tz = (int)z;               // Coord(double...) truncates (openvdb/math/Coord.h:61)
    if (solid_at_r(g, fl, (int)round(__dadd_rn(x, sx)) + g.ox, ty + g.oy, tz + g.oz)) a *= -1.0 * e;    // :1017-1020
    if (solid_at_r(g, fl, tx + g.ox, (int)round(__dadd_rn(y, sy)) + g.oy, tz + g.oz)) b *= -1.0 * e;    // :1021-1024
    if (solid_at_r(g, fl, tx + g.ox, ty + g.oy, (int)round(__dadd_rn(z, sz)) + g.oz)) cc *= -1.0 * e;   // :1025-1028
    ox = __dadd_rn(x, __dmul_rn(a, timestep));                     // :1029
    oy = __dadd_rn(y, __dmul_rn(b, timestep));
    oz = __dadd_rn(z, __dmul_rn(cc, timestep));
  } else {
    ox = nxp; oy = nyp; oz = nzp;                                  // :1033
  }
}

// ---------------------------------------------------------------------------------------------
// G2P (FLIP / PIC blend) + advection with a GIVEN time step, one pass: the new (p, v) of every particle of
// the row go to `out` (the sort's alternate buffers), the inputs stay untouched.  ADVECT = false: G2P only
// (positions are copied), used when dt turns out to differ from the bet.  Max speed -> ds->max_speed_bits.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(RS_GT, 4)
k_g2p_advect_rows(Grid g, const int32_t* __restrict__ start, const uint8_t* __restrict__ fl, Ptrs6 in, OPtrs6 out,
                  const double* __restrict__ dcx, const double* __restrict__ dcy, const double* __restrict__ dcz,
                  DevScalars* __restrict__ ds, double ratio, double timestep, double e_rest) {
  extern __shared__ __align__(128) unsigned char smraw[];
  double* st = reinterpret_cast<double*>(smraw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smraw + RS_NST * RS_STAGE_BYTES);
  __shared__ double red[RS_GT / 32];
  const int j = blockIdx.x, k = g.k0 + blockIdx.y;
  const int tid = threadIdx.x;
  const int q0 = gidx(g, 0, j, k);
  const int cells = g.nx - 1;
  if (start[q0] == start[q0 + cells]) return;                  // empty row (block-uniform)
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < RS_NST; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  RowIter prod, cons;
  prod.init(q0, cells); cons.init(q0, cells);
  bool pmore = true;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < RS_NST; ++s) { pmore = prod.next(start); if (pmore) issue_load<6>(st + s * 6 * RS_LD, &full[s], prod, in.p); else break; }
  }
  double ddx[8], ddy[8], ddz[8];
  unsigned inmask = 0;
  int cur_g = -1, s_c = 0, e_c = 0;
  double nx0 = 0, ny0 = (double)(j - g.oy), nz0 = (double)(k - g.oz);
  // two threads per cell: thread 2c takes the cell's even particles (counted from its first one), 2c + 1 the odd ones --
  // the stage is the same size, the block has twice the warps to hide the fp64 / shared-memory latency with.
  // rot: the pair of cell c starts its walk 2 (c / 2 mod 4) particles in, which spreads a half-warp's 64-bit loads over 16 bank pairs
  const int part = tid & 1;
  const int rot = 2 * (((tid & 31) >> 2) & 3);
  double speed = 0.0;
  for (int item = 0; cons.next(start); ++item) {
    const int s = item % RS_NST;
    if (cons.gi != cur_g) {
      cur_g = cons.gi;
      const int c = cur_g * RS_CELLS + (tid >> 1);
      s_c = e_c = 0;
      if (c < cells) { const int qc = q0 + c; s_c = start[qc]; e_c = start[qc + 1]; }
      nx0 = (double)(c - g.ox);
      if (e_c > s_c) {
        const int qc = q0 + c;
        inmask = 0;
#pragma unroll
        for (int m = 0; m < 8; ++m) {
          const int ii = c + (m & 1), jj = j + ((m >> 1) & 1), kk = k + ((m >> 2) & 1);
          const int qq = qc + (m & 1) + ((m & 2) ? g.px : 0) + ((m & 4) ? g.plane : 0);
          const bool in_s = inside_shell(g, ii, jj, kk);             // fluid.cc:237 (no solid test)
          ddx[m] = in_s ? dcx[qq] : 0.0; ddy[m] = in_s ? dcy[qq] : 0.0; ddz[m] = in_s ? dcz[qq] : 0.0;
          if (in_s) inmask |= 1u << m;
        }
      }
    }
    mbar_wait(&full[s], (uint32_t)((item / RS_NST) & 1));
    double* S = st + s * 6 * RS_LD;
    const int alo = cons.lo & ~1;
    const int lo_t = max(s_c, cons.lo), hi_t = min(e_c, cons.hi), cnt = hi_t - lo_t;
    // elements of the chunk that the aligned bulk store cannot cover (see below) are written by their owner
    const int blo = (cons.lo + 1) & ~1, bhi = cons.hi & ~1;
    if (cnt > 0) {
      // one particle: G2P, advection, results back into the stage (and straight to HBM for the chunk's odd ends)
      auto one = [&](int n, double x, double y, double z, double a, double b, double cc) {
        const int u = n - alo;
        double wx0, wx1, wy0, wy1, wz0, wz1;
        spline_pair(x, nx0, wx0, wx1); spline_pair(y, ny0, wy0, wy1); spline_pair(z, nz0, wz0, wz1);
        double wsum = 0, dx = 0, dy = 0, dz = 0;
#pragma unroll
        for (int m = 0; m < 8; ++m) {
          const double cw = ((m & 1) ? wx1 : wx0) * ((m & 2) ? wy1 : wy0) * ((m & 4) ? wz1 : wz0);
          if (inmask & (1u << m)) wsum += cw;
          dx += ddx[m] * cw; dy += ddy[m] * cw; dz += ddz[m] * cw;
        }
        if (ratio != 1.0) { a *= ratio; b *= ratio; cc *= ratio; }          // PIC/FLIP blend (grid.cu k_cell_delta)
        if (wsum != 0.0) { const double iw = 1.0 / wsum; a += dx * iw; b += dy * iw; cc += dz * iw; }   // fluid.cc:258-262, 981
        speed = fmax(speed, a * a + b * b + cc * cc);                       // before the collision, fluid.cc:982
        advect_one_r(g, fl, timestep, e_rest, x, y, z, a, b, cc, x, y, z);
        S[u] = x; S[RS_LD + u] = y; S[2 * RS_LD + u] = z; S[3 * RS_LD + u] = a; S[4 * RS_LD + u] = b; S[5 * RS_LD + u] = cc;
        if (n < blo || n >= bhi) { out.p[0][n] = x; out.p[1][n] = y; out.p[2][n] = z; out.p[3][n] = a; out.p[4][n] = b; out.p[5][n] = cc; }
      };
      // this thread's share: particles lo_t + part, lo_t + part + 2, ... walked from a rotated start
      const int mine = (cnt - part + 1) >> 1;
      if (mine > 0) {
        int n = lo_t + part + (rot % (2 * mine));
        if (n >= hi_t) n = lo_t + part;
        for (int t = 0; t < mine; ++t) {
          const int u = n - alo;
          one(n, S[u], S[RS_LD + u], S[2 * RS_LD + u], S[3 * RS_LD + u], S[4 * RS_LD + u], S[5 * RS_LD + u]);
          n += 2;
          if (n >= hi_t) n = lo_t + part;
        }
      }
    }
    fence_proxy_async();                                         // my generic-proxy writes to the stage before the bulk store reads it
    __syncthreads();
    if (tid == 0) {
      // [blo, bhi) is 16-byte aligned on both ends; an odd first / last element of the chunk may share its
      // 16 bytes with a particle of another chunk (possibly another block's), so its owner stored it above
      if (bhi > blo) {
#pragma unroll
        for (int a = 0; a < 6; ++a) bulk_store(out.p[a] + blo, S + a * RS_LD + (blo - alo), (uint32_t)(bhi - blo) * 8u);
        bulk_commit();
        bulk_wait_read0();                                       // the stage may be overwritten once the store has READ it
      }
      if (pmore) { pmore = prod.next(start); if (pmore) issue_load<6>(S, &full[s], prod, in.p); }
    }
  }
  if (tid == 0) bulk_wait0();
  speed = sqrt(warp_max(speed));                                 // fluid.cc:982 (Vec3::length)
  if ((tid & 31) == 0) red[tid >> 5] = speed;
  __syncthreads();
  if (tid == 0) {
    double m = red[0];
#pragma unroll
    for (int w = 1; w < RS_GT / 32; ++w) m = fmax(m, red[w]);
    if (m > 0.0) atomicMax(&ds->max_speed_bits, (unsigned long long)__double_as_longlong(m));
  }
}

}  // namespace

static int rows_attr(flipb200_ctx* c) {
  static bool done = false;
  if (!done) {
    CK(cudaFuncSetAttribute(k_p2g_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, RS_SMEM));
    CK(cudaFuncSetAttribute(k_g2p_advect_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, RS_SMEM));
    // four 50 KB blocks per SM need the large shared-memory carve-out (the default heuristic may leave room for one)
    CK(cudaFuncSetAttribute(k_p2g_rows, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    CK(cudaFuncSetAttribute(k_g2p_advect_rows, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    done = true;
  }
  return 0;
}

// splat of the cell-sorted particles [0, n_sorted) (the caller has cleared the accumulators)
int fb_p2g_rows(flipb200_ctx* c) {
  const Grid& g = c->g;
  Particles& pt = c->pt;
  int rc = rows_attr(c);
  if (rc) return rc;
  Ptrs6 in; for (int q = 0; q < 6; ++q) in.p[q] = pt.a[q];
  const int rows_k = min(g.k1, g.gnz - 1 - g.kg0) - g.k0;       // base planes k with k + kg0 <= gnz - 2
  if (rows_k <= 0) return 0;
  k_p2g_rows<<<dim3(g.ny - 1, rows_k), RS_GT, RS_SMEM, c->stream>>>(g, c->cell_start, c->flags_static, in, c->wacc, c->u[0], c->u[1], c->u[2]);
  CK_LAUNCH(); c->launches++;
  return 0;
}

// G2P + advection with time step `timestep` for the cell-sorted particles [0, n_sorted) into pt.b (inputs untouched)
int fb_g2p_advect_rows(flipb200_ctx* c, double timestep) {
  const Grid& g = c->g;
  Particles& pt = c->pt;
  int rc = rows_attr(c);
  if (rc) return rc;
  Ptrs6 in; OPtrs6 out;
  for (int q = 0; q < 6; ++q) { in.p[q] = pt.a[q]; out.p[q] = pt.b[q]; }
  const int rows_k = min(g.k1, g.gnz - 1 - g.kg0) - g.k0;
  if (rows_k <= 0) return 0;
  k_g2p_advect_rows<<<dim3(g.ny - 1, rows_k), RS_GT, RS_SMEM, c->stream>>>(g, c->cell_start, c->flags, in, out, c->r, c->pd, c->Ap, c->ds,
                                                                         c->P.flip_ratio, timestep, c->P.restitution);
  CK_LAUNCH(); c->launches++;
  return 0;
}
