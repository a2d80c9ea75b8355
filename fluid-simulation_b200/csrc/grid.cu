// grid.cu -- grid side of the FLIP substep: normalise + classify, rhs/divergence (+ relaxation
// error), pressure-gradient/gravity velocity update, cell-centred FLIP delta.
//   reference: P2Gtransfer normalise fluid.cc:1131-1146; main :1388-1455; setRHS :414-479;
//              setDiver :566-610; setOnlyB :543-564; velUpdate :612-703; getVelocity :59-70
#include "common.cuh"

// ---------------------------------------------------------------------------------------------
// normalise + classify + velBeforeUpdate copy, whole grid, one pass:
//   wf = float(w)                    weights / containerGrid are FloatGrids (fluid.cc:809,1157)
//   if (wf > 0) u /= wf              fluid.cc:1138-1141
//   ub = u                           velBeforeUpdate = vels->deepCopy(), fluid.cc:1455
//   fluid = !solid && wf > 0         fluid.cc:1423-1428 (inside-shell is implied: the shell is solid)
// The second splat PointList::interpolate (fluid.cc:843-882) accumulates the same cw into
// containerGrid as p2gCatmullRom does into weights, so container := weights (no second pass).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, FB_MINB_CLASSIFY)
k_classify(Grid g, const double* __restrict__ wacc, float* __restrict__ wf,
           double* __restrict__ u0, double* __restrict__ u1, double* __restrict__ u2,
           double* __restrict__ b0, double* __restrict__ b1, double* __restrict__ b2,
           double* __restrict__ xpress,
           const uint8_t* __restrict__ fs, uint8_t* __restrict__ fl, DevScalars* __restrict__ ds, int keep_x, Box R) {
  __shared__ int sred[7][8];
  int i = R.x0 + blockIdx.x * blockDim.x + threadIdx.x;
  int j = R.y0 + blockIdx.y * blockDim.y + threadIdx.y;
  int k = R.z0 + blockIdx.z;                      // owned planes of the active region (fb_active_region)
  int cnt = 0, mnx = 1 << 30, mny = 1 << 30, mnz = 1 << 30, mxx = -1, mxy = -1, mxz = -1;
  if (i < R.x1 && j < R.y1) {
    int q = gidx(g, i, j, k);
    float w = (float)wacc[q];
    wf[q] = w;
    double a = u0[q], b = u1[q], cc = u2[q];
    if (w > 0.0f) {
      double wd = (double)w;
      a = a / wd; b = b / wd; cc = cc / wd;
      u0[q] = a; u1[q] = b; u2[q] = cc;
    }
    b0[q] = a; b1[q] = b; b2[q] = cc;
    uint8_t f = fs[q];
    // a fluid cell whose six neighbours are all solid has Adiag == 0 and no row in the
    // reference's matrix (fluid.cc:498); it is left out of the system here as well.
    bool fluid = !(f & FB_SOLID) && w > 0.0f && (f & FB_SMASK) != FB_SMASK;
    // pressure: 0 off the fluid cells (every solver kernel relies on it); with params.warm_start the value of a
    // cell that stays fluid survives as the next solve's initial guess (pcg.cu)
    if (!(keep_x && fluid)) xpress[q] = 0.0;
    if (fluid) { f |= FB_FLUID; cnt = 1; }
    // bounding box (GLOBAL planes) of the WET nodes -- every fluid cell, and the rare non-solid node with weight that
    // is no row of the system: the next step's active region is grown from it, so it must cover all non-empty nodes
    if (w > 0.0f && !(f & FB_SOLID)) { mnx = mxx = i; mny = mxy = j; mnz = mxz = k + g.kg0; }
    fl[q] = f;
  }
  // block reduction (count, bbox) -> one atomic set per block
  int tid = threadIdx.x + blockDim.x * threadIdx.y, lane = tid & 31, wid = tid >> 5;
  cnt = warp_sum_i(cnt);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mnx = min(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mny = min(mny, __shfl_xor_sync(0xffffffffu, mny, o));
    mnz = min(mnz, __shfl_xor_sync(0xffffffffu, mnz, o)); mxx = max(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
    mxy = max(mxy, __shfl_xor_sync(0xffffffffu, mxy, o)); mxz = max(mxz, __shfl_xor_sync(0xffffffffu, mxz, o));
  }
  if (lane == 0) {
    sred[0][wid] = cnt; sred[1][wid] = mnx; sred[2][wid] = mny; sred[3][wid] = mnz;
    sred[4][wid] = mxx; sred[5][wid] = mxy; sred[6][wid] = mxz;
  }
  __syncthreads();
  if (tid == 0) {
    int nw = (blockDim.x * blockDim.y + 31) >> 5;
    for (int w = 1; w < nw; ++w) {
      cnt += sred[0][w];
      mnx = min(mnx, sred[1][w]); mny = min(mny, sred[2][w]); mnz = min(mnz, sred[3][w]);
      mxx = max(mxx, sred[4][w]); mxy = max(mxy, sred[5][w]); mxz = max(mxz, sred[6][w]);
    }
    if (mxx >= 0) {
      // 65 k blocks on a 256^3 tank: one counter word and six bbox words serialised their atomics in L2
      // (ncu: the kernel's tail was atomic throughput).  Counts go to 64 words picked by block index, bbox
      // atomics are sent only when they would change the (monotone) value just read.
      if (cnt) atomicAdd(&ds->n_fluid_part[(blockIdx.x + 3 * blockIdx.y + 7 * blockIdx.z) & 63], cnt);
      volatile int* bb = ds->bbox;
      if (mnx < bb[0]) atomicMin(&ds->bbox[0], mnx);
      if (mny < bb[1]) atomicMin(&ds->bbox[1], mny);
      if (mnz < bb[2]) atomicMin(&ds->bbox[2], mnz);
      if (mxx > bb[3]) atomicMax(&ds->bbox[3], mxx);
      if (mxy > bb[4]) atomicMax(&ds->bbox[4], mxy);
      if (mxz > bb[5]) atomicMax(&ds->bbox[5], mxz);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// extrapolate (fluid.cc:705-802; the call at :1147 is commented out, params.extrapolate = 1 runs it).
// `defined` of P2Gtransfer (:1109-1125, :1141) = outside 58 | solid | w > 0, so the targets are the
// non-solid in-shell air nodes.  Wave 1 averages, for every such node, the velocities of the
// w > 0 nodes of its 3x3x3 neighbourhood (clamped to the box); wave n + 1 those of the nodes wave n
// defined -- the reference's frontier BFS as a gather: layer[q] = wave that defined q (0: w > 0,
// -1: not yet, -2: never a target), sources of wave n are the nodes of layer n - 1.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_extrap_init(Grid g, const float* __restrict__ wf, const uint8_t* __restrict__ fs, int32_t* __restrict__ layer) {
  int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y, k = blockIdx.z;
  if (i >= g.nx || j >= g.ny) return;
  const int q = gidx(g, i, j, k);
  const bool blocked = !inside_shell(g, i, j, k) || (fs[q] & FB_SOLID);
  layer[q] = blocked ? -2 : (wf[q] > 0.0f ? 0 : -1);
}
__global__ void __launch_bounds__(256)
k_extrap_wave(Grid g, int wave, int32_t* __restrict__ layer, double* __restrict__ u0, double* __restrict__ u1,
              double* __restrict__ u2, int* __restrict__ n_new) {
  int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y, k = blockIdx.z;
  if (i >= g.nx || j >= g.ny) return;
  const int q = gidx(g, i, j, k);
  if (layer[q] != -1) return;
  double a = u0[q], b = u1[q], c = u2[q];                    // the sums start from the node's own value (:739)
  int cnt = 0;
  for (int dk = -1; dk <= 1; ++dk)
    for (int dj = -1; dj <= 1; ++dj)
      for (int di = -1; di <= 1; ++di) {
        const int ii = i + di, jj = j + dj, kk = k + dk;
        if (ii < 0 || jj < 0 || kk < 0 || ii >= g.nx || jj >= g.ny || kk >= g.nz) continue;
        const int n = gidx(g, ii, jj, kk);
        if (layer[n] != wave - 1) continue;                   // wave-(n-1) nodes are read-only during wave n
        a += u0[n]; b += u1[n]; c += u2[n]; ++cnt;
      }
  if (cnt) {
    u0[q] = a / cnt; u1[q] = b / cnt; u2[q] = c / cnt;       // :751, :796
    layer[q] = wave;
    atomicAdd(n_new, 1);
  }
}

int fb_extrapolate(flipb200_ctx* c) {
  const Grid& g = c->g;
  if (c->world > 1) return FLIPB200_ERR_ARG;                  // the waves would need a ghost exchange each: single GPU only
  if (!c->ext_layer) CK(cudaMalloc((void**)&c->ext_layer, (size_t)g.total * sizeof(int32_t) + 256));
  int* n_new = reinterpret_cast<int*>(c->ext_layer + g.total);
  dim3 blk(32, 8, 1), grd(cdiv(g.nx, 32), cdiv(g.ny, 8), g.nz);
  k_extrap_init<<<grd, blk, 0, c->stream>>>(g, c->wf, c->flags_static, c->ext_layer);
  c->launches++;
  const int max_waves = g.nx + g.ny + g.nz;
  for (int wave = 1; wave <= max_waves; wave += 4) {           // one host round trip per four waves
    CK(cudaMemsetAsync(n_new, 0, sizeof(int), c->stream));
    for (int w = wave; w < wave + 4; ++w)
      k_extrap_wave<<<grd, blk, 0, c->stream>>>(g, w, c->ext_layer, c->u[0], c->u[1], c->u[2], n_new);
    c->launches += 4;
    int h = 0;
    CK(cudaMemcpyAsync(&h, n_new, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    if (h == 0) break;
  }
  // velBeforeUpdate = vels->deepCopy() (fluid.cc:1455) comes after P2Gtransfer: take the extrapolated field
  for (int a = 0; a < 3; ++a)
    CK(cudaMemcpyAsync(c->ub[a], c->u[a], (size_t)g.total * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  CK_LAUNCH();
  return 0;
}

__global__ void k_classify_sum(DevScalars* ds) {
  int v = ds->n_fluid_part[threadIdx.x] + ds->n_fluid_part[threadIdx.x + 32];
  v = warp_sum_i(v);
  if (threadIdx.x == 0) ds->n_fluid = v;
}
__global__ void k_reset_classify(DevScalars* ds) {
  ds->n_fluid_part[threadIdx.x] = 0;
  if (threadIdx.x) return;
  ds->n_fluid = 0;
  ds->bbox[0] = ds->bbox[1] = ds->bbox[2] = 1 << 30;
  ds->bbox[3] = ds->bbox[4] = ds->bbox[5] = -1;
}

Box fb_active_region(const flipb200_ctx* c, bool* full) {
  const Grid& g = c->g;
  Box r{0, 0, g.k0, g.nx, g.ny, g.k1};
  const bool sparse = c->opt_sparse && c->sparse_ok && !c->P.extrapolate;
  if (full) *full = !sparse;
  if (!sparse) return r;
  const int* w = c->wet;
  r.x0 = max(0, w[0] - SPARSE_MARGIN) & ~31;           // rows stay sector/line aligned
  r.y0 = max(0, w[1] - SPARSE_MARGIN);
  r.x1 = min(g.nx, w[3] + SPARSE_MARGIN + 1); r.y1 = min(g.ny, w[4] + SPARSE_MARGIN + 1);
  r.z0 = max(max(0, w[2] - SPARSE_MARGIN) - g.kg0, g.k0);
  r.z1 = min(min(g.gnz, w[5] + SPARSE_MARGIN + 1) - g.kg0, g.k1);
  if (r.z1 < r.z0) r.z1 = r.z0;                        // the region does not reach this rank's planes
  return r;
}

int fb_classify(flipb200_ctx* c) {
  const Grid& g = c->g;
  bool full = true;
  const Box R = fb_active_region(c, &full);
  int prev_wet[6];
  for (int a = 0; a < 6; ++a) prev_wet[a] = c->wet[a];
  k_reset_classify<<<1, 64, 0, c->stream>>>(c->ds);
  if (R.z1 > R.z0) {
    dim3 blk(32, 8, 1), grd(cdiv(R.x1 - R.x0, 32), cdiv(R.y1 - R.y0, 8), R.z1 - R.z0);
    k_classify<<<grd, blk, 0, c->stream>>>(g, c->wacc, c->wf, c->u[0], c->u[1], c->u[2], c->ub[0], c->ub[1], c->ub[2], c->x,
                                           c->flags_static, c->flags, c->ds, (c->P.warm_start && c->x_valid) ? 1 : 0, R);
  }
  k_classify_sum<<<1, 32, 0, c->stream>>>(c->ds);
  CK_LAUNCH(); c->launches += 3;
  int rc;
  if (c->P.extrapolate && (rc = fb_extrapolate(c))) return rc;
  if (c->world > 1) {
    // global cell count and bounding box; ghost planes of what the neighbours' stencils read
    if ((rc = fb_allreduce(c, &c->ds->n_fluid, 1, FB_I32, FB_SUM))) return rc;
    if ((rc = fb_allreduce(c, &c->ds->bbox[0], 3, FB_I32, FB_MIN))) return rc;
    if ((rc = fb_allreduce(c, &c->ds->bbox[3], 3, FB_I32, FB_MAX))) return rc;
    if ((rc = fb_halo_fill(c, g, c->flags, 1, 1, 1))) return rc;
    if ((rc = fb_halo_fill(c, g, c->u[2], sizeof(double), 0, 1))) return rc;    // u_z[c+e_z] of setRHS/setDiver/getVelocity
    if ((rc = fb_halo_fill(c, g, c->ub[2], sizeof(double), 0, 1))) return rc;
    // pressure ghosts start at 0 like the owned planes
    CK(cudaMemsetAsync(c->x, 0, (size_t)g.k0 * g.plane * sizeof(double), c->stream));
    CK(cudaMemsetAsync(c->x + (size_t)g.k1 * g.plane, 0, (size_t)(g.nz - g.k1) * g.plane * sizeof(double), c->stream));
  }
  rc = fb_fetch_scalars(c);
  if (rc) return rc;
  c->n_fluid = c->hs->n_fluid;
  {
    // the wet bounding box of THIS step: the next step's active region.  A sparse sweep whose result reaches the rim of
    // its region cannot vouch for what lies beyond (it never should: the margin is one node wider than a particle can
    // travel): start over with the whole grid.
    const int* bb = c->hs->bbox;
    const bool have = bb[3] >= bb[0] && bb[4] >= bb[1] && bb[5] >= bb[2];
    if (!full && have) {
      const int n[3] = {g.nx, g.ny, g.gnz};
      bool rim = false;
      for (int a = 0; a < 3; ++a) {
        const int lo = max(0, prev_wet[a] - SPARSE_MARGIN), hi = min(n[a], prev_wet[a + 3] + SPARSE_MARGIN + 1);   // [lo, hi)
        if ((lo > 0 && bb[a] <= lo) || (hi < n[a] && bb[a + 3] >= hi - 1)) rim = true;
      }
      if (rim) {
        c->sparse_ok = false;
        if ((rc = fb_p2g(c))) return rc;
        return fb_classify(c);
      }
    }
    c->sparse_ok = have;
    for (int a = 0; a < 6; ++a) c->wet[a] = bb[a];
  }
  Box b, gb;
  if (c->n_fluid == 0) {
    b = Box{0, 0, 0, 0, 0, 0}; gb = b;
  } else {
    const int* bb = c->hs->bbox;
    gb.x0 = max(0, bb[0] - 2) & ~31;            // rows stay sector/line aligned
    gb.y0 = max(0, bb[1] - 2); gb.z0 = max(0, bb[2] - 2);
    gb.x1 = min(g.nx, bb[3] + 3); gb.y1 = min(g.ny, bb[4] + 3); gb.z1 = min(g.gnz, bb[5] + 3);
    // with extrapolated (non-zero) air velocities velUpdate's solid loop (fluid.cc:668-700) matters on
    // every solid cell of the domain, not only next to the fluid: sweep the whole grid
    if (c->P.extrapolate) { gb.x0 = gb.y0 = gb.z0 = 0; gb.x1 = g.nx; gb.y1 = g.ny; gb.z1 = g.gnz; }
    b = gb;
    // this rank sweeps the part of the global box that lies in its owned planes (local indices);
    // a rank the box does not reach sweeps one inert plane so that every rank launches the same kernels
    b.z0 = max(gb.z0, c->gz0) - g.kg0; b.z1 = min(gb.z1, c->gz1) - g.kg0;
    if (b.z1 <= b.z0) { b.z0 = g.k0; b.z1 = g.k0 + 1; }
  }
  c->box = b;
  c->gbox = gb;
  c->bcur = 0;
  c->mg_dirty = true;
  return 0;
}

// ---------------------------------------------------------------------------------------------
// rhs + divergence in one sweep (setRHS fluid.cc:414-479, setDiver :566-610).  Arithmetic and
// fp32 stores follow the reference statement by statement (explicit _rn intrinsics: no FMA).
//   MODE 0: bout = b;           ds->bnorm2  = |b|^2
//   MODE 1: bout = b2 (new);    ds->err_num = |bold - b2|^2, ds->err_den = |bold|^2 (fluid.cc:1483)
//           and ds->bnorm2 = |b2|^2 (b2 is the next pass's b: same velocities, same dt)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float rhs_div_cell(uint8_t f, double vx, double vy, double vz, double vix, double vjy, double vkz,
                                              double scale, double dx, double gx, double gy, double gz) {
  float R = 0.0f;
  if (f & FB_SXM) R = __double2float_rn(__dsub_rn((double)R, __dmul_rn(scale, __dadd_rn(vx, gx))));    // :449
  if (f & FB_SXP) R = __double2float_rn(__dadd_rn((double)R, __dmul_rn(scale, __dadd_rn(vix, gx))));   // :453
  if (f & FB_SYM) R = __double2float_rn(__dsub_rn((double)R, __dmul_rn(scale, __dadd_rn(vy, gy))));    // :458
  if (f & FB_SYP) R = __double2float_rn(__dadd_rn((double)R, __dmul_rn(scale, __dadd_rn(vjy, gy))));   // :462
  if (f & FB_SZM) R = __double2float_rn(__dsub_rn((double)R, __dmul_rn(scale, __dadd_rn(vz, gz))));    // :467
  if (f & FB_SZP) R = __double2float_rn(__dadd_rn((double)R, __dmul_rn(scale, __dadd_rn(vkz, gz))));   // :471
  double du = 0.0, dv = 0.0, dw = 0.0;
  // :592-595.  x / 1.0 == x exactly: the reference's dx = 1 skips the three fp64 divisions
  if (!(f & FB_SXP)) { du = __dsub_rn(vix, vx); if (dx != 1.0) du = __ddiv_rn(du, dx); }
  if (!(f & FB_SYP)) { dv = __dsub_rn(vjy, vy); if (dx != 1.0) dv = __ddiv_rn(dv, dx); }
  if (!(f & FB_SZP)) { dw = __dsub_rn(vkz, vz); if (dx != 1.0) dw = __ddiv_rn(dw, dx); }
  return __double2float_rn(__dsub_rn(__dsub_rn(__dsub_rn((double)R, du), dv), dw));   // :605
}

template <int MODE>
__global__ void __launch_bounds__(256, FB_MINB_RHS_DIV)
k_rhs_div(Grid g, Box box, int tiles_x, int tiles_y, int nitems, const uint8_t* __restrict__ fl,
          const double* __restrict__ u0, const double* __restrict__ u1, const double* __restrict__ u2,
          const float* __restrict__ bold, float* __restrict__ bout, double scale, double dx, double gx, double gy,
          double gz, DevScalars* __restrict__ ds) {
  __shared__ double sm[3 * 32];
  double acc[3] = {0.0, 0.0, 0.0};
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
    int tz = item / (tiles_x * tiles_y), rem = item - tz * tiles_x * tiles_y;
    int ty = rem / tiles_x, tx = rem - ty * tiles_x;
    int i = box.x0 + tx * 32 + threadIdx.x, j = box.y0 + ty * 8 + threadIdx.y, k = box.z0 + tz;
    if (i < box.x1 && j < box.y1) {
      int q = gidx(g, i, j, k);
      // every load is issued before the flag is looked at (one round trip instead of two); the +1
      // neighbours of a node on the last row / plane of the array are clamped (such a node is never fluid)
      const int qx = i + 1 < g.px ? q + 1 : q, qy = j + 1 < g.ny ? q + g.px : q, qz = k + 1 < g.nz ? q + g.plane : q;
      const uint8_t f = fl[q];
      const double vx = u0[q], vy = u1[q], vz = u2[q], vix = u0[qx], vjy = u1[qy], vkz = u2[qz];
      float D = 0.0f;
      if (f & FB_FLUID) D = rhs_div_cell(f, vx, vy, vz, vix, vjy, vkz, scale, dx, gx, gy, gz);
      if (MODE == 1) {
        double bo = (double)bold[q];
        double d = bo - (double)D;
        acc[1] += d * d; acc[2] += bo * bo;
      }
      acc[0] += (double)D * (double)D;
      bout[q] = D;
    }
  }
  block_sum<3>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0) {
    atomicAdd(&ds->bnorm2, acc[0]);
    if (MODE == 1) { atomicAdd(&ds->err_num, acc[1]); atomicAdd(&ds->err_den, acc[2]); }
  }
}

__global__ void k_reset_norms(DevScalars* ds) { ds->bnorm2 = 0.0; ds->err_num = 0.0; ds->err_den = 0.0; }

// mode 0: b of the first pass into bf[bcur].  mode 1: b2 into bf[bcur^1] + error sums; then flip.
int fb_rhs_div(flipb200_ctx* c, double dt, int mode) {
  const Grid& g = c->g;
  const Box& b = c->box;
  k_reset_norms<<<1, 1, 0, c->stream>>>(c->ds);
  c->launches++;
  if (b.x1 > b.x0 && b.y1 > b.y0 && b.z1 > b.z0) {
    int tiles_x = cdiv(b.x1 - b.x0, 32), tiles_y = cdiv(b.y1 - b.y0, 8);
    int nitems = tiles_x * tiles_y * (b.z1 - b.z0);
    int grid = min(nitems, c->num_sms * 8);
    dim3 blk(32, 8, 1);
    const double scale = 1.0 / c->P.dx;                                  // fluid.cc:416
    const double gx = c->P.gx * dt, gy = c->P.gy * dt, gz = c->P.gz * dt;   // :420 (full dt)
    if (mode == 0)
      k_rhs_div<0><<<grid, blk, 0, c->stream>>>(g, b, tiles_x, tiles_y, nitems, c->flags, c->u[0], c->u[1], c->u[2],
                                                nullptr, c->bf[c->bcur], scale, c->P.dx, gx, gy, gz, c->ds);
    else
      k_rhs_div<1><<<grid, blk, 0, c->stream>>>(g, b, tiles_x, tiles_y, nitems, c->flags, c->u[0], c->u[1], c->u[2],
                                                c->bf[c->bcur], c->bf[c->bcur ^ 1], scale, c->P.dx, gx, gy, gz, c->ds);
    CK_LAUNCH(); c->launches++;
  }
  if (mode == 1) c->bcur ^= 1;
  int rc = fb_allreduce(c, &c->ds->bnorm2, 3, FB_F64, FB_SUM);   // bnorm2, err_num, err_den are adjacent
  if (rc) return rc;
  return fb_fetch_scalars(c);
}

// ---------------------------------------------------------------------------------------------
// velUpdate (fluid.cc:612-703) as a per-face gather.  The reference scatters, in scan order:
//   loop 1, fluid cell c:  u[c] = u[c] - s*p_c + g'   (all 3 comps);  u_a[c+e_a] += s*p_c
//   loop 2, solid cell c:  u[c] = 0;  u_a[c+e_a] = 0
// so component a at node c receives, in this order, (+ s*p_m) from m = c-e_a if m is fluid, then
// (- s*p_c + g'_a) if c is fluid, and is zeroed if c or m is solid.  Same order, no FMA.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_vel_update(Grid g, Box box, int tiles_x, int tiles_y, int nitems, const uint8_t* __restrict__ fl,
             const double* __restrict__ p, double* __restrict__ u0, double* __restrict__ u1,
             double* __restrict__ u2, double scale, double gx, double gy, double gz) {
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
    int tz = item / (tiles_x * tiles_y), rem = item - tz * tiles_x * tiles_y;
    int ty = rem / tiles_x, tx = rem - ty * tiles_x;
    int i = box.x0 + tx * 32 + threadIdx.x, j = box.y0 + ty * 8 + threadIdx.y, k = box.z0 + tz;
    if (i >= box.x1 || j >= box.y1) continue;
    int q = gidx(g, i, j, k);
    uint8_t f = fl[q];
    if (f & FB_SOLID) { u0[q] = 0.0; u1[q] = 0.0; u2[q] = 0.0; continue; }   // :680-683
    // non-solid nodes are inside the shell, so q-1, q-px, q-plane are in the box
    double sp = 0.0;
    if (f & FB_FLUID) sp = __dmul_rn(scale, p[q]);                            // scale * pre, :637-641
    {
      double v = u0[q];
      if (f & FB_SXM) v = 0.0;                                               // :684-688 seen from c+e_x
      else {
        if (fl[q - 1] & FB_FLUID) v = __dadd_rn(v, __dmul_rn(scale, p[q - 1]));     // :646
        if (f & FB_FLUID) v = __dadd_rn(__dsub_rn(v, sp), gx);                      // :639
      }
      u0[q] = v;
    }
    {
      double v = u1[q];
      if (f & FB_SYM) v = 0.0;
      else {
        if (fl[q - g.px] & FB_FLUID) v = __dadd_rn(v, __dmul_rn(scale, p[q - g.px]));   // :653
        if (f & FB_FLUID) v = __dadd_rn(__dsub_rn(v, sp), gy);                          // :640
      }
      u1[q] = v;
    }
    {
      double v = u2[q];
      if (f & FB_SZM) v = 0.0;
      else {
        if (fl[q - g.plane] & FB_FLUID) v = __dadd_rn(v, __dmul_rn(scale, p[q - g.plane]));   // :660
        if (f & FB_FLUID) v = __dadd_rn(__dsub_rn(v, sp), gz);                                // :641
      }
      u2[q] = v;
    }
  }
}

int fb_vel_update(flipb200_ctx* c, double dt_sub) {
  const Grid& g = c->g;
  const Box& b = c->box;
  if (!(b.x1 > b.x0 && b.y1 > b.y0 && b.z1 > b.z0)) return 0;
  int rc = fb_halo_fill(c, g, c->x, sizeof(double), 1, 0);             // p of the cell below my first plane
  if (rc) return rc;
  int tiles_x = cdiv(b.x1 - b.x0, 32), tiles_y = cdiv(b.y1 - b.y0, 8);
  int nitems = tiles_x * tiles_y * (b.z1 - b.z0);
  int grid = min(nitems, c->num_sms * 8);
  const double scale = dt_sub / (c->P.rho * c->P.dx);                     // fluid.cc:614
  k_vel_update<<<grid, dim3(32, 8, 1), 0, c->stream>>>(g, b, tiles_x, tiles_y, nitems, c->flags, c->x, c->u[0],
                                                       c->u[1], c->u[2], scale, c->P.gx * dt_sub, c->P.gy * dt_sub,
                                                       c->P.gz * dt_sub);   // g = gravity*dt, :638
  CK_LAUNCH(); c->launches++;
  return fb_halo_fill(c, g, c->u[2], sizeof(double), 0, 1);
}

// ---------------------------------------------------------------------------------------------
// cell-centred FLIP delta for every node inside the shell (CatmullRomFLIP fluid.cc:237-240,252):
//   dc_a[c] = (u_a[c] + u_a[c+e_a])/2 - (ub_a[c] + ub_a[c+e_a])/2
// written into the (now idle) PCG vectors r, pd, Ap; zero outside the shell.
// PIC/FLIP blend (params.flip_ratio = r < 1; the reference runs r = 1): the particle update is
//   v_p = r (v_p + interp(dc)) + (1 - r) interp(vc),  vc = getVelocity(c, vels)  (clampedCatmullRom :125-207)
// and interp is linear with the same weights, so the grid carries  r dc + (1 - r) vc  instead.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_cell_delta(Grid g, const double* __restrict__ u0, const double* __restrict__ u1, const double* __restrict__ u2,
             const double* __restrict__ b0, const double* __restrict__ b1, const double* __restrict__ b2,
             double* __restrict__ d0, double* __restrict__ d1, double* __restrict__ d2, double ratio, Box R) {
  int i = R.x0 + blockIdx.x * blockDim.x + threadIdx.x;
  int j = R.y0 + blockIdx.y * blockDim.y + threadIdx.y;
  int k = R.z0 + blockIdx.z;                      // owned planes of the active region: every node a particle's support can touch
  if (i >= R.x1 || j >= R.y1) return;
  int q = gidx(g, i, j, k);
  double a = 0.0, b = 0.0, cc = 0.0;
  if (inside_shell(g, i, j, k)) {
    a = (u0[q] + u0[q + 1]) / 2.0 - (b0[q] + b0[q + 1]) / 2.0;
    b = (u1[q] + u1[q + g.px]) / 2.0 - (b1[q] + b1[q + g.px]) / 2.0;
    cc = (u2[q] + u2[q + g.plane]) / 2.0 - (b2[q] + b2[q + g.plane]) / 2.0;
    if (ratio != 1.0) {
      a = ratio * a + (1.0 - ratio) * ((u0[q] + u0[q + 1]) / 2.0);
      b = ratio * b + (1.0 - ratio) * ((u1[q] + u1[q + g.px]) / 2.0);
      cc = ratio * cc + (1.0 - ratio) * ((u2[q] + u2[q + g.plane]) / 2.0);
    }
  }
  d0[q] = a; d1[q] = b; d2[q] = cc;
}

int fb_cell_delta(flipb200_ctx* c) { return fb_cell_delta_r(c, c->P.flip_ratio); }

// ratio = 0 leaves the plain cell-centred velocity getVelocity(c, vels) in the arrays (PIC gather input)
int fb_cell_delta_r(flipb200_ctx* c, double ratio) {
  const Grid& g = c->g;
  Box R = fb_active_region(c, nullptr);             // after fb_classify: this step's wet box + margin, or the whole grid
  if (R.x1 == g.nx) R.x1 = g.px;                    // the pad columns as well, as the whole-grid sweep always did
  if (R.z1 > R.z0) {
    dim3 blk(32, 8, 1), grd(cdiv(R.x1 - R.x0, 32), cdiv(R.y1 - R.y0, 8), R.z1 - R.z0);
    k_cell_delta<<<grd, blk, 0, c->stream>>>(g, c->u[0], c->u[1], c->u[2], c->ub[0], c->ub[1], c->ub[2], c->r, c->pd,
                                             c->Ap, ratio, R);
  }
  CK_LAUNCH(); c->launches++;
  // particles of my top base plane read the deltas of plane k1
  double* dc[3] = {c->r, c->pd, c->Ap};
  for (int a = 0; a < 3; ++a) { int rc = fb_halo_fill(c, g, dc[a], sizeof(double), 0, 1); if (rc) return rc; }
  return 0;
}
