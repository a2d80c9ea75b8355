// pcg_p32.cu -- params.mixed bit 0 (default ON since round 2).
// The CG search direction p stored as fp32: A.p, p.Ap and x += alpha p are still evaluated in fp64 FROM THE
// STORED p, so x and r stay consistent (r remains the residual of x to fp64 rounding) and only the
// conjugacy of the directions is perturbed, by 6e-8 relative.  Prototype (tools/mg_proto7.py scene, full
// tank 128^3): 21 iterations and the same solution to 2.5e-16 with p in fp32 or fp64.  Ap, r, x stay
// fp64 (an fp32 Ap would cap the attainable residual at 1e-7).  Saves 12 of the ~54 B/node the CG part of
// an iteration moves: k_spmv 17 -> 13 B/node, direction update 37 -> 29.
// The kernels are the four-node shapes of pcg.cu (k_spmv4<false>, k_update_p4, k_pcg_finish_x) with the
// loads / stores of p swapped; p lives in the first half of the ctx's pd array.  Own translation unit so
// that the SASS of the measured default kernels stays identical (tools/sass_diff.py).  Single GPU.
#include "sweep.cuh"
#include "pcg.cuh"
#include "halo.cuh"

// K1: Ap = A p (fp64) from the fp32 p; neighbour sums in fp64 (the stored values are exact in fp64)
template <bool PEER>
__global__ void __launch_bounds__(32 * ST_BY, FB_MINB_SPMV4F(PEER))
k_spmv4f(const __grid_constant__ Sweep S, const __grid_constant__ PcgCoef cf, const uint8_t* __restrict__ fl,
         const float* __restrict__ p, double* __restrict__ Ap, DevScalars* __restrict__ ds, int it_arg, int noexit,
         int dist, const __grid_constant__ PeerHalo H) {
  __shared__ double sm[32];
  if (!noexit && ds->pcg_done) return;
  const int it = pcg_iter(ds, it_arg);
  if (PEER) {                                          // z-slabs: boundary blocks wait for the neighbours' planes of p
    const int kz0 = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * ST_ZP, kz1 = min(kz0 + ST_ZP, S.box.z1);
    halo_wait(H, ds, kz0 <= H.kdn && H.kdn < kz1, kz0 <= H.kup && H.kup < kz1);
  }
  double acc[1] = {0.0};
  StencilMarch4 m;
  if (m.begin(S, p)) {
    const int kend = min(m.k + ST_ZP, S.box.z1);
#pragma unroll
    for (int z = 0; z < ST_ZP; ++z) {
      if (m.k >= kend) break;
      m.load(S, p);
      const uint32_t f = m.act ? *reinterpret_cast<const uint32_t*>(fl + m.q) : 0u;
      const uint32_t f0 = f & 0xffu, f1 = (f >> 8) & 0xffu, f2 = (f >> 16) & 0xffu, f3 = f >> 24;
      const double c0 = m.pc.x, c1 = m.pc.y, c2 = m.pc.z, c3 = m.pc.w;
      const double s0 = ((double)m.pm.x + (double)m.pn.x) + ((double)m.ym.x + (double)m.yp.x) + ((double)m.left + c1);
      const double s1 = ((double)m.pm.y + (double)m.pn.y) + ((double)m.ym.y + (double)m.yp.y) + (c0 + c2);
      const double s2 = ((double)m.pm.z + (double)m.pn.z) + ((double)m.ym.z + (double)m.yp.z) + (c1 + c3);
      const double s3 = ((double)m.pm.w + (double)m.pn.w) + ((double)m.ym.w + (double)m.yp.w) + (c2 + (double)m.right);
      double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
      if (f0 & FB_FLUID) { a0 = cf.diag[6 - __popc(f0 & FB_SMASK)] * c0 + cf.off * s0; acc[0] += c0 * a0; }
      if (f1 & FB_FLUID) { a1 = cf.diag[6 - __popc(f1 & FB_SMASK)] * c1 + cf.off * s1; acc[0] += c1 * a1; }
      if (f2 & FB_FLUID) { a2 = cf.diag[6 - __popc(f2 & FB_SMASK)] * c2 + cf.off * s2; acc[0] += c2 * a2; }
      if (f3 & FB_FLUID) { a3 = cf.diag[6 - __popc(f3 & FB_SMASK)] * c3 + cf.off * s3; acc[0] += c3 * a3; }
      if ((f0 | f1) & FB_FLUID) st2<double>(Ap + m.q, a0, a1);
      if ((f2 | f3) & FB_FLUID) st2<double>(Ap + m.q + 2, a2, a3);
      m.advance(S);
    }
  }
  block_sum<1>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0 && acc[0] != 0.0) atomicAdd(dist ? &ds->part[PART_PAP] : &ds->pAp[it & 3], acc[0]);
}

// K3 (MG path): x += alpha p_old, p = float(z + beta p_old); four nodes per lane (k_update_p4 of pcg.cu)
__global__ void __launch_bounds__(32 * SW_BY, FB_MINB_UPDATE_P4F)
k_update_p4f(const __grid_constant__ Sweep S, const uint8_t* __restrict__ fl, const float* __restrict__ zf,
             float* __restrict__ p, double* __restrict__ x, DevScalars* __restrict__ ds, int it_arg, int noexit,
             double inv_bscale /* 1, or 1 / bscale with params.mixed bit 1 (pcg.cuh) */, const __grid_constant__ PeerHalo H) {
  if (!noexit && ds->pcg_done) return;
  const int it = pcg_iter(ds, it_arg);
  const bool first = it < 0;
  const double beta = first ? 0.0 : ds->rz[it & 3] / ds->rz[(it + 3) & 3];
  const double rrp = ds->rr[(it + 3) & 3];
  const double invsc = (rrp > 0.0 ? sqrt(rrp) : 1.0) * inv_bscale;
  const double alpha = !first ? ds->rz[(it + 3) & 3] / ds->pAp[it & 3] : 0.0;   // as in k_update_xr4 of this iteration
  const int i = S.box.x0 + blockIdx.x * 128 + 4 * threadIdx.x;
  const int j = S.box.y0 + blockIdx.y * SW_BY + threadIdx.y;
  const int k = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp);
  if (i < S.xe && j < S.box.y1) {
    const int q = gidx(S.g, i, j, k);
    const uint32_t f = *reinterpret_cast<const uint32_t*>(fl + q);
    const float4 pv = first ? make_float4(0.f, 0.f, 0.f, 0.f) : *reinterpret_cast<const float4*>(p + q);   // the array holds no floats yet
    const float4 zv = *reinterpret_cast<const float4*>(zf + q);
    const bool f0 = f & FB_FLUID, f1 = (f >> 8) & FB_FLUID, f2 = (f >> 16) & FB_FLUID, f3 = (f >> 24) & FB_FLUID;
    if (!first) {                                      // x += alpha p with the old direction (ConjugateGradient.h:73)
      double2 xa = ld2<double>(x + q), xb = ld2<double>(x + q + 2);
      if (f0) xa.x += alpha * (double)pv.x;
      if (f1) xa.y += alpha * (double)pv.y;
      if (f2) xb.x += alpha * (double)pv.z;
      if (f3) xb.y += alpha * (double)pv.w;
      if (f0 | f1) st2<double>(x + q, xa.x, xa.y);
      if (f2 | f3) st2<double>(x + q + 2, xb.x, xb.y);
    }
    float4 o;
    o.x = f0 ? (float)((double)zv.x * invsc + beta * (double)pv.x) : 0.f;
    o.y = f1 ? (float)((double)zv.y * invsc + beta * (double)pv.y) : 0.f;
    o.z = f2 ? (float)((double)zv.z * invsc + beta * (double)pv.z) : 0.f;
    o.w = f3 ? (float)((double)zv.w * invsc + beta * (double)pv.w) : 0.f;
    *reinterpret_cast<float4*>(p + q) = o;             // every swept node: the stencil reads 0 off the fluid cells
    halo_push4(H, q, k, o);                            // z-slabs: the neighbours' ghost planes of p (fp32)
  }
  halo_signal(H, ds, k == H.kdn, k == H.kup);
}

// the stop test ended the solve before the direction update applied x += alpha p (k_pcg_finish_x of pcg.cu)
__global__ void __launch_bounds__(32 * SW_BY)
k_pcg_finish_xf(const __grid_constant__ Sweep S, const uint8_t* __restrict__ fl, double* __restrict__ x,
                const float* __restrict__ p, const DevScalars* __restrict__ ds) {
  if (!ds->x_pending) return;
  const int it = ds->pcg_iters;
  const double alpha = ds->rz[(it + 3) & 3] / ds->pAp[it & 3];
  STREAM_PROLOGUE(S)
  if (act) {
    uchar2 f = *reinterpret_cast<const uchar2*>(fl + q);
    double2 xv = ld2<double>(x + q);
    const float2 pv = ld2<float>(p + q);
    if (f.x & FB_FLUID) xv.x += alpha * (double)pv.x;
    if (f.y & FB_FLUID) xv.y += alpha * (double)pv.y;
    if ((f.x | f.y) & FB_FLUID) st2<double>(x + q, xv.x, xv.y);
  }
}

void fb_p32_spmv(flipb200_ctx* c, const PcgCoef& cf, int it_arg, int noexit, int dist, const PeerHalo& H) {
  const Sweep S4 = make_sweep(c->g, c->box, 4);
  if (H.mode) k_spmv4f<true><<<stencil_grid(S4, 4), STENCIL_BLOCK, 0, c->stream>>>(S4, cf, c->flags, reinterpret_cast<const float*>(c->pd), c->Ap, c->ds, it_arg, noexit, dist, H);
  else k_spmv4f<false><<<stencil_grid(S4, 4), STENCIL_BLOCK, 0, c->stream>>>(S4, cf, c->flags, reinterpret_cast<const float*>(c->pd), c->Ap, c->ds, it_arg, noexit, dist, H);
}
void fb_p32_update_p(flipb200_ctx* c, const float* zf, int it_arg, int noexit, double inv_bscale, bool down, const PeerHalo& H) {
  const Sweep S4 = make_sweep(c->g, c->box, 4);
  const dim3 g(cdiv(S4.xe - S4.box.x0, 128), cdiv(S4.box.y1 - S4.box.y0, SW_BY), S4.box.z1 - S4.box.z0);
  k_update_p4f<<<g, STREAM_BLOCK, 0, c->stream>>>(down ? sweep_down(S4) : S4, c->flags, zf, reinterpret_cast<float*>(c->pd), c->x, c->ds, it_arg, noexit, inv_bscale, H);
}
void fb_p32_finish_x(flipb200_ctx* c) {
  const Sweep S = make_sweep(c->g, c->box);
  k_pcg_finish_xf<<<stream_grid(S), STREAM_BLOCK, 0, c->stream>>>(S, c->flags, c->x, reinterpret_cast<const float*>(c->pd), c->ds);
}
