// pcg.cu -- matrix-free 7-point preconditioned conjugate gradient on the fluid cells.
//   replaces: setA fluid.cc:304-412, setA2 :481-541 (never materialised), cg.compute(A) :1473,
//             p = cg.solve(b) :1474 (Eigen ConjugateGradient.h:26-90 loop, same stop rule).
// Matrix (from setA): for a fluid cell c with k non-solid neighbours, A_cc = Adiag = k
// fp32-accumulated additions of scale = dt/(rho dx^2) (fluid.cc:333-370), A_cn = float(-scale) for
// every fluid neighbour n (:335); air neighbours are Dirichlet 0, solid neighbours Neumann.
// All vectors hold 0 on non-fluid nodes, so A.p = diag[k]*p_c + off * (sum of the 6 neighbours).
//
// Preconditioner M: Jacobi (params.precond = 0) or a multigrid V-cycle in fp32 (precond = 1, mg.cu).
// One iteration (algorithmic bytes per swept node, fp64 vectors, 1-byte flags):
//   K1 k_spmv      Ap = A p,  pAp += p.Ap                      read p 8, flag 1; write Ap 8        = 17
//   K2 k_update_xr x += a p, r -= a Ap, rr += r.r (+ Jacobi: rz += r.M^-1 r;
//                  + MG: b0 = float(r/|r_prev|), x1 = w D^-1 b0)  read x,p,r,Ap 32, flag 1; write x,r 16 = 49 (+8 MG)
//   [MG: V-cycle, ~74 B/node on the finest level in fp32, r.z fused into its last sweep]
//   K3 k_update_p  p = z + beta p  (Jacobi: z = M^-1 r on the fly; stop test)  read r|z,p, flag; write p = 25 (21 MG)
// Scalars never visit the host inside the loop: dot products land in a 4-slot ring in device
// memory (slot = iteration & 3), alpha/beta are formed by the consuming kernel, convergence sets
// ds->pcg_done and the remaining launches of a batch return immediately.
// Kernel shapes: sweep.cuh (STREAM for K2/K3/init, STENCIL for K1).
#include "sweep.cuh"
#include "pcg.cuh"
#include "halo.cuh"
#include <stdlib.h>

__global__ void k_pcg_reset(DevScalars* ds, double bnorm2, double threshold, int max_iter, int halo_base) {
  ds->threshold = threshold; ds->pcg_max_iter = max_iter; ds->halo_base = halo_base;
  for (int s = 0; s < 4; ++s) { ds->pAp[s] = 0.0; ds->rz[s] = 0.0; ds->rr[s] = 0.0; }
  // |r|^2 of "iteration -1" (= |b|^2): scale of the first preconditioner applications
  ds->rr[2] = bnorm2; ds->rr[3] = bnorm2;
  ds->pcg_done = 0; ds->pcg_iters = 0; ds->pcg_status = 0; ds->rr_final = 0.0; ds->pcg_it = 0; ds->x_pending = 0;
  for (int s = 0; s < 4; ++s) ds->part[s] = 0.0;
}

// x = 0, r = b;  Jacobi: p = M^-1 r, rz[3] = r.p   (Eigen ConjugateGradient.h:43,63,66 with x0 = 0)
//               MG:     b0 = float(r * sc), x1 = w D^-1 b0  (input of the first V-cycle)
template <int MODE, typename TB = float>   // TB: storage type of the V-cycle's right-hand side b0 (fp16: FLIPB200_MG_H16, pcg.cuh)
__global__ void __launch_bounds__(32 * SW_BY)
k_pcg_init(const __grid_constant__ Sweep S, const __grid_constant__ PcgCoef cf, const uint8_t* __restrict__ fl,
           const float* __restrict__ b, double* __restrict__ x, double* __restrict__ r, double* __restrict__ p,
           TB* __restrict__ b0, float* __restrict__ x1, double sc, DevScalars* __restrict__ ds, int dist,
           const __grid_constant__ PeerHalo H) {
  __shared__ double sm[32];
  STREAM_PROLOGUE(S)
  double acc[1] = {0.0};
  if (act) {
    uchar2 f = *reinterpret_cast<const uchar2*>(fl + q);
    float2 bv = *reinterpret_cast<const float2*>(b + q);
    double r0 = (f.x & FB_FLUID) ? (double)bv.x : 0.0, r1 = (f.y & FB_FLUID) ? (double)bv.y : 0.0;
    int k0 = 6 - __popc(f.x & FB_SMASK), k1 = 6 - __popc(f.y & FB_SMASK);
    st2<double>(x + q, 0.0, 0.0);
    st2<double>(r + q, r0, r1);
    if (MODE == 0) {
      double p0 = r0 * cf.inv[k0], p1 = r1 * cf.inv[k1];
      st2<double>(p + q, p0, p1);
      acc[0] = r0 * p0 + r1 * p1;
    } else {
      float s0 = (float)(r0 * sc), s1 = (float)(r1 * sc);
      stq2<TB>(b0 + q, s0, s1);
      st2<float>(x1 + q, cf.winv[k0] * s0, cf.winv[k1] * s1);
      halo_push2<float>(H, q, k, cf.winv[k0] * s0, cf.winv[k1] * s1);
      st2<double>(p + q, 0.0, 0.0);
    }
  }
  if (MODE == 0) {
    block_sum<1>(acc, sm);
    if (threadIdx.x == 0 && threadIdx.y == 0 && acc[0] != 0.0) atomicAdd(dist ? &ds->part[PART_RZ] : &ds->rz[3], acc[0]);
  }
  if (MODE == 1) halo_signal(H, ds, k == H.kdn, k == H.kup);
}

// ---------------------------------------------------------------------------------------------
// Warm start (params.warm_start; Eigen's cg.solve() starts from x = 0, ConjugateGradient.h:233-237, the stop
// rule |r|^2 < tol^2 |b|^2 is unchanged).  The guess g is the previous solution, which k_classify keeps only
// on cells that are still fluid; over the relaxation passes of one step the right-hand sides obey
// b_{k+1} = th b_k + G with th = 1 - 1/relax_div (fluid.cc:1475: velUpdate applies a tenth of the gradient), so
// from pass 3 on g = x_k + th (x_k - x_{k-1}) (k_warm_extrap).  x0 = gamma g with the residual-minimising
// gamma = (b . A g) / (A g . A g); |r0|^2 is summed exactly in a 12 B/node pass before r0 is stored, so that the
// fp32 / fp16 V-cycle input gets its O(1) scale in the pass that writes it.
//   k_warm_extrap   g = x + th (x - xprev), xprev = x                         (passes >= 3)   32 B/node
//   k_warm_ax       Ap = A g, (b . Ap), (Ap . Ap)                                             21 B/node
//   k_warm_scalars  gamma;  k_warm_rr  |b - gamma Ap|^2 exactly (12 B/node) -> ring slots 2, 3 (scale of the first V-cycle, stop test)
//   k_warm_init     x = gamma g, r = b - gamma Ap; MG: b0, x1; Jacobi: p, r.z                          45 B/node
// ---------------------------------------------------------------------------------------------
// guess of pass k of a step (th = 1 - 1/relax_div); all guesses are masked to the current fluid cells
//   MODE 0 (first pass):   if pendG: xG = x - th xprev (the last step ended after two passes);  x = src (x itself or xfirst)
//   MODE 1 (second pass):  xprev = xfirst = x (= x_1);  x = th x_1 + xG
//   MODE 2 (third pass):   xG = x - th xprev;  g = x + th (x - xprev);  xprev = x
//   MODE 3 (later passes): g = x + th (x - xprev);  xprev = x
template <int MODE>
__global__ void __launch_bounds__(32 * SW_BY)
k_warm_guess(const __grid_constant__ Sweep S, const uint8_t* __restrict__ fl, double* __restrict__ x, double* __restrict__ xprev,
             double* __restrict__ xfirst, double* __restrict__ xG, double theta, int pendG, int src_first, int have_G) {
  STREAM_PROLOGUE(S)
  if (!act) return;
  const uchar2 f = *reinterpret_cast<const uchar2*>(fl + q);
  const bool f0 = f.x & FB_FLUID, f1 = f.y & FB_FLUID;
  const double2 xv = ld2<double>(x + q);
  if (MODE == 0) {
    if (pendG) { const double2 pv = ld2<double>(xprev + q); st2<double>(xG + q, xv.x - theta * pv.x, xv.y - theta * pv.y); }
    if (src_first) { const double2 sv = ld2<double>(xfirst + q); st2<double>(x + q, f0 ? sv.x : 0.0, f1 ? sv.y : 0.0); }
  } else if (MODE == 1) {
    st2<double>(xprev + q, xv.x, xv.y);
    st2<double>(xfirst + q, xv.x, xv.y);
    double g0 = theta * xv.x, g1 = theta * xv.y;
    if (have_G) { const double2 gv = ld2<double>(xG + q); g0 += gv.x; g1 += gv.y; }
    st2<double>(x + q, f0 ? g0 : 0.0, f1 ? g1 : 0.0);
  } else {
    const double2 pv = ld2<double>(xprev + q);
    if (MODE == 2) st2<double>(xG + q, xv.x - theta * pv.x, xv.y - theta * pv.y);
    st2<double>(x + q, f0 ? xv.x + theta * (xv.x - pv.x) : 0.0, f1 ? xv.y + theta * (xv.y - pv.y) : 0.0);
    st2<double>(xprev + q, xv.x, xv.y);
  }
}

__global__ void __launch_bounds__(32 * ST_BY)
k_warm_ax(const __grid_constant__ Sweep S, const __grid_constant__ PcgCoef cf, const uint8_t* __restrict__ fl,
          const float* __restrict__ b, const double* __restrict__ x, double* __restrict__ Ap, DevScalars* __restrict__ ds) {
  __shared__ double sm[64];
  double acc[2] = {0.0, 0.0};
  StencilMarch<double> m;
  if (m.begin(S, x)) {
    const int kend = min(m.k + S.zc, S.box.z1);
#pragma unroll
    for (int z = 0; z < ST_ZP; ++z) {
      if (m.k >= kend) break;
      m.load(S, x);
      uchar2 f = m.act ? *reinterpret_cast<const uchar2*>(fl + m.q) : make_uchar2(0, 0);
      const float2 bv = m.act ? ld2<float>(b + m.q) : make_float2(0.f, 0.f);
      double a0 = 0.0, a1 = 0.0;
      if (f.x & FB_FLUID) { a0 = cf.diag[6 - __popc(f.x & FB_SMASK)] * m.c0() + cf.off * m.sum0(); acc[0] += (double)bv.x * a0; acc[1] += a0 * a0; }
      if (f.y & FB_FLUID) { a1 = cf.diag[6 - __popc(f.y & FB_SMASK)] * m.c1() + cf.off * m.sum1(); acc[0] += (double)bv.y * a1; acc[1] += a1 * a1; }
      if (m.act) st2<double>(Ap + m.q, a0, a1);
      m.advance(S);
    }
  }
  block_sum<2>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0) {
    if (acc[0] != 0.0) atomicAdd(&ds->part[PART_PAP], acc[0]);
    if (acc[1] != 0.0) atomicAdd(&ds->part[PART_RZ], acc[1]);
  }
}

__global__ void k_warm_scalars(DevScalars* ds) {
  const double bAg = ds->part[PART_PAP], AgAg = ds->part[PART_RZ];
  ds->part[PART_PAP] = 0.0; ds->part[PART_RZ] = 0.0; ds->part[PART_RR] = 0.0;
  double gamma = AgAg > 0.0 ? bAg / AgAg : 0.0;
  if (!(gamma == gamma) || !(fabs(gamma) < 1e300)) gamma = 0.0;
  ds->warm_gamma = gamma;
}
// |b - gamma A g|^2, exactly (12 B/node).  The closed form |b|^2 - gamma (b . A g) is pure cancellation once the guess is good
// (|r0| / |b| ~ 1e-8 on the extrapolated passes): as the SCALE of the fp16-stored V-cycle input it overflowed fp16 in round 2.
__global__ void __launch_bounds__(32 * SW_BY)
k_warm_rr(const __grid_constant__ Sweep S, const uint8_t* __restrict__ fl, const float* __restrict__ b, const double* __restrict__ Ap,
          DevScalars* __restrict__ ds) {
  __shared__ double sm[32];
  const double gamma = ds->warm_gamma;
  STREAM_PROLOGUE(S)
  double acc[1] = {0.0};
  if (act) {
    const uchar2 f = *reinterpret_cast<const uchar2*>(fl + q);
    const float2 bv = ld2<float>(b + q);
    const double2 av = ld2<double>(Ap + q);
    const double r0 = (f.x & FB_FLUID) ? (double)bv.x - gamma * av.x : 0.0, r1 = (f.y & FB_FLUID) ? (double)bv.y - gamma * av.y : 0.0;
    acc[0] = r0 * r0 + r1 * r1;
  }
  block_sum<1>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0 && acc[0] != 0.0) atomicAdd(&ds->part[PART_RR], acc[0]);
}
// after the (all-reduced) sum: ring slots 2 and 3 = |r0|^2 ("iteration -1": scale of the first V-cycle, stop test of the guess)
__global__ void k_warm_scalars2(DevScalars* ds) {
  const double rr = ds->part[PART_RR];
  ds->part[PART_RR] = 0.0;
  ds->rr[2] = rr; ds->rr[3] = rr;
}

template <int MODE, typename TB = float>
__global__ void __launch_bounds__(32 * SW_BY)
k_warm_init(const __grid_constant__ Sweep S, const __grid_constant__ PcgCoef cf, const uint8_t* __restrict__ fl,
            const float* __restrict__ b, double* __restrict__ x, double* __restrict__ r, double* __restrict__ p,
            const double* __restrict__ Ap, TB* __restrict__ b0, float* __restrict__ x1, double bscale,
            DevScalars* __restrict__ ds, int dist, const __grid_constant__ PeerHalo H) {
  __shared__ double sm[64];
  const double gamma = ds->warm_gamma;
  const double rr_est = ds->rr[2];
  const double sc = (rr_est > 0.0 ? rsqrt(rr_est) : 1.0) * bscale;
  STREAM_PROLOGUE(S)
  double acc[2] = {0.0, 0.0};
  if (act) {
    uchar2 f = *reinterpret_cast<const uchar2*>(fl + q);
    const float2 bv = ld2<float>(b + q);
    const double2 xv = ld2<double>(x + q), av = ld2<double>(Ap + q);
    const bool f0 = f.x & FB_FLUID, f1 = f.y & FB_FLUID;
    const double r0 = f0 ? (double)bv.x - gamma * av.x : 0.0, r1 = f1 ? (double)bv.y - gamma * av.y : 0.0;
    st2<double>(x + q, f0 ? gamma * xv.x : 0.0, f1 ? gamma * xv.y : 0.0);
    st2<double>(r + q, r0, r1);
    acc[0] = r0 * r0 + r1 * r1;
    const int k0 = 6 - __popc(f.x & FB_SMASK), k1 = 6 - __popc(f.y & FB_SMASK);
    if (MODE == 0) {
      const double p0 = r0 * cf.inv[k0], p1 = r1 * cf.inv[k1];
      st2<double>(p + q, p0, p1);
      acc[1] = r0 * p0 + r1 * p1;
    } else {
      const float s0 = (float)(r0 * sc), s1 = (float)(r1 * sc);
      stq2<TB>(b0 + q, s0, s1);
      st2<float>(x1 + q, cf.winv[k0] * s0, cf.winv[k1] * s1);
      halo_push2<float>(H, q, k, cf.winv[k0] * s0, cf.winv[k1] * s1);
      st2<double>(p + q, 0.0, 0.0);
    }
  }
  block_sum<2>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0) {
    if (MODE == 0 && acc[1] != 0.0) atomicAdd(dist ? &ds->part[PART_RZ] : &ds->rz[3], acc[1]);
  }
  if (MODE == 1) halo_signal(H, ds, k == H.kdn, k == H.kup);
}

// stop test on the warm start's own residual (0 iterations) and the scale of iteration 0
__global__ void k_warm_check(DevScalars* ds) {
  const double rr = ds->rr[3];
  ds->warm_rr0 = rr;
  if (rr < ds->threshold) { ds->pcg_done = 1; ds->pcg_iters = 0; ds->rr_final = rr; ds->x_pending = 0; }
}

// K1: Ap = A p and p.Ap (STENCIL shape: z-march in registers, x neighbours by shuffle)
__global__ void __launch_bounds__(32 * ST_BY)
k_spmv(const __grid_constant__ Sweep S, const __grid_constant__ PcgCoef cf, const uint8_t* __restrict__ fl,
       const double* __restrict__ p, double* __restrict__ Ap, DevScalars* __restrict__ ds, int it_arg, int noexit,
       int dist, const __grid_constant__ PeerHalo H) {
  __shared__ double sm[32];
  if (!noexit && ds->pcg_done) return;
  const int it = pcg_iter(ds, it_arg);
  {
    const int kz0 = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * S.zc, kz1 = min(kz0 + S.zc, S.box.z1);
    halo_wait(H, ds, kz0 <= H.kdn && H.kdn < kz1, kz0 <= H.kup && H.kup < kz1);
  }
  double acc[1] = {0.0};
  StencilMarch<double> m;
  if (m.begin(S, p)) {
    const int kend = min(m.k + S.zc, S.box.z1);
#pragma unroll
    for (int z = 0; z < ST_ZP; ++z) {
      if (m.k >= kend) break;
      m.load(S, p);
      uchar2 f = m.act ? *reinterpret_cast<const uchar2*>(fl + m.q) : make_uchar2(0, 0);
      double a0 = 0.0, a1 = 0.0;
      if (f.x & FB_FLUID) { a0 = cf.diag[6 - __popc(f.x & FB_SMASK)] * m.c0() + cf.off * m.sum0(); acc[0] += m.c0() * a0; }
      if (f.y & FB_FLUID) { a1 = cf.diag[6 - __popc(f.y & FB_SMASK)] * m.c1() + cf.off * m.sum1(); acc[0] += m.c1() * a1; }
      if ((f.x | f.y) & FB_FLUID) st2<double>(Ap + m.q, a0, a1);
      m.advance(S);
    }
  }
  block_sum<1>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0 && acc[0] != 0.0) atomicAdd(dist ? &ds->part[PART_PAP] : &ds->pAp[it & 3], acc[0]);
}

// K1 in the STENCIL4 shape (four doubles per lane); used when the sweep's rows are >= 128 nodes wide
// PEER = false (single GPU): the ghost-plane wait of z-slab runs is not compiled in -- its mere
// presence costs 5 us of 65 at 256^3 (tools/kbench.cu T9, FL bit 3).
template <bool PEER>
__global__ void __launch_bounds__(32 * ST_BY)
k_spmv4(const __grid_constant__ Sweep S, const __grid_constant__ PcgCoef cf, const uint8_t* __restrict__ fl,
        const double* __restrict__ p, double* __restrict__ Ap, DevScalars* __restrict__ ds, int it_arg, int noexit,
        int dist, const __grid_constant__ PeerHalo H) {
  __shared__ double sm[32];
  if (!noexit && ds->pcg_done) return;
  const int it = pcg_iter(ds, it_arg);
  if (PEER) {
    const int kz0 = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * ST_ZP, kz1 = min(kz0 + ST_ZP, S.box.z1);
    halo_wait(H, ds, kz0 <= H.kdn && H.kdn < kz1, kz0 <= H.kup && H.kup < kz1);
  }
  double acc[1] = {0.0};
  StencilMarch4D m;
  if (m.begin(S, p)) {
    const int kend = min(m.k + ST_ZP, S.box.z1);
#pragma unroll
    for (int z = 0; z < ST_ZP; ++z) {
      if (m.k >= kend) break;      // predicating the planes of a partial chunk instead measured slower (kbench T9, FL bit 0)
      m.load(S, p);
      const uint32_t f = m.act ? *reinterpret_cast<const uint32_t*>(fl + m.q) : 0u;
      const uint32_t f0 = f & 0xffu, f1 = (f >> 8) & 0xffu, f2 = (f >> 16) & 0xffu, f3 = f >> 24;
      double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
      if (f0 & FB_FLUID) { a0 = cf.diag[6 - __popc(f0 & FB_SMASK)] * m.pca.x + cf.off * m.sum0(); acc[0] += m.pca.x * a0; }
      if (f1 & FB_FLUID) { a1 = cf.diag[6 - __popc(f1 & FB_SMASK)] * m.pca.y + cf.off * m.sum1(); acc[0] += m.pca.y * a1; }
      if (f2 & FB_FLUID) { a2 = cf.diag[6 - __popc(f2 & FB_SMASK)] * m.pcb.x + cf.off * m.sum2(); acc[0] += m.pcb.x * a2; }
      if (f3 & FB_FLUID) { a3 = cf.diag[6 - __popc(f3 & FB_SMASK)] * m.pcb.y + cf.off * m.sum3(); acc[0] += m.pcb.y * a3; }
      if ((f0 | f1) & FB_FLUID) st2<double>(Ap + m.q, a0, a1);
      if ((f2 | f3) & FB_FLUID) st2<double>(Ap + m.q + 2, a2, a3);
      m.advance(S);
    }
  }
  block_sum<1>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0 && acc[0] != 0.0) atomicAdd(dist ? &ds->part[PART_PAP] : &ds->pAp[it & 3], acc[0]);
}

// launch K1 in the shape that fits the sweep (FLIPB200_ST4=0: always two nodes per lane)
static void launch_spmv(flipb200_ctx* c, const PcgCoef& cf, int it_arg, int noexit, int dist, const PeerHalo& H, bool p32 = false) {
  if (p32) { fb_p32_spmv(c, cf, it_arg, noexit, dist, H); return; }       // fp32 search direction (pcg_p32.cu)
  const bool st4_on = c->opt_st4;
  const Sweep S4 = make_sweep(c->g, c->box, 4);
  if (st4_on && S4.xe - S4.box.x0 >= 128) {
    if (H.mode) k_spmv4<true><<<stencil_grid(S4, 4), STENCIL_BLOCK, 0, c->stream>>>(S4, cf, c->flags, c->pd, c->Ap, c->ds, it_arg, noexit, dist, H);
    else k_spmv4<false><<<stencil_grid(S4, 4), STENCIL_BLOCK, 0, c->stream>>>(S4, cf, c->flags, c->pd, c->Ap, c->ds, it_arg, noexit, dist, H);
  } else {
    const Sweep S = make_sweep(c->g, c->box);
    k_spmv<<<stencil_grid(S), STENCIL_BLOCK, 0, c->stream>>>(S, cf, c->flags, c->pd, c->Ap, c->ds, it_arg, noexit, dist, H);
  }
}

// K2: alpha = rz_old / pAp;  x += alpha p;  r -= alpha Ap;  rr = |r|^2
//     MODE 0 (Jacobi): rz = r.M^-1 r.   MODE 1 (MG): b0 = float(r*sc), x1 = w D^-1 b0 with
//     sc = 1/|r_prev| so that the fp32 V-cycle works on O(1) numbers whatever |r| has become;
//     x += alpha p is left to k_update_p<1> of the same iteration, which reads p anyway (8 B/node
//     less); when the stop test ends the solve in between, k_pcg_finish_x applies it.
template <int MODE>
__global__ void __launch_bounds__(32 * SW_BY)
k_update_xr(const __grid_constant__ Sweep S, const __grid_constant__ PcgCoef cf, const uint8_t* __restrict__ fl,
            double* __restrict__ x, double* __restrict__ r, const double* __restrict__ p,
            const double* __restrict__ Ap, float* __restrict__ b0, float* __restrict__ x1,
            DevScalars* __restrict__ ds, int it_arg, int noexit, int dist, const __grid_constant__ PeerHalo H) {
  __shared__ double sm[64];
  if (!noexit && ds->pcg_done) return;
  const int it = pcg_iter(ds, it_arg);
  const double pAp = ds->pAp[it & 3];
  const double rz_old = ds->rz[(it + 3) & 3];
  if (!noexit && !(pAp > 0.0)) {   // breakdown (also catches NaN)
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0 && threadIdx.y == 0) {
      ds->pcg_status = FLIPB200_ERR_BREAKDOWN; ds->pcg_done = 1; ds->pcg_iters = it;
      ds->rr_final = ds->rr[(it + 3) & 3];
    }
    return;
  }
  const double alpha = rz_old / pAp;                   // ConjugateGradient.h:72
  const double rrp = ds->rr[(it + 3) & 3];
  const double sc = MODE == 1 ? (rrp > 0.0 ? rsqrt(rrp) : 1.0) : 0.0;
  STREAM_PROLOGUE(S)
  double acc[2] = {0.0, 0.0};
  if (act) {
    uchar2 f = *reinterpret_cast<const uchar2*>(fl + q);
    double2 rv = ld2<double>(r + q), av = ld2<double>(Ap + q);
    double2 xv = zero2<double>(), pv = zero2<double>();
    if (MODE == 0) { xv = ld2<double>(x + q); pv = ld2<double>(p + q); }
    const bool f0 = f.x & FB_FLUID, f1 = f.y & FB_FLUID;
    int k0 = 6 - __popc(f.x & FB_SMASK), k1 = 6 - __popc(f.y & FB_SMASK);
    double r0 = 0.0, r1 = 0.0;
    if (f0) { xv.x += alpha * pv.x; r0 = rv.x - alpha * av.x; }      // :73-74
    if (f1) { xv.y += alpha * pv.y; r1 = rv.y - alpha * av.y; }
    if (f0 | f1) { if (MODE == 0) st2<double>(x + q, xv.x, xv.y); st2<double>(r + q, r0, r1); }
    acc[0] = r0 * r0 + r1 * r1;                                        // :76
    if (MODE == 0) acc[1] = r0 * (r0 * cf.inv[k0]) + r1 * (r1 * cf.inv[k1]);   // :80,83
    else {
      float s0 = (float)(r0 * sc), s1 = (float)(r1 * sc);
      st2<float>(b0 + q, s0, s1);
      st2<float>(x1 + q, cf.winv[k0] * s0, cf.winv[k1] * s1);
      halo_push2<float>(H, q, k, cf.winv[k0] * s0, cf.winv[k1] * s1);
    }
  }
  block_sum<2>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0) {
    if (acc[0] != 0.0) atomicAdd(dist ? &ds->part[PART_RR] : &ds->rr[it & 3], acc[0]);
    if (MODE == 0 && acc[1] != 0.0) atomicAdd(dist ? &ds->part[PART_RZ] : &ds->rz[it & 3], acc[1]);
  }
  if (MODE == 1) halo_signal(H, ds, k == H.kdn, k == H.kup);
}

// k_update_xr<1> with FOUR nodes per lane (block = 8 rows x 128 nodes).  Once x += alpha p had moved
// to k_update_p the two-node shape had too few bytes in flight: 120 us = 0.70 of the copy rate for
// 33 B/node at 256^3.  Same arithmetic per node: identical bits.
// TB = __half (FLIPB200_MG_H16): b0 is stored as fp16 and scaled by bscale (pcg.cuh); TB = float ignores bscale.
template <typename TB = float>
__global__ void __launch_bounds__(32 * SW_BY, FB_MINB_UPDATE_XR4)
k_update_xr4(const __grid_constant__ Sweep S, const __grid_constant__ PcgCoef cf, const uint8_t* __restrict__ fl,
             double* __restrict__ r, const double* __restrict__ Ap, TB* __restrict__ b0, float* __restrict__ x1,
             DevScalars* __restrict__ ds, int it_arg, int noexit, int dist, const __grid_constant__ PeerHalo H, double bscale) {
  __shared__ double sm[32];
  if (!noexit && ds->pcg_done) return;
  const int it = pcg_iter(ds, it_arg);
  const double pAp = ds->pAp[it & 3];
  const double rz_old = ds->rz[(it + 3) & 3];
  if (!noexit && !(pAp > 0.0)) {   // breakdown (also catches NaN)
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0 && threadIdx.y == 0) {
      ds->pcg_status = FLIPB200_ERR_BREAKDOWN; ds->pcg_done = 1; ds->pcg_iters = it;
      ds->rr_final = ds->rr[(it + 3) & 3];
    }
    return;
  }
  const double alpha = rz_old / pAp;                   // ConjugateGradient.h:72
  const double rrp = ds->rr[(it + 3) & 3];
  double sc = rrp > 0.0 ? rsqrt(rrp) : 1.0;
  if (!std::is_same<TB, float>::value) sc *= bscale;
  const int i = S.box.x0 + blockIdx.x * 128 + 4 * threadIdx.x;
  const int j = S.box.y0 + blockIdx.y * SW_BY + threadIdx.y;
  const int k = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp);
  const bool act = i < S.xe && j < S.box.y1;
  const int q = gidx(S.g, i, j, k);
  double acc[1] = {0.0};
  if (act) {
    const uint32_t f = *reinterpret_cast<const uint32_t*>(fl + q);
    const double2 ra = ld2<double>(r + q), rb = ld2<double>(r + q + 2);
    const double2 aa = ld2<double>(Ap + q), ab = ld2<double>(Ap + q + 2);
    const uint32_t fa = f & 0xffu, fb = (f >> 8) & 0xffu, fc = (f >> 16) & 0xffu, fd = f >> 24;
    const double r0 = (fa & FB_FLUID) ? ra.x - alpha * aa.x : 0.0;      // :74
    const double r1 = (fb & FB_FLUID) ? ra.y - alpha * aa.y : 0.0;
    const double r2 = (fc & FB_FLUID) ? rb.x - alpha * ab.x : 0.0;
    const double r3 = (fd & FB_FLUID) ? rb.y - alpha * ab.y : 0.0;
    if ((fa | fb) & FB_FLUID) st2<double>(r + q, r0, r1);
    if ((fc | fd) & FB_FLUID) st2<double>(r + q + 2, r2, r3);
    acc[0] = (r0 * r0 + r1 * r1) + (r2 * r2 + r3 * r3);                 // :76
    float4 s, y;
    s.x = (float)(r0 * sc); s.y = (float)(r1 * sc); s.z = (float)(r2 * sc); s.w = (float)(r3 * sc);
    y.x = cf.winv[6 - __popc(fa & FB_SMASK)] * s.x; y.y = cf.winv[6 - __popc(fb & FB_SMASK)] * s.y;
    y.z = cf.winv[6 - __popc(fc & FB_SMASK)] * s.z; y.w = cf.winv[6 - __popc(fd & FB_SMASK)] * s.w;
    stq4<TB>(b0 + q, s);
    *reinterpret_cast<float4*>(x1 + q) = y;
    halo_push4(H, q, k, y);
  }
  block_sum<1>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0 && acc[0] != 0.0) atomicAdd(dist ? &ds->part[PART_RR] : &ds->rr[it & 3], acc[0]);
  halo_signal(H, ds, k == H.kdn, k == H.kup);
}

// stop test of ConjugateGradient.h:77-78 (MG path: evaluated before the preconditioner is applied)
__global__ void k_pcg_check(DevScalars* __restrict__ ds, int it_arg) {
  if (ds->pcg_done) return;
  const int it = pcg_iter(ds, it_arg);
  const double rr = ds->rr[it & 3];
  if (rr < ds->threshold) { ds->pcg_done = 1; ds->pcg_iters = it; ds->rr_final = rr; ds->x_pending = 1; }
}

// z-slab runs: after the all-reduce of ds->part[], move the global sums into the ring slot of this
// iteration and clear the partials (it = -1 addresses slot 3, the "previous" slot of iteration 0).
__global__ void k_pcg_publish(DevScalars* __restrict__ ds, int mask, int it_arg) {
  const int it = pcg_iter(ds, it_arg);
  const int s = it & 3;
  if (mask & (1 << PART_PAP)) { ds->pAp[s] = ds->part[PART_PAP]; ds->part[PART_PAP] = 0.0; }
  if (mask & (1 << PART_RZ)) { ds->rz[s] = ds->part[PART_RZ]; ds->part[PART_RZ] = 0.0; }
  if (mask & (1 << PART_RR)) { ds->rr[s] = ds->part[PART_RR]; ds->part[PART_RR] = 0.0; }
}
// all-reduce of ds->part[id] over the peer-mapped mailboxes (halo.cuh).  One warp; lane r talks to rank r.
__global__ void k_pcg_exchange(DevScalars* __restrict__ ds, const __grid_constant__ RedPub R, int id, int it_arg, int noexit) {
  if (!noexit && ds->pcg_done) return;
  const int it = pcg_iter(ds, it_arg);
  const int lane = threadIdx.x;
  const unsigned seq = ds->red_seq[id] + 1u;
  const int par = (int)(seq & 1u);
  const double mine = ds->part[id];
  double v = 0.0;
  bool bad = false;
  if (lane < R.world) {
    const unsigned long long bits = (unsigned long long)__double_as_longlong(mine);
    const unsigned long long tag = (unsigned long long)seq << 32;
    unsigned long long* dst = R.peers[lane]->w[id][par][R.rank];
    st_relaxed_sys_u64(dst, (bits & 0xffffffffull) | tag);
    st_relaxed_sys_u64(dst + 1, (bits >> 32) | tag);
    const unsigned long long* src = R.peers[R.rank]->w[id][par][lane];
    const long long t0 = clock64();
    unsigned long long lo, hi;
    for (;;) {
      lo = ld_relaxed_sys_u64(src); hi = ld_relaxed_sys_u64(src + 1);
      if ((unsigned)(lo >> 32) == seq && (unsigned)(hi >> 32) == seq) break;
      if (clock64() - t0 > HALO_SPIN_LIMIT) { bad = true; break; }
    }
    v = __longlong_as_double((long long)((lo & 0xffffffffull) | (hi << 32)));
  }
  if (__any_sync(0xffffffffu, bad)) { if (lane == 0) { ds->pcg_status = FLIPB200_ERR_COMM; ds->pcg_done = 1; } return; }
  double sum = 0.0;
  for (int r = 0; r < R.world; ++r) sum += __shfl_sync(0xffffffffu, v, r);     // rank order: same bits on every rank
  if (lane == 0) {
    const int s = it & 3;
    if (id == PART_PAP) ds->pAp[s] = sum; else if (id == PART_RZ) ds->rz[s] = sum; else ds->rr[s] = sum;
    ds->part[id] = 0.0;
    ds->red_seq[id] = seq;
  }
}
// all-reduce + publish of one dot product: fused (mailboxes over NVLink) or NCCL
static int reduce_dot(flipb200_ctx* c, int id, int it_arg, int noexit) {
  if (c->halo.red_on) {
    k_pcg_exchange<<<1, 32, 0, c->stream>>>(c->ds, fb_red_make(c), id, it_arg, noexit);
    c->launches++;
    return 0;
  }
  int rc = fb_allreduce(c, &c->ds->part[id], 1, FB_F64, FB_SUM);
  if (rc) return rc;
  fb_pcg_publish(c, 1 << id, it_arg);
  return 0;
}
int fb_pcg_reduce_dot(flipb200_ctx* c, int id, int it_arg) { return reduce_dot(c, id, it_arg, 0); }

void fb_pcg_publish(flipb200_ctx* c, int mask, int it_arg) {
  k_pcg_publish<<<1, 1, 0, c->stream>>>(c->ds, mask, it_arg);
  c->launches++;
}

// K3: beta = rz/rz_old (:84), p = z + beta p (:85).
//     MODE 0: z = M^-1 r (Jacobi) on the fly, includes the stop test.
//     MODE 1: z = zf / sc from the V-cycle (fp32, scaled); it = -1 sets the first direction p = z.
template <int MODE>
__global__ void __launch_bounds__(32 * SW_BY)
k_update_p(const __grid_constant__ Sweep S, const __grid_constant__ PcgCoef cf, const uint8_t* __restrict__ fl,
           const double* __restrict__ r, const float* __restrict__ zf, double* __restrict__ p, double* __restrict__ x,
           DevScalars* __restrict__ ds, int it_arg, int noexit, const __grid_constant__ PeerHalo H) {
  if (!noexit && ds->pcg_done) return;
  const int it = pcg_iter(ds, it_arg);
  const bool lead = blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0 && threadIdx.y == 0;
  const bool first = it < 0;
  const double rr = first ? 0.0 : ds->rr[it & 3];
  if (MODE == 0 && !noexit && rr < ds->threshold) {
    if (lead) { ds->pcg_done = 1; ds->pcg_iters = it; ds->rr_final = rr; }
    return;
  }
  const double beta = first ? 0.0 : ds->rz[it & 3] / ds->rz[(it + 3) & 3];
  const double rrp = ds->rr[(it + 3) & 3];
  const double invsc = MODE == 1 ? (rrp > 0.0 ? sqrt(rrp) : 1.0) : 0.0;
  const double alpha = (MODE == 1 && !first) ? ds->rz[(it + 3) & 3] / ds->pAp[it & 3] : 0.0;   // as in k_update_xr of this iteration
  STREAM_PROLOGUE(S)
  if (act) {
    uchar2 f = *reinterpret_cast<const uchar2*>(fl + q);
    double2 pv = ld2<double>(p + q);
    if (MODE == 1 && !first) {                       // x += alpha p with the old direction (ConjugateGradient.h:73)
      double2 xv = ld2<double>(x + q);
      if (f.x & FB_FLUID) xv.x += alpha * pv.x;
      if (f.y & FB_FLUID) xv.y += alpha * pv.y;
      if ((f.x | f.y) & FB_FLUID) st2<double>(x + q, xv.x, xv.y);
    }
    double z0, z1;
    if (MODE == 0) {
      double2 rv = ld2<double>(r + q);
      z0 = rv.x * cf.inv[6 - __popc(f.x & FB_SMASK)]; z1 = rv.y * cf.inv[6 - __popc(f.y & FB_SMASK)];
    } else {
      float2 zv = ld2<float>(zf + q);
      z0 = (double)zv.x * invsc; z1 = (double)zv.y * invsc;
    }
    const double p0 = (f.x & FB_FLUID) ? z0 + beta * pv.x : 0.0;
    const double p1 = (f.y & FB_FLUID) ? z1 + beta * pv.y : 0.0;
    if ((f.x | f.y) & FB_FLUID) st2<double>(p + q, p0, p1);
    halo_push2<double>(H, q, k, p0, p1);          // the neighbours' ghost planes see 0 off the fluid cells
  }
  halo_signal(H, ds, k == H.kdn, k == H.kup);
}

// k_update_p<1> with FOUR nodes per lane (block = 8 rows x 128 nodes), like k_update_xr4: since it also
// carries x += alpha p the kernel moves 37 B/node and the two-node shape keeps ~42 B per lane in flight.
// FLIPB200_UP4=1; NOT the default and not yet measured (written after the GPU budget of round 1 was
// spent).  Same expressions per node as k_update_p<1>: identical bits.
__global__ void __launch_bounds__(32 * SW_BY)
k_update_p4(const __grid_constant__ Sweep S, const uint8_t* __restrict__ fl, const float* __restrict__ zf,
            double* __restrict__ p, double* __restrict__ x, DevScalars* __restrict__ ds, int it_arg, int noexit,
            const __grid_constant__ PeerHalo H, double inv_bscale /* 1, or 1 / bscale with FLIPB200_MG_H16 (pcg.cuh) */) {
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
  const bool act = i < S.xe && j < S.box.y1;
  const int q = gidx(S.g, i, j, k);
  if (act) {
    const uint32_t f = *reinterpret_cast<const uint32_t*>(fl + q);
    const double2 pa = ld2<double>(p + q), pb = ld2<double>(p + q + 2);
    const float4 zv = *reinterpret_cast<const float4*>(zf + q);
    const bool f0 = f & FB_FLUID, f1 = (f >> 8) & FB_FLUID, f2 = (f >> 16) & FB_FLUID, f3 = (f >> 24) & FB_FLUID;
    if (!first) {                                    // x += alpha p with the old direction (ConjugateGradient.h:73)
      double2 xa = ld2<double>(x + q), xb = ld2<double>(x + q + 2);
      if (f0) xa.x += alpha * pa.x;
      if (f1) xa.y += alpha * pa.y;
      if (f2) xb.x += alpha * pb.x;
      if (f3) xb.y += alpha * pb.y;
      if (f0 | f1) st2<double>(x + q, xa.x, xa.y);
      if (f2 | f3) st2<double>(x + q + 2, xb.x, xb.y);
    }
    const double z0 = (double)zv.x * invsc, z1 = (double)zv.y * invsc, z2 = (double)zv.z * invsc, z3 = (double)zv.w * invsc;
    const double p0 = f0 ? z0 + beta * pa.x : 0.0;
    const double p1 = f1 ? z1 + beta * pa.y : 0.0;
    const double p2 = f2 ? z2 + beta * pb.x : 0.0;
    const double p3 = f3 ? z3 + beta * pb.y : 0.0;
    if (f0 | f1) st2<double>(p + q, p0, p1);
    if (f2 | f3) st2<double>(p + q + 2, p2, p3);
    halo_push2<double>(H, q, k, p0, p1);             // the neighbours' ghost planes see 0 off the fluid cells
    halo_push2<double>(H, q + 2, k, p2, p3);
  }
  halo_signal(H, ds, k == H.kdn, k == H.kup);
}

// The stop test ended the solve between k_update_xr and k_update_p of iteration pcg_iters: apply that
// iteration's x += alpha p (the ring slots are intact, k_pcg_advance did not run).
__global__ void __launch_bounds__(32 * SW_BY)
k_pcg_finish_x(const __grid_constant__ Sweep S, const uint8_t* __restrict__ fl, double* __restrict__ x,
               const double* __restrict__ p, const DevScalars* __restrict__ ds) {
  if (!ds->x_pending) return;
  const int it = ds->pcg_iters;
  const double alpha = ds->rz[(it + 3) & 3] / ds->pAp[it & 3];
  STREAM_PROLOGUE(S)
  if (act) {
    uchar2 f = *reinterpret_cast<const uchar2*>(fl + q);
    double2 xv = ld2<double>(x + q), pv = ld2<double>(p + q);
    if (f.x & FB_FLUID) xv.x += alpha * pv.x;
    if (f.y & FB_FLUID) xv.y += alpha * pv.y;
    if ((f.x | f.y) & FB_FLUID) st2<double>(x + q, xv.x, xv.y);
  }
}

// end of an iteration: clear the ring slot the next iteration accumulates into, enforce the
// iteration limit (IterativeSolverBase.h:283) and advance the device-side iteration counter
__global__ void k_pcg_advance(DevScalars* __restrict__ ds, int it_arg, int noexit, int nprod) {
  if (!noexit && ds->pcg_done) return;
  ds->halo_base += nprod;                      // fused ghost pushes: sequence base of the next segment (halo.cuh)
  const int it = pcg_iter(ds, it_arg);
  const int s = (it + 1) & 3;
  ds->pAp[s] = 0.0; ds->rz[s] = 0.0; ds->rr[s] = 0.0;
  const int max_iter = ds->pcg_max_iter;
  if (!noexit && it >= 0 && it + 1 >= max_iter) {
    ds->pcg_done = 1; ds->pcg_iters = max_iter; ds->rr_final = ds->rr[it & 3]; ds->pcg_status = FLIPB200_ERR_NOCONV;
  }
  if (it >= 0) ds->pcg_it = it + 1;
}

void fb_make_coef(const flipb200_ctx* c, double dt, PcgCoef* cf) {
  const double scale = dt / (c->P.rho * c->P.dx * c->P.dx);       // fluid.cc:306
  float acc = 0.0f;
  cf->diag[0] = 0.0; cf->inv[0] = 0.0; cf->winv[0] = 0.f;
  for (int k = 1; k <= 6; ++k) {
    acc = (float)((double)acc + scale);                           // daccessor.setValue(c, getValue(c) + scale), :333
    cf->diag[k] = (double)acc;
    cf->inv[k] = 1.0 / (double)acc;
    cf->winv[k] = (float)(MG_OMEGA1 / (k * scale));
  }
  cf->off = (double)(float)(-1 * scale);                          // :335
}

// one CG iteration on the stream (it_arg = explicit iteration or IT_DEV for graph replay).
// z-slab runs add: a one-plane halo fill of the search direction before A.p and an all-reduce +
// publish after every kernel that ends in a dot product.
// bookkeeping of the fused ghost pushes (halo.cuh): a "segment" is what runs between two
// k_pcg_advance launches; sequence numbers inside kernels are relative to its base.
static inline int halo_seg_close(flipb200_ctx* c) {
  HaloState& hs = c->halo;
  const int n = (int)(hs.abs_seq - hs.seg_base);
  hs.seg_base = hs.abs_seq;
  return n;
}
static inline void halo_replayed(flipb200_ctx* c, int n) {   // a captured iteration ran again: shift the host mirror
  HaloState& hs = c->halo;
  if (hs.mode == 0) return;
  hs.abs_seq += n; hs.seg_base += n;
  for (int l = 0; l < MG_MAX_LEVELS; ++l) for (int a = 0; a < 3; ++a) hs.last_prod[l][a] += n;
}

static int enqueue_iteration(flipb200_ctx* c, const Sweep& S, const PcgCoef& cf, bool use_mg, int it_arg, int noexit) {
  const dim3 sg = stream_grid(S), tg = stencil_grid(S);
  MgLevel& L0 = c->mg[0];
  const int dist = c->world > 1 ? 1 : 0;
  const bool fused = use_mg && c->halo.mode != 0;
  const int k0 = c->g.k0, k1 = c->g.k1 - 1;
  const bool h16 = use_mg && mg_h16_active(c);              // experimental fp16 storage of b0 / the level-0 residual (pcg.cuh)
  const double bscale = h16 ? mg_h16_bscale(c) : 1.0;
  int rc;
  fb_prof(c, "start");
  const bool p32 = use_mg && pcg_p32_active(c);
  const size_t pel = p32 ? sizeof(float) : sizeof(double);      // storage width of the search direction
  if (fused) rc = fb_halo_point(c); else rc = fb_halo_fill(c, c->g, c->pd, pel, 1, 1);
  if (rc) return rc;
  fb_prof(c, "halo pd");
  launch_spmv(c, cf, it_arg, noexit, dist, fused ? fb_halo_make(c, 0, HA_PD, -1, -1, 0, k0, k1, tg.x * tg.y) : halo_off(), p32);
  fb_prof(c, "k_spmv");
  if (dist) {
    if (use_mg) { if ((rc = reduce_dot(c, PART_PAP, it_arg, noexit))) return rc; }
    else {
      if ((rc = fb_allreduce(c, &c->ds->part[PART_PAP], 1, FB_F64, FB_SUM))) return rc;
      fb_pcg_publish(c, 1 << PART_PAP, it_arg);
    }
  }
  fb_prof(c, "allreduce pAp");
  if (use_mg) {
    const bool st4_on = c->opt_st4;
    const Sweep S4 = make_sweep(c->g, c->box, 4);
    if (st4_on && S4.xe - S4.box.x0 >= 128) {
      const dim3 sg4(cdiv(S4.xe - S4.box.x0, 128), sg.y, sg.z);
      if (h16) k_update_xr4<__half><<<sg4, STREAM_BLOCK, 0, c->stream>>>(sweep_down(S4), cf, c->flags, c->r, c->Ap, reinterpret_cast<__half*>(L0.b), L0.xa, c->ds,
                                                                         it_arg, noexit, dist,
                                                                         fused ? fb_halo_make(c, -1, -1, 0, HA_XA, sizeof(float), k0, k1, sg4.x * sg4.y) : halo_off(), bscale);
      else
      k_update_xr4<<<sg4, STREAM_BLOCK, 0, c->stream>>>(sweep_down(S4), cf, c->flags, c->r, c->Ap, L0.b, L0.xa, c->ds, it_arg, noexit, dist,
                                                        fused ? fb_halo_make(c, -1, -1, 0, HA_XA, sizeof(float), k0, k1, sg4.x * sg4.y) : halo_off(), 1.0);
    } else
      k_update_xr<1><<<sg, STREAM_BLOCK, 0, c->stream>>>(sweep_down(S), cf, c->flags, c->x, c->r, c->pd, c->Ap, L0.b, L0.xa, c->ds, it_arg, noexit, dist,
                                                         fused ? fb_halo_make(c, -1, -1, 0, HA_XA, sizeof(float), k0, k1, sg.x * sg.y) : halo_off());
    fb_prof(c, "k_update_xr");
    if (dist && (rc = reduce_dot(c, PART_RR, it_arg, noexit))) return rc;
    if (!noexit) { k_pcg_check<<<1, 1, 0, c->stream>>>(c->ds, it_arg); c->launches++; }
    fb_prof(c, "allreduce rr");
    rc = fb_mg_vcycle(c, it_arg);
    if (rc) return rc;
    const bool up4_on = true;                       // four nodes per lane whenever the rows are wide enough (k_update_p4)
    if (p32) {
      const dim3 sg4(cdiv(S4.xe - S4.box.x0, 128), sg.y, sg.z);
      fb_p32_update_p(c, L0.xb, it_arg, noexit, h16 ? 1.0 / bscale : 1.0, true,
                      fused ? fb_halo_make(c, -1, -1, 0, HA_PD, sizeof(float), k0, k1, sg4.x * sg4.y) : halo_off());
    } else if ((up4_on || h16) && st4_on && S4.xe - S4.box.x0 >= 128) {
      const dim3 sg4(cdiv(S4.xe - S4.box.x0, 128), sg.y, sg.z);
      k_update_p4<<<sg4, STREAM_BLOCK, 0, c->stream>>>(sweep_down(S4), c->flags, L0.xb, c->pd, c->x, c->ds, it_arg, noexit,
                                                       fused ? fb_halo_make(c, -1, -1, 0, HA_PD, sizeof(double), k0, k1, sg4.x * sg4.y) : halo_off(),
                                                       h16 ? 1.0 / bscale : 1.0);
    } else
    k_update_p<1><<<sg, STREAM_BLOCK, 0, c->stream>>>(sweep_down(S), cf, c->flags, c->r, L0.xb, c->pd, c->x, c->ds, it_arg, noexit,
                                                      fused ? fb_halo_make(c, -1, -1, 0, HA_PD, sizeof(double), k0, k1, sg.x * sg.y) : halo_off());
  } else {
    k_update_xr<0><<<sg, STREAM_BLOCK, 0, c->stream>>>(sweep_down(S), cf, c->flags, c->x, c->r, c->pd, c->Ap, nullptr, nullptr, c->ds, it_arg, noexit, dist, halo_off());
    if (dist) {
      if ((rc = fb_allreduce(c, &c->ds->part[PART_RZ], 2, FB_F64, FB_SUM))) return rc;   // r.z and r.r are adjacent
      fb_pcg_publish(c, (1 << PART_RZ) | (1 << PART_RR), it_arg);
    }
    k_update_p<0><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, cf, c->flags, c->r, nullptr, c->pd, nullptr, c->ds, it_arg, noexit, halo_off());
  }
  fb_prof(c, "k_update_p");
  k_pcg_advance<<<1, 1, 0, c->stream>>>(c->ds, it_arg, noexit, halo_seg_close(c));
  c->launches += 4;
  return 0;
}

// The iteration is captured into a CUDA graph (36 kernels with the V-cycle, most of them
// launch-latency bound on the coarse levels) and replayed; kernels read the iteration number, the
// threshold and the iteration limit from DevScalars, so the graph stays valid across solves and steps
// until the sweep box, dt or the mode change.  FLIPB200_NO_GRAPH=1 falls back to plain launches.
void fb_graph_invalidate(flipb200_ctx* c) {
  if (c->it_exec) cudaGraphExecDestroy(c->it_exec);
  if (c->it_graph) cudaGraphDestroy(c->it_graph);
  c->it_exec = nullptr; c->it_graph = nullptr;
}
static bool same_box(const Box& a, const Box& b) {
  return a.x0 == b.x0 && a.y0 == b.y0 && a.z0 == b.z0 && a.x1 == b.x1 && a.y1 == b.y1 && a.z1 == b.z1;
}
static bool get_iteration_graph(flipb200_ctx* c, const Sweep& S, const PcgCoef& cf, bool use_mg, double dt, int noexit) {
  if (c->opt_no_graph) return false;
  if (!fb_comm_capturable(c)) return false;      // in-process test transport synchronises on the host
  // The NCCL-copy fallback of the ghost planes (FLIPB200_FUSED_HALO=0) runs with plain launches: replaying its
  // send/recv pairs from the captured iteration ended solves early with a wrong pressure on 2 GPUs
  // (tests/test_gpu_nccl.py [2], round 2; plain launches of the same calls are correct).  The default path -- ghost
  // planes pushed by the producing kernels -- replays from the graph.
  if (c->world > 1 && c->halo.mode == 0) return false;
  const int mode = (use_mg ? 1 : 0) | (noexit ? 2 : 0);
  if (c->it_exec && same_box(c->it_box, c->box) && same_box(c->it_gbox, c->gbox) && c->it_dt == dt && c->it_mode == mode) return true;
  fb_graph_invalidate(c);
  if (cudaStreamBeginCapture(c->stream, c->world > 1 ? cudaStreamCaptureModeRelaxed : cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError(); return false; }
  const int64_t l0 = c->launches;
  const long long seq0 = c->halo.abs_seq;
  int rc = enqueue_iteration(c, S, cf, use_mg, IT_DEV, noexit);
  c->it_nprod = (int)(c->halo.abs_seq - seq0);      // producers per iteration (the capture itself did not run them)
  c->it_kernels = c->launches - l0;
  c->launches = l0;
  cudaError_t e = cudaStreamEndCapture(c->stream, &c->it_graph);
  if (rc || e != cudaSuccess || !c->it_graph) { cudaGetLastError(); fb_graph_invalidate(c); return false; }
  if (cudaGraphInstantiate(&c->it_exec, c->it_graph, 0) != cudaSuccess) { cudaGetLastError(); fb_graph_invalidate(c); return false; }
  c->it_box = c->box; c->it_gbox = c->gbox; c->it_dt = dt; c->it_mode = mode;
  return true;
}

int fb_pcg_solve(flipb200_ctx* c, double dt, int* iters_out, double* rel_out) {
  if (iters_out) *iters_out = 0;
  if (rel_out) *rel_out = 0.0;
  const Sweep S = make_sweep(c->g, c->box);
  if (c->n_fluid == 0 || sweep_empty(S)) return 0;
  PcgCoef cf; fb_make_coef(c, dt, &cf);
  const dim3 sg = stream_grid(S);
  const double bnorm2 = c->hs->bnorm2;             // fetched by fb_rhs_div
  const float* bvec = c->bf[c->bcur];
  const bool use_mg = c->mg_levels >= 2;
  const int dist = c->world > 1 ? 1 : 0;
  MgLevel& L0 = c->mg[0];
  const double tol = c->P.pcg_tol;
  const double threshold = tol * tol * bnorm2;     // ConjugateGradient.h:53
  const int max_iter = c->P.pcg_max_iter < 0 ? 2 * c->n_fluid : c->P.pcg_max_iter;   // IterativeSolverBase.h:283
  if (c->world > 1 && !c->halo_ready) {            // first collective solve: map the neighbours' arrays (halo.cuh)
    int rc = fb_halo_setup(c);
    if (rc) return rc;
    c->halo_ready = true;
  }
  const bool fused = use_mg && c->halo.mode != 0;
  const int k0 = c->g.k0, k1 = c->g.k1 - 1;
  if (use_mg) {
    int rc = fb_mg_setup(c, dt);                   // hierarchy: rebuilt only when the cell classes changed
    if (rc) return rc;
  }
  c->halo.seg_base = c->halo.abs_seq;
  k_pcg_reset<<<1, 1, 0, c->stream>>>(c->ds, bnorm2, threshold, max_iter, (int)c->halo.seg_base);
  const bool h16 = use_mg && mg_h16_active(c);              // fp16 storage of b0 / the level-0 residual (pcg.cuh)
  const double bscale = h16 ? mg_h16_bscale(c) : 1.0;
  // warm start: only from a converged previous solution, and never for b = 0 (Eigen returns x = 0 there)
  const bool warm = c->P.warm_start && c->x_valid && c->xprev && bnorm2 > 0.0 && bnorm2 == bnorm2 && max_iter > 0;
  c->x_valid = false;
  if (warm) {
    const double theta = 1.0 - 1.0 / (double)c->P.relax_div;
    // guess (see k_warm_guess): first pass from the previous step's first solution, second pass th x_1 + A^-1 G of the
    // previous step, later passes extrapolated along b_{k+1} - b_k = th (b_k - b_{k-1})
    const int pk = c->pass_idx;
    if (pk == 0) {
      const int pendG = c->last_step_passes == 2 ? 1 : 0;
      const int src_first = (c->last_step_passes != 1 && c->have_first) ? 1 : 0;
      if (pendG || src_first) {
        k_warm_guess<0><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, c->flags, c->x, c->xprev, c->xfirst, c->xG, theta, pendG, src_first, 0);
        c->launches++;
      }
      if (pendG) c->have_G = true;
      c->last_step_passes = 0;
      c->xprev_pass = -1;
    } else if (pk == 1) {
      k_warm_guess<1><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, c->flags, c->x, c->xprev, c->xfirst, c->xG, theta, 0, 0, c->have_G ? 1 : 0);
      c->have_first = true; c->xprev_pass = 0; c->launches++;
    } else if (c->xprev_pass == pk - 2) {
      if (pk == 2) { k_warm_guess<2><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, c->flags, c->x, c->xprev, c->xfirst, c->xG, theta, 0, 0, 0); c->have_G = true; }
      else k_warm_guess<3><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, c->flags, c->x, c->xprev, c->xfirst, c->xG, theta, 0, 0, 0);
      c->xprev_pass = pk - 1; c->launches++;
    }
    int rc = fb_halo_fill(c, c->g, c->x, sizeof(double), 1, 1);     // z-slabs: ghost planes of the guess
    if (rc) return rc;
    k_warm_ax<<<stencil_grid(S), STENCIL_BLOCK, 0, c->stream>>>(S, cf, c->flags, bvec, c->x, c->Ap, c->ds);
    if (dist && (rc = fb_allreduce(c, &c->ds->part[PART_PAP], 2, FB_F64, FB_SUM))) return rc;    // (b.Ag, Ag.Ag) are adjacent
    k_warm_scalars<<<1, 1, 0, c->stream>>>(c->ds);
    k_warm_rr<<<sg, STREAM_BLOCK, 0, c->stream>>>(S, c->flags, bvec, c->Ap, c->ds);
    if (dist && (rc = fb_allreduce(c, &c->ds->part[PART_RR], 1, FB_F64, FB_SUM))) return rc;
    k_warm_scalars2<<<1, 1, 0, c->stream>>>(c->ds);
    const PeerHalo Hw = (use_mg && fused) ? fb_halo_make(c, -1, -1, 0, HA_XA, sizeof(float), k0, k1, sg.x * sg.y) : halo_off();
    if (h16) k_warm_init<1, __half><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, cf, c->flags, bvec, c->x, c->r, c->pd, c->Ap, reinterpret_cast<__half*>(L0.b), L0.xa, bscale, c->ds, dist, Hw);
    else if (use_mg) k_warm_init<1><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, cf, c->flags, bvec, c->x, c->r, c->pd, c->Ap, L0.b, L0.xa, 1.0, c->ds, dist, Hw);
    else {
      k_warm_init<0><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, cf, c->flags, bvec, c->x, c->r, c->pd, c->Ap, (float*)nullptr, nullptr, 1.0, c->ds, dist, halo_off());
      if (dist) {
        if ((rc = fb_allreduce(c, &c->ds->part[PART_RZ], 1, FB_F64, FB_SUM))) return rc;
        fb_pcg_publish(c, 1 << PART_RZ, -1);
      }
    }
    k_warm_check<<<1, 1, 0, c->stream>>>(c->ds);
    CK_LAUNCH(); c->launches += 7;
    if ((rc = fb_fetch_scalars(c))) return rc;          // one round trip: |r0|^2 sizes the first batch, and the guess may already pass
  } else if (h16)
    k_pcg_init<1, __half><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, cf, c->flags, bvec, c->x, c->r, c->pd, reinterpret_cast<__half*>(L0.b), L0.xa,
                                                              (bnorm2 > 0.0 ? 1.0 / sqrt(bnorm2) : 1.0) * bscale, c->ds, dist,
                                                              fused ? fb_halo_make(c, -1, -1, 0, HA_XA, sizeof(float), k0, k1, sg.x * sg.y) : halo_off());
  else if (use_mg)
    k_pcg_init<1><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, cf, c->flags, bvec, c->x, c->r, c->pd, L0.b, L0.xa,
                                                      bnorm2 > 0.0 ? 1.0 / sqrt(bnorm2) : 1.0, c->ds, dist,
                                                      fused ? fb_halo_make(c, -1, -1, 0, HA_XA, sizeof(float), k0, k1, sg.x * sg.y) : halo_off());
  else {
    k_pcg_init<0><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, cf, c->flags, bvec, c->x, c->r, c->pd, (float*)nullptr, nullptr, 0.0, c->ds, dist, halo_off());
    if (dist) {
      int rc = fb_allreduce(c, &c->ds->part[PART_RZ], 1, FB_F64, FB_SUM);
      if (rc) return rc;
      fb_pcg_publish(c, 1 << PART_RZ, -1);
    }
  }
  CK_LAUNCH(); c->launches += 2;
  if (bnorm2 == 0.0) return 0;                     // rhsNorm2 == 0 => x = 0, 0 iterations (ConjugateGradient.h:46-52)
  if (!(bnorm2 == bnorm2)) return FLIPB200_ERR_BREAKDOWN;
  bool done = max_iter == 0 || (warm && c->hs->pcg_done != 0);
  const double rr_start = warm ? c->hs->warm_rr0 : bnorm2;
  if (use_mg && !done) {
    int rc = fb_mg_vcycle(c, -1);                  // z0 = M^-1 r0, rz[3] = r0.z0
    if (rc) return rc;
    const Sweep S4p = make_sweep(c->g, c->box, 4);
    const dim3 sg4p(cdiv(S4p.xe - S4p.box.x0, 128), sg.y, sg.z);
    if (pcg_p32_active(c))
      fb_p32_update_p(c, L0.xb, -1, 0, h16 ? 1.0 / bscale : 1.0, false,
                      fused ? fb_halo_make(c, -1, -1, 0, HA_PD, sizeof(float), k0, k1, sg4p.x * sg4p.y) : halo_off());
    else if (h16)
      k_update_p4<<<sg4p, STREAM_BLOCK, 0, c->stream>>>(S4p, c->flags, L0.xb, c->pd, c->x, c->ds, -1, 0,
                                                        fused ? fb_halo_make(c, -1, -1, 0, HA_PD, sizeof(double), k0, k1, sg4p.x * sg4p.y) : halo_off(), 1.0 / bscale);
    else
    k_update_p<1><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, cf, c->flags, c->r, L0.xb, c->pd, c->x, c->ds, -1, 0,
                                                      fused ? fb_halo_make(c, -1, -1, 0, HA_PD, sizeof(double), k0, k1, sg.x * sg.y) : halo_off());
    k_pcg_advance<<<1, 1, 0, c->stream>>>(c->ds, -1, 0, halo_seg_close(c));
    c->launches += 2;
  }
  const bool graph = !done && get_iteration_graph(c, S, cf, use_mg, dt, 0);
  // first batch: as many iterations as the previous solve needed (consecutive solves of a scene need
  // the same number +-1: usually one batch, one host round trip), then small batches
  int batch = c->last_iters > 4 ? c->last_iters : (use_mg ? 8 : 32);
  if (warm && c->last_rate > 0.0 && rr_start > threshold) {      // digits still to gain / digits per iteration of the last solve
    const int need = (int)ceil(log10(rr_start / threshold) / c->last_rate) + 1;
    batch = need < 2 ? 2 : need;
  }
  int it = 0;
  while (!done) {
    int stop = min(it + batch, max_iter);
    for (; it < stop; ++it) {
      if (graph) {
        if (cudaGraphLaunch(c->it_exec, c->stream) != cudaSuccess) { fb_graph_invalidate(c); return FLIPB200_ERR_CUDA; }
        c->launches += c->it_kernels;
        halo_replayed(c, c->it_nprod);
      } else {
        int rc = enqueue_iteration(c, S, cf, use_mg, it, 0);
        if (rc) return rc;
      }
    }
    cudaError_t le = cudaGetLastError();
    int rc = le == cudaSuccess ? fb_fetch_scalars(c) : FLIPB200_ERR_CUDA;
    if (rc) return rc;
    done = c->hs->pcg_done != 0 || it >= max_iter;
    batch = use_mg ? 2 : 8;
  }
  if (c->hs->x_pending) {
    if (use_mg && pcg_p32_active(c)) fb_p32_finish_x(c);
    else
    k_pcg_finish_x<<<sg, STREAM_BLOCK, 0, c->stream>>>(S, c->flags, c->x, c->pd, c->ds);
    CK_LAUNCH(); c->launches++;
  }
  c->last_iters = c->hs->pcg_iters;
  if (c->hs->pcg_iters > 0 && c->hs->rr_final > 0.0 && rr_start > c->hs->rr_final)
    c->last_rate = log10(rr_start / c->hs->rr_final) / c->hs->pcg_iters;
  c->x_valid = c->hs->pcg_status == 0 && c->hs->pcg_done != 0;
  if (c->opt_warm_debug)
    fprintf(stderr, "flipb200 solve: pass %d warm %d gamma %.6g |r0|/|b| %.3e iters %d rate %.3f digits/it\n", c->pass_idx, (int)warm,
            warm ? c->hs->warm_gamma : 0.0, sqrt(rr_start / bnorm2), c->hs->pcg_iters, c->last_rate / 2);
  if (iters_out) *iters_out = c->hs->pcg_iters;
  if (rel_out) *rel_out = sqrt(c->hs->rr_final / bnorm2);          // cg.error(), ConjugateGradient.h:88
  if (c->hs->pcg_status == FLIPB200_ERR_BREAKDOWN) return FLIPB200_ERR_BREAKDOWN;
  if (c->hs->pcg_status == FLIPB200_ERR_COMM) return FLIPB200_ERR_COMM;     // a ghost-plane wait timed out (halo.cuh)
  return 0;   // hitting max_iter is reported through iters == max_iter, as Eigen does
}

// Timing hook: `iters` launches of K1 alone (which == SPMV) or of one whole PCG iteration
// (which == PCG_ITER; includes the V-cycle when precond = 1) on the current system, no convergence
// exit.  units = swept nodes per launch.
int fb_pcg_time(flipb200_ctx* c, int which, int iters, float* ms, int64_t* units) {
  const Sweep S = make_sweep(c->g, c->box);
  if (sweep_empty(S)) return FLIPB200_ERR_STATE;
  c->x_valid = false;                              // the timing loop overwrites x
  const double dt = c->P.dt_max;
  PcgCoef cf; fb_make_coef(c, dt, &cf);
  const dim3 sg = stream_grid(S), tg = stencil_grid(S);
  const bool use_mg = c->mg_levels >= 2 && which == FLIPB200_KERNEL_PCG_ITER;
  const int dist = c->world > 1 ? 1 : 0;
  MgLevel& L0 = c->mg[0];
  const double bn = c->hs->bnorm2 > 0.0 ? c->hs->bnorm2 : 1.0;
  if (c->world > 1 && !c->halo_ready) {
    int rc = fb_halo_setup(c);
    if (rc) return rc;
    c->halo_ready = true;
  }
  const bool fused = use_mg && c->halo.mode != 0;
  const int k0 = c->g.k0, k1 = c->g.k1 - 1;
  if (use_mg) { int rc = fb_mg_setup(c, dt); if (rc) return rc; }
  c->halo.seg_base = c->halo.abs_seq;
  k_pcg_reset<<<1, 1, 0, c->stream>>>(c->ds, bn, 0.0, 1 << 30, (int)c->halo.seg_base);
  const bool h16 = use_mg && mg_h16_active(c);
  const double bscale = h16 ? mg_h16_bscale(c) : 1.0;
  if (h16) {
    k_pcg_init<1, __half><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, cf, c->flags, c->bf[c->bcur], c->x, c->r, c->pd, reinterpret_cast<__half*>(L0.b), L0.xa,
                                                              bscale / sqrt(bn), c->ds, dist,
                                                              fused ? fb_halo_make(c, -1, -1, 0, HA_XA, sizeof(float), k0, k1, sg.x * sg.y) : halo_off());
    int rc = fb_mg_vcycle(c, -1); if (rc) return rc;
    const Sweep S4 = make_sweep(c->g, c->box, 4);
    const dim3 sg4(cdiv(S4.xe - S4.box.x0, 128), sg.y, sg.z);
    if (pcg_p32_active(c)) fb_p32_update_p(c, L0.xb, -1, 1, 1.0 / bscale, false,
                                           fused ? fb_halo_make(c, -1, -1, 0, HA_PD, sizeof(float), k0, k1, sg4.x * sg4.y) : halo_off());
    else
    k_update_p4<<<sg4, STREAM_BLOCK, 0, c->stream>>>(S4, c->flags, L0.xb, c->pd, c->x, c->ds, -1, 1,
                                                     fused ? fb_halo_make(c, -1, -1, 0, HA_PD, sizeof(double), k0, k1, sg4.x * sg4.y) : halo_off(), 1.0 / bscale);
    k_pcg_advance<<<1, 1, 0, c->stream>>>(c->ds, -1, 1, halo_seg_close(c));
  } else if (use_mg) {
    k_pcg_init<1><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, cf, c->flags, c->bf[c->bcur], c->x, c->r, c->pd, L0.b, L0.xa, 1.0 / sqrt(bn), c->ds, dist,
                                                      fused ? fb_halo_make(c, -1, -1, 0, HA_XA, sizeof(float), k0, k1, sg.x * sg.y) : halo_off());
    int rc = fb_mg_vcycle(c, -1); if (rc) return rc;
    if (pcg_p32_active(c)) {
      const Sweep S4 = make_sweep(c->g, c->box, 4);
      const dim3 sg4(cdiv(S4.xe - S4.box.x0, 128), sg.y, sg.z);
      fb_p32_update_p(c, L0.xb, -1, 1, 1.0, false, fused ? fb_halo_make(c, -1, -1, 0, HA_PD, sizeof(float), k0, k1, sg4.x * sg4.y) : halo_off());
    } else
    k_update_p<1><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, cf, c->flags, c->r, L0.xb, c->pd, c->x, c->ds, -1, 1,
                                                      fused ? fb_halo_make(c, -1, -1, 0, HA_PD, sizeof(double), k0, k1, sg.x * sg.y) : halo_off());
    k_pcg_advance<<<1, 1, 0, c->stream>>>(c->ds, -1, 1, halo_seg_close(c));
  } else {
    k_pcg_init<0><<<sg, STREAM_BLOCK, 0, c->stream>>>(S, cf, c->flags, c->bf[c->bcur], c->x, c->r, c->pd, (float*)nullptr, nullptr, 0.0, c->ds, dist, halo_off());
    if (dist) {
      int rc = fb_allreduce(c, &c->ds->part[PART_RZ], 1, FB_F64, FB_SUM);
      if (rc) return rc;
      fb_pcg_publish(c, 1 << PART_RZ, -1);
    }
  }
  bool graph = false;
  const char* pe = getenv("FLIPB200_PROFILE_ITER");
  const bool prof_env = pe && pe[0] == '1';
  const bool prof = (prof_env || (c->prof && c->prof->request)) && which == FLIPB200_KERNEL_PCG_ITER;
  if (prof && !c->prof) c->prof = (IterProf*)calloc(1, sizeof(IterProf));
  if (which == FLIPB200_KERNEL_PCG_ITER && !prof) graph = get_iteration_graph(c, S, cf, use_mg, dt, 1);
  for (int it = 0; it < 3 + iters; ++it) {
    if (prof && it >= 3) { c->prof->on = true; c->prof->n = 0; c->halo.dbg_n = 0; }
    if (it == 3) CK(cudaEventRecord(c->ev[6], c->stream));
    if (which != FLIPB200_KERNEL_PCG_ITER) {
      launch_spmv(c, cf, it, 1, 0, halo_off(), c->mg_levels >= 2 && pcg_p32_active(c));   // the stencil the product runs
      c->launches++;
    } else if (graph) {
      if (cudaGraphLaunch(c->it_exec, c->stream) != cudaSuccess) { fb_graph_invalidate(c); return FLIPB200_ERR_CUDA; }
      c->launches += c->it_kernels;
      halo_replayed(c, c->it_nprod);
    } else {
      int rc = enqueue_iteration(c, S, cf, use_mg, it, 1);
      if (rc) return rc;
    }
    if (prof && it >= 3) {
      IterProf* p = c->prof;
      p->on = false;
      CK(cudaStreamSynchronize(c->stream));
      for (int m = 1; m < p->n; ++m) { float t = 0.f; cudaEventElapsedTime(&t, p->ev[m - 1], p->ev[m]); p->acc[m] += t; }
      p->iters++;
    }
  }
  if (prof && c->prof->iters > 0) {
    IterProf* p = c->prof;
    double tot = 0.0;
    for (int m = 1; m < p->n; ++m) tot += p->acc[m] / p->iters;
    p->out_n = 0;
    for (int m = 1; m < p->n; ++m) { p->out_us[p->out_n] = (float)(p->acc[m] / p->iters * 1e3); p->out_name[p->out_n] = p->name[m]; p->out_n++; }
    p->request = false;
    if (prof_env) {
    fprintf(stderr, "flipb200 rank %d: PCG iteration profile over %d iterations (plain launches), total %.1f us\n", c->rank, p->iters, tot * 1e3);
    for (int m = 1; m < p->n; ++m) fprintf(stderr, "  r%d %-28s %8.1f us\n", c->rank, p->name[m], p->acc[m] / p->iters * 1e3);
    }
    for (int m = 0; m < PROF_MAX; ++m) p->acc[m] = 0.0;
    p->iters = 0;
    if (c->halo.dbg && prof_env) {   // boundary-block timestamps of the last profiled iteration (ns, relative to the first)
      unsigned long long h[8 * HALO_DBG_MAX];
      CK(cudaMemcpy(h, c->halo.dbg, sizeof(h), cudaMemcpyDeviceToHost));
      unsigned long long t0 = ~0ull;
      const int n = c->halo.dbg_n < HALO_DBG_MAX ? c->halo.dbg_n : HALO_DBG_MAX;
      for (int m = 0; m < 8 * n; ++m) if (h[m] && h[m] < t0) t0 = h[m];
      fprintf(stderr, "flipb200 rank %d: halo timestamps (us): launch | wait_dn start,end | wait_up start,end | computed | fenced | flag_dn | flag_up\n", c->rank);
      for (int m = 0; m < n; ++m) {
        double v[8];
        for (int q = 0; q < 8; ++q) v[q] = h[8 * m + q] ? (double)(h[8 * m + q] - t0) * 1e-3 : -1.0;
        fprintf(stderr, "  h%d %2d | %7.1f %7.1f | %7.1f %7.1f | %7.1f | %7.1f | %7.1f | %7.1f\n", c->rank, m, v[0], v[2], v[1], v[3], v[4], v[5], v[6], v[7]);
      }
    }
  }
  CK(cudaEventRecord(c->ev[7], c->stream));
  CK(cudaStreamSynchronize(c->stream));
  CK_LAUNCH();
  float e = 0.f;
  CK(cudaEventElapsedTime(&e, c->ev[6], c->ev[7]));
  *ms = e / iters;
  *units = (int64_t)(S.box.x1 - S.box.x0) * (S.box.y1 - S.box.y0) * (S.box.z1 - S.box.z0);
  return 0;
}
