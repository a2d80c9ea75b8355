// mg_h16.cu -- params.mixed bit 1 (default ON since round 2; see pcg.cuh).
// Level-0 V-cycle sweeps that keep the right-hand side b0 and the residual handed to the restriction
// in fp16 (arithmetic in fp32): copies of k_mg_smooth4 / k_mg_resid4 / k_mg_restrict4 of mg.cu with the
// loads / stores of those two arrays swapped.  They live in their own translation unit, behind host
// wrappers, so that the device code of the measured default kernels stays byte-identical (adding them
// to mg.cu moved its register allocation); merge them into templates once the variant has been measured.
// Prototype: tools/mg_proto7.py (full tank 128^3: 21 CG iterations with fp32 and with fp16 storage of b0 and
// the residual; fp16 iterates cost 2 iterations, bf16 anything more).
#include "sweep.cuh"
#include "pcg.cuh"
#include "halo.cuh"

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
template <int DOT, bool PEER>
__global__ void __launch_bounds__(32 * ST_BY, FB_MINB_SMOOTH4H(DOT, PEER))
k_mg_smooth4h(const __grid_constant__ Sweep S, const __grid_constant__ MgCoef mc, const uint8_t* __restrict__ fl,
             const __half* __restrict__ b, const float* __restrict__ xi, float* __restrict__ xo,
             const double* __restrict__ r64, DevScalars* __restrict__ ds, int it_arg, int dist,
             const __grid_constant__ PeerHalo H, float rzs /* 1 / bscale^2 */) {
  __shared__ double sm[32];
  if (ds->pcg_done) return;
  const int it = pcg_iter(ds, it_arg);
  const int kz0 = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * ST_ZP, kz1 = min(kz0 + ST_ZP, S.box.z1);
  const bool bdn = kz0 <= H.kdn && H.kdn < kz1, bup = kz0 <= H.kup && H.kup < kz1;
  if (PEER) halo_wait(H, ds, bdn, bup);
  double acc[1] = {0.0};
  StencilMarch4 m;
  if (m.begin(S, xi)) {
    const int kend = min(m.k + ST_ZP, S.box.z1);
#pragma unroll
    for (int z = 0; z < ST_ZP; ++z) {
      if (m.k >= kend) break;
      m.load(S, xi);
      uint32_t f = 0;
      float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
      double2 ra = make_double2(0.0, 0.0), rb = ra;
      if (m.act) {
        f = *reinterpret_cast<const uint32_t*>(fl + m.q);
        bv = ldq4<__half>(b + m.q);
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
      atomicAdd(dist ? &ds->part[PART_RZ] : &ds->rz[it & 3], acc[0] * (rrp > 0.0 ? (DOT == 2 ? rrp : sqrt(rrp)) : 1.0) * (double)rzs);   // b0 and z both carry bscale (DOT 2)
    }
  }
}
template <bool PEER>
__global__ void __launch_bounds__(32 * ST_BY, FB_MINB_RESID4H)
k_mg_resid4h(const __grid_constant__ Sweep S, const __grid_constant__ MgCoef mc, const uint8_t* __restrict__ fl,
            const uint8_t* __restrict__ nrm, const __half* __restrict__ b, const float* __restrict__ x,
            __half* __restrict__ rn, DevScalars* __restrict__ ds, const __grid_constant__ PeerHalo H) {
  if (ds->pcg_done) return;
  const int kz0 = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * ST_ZP, kz1 = min(kz0 + ST_ZP, S.box.z1);
  const bool bdn = kz0 <= H.kdn && H.kdn < kz1, bup = kz0 <= H.kup && H.kup < kz1;
  if (PEER) halo_wait(H, ds, bdn, bup);
  StencilMarch4 m;
  if (m.begin(S, x)) {
    const int kend = min(m.k + ST_ZP, S.box.z1);
#pragma unroll
    for (int z = 0; z < ST_ZP; ++z) {
      if (m.k >= kend) break;
      m.load(S, x);
      uint32_t f = 0, nn = 0x40404040u;
      float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m.act) {
        f = *reinterpret_cast<const uint32_t*>(fl + m.q);
        nn = *reinterpret_cast<const uint32_t*>(nrm + m.q);
        bv = ldq4<__half>(b + m.q);
      }
      float4 o;
      o.x = resid_node(mc, f & 0xffu, nn & 0xffu, m.pc.x, bv.x, m.sum0());
      o.y = resid_node(mc, (f >> 8) & 0xffu, (nn >> 8) & 0xffu, m.pc.y, bv.y, m.sum1());
      o.z = resid_node(mc, (f >> 16) & 0xffu, (nn >> 16) & 0xffu, m.pc.z, bv.z, m.sum2());
      o.w = resid_node(mc, f >> 24, nn >> 24, m.pc.w, bv.w, m.sum3());
      if (m.act) { stq4<__half>(rn + m.q, o); if (PEER) halo_push4h(H, m.q, m.k, o); }
      m.advance(S);
    }
  }
  if (PEER) halo_signal(H, ds, bdn, bup);
}

__global__ void __launch_bounds__(256, FB_MINB_RESTRICT4H)
k_mg_restrict4h(Grid gf, Box fbox, Grid gc, Box cbox, MgCoef mcc, const uint8_t* __restrict__ fc,
               const __half* __restrict__ rn, float* __restrict__ bc, float* __restrict__ x1c,
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
        const __half* row = rn + gidx(gf, 0, jr[t], kc);
        v[t] = ldq4<__half>(row + xc);
        ea[t] = eL ? __half2float(row[xf - 1]) : 0.f;
        ed[t] = eR ? __half2float(row[xf + 4]) : 0.f;
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


// host wrappers; H.mode != 0 on z-slab runs (ghost planes pushed by the producing kernel, halo.cuh)
void fb_h16_smooth4(flipb200_ctx* c, int dot, dim3 g, const Sweep& S, const MgCoef& m, const uint8_t* fl, const float* b0,
                    const float* xi, float* xo, int it, float rzs, int dist, const PeerHalo& H) {
  const __half* b = reinterpret_cast<const __half*>(b0);
  if (H.mode) {
    if (dot == 2) k_mg_smooth4h<2, true><<<g, STENCIL_BLOCK, 0, c->stream>>>(S, m, fl, b, xi, xo, nullptr, c->ds, it, dist, H, rzs);
    else k_mg_smooth4h<0, true><<<g, STENCIL_BLOCK, 0, c->stream>>>(S, m, fl, b, xi, xo, nullptr, c->ds, it, dist, H, rzs);
  } else {
    if (dot == 2) k_mg_smooth4h<2, false><<<g, STENCIL_BLOCK, 0, c->stream>>>(S, m, fl, b, xi, xo, nullptr, c->ds, it, dist, H, rzs);
    else k_mg_smooth4h<0, false><<<g, STENCIL_BLOCK, 0, c->stream>>>(S, m, fl, b, xi, xo, nullptr, c->ds, it, dist, H, rzs);
  }
}
void fb_h16_resid4(flipb200_ctx* c, dim3 g, const Sweep& S, const MgCoef& m, const uint8_t* fl, const uint8_t* nrm, const float* b0,
                   const float* x, float* rn, const PeerHalo& H) {
  if (H.mode) k_mg_resid4h<true><<<g, STENCIL_BLOCK, 0, c->stream>>>(S, m, fl, nrm, reinterpret_cast<const __half*>(b0), x, reinterpret_cast<__half*>(rn), c->ds, H);
  else k_mg_resid4h<false><<<g, STENCIL_BLOCK, 0, c->stream>>>(S, m, fl, nrm, reinterpret_cast<const __half*>(b0), x, reinterpret_cast<__half*>(rn), c->ds, H);
}
void fb_h16_restrict4(flipb200_ctx* c, dim3 g, dim3 blk, const Grid& gf, const Box& fbox, const Grid& gc, const Box& cbox, const MgCoef& mcc,
                      const uint8_t* fc, const float* rn, float* bc, float* x1c, int zc, const PeerHalo& H, const GatherPush& G) {
  k_mg_restrict4h<<<g, blk, 0, c->stream>>>(gf, fbox, gc, cbox, mcc, fc, reinterpret_cast<const __half*>(rn), bc, x1c, c->ds, zc, H, G);
}
