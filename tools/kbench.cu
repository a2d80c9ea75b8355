// kbench.cu -- design notebook (not product): which kernel shapes reach HBM peak on B200 for the
// two patterns the pressure solver is made of: (S) pure streaming axpy-type updates and (T) 7-point
// stencil sweeps, on 256^3 fp64 arrays with a 1-byte flag array.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/kbench tools/kbench.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
#include <vector>
// T9 replays the product's k_spmv4 on this notebook's data: the product headers give the same march
#include "../fluid-simulation_b200/csrc/halo.cuh"
#include "../fluid-simulation_b200/csrc/pcg.cuh"
#undef CK
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

static int N = 256;
__device__ int g_spread = 32;   // partial sums go to out[(bx+by+bz) % g_spread]
static size_t T;
static cudaEvent_t e0, e1;

template <typename F> float timeit(F f, int iters = 20) {
  for (int i = 0; i < 3; ++i) f();
  CK(cudaEventRecord(e0));
  for (int i = 0; i < iters; ++i) f();
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
  CK(cudaGetLastError());
  return ms / iters * 1e3f;   // us
}

// ---------------- streaming: x += a p ; r -= a Ap ; rr += r.r  (49 B/node) ----------------
__global__ void s1_scalar(size_t n, const uint8_t* __restrict__ fl, double* __restrict__ x, double* __restrict__ r,
                          const double* __restrict__ p, const double* __restrict__ Ap, double a, double* out) {
  double acc = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    uint8_t f = fl[i];
    if (f & 64) { double xn = x[i] + a * p[i], rn = r[i] - a * Ap[i]; x[i] = xn; r[i] = rn; acc += rn * rn; }
  }
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(~0u, acc, o);
  if ((threadIdx.x & 31) == 0 && acc != 0) atomicAdd(out, acc);
}
// unconditional loads, double2, UNROLL independent vectors in flight
template <int U>
__global__ void s2_vec(size_t n2, const uchar2* __restrict__ fl, double2* __restrict__ x, double2* __restrict__ r,
                       const double2* __restrict__ p, const double2* __restrict__ Ap, double a, double* out) {
  double acc = 0;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n2; i += stride * U) {
    uchar2 f[U]; double2 xv[U], rv[U], pv[U], av[U];
#pragma unroll
    for (int u = 0; u < U; ++u) { size_t k = i + u * stride; if (k < n2) { f[u] = fl[k]; xv[u] = x[k]; rv[u] = r[k]; pv[u] = p[k]; av[u] = Ap[k]; } }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      size_t k = i + u * stride;
      if (k < n2) {
        bool f0 = f[u].x & 64, f1 = f[u].y & 64;
        double2 xn = xv[u], rn = rv[u];
        if (f0) { xn.x += a * pv[u].x; rn.x -= a * av[u].x; acc += rn.x * rn.x; }
        if (f1) { xn.y += a * pv[u].y; rn.y -= a * av[u].y; acc += rn.y * rn.y; }
        if (f0 | f1) { x[k] = xn; r[k] = rn; }
      }
    }
  }
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(~0u, acc, o);
  if ((threadIdx.x & 31) == 0 && acc != 0) atomicAdd(out, acc);
}
__global__ void copy2(size_t n2, const double2* __restrict__ a, double2* __restrict__ b) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n2; i += (size_t)gridDim.x * blockDim.x) b[i] = a[i];
}

// ---------------- stencil: Ap = A p, pAp (17 B/node) ----------------
// T3: one thread per node, block (BX,BY), grid.z = planes; all 7 loads straight from L1/L2
template <int BX, int BY>
__global__ void t3_node(int n, const uint8_t* __restrict__ fl, const double* __restrict__ p, double* __restrict__ Ap, double* out) {
  int i = blockIdx.x * BX + threadIdx.x, j = blockIdx.y * BY + threadIdx.y, k = blockIdx.z;
  int plane = n * n, q = k * plane + j * n + i;
  double acc = 0;
  uint8_t f = fl[q];
  if (f & 64) {
    int kk = 6 - __popc(f & 63);
    double pc = p[q];
    double s = (p[q - 1] + p[q + 1]) + (p[q - n] + p[q + n]) + (p[q - plane] + p[q + plane]);
    double ap = kk * 0.1 * pc - 0.1 * s;
    Ap[q] = ap; acc = pc * ap;
  }
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(~0u, acc, o);
  __shared__ double sm[32];
  int tid = threadIdx.y * BX + threadIdx.x;
  if ((tid & 31) == 0) sm[tid >> 5] = acc;
  __syncthreads();
  if (tid == 0) { double t = 0; for (int w = 0; w < BX * BY / 32; ++w) t += sm[w]; if (t != 0) atomicAdd(out + (blockIdx.x + blockIdx.y + blockIdx.z) % g_spread, t); }
}
// T4: 2 nodes per thread (double2), x neighbours by shuffle, ZP planes per block marched with
// registers; unconditional loads
template <int BY, int ZP>
__global__ void t4_vec2(int n, const uint8_t* __restrict__ fl, const double* __restrict__ p, double* __restrict__ Ap, double* out) {
  const int lane = threadIdx.x;
  int i = blockIdx.x * 64 + 2 * lane, j = blockIdx.y * BY + threadIdx.y, k0 = blockIdx.z * ZP;
  int plane = n * n;
  int q = k0 * plane + j * n + i;
  const bool jlo = j > 0, jhi = j < n - 1;
  double acc = 0;
  double2 pm = k0 > 0 ? *(const double2*)(p + q - plane) : make_double2(0, 0);
  double2 pc = *(const double2*)(p + q);
#pragma unroll
  for (int z = 0; z < ZP; ++z, q += plane) {
    int k = k0 + z;
    double2 pn = k < n - 1 ? *(const double2*)(p + q + plane) : make_double2(0, 0);
    double2 ym = jlo ? *(const double2*)(p + q - n) : make_double2(0, 0);
    double2 yp = jhi ? *(const double2*)(p + q + n) : make_double2(0, 0);
    uchar2 f = *(const uchar2*)(fl + q);
    double left = __shfl_up_sync(~0u, pc.y, 1), right = __shfl_down_sync(~0u, pc.x, 1);
    if (lane == 0) left = i > 0 ? p[q - 1] : 0.0;
    if (lane == 31) right = i + 2 < n ? p[q + 2] : 0.0;
    double2 ap = make_double2(0, 0);
    if (f.x & 64) { int kk = 6 - __popc(f.x & 63); ap.x = kk * 0.1 * pc.x - 0.1 * ((pm.x + pn.x) + (ym.x + yp.x) + (left + pc.y)); acc += pc.x * ap.x; }
    if (f.y & 64) { int kk = 6 - __popc(f.y & 63); ap.y = kk * 0.1 * pc.y - 0.1 * ((pm.y + pn.y) + (ym.y + yp.y) + (pc.x + right)); acc += pc.y * ap.y; }
    if ((f.x | f.y) & 64) *(double2*)(Ap + q) = ap;
    pm = pc; pc = pn;
  }
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(~0u, acc, o);
  __shared__ double sm[32];
  if (lane == 0) sm[threadIdx.y] = acc;
  __syncthreads();
  if (lane == 0 && threadIdx.y == 0) { double t = 0; for (int w = 0; w < BY; ++w) t += sm[w]; if (t != 0) atomicAdd(out + (blockIdx.x + blockIdx.y + blockIdx.z) % g_spread, t); }
}

// T5: T4 with the centre column fetched PD = 2 planes ahead and the flags one plane ahead
template <int BY, int ZP>
__global__ void t5_pref(int n, const uint8_t* __restrict__ fl, const double* __restrict__ p, double* __restrict__ Ap, double* out) {
  const int lane = threadIdx.x;
  int i = blockIdx.x * 64 + 2 * lane, j = blockIdx.y * BY + threadIdx.y, k0 = blockIdx.z * ZP;
  int plane = n * n;
  int q = k0 * plane + j * n + i;
  const bool jlo = j > 0, jhi = j < n - 1;
  double acc = 0;
  const double2 z2 = make_double2(0, 0);
  double2 pm = k0 > 0 ? *(const double2*)(p + q - plane) : z2;
  double2 pc = *(const double2*)(p + q);
  double2 pn = k0 < n - 1 ? *(const double2*)(p + q + plane) : z2;
  uchar2 f = *(const uchar2*)(fl + q);
#pragma unroll
  for (int z = 0; z < ZP; ++z, q += plane) {
    int k = k0 + z;
    double2 pnn = (z + 1 < ZP && k + 2 < n) ? *(const double2*)(p + q + 2 * plane) : z2;
    uchar2 fn = (z + 1 < ZP) ? *(const uchar2*)(fl + q + plane) : make_uchar2(0, 0);
    double2 ym = jlo ? *(const double2*)(p + q - n) : z2;
    double2 yp = jhi ? *(const double2*)(p + q + n) : z2;
    double left = __shfl_up_sync(~0u, pc.y, 1), right = __shfl_down_sync(~0u, pc.x, 1);
    if (lane == 0) left = i > 0 ? p[q - 1] : 0.0;
    if (lane == 31) right = i + 2 < n ? p[q + 2] : 0.0;
    double2 ap = z2;
    if (f.x & 64) { int kk = 6 - __popc(f.x & 63); ap.x = kk * 0.1 * pc.x - 0.1 * ((pm.x + pn.x) + (ym.x + yp.x) + (left + pc.y)); acc += pc.x * ap.x; }
    if (f.y & 64) { int kk = 6 - __popc(f.y & 63); ap.y = kk * 0.1 * pc.y - 0.1 * ((pm.y + pn.y) + (ym.y + yp.y) + (pc.x + right)); acc += pc.y * ap.y; }
    if ((f.x | f.y) & 64) *(double2*)(Ap + q) = ap;
    pm = pc; pc = pn; pn = pnn; f = fn;
  }
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(~0u, acc, o);
  __shared__ double sm[32];
  if (lane == 0) sm[threadIdx.y] = acc;
  __syncthreads();
  if (lane == 0 && threadIdx.y == 0) { double t = 0; for (int w = 0; w < BY; ++w) t += sm[w]; if (t != 0) atomicAdd(out + (blockIdx.x + blockIdx.y + blockIdx.z) % g_spread, t); }
}

// T7: T4 with FOUR doubles per lane (a warp covers 128 nodes)
template <int BY, int ZP>
__global__ void __launch_bounds__(32 * BY) t7_vec4(int n, const uint8_t* __restrict__ fl, const double* __restrict__ p, double* __restrict__ Ap, double* out) {
  const int lane = threadIdx.x;
  int i = blockIdx.x * 128 + 4 * lane, j = blockIdx.y * BY + threadIdx.y, k0 = blockIdx.z * ZP;
  int plane = n * n;
  int q = k0 * plane + j * n + i;
  const bool jlo = j > 0, jhi = j < n - 1;
  double acc = 0;
  const double2 z2 = make_double2(0, 0);
  double2 pma = k0 > 0 ? *(const double2*)(p + q - plane) : z2, pmb = k0 > 0 ? *(const double2*)(p + q - plane + 2) : z2;
  double2 pca = *(const double2*)(p + q), pcb = *(const double2*)(p + q + 2);
#pragma unroll
  for (int z = 0; z < ZP; ++z, q += plane) {
    int k = k0 + z;
    double2 pna = k < n - 1 ? *(const double2*)(p + q + plane) : z2, pnb = k < n - 1 ? *(const double2*)(p + q + plane + 2) : z2;
    double2 yma = jlo ? *(const double2*)(p + q - n) : z2, ymb = jlo ? *(const double2*)(p + q - n + 2) : z2;
    double2 ypa = jhi ? *(const double2*)(p + q + n) : z2, ypb = jhi ? *(const double2*)(p + q + n + 2) : z2;
    uint32_t f = *(const uint32_t*)(fl + q);
    double left = __shfl_up_sync(~0u, pcb.y, 1), right = __shfl_down_sync(~0u, pca.x, 1);
    if (lane == 0) left = i > 0 ? p[q - 1] : 0.0;
    if (lane == 31) right = i + 4 < n ? p[q + 4] : 0.0;
    double2 a0 = z2, a1 = z2;
    if (f & 64) { int kk = 6 - __popc(f & 63); a0.x = kk * 0.1 * pca.x - 0.1 * ((pma.x + pna.x) + (yma.x + ypa.x) + (left + pca.y)); acc += pca.x * a0.x; }
    if ((f >> 8) & 64) { int kk = 6 - __popc((f >> 8) & 63); a0.y = kk * 0.1 * pca.y - 0.1 * ((pma.y + pna.y) + (yma.y + ypa.y) + (pca.x + pcb.x)); acc += pca.y * a0.y; }
    if ((f >> 16) & 64) { int kk = 6 - __popc((f >> 16) & 63); a1.x = kk * 0.1 * pcb.x - 0.1 * ((pmb.x + pnb.x) + (ymb.x + ypb.x) + (pca.y + pcb.y)); acc += pcb.x * a1.x; }
    if ((f >> 24) & 64) { int kk = 6 - __popc((f >> 24) & 63); a1.y = kk * 0.1 * pcb.y - 0.1 * ((pmb.y + pnb.y) + (ymb.y + ypb.y) + (pcb.x + right)); acc += pcb.y * a1.y; }
    if (f & 0x40404040u) { *(double2*)(Ap + q) = a0; *(double2*)(Ap + q + 2) = a1; }
    pma = pca; pmb = pcb; pca = pna; pcb = pnb;
  }
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(~0u, acc, o);
  __shared__ double sm[32];
  if (lane == 0) sm[threadIdx.y] = acc;
  __syncthreads();
  if (lane == 0 && threadIdx.y == 0) { double t = 0; for (int w = 0; w < BY; ++w) t += sm[w]; if (t != 0) atomicAdd(out + (blockIdx.x + blockIdx.y + blockIdx.z) % g_spread, t); }
}

// T9: the product kernel k_spmv4 (csrc/pcg.cu) with its features switchable, to find what separates it
// from T7.  FL bit 0: leave a partial chunk with a break; bit 1: diagonal from the coefficient table
// (else kk * 0.1); bit 2: two predicated half stores (else one test for the four nodes); bit 3: the
// ghost-plane wait of z-slab runs compiled in; bit 4: done flag / iteration read from DevScalars;
// bit 5: no x-range predicate (every lane inside the sweep); bit 6: T7's plain block reduction.
template <int FL>
__global__ void __launch_bounds__(32 * ST_BY)
t9_prod(const __grid_constant__ Sweep S, const __grid_constant__ PcgCoef cf, const uint8_t* __restrict__ fl,
        const double* __restrict__ p, double* __restrict__ Ap, DevScalars* __restrict__ ds, int it_arg, int noexit,
        int dist, const __grid_constant__ PeerHalo H) {
  __shared__ double sm[32];
  if (FL & 16) { if (!noexit && ds->pcg_done) return; }
  const int it = (FL & 16) ? pcg_iter(ds, it_arg) : it_arg;
  if (FL & 8) {
    const int kz0 = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * ST_ZP, kz1 = min(kz0 + ST_ZP, S.box.z1);
    halo_wait(H, ds, kz0 <= H.kdn && H.kdn < kz1, kz0 <= H.kup && H.kup < kz1);
  }
  double acc[1] = {0.0};
  StencilMarch4DT<(FL & 32) != 0> m;
  if (m.begin(S, p)) {
    const int kend = min(m.k + ST_ZP, S.box.z1);
#pragma unroll
    for (int z = 0; z < ST_ZP; ++z) {
      if (FL & 1) { if (m.k >= kend) break; }
      const bool live = (FL & 1) || m.k < kend;
      m.load(S, p, live);
      const uint32_t f = (m.act && live) ? *reinterpret_cast<const uint32_t*>(fl + m.q) : 0u;
      const uint32_t f0 = f & 0xffu, f1 = (f >> 8) & 0xffu, f2 = (f >> 16) & 0xffu, f3 = f >> 24;
      double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#define T9_DIAG(fx) ((FL & 2) ? cf.diag[6 - __popc((fx) & FB_SMASK)] : (6 - __popc((fx) & FB_SMASK)) * 0.1)
      if (f0 & FB_FLUID) { a0 = T9_DIAG(f0) * m.pca.x + cf.off * m.sum0(); acc[0] += m.pca.x * a0; }
      if (f1 & FB_FLUID) { a1 = T9_DIAG(f1) * m.pca.y + cf.off * m.sum1(); acc[0] += m.pca.y * a1; }
      if (f2 & FB_FLUID) { a2 = T9_DIAG(f2) * m.pcb.x + cf.off * m.sum2(); acc[0] += m.pcb.x * a2; }
      if (f3 & FB_FLUID) { a3 = T9_DIAG(f3) * m.pcb.y + cf.off * m.sum3(); acc[0] += m.pcb.y * a3; }
      if (FL & 4) {
        if ((f0 | f1) & FB_FLUID) st2<double>(Ap + m.q, a0, a1);
        if ((f2 | f3) & FB_FLUID) st2<double>(Ap + m.q + 2, a2, a3);
      } else if ((f0 | f1 | f2 | f3) & FB_FLUID) { st2<double>(Ap + m.q, a0, a1); st2<double>(Ap + m.q + 2, a2, a3); }
      m.advance(S);
    }
  }
  if (FL & 64) {
    double a = acc[0];
    for (int o = 16; o; o >>= 1) a += __shfl_xor_sync(~0u, a, o);
    if (threadIdx.x == 0) sm[threadIdx.y] = a;
    __syncthreads();
    if (threadIdx.x == 0 && threadIdx.y == 0) { double t = 0; for (int w = 0; w < ST_BY; ++w) t += sm[w]; if (t != 0) atomicAdd(&ds->pAp[it & 3], t); }
    return;
  }
  block_sum<1>(acc, sm);
  if (threadIdx.x == 0 && threadIdx.y == 0 && acc[0] != 0.0) atomicAdd(dist ? &ds->part[PART_PAP] : &ds->pAp[it & 3], acc[0]);
}

// T8: T7 + prefetch.global.L2 of every plane of the march at block start (PF = 1) or two planes ahead (PF = 2)
template <int BY, int ZP, int PF>
__global__ void __launch_bounds__(32 * BY) t8_pf(int n, const uint8_t* __restrict__ fl, const double* __restrict__ p, double* __restrict__ Ap, double* out) {
  const int lane = threadIdx.x;
  int i = blockIdx.x * 128 + 4 * lane, j = blockIdx.y * BY + threadIdx.y, k0 = blockIdx.z * ZP;
  int plane = n * n;
  int q = k0 * plane + j * n + i;
  const bool jlo = j > 0, jhi = j < n - 1;
  double acc = 0;
  const double2 z2 = make_double2(0, 0);
  double2 pma = k0 > 0 ? *(const double2*)(p + q - plane) : z2, pmb = k0 > 0 ? *(const double2*)(p + q - plane + 2) : z2;
  double2 pca = *(const double2*)(p + q), pcb = *(const double2*)(p + q + 2);
  if (PF == 1) {
#pragma unroll
    for (int z = 1; z <= ZP; ++z)
      if (k0 + z < n) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + q + (size_t)z * plane));
  }
#pragma unroll
  for (int z = 0; z < ZP; ++z, q += plane) {
    int k = k0 + z;
    if (PF == 2 && z + 2 <= ZP && k + 2 < n) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + q + 2 * (size_t)plane));
    double2 pna = k < n - 1 ? *(const double2*)(p + q + plane) : z2, pnb = k < n - 1 ? *(const double2*)(p + q + plane + 2) : z2;
    double2 yma = jlo ? *(const double2*)(p + q - n) : z2, ymb = jlo ? *(const double2*)(p + q - n + 2) : z2;
    double2 ypa = jhi ? *(const double2*)(p + q + n) : z2, ypb = jhi ? *(const double2*)(p + q + n + 2) : z2;
    uint32_t f = *(const uint32_t*)(fl + q);
    double left = __shfl_up_sync(~0u, pcb.y, 1), right = __shfl_down_sync(~0u, pca.x, 1);
    if (lane == 0) left = i > 0 ? p[q - 1] : 0.0;
    if (lane == 31) right = i + 4 < n ? p[q + 4] : 0.0;
    double2 a0 = z2, a1 = z2;
    if (f & 64) { int kk = 6 - __popc(f & 63); a0.x = kk * 0.1 * pca.x - 0.1 * ((pma.x + pna.x) + (yma.x + ypa.x) + (left + pca.y)); acc += pca.x * a0.x; }
    if ((f >> 8) & 64) { int kk = 6 - __popc((f >> 8) & 63); a0.y = kk * 0.1 * pca.y - 0.1 * ((pma.y + pna.y) + (yma.y + ypa.y) + (pca.x + pcb.x)); acc += pca.y * a0.y; }
    if ((f >> 16) & 64) { int kk = 6 - __popc((f >> 16) & 63); a1.x = kk * 0.1 * pcb.x - 0.1 * ((pmb.x + pnb.x) + (ymb.x + ypb.x) + (pca.y + pcb.y)); acc += pcb.x * a1.x; }
    if ((f >> 24) & 64) { int kk = 6 - __popc((f >> 24) & 63); a1.y = kk * 0.1 * pcb.y - 0.1 * ((pmb.y + pnb.y) + (ymb.y + ypb.y) + (pcb.x + right)); acc += pcb.y * a1.y; }
    if (f & 0x40404040u) { *(double2*)(Ap + q) = a0; *(double2*)(Ap + q + 2) = a1; }
    pma = pca; pmb = pcb; pca = pna; pcb = pnb;
  }
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(~0u, acc, o);
  __shared__ double sm[32];
  if (lane == 0) sm[threadIdx.y] = acc;
  __syncthreads();
  if (lane == 0 && threadIdx.y == 0) { double t = 0; for (int w = 0; w < BY; ++w) t += sm[w]; if (t != 0) atomicAdd(out + (blockIdx.x + blockIdx.y + blockIdx.z) % g_spread, t); }
}

// T6: planes staged in shared memory by TMA (box 68 x (BY+2) x 1 doubles, out-of-range -> 0), NST-deep
// ring with one mbarrier per slot; the block marches ZP planes.  Centre column rotates in registers.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int cnt) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(cnt)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* m, int c0, int c1, int c2, uint64_t* b) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
               ::"r"(smem_u32(dst)), "l"(m), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(b)) : "memory");
}
template <int BY, int ZP, int NST>
__global__ void __launch_bounds__(32 * BY)
t6_tma(const __grid_constant__ CUtensorMap mp, int n, const uint8_t* __restrict__ fl, double* __restrict__ Ap, double* out) {
  constexpr int TX = 68, TY = BY + 2;
  constexpr int STAGE = ((TX * TY * 8 + 127) / 128) * 128;
  extern __shared__ __align__(128) unsigned char smraw[];
  __shared__ __align__(8) uint64_t bar[NST];
  __shared__ double sm[32];
  const int lane = threadIdx.x, ty = threadIdx.y, tid = ty * 32 + lane;
  const int i0 = blockIdx.x * 64, j0 = blockIdx.y * BY, k0 = blockIdx.z * ZP;
  if (tid == 0) {
    for (int s = 0; s < NST; ++s) mbar_init(&bar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // plane t (0 .. ZP+1) is z = k0 - 1 + t, slot t % NST, parity (t / NST) & 1
  auto issue = [&](int t) {
    const int s = t % NST;
    mbar_expect(&bar[s], TX * TY * 8);
    tma_load_3d(smraw + s * STAGE, &mp, i0 - 2, j0 - 1, k0 - 1 + t, &bar[s]);
  };
  if (tid == 0)
    for (int t = 0; t < NST && t < ZP + 2; ++t) issue(t);
  const int i = i0 + 2 * lane, j = j0 + ty;
  const int plane = n * n;
  int q = k0 * plane + j * n + i;
  const int off = (ty + 1) * TX + 2 * lane + 2;             // this lane's pair inside a stage
  auto tile = [&](int t) -> const double* { return (const double*)(smraw + (t % NST) * STAGE); };
  mbar_wait(&bar[0], 0);
  double2 pm = *(const double2*)(tile(0) + off);
  mbar_wait(&bar[1 % NST], (1 / NST) & 1);
  double2 pc = *(const double2*)(tile(1) + off);
  uchar2 f = *(const uchar2*)(fl + q);
  double acc = 0;
#pragma unroll 1
  for (int z = 0; z < ZP; ++z, q += plane) {
    const int t = z + 1;                                     // plane of the centre
    uchar2 fn = (z + 1 < ZP) ? *(const uchar2*)(fl + q + plane) : make_uchar2(0, 0);
    mbar_wait(&bar[(t + 1) % NST], ((t + 1) / NST) & 1);
    const double* tc = tile(t);
    double2 pn = *(const double2*)(tile(t + 1) + off);
    double2 ym = *(const double2*)(tc + off - TX);
    double2 yp = *(const double2*)(tc + off + TX);
    double left = __shfl_up_sync(~0u, pc.y, 1), right = __shfl_down_sync(~0u, pc.x, 1);
    if (lane == 0) left = tc[off - 1];
    if (lane == 31) right = tc[off + 2];
    double2 ap = make_double2(0, 0);
    if (f.x & 64) { int kk = 6 - __popc(f.x & 63); ap.x = kk * 0.1 * pc.x - 0.1 * ((pm.x + pn.x) + (ym.x + yp.x) + (left + pc.y)); acc += pc.x * ap.x; }
    if (f.y & 64) { int kk = 6 - __popc(f.y & 63); ap.y = kk * 0.1 * pc.y - 0.1 * ((pm.y + pn.y) + (ym.y + yp.y) + (pc.x + right)); acc += pc.y * ap.y; }
    if ((f.x | f.y) & 64) *(double2*)(Ap + q) = ap;
    pm = pc; pc = pn; f = fn;
    __syncthreads();                                         // everyone is done with plane t-1 (slot (t-1) % NST)
    if (tid == 0 && t - 1 + NST < ZP + 2) issue(t - 1 + NST);
  }
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(~0u, acc, o);
  if (lane == 0) sm[ty] = acc;
  __syncthreads();
  if (tid == 0) { double t = 0; for (int w = 0; w < BY; ++w) t += sm[w]; if (t != 0) atomicAdd(out + (blockIdx.x + blockIdx.y + blockIdx.z) % g_spread, t); }
}

typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static CUtensorMap make_map(void* base, int n, int by) {
  void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
  CUtensorMap m;
  cuuint64_t dims[3] = {(cuuint64_t)n, (cuuint64_t)n, (cuuint64_t)n};
  cuuint64_t strides[2] = {(cuuint64_t)n * 8, (cuuint64_t)n * n * 8};
  cuuint32_t box[3] = {68, (cuuint32_t)by + 2, 1};
  cuuint32_t es[3] = {1, 1, 1};
  CUresult r = ((EncodeTiled)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                 CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("cuTensorMapEncodeTiled failed %d\n", (int)r); exit(1); }
  return m;
}

__global__ void fill_rand(size_t n, double* p) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    unsigned long long h = i * 0x9E3779B97F4A7C15ull; h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
    p[i] = (double)(h & 0xffffff) * (1.0 / 16777216.0) - 0.5;
  }
}
int main(int argc, char** argv) {
  if (argc > 1) N = atoi(argv[1]);
  T = (size_t)N * N * N;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  double *x, *r, *p, *Ap, *out; uint8_t* fl;
  CK(cudaMalloc(&x, T * 8)); CK(cudaMalloc(&r, T * 8)); CK(cudaMalloc(&p, T * 8)); CK(cudaMalloc(&Ap, T * 8));
  CK(cudaMalloc(&fl, T)); CK(cudaMalloc(&out, 64 * 8));
  CK(cudaMemset(x, 0, T * 8)); CK(cudaMemset(r, 0, T * 8)); CK(cudaMemset(p, 0, T * 8)); CK(cudaMemset(Ap, 0, T * 8));
  CK(cudaMemset(out, 0, 64 * 8));
  {
    std::vector<uint8_t> h(T, 0);
    for (int k = 0; k < N; ++k) for (int j = 0; j < N; ++j) for (int i = 0; i < N; ++i) {
      bool in = i >= 2 && j >= 2 && k >= 2 && i < N - 2 && j < N - 4 && k < N - 2;
      h[((size_t)k * N + j) * N + i] = in ? 64 : 128;
    }
    CK(cudaMemcpy(fl, h.data(), T, cudaMemcpyHostToDevice));
  }
  if (argc > 4) { fill_rand<<<148 * 8, 256>>>(T, p); CK(cudaDeviceSynchronize()); printf("p = random\n"); }
  int sms = 148;
  double gb49 = 49.0 * T / 1e9, gb17 = 17.0 * T / 1e9, gb16 = 16.0 * T / 1e9;
  float us;
  us = timeit([&] { CK(cudaMemcpyAsync(Ap, p, T * 8, cudaMemcpyDeviceToDevice)); });
  printf("memcpy d2d                 %8.1f us  %7.1f GB/s\n", us, gb16 / us * 1e6);
  for (int bps : {4, 8, 16}) {
    us = timeit([&] { copy2<<<sms * bps, 256>>>(T / 2, (double2*)p, (double2*)Ap); });
    printf("copy2 grid=%2dxSM           %8.1f us  %7.1f GB/s\n", bps, us, gb16 / us * 1e6);
  }
  for (int bps : {4, 8}) {
    us = timeit([&] { s1_scalar<<<sms * bps, 256>>>(T, fl, x, r, p, Ap, 1e-3, out); });
    printf("S1 scalar grid=%dxSM        %8.1f us  %7.1f GB/s (49B)\n", bps, us, gb49 / us * 1e6);
  }
  for (int bps : {2, 4, 8}) {
    us = timeit([&] { s2_vec<1><<<sms * bps, 256>>>(T / 2, (uchar2*)fl, (double2*)x, (double2*)r, (double2*)p, (double2*)Ap, 1e-3, out); });
    printf("S2 vec2 U1 grid=%dxSM       %8.1f us  %7.1f GB/s\n", bps, us, gb49 / us * 1e6);
    us = timeit([&] { s2_vec<2><<<sms * bps, 256>>>(T / 2, (uchar2*)fl, (double2*)x, (double2*)r, (double2*)p, (double2*)Ap, 1e-3, out); });
    printf("S2 vec2 U2 grid=%dxSM       %8.1f us  %7.1f GB/s\n", bps, us, gb49 / us * 1e6);
    us = timeit([&] { s2_vec<4><<<sms * bps, 256>>>(T / 2, (uchar2*)fl, (double2*)x, (double2*)r, (double2*)p, (double2*)Ap, 1e-3, out); });
    printf("S2 vec2 U4 grid=%dxSM       %8.1f us  %7.1f GB/s\n", bps, us, gb49 / us * 1e6);
  }
  us = timeit([&] { s2_vec<1><<<(unsigned)((T / 2 + 255) / 256), 256>>>(T / 2, (uchar2*)fl, (double2*)x, (double2*)r, (double2*)p, (double2*)Ap, 1e-3, out); });
  printf("S2 vec2 one-shot grid      %8.1f us  %7.1f GB/s\n", us, gb49 / us * 1e6);
  // stencils
  us = timeit([&] { t3_node<32, 8><<<dim3(N / 32, N / 8, N), dim3(32, 8)>>>(N, fl, p, Ap, out); });
  printf("T3 node 32x8               %8.1f us  %7.1f GB/s (17B)\n", us, gb17 / us * 1e6);
  us = timeit([&] { t3_node<64, 4><<<dim3(N / 64, N / 4, N), dim3(64, 4)>>>(N, fl, p, Ap, out); });
  printf("T3 node 64x4               %8.1f us  %7.1f GB/s\n", us, gb17 / us * 1e6);
  us = timeit([&] { t3_node<128, 2><<<dim3(N / 128, N / 2, N), dim3(128, 2)>>>(N, fl, p, Ap, out); });
  printf("T3 node 128x2              %8.1f us  %7.1f GB/s\n", us, gb17 / us * 1e6);
  us = timeit([&] { t3_node<32, 16><<<dim3(N / 32, N / 16, N), dim3(32, 16)>>>(N, fl, p, Ap, out); });
  printf("T3 node 32x16              %8.1f us  %7.1f GB/s\n", us, gb17 / us * 1e6);
#define RUN_T4(BY, ZP)                                                                                       \
  us = timeit([&] { t4_vec2<BY, ZP><<<dim3(N / 64, N / BY, N / ZP), dim3(32, BY)>>>(N, fl, p, Ap, out); }); \
  printf("T4 vec2 BY=%d ZP=%d          %8.1f us  %7.1f GB/s\n", BY, ZP, us, gb17 / us * 1e6);
#define RUN_T5(BY, ZP)                                                                                       \
  us = timeit([&] { t5_pref<BY, ZP><<<dim3(N / 64, N / BY, N / ZP), dim3(32, BY)>>>(N, fl, p, Ap, out); }); \
  printf("T5 pref BY=%d ZP=%d          %8.1f us  %7.1f GB/s\n", BY, ZP, us, gb17 / us * 1e6);
#define RUN_T6(BY, ZP, NST)                                                                                  \
  {                                                                                                          \
    CUtensorMap mp = make_map(p, N, BY);                                                                     \
    size_t smem = (size_t)NST * (((68 * (BY + 2) * 8 + 127) / 128) * 128);                                   \
    CK(cudaFuncSetAttribute(t6_tma<BY, ZP, NST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));   \
    us = timeit([&] { t6_tma<BY, ZP, NST><<<dim3(N / 64, N / BY, N / ZP), dim3(32, BY), smem>>>(mp, N, fl, Ap, out); }); \
    printf("T6 tma  BY=%d ZP=%d NST=%d   %8.1f us  %7.1f GB/s\n", BY, ZP, NST, us, gb17 / us * 1e6);       \
  }
  if (argc > 2) {   // stencil variants only
    RUN_T4(4, 4) RUN_T4(8, 4) RUN_T4(4, 8)
#define RUN_T7(BY, ZP)                                                                                       \
  us = timeit([&] { t7_vec4<BY, ZP><<<dim3(N / 128, N / BY, N / ZP), dim3(32, BY)>>>(N, fl, p, Ap, out); }); \
  printf("T7 vec4 BY=%d ZP=%d          %8.1f us  %7.1f GB/s\n", BY, ZP, us, gb17 / us * 1e6);
    RUN_T7(4, 4) RUN_T7(8, 4) RUN_T7(4, 8) RUN_T7(2, 4) RUN_T7(4, 2) RUN_T7(8, 2)
#define RUN_T8(BY, ZP, PF)                                                                                   \
  us = timeit([&] { t8_pf<BY, ZP, PF><<<dim3(N / 128, N / BY, N / ZP), dim3(32, BY)>>>(N, fl, p, Ap, out); }); \
  printf("T8 pf%d  BY=%d ZP=%d          %8.1f us  %7.1f GB/s\n", PF, BY, ZP, us, gb17 / us * 1e6);
    RUN_T8(4, 4, 1) RUN_T8(4, 8, 1) RUN_T8(4, 4, 2) RUN_T8(4, 8, 2) RUN_T8(4, 16, 2) RUN_T8(2, 8, 1)
    {   // T9: product kernel, product flag encoding (FB_FLUID = bit 6 here too), whole box or the product's 255-row box
      static_assert(FB_FLUID == 64, "flag encoding");
      DevScalars* ds; CK(cudaMalloc(&ds, sizeof(DevScalars))); CK(cudaMemset(ds, 0, sizeof(DevScalars)));
      PcgCoef cf; memset(&cf, 0, sizeof(cf));
      for (int k = 1; k <= 6; ++k) { cf.diag[k] = 0.1 * k; cf.inv[k] = 1.0 / cf.diag[k]; }
      cf.off = -0.1;
      Grid g; memset(&g, 0, sizeof(g));
      g.nx = N; g.ny = N; g.nz = N; g.px = N; g.plane = N * N; g.total = N * N * N; g.gnz = N; g.k1 = N;
      for (int yb = 0; yb < 1; ++yb) {
        Box b = {0, yb, 0, N, N, N};      // yb = 1: rows 1..N-1 as in the bench scene
        Sweep S4 = make_sweep(g, b, 4);
#define RUN_T9(FL) \
        us = timeit([&] { t9_prod<FL><<<stencil_grid(S4, 4), STENCIL_BLOCK>>>(S4, cf, fl, p, Ap, ds, 5, 1, 0, halo_off()); }); \
        printf("T9 prod FL=%2d y0=%d          %8.1f us  %7.1f GB/s\n", FL, yb, us, gb17 / us * 1e6);
        RUN_T9(1) RUN_T9(33) RUN_T9(65) RUN_T9(97) RUN_T9(23) RUN_T9(55) RUN_T9(119) RUN_T9(31)
      }
      RUN_T7(4, 4)
    }
    { int one = 1; CK(cudaMemcpyToSymbol(g_spread, &one, sizeof(int))); printf("single-address atomics:\n"); }
    RUN_T7(4, 4) RUN_T4(4, 4)
    { int v = 32; CK(cudaMemcpyToSymbol(g_spread, &v, sizeof(int))); }
    if (argc > 3) return 0;
    RUN_T5(4, 4) RUN_T5(8, 4) RUN_T5(4, 8) RUN_T5(8, 8) RUN_T5(4, 16) RUN_T5(8, 16)
    RUN_T6(8, 16, 4) RUN_T6(8, 32, 4) RUN_T6(8, 64, 4) RUN_T6(16, 32, 4) RUN_T6(16, 64, 4) RUN_T6(8, 32, 6) RUN_T6(4, 32, 4) RUN_T6(4, 64, 6) RUN_T6(8, 256, 4)
    return 0;
  }
  RUN_T4(4, 1) RUN_T4(8, 1) RUN_T4(4, 2) RUN_T4(8, 2) RUN_T4(4, 4) RUN_T4(8, 4) RUN_T4(4, 8) RUN_T4(8, 8) RUN_T4(16, 4) RUN_T4(4, 16)
  return 0;
}
