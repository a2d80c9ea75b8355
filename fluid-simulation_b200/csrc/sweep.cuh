// sweep.cuh -- the two kernel shapes every grid-side solver kernel is built from.  Both were
// chosen by measurement on B200 (tools/kbench.cu, 256^3 fp64): all loads unconditional and issued
// before any use, results masked by the flag byte, two x-nodes per lane.
//
//  STREAM shape: block (32,8) = 8 rows x 64 nodes, one z-plane per blockIdx.z, one shot.  A pure
//                axpy-type update in this shape runs at the device-copy rate (6.7 TB/s counted on its
//                algorithmic bytes); the earlier z-marching tile loop with a flag test in front of
//                the loads reached 4.3 TB/s.
//  STENCIL shape: block (32,4), each lane owns 2 x-nodes, the block marches ST_ZP z-planes keeping
//                z-1/z/z+1 in registers; x-neighbours come from warp shuffles (lane 0/31 load the
//                two edge values), y-neighbours are vector loads that hit L1/L2.  7-point A.p runs
//                at 4.3 TB/s (65.7 us at 256^3) against 2.7 TB/s for one-thread-per-node.
#pragma once
#include "common.cuh"
#include <cuda_fp16.h>
#include <type_traits>

#define ST_BY 4
#define ST_ZP 4
#define SW_BY 8

template <typename T> struct Vec2T;
template <> struct Vec2T<double> { typedef double2 type; };
template <> struct Vec2T<float> { typedef float2 type; };

template <typename T> __device__ __forceinline__ typename Vec2T<T>::type ld2(const T* p) {
  return *reinterpret_cast<const typename Vec2T<T>::type*>(p);
}
template <typename T> __device__ __forceinline__ void st2(T* p, T a, T b) {
  typename Vec2T<T>::type v; v.x = a; v.y = b;
  *reinterpret_cast<typename Vec2T<T>::type*>(p) = v;
}
template <typename T> __device__ __forceinline__ typename Vec2T<T>::type zero2() {
  typename Vec2T<T>::type v; v.x = T(0); v.y = T(0); return v;
}

// Storage-type helpers: a vector of fp32 values kept in memory as fp32 or fp16 (FLIPB200_MG_H16, mg.cu).
template <typename T> __device__ __forceinline__ float4 ldq4(const T* p);
template <> __device__ __forceinline__ float4 ldq4<float>(const float* p) { return *reinterpret_cast<const float4*>(p); }
template <> __device__ __forceinline__ float4 ldq4<__half>(const __half* p) {
  const uint2 raw = *reinterpret_cast<const uint2*>(p);
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&raw.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
  return make_float4(a.x, a.y, b.x, b.y);
}
template <typename T> __device__ __forceinline__ void stq4(T* p, float4 v);
template <> __device__ __forceinline__ void stq4<float>(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
template <> __device__ __forceinline__ void stq4<__half>(__half* p, float4 v) {
  const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
  uint2 raw; raw.x = *reinterpret_cast<const unsigned*>(&a); raw.y = *reinterpret_cast<const unsigned*>(&b);
  *reinterpret_cast<uint2*>(p) = raw;
}
template <typename T> __device__ __forceinline__ void stq2(T* p, float a, float b);
template <> __device__ __forceinline__ void stq2<float>(float* p, float a, float b) { st2<float>(p, a, b); }
template <> __device__ __forceinline__ void stq2<__half>(__half* p, float a, float b) { *reinterpret_cast<__half2*>(p) = __floats2half2_rn(a, b); }
template <typename T> __device__ __forceinline__ float ldq1(const T* p);
template <> __device__ __forceinline__ float ldq1<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldq1<__half>(const __half* p) { return __half2float(*p); }

// geometry of one sweep over a box of a level
struct Sweep {
  Grid g;
  Box box;
  int xe;        // exclusive x limit for lanes (even, <= px)
  int zp;        // block order in z (zperm): 0 ascending, 1 boundary chunks first (z-slab runs), 2 descending
  int zc;        // planes a block of the STENCIL shapes marches (<= ST_ZP; ST_ZP on level 0, fewer on the latency-bound coarse levels)
};
#define ZP_UP 0
#define ZP_EDGE 1
#define ZP_DOWN 2
// W = x-nodes per lane (2, or 4 for the fp32 STENCIL4 shape): the lane range is aligned to W
static inline Sweep make_sweep(const Grid& g, const Box& b, int W = 2) {
  Sweep s; s.g = g; s.box = b; s.zp = (g.k0 > 0) ? 1 : 0;   // slabs have ghost planes below k0
  s.zc = ST_ZP;
  s.box.x0 = b.x0 & ~(W - 1);
  int xe = (b.x1 + W - 1) & ~(W - 1);
  s.xe = xe < g.px ? xe : g.px;
  return s;
}
static inline bool sweep_empty(const Sweep& s) {
  return !(s.box.x1 > s.box.x0 && s.box.y1 > s.box.y0 && s.box.z1 > s.box.z0);
}
static inline dim3 stream_grid(const Sweep& s) {
  return dim3(cdiv(s.xe - s.box.x0, 64), cdiv(s.box.y1 - s.box.y0, SW_BY), s.box.z1 - s.box.z0);
}
static inline dim3 stencil_grid(const Sweep& s, int W = 2) {
  return dim3(cdiv(s.xe - s.box.x0, 32 * W), cdiv(s.box.y1 - s.box.y0, ST_BY), cdiv(s.box.z1 - s.box.z0, s.zc));
}
// The coarse multigrid levels (rows < 128 nodes, 2-node shape) are latency-bound: a 64^3 sweep is 256 blocks of 4 warps
// on 148 SMs, each thread walking ST_ZP planes one dependent round trip after the other.  Fewer planes per block =
// more blocks and a shorter chain per thread.  FB_SMALL_ZC: chosen by A/B build.
#ifndef FB_SMALL_ZC
#define FB_SMALL_ZC 1
#endif
#ifndef FB_L1_ZC
#define FB_L1_ZC 4       // fp32 four-node sweeps of the levels >= 1 (128^3 below a 256^3 grid)
#endif
static inline int small_level_zc(const Sweep& s) {
  const long cells = (long)(s.xe - s.box.x0) * (s.box.y1 - s.box.y0) * (s.box.z1 - s.box.z0);
  return cells <= (1L << 19) ? FB_SMALL_ZC : ST_ZP;
}
#define STREAM_BLOCK dim3(32, SW_BY, 1)
#define STENCIL_BLOCK dim3(32, ST_BY, 1)

// Block z index -> plane (chunk) index: the first and the last chunk of the sweep are issued FIRST.
// z-slab runs push those boundary planes to the neighbours (halo.cuh); doing them at the start of
// the kernel hides the NVLink signalling latency behind the interior of the sweep.
// Only z-slab runs use it (Sweep::zp / the PeerHalo mode).
// ZP_DOWN issues the chunks from the top: on a single GPU consecutive sweeps of the PCG iteration
// alternate their direction (sweep_down() below), so each one starts on the planes the previous
// sweep left in the 126 MB L2 (a 256^3 level-0 sweep streams 150-950 MB).
__device__ __forceinline__ int zperm(int bz, int nz, int on) {
  return on == ZP_UP ? bz : (on == ZP_DOWN ? nz - 1 - bz : (bz == 0 ? 0 : (bz == 1 ? nz - 1 : bz - 1)));
}
// the same sweep issued from the top (no-op for z-slab sweeps and with FLIPB200_ZIGZAG=0)
static inline Sweep sweep_down(Sweep s) {
  static const bool on = [] { const char* e = getenv("FLIPB200_ZIGZAG"); return !(e && e[0] == '0'); }();
  if (on && s.zp == ZP_UP) s.zp = ZP_DOWN;
  return s;
}

// STREAM prologue: i (even x of this lane), j, k, q and `act`
#define STREAM_PROLOGUE(S)                                                      \
  const int i = (S).box.x0 + blockIdx.x * 64 + 2 * threadIdx.x;                 \
  const int j = (S).box.y0 + blockIdx.y * SW_BY + threadIdx.y;                  \
  const int k = (S).box.z0 + zperm(blockIdx.z, gridDim.z, (S).zp);              \
  const bool act = i < (S).xe && j < (S).box.y1;                                \
  const int q = gidx((S).g, i, j, k);

// Z-marching 7-point neighbourhood for the two nodes (i, i+1) of a lane.
// Usage:  StencilMarch<T> m; m.begin(S, src); for (z...) { m.load(S, src); ... m.sum0()/sum1()/c0()/c1(); m.advance(); }
template <typename T>
struct StencilMarch {
  typedef typename Vec2T<T>::type V;
  V pm, pc, pn, ym, yp;
  T left, right;
  int i, j, k, q, lane;
  bool act, jlo, jhi;

  __device__ __forceinline__ bool begin(const Sweep& S, const T* __restrict__ src) {
    lane = threadIdx.x;
    i = S.box.x0 + blockIdx.x * 64 + 2 * lane;
    j = S.box.y0 + blockIdx.y * ST_BY + threadIdx.y;
    k = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * S.zc;
    if (j >= S.box.y1) return false;                 // whole warp (one row) leaves together
    act = i < S.xe;
    jlo = j > 0; jhi = j < S.g.ny - 1;
    q = gidx(S.g, i, j, k);
    pm = (act && k > 0) ? ld2<T>(src + q - S.g.plane) : zero2<T>();
    pc = act ? ld2<T>(src + q) : zero2<T>();
    return true;
  }
  // issue every load of plane k (call before using the values)
  __device__ __forceinline__ void load(const Sweep& S, const T* __restrict__ src) {
    pn = (act && k < S.g.nz - 1) ? ld2<T>(src + q + S.g.plane) : zero2<T>();
    ym = (act && jlo) ? ld2<T>(src + q - S.g.px) : zero2<T>();
    yp = (act && jhi) ? ld2<T>(src + q + S.g.px) : zero2<T>();
    left = __shfl_up_sync(0xffffffffu, pc.y, 1);
    right = __shfl_down_sync(0xffffffffu, pc.x, 1);
    if (lane == 0) left = (act && i > 0) ? src[q - 1] : T(0);
    if (lane == 31) right = (act && i + 2 < S.g.px) ? src[q + 2] : T(0);
  }
  __device__ __forceinline__ T c0() const { return pc.x; }
  __device__ __forceinline__ T c1() const { return pc.y; }
  __device__ __forceinline__ T sum0() const { return (pm.x + pn.x) + (ym.x + yp.x) + (left + pc.y); }
  __device__ __forceinline__ T sum1() const { return (pm.y + pn.y) + (ym.y + yp.y) + (pc.x + right); }
  __device__ __forceinline__ void advance(const Sweep& S) { pm = pc; pc = pn; q += S.g.plane; ++k; }
};

// STENCIL4: the same march with FOUR fp32 x-nodes per lane (float4 loads, a warp covers 128 nodes).
// The fp32 multigrid sweeps of the big levels are issue-bound in the 2-node shape (80 % of the issue
// slots at 49 % DRAM, profiles/r1c_summary.txt); four nodes per lane halve the per-node overhead of
// the index math, shuffles and predicates.  Same association of the sums as StencilMarch: identical bits.
struct StencilMarch4 {
  float4 pm, pc, pn, ym, yp;
  float left, right;
  int i, j, k, q, lane;
  bool act, jlo, jhi;
  static __device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
  static __device__ __forceinline__ float4 zero4() { return make_float4(0.f, 0.f, 0.f, 0.f); }

  __device__ __forceinline__ bool begin(const Sweep& S, const float* __restrict__ src) {
    lane = threadIdx.x;
    i = S.box.x0 + blockIdx.x * 128 + 4 * lane;
    j = S.box.y0 + blockIdx.y * ST_BY + threadIdx.y;
    k = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * S.zc;
    if (j >= S.box.y1) return false;                 // whole warp (one row) leaves together
    act = i < S.xe;
    jlo = j > 0; jhi = j < S.g.ny - 1;
    q = gidx(S.g, i, j, k);
    pm = (act && k > 0) ? ld4(src + q - S.g.plane) : zero4();
    pc = act ? ld4(src + q) : zero4();
    return true;
  }
  // live = false (block-uniform: plane k lies beyond the block's chunk) turns every load off: a
  // march can predicate the planes of a partial last chunk this way instead of leaving the unrolled
  // loop with a break.  Measured slower (tools/kbench.cu T9: 64.7 vs 61.4 us); the kernels break.
  __device__ __forceinline__ void load(const Sweep& S, const float* __restrict__ src, bool live = true) {
    const bool a = act && live;
    pn = (a && k < S.g.nz - 1) ? ld4(src + q + S.g.plane) : zero4();
    ym = (a && jlo) ? ld4(src + q - S.g.px) : zero4();
    yp = (a && jhi) ? ld4(src + q + S.g.px) : zero4();
    left = __shfl_up_sync(0xffffffffu, pc.w, 1);
    right = __shfl_down_sync(0xffffffffu, pc.x, 1);
    if (lane == 0) left = (a && i > 0) ? src[q - 1] : 0.f;
    if (lane == 31) right = (a && i + 4 < S.g.px) ? src[q + 4] : 0.f;
  }
  __device__ __forceinline__ float sum0() const { return (pm.x + pn.x) + (ym.x + yp.x) + (left + pc.y); }
  __device__ __forceinline__ float sum1() const { return (pm.y + pn.y) + (ym.y + yp.y) + (pc.x + pc.z); }
  __device__ __forceinline__ float sum2() const { return (pm.z + pn.z) + (ym.z + yp.z) + (pc.y + pc.w); }
  __device__ __forceinline__ float sum3() const { return (pm.w + pn.w) + (ym.w + yp.w) + (pc.z + right); }
  __device__ __forceinline__ void advance(const Sweep& S) { pm = pc; pc = pn; q += S.g.plane; ++k; }
};

// fp64 STENCIL4: four doubles per lane as two double2 halves (a, b).  k_spmv in this shape sweeps
// 256^3 in 55.9 us against 65.6 us for two doubles per lane (tools/kbench.cu T7 vs T4).
// FULL = true: the caller guarantees that every lane of the block lies inside the sweep's x range
template <bool FULL>
struct StencilMarch4DT {
  double2 pma, pmb, pca, pcb, pna, pnb, yma, ymb, ypa, ypb;
  double left, right;
  const double* ptr;          // &src[q]: advanced by one plane per step (keeps the address math out of the loop)
  int i, j, k, q, lane;
  bool act, jlo, jhi, xlo, xhi;

  __device__ __forceinline__ bool begin(const Sweep& S, const double* __restrict__ src) {
    lane = threadIdx.x;
    i = S.box.x0 + blockIdx.x * 128 + 4 * lane;
    j = S.box.y0 + blockIdx.y * ST_BY + threadIdx.y;
    k = S.box.z0 + zperm(blockIdx.z, gridDim.z, S.zp) * ST_ZP;
    if (j >= S.box.y1) return false;                 // whole warp (one row) leaves together
    act = FULL || i < S.xe;
    jlo = act && j > 0; jhi = act && j < S.g.ny - 1;
    xlo = lane == 0 && act && i > 0; xhi = lane == 31 && act && i + 4 < S.g.px;
    q = gidx(S.g, i, j, k);
    ptr = src + q;
    const bool lo = act && k > 0;
    const ptrdiff_t P = S.g.plane;
    pma = lo ? ld2<double>(ptr - P) : zero2<double>();
    pmb = lo ? ld2<double>(ptr - P + 2) : zero2<double>();
    pca = act ? ld2<double>(ptr) : zero2<double>();
    pcb = act ? ld2<double>(ptr + 2) : zero2<double>();
    return true;
  }
  __device__ __forceinline__ void load(const Sweep& S, const double* __restrict__, bool live = true) {
    const bool hi = live && act && k < S.g.nz - 1, lo_y = live && jlo, hi_y = live && jhi;
    const ptrdiff_t P = S.g.plane, X = S.g.px;
    pna = hi ? ld2<double>(ptr + P) : zero2<double>();
    pnb = hi ? ld2<double>(ptr + P + 2) : zero2<double>();
    yma = lo_y ? ld2<double>(ptr - X) : zero2<double>();
    ymb = lo_y ? ld2<double>(ptr - X + 2) : zero2<double>();
    ypa = hi_y ? ld2<double>(ptr + X) : zero2<double>();
    ypb = hi_y ? ld2<double>(ptr + X + 2) : zero2<double>();
    left = __shfl_up_sync(0xffffffffu, pcb.y, 1);
    right = __shfl_down_sync(0xffffffffu, pca.x, 1);
    if (lane == 0) left = (live && xlo) ? ptr[-1] : 0.0;
    if (lane == 31) right = (live && xhi) ? ptr[4] : 0.0;
  }
  __device__ __forceinline__ double sum0() const { return (pma.x + pna.x) + (yma.x + ypa.x) + (left + pca.y); }
  __device__ __forceinline__ double sum1() const { return (pma.y + pna.y) + (yma.y + ypa.y) + (pca.x + pcb.x); }
  __device__ __forceinline__ double sum2() const { return (pmb.x + pnb.x) + (ymb.x + ypb.x) + (pca.y + pcb.y); }
  __device__ __forceinline__ double sum3() const { return (pmb.y + pnb.y) + (ymb.y + ypb.y) + (pcb.x + right); }
  __device__ __forceinline__ void advance(const Sweep& S) { pma = pca; pmb = pcb; pca = pna; pcb = pnb; q += S.g.plane; ptr += S.g.plane; ++k; }
};
typedef StencilMarch4DT<false> StencilMarch4D;

// block reduction for the (32, BY) shapes: thread (0,0) ends with the sum
template <int NV>
__device__ __forceinline__ void block_sum_xy(double (&v)[NV], double* smem /* >= 32*NV */) {
  block_sum<NV>(v, smem);
}
