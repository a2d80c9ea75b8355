// noalloc.cu -- flipb200_p2g_noalloc: the particle -> grid transfer as ONE call that allocates nothing.
//
// No counterpart in the reference: callers that cannot call cudaMalloc
// -- code under CUDA graph capture, tight serving loops, harnesses that forbid it -- get the P2G + normalisation of
// flipb200_p2g / flipb200_classify (PointList::P2Gtransfer fluid.cc:1106-1148, p2gCatmullRom :265-299) on buffers
// they own.  The only scratch the transfer needs, the fp64 weight accumulator, lives in a file-scope __device__
// array of FLIPB200_NOALLOC_MAX_NODES doubles (16 MiB) that every GPU loading the library holds until exit.
//
// The array's NAME is a host-side stand-in in host code, not a device address; the device address comes from
// cudaGetSymbolAddress, once per device with that device current (na_scratch below).
// Every call shares the array and it keeps the last call's contents: each call clears the range it uses first, and two
// calls that overlap on the GPU (different streams, no dependency) race on it.
#include "common.cuh"
#include <mutex>

static __device__ __align__(256) double g_na_wacc[FLIPB200_NOALLOC_MAX_NODES];

// this file has no context to hang an error message on: report on stderr and return the status
#define NA_CK(call)                                                                                             \
  do {                                                                                                          \
    cudaError_t e__ = (call);                                                                                   \
    if (e__ != cudaSuccess) {                                                                                   \
      fprintf(stderr, "flipb200_p2g_noalloc: CUDA error %s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e__)); \
      return FLIPB200_ERR_CUDA;                                                                                 \
    }                                                                                                           \
  } while (0)

namespace {

constexpr int NA_MAX_DEV = 64;
double* g_na_addr[NA_MAX_DEV];          // device address of g_na_wacc on each device (nullptr: not asked yet)
std::mutex g_na_mu;

// device address of the scratch array on the CURRENT device
int na_scratch(double** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= NA_MAX_DEV) return FLIPB200_ERR_CUDA;
  std::lock_guard<std::mutex> lk(g_na_mu);
  if (!g_na_addr[dev]) {
    double* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_na_wacc);
    if (e != cudaSuccess || !p) {
      fprintf(stderr, "flipb200_p2g_noalloc: cudaGetSymbolAddress failed on device %d: %s\n", dev, cudaGetErrorString(e));
      return FLIPB200_ERR_CUDA;
    }
    g_na_addr[dev] = p;
  }
  *out = g_na_addr[dev];
  return 0;
}

struct NaGrid { int nx, ny, nz, ox, oy, oz, wall; };

__device__ __forceinline__ bool na_solid(const NaGrid& g, const uint8_t* __restrict__ solid, int i, int j, int k, int q) {
  const bool shell = i < g.wall || j < g.wall || k < g.wall || i > g.nx - 1 - g.wall || j > g.ny - 1 - g.wall || k > g.nz - 1 - g.wall;
  return shell || (solid && solid[q]);
}

// one thread per particle, 8 nodes x 4 REDs (the reference's loop shape; particles need not be sorted)
__global__ void __launch_bounds__(128)
k_na_splat(NaGrid g, int64_t n, const uint8_t* __restrict__ solid,
           const double* __restrict__ px, const double* __restrict__ py, const double* __restrict__ pz,
           const double* __restrict__ vx, const double* __restrict__ vy, const double* __restrict__ vz,
           double* __restrict__ wacc, double* __restrict__ ux, double* __restrict__ uy, double* __restrict__ uz) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const double x = px[t], y = py[t], z = pz[t];
  if (!(fabs(x) < 1.0e9) || !(fabs(y) < 1.0e9) || !(fabs(z) < 1.0e9)) return;
  const int i = base_cell(x, g.ox), j = base_cell(y, g.oy), k = base_cell(z, g.oz);
  if (i < 0 || j < 0 || k < 0 || i > g.nx - 2 || j > g.ny - 2 || k > g.nz - 2) return;    // support leaves the box
  const double a = vx[t], b = vy[t], c = vz[t];
  double wx0, wx1, wy0, wy1, wz0, wz1;
  spline_pair(x, (double)(i - g.ox), wx0, wx1); spline_pair(y, (double)(j - g.oy), wy0, wy1); spline_pair(z, (double)(k - g.oz), wz0, wz1);
  const int plane = g.nx * g.ny, q = (k * g.ny + j) * g.nx + i;
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int ii = i + (m & 1), jj = j + ((m >> 1) & 1), kk = k + ((m >> 2) & 1);
    const int qq = q + (m & 1) + ((m & 2) ? g.nx : 0) + ((m & 4) ? plane : 0);
    if (na_solid(g, solid, ii, jj, kk, qq)) continue;                                     // fluid.cc:283
    const double cw = ((m & 1) ? wx1 : wx0) * ((m & 2) ? wy1 : wy0) * ((m & 4) ? wz1 : wz0);
    atomicAdd(&wacc[qq], cw);
    atomicAdd(&ux[qq], cw * a);
    atomicAdd(&uy[qq], cw * b);
    atomicAdd(&uz[qq], cw * c);
  }
}

// weights are a FloatGrid in the reference (fluid.cc:1183): the division uses the float-rounded sum (:1138-1141)
__global__ void __launch_bounds__(256)
k_na_normalise(int nodes, const double* __restrict__ wacc, float* __restrict__ wout,
               double* __restrict__ ux, double* __restrict__ uy, double* __restrict__ uz) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nodes) return;
  const float w = (float)wacc[q];
  wout[q] = w;
  if (w > 0.0f) {
    const double wd = (double)w;
    ux[q] = ux[q] / wd; uy[q] = uy[q] / wd; uz[q] = uz[q] / wd;
  }
}

}  // namespace

extern "C" int flipb200_p2g_noalloc(int32_t nx, int32_t ny, int32_t nz, int32_t ox, int32_t oy, int32_t oz, int32_t wall,
                                    const uint8_t* solid_dev,
                                    const double* px_dev, const double* py_dev, const double* pz_dev,
                                    const double* vx_dev, const double* vy_dev, const double* vz_dev, int64_t n,
                                    double* u_dev, double* v_dev, double* w_dev, float* weight_dev, void* stream) {
  if (wall < 1 || nx < 2 * wall + 2 || ny < 2 * wall + 2 || nz < 2 * wall + 2) return FLIPB200_ERR_ARG;
  const int64_t nodes = (int64_t)nx * ny * nz;
  if (nodes > (int64_t)FLIPB200_NOALLOC_MAX_NODES) {
    fprintf(stderr, "flipb200_p2g_noalloc: grid of %d x %d x %d = %lld nodes exceeds FLIPB200_NOALLOC_MAX_NODES = %d "
                    "(the static scratch array); use flipb200_create + flipb200_p2g for larger grids\n",
            nx, ny, nz, (long long)nodes, (int)FLIPB200_NOALLOC_MAX_NODES);
    return FLIPB200_ERR_LIMIT;
  }
  if (n < 0 || !u_dev || !v_dev || !w_dev || !weight_dev) return FLIPB200_ERR_ARG;
  if (n > 0 && (!px_dev || !py_dev || !pz_dev || !vx_dev || !vy_dev || !vz_dev)) return FLIPB200_ERR_ARG;
  double* wacc = nullptr;
  int rc = na_scratch(&wacc);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t bytes = (size_t)nodes * sizeof(double);
  // the scratch keeps the previous call's sums, the outputs whatever the caller left there: both start at zero
  NA_CK(cudaMemsetAsync(wacc, 0, bytes, st));
  NA_CK(cudaMemsetAsync(u_dev, 0, bytes, st));
  NA_CK(cudaMemsetAsync(v_dev, 0, bytes, st));
  NA_CK(cudaMemsetAsync(w_dev, 0, bytes, st));
  const NaGrid g{nx, ny, nz, ox, oy, oz, wall};
  if (n > 0) {
    const int64_t blocks = (n + 127) / 128;
    if (blocks > 0x7fffffffLL) return FLIPB200_ERR_ARG;
    k_na_splat<<<(unsigned)blocks, 128, 0, st>>>(g, n, solid_dev, px_dev, py_dev, pz_dev, vx_dev, vy_dev, vz_dev, wacc, u_dev, v_dev, w_dev);
  }
  k_na_normalise<<<(unsigned)((nodes + 255) / 256), 256, 0, st>>>((int)nodes, wacc, weight_dev, u_dev, v_dev, w_dev);
  NA_CK(cudaGetLastError());
  return 0;
}
