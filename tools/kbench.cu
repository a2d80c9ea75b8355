// kbench.cu -- design notebook (not product): which kernel shapes reach HBM peak on B200 for the
// two patterns the pressure solver is made of: (S) pure streaming axpy-type updates and (T) 7-point
// stencil sweeps, on 256^3 fp64 arrays with a 1-byte flag array.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/kbench tools/kbench.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

static int N = 256;
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
  if (tid == 0) { double t = 0; for (int w = 0; w < BX * BY / 32; ++w) t += sm[w]; if (t != 0) atomicAdd(out + (blockIdx.x + blockIdx.y + blockIdx.z) % 32, t); }
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
  if (lane == 0 && threadIdx.y == 0) { double t = 0; for (int w = 0; w < BY; ++w) t += sm[w]; if (t != 0) atomicAdd(out + (blockIdx.x + blockIdx.y + blockIdx.z) % 32, t); }
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
  RUN_T4(4, 1) RUN_T4(8, 1) RUN_T4(4, 2) RUN_T4(8, 2) RUN_T4(4, 4) RUN_T4(8, 4) RUN_T4(4, 8) RUN_T4(8, 8) RUN_T4(16, 4) RUN_T4(4, 16)
  return 0;
}
