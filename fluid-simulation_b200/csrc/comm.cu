// comm.cu -- exchanges of the z-slab decomposition.  The reference is single-process
// (its only parallelism is tbb::parallel_for over particles, fluid.cc:845-1126); this is new design.
//
// Two transports behind the same five primitives (halo fill, P2G ghost sum, all-reduce, plane
// all-gather, particle migration):
//   NCCL   one process per GPU (torch.distributed / mpirun bootstrap hands the 128-byte unique id
//          through flipb200_comm_init_nccl); point-to-point ncclSend/ncclRecv pairs between z
//          neighbours and small ncclAllReduce calls, all enqueued on the ctx stream so they can
//          be captured into the CUDA graph of the PCG iteration.  libnccl.so.2 is dlopen'ed on
//          first use: a single-GPU user of libflipb200.so needs no NCCL at all.
//   LOCAL  several ranks as host threads of ONE process (any devices, also all on one GPU): each
//          primitive is stream-sync + host barrier + cudaMemcpyAsync pulls.  Slow by design; it
//          exists so that the slab logic (ghost planes, distributed V-cycle, migration) is
//          parity-tested on a single GPU (tests/test_gpu_slab.py).
#include "common.cuh"
#include <dlfcn.h>
#include <string.h>
#include <stdlib.h>
#include <mutex>
#include <condition_variable>
#include <map>
#include <vector>
#include <nccl.h>

#define MAXR 16

// ---- NCCL entry points (dlsym) -----------------------------------------------------------------
struct NcclApi {
  void* h;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*GroupStart)();
  ncclResult_t (*GroupEnd)();
  const char* (*GetErrorString)(ncclResult_t);
};
static NcclApi g_nccl;
static std::mutex g_nccl_mu;

static bool nccl_load() {
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl.h) return true;
  const char* names[] = {getenv("FLIPB200_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* n : names) {
    if (!n || !n[0]) continue;
    h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) { fprintf(stderr, "flipb200: cannot dlopen libnccl.so.2: %s\n", dlerror()); return false; }
#define SYM(field, name)                                                   \
  *(void**)(&g_nccl.field) = dlsym(h, name);                               \
  if (!g_nccl.field) { fprintf(stderr, "flipb200: %s missing in NCCL\n", name); return false; }
  SYM(GetUniqueId, "ncclGetUniqueId") SYM(CommInitRank, "ncclCommInitRank") SYM(CommDestroy, "ncclCommDestroy")
  SYM(Send, "ncclSend") SYM(Recv, "ncclRecv") SYM(AllReduce, "ncclAllReduce") SYM(Broadcast, "ncclBroadcast") SYM(AllGather, "ncclAllGather")
  SYM(GroupStart, "ncclGroupStart") SYM(GroupEnd, "ncclGroupEnd") SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
  g_nccl.h = h;
  return true;
}

// ---- LOCAL transport: ranks are threads of this process ----------------------------------------
struct LocalGroup {
  int world = 0, joined = 0;
  std::mutex mu;
  std::condition_variable cv;
  int waiting = 0; unsigned phase = 0;
  flipb200_ctx* member[MAXR] = {};
  void* ptr[MAXR][8] = {};             // published device pointers of the current primitive
  Grid grid[MAXR];
  int64_t cnt[MAXR][4] = {};
  unsigned char red[MAXR][128] = {};   // all-reduce staging (host)
  void barrier() {
    std::unique_lock<std::mutex> lk(mu);
    unsigned ph = phase;
    if (++waiting == world) { waiting = 0; ++phase; cv.notify_all(); }
    else cv.wait(lk, [&] { return phase != ph; });
  }
};
static std::map<int, LocalGroup*> g_groups;
static std::mutex g_groups_mu;

struct Comm {
  int kind;            // 1 NCCL, 2 LOCAL
  ncclComm_t nc;
  LocalGroup* lg;
};

#define NC(call)                                                                                    \
  do {                                                                                              \
    ncclResult_t r__ = (call);                                                                      \
    if (r__ != ncclSuccess) {                                                                       \
      snprintf(c->errmsg, sizeof(c->errmsg), "%s:%d NCCL %s", __FILE__, __LINE__, g_nccl.GetErrorString(r__)); \
      fprintf(stderr, "flipb200: %s\n", c->errmsg);                                                 \
      return FLIPB200_ERR_COMM;                                                                     \
    }                                                                                               \
  } while (0)

extern "C" {

int flipb200_comm_unique_id(void* id128) {
  if (!id128) return FLIPB200_ERR_ARG;
  if (!nccl_load()) return FLIPB200_ERR_COMM;
  ncclUniqueId id;
  if (g_nccl.GetUniqueId(&id) != ncclSuccess) return FLIPB200_ERR_COMM;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  memcpy(id128, &id, 128);
  return FLIPB200_OK;
}

int flipb200_comm_init_nccl(flipb200_ctx* c, const void* id128) {
  if (!c || !id128) return FLIPB200_ERR_ARG;
  if (c->world == 1) return FLIPB200_OK;
  if (c->comm) return FLIPB200_ERR_STATE;
  if (!nccl_load()) return FLIPB200_ERR_COMM;
  CK(cudaSetDevice(c->device));
  ncclUniqueId id; memcpy(&id, id128, 128);
  Comm* cm = (Comm*)calloc(1, sizeof(Comm));
  cm->kind = 1;
  ncclResult_t r = g_nccl.CommInitRank(&cm->nc, c->world, id, c->rank);
  if (r != ncclSuccess) {
    fprintf(stderr, "flipb200: ncclCommInitRank: %s\n", g_nccl.GetErrorString(r));
    free(cm);
    return FLIPB200_ERR_COMM;
  }
  c->comm = cm;
  return FLIPB200_OK;
}

int flipb200_comm_init_local(flipb200_ctx* c, int group_id) {
  if (!c) return FLIPB200_ERR_ARG;
  if (c->world == 1) return FLIPB200_OK;
  if (c->comm || c->world > MAXR) return FLIPB200_ERR_STATE;
  std::lock_guard<std::mutex> lk(g_groups_mu);
  LocalGroup*& g = g_groups[group_id];
  if (!g) { g = new LocalGroup(); g->world = c->world; }
  if (g->world != c->world || g->member[c->rank]) return FLIPB200_ERR_ARG;
  g->member[c->rank] = c; g->joined++;
  Comm* cm = (Comm*)calloc(1, sizeof(Comm));
  cm->kind = 2; cm->lg = g;
  c->comm = cm;
  return FLIPB200_OK;
}

}  // extern "C"

void fb_comm_free(flipb200_ctx* c) {
  if (!c->comm) return;
  if (c->comm->kind == 1 && g_nccl.h) g_nccl.CommDestroy(c->comm->nc);
  if (c->comm->kind == 2) {
    std::lock_guard<std::mutex> lk(g_groups_mu);
    LocalGroup* g = c->comm->lg;
    g->member[c->rank] = nullptr;
    if (--g->joined == 0) {
      for (auto it = g_groups.begin(); it != g_groups.end(); ++it)
        if (it->second == g) { g_groups.erase(it); break; }
      delete g;
    }
  }
  free(c->comm);
  c->comm = nullptr;
}

static inline bool comm_missing(flipb200_ctx* c) {
  if (c->comm) return false;
  snprintf(c->errmsg, sizeof(c->errmsg), "slab context used before flipb200_comm_init_*");
  fprintf(stderr, "flipb200: %s\n", c->errmsg);
  return true;
}

int fb_comm_barrier(flipb200_ctx* c) {
  if (c->world == 1) return 0;
  if (comm_missing(c)) return FLIPB200_ERR_COMM;
  CK(cudaStreamSynchronize(c->stream));
  if (c->comm->kind == 2) c->comm->lg->barrier();
  return 0;
}

// ---- halo fill ---------------------------------------------------------------------------------
int fb_halo_fill(flipb200_ctx* c, const Grid& g, void* base, size_t elem, int nlo, int nhi) {
  if (c->world == 1) return 0;
  if (comm_missing(c)) return FLIPB200_ERR_COMM;
  const bool up = c->rank + 1 < c->world, dn = c->rank > 0;
  const size_t pb = (size_t)g.plane * elem;
  char* b = (char*)base;
  if (c->comm->kind == 1) {
    ncclComm_t nc = c->comm->nc;
    NC(g_nccl.GroupStart());
    if (up) {
      if (nlo) NC(g_nccl.Send(b + (size_t)(g.k1 - nlo) * pb, nlo * pb, ncclChar, c->rank + 1, nc, c->stream));
      if (nhi) NC(g_nccl.Recv(b + (size_t)g.k1 * pb, nhi * pb, ncclChar, c->rank + 1, nc, c->stream));
    }
    if (dn) {
      if (nhi) NC(g_nccl.Send(b + (size_t)g.k0 * pb, nhi * pb, ncclChar, c->rank - 1, nc, c->stream));
      if (nlo) NC(g_nccl.Recv(b + (size_t)(g.k0 - nlo) * pb, nlo * pb, ncclChar, c->rank - 1, nc, c->stream));
    }
    NC(g_nccl.GroupEnd());
    return 0;
  }
  LocalGroup* lg = c->comm->lg;
  lg->ptr[c->rank][0] = base; lg->grid[c->rank] = g;
  CK(cudaStreamSynchronize(c->stream));
  lg->barrier();
  if (dn && nlo) {
    const Grid& pg = lg->grid[c->rank - 1];
    CK(cudaMemcpyAsync(b + (size_t)(g.k0 - nlo) * pb, (char*)lg->ptr[c->rank - 1][0] + (size_t)(pg.k1 - nlo) * pb,
                       nlo * pb, cudaMemcpyDefault, c->stream));
  }
  if (up && nhi) {
    const Grid& pg = lg->grid[c->rank + 1];
    CK(cudaMemcpyAsync(b + (size_t)g.k1 * pb, (char*)lg->ptr[c->rank + 1][0] + (size_t)pg.k0 * pb, nhi * pb,
                       cudaMemcpyDefault, c->stream));
  }
  CK(cudaStreamSynchronize(c->stream));
  lg->barrier();
  return 0;
}

// ---- P2G ghost sum -------------------------------------------------------------------------------
__global__ void k_add_planes(int n, const double* __restrict__ src, double* __restrict__ d0, double* __restrict__ d1,
                             double* __restrict__ d2, double* __restrict__ d3) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  d0[i] += src[i]; d1[i] += src[n + i]; d2[i] += src[2 * n + i]; d3[i] += src[3 * n + i];
}

int fb_halo_sum_up(flipb200_ctx* c, double* const arr[4]) {
  if (c->world == 1) return 0;
  if (comm_missing(c)) return FLIPB200_ERR_COMM;
  const Grid& g = c->g;
  const bool up = c->rank + 1 < c->world, dn = c->rank > 0;
  const size_t pn = (size_t)g.plane, pb = pn * sizeof(double);
  if (c->comm->kind == 1) {
    ncclComm_t nc = c->comm->nc;
    NC(g_nccl.GroupStart());
    for (int a = 0; a < 4; ++a) {
      if (up) NC(g_nccl.Send(arr[a] + (size_t)g.k1 * pn, pn, ncclDouble, c->rank + 1, nc, c->stream));
      if (dn) NC(g_nccl.Recv(c->halo_tmp + a * pn, pn, ncclDouble, c->rank - 1, nc, c->stream));
    }
    NC(g_nccl.GroupEnd());
  } else {
    LocalGroup* lg = c->comm->lg;
    for (int a = 0; a < 4; ++a) lg->ptr[c->rank][a] = arr[a];
    lg->grid[c->rank] = g;
    CK(cudaStreamSynchronize(c->stream));
    lg->barrier();
    if (dn) {
      const Grid& pg = lg->grid[c->rank - 1];
      for (int a = 0; a < 4; ++a)
        CK(cudaMemcpyAsync(c->halo_tmp + a * pn, (double*)lg->ptr[c->rank - 1][a] + (size_t)pg.k1 * pn, pb,
                           cudaMemcpyDefault, c->stream));
    }
    CK(cudaStreamSynchronize(c->stream));
    lg->barrier();
  }
  if (dn) {
    const size_t o = (size_t)g.k0 * pn;
    k_add_planes<<<cdiv(g.plane, 256), 256, 0, c->stream>>>(g.plane, c->halo_tmp, arr[0] + o, arr[1] + o, arr[2] + o,
                                                            arr[3] + o);
    CK_LAUNCH(); c->launches++;
  }
  return 0;
}

// ---- all-reduce (in place, device memory, <= 16 elements) ---------------------------------------
template <typename T>
static void host_reduce(T* out, const unsigned char (*in)[128], int world, int count, int op) {
  for (int e = 0; e < count; ++e) {
    T v = ((const T*)in[0])[e];
    for (int r = 1; r < world; ++r) {
      T w = ((const T*)in[r])[e];
      v = op == FB_SUM ? v + w : (op == FB_MIN ? (w < v ? w : v) : (w > v ? w : v));
    }
    out[e] = v;
  }
}

int fb_allreduce(flipb200_ctx* c, void* dev, int count, int type, int op) {
  if (c->world == 1) return 0;
  if (comm_missing(c)) return FLIPB200_ERR_COMM;
  const size_t esz = type == FB_I32 ? 4 : 8;
  if (count < 1 || count * esz > 128) return FLIPB200_ERR_ARG;
  if (c->comm->kind == 1) {
    const ncclDataType_t t = type == FB_F64 ? ncclDouble : (type == FB_I32 ? ncclInt32 : ncclUint64);
    const ncclRedOp_t o = op == FB_SUM ? ncclSum : (op == FB_MIN ? ncclMin : ncclMax);
    NC(g_nccl.AllReduce(dev, dev, (size_t)count, t, o, c->comm->nc, c->stream));
    return 0;
  }
  LocalGroup* lg = c->comm->lg;
  CK(cudaMemcpyAsync(lg->red[c->rank], dev, count * esz, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  lg->barrier();
  unsigned char out[128];
  if (type == FB_F64) host_reduce<double>((double*)out, lg->red, c->world, count, op);
  else if (type == FB_I32) host_reduce<int32_t>((int32_t*)out, lg->red, c->world, count, op);
  else host_reduce<unsigned long long>((unsigned long long*)out, lg->red, c->world, count, op);
  CK(cudaMemcpyAsync(dev, out, count * esz, cudaMemcpyHostToDevice, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  lg->barrier();
  return 0;
}

// ---- all-gather of the owned planes of a replicated grid ----------------------------------------
// All-gather of the z-planes [zs[r], zs[r+1]) that rank r owns of up to two arrays of one level
// (one NCCL launch: in-place ncclAllGather when the ranks own equally many planes, grouped
// broadcasts otherwise).
int fb_allgather_planes2(flipb200_ctx* c, const Grid& gg, void* base0, void* base1, size_t elem, const int* zs) {
  if (c->world == 1) return 0;
  if (comm_missing(c)) return FLIPB200_ERR_COMM;
  const size_t pb = (size_t)gg.plane * elem;
  void* bases[2] = {base0, base1};
  if (c->comm->kind == 1) {
    bool uniform = zs[0] == 0;
    for (int r = 1; r < c->world; ++r) uniform = uniform && (zs[r + 1] - zs[r] == zs[1] - zs[0]);
    NC(g_nccl.GroupStart());
    for (int a = 0; a < 2; ++a) {
      char* b = (char*)bases[a];
      if (!b) continue;
      if (uniform && g_nccl.AllGather) {
        const size_t n = (size_t)(zs[1] - zs[0]) * pb;
        NC(g_nccl.AllGather(b + (size_t)c->rank * n, b, n, ncclChar, c->comm->nc, c->stream));
        continue;
      }
      for (int r = 0; r < c->world; ++r) {
        const size_t n = (size_t)(zs[r + 1] - zs[r]) * pb;
        if (n) NC(g_nccl.Broadcast(b + (size_t)zs[r] * pb, b + (size_t)zs[r] * pb, n, ncclChar, r, c->comm->nc, c->stream));
      }
    }
    NC(g_nccl.GroupEnd());
    return 0;
  }
  LocalGroup* lg = c->comm->lg;
  lg->ptr[c->rank][0] = base0;
  lg->ptr[c->rank][1] = base1;
  CK(cudaStreamSynchronize(c->stream));
  lg->barrier();
  for (int a = 0; a < 2; ++a) {
    char* b = (char*)bases[a];
    if (!b) continue;
    for (int r = 0; r < c->world; ++r) {
      const size_t n = (size_t)(zs[r + 1] - zs[r]) * pb;
      if (r == c->rank || !n) continue;
      CK(cudaMemcpyAsync(b + (size_t)zs[r] * pb, (char*)lg->ptr[r][a] + (size_t)zs[r] * pb, n, cudaMemcpyDefault, c->stream));
    }
  }
  CK(cudaStreamSynchronize(c->stream));
  lg->barrier();
  return 0;
}
int fb_allgather_planes(flipb200_ctx* c, const Grid& gg, void* base, size_t elem, const int* zs) {
  return fb_allgather_planes2(c, gg, base, nullptr, elem, zs);
}

// ---- particle migration --------------------------------------------------------------------------
// index 0 = the lower neighbour (rank-1), 1 = the upper neighbour (rank+1)
int fb_comm_exchange_counts(flipb200_ctx* c, const int64_t send[2], int64_t recv[2]) {
  recv[0] = recv[1] = 0;
  if (c->world == 1) return 0;
  if (comm_missing(c)) return FLIPB200_ERR_COMM;
  const bool up = c->rank + 1 < c->world, dn = c->rank > 0;
  if (c->comm->kind == 1) {
    // staged through DevScalars::part (idle outside the PCG): [0..1] send, [2..3] receive
    double h[4] = {(double)send[0], (double)send[1], 0.0, 0.0};
    CK(cudaMemcpyAsync(c->ds->part, h, sizeof(h), cudaMemcpyHostToDevice, c->stream));
    ncclComm_t nc = c->comm->nc;
    NC(g_nccl.GroupStart());
    if (dn) { NC(g_nccl.Send(&c->ds->part[0], 1, ncclDouble, c->rank - 1, nc, c->stream));
              NC(g_nccl.Recv(&c->ds->part[2], 1, ncclDouble, c->rank - 1, nc, c->stream)); }
    if (up) { NC(g_nccl.Send(&c->ds->part[1], 1, ncclDouble, c->rank + 1, nc, c->stream));
              NC(g_nccl.Recv(&c->ds->part[3], 1, ncclDouble, c->rank + 1, nc, c->stream)); }
    NC(g_nccl.GroupEnd());
    CK(cudaMemcpyAsync(h, c->ds->part, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    recv[0] = (int64_t)h[2]; recv[1] = (int64_t)h[3];
    CK(cudaMemsetAsync(c->ds->part, 0, sizeof(h), c->stream));
    return 0;
  }
  LocalGroup* lg = c->comm->lg;
  lg->cnt[c->rank][0] = send[0]; lg->cnt[c->rank][1] = send[1];
  lg->barrier();
  if (dn) recv[0] = lg->cnt[c->rank - 1][1];
  if (up) recv[1] = lg->cnt[c->rank + 1][0];
  lg->barrier();
  return 0;
}

int fb_comm_exchange_particles(flipb200_ctx* c, int narr, void* const src[], void* const dst[], const size_t elem[],
                               const int64_t send_off[2], const int64_t send_n[2],
                               const int64_t recv_off[2], const int64_t recv_n[2]) {
  if (c->world == 1) return 0;
  if (comm_missing(c)) return FLIPB200_ERR_COMM;
  if (narr > 8) return FLIPB200_ERR_ARG;
  const bool up = c->rank + 1 < c->world, dn = c->rank > 0;
  if (c->comm->kind == 1) {
    ncclComm_t nc = c->comm->nc;
    NC(g_nccl.GroupStart());
    for (int a = 0; a < narr; ++a) {
      char* s = (char*)src[a]; char* d = (char*)dst[a];
      if (dn && send_n[0]) NC(g_nccl.Send(s + send_off[0] * elem[a], send_n[0] * elem[a], ncclChar, c->rank - 1, nc, c->stream));
      if (up && send_n[1]) NC(g_nccl.Send(s + send_off[1] * elem[a], send_n[1] * elem[a], ncclChar, c->rank + 1, nc, c->stream));
      if (dn && recv_n[0]) NC(g_nccl.Recv(d + recv_off[0] * elem[a], recv_n[0] * elem[a], ncclChar, c->rank - 1, nc, c->stream));
      if (up && recv_n[1]) NC(g_nccl.Recv(d + recv_off[1] * elem[a], recv_n[1] * elem[a], ncclChar, c->rank + 1, nc, c->stream));
    }
    NC(g_nccl.GroupEnd());
    return 0;
  }
  LocalGroup* lg = c->comm->lg;
  for (int a = 0; a < narr; ++a) lg->ptr[c->rank][a] = src[a];
  lg->cnt[c->rank][2] = send_off[0]; lg->cnt[c->rank][3] = send_off[1];
  CK(cudaStreamSynchronize(c->stream));
  lg->barrier();
  for (int a = 0; a < narr; ++a) {
    char* d = (char*)dst[a];
    if (dn && recv_n[0])   // the lower neighbour's "up" bucket
      CK(cudaMemcpyAsync(d + recv_off[0] * elem[a], (char*)lg->ptr[c->rank - 1][a] + lg->cnt[c->rank - 1][3] * elem[a],
                         recv_n[0] * elem[a], cudaMemcpyDefault, c->stream));
    if (up && recv_n[1])   // the upper neighbour's "down" bucket
      CK(cudaMemcpyAsync(d + recv_off[1] * elem[a], (char*)lg->ptr[c->rank + 1][a] + lg->cnt[c->rank + 1][2] * elem[a],
                         recv_n[1] * elem[a], cudaMemcpyDefault, c->stream));
  }
  CK(cudaStreamSynchronize(c->stream));
  lg->barrier();
  return 0;
}

extern "C" int flipb200_halo_mode(flipb200_ctx* c) { return c ? c->halo.mode : FLIPB200_ERR_ARG; }
// diagnostics for bench.py: [0] halo mode [1] PCG iteration replayed from a CUDA graph [2] multigrid levels
// [3] distributed levels [4] kernels per iteration [5] ghost-pushing kernels per iteration [6] rank [7] world
extern "C" int flipb200_debug_info(flipb200_ctx* c, int32_t* out8) {
  if (!c || !out8) return FLIPB200_ERR_ARG;
  out8[0] = c->halo.mode + (c->halo.red_on ? 10 : 0) + (c->halo.gather_on ? 20 : 0);   // +10: all-reduce fused too, +20: all-gather by peer stores
  out8[1] = c->it_exec ? 1 : 0; out8[2] = c->mg_levels; out8[3] = c->mg_nd;
  out8[4] = (int32_t)c->it_kernels; out8[5] = c->it_nprod; out8[6] = c->rank; out8[7] = c->world;
  return FLIPB200_OK;
}

bool fb_comm_capturable(flipb200_ctx* c) { return c->world == 1 || (c->comm && c->comm->kind == 1); }

// ---- ghost pushes fused into the PCG kernels (halo.cuh) -------------------------------------------
#include "halo.cuh"

static bool link_active(const flipb200_ctx* c, const MgLevel& L, int dir) {
  if (!L.dist) return false;
  const Box& gb = L.gbox;
  if (!(gb.x1 > gb.x0 && gb.z1 > gb.z0)) return false;
  if (dir == 0) { if (c->rank == 0) return false; const int zb = L.zs[c->rank]; return gb.z0 < zb && gb.z1 > zb; }
  if (c->rank + 1 >= c->world) return false;
  const int zb = L.zs[c->rank + 1];
  return gb.z0 < zb && gb.z1 > zb;       // the sweep box straddles the slab boundary: both sides sweep their boundary plane
}

// Describe one launch: it consumes array (lin, ain) -- ghost planes pushed by the neighbours' latest
// producer of that array -- and/or produces array (lout, aout).  -1 = none.  kdn / kup: first / last
// owned plane in the index space of the sweep; nb: blocks per boundary plane.
PeerHalo fb_halo_make(flipb200_ctx* c, int lin, int ain, int lout, int aout, size_t elem_out, int kdn, int kup, int nb) {
  PeerHalo h = halo_off();
  HaloState& hs = c->halo;
  if (hs.mode == 0) return h;
  h.mode = hs.mode; h.kdn = kdn; h.kup = kup; h.nb = nb;
  h.acq_sys = hs.acq_sys;
  if (hs.dbg) { h.dbg = hs.dbg + 8 * (hs.dbg_n % HALO_DBG_MAX); hs.dbg_n++; }
  h.my_flags = hs.flags; h.cnt = hs.flags + 2;
  h.sig_dn = hs.peer_flags[0] ? hs.peer_flags[0] + 1 : nullptr;
  h.sig_up = hs.peer_flags[1] ? hs.peer_flags[1] + 0 : nullptr;
  if (ain >= 0) {
    const MgLevel& L = c->mg[lin];
    h.wait_dn = link_active(c, L, 0); h.wait_up = link_active(c, L, 1);
    h.wrel_dn = h.wrel_up = (int)(hs.last_prod[lin][ain] - hs.seg_base);
  }
  if (aout >= 0) {
    const MgLevel& L = c->mg[lout];
    h.push_dn = link_active(c, L, 0); h.push_up = link_active(c, L, 1);
    h.srel = (int)(hs.abs_seq - hs.seg_base);
    hs.last_prod[lout][aout] = hs.abs_seq;
    hs.abs_seq++;
    const Grid& g = L.g;
    const size_t pb = (size_t)g.plane * elem_out;
    if (h.push_dn) {       // lower neighbour's upper ghost plane k1' <- my plane k0
      const int k1p = SLAB_GHOST + (L.zs[c->rank] - L.zs[c->rank - 1]);
      h.dst_dn = (char*)hs.peer[0][lout][aout] + ((ptrdiff_t)k1p - g.k0) * (ptrdiff_t)pb;
    }
    if (h.push_up)         // upper neighbour's lower ghost plane k0' - 1 <- my plane k1 - 1
      h.dst_up = (char*)hs.peer[1][lout][aout] + ((ptrdiff_t)(SLAB_GHOST - 1) - (g.k1 - 1)) * (ptrdiff_t)pb;
  }
  return h;
}

RedPub fb_red_make(flipb200_ctx* c) {
  RedPub r = red_off();
  if (!c->halo.red_on) return r;
  r.on = 1; r.rank = c->rank; r.world = c->world;
  for (int q = 0; q < c->world; ++q) r.peers[q] = c->halo.peer_mail[q];
  return r;
}

// All-gather of the first replicated level by peer stores (halo.cuh GatherPush): wanted when the level
// is fed by a distributed one and has its own kernels (not inside the one-block tail);
// FLIPB200_FUSED_GATHER=0 keeps the NCCL all-gather.
static bool gather_wanted(const flipb200_ctx* c) {
  const char* e = getenv("FLIPB200_FUSED_GATHER");
  if (e && e[0] == '0') return false;
  const int nd = c->mg_nd;
  return c->world > 1 && c->world <= RED_MAXR && nd >= 1 && nd < c->mg_levels && nd < c->mg_tail0 && !c->mg[nd].dist;
}

GatherPush fb_gather_make(flipb200_ctx* c, int nblocks) {
  GatherPush g = gather_off();
  HaloState& hs = c->halo;
  if (!hs.gather_on || hs.mode == 0) return g;
  g.on = 1; g.mode = hs.mode; g.rank = c->rank; g.world = c->world; g.nblocks = nblocks;
  g.srel = (int)(hs.abs_seq - hs.seg_base);
  hs.gather_seq = hs.abs_seq;
  hs.abs_seq++;
  for (int r = 0; r < c->world; ++r) {
    g.b[r] = hs.gather_b[r]; g.xa[r] = hs.gather_xa[r];
    g.flag[r] = hs.mode == 2 && hs.peer_mail[r] ? hs.peer_mail[r]->gflag_ptr() : nullptr;
  }
  g.cnt = hs.flags + 4;
  return g;
}

void fb_gather_wait(flipb200_ctx* c, PeerHalo* h) {
  HaloState& hs = c->halo;
  if (!hs.gather_on || hs.mode != 2) return;
  h->mode = 2;
  h->acq_sys = hs.acq_sys;
  h->gat_flags = hs.mail->gflag_ptr();
  h->gat_n = c->world; h->gat_self = c->rank;
  h->gat_rel = (int)(hs.gather_seq - hs.seg_base);
}

// where a halo copy used to be: nothing to do with device flags; the in-process transport orders
// producers and consumers with its host barrier
int fb_halo_point(flipb200_ctx* c) {
  if (c->halo.mode == 1) return fb_comm_barrier(c);
  return 0;
}

// exchange `n` bytes with both neighbours over NCCL (setup only): out_dn/out_up receive theirs
static int nccl_swap_bytes(flipb200_ctx* c, const void* mine, void* from_dn, void* from_up, size_t n) {
  char* d = nullptr;
  CK(cudaMalloc((void**)&d, 3 * n));
  CK(cudaMemcpyAsync(d, mine, n, cudaMemcpyHostToDevice, c->stream));
  const bool up = c->rank + 1 < c->world, dn = c->rank > 0;
  ncclComm_t nc = c->comm->nc;
  NC(g_nccl.GroupStart());
  if (dn) { NC(g_nccl.Send(d, n, ncclChar, c->rank - 1, nc, c->stream)); NC(g_nccl.Recv(d + n, n, ncclChar, c->rank - 1, nc, c->stream)); }
  if (up) { NC(g_nccl.Send(d, n, ncclChar, c->rank + 1, nc, c->stream)); NC(g_nccl.Recv(d + 2 * n, n, ncclChar, c->rank + 1, nc, c->stream)); }
  NC(g_nccl.GroupEnd());
  if (dn) CK(cudaMemcpyAsync(from_dn, d + n, n, cudaMemcpyDeviceToHost, c->stream));
  if (up) CK(cudaMemcpyAsync(from_up, d + 2 * n, n, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  cudaFree(d);
  return 0;
}

// Map the neighbours' exchanged arrays.  NCCL transport: cudaIpc handles travel over NCCL itself;
// in-process transport: the other contexts' pointers are directly usable (one address space).
// FLIPB200_FUSED_HALO=0 keeps the separate halo copies.
int fb_halo_setup(flipb200_ctx* c) {
  HaloState& hs = c->halo;
  hs.mode = 0;
  if (c->world == 1 || !c->comm || c->mg_levels < 2 || c->mg_nd < 1) return 0;
  const char* e = getenv("FLIPB200_FUSED_HALO");
  if (e && e[0] == '0') return 0;
  { const char* a = getenv("FLIPB200_HALO_ACQ"); hs.acq_sys = (a && a[0] == 'g') ? 0 : 1; }   // halo.cuh: consumer-side acquire scope
  CK(cudaMalloc((void**)&hs.flags, 8 * sizeof(unsigned)));   // [0,1] flags, [2,3] push counters, [4] gather counter
  {
    const char* de = getenv("FLIPB200_HALO_DEBUG");
    if (de && de[0] == '1' && !hs.dbg) {
      CK(cudaMalloc((void**)&hs.dbg, 8 * HALO_DBG_MAX * sizeof(unsigned long long)));
      CK(cudaMemset(hs.dbg, 0, 8 * HALO_DBG_MAX * sizeof(unsigned long long)));
    }
  }
  CK(cudaMemsetAsync(hs.flags, 0, 8 * sizeof(unsigned), c->stream));
  CK(cudaStreamSynchronize(c->stream));
  const int nd = c->mg_nd;
  // my exported pointers: flags, pd, then (xa, xb) of every distributed level
  void* mine[2 + 2 * MG_MAX_LEVELS]; int n = 0;
  mine[n++] = hs.flags; mine[n++] = c->pd;
  for (int l = 0; l < nd; ++l) { mine[n++] = c->mg[l].xa; mine[n++] = c->mg[l].xb; }
  void* theirs[2][2 + 2 * MG_MAX_LEVELS] = {};
  const bool up = c->rank + 1 < c->world, dn = c->rank > 0;
  if (c->comm->kind == 2) {
    LocalGroup* lg = c->comm->lg;
    lg->barrier();                                   // every rank has allocated its arrays
    for (int dir = 0; dir < 2; ++dir) {
      if (!(dir ? up : dn)) continue;
      flipb200_ctx* o = lg->member[c->rank + (dir ? 1 : -1)];
      if (!o || o->mg_nd != nd) return FLIPB200_ERR_STATE;
      int m = 0;
      // the neighbour may not have run its own setup yet: its flag block is published below
      theirs[dir][m++] = nullptr; theirs[dir][m++] = o->pd;
      for (int l = 0; l < nd; ++l) { theirs[dir][m++] = o->mg[l].xa; theirs[dir][m++] = o->mg[l].xb; }
    }
    hs.mode = 1;                                     // flags unused: host barriers order the kernels
    if (gather_wanted(c)) {
      bool okg = true;
      for (int r = 0; r < c->world; ++r) {
        flipb200_ctx* o = lg->member[r];
        if (!o || o->mg_nd != nd || o->mg_levels != c->mg_levels) { okg = false; break; }
        hs.gather_b[r] = o->mg[nd].b; hs.gather_xa[r] = o->mg[nd].xa;
      }
      hs.gather_on = okg ? 1 : 0;
    }
  } else {
    cudaIpcMemHandle_t hm[2 + 2 * MG_MAX_LEVELS], hd[2 + 2 * MG_MAX_LEVELS], hu[2 + 2 * MG_MAX_LEVELS];
    memset(hm, 0, sizeof(hm));
    bool ok = true;
    for (int i = 0; i < n; ++i) if (cudaIpcGetMemHandle(&hm[i], mine[i]) != cudaSuccess) { ok = false; cudaGetLastError(); }
    unsigned char okb = ok ? 1 : 0;
    memcpy(&hm[n], &okb, 1);                         // slot n carries the "handles valid" byte
    int rc = nccl_swap_bytes(c, hm, hd, hu, sizeof(cudaIpcMemHandle_t) * (n + 1));
    if (rc) return rc;
    for (int dir = 0; dir < 2 && ok; ++dir) {
      if (!(dir ? up : dn)) continue;
      cudaIpcMemHandle_t* hh = dir ? hu : hd;
      if (*(unsigned char*)&hh[n] != 1) { ok = false; break; }
      for (int i = 0; i < n; ++i)
        if (cudaIpcOpenMemHandle(&theirs[dir][i], hh[i], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
          fprintf(stderr, "flipb200: cudaIpcOpenMemHandle failed (%s): ghost planes go through NCCL copies\n",
                  cudaGetErrorString(cudaGetLastError()));
          ok = false; break;
        }
    }
    // all ranks must agree on the mode
    int* flag = nullptr; int hflag = ok ? 1 : 0;
    CK(cudaMalloc((void**)&flag, sizeof(int)));
    CK(cudaMemcpyAsync(flag, &hflag, sizeof(int), cudaMemcpyHostToDevice, c->stream));
    NC(g_nccl.AllReduce(flag, flag, 1, ncclInt32, ncclMin, c->comm->nc, c->stream));
    CK(cudaMemcpyAsync(&hflag, flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    cudaFree(flag);
    if (!hflag) return 0;
    hs.mode = 2;
    // mailboxes of the fused all-reduce: every rank maps every rank's (FLIPB200_FUSED_REDUCE=0: NCCL all-reduce)
    const char* fr = getenv("FLIPB200_FUSED_REDUCE");
    if (!(fr && fr[0] == '0') && c->world <= RED_MAXR) {
      CK(cudaMalloc((void**)&hs.mail, sizeof(RedMail)));
      CK(cudaMemsetAsync(hs.mail, 0, sizeof(RedMail), c->stream));
      // every rank maps every rank's mailbox and -- for the all-gather by peer stores -- its (b, xa) of
      // the first replicated level: three handles per rank travel in one grouped broadcast
      const bool wantg = gather_wanted(c);
      struct H3 { cudaIpcMemHandle_t h[3]; int ok[3]; };
      H3 mh; memset(&mh, 0, sizeof(mh));
      mh.ok[0] = cudaIpcGetMemHandle(&mh.h[0], hs.mail) == cudaSuccess ? 1 : 0;
      if (wantg) {
        mh.ok[1] = cudaIpcGetMemHandle(&mh.h[1], c->mg[nd].b) == cudaSuccess ? 1 : 0;
        mh.ok[2] = cudaIpcGetMemHandle(&mh.h[2], c->mg[nd].xa) == cudaSuccess ? 1 : 0;
      }
      cudaGetLastError();
      int okm = mh.ok[0], okg = wantg && mh.ok[1] && mh.ok[2] ? 1 : 0;
      H3* dh = nullptr;
      std::vector<H3> all(c->world);
      CK(cudaMalloc((void**)&dh, sizeof(H3) * c->world));
      CK(cudaMemcpyAsync(dh + c->rank, &mh, sizeof(H3), cudaMemcpyHostToDevice, c->stream));
      NC(g_nccl.GroupStart());
      for (int r = 0; r < c->world; ++r) NC(g_nccl.Broadcast(dh + r, dh + r, sizeof(H3), ncclChar, r, c->comm->nc, c->stream));
      NC(g_nccl.GroupEnd());
      CK(cudaMemcpyAsync(all.data(), dh, sizeof(H3) * c->world, cudaMemcpyDeviceToHost, c->stream));
      CK(cudaStreamSynchronize(c->stream));
      cudaFree(dh);
      for (int r = 0; r < c->world; ++r) {
        if (r == c->rank) { hs.peer_mail[r] = hs.mail; hs.gather_b[r] = c->mg[nd].b; hs.gather_xa[r] = c->mg[nd].xa; continue; }
        void* p = nullptr;
        if (okm) {
          if (!all[r].ok[0] || cudaIpcOpenMemHandle(&p, all[r].h[0], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { okm = 0; cudaGetLastError(); }
          hs.peer_mail[r] = (RedMail*)p;
        }
        for (int a = 1; a <= 2 && okg; ++a) {
          p = nullptr;
          if (!all[r].ok[a] || cudaIpcOpenMemHandle(&p, all[r].h[a], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { okg = 0; cudaGetLastError(); }
          if (a == 1) hs.gather_b[r] = (float*)p; else hs.gather_xa[r] = (float*)p;
        }
      }
      int* flag2 = nullptr;
      int both[2] = {okm, okm && okg ? 1 : 0};
      CK(cudaMalloc((void**)&flag2, 2 * sizeof(int)));
      CK(cudaMemcpyAsync(flag2, both, 2 * sizeof(int), cudaMemcpyHostToDevice, c->stream));
      NC(g_nccl.AllReduce(flag2, flag2, 2, ncclInt32, ncclMin, c->comm->nc, c->stream));
      CK(cudaMemcpyAsync(both, flag2, 2 * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
      CK(cudaStreamSynchronize(c->stream));
      cudaFree(flag2);
      okm = both[0];
      hs.gather_on = both[1];
      hs.red_on = okm;
    }
  }
  for (int dir = 0; dir < 2; ++dir) {
    int m = 0;
    hs.peer_flags[dir] = (unsigned*)theirs[dir][m++];
    hs.peer[dir][0][HA_PD] = theirs[dir][m++];
    for (int l = 0; l < nd; ++l) { hs.peer[dir][l][HA_XA] = theirs[dir][m++]; hs.peer[dir][l][HA_XB] = theirs[dir][m++]; }
  }
  hs.abs_seq = 1; hs.seg_base = 1;
  return 0;
}
