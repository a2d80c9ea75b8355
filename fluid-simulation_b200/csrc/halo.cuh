// halo.cuh -- ghost-plane exchange FUSED into the producing kernels (z-slab runs).
//
// Inside the PCG iteration every sweep of a distributed level needs one ghost plane of its input
// from each z neighbour (~20 exchanges per iteration with the V-cycle).  As separate NCCL
// send/recv launches they cost ~14 us each, a third of the iteration.  Here the kernel that
// PRODUCES an exchanged array also stores its two boundary planes straight into the neighbours'
// ghost planes through NVLink peer mappings (cudaIpc), and the kernel that CONSUMES the array makes
// only its boundary blocks wait on a flag -- interior blocks run while the planes are in flight.
//
//   producer block that owns boundary plane kdn / kup:   st.global to the peer pointer, then
//       one system fence per block, block counter; the LAST boundary block release-stores the sequence
//       number of this producer kernel into the neighbour's flag word.
//   consumer block that reads ghost plane kdn-1 / kup+1: thread 0 acquire-spins until the flag
//       reaches the sequence number of the latest producer of that array (bounded: a time-out sets
//       ds->pcg_status = FLIPB200_ERR_COMM instead of hanging the GPU).
//
// Sequence numbers: every producer launch gets the next number; all ranks launch the same kernels
// in the same order, so "my producer #s" pairs with "the neighbour's producer #s".  Kernels read
// seq = ds->halo_base + rel (rel is a launch constant), which keeps a captured CUDA graph valid:
// k_pcg_advance adds the number of producers of the iteration to halo_base.
// Write-after-read safety (a push must not overwrite a ghost plane the neighbour still reads) follows
// from the ping-pong of xa/xb: the blocks that push array X have first waited on the neighbour's
// producer of their own input, which runs after the neighbour's last reader of X.
//
// mode 0: off (separate halo copies);  mode 1: push only, the host barrier of the in-process test
// transport orders producers and consumers;  mode 2: push + device flags (one process per GPU).
#pragma once
#include "sweep.cuh"
#include <string.h>

#define RED_MAXR 16   // ranks the peer-mapped mailboxes / gather tables are sized for

struct PeerHalo {
  int mode;
  int kdn, kup;              // planes (sweep index space) whose blocks are the boundary blocks: first / last owned
  int wait_dn, wait_up;      // consumer side: wait for the lower / upper neighbour's producer
  int push_dn, push_up;      // producer side: push + signal to the lower / upper neighbour
  int wrel_dn, wrel_up;      // seq to wait for   = ds->halo_base + wrel
  int srel;                  // seq to signal     = ds->halo_base + srel
  int nb;                    // boundary blocks per direction (gridDim.x * gridDim.y)
  void* dst_dn;              // peer array rebased so that MY element index addresses the peer's ghost plane
  void* dst_up;
  unsigned* my_flags;        // [0] written by the lower neighbour, [1] by the upper neighbour
  unsigned* sig_dn;          // lower neighbour's my_flags[1]
  unsigned* sig_up;          // upper neighbour's my_flags[0]
  unsigned* cnt;             // [2] block counters (own memory)
  unsigned long long* dbg;   // FLIPB200_HALO_DEBUG=1: timestamps of this launch's boundary blocks (8 per launch)
  // consumer of the first replicated level: EVERY block waits until every rank's restriction has
  // pushed its planes (GatherPush below): gat_flags[r] >= halo_base + gat_rel for r != gat_self
  const unsigned* gat_flags;
  int gat_n, gat_self, gat_rel;
  // consumer-side acquire: 1 (default) = after the relaxed spin ONE ld.acquire.sys of the flag, so that the
  // acquire's scope covers the releasing thread on the other GPU (PTX memory model: synchronizes-with needs
  // a scope that includes both threads); 0 (FLIPB200_HALO_ACQ=gpu) = the round-1 gpu-scope fence.
  int acq_sys;
};
__device__ __forceinline__ unsigned long long halo_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned ld_relaxed_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

#define HALO_SPIN_LIMIT (4000000000ll)   // ~2 s of SM clocks

// Call at kernel start (after the uniform early exits), by ALL threads of the block.
// bdn / bup: this block reads the ghost plane below kdn / above kup.
__device__ __forceinline__ void halo_wait(const PeerHalo& H, DevScalars* ds, bool bdn, bool bup) {
  if (H.mode != 2) return;
  if (H.gat_n) {                                     // block-uniform
    const int t = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
    if (t < H.gat_n && t != H.gat_self) {
      const unsigned need = (unsigned)ds->halo_base + (unsigned)H.gat_rel;
      const long long t0 = clock64();
      while ((int)(ld_relaxed_sys(&H.gat_flags[t]) - need) < 0)
        if (clock64() - t0 > HALO_SPIN_LIMIT) { ds->pcg_status = FLIPB200_ERR_COMM; break; }
      if (H.acq_sys) (void)ld_acquire_sys(&H.gat_flags[t]); else fence_acq_rel_gpu();
    }
    __syncthreads();
  }
  const bool w0 = bdn && H.wait_dn, w1 = bup && H.wait_up;
  if (!(w0 || w1)) return;
  const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  if (tid == 0) {
    // relaxed polls, ONE acquire fence after the flags have arrived (an acquire load per poll is a
    // system-scope round trip each time)
    if (H.dbg && blockIdx.x == 0 && blockIdx.y == 0) H.dbg[w0 ? 0 : 1] = halo_now();
    const unsigned base = (unsigned)ds->halo_base;
    const long long t0 = clock64();
    if (w0) {
      const unsigned need = base + (unsigned)H.wrel_dn;
      while ((int)(ld_relaxed_sys(&H.my_flags[0]) - need) < 0)
        if (clock64() - t0 > HALO_SPIN_LIMIT) { ds->pcg_status = FLIPB200_ERR_COMM; break; }
    }
    if (w1) {
      const unsigned need = base + (unsigned)H.wrel_up;
      while ((int)(ld_relaxed_sys(&H.my_flags[1]) - need) < 0)
        if (clock64() - t0 > HALO_SPIN_LIMIT) { ds->pcg_status = FLIPB200_ERR_COMM; break; }
    }
    // The neighbour (another GPU) ordered its plane stores before the flag with a system-scope release.
    // The acquire that pairs with it must be at a scope that includes that thread: ONE ld.acquire.sys of
    // the flag after the relaxed spin (the polls themselves stay relaxed: an acquire per poll is a
    // system-scope round trip each time).  It reads the value the spin saw (or a later one of the same
    // release sequence), so it synchronizes-with the release; bar.sync below extends that to the block.
    // FLIPB200_HALO_ACQ=gpu keeps round 1's gpu-scope fence (works on this box, not guaranteed by the model).
    if (H.acq_sys) {
      if (w0) (void)ld_acquire_sys(&H.my_flags[0]);
      if (w1) (void)ld_acquire_sys(&H.my_flags[1]);
    } else
      fence_acq_rel_gpu();
    if (H.dbg && blockIdx.x == 0 && blockIdx.y == 0) H.dbg[w0 ? 2 : 3] = halo_now();
  }
  __syncthreads();
}

// Store the pair (v0, v1) written at element q of plane k into the neighbours' ghost planes.
template <typename T>
__device__ __forceinline__ void halo_push2(const PeerHalo& H, int q, int k, T v0, T v1) {
  if (H.mode == 0) return;
  if (H.push_dn && k == H.kdn) st2<T>(reinterpret_cast<T*>(H.dst_dn) + q, v0, v1);
  if (H.push_up && k == H.kup) st2<T>(reinterpret_cast<T*>(H.dst_up) + q, v0, v1);
}
__device__ __forceinline__ void halo_push4(const PeerHalo& H, int q, int k, float4 v) {
  if (H.mode == 0) return;
  if (H.push_dn && k == H.kdn) *reinterpret_cast<float4*>(reinterpret_cast<float*>(H.dst_dn) + q) = v;
  if (H.push_up && k == H.kup) *reinterpret_cast<float4*>(reinterpret_cast<float*>(H.dst_up) + q) = v;
}
// four values stored as fp16 (the level-0 residual of the V-cycle with params.mixed bit 1; q counts halves)
__device__ __forceinline__ void halo_push4h(const PeerHalo& H, int q, int k, float4 v) {
  if (H.mode == 0) return;
  if (H.push_dn && k == H.kdn) stq4<__half>(reinterpret_cast<__half*>(H.dst_dn) + q, v);
  if (H.push_up && k == H.kup) stq4<__half>(reinterpret_cast<__half*>(H.dst_up) + q, v);
}
template <typename T>
__device__ __forceinline__ void halo_push1(const PeerHalo& H, int q, int k, T v) {
  if (H.mode == 0) return;
  if (H.push_dn && k == H.kdn) reinterpret_cast<T*>(H.dst_dn)[q] = v;
  if (H.push_up && k == H.kup) reinterpret_cast<T*>(H.dst_up)[q] = v;
}

// Call at kernel end by ALL threads of every block (no early returns before it).
// bdn / bup: this block owns (part of) boundary plane kdn / kup.
__device__ __forceinline__ void halo_signal(const PeerHalo& H, DevScalars* ds, bool bdn, bool bup) {
  if (H.mode != 2) return;
  const bool s0 = bdn && H.push_dn, s1 = bup && H.push_up;
  if (!(s0 || s1)) return;
  // bar.sync orders the block's peer stores before thread 0's fence (cumulativity), so one thread
  // fences for the block -- at gpu scope: it only has to order the stores before the block counter
  // (a gpu-scope atomic).  The last block observes every other block through that counter (atomic,
  // then the fence inside its release = acquire side) and its system-scope release-store of the
  // sequence number is cumulative over all of them.  Measured (FLIPB200_HALO_DEBUG): a system-scope
  // fence is a 3.5 us round trip; with one per block plus one in front of the release the signalling
  // path of a boundary block (17 us) was longer than a whole level-1 sweep (12 us).
  __syncthreads();
  const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  if (tid == 0) {
    const bool dbg = H.dbg && blockIdx.x == 0 && blockIdx.y == 0;
    if (dbg) H.dbg[4] = halo_now();                  // block done computing + pushing
    fence_acq_rel_gpu();
    if (dbg) H.dbg[5] = halo_now();
    const unsigned seq = (unsigned)ds->halo_base + (unsigned)H.srel;
    if (s0 && atomicAdd(&H.cnt[0], 1u) == (unsigned)H.nb - 1u) { H.cnt[0] = 0u; st_release_sys(H.sig_dn, seq); if (H.dbg) H.dbg[6] = halo_now(); }
    if (s1 && atomicAdd(&H.cnt[1], 1u) == (unsigned)H.nb - 1u) { H.cnt[1] = 0u; st_release_sys(H.sig_up, seq); if (H.dbg) H.dbg[7] = halo_now(); }
  }
}

// ---- all-gather of the first replicated multigrid level by peer stores ----------------------------
// The restriction that feeds the first replicated level writes its planes of (b, xa) into EVERY
// rank's copy of the level (same layout everywhere), so no collective launch is needed: an NCCL
// all-gather of those two small arrays costs 26 us (2 GPUs) to 41 us (8 GPUs) of launch latency per
// V-cycle.  Every block pushes; the last block to finish fences once at system scope and stores the
// sequence number into slot [rank] of every rank's gflag array (in the mailbox allocation, which
// every rank has mapped).  Write-after-read: a rank reaches the next restriction only after the
// r.z exchange of this iteration, which every rank enters after its last reader of the level.
struct GatherPush {
  int on, mode, rank, world;
  int srel;                          // sequence number = ds->halo_base + srel
  int nblocks;                       // blocks of the launch (all of them push)
  float* b[RED_MAXR];                // [rank] level arrays (own entries unused)
  float* xa[RED_MAXR];
  unsigned* flag[RED_MAXR];          // [rank] -> that rank's gflag array
  unsigned* cnt;                     // block counter (own memory)
};
static inline GatherPush gather_off() { GatherPush g; memset(&g, 0, sizeof(g)); return g; }
__device__ __forceinline__ void gather_push1(const GatherPush& G, int q, float r, float x1) {
  if (!G.on) return;
  for (int p = 0; p < G.world; ++p)
    if (p != G.rank) { G.b[p][q] = r; G.xa[p][q] = x1; }
}
__device__ __forceinline__ void gather_push2(const GatherPush& G, int q, float r0, float r1, float x0, float x1) {
  if (!G.on) return;
  for (int p = 0; p < G.world; ++p)
    if (p != G.rank) { st2<float>(G.b[p] + q, r0, r1); st2<float>(G.xa[p] + q, x0, x1); }
}
// by ALL threads of EVERY block, at the end of the kernel
__device__ __forceinline__ void gather_signal(const GatherPush& G, DevScalars* ds) {
  if (!G.on || G.mode != 2) return;
  __syncthreads();
  const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  if (tid == 0) {
    fence_acq_rel_gpu();
    if (atomicAdd(G.cnt, 1u) == (unsigned)G.nblocks - 1u) {
      *G.cnt = 0u;
      fence_acq_rel_sys();                                     // cumulative over the block counter: every block's stores
      const unsigned seq = (unsigned)ds->halo_base + (unsigned)G.srel;
      for (int p = 0; p < G.world; ++p)
        if (p != G.rank) asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(G.flag[p] + G.rank), "r"(seq) : "memory");
    }
  }
}

// ---- all-reduce of a dot product over peer memory ------------------------------------------------
// Blocks add their partial sums into ds->part[id] as on one GPU.  One 32-thread kernel
// (k_pcg_exchange, pcg.cu) then does the whole all-reduce: lane r stores this rank's sum into slot
// [rank] of rank r's mailbox (peer store over NVLink), waits for slot [r] of its own mailbox, and the
// warp adds the `world` values in rank order (identical bits on every rank) into the ring slot the
// next kernel reads.  The mailbox words are self-validating (the NCCL "LL" idea): the double travels
// as two 8-byte words {32 value bits, 32-bit sequence number}; an aligned 8-byte store is single-copy
// atomic, so the reader just polls both words until they carry this exchange's sequence number --
// no fence on either side (a system-scope fence costs ~3.5 us on B200, measured with
// FLIPB200_HALO_DEBUG).  Replaces an 8-byte ncclAllReduce (17-23 us) plus a publish kernel by one
// launch and one NVLink one-way trip.  Slots are double-buffered by sequence parity; reuse is safe
// because a rank reaches exchange n+2 of an id only after every rank has finished n (the other dot
// products of the iteration depend on it).
// (Publishing from the last block of the producing kernel was tried first: the per-block
// fence + returning atomic made k_update_xr / k_spmv 5 % slower, a net loss.)
struct RedMail {
  unsigned long long w[4][2][RED_MAXR][2];   // [PART_* id][seq parity][source rank][lo, hi]: value half | seq << 32
  unsigned gflag[RED_MAXR];                  // all-gather by peer stores (GatherPush): [source rank] = sequence number
  // address of gflag through a (possibly peer-mapped) mailbox pointer, computed on the host
  unsigned* gflag_ptr() { return reinterpret_cast<unsigned*>(reinterpret_cast<char*>(this) + sizeof(w)); }
};
struct RedPub {
  int on, rank, world;
  RedMail* peers[RED_MAXR];          // [rank]; own mailbox included
};
static inline RedPub red_off() { RedPub r; memset(&r, 0, sizeof(r)); return r; }
RedPub fb_red_make(flipb200_ctx* c);   // comm.cu

// host (comm.cu): describe one launch as consumer of (lin, ain) and/or producer of (lout, aout)
PeerHalo fb_halo_make(flipb200_ctx* c, int lin, int ain, int lout, int aout, size_t elem_out, int kdn, int kup, int nb);
GatherPush fb_gather_make(flipb200_ctx* c, int nblocks);   // takes a sequence number
void fb_gather_wait(flipb200_ctx* c, PeerHalo* h);         // consumer of the latest gather

static inline PeerHalo halo_off() { PeerHalo h; memset(&h, 0, sizeof(h)); return h; }
