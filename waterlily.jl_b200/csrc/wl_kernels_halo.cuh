// wl_kernels_halo.cuh -- halo exchange between neighbouring slabs as direct stores into the
// neighbour's memory (NVLink peer access through CUDA IPC), one kernel per exchange.
//
// Every rank carves its fields from an arena with the same layout, so the ghost plane that must
// receive my boundary plane lives at the same offset in the neighbour's arena.  Per link (lower /
// upper neighbour) two monotonically increasing flags live in the RECEIVER's arena:
//    ready: "my ghost planes may be overwritten for exchange n"  (set when the rank's stream reaches
//           exchange n, i.e. after every kernel that read the ghosts of exchange n-1)
//    done : "exchange n has landed in your ghost planes"
// The kernel announces ready, waits for the neighbours' ready, copies, fences, and its last block
// raises the neighbours' done flags and waits for its own: when the kernel retires, the rank's ghost
// planes are current.  ~2 NVLink round trips instead of an NCCL send/recv group (~18 µs measured).
// All spins give up after 10 s and raise an error flag instead of hanging the GPU.
#pragma once
#include "wl_common.cuh"

namespace wl {

constexpr int HALO_MAXSEG = 8;
struct HaloSeg { const float* src; float* dst; };
struct HaloP2PArgs {
  HaloSeg seg[HALO_MAXSEG];
  int nseg;
  long long cnt;                       // floats per segment
  unsigned int* my;                    // local flags: [0] ready from lower, [1] ready from upper, [2] done from lower,
                                       // [3] done from upper (unused), [5] error, [6]/[7] arrival counters
  unsigned int* nb_lo;                 // lower neighbour's flag block (I am ITS upper neighbour: slots 1 and 3)
  unsigned int* nb_hi;                 // upper neighbour's flag block (slots 0 and 2)
  unsigned int seq_lo, seq_hi;
  int use_lo, use_hi;
                                       // flags: measured slower (TGV 128³ x2: 1.12 vs 1.04 ms/step), kept for reference
  int skip_lo, skip_hi;                // per link: the caller proved the ghost planes idle (see halo() in wl_b200.cu)
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_sys_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long wl_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// spin until *p has reached seq (wrap-safe); false on timeout
__device__ __forceinline__ bool wait_flag(const unsigned int* p, unsigned int seq, unsigned int* err) {
  const unsigned long long t0 = wl_now_ns();
  while ((int)(ld_acquire_sys(p) - seq) < 0) {
    if (wl_now_ns() - t0 > 10000000000ull) { atomicExch(err, 1u); return false; }
    __nanosleep(20);
  }
  return true;
}

// a.my + 16 (as unsigned long long[4]): accumulated ns spent waiting for ready / copying+fencing / waiting for
// done, and the number of exchanges (read by wl_halo_stats; block 0 / the last block only)
__global__ void __launch_bounds__(256) k_halo_p2p(const HaloP2PArgs a) {
  pdl_entry();
  const int tid = threadIdx.x;
  unsigned long long* stat = reinterpret_cast<unsigned long long*>(a.my + 16);
  const unsigned long long t_start = wl_now_ns();
  const bool rdy_lo = a.use_lo && !a.skip_lo, rdy_hi = a.use_hi && !a.skip_hi;
  if (rdy_lo || rdy_hi) {
    if (tid == 0) {
      if (blockIdx.x == 0) {
        if (rdy_lo) st_release_sys(a.nb_lo + 1, a.seq_lo);
        if (rdy_hi) st_release_sys(a.nb_hi + 0, a.seq_hi);
      }
      if (rdy_lo) wait_flag(a.my + 0, a.seq_lo, a.my + 5);
      if (rdy_hi) wait_flag(a.my + 1, a.seq_hi, a.my + 5);
    }
    __syncthreads();
  }
  const unsigned long long t_ready = wl_now_ns();
  if (tid == 0 && blockIdx.x == 0) atomicAdd(stat + 0, t_ready - t_start);
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long gt = (long long)blockIdx.x * blockDim.x + tid;
  for (int sgi = 0; sgi < a.nseg; ++sgi) {
    const float* __restrict__ src = a.seg[sgi].src;
    float* __restrict__ dst = a.seg[sgi].dst;
    // planes start 124 bytes past a 128-byte boundary (wl_common.cuh): one scalar, then aligned 16-byte
    // stores (source and destination share the layout, hence the alignment), then the tail
    const long long head = (4 - (((uintptr_t)src >> 2) & 3)) & 3;
    const long long nv = (a.cnt - head) / 4, tail0 = head + 4 * nv;
    if (((uintptr_t)src & 3) == 0 && ((((uintptr_t)src >> 2) & 3) == (((uintptr_t)dst >> 2) & 3))) {
      if (gt < head) dst[gt] = src[gt];
      const float4* s4 = reinterpret_cast<const float4*>(src + head);
      float4* d4 = reinterpret_cast<float4*>(dst + head);
      for (long long e = gt; e < nv; e += stride) d4[e] = s4[e];
      if (gt < a.cnt - tail0) dst[tail0 + gt] = src[tail0 + gt];
    } else {
      for (long long e = gt; e < a.cnt; e += stride) dst[e] = src[e];
    }
  }
  __syncthreads();
  // completion: one system-scope fence per block, a local ticket, and the last block raises the done flags
  __shared__ int last;
  if (tid == 0) __threadfence_system();
  if (tid == 0) last = atomicAdd(a.my + 4, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!last || tid != 0) return;
  a.my[4] = 0u;
  // every block fenced (system scope) before taking its ticket; this fence pairs with those (acquire side)
  // and orders the flag stores after the whole exchange (release side): one fence, two relaxed stores
  __threadfence_system();
  if (a.use_lo) st_relaxed_sys(a.nb_lo + 3, a.seq_lo);
  if (a.use_hi) st_relaxed_sys(a.nb_hi + 2, a.seq_hi);
  const unsigned long long t_copied = wl_now_ns();
  if (a.use_lo) wait_flag(a.my + 2, a.seq_lo, a.my + 5);
  if (a.use_hi) wait_flag(a.my + 3, a.seq_hi, a.my + 5);
  const unsigned long long t_done = wl_now_ns();
  atomicAdd(stat + 1, t_copied - t_ready);
  atomicAdd(stat + 2, t_done - t_copied);
  atomicAdd(stat + 3, 1ull);
}

// ---------------------------------------------------------------- reduction slots of all ranks, by peer stores
// Every rank stores its NRED reduction slots into each peer's gather table (arena header), raises a
// per-source flag there, and waits until all sources have raised its own: one small kernel (~2 NVLink
// latencies) in place of an NCCL all-gather.  Two tables alternate by sequence parity: a peer can be
// at most one gather ahead (it needs my flag to finish the current one), so the table it writes next
// is never the one my kernels are still reading.
constexpr int P2P_MAXRANKS = 8;
constexpr int GATHER_FLAG_WORD = 64;          // unsigned[P2P_MAXRANKS] in the arena header
constexpr int GATHER_TABLE_BYTE = 4096;       // Red[2][P2P_MAXRANKS][NRED]
constexpr int ARENA_HDR = 8192;
struct GatherArgs {
  const Red* mine;
  char* peer[P2P_MAXRANKS];                   // arenas of all ranks (own included)
  int n, rank;
  unsigned int seq;
};
__global__ void __launch_bounds__(128) k_gather_p2p(const GatherArgs a) {
  pdl_entry();
  const int tid = threadIdx.x;
  const int par = a.seq & 1;
  for (int e = tid; e < a.n * NRED; e += blockDim.x) {
    const int r = e / NRED, q = e - r * NRED;
    Red* dst = reinterpret_cast<Red*>(a.peer[r] + GATHER_TABLE_BYTE) + (par * P2P_MAXRANKS + a.rank) * NRED + q;
    *dst = a.mine[q];
  }
  __syncthreads();
  if (tid < a.n) {
    unsigned int* my = reinterpret_cast<unsigned int*>(a.peer[a.rank]);
    __threadfence_system();
    st_release_sys(reinterpret_cast<unsigned int*>(a.peer[tid]) + GATHER_FLAG_WORD + a.rank, a.seq);
    wait_flag(my + GATHER_FLAG_WORD + tid, a.seq, my + 5);
  }
}

}  // namespace wl
