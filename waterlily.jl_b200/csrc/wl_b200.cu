// wl_b200.cu -- host side of libwaterlily_b200.so: device state, kernel orchestration of
// mom_step! / solver! / update! / measure!, and the C ABI declared in include/wl_b200.h.
// There is no CPU fallback anywhere in this file: without an sm_100 device wl_create fails.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <utility>
#include <vector>

#include "../../include/wl_b200.h"
#include "wl_common.cuh"
#include "wl_nccl.cuh"
#include "wl_kernels_halo.cuh"
#include "wl_kernels_body.cuh"
#include "wl_kernels_mg.cuh"
#include "wl_kernels_mg2.cuh"
#include "wl_kernels_mom.cuh"
#include "wl_kernels_mom3.cuh"
#include "wl_kernels_mg4.cuh"
#include "wl_kernels_gs2.cuh"
#include "wl_kernels_pcg.cuh"

using namespace wl;

namespace {

std::string g_create_error;

struct DevField {
  float* alloc = nullptr;
  float* base = nullptr;   // element (0,0,0) of component 0
  Lay lay{};
  int ncomp = 0;
  size_t bytes = 0;
  bool in_arena = false;   // carved from the sim's peer-mapped arena (slab decomposition): not freed one by one
};

struct Level {
  Lay lay{};
  DevField L, x, z, r, eps, Dg, iD;
  DevField tmp;              // second ϵ buffer of the paired sweeps (K-GS2 cannot update ϵ in place)
  unsigned char* ucore = nullptr; unsigned char* uflag = nullptr;   // pattern-tile classes (TileFlags, wl_common.cuh)
  int* ulist = nullptr; int* ucount = nullptr; int ulist_n = 0;      // class-0/2 tiles: ids, device counter, count on the host
  int unx = 0, uny = 0, unz = 0;
  bool uvalid = false;       // classes match the level's current L (set by set_diag_h, cleared by uploads / measure!)
  float lu[3] = {1.f, 1.f, 1.f};   // the level's body-free face coefficient per direction (restrictL of all-ones)
  bool owns = false;
  int c[3] = {0, 0, 0};
  // multi-GPU: a level is either DISTRIBUTED (each rank holds its slab; `per` marks the interfaces
  // and koff is the global plane offset) or REPLICATED (every rank holds and solves the whole level)
  bool dist = false;
  int koff = 0;
  Per per{};
  long long ncells = 0;      // interior cells of the global level
};

constexpr int MAXRANKS = 64;

// Slab decomposition along the last dimension, one process per GPU.
struct Dist {
  int rank = 0, n = 1;
  ncclComm_t comm = nullptr;
  int lo = -1, hi = -1;                    // neighbour ranks, -1 at a physical (non-periodic) boundary
  bool lo_wrap = false, hi_wrap = false;   // neighbour reached through the periodic wrap
  int nzg = 0, nzl = 0;                    // global / local interior planes
  int min_planes = 4;                      // levels with fewer local planes are replicated
  long long min_cells = 1 << 20;           // ... and so are levels with at most this many cells in total
};

struct ProfRec { const char* name; cudaEvent_t a, b; long long blocks; };

}  // namespace

struct wl_sim {
  wl_config cfg{};
  int D = 3;
  Lay lay{};
  Per per{};                  // what the kernels of the (local) fine grid see
  Per per_glob{};             // true periodicity of the global grid (replicated levels)
  Dist dist{};
  // peer-mapped arena (slab decomposition): every rank carves the same fields at the same offsets, so
  // a neighbour's copy of any field is peer_base + (ptr - arena).  Halo planes are then stored
  // straight into the neighbour's ghost planes over NVLink (k_halo_p2p) instead of NCCL send/recv.
  char* arena = nullptr;
  size_t arena_cap = 0, arena_used = 0, arena_mark = 0;
  char* peer_lo = nullptr; char* peer_hi = nullptr;   // neighbours' arenas mapped with cudaIpcOpenMemHandle
  char* peer_all[P2P_MAXRANKS] = {};                  // every rank's arena (own included) when nranks <= P2P_MAXRANKS
  bool p2p_gather = false;                            // reduction slots travel by k_gather_p2p instead of NCCL
  unsigned int gseq = 0;
  DevField accf[2];                                   // accelerate! fields at t₀ / t₁ (wl_set_accel_field), else unallocated
  bool accf_on = false;
  float* bc_tab = nullptr; int bc_off[9] = {0}; int bc_psz[3] = {0, 0, 0};   // Function-uBC tables (wl_set_bc_tables)
  unsigned int gs_seq_lo = 0, gs_seq_hi = 0;          // fused half-sweep pushes per link (k_gs_half_cx)
  bool p2p = false;
  unsigned int seq_lo = 0, seq_hi = 0;                // exchanges done on the lower / upper link
  float* stage = nullptr;     // packed device buffer for upload/download (grown on demand)
  size_t stage_bytes = 0;
  Red* d_redg = nullptr;      // all-gathered reduction slots [rank][NRED]
  Red* h_redg = nullptr;
  int64_t halo_msgs = 0, halo_bytes = 0;
  DevField u, u0, f, p, sigma, V, mu0, mu1;
  DevField w;                 // third rotating velocity buffer of the fused 3D K-MOM path
  int* band_list = nullptr;   // interior cells with μ₁ ≠ 0 (patched after K-MOM)
  int* band_count = nullptr;
  int band_cap = 0;
  DevField mf_P, mf_U, mf_UU;  // MeanFlow (Metrics.jl:205-263): allocated by wl_meanflow_init
  int* band_mark = nullptr;   // per cell: epoch of the last list2 build that took it (de-duplication)
  int* band_list2 = nullptr;  // band cells and their face neighbours: the cells whose f the patch needs
  int* band_count2 = nullptr;
  int band_cap2 = 0, band_epoch = 0;
  bool band_dirty = false;
  bool band_from_measure = false;   // k_measure appended the band cells itself: only the count has to be checked
  std::map<const float*, CUtensorMap> tmaps;       // 4-D tensor maps of the vector fields K-MOM fetches by TMA, keyed by allocation
  int tma_ok = -1;                                 // -1 not probed, 0 unavailable (encoder missing / WL_NO_TMA), 1 in use
  int* fw_off = nullptr; float* fw_w = nullptr; int fw_n = -1;   // pressure_force weights of a host-measured body (wl_set_force_weights)
  bool free_flow = true;      // V ≡ 0, μ₀ ≡ 1, μ₁ ≡ 0 as the constructor left them (no measure!, no upload of them)
  std::vector<Level> lv;
  cudaStream_t st = nullptr;
  RedWork redw{};
  Red* d_red = nullptr;
  Red* h_red = nullptr;
  float dt = 0.25f;
  std::string err;
  int64_t launches = 0;
  int device = 0;
  bool timing = false;
  cudaEvent_t ev[8] = {};
  double phase_s[8] = {};
  int status = WL_OK;
  // profiling hooks
  bool prof_on = false, prof_split = false;
  std::string prof_filter;
  std::vector<cudaEvent_t> evpool;
  std::vector<ProfRec> recs;
  cudaEvent_t uev[8] = {};
};

namespace {

#define WL_CUDA(s, call)                                                                   \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess) {                                                               \
      (s)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                       \
      (s)->status = WL_ERR_CUDA;                                                           \
      return WL_ERR_CUDA;                                                                  \
    }                                                                                      \
  } while (0)

int alloc_field(wl_sim* s, DevField& f, const Lay& lay, int ncomp) {
  f.lay = lay; f.ncomp = ncomp;
  f.bytes = (size_t)lay.cstride * ncomp * sizeof(float) + 256;
  if (s->arena) {
    const size_t need = (f.bytes + 255) & ~(size_t)255;
    if (s->arena_used + need > s->arena_cap) {
      s->err = "peer arena exhausted (layout plan and allocation disagree)";
      s->status = WL_ERR_CUDA;
      return WL_ERR_CUDA;
    }
    f.alloc = reinterpret_cast<float*>(s->arena + s->arena_used);
    s->arena_used += need;
    f.in_arena = true;
  } else {
    WL_CUDA(s, cudaMalloc(&f.alloc, f.bytes));
  }
  WL_CUDA(s, cudaMemsetAsync(f.alloc, 0, f.bytes, s->st));
  f.base = f.alloc + LPAD + (size_t)lay.zpad * lay.plane;
  return WL_OK;
}
void free_field(DevField& f) { if (f.alloc && !f.in_arena) cudaFree(f.alloc); f.alloc = f.base = nullptr; f.in_arena = false; }

dim3 blk(int D) { return D == 3 ? dim3(32, 4, 2) : dim3(32, 8, 1); }   // even extents: K-DOWN keeps the children of a coarse cell in one block
dim3 grid_inside(const Lay& l) {
  dim3 b = blk(l.D);
  return dim3((l.n0 - 2 + b.x - 1) / b.x, (l.n1 - 2 + b.y - 1) / b.y, l.D == 3 ? (l.n2 - 2 + b.z - 1) / b.z : 1);
}
dim3 grid_all(const Lay& l) {
  dim3 b = blk(l.D);
  return dim3((l.n0 + LPAD + 31) / 32, (l.n1 + b.y - 1) / b.y, l.D == 3 ? (l.n2 + b.z - 1) / b.z : 1);
}
// 1-D persistent grid for the kernels built on for_tiles_inside / for_tiles_all: at most 8 blocks per SM
dim3 grid_tiles(const Lay& l, bool all) {
  const dim3 g = all ? grid_all(l) : grid_inside(l);
  const long long nt = (long long)g.x * g.y * g.z;
  return dim3((unsigned)(nt < 148 * 8 ? nt : 148 * 8));
}
// plane kernels: threads over (fast in-plane dim, slow in-plane dim); sized for the largest plane
void plane_launch(const Lay& l, dim3& g, dim3& b) {
  int m = l.n0 > l.n1 ? l.n0 : l.n1;
  int m2 = l.D == 3 ? (l.n1 > l.n2 ? l.n1 : l.n2) : 1;
  b = l.D == 3 ? dim3(32, 8, 1) : dim3(128, 1, 1);
  g = dim3((m + b.x - 1) / b.x, (m2 + b.y - 1) / b.y, 1);
}

// Every kernel goes through here: a launch that may overlap its scheduling with the tail of the
// previous kernel of the stream (see pdl_entry in wl_common.cuh).  WL_NO_PDL=1 restores plain launches.
template <class... KArgs, class... Args>
inline void launch_pdl_smem(cudaStream_t st, dim3 g, dim3 b, size_t smem, void (*kern)(KArgs...), Args&&... args) {
  static const bool off = getenv("WL_NO_PDL") != nullptr;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = g; cfg.blockDim = b; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = off ? 0 : 1;
  cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}
template <class... KArgs, class... Args>
inline void launch_pdl(cudaStream_t st, dim3 g, dim3 b, void (*kern)(KArgs...), Args&&... args) {
  static const bool off = getenv("WL_NO_PDL") != nullptr;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = g; cfg.blockDim = b; cfg.dynamicSmemBytes = 0; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = off ? 0 : 1;
  cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

// Optional per-kernel CUDA-event timing on the launching stream (wl_prof_begin/wl_prof_end).
inline bool prof_pre(wl_sim* s, const char* name, cudaEvent_t* a, cudaEvent_t* b) {
  if (!s->prof_on) return false;
  if (!s->prof_filter.empty() && !strstr(name, s->prof_filter.c_str())) return false;
  if (s->evpool.size() < 2) {
    for (int i = 0; i < 64; ++i) { cudaEvent_t e; cudaEventCreate(&e); s->evpool.push_back(e); }
  }
  *a = s->evpool.back(); s->evpool.pop_back();
  *b = s->evpool.back(); s->evpool.pop_back();
  cudaEventRecord(*a, s->st);
  return true;
}
inline void prof_post(wl_sim* s, const char* name, cudaEvent_t a, cudaEvent_t b, const dim3& g = dim3(0)) {
  cudaEventRecord(b, s->st);
  s->recs.push_back({name, a, b, (long long)g.x * g.y * g.z});
}

#define LAUNCH(s, kern, g, b, ...)                         \
  do {                                                     \
    cudaEvent_t pa_, pb_;                                  \
    const bool prof_ = prof_pre(s, #kern, &pa_, &pb_);     \
    const dim3 g_ = (g);                                   \
    launch_pdl((s)->st, g_, b, kern, __VA_ARGS__);         \
    if (prof_) prof_post(s, #kern, pa_, pb_, g_);          \
    ++(s)->launches;                                       \
  } while (0)

#define LAUNCH_SMEM(s, kern, g, b, smem, ...)                          \
  do {                                                                \
    cudaEvent_t pa_, pb_;                                             \
    const bool prof_ = prof_pre(s, #kern, &pa_, &pb_);                \
    const dim3 g_ = (g);                                              \
    launch_pdl_smem((s)->st, g_, b, smem, kern, __VA_ARGS__);         \
    if (prof_) prof_post(s, #kern, pa_, pb_, g_);                     \
    ++(s)->launches;                                                  \
  } while (0)

#define DISPATCH_D(s, kern, g, b, ...)                                    \
  do {                                                                    \
    if ((s)->D == 3) LAUNCH(s, kern<3>, g, b, __VA_ARGS__);               \
    else LAUNCH(s, kern<2>, g, b, __VA_ARGS__);                           \
  } while (0)

#define DISPATCH_DB(s, kern, B, g, b, ...)                                \
  do {                                                                    \
    if ((s)->D == 3) LAUNCH(s, (kern<3, B>), g, b, __VA_ARGS__);          \
    else LAUNCH(s, (kern<2, B>), g, b, __VA_ARGS__);                      \
  } while (0)

#define DISPATCH_DS(s, kern, g, b, ...)                                                   \
  do {                                                                                    \
    const int sc_ = (s)->cfg.scheme;                                                      \
    if ((s)->D == 3) {                                                                    \
      if (sc_ == 0) LAUNCH(s, (kern<3, 0>), g, b, __VA_ARGS__);                           \
      else if (sc_ == 1) LAUNCH(s, (kern<3, 1>), g, b, __VA_ARGS__);                      \
      else LAUNCH(s, (kern<3, 2>), g, b, __VA_ARGS__);                                    \
    } else {                                                                              \
      if (sc_ == 0) LAUNCH(s, (kern<2, 0>), g, b, __VA_ARGS__);                           \
      else if (sc_ == 1) LAUNCH(s, (kern<2, 1>), g, b, __VA_ARGS__);                      \
      else LAUNCH(s, (kern<2, 2>), g, b, __VA_ARGS__);                                    \
    }                                                                                     \
  } while (0)

#define WL_NCCL(s, call)                                                                   \
  do {                                                                                     \
    ncclResult_t r_ = (call);                                                              \
    if (r_ != ncclSuccess) {                                                               \
      (s)->err = std::string(#call) + ": " + ncclGetErrorString(r_);                       \
      (s)->status = WL_ERR_CUDA;                                                           \
      return WL_ERR_CUDA;                                                                  \
    }                                                                                      \
  } while (0)

static bool getenv_flag(const char* name) {
  static std::map<std::string, bool> cache;
  auto it = cache.find(name);
  if (it != cache.end()) return it->second;
  const bool v = getenv(name) != nullptr;
  cache[name] = v;
  return v;
}
// ---------------------------------------------------------------- slab halos and global scalars
// Exchange `depth` planes per side of every component of `f` with the slab neighbours, on the
// sim's stream: interior planes [1,1+depth) go to the lower neighbour's [n2-1,n2-1+depth) and
// [n2-1-depth,n2-1) to the upper neighbour's [1-depth,1).  Whole planes travel (x/y ghosts
// included), which is what the sequential BC!/perBC! passes of the reference produce for the last
// direction (core.jl:207-217,239-243).  wrap=false skips the pair that closes a periodic ring (the
// reference's red-black sweeps never refresh that ghost, Poisson.jl:141-148).
// idle=true: no kernel since the previous exchange of `f` with the same neighbours can observe a change in
// the ghost planes this exchange rewrites, so the "ready" handshake is skipped.  Used between red-black
// half-sweeps: the sweep that ran since the last exchange read only ghost cells of the OTHER colour, whose
// values this exchange rewrites unchanged, and the neighbour cannot be further behind than that sweep
// (it left the previous exchange only after receiving this rank's planes).
// only_wrap: just the links that close a periodic ring (the others were served by fused kernels).
int halo(wl_sim* s, const DevField& f, int depth, bool wrap = true, bool idle = false, bool only_wrap = false) {
  const Dist& d = s->dist;
  if (d.n == 1) return WL_OK;
  const Lay& l = f.lay;
  const size_t cnt = (size_t)l.plane * depth;
  const bool use_lo = d.lo >= 0 && (wrap || !d.lo_wrap) && (!only_wrap || d.lo_wrap),
             use_hi = d.hi >= 0 && (wrap || !d.hi_wrap) && (!only_wrap || d.hi_wrap);
  if (!use_lo && !use_hi) return WL_OK;
  s->halo_msgs += 1;
  s->halo_bytes += (int64_t)cnt * 4 * f.ncomp * ((use_lo ? 1 : 0) + (use_hi ? 1 : 0));
  if (s->p2p && f.in_arena && 2 * f.ncomp <= HALO_MAXSEG) {
    // direct stores into the neighbours' ghost planes (wl_kernels_halo.cuh)
    HaloP2PArgs a{};
    auto peer = [&](char* pb, const float* mine) { return reinterpret_cast<float*>(pb + (reinterpret_cast<const char*>(mine) - s->arena)); };
    for (int c = 0; c < f.ncomp; ++c) {
      float* b = f.base + l.cstride * c;
      if (use_lo) a.seg[a.nseg++] = HaloSeg{b + (size_t)l.plane * 1, peer(s->peer_lo, b + (size_t)l.plane * (l.n2 - 1))};
      if (use_hi) a.seg[a.nseg++] = HaloSeg{b + (size_t)l.plane * (l.n2 - 1 - depth), peer(s->peer_hi, b + (ptrdiff_t)l.plane * (1 - depth))};
    }
    a.cnt = (long long)cnt;
    a.my = reinterpret_cast<unsigned int*>(s->arena);
    a.nb_lo = reinterpret_cast<unsigned int*>(s->peer_lo);
    a.nb_hi = reinterpret_cast<unsigned int*>(s->peer_hi);
    a.use_lo = use_lo; a.use_hi = use_hi;
    // (a link that closes a periodic ring sat out the exchanges between the sweeps: it keeps its handshake)
    a.skip_lo = idle && !d.lo_wrap; a.skip_hi = idle && !d.hi_wrap;
    if (use_lo) a.seq_lo = ++s->seq_lo;
    if (use_hi) a.seq_hi = ++s->seq_hi;
    long long nb = ((long long)cnt * a.nseg + 2047) / 2048;
    nb = nb < 1 ? 1 : (nb > 296 ? 296 : nb);
    LAUNCH(s, k_halo_p2p, dim3((unsigned)nb), dim3(256), a);
    return WL_OK;
  }
  WL_NCCL(s, ncclGroupStart());
  for (int c = 0; c < f.ncomp; ++c) {
    float* b = f.base + l.cstride * c;
    // sends in (lower, upper) order, receives in (upper, lower) order: with two ranks on a periodic
    // ring both neighbours are the same peer and messages match in posting order
    if (use_lo) WL_NCCL(s, ncclSend(b + (size_t)l.plane * 1, cnt, ncclFloat, d.lo, d.comm, s->st));
    if (use_hi) WL_NCCL(s, ncclSend(b + (size_t)l.plane * (l.n2 - 1 - depth), cnt, ncclFloat, d.hi, d.comm, s->st));
    if (use_hi) WL_NCCL(s, ncclRecv(b + (size_t)l.plane * (l.n2 - 1), cnt, ncclFloat, d.hi, d.comm, s->st));
    if (use_lo) WL_NCCL(s, ncclRecv(b + (ptrdiff_t)l.plane * (1 - depth), cnt, ncclFloat, d.lo, d.comm, s->st));
  }
  WL_NCCL(s, ncclGroupEnd());
  return WL_OK;
}
// The reduction slots of all ranks, gathered so that every consumer folds them in rank order
// (results do not depend on the collective's algorithm).  Single GPU: the slots themselves.
Red* gather_table(wl_sim* s) {      // the table the LAST gather filled (consumers are enqueued after it)
  if (!s->p2p_gather) return s->d_redg;
  return reinterpret_cast<Red*>(s->arena + GATHER_TABLE_BYTE) + (size_t)(s->gseq & 1) * P2P_MAXRANKS * NRED;
}
RedIn red_in(wl_sim* s, int slot) {
  if (s->dist.n == 1) return RedIn{s->d_red + slot, 1, NRED};
  return RedIn{gather_table(s) + slot, s->dist.n, NRED};
}
int gather_red(wl_sim* s) {
  if (s->dist.n == 1) return WL_OK;
  if (s->p2p_gather) {
    GatherArgs a{};
    a.mine = s->d_red; a.n = s->dist.n; a.rank = s->dist.rank; a.seq = ++s->gseq;
    for (int r = 0; r < s->dist.n; ++r) a.peer[r] = s->peer_all[r];
    LAUNCH(s, k_gather_p2p, dim3(1), dim3(128), a);
    return WL_OK;
  }
  WL_NCCL(s, ncclAllGather(s->d_red, s->d_redg, sizeof(Red) * NRED, ncclChar, s->dist.comm, s->st));
  return WL_OK;
}

// ---------------------------------------------------------------- boundary conditions
void bc_vec(wl_sim* s, DevField& a, const float* U, bool saveexit, const Per* per = nullptr) {
  if (U && s->bc_tab && &a == &s->u) {                 // Function uBC sampled into tables (wl_set_bc_tables)
    BcTabArgs t;
    t.l = a.lay; t.a = a.base; t.per = s->per; t.saveexit = saveexit ? 1 : 0; t.tab = s->bc_tab;
    for (int q = 0; q < 9; ++q) t.off[q] = s->bc_off[q];
    for (int d = 0; d < 3; ++d) t.psz[d] = s->bc_psz[d];
    dim3 g, b;
    plane_launch(a.lay, g, b);
    g.z = 3 * s->D;
    DISPATCH_D(s, k_bc_vec_tab, g, b, t);
    return;
  }
  BcArgs args;
  args.l = a.lay; args.a = a.base; args.per = per ? *per : s->per; args.saveexit = saveexit ? 1 : 0;
  for (int d = 0; d < 3; ++d) args.U.v[d] = U ? U[d] : 0.f;
  dim3 g, b;
  plane_launch(a.lay, g, b);
  g.z = 3 * s->D;
  DISPATCH_D(s, k_bc_vec, g, b, args);
}
void perbc(wl_sim* s, const Lay& l, float* a, const Per* per = nullptr) {
  const Per& pp = per ? *per : s->per;
  for (int d = 0; d < s->D; ++d) {
    if (!pp.p[d]) continue;
    PerArgs args{l, a, d};
    dim3 g, b;
    plane_launch(l, g, b);
    DISPATCH_D(s, k_perbc, g, b, args);
  }
}
int exitbc(wl_sim* s, DevField& u, DevField& u0, float dt) {
  if (s->dist.n == 1) {
    ExitArgs args{u.lay, u.base, u0.base, dt};
    DISPATCH_D(s, k_exitbc, dim3(1), dim3(1024), args);
    return WL_OK;
  }
  const long long len = (long long)(u.lay.n1 - 2) * s->dist.nzg;
  for (int stage = 0; stage < 3; ++stage) {
    ExitDistArgs a{u.lay, u.base, u0.base, dt, stage, len, red_in(s, 8), red_in(s, 9), s->d_red + 8 + (stage == 1)};
    LAUNCH(s, k_exitbc_dist, dim3(1), dim3(1024), a);
    if (stage < 2) { if (int e = gather_red(s)) return e; }
  }
  return WL_OK;
}

// ---------------------------------------------------------------- multigrid pieces
TileFlags tile_flags(const Level& p) {
  static const bool off = getenv("WL_NO_TILE_FLAGS") != nullptr;      // A/B switch: every tile takes the general path
  TileFlags t{nullptr, 0, 0, nullptr, {p.lu[0], p.lu[1], p.lu[2]}, 0.f, 0.f};
  float dg = 0.f;                                                    // set_diag! on an all-pattern interior cell
  dg -= p.lu[0] + p.lu[0]; dg -= p.lu[1] + p.lu[1]; dg -= p.lu[2] + p.lu[2];
  t.idu = dg == 0.f ? 0.f : 1.f / dg;
  float d1 = 0.f;                                                    // ... and on one next to an x boundary: (0 + lu) or
  d1 -= p.lu[0]; d1 -= p.lu[1] + p.lu[1]; d1 -= p.lu[2] + p.lu[2];   // (lu + 0) is lu exactly
  t.idx1 = d1 == 0.f ? 0.f : 1.f / d1;
  if (p.uvalid && p.uflag && !off) { t.f = p.uflag; t.nx = p.unx; t.ny = p.uny; t.list = p.ulist; }
  return t;
}
LevArgs lev_args(wl_sim* s, Level& p) {
  LevArgs a{};
  a.l = p.lay; a.L = p.L.base; a.x = p.x.base; a.r = p.r.base; a.eps = p.eps.base; a.z = p.z.base;
  a.omega = 1.f; a.k0 = 0; a.w = s->redw; a.out = s->d_red; a.in = red_in(s, 0); a.want_norms = 0;
  a.ncells = p.ncells;
  a.uf = tile_flags(p); a.per = p.per;
  return a;
}
void set_diag(wl_sim* s, Level& p) {
  DiagArgs a{p.lay, p.L.base, p.Dg.base, p.iD.base};
  DISPATCH_D(s, k_set_diag, grid_inside(p.lay), blk(s->D), a);
}
void increment(wl_sim* s, Level& p, float w, Red* norms) {
  perbc(s, p.lay, p.eps.base, &p.per);
  LevArgs a = lev_args(s, p);
  a.omega = w; a.want_norms = norms ? 1 : 0; a.out = norms ? norms : s->d_red;
  DISPATCH_D(s, k_increment, grid_inside(p.lay), blk(s->D), a);
}
// Jacobi!(p) with it=1, ω=1 (Poisson.jl:111-114)
void jacobi(wl_sim* s, Level& p) {
  LevArgs a = lev_args(s, p);
  DISPATCH_D(s, k_eps_init, grid_inside(p.lay), blk(s->D), a);
  increment(s, p, 1.f, nullptr);
}
// GaussSeidelRB!(p;it=4,ω) (Poisson.jl:141-148)
void smooth(wl_sim* s, Level& p, float w, Red* norms) {
  LevArgs a = lev_args(s, p);
  DISPATCH_D(s, k_eps_init, grid_inside(p.lay), blk(s->D), a);
  perbc(s, p.lay, p.eps.base, &p.per);
  for (int k0 = 1; k0 <= 4; ++k0) {
    a.k0 = k0;
    DISPATCH_D(s, k_gs_half, grid_inside(p.lay), blk(s->D), a);
  }
  increment(s, p, w, norms);
}
void restrict_r(wl_sim* s, Level& fi, Level& co) {
  XferArgs a{fi.lay, co.lay, fi.r.base, co.r.base, {co.c[0], co.c[1], co.c[2]}};
  DISPATCH_D(s, k_restrict, grid_inside(co.lay), blk(s->D), a);
}
void prolongate(wl_sim* s, Level& fi, Level& co) {
  XferArgs a{fi.lay, co.lay, co.x.base, fi.eps.base, {co.c[0], co.c[1], co.c[2]}};
  DISPATCH_D(s, k_prolongate, grid_inside(fi.lay), blk(s->D), a);
}
// restrictL! (MultiLevelPoisson.jl:42-47).  Distributed fine level over a replicated coarse one: every
// rank fills its plane range of the coarse array, the ranges are all-gathered in place, then BC!.
int restrictL(wl_sim* s, Level& fi, Level& co) {
  const bool gather = fi.dist && !co.dist;
  const int nzc = co.c[2] ? s->dist.nzl >> 0 : 0;   // (unused; the local coarse extent follows from fi.lay)
  (void)nzc;
  Lay lc = co.lay;
  float* Lc = co.L.base;
  int nloc = 0;
  if (gather) {                                     // local view of the replicated coarse array
    nloc = co.c[2] ? (fi.lay.n2 - 2) / 2 : fi.lay.n2 - 2;
    lc.n2 = nloc + 2;
    Lc += (size_t)co.lay.plane * (s->dist.rank * nloc);
  }
  XferLArgs a{fi.lay, lc, fi.L.base, Lc, {co.c[0], co.c[1], co.c[2]}};
  DISPATCH_D(s, k_restrictL, grid_inside(lc), blk(s->D), a);
  if (gather) {
    const size_t cnt = (size_t)co.lay.plane * nloc;
    for (int c = 0; c < s->D; ++c) {
      float* b = co.L.base + co.lay.cstride * c + co.lay.plane;
      WL_NCCL(s, ncclAllGather(b + cnt * s->dist.rank, b, cnt, ncclFloat, s->dist.comm, s->st));
    }
  }
  bc_vec(s, co.L, nullptr, false, &co.per);
  if (co.dist) return halo(s, co.L, 1);
  return WL_OK;
}
// set_diag! + the halo of iD that the fused kernels read at neighbour cells
int set_diag_h(wl_sim* s, Level& p) {
  set_diag(s, p);
  // pattern-tile classes of the level (its L, ghosts and slab halos included, is final here)
  p.uvalid = false;
  if (s->D == 3 && vec4_ok(p.lay)) {
    const int nx = (p.lay.n0 - 2 + UT_X - 1) / UT_X, ny = (p.lay.n1 - 2 + UT_Y - 1) / UT_Y, nz = (p.lay.n2 - 2 + UT_Z - 1) / UT_Z;
    if (!p.uflag || nx != p.unx || ny != p.uny || nz != p.unz) {
      if (p.ucore) cudaFree(p.ucore);
      if (p.uflag) cudaFree(p.uflag);
      if (p.ulist) cudaFree(p.ulist);
      if (p.ucount) cudaFree(p.ucount);
      p.ucore = p.uflag = nullptr; p.ulist = p.ucount = nullptr;
      WL_CUDA(s, cudaMalloc(&p.ucore, (size_t)nx * ny * nz));
      WL_CUDA(s, cudaMalloc(&p.uflag, (size_t)nx * ny * nz));
      WL_CUDA(s, cudaMalloc(&p.ulist, sizeof(int) * (size_t)nx * ny * nz));
      WL_CUDA(s, cudaMalloc(&p.ucount, sizeof(int)));
      p.unx = nx; p.uny = ny; p.unz = nz;
    }
    TileArgs t{p.lay, p.L.base, p.ucore, p.uflag, nx, ny, nz, p.per, {p.lu[0], p.lu[1], p.lu[2]}, p.ulist, p.ucount};
    LAUNCH(s, k_tile_core, dim3(nx, ny, nz), dim3(256), t);
    LAUNCH(s, k_tile_class, dim3((unsigned)std::min((nx * ny * nz + 255) / 256, 1184)), dim3(256), t);
    LAUNCH(s, k_tile_list, dim3(1), dim3(1024), t);
    // the number of class-0/2 tiles sizes the second launch of the split kernels: one small read-back per update!
    WL_CUDA(s, cudaMemcpyAsync(&p.ulist_n, p.ucount, sizeof(int), cudaMemcpyDeviceToHost, s->st));
    WL_CUDA(s, cudaStreamSynchronize(s->st));
    p.uvalid = true;
  }
  if (p.dist) return halo(s, p.iD, 1);
  return WL_OK;
}
// Vcycle!(ml;l,ω) (MultiLevelPoisson.jl:88-101), v1 kernels (kept for WL_FLAG_NO_FUSED)
void vcycle_v1(wl_sim* s, size_t l, float w) {
  Level &fine = s->lv[l], &coarse = s->lv[l + 1];
  jacobi(s, fine);
  restrict_r(s, fine, coarse);
  cudaMemsetAsync(coarse.x.alloc, 0, coarse.x.bytes, s->st);
  if (l + 2 < s->lv.size()) vcycle_v1(s, l + 1, w);
  smooth(s, coarse, w, nullptr);
  prolongate(s, fine, coarse);
  increment(s, fine, w, nullptr);
}

// ---- v2: fused kernels (wl_kernels_mg2.cuh)
// L2 prefetch distance of the 4-cells-per-thread kernels (pf4, wl_kernels_mg4.cuh): ~1.5 waves of blocks (5 resident per
// SM) in planes (a block layer is 2 planes thick); off on levels that fit the L2 anyway.  WL_PREFETCH=<planes> overrides.
int prefetch_planes(const Lay& l) {
  static const int env = getenv("WL_PREFETCH") ? atoi(getenv("WL_PREFETCH")) : -1;
  if (l.D != 3 || !vec4_ok(l)) return 0;
  const long long per_layer = (long long)((l.n0 - 2 + 127) / 128) * ((l.n1 - 2 + 3) / 4), layers = (l.n2 - 2 + 1) / 2;
  if (per_layer * layers < 148 * 5 * 4) return 0;
  if (env >= 0) return env;
  return (int)std::min<long long>(16, 2 * ((148 * 5 * 3 / 2 + per_layer - 1) / per_layer));
}
LevDev levdev(const Level& p) {
  LevDev d{};
  d.l = p.lay; d.L = p.L.base; d.iD = p.iD.base; d.x = p.x.base; d.r = p.r.base; d.eps = p.eps.base;
  for (int q = 0; q < 3; ++q) d.c[q] = p.c[q];
  d.uf = tile_flags(p);
  d.pf = prefetch_planes(p.lay);
  return d;
}
// the rank's window into a REPLICATED coarse level as seen from the DISTRIBUTED level above it:
// local plane k of the window is global plane k + koff of the coarse array
LevDev levdev_window(wl_sim* s, const Level& fine, const Level& co) {
  LevDev d = levdev(co);
  if (fine.dist && !co.dist) {
    const int nloc = co.c[2] ? (fine.lay.n2 - 2) / 2 : fine.lay.n2 - 2;
    const size_t sh = (size_t)co.lay.plane * (s->dist.rank * nloc);
    d.l.n2 = nloc + 2;
    d.x += sh; d.r += sh; d.eps += sh;
  }
  return d;
}
// one block of 1024 threads sweeps a level in (cells/1024) dependent round trips to L2 per phase (~1 µs
// per phase whatever the size): with dependent launches costing ~2.5 µs per kernel it pays up to ~10³
// (measured on TGV 128³: limit 8000 -> 0.311, 2000 -> 0.298, 300 -> 0.303 ns/cell/step)
inline bool level_is_small(const Level& p) {
  static const long long lim = getenv("WL_SMALL_CELLS") ? atoll(getenv("WL_SMALL_CELLS")) : 2000;
  return !p.dist && (long long)p.lay.n0 * p.lay.n1 * p.lay.n2 <= lim;
}
// GaussSeidelRB!(p;it=4,ω): ϵ init + ghost images, 4 half sweeps on stale ghosts, wrap-read increment.
// Distributed levels: the ϵ ghost planes at slab interfaces are refreshed after every half sweep (a
// colour only reads the other colour, so this reproduces the single-domain sweep exactly); the
// planes across a periodic wrap are sent once after the init and once before the increment, which
// is when the reference's perBC!(ϵ) runs.
// x-vectorised pressure kernels (wl_kernels_mg4.cuh): fixed 32x4x2 blocks of quads
bool use_vec4(const Lay& l) {
  static int off = -1;
  if (off < 0) off = getenv("WL_NO_VEC4") ? 1 : 0;
  return !off && vec4_ok(l);
}
dim3 grid_quads(const Lay& l, int cells_per_thread = 4) {
  return dim3((l.n0 - 2 + 32 * cells_per_thread - 1) / (32 * cells_per_thread), (l.n1 - 2 + 3) / 4, (l.n2 - 2 + 1) / 2);
}
// Reductions of the vectorised kernels: every block leaves one partial, a k_fold_partials launch folds them in
// index order (with dependent launches that is cheaper than a fence + atomic ticket in every block: TGV 128³ 0.301 vs
// 0.314, sphere 512x256² 0.182 vs 0.189 ns/cell/step, round 1).
void fold_partials_to(wl_sim* s, const dim3& g, Red* out) {
  FoldArgs fa{s->redw, (int)(g.x * g.y * g.z), out};
  LAUNCH(s, k_fold_partials, dim3((unsigned)((fa.n + FOLD_CHUNK - 1) / FOLD_CHUNK)), dim3(1024), fa);
}

// K-DIVRES / K-DOWN as two launches when the level has tile classes (see split_tile in wl_kernels_mg4.cuh): the lean
// instantiation over the grid of tiles, then the general one over the class-0/2 list.  WL_NO_SPLIT: one launch (A/B).
inline bool split_ok(const Level& p) {
  static const bool off = getenv("WL_NO_SPLIT") != nullptr;
  // (below ~16 k tiles the second launch costs what the leaner first one saves: measured on the 256x128x128 level)
  static const long long min_tiles = getenv("WL_SPLIT_MIN_TILES") ? atoll(getenv("WL_SPLIT_MIN_TILES")) : 16384;
  return !off && p.uvalid && p.uflag && p.ulist && tile_flags(p).f != nullptr && (long long)p.unx * p.uny * p.unz >= min_tiles;
}
#define LAUNCH_SPLIT(s, kern, level, grid, args)                                            \
  do {                                                                                      \
    if (split_ok(level)) {                                                                  \
      LAUNCH(s, kern<1>, grid, dim3(32, 4, 2), args);                                       \
      if ((level).ulist_n > 0) LAUNCH(s, kern<2>, dim3((unsigned)(level).ulist_n), dim3(32, 4, 2), args); \
    } else {                                                                                \
      LAUNCH(s, kern<0>, grid, dim3(32, 4, 2), args);                                       \
    }                                                                                       \
  } while (0)

// K-GS2 (wl_kernels_gs2.cuh): the four colour sweeps as two passes.  Single-GPU levels that the 4-cells-per-thread
// kernels serve (row length a multiple of 4) with an even last dimension (half_rangek then sweeps every plane).
// Opt-in (WL_FLAG_FUSED_SMOOTH): bit-identical to the per-sweep kernels and at the traffic goal (27 + 32 B/cell instead of
// 4 x 24), but issue-bound and slower for now (profiles/r2_smoother_experiments.md).
bool pair_ok(wl_sim* s, const Level& p) {
  return (s->cfg.flags & WL_FLAG_FUSED_SMOOTH) && s->D == 3 && !p.dist && p.tmp.alloc && use_vec4(p.lay) && p.lay.n2 % 2 == 0 && p.lay.n2 >= 4;
}
void gs_pairs(wl_sim* s, Level& p) {
  const Lay& l = p.lay;
  Gs2Args g{l, p.L.base, p.iD.base, p.r.base, p.eps.base, p.eps.base, p.per, l.n2 - 2};
  const long long bxy = (long long)((l.n0 - 2 + G2_XS - 1) / G2_XS) * ((l.n1 - 2 + G2_TY - 1) / G2_TY);
  const int nz = l.n2 - 2;
  {                                                   // chunk length: whole waves of 2 blocks per SM, 3 extra planes per chunk
    long long best = -1;
    for (int n = 1; n <= (nz + 3) / 4; ++n) {
      const int kz = (nz + n - 1) / n;
      const long long chunks = (nz + kz - 1) / kz;
      const long long cost = ((bxy * chunks + 295) / 296) * (kz + 3);
      if (best < 0 || cost < best) { best = cost; g.KZ = kz; }
    }
  }
  const dim3 grid((l.n0 - 2 + G2_XS - 1) / G2_XS, (l.n1 - 2 + G2_TY - 1) / G2_TY, (nz + g.KZ - 1) / g.KZ), blk2(32, G2_NR);
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(k_gs_pair<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Gs2Ring));
    cudaFuncSetAttribute(k_gs_pair<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Gs2Ring));
    attr = true;
  }
  {
    cudaEvent_t pa_, pb_;
    bool prof_ = prof_pre(s, "k_gs_pair<true>", &pa_, &pb_);
    launch_pdl_smem(s->st, grid, blk2, sizeof(Gs2Ring), k_gs_pair<true>, g);          // ϵ = r·iD and sweeps 1, 2
    if (prof_) prof_post(s, "k_gs_pair<true>", pa_, pb_, grid);
    g.ei = p.eps.base; g.eo = p.tmp.base;
    prof_ = prof_pre(s, "k_gs_pair<false>", &pa_, &pb_);
    launch_pdl_smem(s->st, grid, blk2, sizeof(Gs2Ring), k_gs_pair<false>, g);         // sweeps 3, 4 into the second buffer
    if (prof_) prof_post(s, "k_gs_pair<false>", pa_, pb_, grid);
    s->launches += 2;
  }
  std::swap(p.eps, p.tmp);
}
int smooth2(wl_sim* s, Level& p, float w, Red* norms, bool eps_ready = false) {
  SmArgs a{levdev(p), p.per, w, 0, s->redw, norms ? norms : s->d_red, norms ? 1 : 0};
  if (pair_ok(s, p)) {
    gs_pairs(s, p);
    a.p = levdev(p);                                  // ϵ now lives in the other buffer
    const dim3 gq = grid_quads(p.lay);
    LAUNCH(s, k_incr4, gq, dim3(32, 4, 2), a);
    if (norms) fold_partials_to(s, gq, norms);
    return WL_OK;
  }
  if (!eps_ready) DISPATCH_D(s, k_eps_init_g, grid_inside(p.lay), blk(s->D), a);   // else K-UP left ϵ = r·iD behind
  if (p.dist) { if (int e = halo(s, p.eps, 1, true)) return e; }
  LevArgs g = lev_args(s, p);
  dim3 gg = grid_inside(p.lay);
  gg.x = (p.lay.n0 - 2 + 2 * blk(s->D).x - 1) / (2 * blk(s->D).x);   // colour-compacted: a thread per x-pair
  const bool v4 = use_vec4(p.lay);
  // slab interfaces: exchange fused into the sweeps (k_gs_half_cx) when the planes can be pushed to peers
  const Dist& d = s->dist;
  const bool link_lo = d.lo >= 0 && !d.lo_wrap, link_hi = d.hi >= 0 && !d.hi_wrap;
  const bool fuse = p.dist && s->p2p && p.eps.in_arena && s->D == 3 && v4 && blk(3).z == 2 &&
                    (link_lo || link_hi) && !getenv_flag("WL_NO_GS_FUSE");
  GsLink x{};
  if (fuse) {
    auto peer = [&](char* pb, float* mine) { return reinterpret_cast<float*>(pb + (reinterpret_cast<char*>(mine) - s->arena)); };
    if (link_lo) x.peer_lo = peer(s->peer_lo, p.eps.base);
    if (link_hi) x.peer_hi = peer(s->peer_hi, p.eps.base);
    x.my = reinterpret_cast<unsigned int*>(s->arena);
    x.nb_lo = reinterpret_cast<unsigned int*>(s->peer_lo);
    x.nb_hi = reinterpret_cast<unsigned int*>(s->peer_hi);
  }
  {
    // prefetch distance: ~1.5 waves of blocks (8 resident per SM) expressed in planes of 2-plane block layers
    static const int pf_env = getenv("WL_PREFETCH") ? atoi(getenv("WL_PREFETCH")) : -1;
    const long long per_layer = (long long)gg.x * gg.y;
    g.pf = pf_env >= 0 ? pf_env : (per_layer > 0 ? (int)std::min<long long>(16, 2 * ((148 * 8 * 3 / 2 + per_layer - 1) / per_layer)) : 0);
    if (per_layer * gg.z < 148 * 8 * 2) g.pf = 0;            // small levels live in L2 anyway
  }
  for (int k0 = 1; k0 <= 4; ++k0) {
    g.k0 = k0;
    if (fuse) {
      x.wait_lo = link_lo && k0 > 1; x.wait_hi = link_hi && k0 > 1;
      x.wait_seq_lo = s->gs_seq_lo; x.wait_seq_hi = s->gs_seq_hi;
      x.push_lo = link_lo; x.push_hi = link_hi;
      if (link_lo) x.push_seq_lo = ++s->gs_seq_lo;
      if (link_hi) x.push_seq_hi = ++s->gs_seq_hi;
      static const bool cx2_off = getenv("WL_NO_GS_C2") != nullptr;
      static const long long cx2_min = getenv("WL_GS_C2_MIN_BLOCKS") ? atoll(getenv("WL_GS_C2_MIN_BLOCKS")) : 2048;
      if (!cx2_off && (long long)gg.x * gg.y * gg.z >= cx2_min) LAUNCH(s, (k_gs_half_cx<3, true>), gg, dim3(32, 2, 2), g, x);
      else LAUNCH(s, k_gs_half_cx<3>, gg, blk(3), g, x);
      if (k0 == 4 && (d.lo_wrap || d.hi_wrap)) { if (int e = halo(s, p.eps, 1, true, false, true)) return e; }
      continue;
    }
    // two rows per thread (k_gs_half_c2): same tiles, half the y threads, four planes per block; levels of a few dozen
    // blocks keep the one-cell kernel (more, smaller blocks fill the GPU better there).  WL_NO_GS_C2 / WL_GS_C2_MIN_BLOCKS: A/B
    static const bool c2_off = getenv("WL_NO_GS_C2") != nullptr;
    static const long long c2_min = getenv("WL_GS_C2_MIN_BLOCKS") ? atoll(getenv("WL_GS_C2_MIN_BLOCKS")) : 2048;
    if (!c2_off && s->D == 3 && !p.dist && blk(3).y == 4 && blk(3).z == 2 && (long long)gg.x * gg.y * gg.z >= c2_min) {
      LAUNCH(s, k_gs_half_c2, dim3(gg.x, gg.y, (gg.z + 1) / 2), dim3(32, 2, 4), g);
      continue;
    }
    DISPATCH_D(s, k_gs_half_c, gg, blk(s->D), g);
    if (p.dist) { if (int e = halo(s, p.eps, 1, k0 == 4, true)) return e; }
  }
  if (fuse) {
    a.wflag = x.my; a.wait_lo = link_lo; a.wait_hi = link_hi;
    a.wait_seq_lo = s->gs_seq_lo; a.wait_seq_hi = s->gs_seq_hi;
  }
  if (v4) {
    const dim3 gq = grid_quads(p.lay);
    LAUNCH(s, k_incr4, gq, dim3(32, 4, 2), a);
    if (norms) fold_partials_to(s, gq, norms);
  } else {
    DISPATCH_D(s, k_increment_w, grid_tiles(p.lay, false), blk(s->D), a);
  }
  return WL_OK;
}
// levels first..end in ONE single-block launch: Vcycle!(l=first) (if any coarser level) + smooth!(first)
void coarse_block(wl_sim* s, size_t first, float w) {
  CoarseArgs a{};
  for (size_t l = 0; l < s->lv.size() && l < (size_t)WL_MAXLEV; ++l) a.lv[l] = levdev(s->lv[l]);
  a.first = (int)first; a.nlev = (int)s->lv.size(); a.per = s->per_glob; a.omega = w;
  // (staging the smallest levels in shared memory measured slower than leaving them in L2 -- 52 -> 75 µs on the 18³..4³
  //  tail of a 128³ grid: the recursion is bound by ~40 dependent phases and one SM's instruction issue; removed in round 2)
  DISPATCH_D(s, k_coarse, dim3(1), dim3(1024), a);
}
// first: the top level's residual still needs its mean removed; K-DOWN does it on the fly and leaves the
// norms of the shifted residual in slot 10
int vcycle(wl_sim* s, size_t l, float w, bool first = false) {
  if (s->cfg.flags & WL_FLAG_NO_FUSED) { vcycle_v1(s, l, w); return WL_OK; }
  Level &fine = s->lv[l], &coarse = s->lv[l + 1];
  const bool gather = fine.dist && !coarse.dist;
  {
    if (fine.dist) { if (int e = halo(s, fine.r, 1)) return e; }       // K-DOWN reads ϵ = r·iD at the neighbours
    DownArgs a{levdev(fine), levdev_window(s, fine, coarse), fine.per, first ? 1 : 0, red_in(s, 0), fine.ncells, s->redw, s->d_red + 10};
    {
      const dim3 g3 = grid_inside(fine.lay);
      const long long nt = (long long)g3.x * g3.y * g3.z;
      if (use_vec4(fine.lay) && coarse.c[0] && coarse.c[1] && coarse.c[2]) {
        const dim3 gq = grid_quads(fine.lay);
        LAUNCH_SPLIT(s, k_down4, fine, gq, a);
        if (first) fold_partials_to(s, gq, s->d_red + 10);
      } else {
        DISPATCH_D(s, k_down, dim3((unsigned)nt), blk(s->D), a);
        if (first) fold_partials_to(s, dim3((unsigned)nt), s->d_red + 10);
      }
    }
    std::swap(fine.r, fine.eps);                     // K-DOWN staged the new residual in ϵ
    if (gather) {                                    // assemble the replicated coarse residual
      const int nloc = a.c.l.n2 - 2;
      const size_t cnt = (size_t)coarse.lay.plane * nloc;
      float* b = coarse.r.base + coarse.lay.plane;
      WL_NCCL(s, ncclAllGather(b + cnt * s->dist.rank, b, cnt, ncclFloat, s->dist.comm, s->st));
      WL_CUDA(s, cudaMemsetAsync(coarse.x.alloc, 0, coarse.x.bytes, s->st));   // fill!(coarse.x,0) on every rank's copy
    }
  }
  if (level_is_small(coarse)) {
    coarse_block(s, l + 1, w);
  } else {
    const bool deeper = l + 2 < s->lv.size();
    if (deeper) { if (int e = vcycle(s, l + 1, w)) return e; }
    if (int e = smooth2(s, coarse, w, nullptr, deeper)) return e;
  }
  if (coarse.dist) { if (int e = halo(s, coarse.x, 1)) return e; }     // K-UP prolongates the neighbours' coarse cells too
  else if (gather && s->per_glob.p[2]) {                               // rank windows read the wrap planes as ghosts
    PerArgs pa{coarse.lay, coarse.x.base, 2};
    dim3 g, b;
    plane_launch(coarse.lay, g, b);
    DISPATCH_D(s, k_perbc, g, b, pa);
  }
  UpArgs u{levdev(fine), levdev_window(s, fine, coarse), fine.per, w, s->redw, s->d_red, 0, pair_ok(s, fine) ? 0 : 1};
  if (use_vec4(fine.lay)) LAUNCH(s, k_up4, grid_quads(fine.lay), dim3(32, 4, 2), u);
  else DISPATCH_D(s, k_up, grid_inside(fine.lay), blk(s->D), u);
  return WL_OK;
}

// reduction slots -> pinned host memory (all ranks' slots folded in rank order)
int fetch_red(wl_sim* s) {
  if (s->dist.n == 1) {
    WL_CUDA(s, cudaMemcpyAsync(s->h_red, s->d_red, sizeof(Red) * NRED, cudaMemcpyDeviceToHost, s->st));
    WL_CUDA(s, cudaStreamSynchronize(s->st));
    return WL_OK;
  }
  if (int e = gather_red(s)) return e;
  WL_CUDA(s, cudaMemcpyAsync(s->h_redg, gather_table(s), sizeof(Red) * NRED * s->dist.n, cudaMemcpyDeviceToHost, s->st));
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  for (int q = 0; q < NRED; ++q) {
    double sum = 0.0; float m = -INFINITY;
    for (int r = 0; r < s->dist.n; ++r) { sum += s->h_redg[r * NRED + q].sum; m = std::fmax(m, s->h_redg[r * NRED + q].maxv); }
    s->h_red[q].sum = sum; s->h_red[q].maxv = m;
  }
  return WL_OK;
}

struct Trace { int n = 0; int flags = 0; float rinf[34], r1[34], omega[34]; };

// solver!(ml;tol,itmx) (MultiLevelPoisson.jl:108-128).  The loop exit and ω are steered by two
// scalars per iteration, exactly as in the reference.
int mlp_solver(wl_sim* s, Trace* tr, bool residual_done = false) {
  Level& p = s->lv[0];
  bool lazy = false;
  const double tol = s->cfg.tol;
  const double r1tol = (tol / 10) * (double)p.ncells, rinftol = tol;
  const int itmx = s->cfg.itmx;
  const bool fused = !(s->cfg.flags & WL_FLAG_NO_FUSED);
  {
    LevArgs a = lev_args(s, p);
    if (!residual_done) {                       // K-DIVRES already produced r and Σr in slot 0
      perbc(s, p.lay, p.x.base, &p.per);
      if (p.dist) { if (int e = halo(s, p.x, 1)) return e; }
      a.out = s->d_red + 0;
      DISPATCH_D(s, k_residual, grid_inside(p.lay), blk(s->D), a);
    }
    if (int e = gather_red(s)) return e;
    a.in = red_in(s, 0); a.out = s->d_red + 1;
    // fused solve on a multi-kernel hierarchy: the first K-DOWN removes the mean and takes the norms
    lazy = fused && residual_done && !level_is_small(p);
    if (!lazy) DISPATCH_D(s, k_shift_norms, grid_tiles(p.lay, false), blk(s->D), a);
  }
  float r1 = 0.f, rinf = 0.f, w = 1.f;
  int np = 0, flags = 0;
  if (!lazy) {
    if (int e = fetch_red(s)) return e;
    r1 = (float)s->h_red[1].sum; rinf = s->h_red[1].maxv;
    if (tr) { tr->rinf[0] = rinf; tr->r1[0] = r1; tr->omega[0] = w; }
  }
  while (np < itmx) {
    if (fused && level_is_small(p)) {            // the whole hierarchy fits the single-block kernel
      coarse_block(s, 0, w);
      LevArgs a = lev_args(s, p);
      a.in = red_in(s, 7); a.out = s->d_red + 1; // slot 7 stays zero: no shift, norms only
      DISPATCH_D(s, k_shift_norms, grid_tiles(p.lay, false), blk(s->D), a);
    } else {
      if (int e = vcycle(s, 0, w, lazy && np == 0)) return e;
      if (fused) { if (int e = smooth2(s, p, w, s->d_red + 1, true)) return e; }
      else smooth(s, p, w, s->d_red + 1);
    }
    if (int e = fetch_red(s)) return e;
    if (lazy && np == 0) {                        // the initial norms arrive with the first iteration's
      r1 = (float)s->h_red[10].sum; rinf = s->h_red[10].maxv;
      if (tr) { tr->rinf[0] = rinf; tr->r1[0] = r1; tr->omega[0] = w; }
    }
    const float rnew = (float)s->h_red[1].sum;
    rinf = s->h_red[1].maxv; ++np;
    if (tr && np < 34) { tr->rinf[np] = rinf; tr->r1[np] = rnew; tr->omega[np] = w; }
    if (rnew >= r1) w = (float)std::fmax(0.2, 0.9 * (double)w);
    else if (rnew < r1) w = (float)std::fmin(1.0, 1.02 * (double)w);
    r1 = rnew;
    if (!std::isfinite(rnew)) flags |= 1;
    if ((double)r1 < r1tol && (double)rinf < rinftol) break;
  }
  if (np >= itmx) flags |= 2;
  if (!residual_done) {                             // fused projection wrap-reads x and stores p's images itself
    perbc(s, p.lay, p.x.base, &p.per);
    if (p.dist) { if (int e = halo(s, p.x, 1)) return e; }
  }
  if (tr) { tr->n = np; tr->flags = flags; }
  return WL_OK;
}

// pcg!(p;it=6) (Poisson.jl:166-186): host control flow of the reference, one read-back per dot product.  On slabs the
// dot products are the reference's `perdot` (Poisson.jl:150-157): every rank's partial sum is gathered and folded in
// rank order by fetch_red, so all ranks take the same branches
int pcg(wl_sim* s, Level& p, int it = 6) {
  const float epsT = 1.1920929e-07f;
  PcgArgs a{p.lay, p.L.base, p.iD.base, p.x.base, p.r.base, p.eps.base, p.z.base, 0.f, 0, s->redw, s->d_red + 11};
  const dim3 g = grid_inside(p.lay), b = blk(s->D);
  DISPATCH_D(s, k_pcg_init, g, b, a);
  if (int e = fetch_red(s)) return e;
  float rho = (float)s->h_red[11].sum;
  if (std::fabs(rho) < 10 * epsT) return WL_OK;
  for (int i = 1; i <= it; ++i) {
    perbc(s, p.lay, p.eps.base, &p.per);
    if (p.dist) { if (int e = halo(s, p.eps, 1)) return e; }   // the search direction's interface planes
    DISPATCH_D(s, k_pcg_mult, g, b, a);
    if (int e = fetch_red(s)) return e;
    const float alpha = rho / (float)s->h_red[11].sum;
    if (std::fabs(alpha) < 1e-2 || std::fabs(alpha) > 1e2) return WL_OK;   // also leaves on NaN-free garbage only
    a.coef = alpha; a.more = i == it ? 0 : 1;
    DISPATCH_D(s, k_pcg_update, g, b, a);
    if (i == it) return WL_OK;
    if (int e = fetch_red(s)) return e;
    const float rho2 = (float)s->h_red[11].sum;
    if (std::fabs(rho2) < 10 * epsT) return WL_OK;
    a.coef = rho2 / rho;
    DISPATCH_D(s, k_pcg_dir, g, b, a);
    rho = rho2;
  }
  return WL_OK;
}

// stand-alone solver!(p::Poisson;tol,itmx) (Poisson.jl:212-223) on level 1: residual!, then pcg! until
// Σ|r| < (tol/10)·N and max|r| < tol
int poisson_solver(wl_sim* s, int itmx, int* n_out) {
  Level& p = s->lv[0];
  const double tol = s->cfg.tol;
  const double r1tol = (tol / 10) * (double)p.ncells, rinftol = tol;
  LevArgs a = lev_args(s, p);
  perbc(s, p.lay, p.x.base, &p.per);
  if (p.dist) { if (int e = halo(s, p.x, 1)) return e; }
  a.out = s->d_red + 0;
  DISPATCH_D(s, k_residual, grid_inside(p.lay), blk(s->D), a);
  if (int e = gather_red(s)) return e;
  a.in = red_in(s, 0); a.out = s->d_red + 1;
  DISPATCH_D(s, k_shift_norms, grid_tiles(p.lay, false), blk(s->D), a);
  int np = 0;
  while (np < itmx) {
    if (int e = pcg(s, p)) return e;
    a.in = red_in(s, 7); a.out = s->d_red + 1;                      // slot 7 stays zero: norms only
    DISPATCH_D(s, k_shift_norms, grid_tiles(p.lay, false), blk(s->D), a);
    if (int e = fetch_red(s)) return e;
    ++np;
    const float r1 = (float)s->h_red[1].sum, rinf = s->h_red[1].maxv;
    if (!std::isfinite(r1)) break;
    if ((double)r1 < r1tol && (double)rinf < rinftol) break;
  }
  perbc(s, p.lay, p.x.base, &p.per);
  if (p.dist) { if (int e = halo(s, p.x, 1)) return e; }
  if (n_out) *n_out = np;
  return WL_OK;
}

// ---------------------------------------------------------------- momentum pieces
void mom_rhs(wl_sim* s, DevField& uadv, const float* acc, const float* accf = nullptr) {
  MomArgs a{};
  a.l = s->lay; a.ua = uadv.base; a.u0 = s->u0.base; a.V = s->V.base; a.f = s->f.base;
  a.per = s->per; a.nu = s->cfg.nu; a.dt = s->dt;
  for (int d = 0; d < 3; ++d) a.acc.v[d] = acc ? acc[d] : 0.f;
  if (accf) DISPATCH_DS(s, k_mom_rhs_acc, grid_all(s->lay), blk(s->D), a, accf);
  else DISPATCH_DS(s, k_mom_rhs, grid_all(s->lay), blk(s->D), a);
}
// list of band cells for k_mom_patch; rebuilt when μ₁ changed (measure!, uploads)
int band_rebuild(wl_sim* s) {
  bool have = false;
  if (s->band_from_measure) {                       // list written by k_measure: valid unless it overflowed
    int n = 0;
    WL_CUDA(s, cudaMemcpyAsync(&n, s->band_count, sizeof(int), cudaMemcpyDeviceToHost, s->st));
    WL_CUDA(s, cudaStreamSynchronize(s->st));
    have = n <= s->band_cap;
    if (!have) {
      if (s->band_list) cudaFree(s->band_list);
      s->band_cap = n + n / 4 + 1024;
      WL_CUDA(s, cudaMalloc(&s->band_list, sizeof(int) * (size_t)s->band_cap));
    }
    s->band_from_measure = false;
  }
  for (int pass = 0; pass < 2 && !have; ++pass) {
    WL_CUDA(s, cudaMemsetAsync(s->band_count, 0, sizeof(int), s->st));
    BandArgs a{s->lay, s->mu1.base, s->band_list, s->band_count, s->band_cap};
    DISPATCH_D(s, k_band_scan, grid_inside(s->lay), blk(s->D), a);
    int n = 0;
    WL_CUDA(s, cudaMemcpyAsync(&n, s->band_count, sizeof(int), cudaMemcpyDeviceToHost, s->st));
    WL_CUDA(s, cudaStreamSynchronize(s->st));
    if (n <= s->band_cap) break;
    if (s->band_list) cudaFree(s->band_list);
    s->band_cap = n + n / 4 + 1024;
    WL_CUDA(s, cudaMalloc(&s->band_list, sizeof(int) * (size_t)s->band_cap));
  }
  // the cells whose f the two-pass patch needs: the band and its face neighbours, each once
  if (s->band_cap > 0) {
    if (!s->band_mark) {
      WL_CUDA(s, cudaMalloc(&s->band_mark, sizeof(int) * (size_t)s->lay.cstride));
      WL_CUDA(s, cudaMemsetAsync(s->band_mark, 0, sizeof(int) * (size_t)s->lay.cstride, s->st));
      WL_CUDA(s, cudaMalloc(&s->band_count2, sizeof(int)));
    }
    if (s->band_cap2 < 7 * s->band_cap) {
      if (s->band_list2) cudaFree(s->band_list2);
      s->band_cap2 = 7 * s->band_cap;
      WL_CUDA(s, cudaMalloc(&s->band_list2, sizeof(int) * (size_t)s->band_cap2));
    }
    WL_CUDA(s, cudaMemsetAsync(s->band_count2, 0, sizeof(int), s->st));
    Band2Args b{s->lay, s->band_list, s->band_count, s->band_cap, s->band_mark, ++s->band_epoch,
                s->band_list2, s->band_count2, s->band_cap2};
    LAUNCH(s, k_band_nbrs, dim3(592), dim3(256), b);
  }
  s->band_dirty = false;
  return WL_OK;
}
// 4-D tensor map (x+3, y, z+zpad, component) over the padded device array of a 3-component field, for the TMA tile
// loads of K-MOM: base = alloc + 28 floats (16-byte aligned; element i sits at x = i + 3), rows of `pitch` floats.
// The encoder comes from the driver through the runtime (no libcuda link); nullptr when it is not available.
const CUtensorMap* tensor_map(wl_sim* s, const DevField& f, int box_x, int box_y) {
  typedef CUresult (*Encode)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static Encode enc = nullptr;
  if (s->tma_ok < 0) {
    s->tma_ok = 0;
    if (!getenv("WL_NO_TMA")) {
      void* fn = nullptr;
      cudaDriverEntryPointQueryResult qr;
      if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) == cudaSuccess && fn &&
          qr == cudaDriverEntryPointSuccess) { enc = (Encode)fn; s->tma_ok = 1; }
      else cudaGetLastError();
    }
  }
  if (s->tma_ok != 1) return nullptr;
  const float* key = f.alloc + (box_x == 32 ? 1 : 0);     // two box shapes per allocation (tile / own-cell operands)
  auto it = s->tmaps.find(key);
  if (it != s->tmaps.end()) return &it->second;
  const Lay& l = f.lay;
  CUtensorMap tm;
  const cuuint64_t dims[4] = {(cuuint64_t)l.pitch, (cuuint64_t)l.n1, (cuuint64_t)(l.n2 + 2 * l.zpad), 3};
  const cuuint64_t strides[3] = {(cuuint64_t)l.pitch * 4, (cuuint64_t)l.plane * 4, (cuuint64_t)l.cstride * 4};
  const cuuint32_t box[4] = {(cuuint32_t)box_x, (cuuint32_t)box_y, 1, 3}, es[4] = {1, 1, 1, 1};
  const CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (void*)(f.alloc + LPAD - 3), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { s->tma_ok = 0; return nullptr; }
  return &(s->tmaps[key] = tm);
}
// K-MOM (3D fused): out = new velocity from ua (advecting) and u0
void mom3(wl_sim* s, DevField& ua, DevField& u0, DevField& out, int corrector, const float* acc, const float* accf = nullptr) {
  Mom3Args a{};
  {
    const CUtensorMap* t0 = tensor_map(s, ua, M3_TW, M3_TH);
    const CUtensorMap* t1 = tensor_map(s, u0, 32, M3_TY - 1);
    const CUtensorMap* t2 = tensor_map(s, s->V, 32, M3_TY - 1);
    const CUtensorMap* t3 = tensor_map(s, s->mu0, 32, M3_TY - 1);
    if (t0 && t1 && t2 && t3) { a.use_tma = 1; a.tm_ua = *t0; a.tm_u0 = *t1; a.tm_V = *t2; a.tm_m0 = *t3; }
  }
  a.l = s->lay; a.ua = ua.base; a.u0 = u0.base; a.V = s->V.base; a.mu0 = s->mu0.base; a.un = out.base;
  a.per = s->per; a.nu = s->cfg.nu; a.dt = s->dt; a.corrector = corrector;
  for (int d = 0; d < 3; ++d) a.acc.v[d] = acc ? acc[d] : 0.f;
  const Lay& l = s->lay;
  for (int c = 0; c < 3; ++c) {
    const long long oc = l.cstride * c;
    a.uac[c] = a.ua + oc; a.u0c[c] = a.u0 + oc; a.Vc[c] = a.V + oc; a.m0c[c] = a.mu0 + oc; a.unc[c] = a.un + oc;
    a.accfc[c] = accf ? accf + oc : nullptr;
  }
  // 32-column tiles (the block's last row supplies lane 31's upper x faces); WL_M3_W32=0 selects the 31-column
  // tiling.  Measured (profiles/r1_v14_kmom_march_sweep.tsv): TGV 128³ 67.3 -> 58.7 us, sphere 512x256² 810 -> 795 us.
  const int bw = 32;
  // march length: the chunk count that minimises waves x (planes per chunk + ~4 planes of pipeline fill and
  // the extra flux plane), a wave being 2 blocks on each of the 148 SMs (measured on 128³ and 512x256²:
  // a grid just under a whole number of waves beats a finer split by up to 8%)
  {
    const long long bxy = (long long)((l.n0 - 2 + bw - 1) / bw) * ((l.n1 - 2 + M3_TY - 2) / (M3_TY - 1));
    const int nz = l.n2 - 2;
    long long best = -1;
    a.KZ = nz;
    // chunks longer than 64 planes lose more to the ragged finish of the last blocks than they save in
    // pipeline fill (sphere 512x256², 32-column tiles, planes per chunk -> us: 256 848, 128 821, 64 795, 43 786)
    for (int n = (nz + 63) / 64; n <= (nz + 3) / 4; ++n) {
      const int kz = (nz + n - 1) / n;
      const long long chunks = (nz + kz - 1) / kz;
      const long long cost = ((bxy * chunks + 295) / 296) * (kz + 4);
      if (best < 0 || cost < best) { best = cost; a.KZ = kz; }
    }
  }
  dim3 g((l.n0 - 2 + bw - 1) / bw, (l.n1 - 2 + M3_TY - 2) / (M3_TY - 1), (l.n2 - 2 + a.KZ - 1) / a.KZ);
  dim3 b(32, M3_TY, 1);
  const int sc = s->cfg.scheme;
  // body-free, fully periodic flow: V ≡ 0 and μ₀ ≡ 1 everywhere, the kernel need not read them
  static const bool nb_off = getenv("WL_NO_NOBODY") != nullptr;
  const bool nb = !nb_off && s->free_flow && s->per_glob.p[0] && s->per_glob.p[1] && s->per_glob.p[2];
  static bool attr = false;
  if (!attr) {                                                       // Mom3Smem exceeds the 48 KB static limit
#define WL_MOM3_ATTR(S_, NB, F) cudaFuncSetAttribute(k_mom3<S_, NB, true, F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Mom3Smem))
    WL_MOM3_ATTR(0, false, false); WL_MOM3_ATTR(0, true, false); WL_MOM3_ATTR(1, false, false); WL_MOM3_ATTR(1, true, false);
    WL_MOM3_ATTR(2, false, false); WL_MOM3_ATTR(2, true, false);
    WL_MOM3_ATTR(0, false, true); WL_MOM3_ATTR(0, true, true); WL_MOM3_ATTR(1, false, true); WL_MOM3_ATTR(1, true, true);
    WL_MOM3_ATTR(2, false, true); WL_MOM3_ATTR(2, true, true);
#undef WL_MOM3_ATTR
    attr = true;
  }
  // NB: body-free variant; F: with the caller's acceleration field (a separate instantiation: the hot variant pays nothing)
#define WL_MOM3_LAUNCH(NB, F)                                                                 \
  do {                                                                                       \
    if (sc == 0) LAUNCH_SMEM(s, (k_mom3<0, NB, true, F>), g, b, sizeof(Mom3Smem), a);        \
    else if (sc == 1) LAUNCH_SMEM(s, (k_mom3<1, NB, true, F>), g, b, sizeof(Mom3Smem), a);   \
    else LAUNCH_SMEM(s, (k_mom3<2, NB, true, F>), g, b, sizeof(Mom3Smem), a);                \
  } while (0)
  if (accf) { if (nb) WL_MOM3_LAUNCH(true, true); else WL_MOM3_LAUNCH(false, true); }
  else { if (nb) WL_MOM3_LAUNCH(true, false); else WL_MOM3_LAUNCH(false, false); }
#undef WL_MOM3_LAUNCH
  PatchArgs pa{};
  pa.m.l = s->lay; pa.m.ua = ua.base; pa.m.u0 = u0.base; pa.m.V = s->V.base; pa.m.f = nullptr;
  pa.m.per = s->per; pa.m.nu = s->cfg.nu; pa.m.dt = s->dt; pa.m.acc = a.acc;
  pa.mu0 = s->mu0.base; pa.mu1 = s->mu1.base; pa.un = out.base; pa.accf = accf;
  pa.list = s->band_list; pa.count = s->band_count; pa.cap = s->band_cap; pa.corrector = corrector;
  if (s->band_cap > 0 && s->band_list2) {
    pa.m.f = s->f.base;
    if (sc == 0) LAUNCH(s, k_patch_f<0>, dim3(1184), dim3(128), pa, s->band_list2, s->band_count2, s->band_cap2);
    else if (sc == 1) LAUNCH(s, k_patch_f<1>, dim3(1184), dim3(128), pa, s->band_list2, s->band_count2, s->band_cap2);
    else LAUNCH(s, k_patch_f<2>, dim3(1184), dim3(128), pa, s->band_list2, s->band_count2, s->band_cap2);
    LAUNCH(s, k_patch_apply, dim3(1184), dim3(128), pa);
  }
  if (s->cfg.exitBC) {
    ExitCopyArgs e{s->lay, ua.base, out.base};
    LAUNCH(s, k_copy_exit_plane, dim3((l.n1 + 127) / 128, l.n2), dim3(128), e);
  }
}
void bdim(wl_sim* s, int corrector) {
  BdimArgs a{s->lay, s->u.base, s->f.base, s->V.base, s->mu0.base, s->mu1.base, corrector};
  DISPATCH_D(s, k_bdim, grid_inside(s->lay), blk(s->D), a);
}
// mom_project!(a,b,w,t) (Flow.jl:223-232)
int mom_project(wl_sim* s, float w, const float* U, Trace* tr) {
  if (!(s->cfg.flags & WL_FLAG_NO_FUSED)) {
    Level& p0 = s->lv[0];
    const float dt = w * s->dt;
    DivResArgs d{s->lay, s->u.base, s->p.base, p0.eps.base, s->mu0.base, p0.r.base, s->per, dt, s->redw, s->d_red + 0, tile_flags(p0), prefetch_planes(s->lay)};
    if (use_vec4(s->lay)) {
      const dim3 gq = grid_quads(s->lay);
      LAUNCH_SPLIT(s, k_divres4, p0, gq, d);
      fold_partials_to(s, gq, s->d_red + 0);
    } else {
      DISPATCH_D(s, k_divres, grid_tiles(s->lay, false), blk(s->D), d);
    }
    std::swap(s->p, p0.eps); p0.x = s->p;                       // scaled x lives in the former ϵ buffer
    if (int e = mlp_solver(s, tr, true)) return e;
    if (int e = halo(s, p0.x, 1)) return e;                     // the solver's final perBC!(x) across slabs
    ProjArgs2 q{s->lay, s->u.base, s->p.base, p0.eps.base, s->mu0.base, s->per, dt, tile_flags(p0)};
    if (use_vec4(s->lay)) LAUNCH(s, k_proj4, grid_quads(s->lay), dim3(32, 4, 2), q);
    else DISPATCH_D(s, k_proj, grid_inside(s->lay), blk(s->D), q);
    std::swap(s->p, p0.eps); p0.x = s->p;                       // (K-PROJ also wrote p's interface ghost planes)
    bc_vec(s, s->u, U, s->cfg.exitBC != 0);
    return WL_OK;
  }
  ProjArgs a{s->lay, s->u.base, s->p.base, s->sigma.base, s->mu0.base, w * s->dt};
  DISPATCH_D(s, k_div_scale, grid_all(s->lay), blk(s->D), a);
  if (int e = mlp_solver(s, tr)) return e;
  DISPATCH_D(s, k_project, grid_inside(s->lay), blk(s->D), a);
  DISPATCH_D(s, k_unscale, grid_all(s->lay), blk(s->D), a);
  bc_vec(s, s->u, U, s->cfg.exitBC != 0);
  return WL_OK;
}
int cfl(wl_sim* s, float* out) {
  CflArgs a{s->lay, s->u.base, s->sigma.base, s->redw, s->d_red + 2, s->per.zlo, s->per.zhi};
  if (use_vec4(s->lay)) {                          // all rows and planes, ghosts included
    const dim3 g((s->lay.n0 - 2 + 127) / 128, (s->lay.n1 + 3) / 4, (s->lay.n2 + 1) / 2);
    LAUNCH(s, k_cfl4, g, dim3(32, 4, 2), a);
    fold_partials_to(s, g, s->d_red + 2);
  } else {
    DISPATCH_D(s, k_cfl, grid_tiles(s->lay, true), blk(s->D), a);
  }
  if (int e = fetch_red(s)) return e;
  const float m = s->h_red[2].maxv;
  *out = std::fmin(10.f, 1.f / (m + 5.f * s->cfg.nu));
  return WL_OK;
}
void ghost_phi(wl_sim* s) {
  GhostPhiArgs a{s->lay, s->u.base, s->sigma.base, s->per, s->cfg.nu};
  dim3 g, b;
  plane_launch(s->lay, g, b);
  g.z = s->D;
  DISPATCH_DS(s, k_ghost_phi, g, b, a);
}

BodyDev body_at(const wl_body* b, int D, float t) {
  BodyDev d{};
  d.kind = b->kind; d.axis = b->axis; d.radius = b->radius; d.minor = b->minor;
  for (int k = 0; k < 3; ++k) { d.c[k] = 0.f; d.Vb[k] = 0.f; }
  for (int k = 0; k < D; ++k) {
    d.c[k] = b->center[k] + b->vel[k] * t + b->acc[k] * t * t + b->amp[k] * std::sin(b->omega * t);
    d.Vb[k] = b->vel[k] + 2 * b->acc[k] * t + b->amp[k] * b->omega * std::cos(b->omega * t);
  }
  return d;
}

// wl_body_node[] -> BodyProg: rotation(θ) (RigidMap.jl:45-49) in Float32 on the host, stack discipline checked here
// so the device loop can trust it.  Returns nullptr or a message.
const char* prog_from(const wl_body_node* nd, int n, int D, BodyProg* out) {
  if (!nd || n < 1 || n > CSG_MAX) return "a body program needs 1..WL_CSG_MAX nodes";
  int sp = 0;
  out->n = n;
  for (int i = 0; i < n; ++i) {
    const wl_body_node& b = nd[i];
    NodeDev& q = out->nd[i];
    q = NodeDev{};
    q.op = b.op;
    if (b.op == 0) {
      if (b.kind < 0 || b.kind > 4) return "unknown leaf kind";
      if (b.kind >= 2 && (b.axis < 0 || b.axis >= D)) return "leaf axis out of range";
      ++sp;
      q.kind = b.kind; q.axis = b.axis; q.radius = b.radius; q.minor = b.minor;
      const float t1 = D == 3 ? b.theta[0] : 0.f, t2 = D == 3 ? b.theta[1] : 0.f, t3 = b.theta[2];
      const float c1 = std::cos(t1), s1 = std::sin(t1), c2 = std::cos(t2), s2 = std::sin(t2), c3 = std::cos(t3), s3 = std::sin(t3);
      float* R = q.R;
      if (D == 2) { R[0] = c3; R[1] = s3; R[3] = -s3; R[4] = c3; R[8] = 1.f; }
      else {
        R[0] = c3 * c2;  R[1] = c3 * s2 * s1 + s3 * c1;   R[2] = -c3 * s2 * c1 + s3 * s1;
        R[3] = -s3 * c2; R[4] = -s3 * s2 * s1 + c3 * c1;  R[5] = s3 * s2 * c1 + c3 * s1;
        R[6] = s2;       R[7] = -c2 * s1;                 R[8] = c2 * c1;
      }
      for (int k = 0; k < 3; ++k) { q.x0[k] = b.x0[k]; q.xp[k] = b.xp[k]; q.V[k] = b.V[k]; q.w[k] = b.omega[k]; }
    } else if (b.op == 3) {
      if (sp < 1) return "complement of an empty stack";
    } else if (b.op == 1 || b.op == 2) {
      if (sp < 2) return "set operation needs two operands";
      --sp;
    } else return "unknown node op";
  }
  return sp == 1 ? nullptr : "body program leaves more than one body on the stack";
}

DevField* flow_field(wl_sim* s, int field, int level, Lay* lay) {
  DevField* f = nullptr;
  if (field < 10) {
    switch (field) {
      case WL_U: f = &s->u; break;
      case WL_U0: f = &s->u0; break;
      case WL_F: f = &s->f; break;
      case WL_P: f = &s->p; break;
      case WL_SIGMA: f = &s->sigma; break;
      case WL_V: f = &s->V; break;
      case WL_MU0: f = &s->mu0; break;
      case WL_MU1: f = &s->mu1; break;
      default: return nullptr;
    }
  } else if (field == WL_ACCEL_T0 || field == WL_ACCEL_T1) {
    f = &s->accf[field - WL_ACCEL_T0];
    if (!f->alloc) return nullptr;
  } else if (field >= WL_MEAN_P && field <= WL_MEAN_UU) {
    f = field == WL_MEAN_P ? &s->mf_P : (field == WL_MEAN_U ? &s->mf_U : &s->mf_UU);
    if (!f->alloc) return nullptr;
  } else {
    if (level < 0 || level >= (int)s->lv.size()) return nullptr;
    Level& l = s->lv[level];
    switch (field) {
      case WL_X: f = &l.x; break;
      case WL_R: f = &l.r; break;
      case WL_EPS: f = &l.eps; break;
      case WL_Z: f = &l.z; break;
      case WL_L: f = &l.L; break;
      case WL_DIAG: f = &l.Dg; break;
      case WL_IDIAG: f = &l.iD; break;
      default: return nullptr;
    }
  }
  if (lay) *lay = f->lay;
  return f;
}

inline bool divisible(int N) { return N % 2 == 0 && N > 4; }

}  // namespace

#include "wl_abi.cuh"
#include "wl_noalloc.cuh"
