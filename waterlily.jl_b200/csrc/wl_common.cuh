// wl_common.cuh -- device layout and small helpers shared by the sm_100a kernels.
//
// Device layout of one scalar field of Julia size (n0,n1,n2) (ghosts included; n2==1 in 2D):
//   element (i,j,k) lives at  base[i + pitch*j + plane*k],  plane = pitch*n1.
// `pitch` is a multiple of 32 floats and `base` is chosen so that the FIRST INTERIOR element
// (i==1) of every row starts a 128-byte line: a warp that owns 32 consecutive interior cells
// stores exactly one full line.  Vector fields are `ncomp` such scalars `cstride` floats apart.
// The padding is invisible through the C ABI (upload/download repack).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace wl {

// Programmatic dependent launch (sm_90+): first statement of every kernel.  launch_dependents lets the
// NEXT kernel of the stream be scheduled while this one still runs (its blocks become resident and
// park in their own wait); wait returns once the PREVIOUS kernel has completed and its writes are
// visible.  Nothing a kernel does precedes its wait, so ordering is exactly that of a plain stream;
// what overlaps is launch latency and block scheduling (~1 µs per kernel boundary, ~60 boundaries per
// step).  Without the launch attribute (host side, LAUNCH in wl_b200.cu) both are no-ops.
__device__ __forceinline__ void pdl_entry() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

constexpr int LPAD = 31;   // floats in front of i==0 so that i==1 is 128-byte aligned

struct Lay {
  int n0, n1, n2;    // sizes including ghosts (n2 == 1 for D == 2)
  int D;
  int pitch, plane;  // floats
  long long cstride; // floats between components of a vector field
  int zpad;          // planes allocated beyond each z ghost plane (host bookkeeping)
  __host__ __device__ int off(int i, int j, int k) const { return i + pitch * j + plane * k; }
  __host__ __device__ int n(int d) const { return d == 0 ? n0 : d == 1 ? n1 : n2; }
  __host__ __device__ int stride(int d) const { return d == 0 ? 1 : d == 1 ? pitch : plane; }
  __host__ __device__ long long ninside() const {
    long long c = (long long)(n0 - 2) * (n1 - 2);
    return D == 3 ? c * (n2 - 2) : c;
  }
};

// zpad: extra planes allocated below k==0 and above k==n2-1 (slab decomposition: the velocity
// needs 3 halo planes per side, the first of which is the ordinary ghost plane)
inline Lay make_lay(int D, const int* ng, int zpad = 0) {
  Lay l;
  l.D = D; l.zpad = zpad;
  l.n0 = ng[0]; l.n1 = ng[1]; l.n2 = D == 3 ? ng[2] : 1;
  l.pitch = ((l.n0 + 31) / 32) * 32;
  if (l.pitch < l.n0 + 0) l.pitch += 32;
  l.plane = l.pitch * l.n1;
  long long vol = (long long)l.plane * (l.n2 + 2 * zpad);
  l.cstride = ((vol + 64 + 31) / 32) * 32;   // +LPAD front, slack behind for vector tails
  return l;
}

// p[d]: direction d is periodic AND handled inside the kernel (wrap-reads / image stores).
// zlo/zhi: the lower/upper end of the last dimension is an INTERFACE to a neighbouring slab
// (multi-GPU): its ghost planes hold the neighbour's cells (filled by halo exchange), so no
// boundary variant applies there and p[2] is 0.  kz0: global index of local plane k==0.
struct Per { int p[3]; int zlo, zhi; int kz0; };
struct Vec3 { float v[3]; };

// Pattern tiles (pressure side).  Away from bodies the face coefficients L of a level follow a pattern known without
// reading them: L_d = lu[d] (1 on the finest level, the restricted constant 2^k... on coarser ones), except 0 on the faces
// of a physical non-periodic boundary (BC!(L,0): indices 0, 1 and N-1 of the normal component, core.jl:208-212); then
// iD = 1/-(sum of the six faces).  The pressure kernels would otherwise re-read those values from HBM in every pass
// (12-20 of their 24-36 B/cell).  update!/build classify each 128x4x2-cell tile (= one block of the 4-cells-per-thread
// kernels): every cell of the tile AND of its neighbouring tiles matches the pattern -> class 1 (no boundary, no slab
// interface next to the tile: all values are the constants lu / idu) or class 2 (boundary or interface tile: L is
// synthesised with index compares, iD of neighbour cells is still loaded; rows that touch no y/z boundary -- all but
// the first and last of the array -- only need two selects on the x ends and the constant idx1); else class 0
// (general path, loads).
// Same values, same operations: bit-identical.  f == nullptr: no classes for this level (2D, rows not a multiple of 4,
// switched off by an upload until the next update!).
// Class 3 = a class-2 tile whose only boundary is an x end: every row of it (and of its y/z neighbours) is a plain row,
// so a kernel compiled for classes 1 and 3 alone needs no coefficient or iD arrays at all -- K-DIVRES and K-DOWN run as
// two launches, that lean one over the whole grid (class 0/2 blocks leave at once) and the general one over `list`.
struct TileFlags {
  const unsigned char* f; int nx, ny;
  const int* list;   // linear tile ids of the class-0 and class-2 tiles, ascending (the "list" launch of the split kernels), or null
  float lu[3];
  float idu;    // iD of a cell whose six faces hold the pattern constants
  float idx1;   // iD of a cell whose lower OR upper x face is a boundary face (0) and the other five hold the constants
};
constexpr int UT_X = 128, UT_Y = 4, UT_Z = 2;
__device__ __forceinline__ int tile_class(const TileFlags& t, int tx, int ty, int tz) {
  return t.f == nullptr ? 0 : t.f[tx + t.nx * (ty + t.ny * tz)];
}
// is index c of direction d a face the boundary pattern zeroes (normal component only)?
__device__ __forceinline__ bool pattern_zero(const Per& per, int d, int c, int n) {
  if (per.p[d]) return false;
  const bool lo_phys = !(d == 2 && per.zlo), hi_phys = !(d == 2 && per.zhi);
  return (lo_phys && c <= 1) || (hi_phys && c == n - 1);
}

// ---- deterministic block reduction helpers (warp shuffle -> smem -> thread 0)
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_down_sync(0xffffffffu, v, o));
  return v;
}

// Result slots of the scalar reductions (device memory mirrored into pinned host memory).
constexpr int NRED = 12;     // reduction slots per sim: 0 Σr, 1 norms, 2 CFL, 3-5 force, 7 zero, 8-9 exitBC!, 10 first-cycle norms
struct Red {
  double sum;       // Σ r   (residual!)  or Σ|r| (L₁)
  float  maxv;      // max|r| (L∞) or max σ (CFL)
  float  pad;
};
// A reduction result as its consumers see it: one Red per rank, `stride` Reds apart, folded in rank
// order (multi-GPU: the slots of all ranks are all-gathered; single GPU: n == 1).
struct RedIn { const Red* p; int n, stride; };
__host__ __device__ inline double redin_sum(const RedIn& r) {
  double s = 0.0;
  for (int q = 0; q < r.n; ++q) s += r.p[(size_t)q * r.stride].sum;
  return s;
}

// Two-stage, order-fixed reduction: every block writes one partial; the last block to finish
// (ticket counter) folds the partials in index order, so the result does not depend on the
// order blocks ran in (iteration counts must not depend on scheduling).
struct RedWork {
  double* psum;          // [maxblocks]
  float* pmax;           // [maxblocks]
  unsigned int* ticket;  // zero before and after every use
};

__device__ __forceinline__ int linear_tid() {
  return threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
}
__device__ __forceinline__ int linear_bid() {
  return blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
}

// Persistent tile loop for kernels that end in a reduction: a 1-D grid of at most a few blocks per
// SM walks the interior tiles, so the fence + ticket round trip of block_reduce_finish is paid once
// per block instead of once per 256 cells.  f(i,j,k) is called for every interior cell of the tiles
// this thread owns (tile = blockDim cells, x fastest).
template <int D, class F>
__device__ __forceinline__ void for_tiles_inside(const Lay& l, F f) {
  const int nbx = (l.n0 - 2 + blockDim.x - 1) / blockDim.x, nby = (l.n1 - 2 + blockDim.y - 1) / blockDim.y;
  const int nbz = D == 3 ? (l.n2 - 2 + blockDim.z - 1) / blockDim.z : 1;
  const int nt = nbx * nby * nbz;
  for (int t = blockIdx.x; t < nt; t += gridDim.x) {
    const int tx = t % nbx, q = t / nbx, ty = q % nby, tz = q / nby;
    const int i = tx * blockDim.x + threadIdx.x + 1, j = ty * blockDim.y + threadIdx.y + 1;
    const int k = D == 3 ? tz * blockDim.z + threadIdx.z + 1 : 0;
    if (i <= l.n0 - 2 && j <= l.n1 - 2 && (D == 2 || k <= l.n2 - 2)) f(i, j, k);
  }
}
// same over ALL cells (ghosts included); i starts LPAD before the row so that interior cells stay aligned
template <int D, class F>
__device__ __forceinline__ void for_tiles_all(const Lay& l, F f) {
  const int nbx = (l.n0 + LPAD + blockDim.x - 1) / blockDim.x, nby = (l.n1 + blockDim.y - 1) / blockDim.y;
  const int nbz = D == 3 ? (l.n2 + blockDim.z - 1) / blockDim.z : 1;
  const int nt = nbx * nby * nbz;
  for (int t = blockIdx.x; t < nt; t += gridDim.x) {
    const int tx = t % nbx, q = t / nbx, ty = q % nby, tz = q / nby;
    const int i = tx * blockDim.x + threadIdx.x - LPAD, j = ty * blockDim.y + threadIdx.y;
    const int k = D == 3 ? tz * blockDim.z + threadIdx.z : 0;
    if (i >= 0 && i < l.n0 && j < l.n1 && k < l.n2) f(i, j, k);
  }
}

constexpr int RED_GROUP = 256;   // blocks per first-level ticket

// fold n partials starting at `base` in a fixed order (thread t takes t, t+nthr, ...; warp tree;
// then one warp over the warp results).  Result valid in thread 0.  Contains __syncthreads.
__device__ __forceinline__ void fold_partials(const double* psum, const float* pmax, int base, int n,
                                              double* ssum, float* smax, double& a, float& mm) {
  const int tid = linear_tid();
  const int nthr = blockDim.x * blockDim.y * blockDim.z;
  const int lane = tid & 31, wid = tid >> 5, nw = (nthr + 31) >> 5;
  a = 0.0; mm = -INFINITY;
  for (int i = tid; i < n; i += nthr) { a += __ldcg(&psum[base + i]); mm = fmaxf(mm, __ldcg(&pmax[base + i])); }
  a = warp_sum(a); mm = warp_max(mm);
  __syncthreads();
  if (lane == 0) { ssum[wid] = a; smax[wid] = mm; }
  __syncthreads();
  if (wid == 0) {
    a = lane < nw ? ssum[lane] : 0.0;
    mm = lane < nw ? smax[lane] : -INFINITY;
    a = warp_sum(a); mm = warp_max(mm);
  }
}

// Reduction in two launches for kernels with ~10^5 blocks: every block leaves one partial
// (block_reduce_partial, no fence, no atomic), k_fold_partials folds them in index order.
__device__ __forceinline__ void block_reduce_partial(double s, float m, const RedWork w, int bid = -1) {
  __shared__ double ssum[32];
  __shared__ float smax[32];
  const int tid = linear_tid();
  const int nthr = blockDim.x * blockDim.y * blockDim.z;
  const int lane = tid & 31, wid = tid >> 5, nw = (nthr + 31) >> 5;
  s = warp_sum(s); m = warp_max(m);
  if (lane == 0) { ssum[wid] = s; smax[wid] = m; }
  __syncthreads();
  if (wid == 0) {
    s = lane < nw ? ssum[lane] : 0.0;
    m = lane < nw ? smax[lane] : -INFINITY;
    s = warp_sum(s); m = warp_max(m);
    if (lane == 0) { const int b = bid >= 0 ? bid : linear_bid(); w.psum[b] = s; w.pmax[b] = m; }
  }
}
struct FoldArgs { RedWork w; int n; Red* out; };
constexpr int FOLD_CHUNK = 4096;   // partials per block of k_fold_partials (one block when n <= FOLD_CHUNK)
// gridDim.x = ceil(n / FOLD_CHUNK) blocks fold one contiguous chunk each into psum[n + b]; the last one to finish
// (ticket 0) folds those in index order: the result does not depend on which block that is.
__global__ void __launch_bounds__(1024) k_fold_partials(const FoldArgs a) {
  pdl_entry();
  __shared__ double ssum[32];
  __shared__ float smax[32];
  __shared__ int last;
  double s; float m;
  if (gridDim.x == 1) {
    fold_partials(a.w.psum, a.w.pmax, 0, a.n, ssum, smax, s, m);
    if (threadIdx.x == 0) { a.out->sum = s; a.out->maxv = m; }
    return;
  }
  const int b = blockIdx.x, base = b * FOLD_CHUNK;
  fold_partials(a.w.psum, a.w.pmax, base, min(FOLD_CHUNK, a.n - base), ssum, smax, s, m);
  if (threadIdx.x == 0) {
    a.w.psum[a.n + b] = s; a.w.pmax[a.n + b] = m;
    __threadfence();
    last = atomicAdd(&a.w.ticket[0], 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  fold_partials(a.w.psum, a.w.pmax, a.n, (int)gridDim.x, ssum, smax, s, m);
  if (threadIdx.x == 0) { a.out->sum = s; a.out->maxv = m; a.w.ticket[0] = 0u; }
}

// Call from ALL threads of the block (contains __syncthreads).  `out` receives the result.
// Two ticket levels (groups of RED_GROUP blocks, then the groups): with >10^5 blocks a single ticket
// serialises that many atomics on one L2 address and the last block would fold 10^5 partials alone.
// w.psum/w.pmax hold the block partials followed (from index `nblocks`) by the group partials;
// w.ticket[0] is the top ticket, w.ticket[1+g] the ticket of group g.  All tickets return to zero.
__device__ __forceinline__ void block_reduce_finish(double s, float m, const RedWork w, Red* out) {
  __shared__ double ssum[32];
  __shared__ float smax[32];
  __shared__ int last;
  const int tid = linear_tid();
  const int nthr = blockDim.x * blockDim.y * blockDim.z;
  const int lane = tid & 31, wid = tid >> 5, nw = (nthr + 31) >> 5;
  s = warp_sum(s); m = warp_max(m);
  if (lane == 0) { ssum[wid] = s; smax[wid] = m; }
  __syncthreads();
  const int nblocks = gridDim.x * gridDim.y * gridDim.z;
  const int b = linear_bid();
  const int g = b / RED_GROUP, ng = (nblocks + RED_GROUP - 1) / RED_GROUP;
  const int gsize = min(RED_GROUP, nblocks - g * RED_GROUP);
  if (wid == 0) {
    s = lane < nw ? ssum[lane] : 0.0;
    m = lane < nw ? smax[lane] : -INFINITY;
    s = warp_sum(s); m = warp_max(m);
    if (lane == 0) {
      w.psum[b] = s; w.pmax[b] = m;
      __threadfence();
      const unsigned int t = atomicAdd(&w.ticket[1 + g], 1u);
      last = (t == (unsigned)gsize - 1);
    }
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  double a; float mm;
  fold_partials(w.psum, w.pmax, g * RED_GROUP, gsize, ssum, smax, a, mm);
  if (tid == 0) {
    w.psum[nblocks + g] = a; w.pmax[nblocks + g] = mm;
    w.ticket[1 + g] = 0u;
    __threadfence();
    const unsigned int t = atomicAdd(&w.ticket[0], 1u);
    last = (t == (unsigned)ng - 1);
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  fold_partials(w.psum, w.pmax, nblocks, ng, ssum, smax, a, mm);
  if (tid == 0) { out->sum = a; out->maxv = mm; w.ticket[0] = 0u; }
}

}  // namespace wl
