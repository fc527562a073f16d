// wl_noalloc.cuh -- wl_cfl_noalloc: an entry point that ALLOCATES NOTHING (included by wl_b200.cu: same translation unit
// as the scratch array, which cudaGetSymbolAddress requires).
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA-graph capture, tight serving loops, harnesses that forbid allocation).  It is the allocation-free
// variant of the CFL kernel (k_cfl / k_cfl4, reference src/Flow.jl:234-244): those reduce through per-simulation scratch
// (block partials + a ticket counter, cudaMalloc'ed by wl_create); this one reduces through a file-scope
// `static __device__` array of fixed size that every GPU holds from the moment the library is loaded until exit.
//
//  * The caller passes the input (velocity, device pointer, the reference's dense layout) and the output (one float,
//    device pointer) and a stream; the call only enqueues work on that stream (one 4-byte memset node + one kernel).
//  * Limit: one partial per grid row, at most WL_NOALLOC_MAX_ROWS rows (n[1]·n[2]); larger grids are rejected.
//  * Every call on a device shares the array and finds the previous call's contents in it.  What must start at zero (the
//    ticket) is cleared by the call itself, partials are overwritten before they are read.  Calls that OVERLAP on one GPU
//    (different streams, or graphs replayed concurrently) race on it: serialise them.
//  * In host code the array's NAME is a host-side stand-in, not a device address (a kernel handed it would fault or, where
//    the GPU can reach host memory, silently use the wrong memory): the address is resolved once per device with
//    cudaGetSymbolAddress, with that device current, and passed to the kernel as an ordinary pointer.
#pragma once

namespace wl {

constexpr int WL_NOALLOC_MAX_ROWS = 262144;                    // e.g. 512 x 512 rows; scratch = 1 MiB + 256 B per GPU
struct NoallocScratch {
  unsigned int ticket;                                         // rows finished; zero between calls
  unsigned int pad[63];
  float rowmax[WL_NOALLOC_MAX_ROWS];
};
static __device__ __align__(256) NoallocScratch wl_noalloc_scratch;

struct CflNaArgs {
  const float* u;      // (n0, n1, n2, D) dense, first dimension fastest, ghosts included
  int n0, n1, n2, D;
  float nu;
  NoallocScratch* sc;  // device address of wl_noalloc_scratch on the launching device
  float* dt_out;
};
// One block per interior row (j,k): σ = flux_out(I,u) (Flow.jl:238-244) for the row's interior cells, block maximum into
// rowmax[row]; the last block to finish folds the rows and writes Δt = min(10, 1/(max σ + 5ν)) (Flow.jl:236).  The
// maximum starts at 0: the reference's `maximum(a.σ)` runs over the whole array, whose lower ghost cells hold 0.
__global__ void __launch_bounds__(128) k_cfl_noalloc(const CflNaArgs a) {
  __shared__ float smax[4];
  __shared__ int last;
  const int D = a.D, n0 = a.n0, n1 = a.n1;
  const int j = 1 + blockIdx.x % (n1 - 2), k = D == 3 ? 1 + blockIdx.x / (n1 - 2) : 0;
  const long long sy = n0, sz = (long long)n0 * n1, sc = sz * a.n2;
  const float* u0 = a.u + sy * j + sz * k;
  float m = 0.f;
  for (int i = 1 + threadIdx.x; i <= n0 - 2; i += blockDim.x) {
    float s = 0.f;
    s += fmaxf(0.f, u0[i + 1]) + fmaxf(0.f, -u0[i]);
    s += fmaxf(0.f, u0[sc + i + sy]) + fmaxf(0.f, -u0[sc + i]);
    if (D == 3) s += fmaxf(0.f, u0[2 * sc + i + sz]) + fmaxf(0.f, -u0[2 * sc + i]);
    m = fmaxf(m, s);
  }
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    m = fmaxf(fmaxf(smax[0], smax[1]), fmaxf(smax[2], smax[3]));
    a.sc->rowmax[blockIdx.x] = m;
    __threadfence();
    last = atomicAdd(&a.sc->ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  m = 0.f;
  for (int r = threadIdx.x; r < (int)gridDim.x; r += blockDim.x) m = fmaxf(m, __ldcg(&a.sc->rowmax[r]));
  m = warp_max(m);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    m = fmaxf(fmaxf(smax[0], smax[1]), fmaxf(smax[2], smax[3]));
    *a.dt_out = fminf(10.f, 1.f / (m + 5.f * a.nu));
    a.sc->ticket = 0u;                                         // leave the array ready for the next call
  }
}

}  // namespace wl

namespace {
// device address of the scratch array, per device ordinal (resolved on the first call for that device)
NoallocScratch* g_noalloc_ptr[64] = {};
}

extern "C" int wl_cfl_noalloc(const float* u_dev, int32_t D, const int32_t n[3], float nu, float* dt_dev, void* cuda_stream) {
  if (!u_dev || !dt_dev || !n) return fail(nullptr, WL_ERR_INVALID, "wl_cfl_noalloc: null argument");
  if (D != 2 && D != 3) return fail(nullptr, WL_ERR_INVALID, "wl_cfl_noalloc: D must be 2 or 3");
  const int n0 = n[0], n1 = n[1], n2 = D == 3 ? n[2] : 1;
  if (n0 < 3 || n1 < 3 || (D == 3 && n2 < 3)) return fail(nullptr, WL_ERR_INVALID, "wl_cfl_noalloc: every extent needs at least one interior cell (n includes the two ghost cells)");
  const long long rows = (long long)(n1 - 2) * (D == 3 ? n2 - 2 : 1);
  if (rows > WL_NOALLOC_MAX_ROWS)
    return fail(nullptr, WL_ERR_INVALID, "wl_cfl_noalloc: the grid has " + std::to_string(rows) + " interior rows, the static scratch holds at most " +
                                             std::to_string(WL_NOALLOC_MAX_ROWS) + " (WL_NOALLOC_MAX_ROWS); use wl_cfl on a wl_sim for larger grids");
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return fail(nullptr, WL_ERR_NO_DEVICE, "wl_cfl_noalloc: no current CUDA device");
  if (!g_noalloc_ptr[dev]) {                                     // set-up, once per device, with that device current
    NoallocScratch* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, wl_noalloc_scratch);
    if (e != cudaSuccess || !p) return fail(nullptr, WL_ERR_CUDA, std::string("wl_cfl_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    g_noalloc_ptr[dev] = p;
  }
  NoallocScratch* sc = g_noalloc_ptr[dev];
  cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
  // the ticket must start at zero whatever an earlier (possibly aborted) call left behind
  cudaError_t e = cudaMemsetAsync(&sc->ticket, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return fail(nullptr, WL_ERR_CUDA, std::string("wl_cfl_noalloc: cudaMemsetAsync: ") + cudaGetErrorString(e));
  CflNaArgs a{u_dev, n0, n1, n2, D, nu, sc, dt_dev};
  k_cfl_noalloc<<<dim3((unsigned)rows), dim3(128), 0, st>>>(a);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail(nullptr, WL_ERR_CUDA, std::string("wl_cfl_noalloc: launch: ") + cudaGetErrorString(e));
  return WL_OK;
}
