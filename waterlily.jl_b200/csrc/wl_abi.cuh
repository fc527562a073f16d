// wl_abi.cuh -- the extern "C" entry points of include/wl_b200.h (included by wl_b200.cu).
#pragma once

namespace {

int fail(wl_sim* s, int code, const std::string& msg) {
  if (s) { s->err = msg; s->status = code; } else g_create_error = msg;
  return code;
}

// MultiLevelPoisson(x,L,z;maxlevels,perdir) (MultiLevelPoisson.jl:68-76) + restrictML (:33-41).
// The hierarchy follows the GLOBAL grid; with several ranks a level stays distributed while every
// rank keeps an even number (>= dist.min_planes) of planes, coarser levels are replicated.
int build_levels(wl_sim* s) {
  for (size_t l = 1; l < s->lv.size(); ++l) {
    Level& v = s->lv[l];
    free_field(v.L); free_field(v.x); free_field(v.z);
  }
  for (auto& v : s->lv) {
    free_field(v.r); free_field(v.eps); free_field(v.Dg); free_field(v.iD); free_field(v.tmp);
    if (v.ucore) cudaFree(v.ucore);
    if (v.uflag) cudaFree(v.uflag);
    if (v.ulist) cudaFree(v.ulist);
    if (v.ucount) cudaFree(v.ucount);
  }
  s->lv.clear();
  s->lv.reserve(s->cfg.maxlevels + 2);
  s->lv.emplace_back();
  if (s->arena) {                               // level arrays are carved after the flow fields; a rebuild starts over
    if (!s->arena_mark) s->arena_mark = s->arena_used;
    s->arena_used = s->arena_mark;
  }
  const int nr = s->dist.n;
  int gn[3] = {s->lay.n0, s->lay.n1, s->D == 3 ? s->dist.nzg + 2 : 1};   // global sizes incl. ghosts
  auto gcells = [&](const int* g) {
    long long c = (long long)(g[0] - 2) * (g[1] - 2);
    return s->D == 3 ? c * (g[2] - 2) : c;
  };
  {
    Level& l0 = s->lv.back();
    l0.lay = s->lay; l0.L = s->mu0; l0.x = s->p; l0.z = s->sigma; l0.owns = false;
    l0.dist = nr > 1; l0.per = s->per; l0.koff = s->per.kz0; l0.ncells = gcells(gn);
    if (int e = alloc_field(s, l0.r, s->lay, 1)) return e;
    if (int e = alloc_field(s, l0.eps, s->lay, 1)) return e;
    if (int e = alloc_field(s, l0.Dg, s->lay, 1)) return e;
    if (int e = alloc_field(s, l0.iD, s->lay, 1)) return e;
    if (int e = alloc_field(s, l0.tmp, s->lay, 1)) return e;
    if (int e = set_diag_h(s, l0)) return e;
  }
  auto anydiv = [&]() {
    for (int d = 0; d < s->D; ++d) if (divisible(gn[d])) return true;
    return false;
  };
  while (anydiv() && (int)s->lv.size() <= s->cfg.maxlevels) {
    const Level& f = s->lv.back();
    Level c;
    int ng[3];
    for (int d = 0; d < 3; ++d) {
      c.c[d] = d < s->D && divisible(gn[d]);
      ng[d] = c.c[d] ? 1 + gn[d] / 2 : gn[d];
    }
    for (int d = 0; d < 3; ++d) gn[d] = ng[d];
    c.ncells = gcells(gn);
    c.per = s->per_glob;
    if (f.dist) {
      const int nzf = f.lay.n2 - 2, nloc = c.c[2] ? nzf / 2 : nzf;
      // small levels are launch/latency bound: solving them redundantly on every rank is cheaper than
      // ~8 halo exchanges per level per V-cycle
      if (nloc >= s->dist.min_planes && nloc % 2 == 0 && c.ncells > s->dist.min_cells) {
        c.dist = true;
        ng[2] = nloc + 2;
        c.per = s->per; c.koff = s->dist.rank * nloc; c.per.kz0 = c.koff;
      }
    }
    c.lay = make_lay(s->D, ng);
    c.owns = true;
    for (int i = 0; i < 3; ++i) {                       // restrictL (MultiLevelPoisson.jl:20-26) of a body-free fine level
      float v = f.lu[i];
      for (int d = 0; d < s->D; ++d) if (d != i && c.c[d]) v = v + v;
      c.lu[i] = (i < s->D && c.c[i]) ? v / 2.f : v;
    }
    s->lv.push_back(c);
    Level& co = s->lv.back();
    if (int e = alloc_field(s, co.L, co.lay, s->D)) return e;
    if (int e = alloc_field(s, co.x, co.lay, 1)) return e;
    if (int e = alloc_field(s, co.z, co.lay, 1)) return e;
    if (int e = alloc_field(s, co.r, co.lay, 1)) return e;
    if (int e = alloc_field(s, co.eps, co.lay, 1)) return e;
    if (int e = alloc_field(s, co.Dg, co.lay, 1)) return e;
    if (int e = alloc_field(s, co.iD, co.lay, 1)) return e;
    if (int e = alloc_field(s, co.tmp, co.lay, 1)) return e;
    if (int e = restrictL(s, s->lv[s->lv.size() - 2], co)) return e;
    if (int e = set_diag_h(s, co)) return e;
  }
  if (s->lv.size() <= 2) return fail(s, WL_ERR_INVALID, "MultiLevelPoisson requires size=a2^n, where n>2");
  return WL_OK;
}

// Bytes the fields of a slab-decomposed sim need (same walk as create_impl + build_levels), so that
// they can all be carved from ONE peer-mapped allocation.
size_t plan_arena_bytes(wl_sim* s) {
  auto fld = [](const Lay& l, int ncomp) { return (((size_t)l.cstride * ncomp * sizeof(float) + 256) + 255) & ~(size_t)255; };
  const int D = s->D;
  size_t tot = ARENA_HDR;                                               // flags, gather tables
  tot += fld(s->lay, D) * 6 + fld(s->lay, 1) * 2 + fld(s->lay, D * D);   // u,u0,f,V,mu0,w ; p,sigma ; mu1
  tot += fld(s->lay, 1) * 5;                                            // level 0: r, eps, D, iD, tmp
  int gn[3] = {s->lay.n0, s->lay.n1, s->dist.nzg + 2};
  bool fdist = true;
  int nzf = s->lay.n2 - 2;
  int nlev = 1;
  auto anydiv = [&]() { for (int d = 0; d < D; ++d) if (divisible(gn[d])) return true; return false; };
  while (anydiv() && nlev <= s->cfg.maxlevels) {
    int ng[3], c[3];
    for (int d = 0; d < 3; ++d) { c[d] = d < D && divisible(gn[d]); ng[d] = c[d] ? 1 + gn[d] / 2 : gn[d]; }
    for (int d = 0; d < 3; ++d) gn[d] = ng[d];
    const long long cells = (long long)(gn[0] - 2) * (gn[1] - 2) * (gn[2] - 2);
    bool cdist = false;
    if (fdist) {
      const int nloc = c[2] ? nzf / 2 : nzf;
      if (nloc >= s->dist.min_planes && nloc % 2 == 0 && cells > s->dist.min_cells) { cdist = true; ng[2] = nloc + 2; nzf = nloc; }
    }
    fdist = cdist;
    const Lay l = make_lay(D, ng);
    tot += fld(l, D) + fld(l, 1) * 7;
    ++nlev;
  }
  return tot + tot / 64 + (1 << 20);
}

// Allocate the arena, exchange its IPC handle with the slab neighbours and map theirs.
int setup_p2p(wl_sim* s) {
  if (getenv("WL_NO_P2P")) return WL_OK;
  Dist& d = s->dist;
  const size_t cap = plan_arena_bytes(s);
  if (cudaMalloc(&s->arena, cap) != cudaSuccess) { cudaGetLastError(); s->arena = nullptr; return WL_OK; }   // fall back to NCCL halos
  s->arena_cap = cap; s->arena_used = ARENA_HDR;
  cudaMemsetAsync(s->arena, 0, ARENA_HDR, s->st);
  cudaIpcMemHandle_t mine;
  if (cudaIpcGetMemHandle(&mine, s->arena) != cudaSuccess) { cudaGetLastError(); return WL_OK; }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  char* dbuf = nullptr;
  WL_CUDA(s, cudaMalloc(&dbuf, 64 * (size_t)d.n));
  WL_CUDA(s, cudaMemcpyAsync(dbuf + 64 * d.rank, &mine, 64, cudaMemcpyHostToDevice, s->st));
  WL_NCCL(s, ncclAllGather(dbuf + 64 * d.rank, dbuf, 64, ncclChar, d.comm, s->st));
  std::vector<cudaIpcMemHandle_t> all(d.n);
  WL_CUDA(s, cudaMemcpyAsync(all.data(), dbuf, 64 * (size_t)d.n, cudaMemcpyDeviceToHost, s->st));
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  cudaFree(dbuf);
  bool ok = true;
  const bool all_peers = d.n <= P2P_MAXRANKS && !getenv("WL_NO_P2P_GATHER");   // one NVSwitch box: map everyone
  if (all_peers) {
    for (int r = 0; r < d.n && ok; ++r) {
      if (r == d.rank) { s->peer_all[r] = s->arena; continue; }
      ok = cudaIpcOpenMemHandle((void**)&s->peer_all[r], all[r], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
      if (!ok) s->peer_all[r] = nullptr;
    }
    if (ok) {
      if (d.lo >= 0) s->peer_lo = s->peer_all[d.lo];
      if (d.hi >= 0) s->peer_hi = s->peer_all[d.hi];
    }
  } else {
    if (d.lo >= 0) ok = ok && cudaIpcOpenMemHandle((void**)&s->peer_lo, all[d.lo], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
    if (d.hi >= 0) {
      if (d.hi == d.lo) s->peer_hi = s->peer_lo;
      else ok = ok && cudaIpcOpenMemHandle((void**)&s->peer_hi, all[d.hi], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
    }
  }
  if (!ok) cudaGetLastError();
  // every rank must agree on the transport (the flag protocol needs both ends): gather the verdicts
  int* dv = nullptr;
  WL_CUDA(s, cudaMalloc(&dv, sizeof(int) * (size_t)d.n));
  const int v = ok ? 1 : 0;
  WL_CUDA(s, cudaMemcpyAsync(dv + d.rank, &v, sizeof(int), cudaMemcpyHostToDevice, s->st));
  WL_NCCL(s, ncclAllGather(dv + d.rank, dv, sizeof(int), ncclChar, d.comm, s->st));
  std::vector<int> vs(d.n);
  WL_CUDA(s, cudaMemcpyAsync(vs.data(), dv, sizeof(int) * (size_t)d.n, cudaMemcpyDeviceToHost, s->st));
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  cudaFree(dv);
  s->p2p = true;
  for (int q = 0; q < d.n; ++q) s->p2p = s->p2p && vs[q] == 1;
  s->p2p_gather = s->p2p && all_peers;
  return WL_OK;
}

int check_async(wl_sim* s, const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(s, WL_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
  return WL_OK;
}

}  // namespace

extern "C" {

int wl_version(void) { return 100; }

const char* wl_last_error(const wl_sim* s) { return s ? s->err.c_str() : g_create_error.c_str(); }

static int create_impl(const wl_config* cfg, const wl_dist* dd, wl_sim** out) {
  if (!cfg || !out) return fail(nullptr, WL_ERR_INVALID, "null argument");
  *out = nullptr;
  const int nranks = dd ? dd->nranks : 1, rank = dd ? dd->rank : 0;
  if (nranks < 1 || rank < 0 || rank >= nranks || nranks > MAXRANKS) return fail(nullptr, WL_ERR_INVALID, "bad rank / nranks");
  if (nranks > 1) {
    if (cfg->D != 3) return fail(nullptr, WL_ERR_UNSUPPORTED, "slab decomposition needs a 3D grid");
    if (cfg->flags & WL_FLAG_NO_FUSED) return fail(nullptr, WL_ERR_UNSUPPORTED, "slab decomposition runs the fused kernels only");
    if (cfg->N[2] % nranks != 0 || (cfg->N[2] / nranks) % 2 != 0 || cfg->N[2] / nranks < 4)
      return fail(nullptr, WL_ERR_INVALID, "the last dimension must split into an even number (>= 4) of planes per rank");
  }
  if (cfg->D != 2 && cfg->D != 3) return fail(nullptr, WL_ERR_INVALID, "D must be 2 or 3");
  for (int d = 0; d < cfg->D; ++d)
    if (cfg->N[d] < 2) return fail(nullptr, WL_ERR_INVALID, "each dimension needs at least 2 cells");
  if (cfg->scheme < 0 || cfg->scheme > 2) return fail(nullptr, WL_ERR_UNSUPPORTED, "convective scheme must be quick, cds or vanLeer");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || cfg->device >= ndev)
    return fail(nullptr, WL_ERR_NO_DEVICE, "no CUDA device: libwaterlily_b200 has no CPU fallback");
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, cfg->device);
  if (prop.major < 10)
    return fail(nullptr, WL_ERR_NO_DEVICE, "device is not sm_100 (Blackwell); this library ships sm_100a code only");
  wl_sim* s = new wl_sim();
  s->cfg = *cfg;
  s->D = cfg->D;
  s->device = cfg->device;
  s->dt = cfg->dt0;
  if (s->cfg.maxlevels <= 0) s->cfg.maxlevels = 10;
  if (s->cfg.itmx <= 0) s->cfg.itmx = 32;
  if (s->cfg.itmx > 32) s->cfg.itmx = 32;
  if (s->cfg.tol <= 0) s->cfg.tol = 2e-3f;
  auto bail = [&](int code) { g_create_error = s->err; wl_destroy(s); return code; };
  if (cudaSetDevice(cfg->device) != cudaSuccess) { s->err = "cudaSetDevice failed"; return bail(WL_ERR_CUDA); }
  if (cudaStreamCreateWithFlags(&s->st, cudaStreamNonBlocking) != cudaSuccess) { s->err = "stream create failed"; return bail(WL_ERR_CUDA); }
  int ng[3];
  for (int d = 0; d < 3; ++d) { ng[d] = cfg->N[d] + 2; s->per.p[d] = d < cfg->D ? (cfg->perdir[d] != 0) : 0; }
  s->per.zlo = s->per.zhi = s->per.kz0 = 0;
  s->per_glob = s->per;
  s->dist.rank = rank; s->dist.n = nranks; s->dist.nzg = cfg->N[2]; s->dist.nzl = cfg->N[2];
  if (const char* e = getenv("WL_DIST_MIN_PLANES")) s->dist.min_planes = std::max(2, atoi(e));
  if (const char* e = getenv("WL_DIST_MIN_CELLS")) s->dist.min_cells = atoll(e);
  if (nranks > 1) {
    Dist& d = s->dist;
    const bool pz = s->per.p[2] != 0;
    d.nzl = cfg->N[2] / nranks;
    d.lo = rank > 0 ? rank - 1 : (pz ? nranks - 1 : -1);
    d.hi = rank < nranks - 1 ? rank + 1 : (pz ? 0 : -1);
    d.lo_wrap = rank == 0 && pz; d.hi_wrap = rank == nranks - 1 && pz;
    s->per.p[2] = 0; s->per.zlo = d.lo >= 0; s->per.zhi = d.hi >= 0; s->per.kz0 = rank * d.nzl;
    ng[2] = d.nzl + 2;
    { const std::string why = nccl_load(); if (!why.empty()) { s->err = why; return bail(WL_ERR_CUDA); } }
    ncclUniqueId id;
    static_assert(sizeof(id) <= sizeof(dd->nccl_id), "wl_dist.nccl_id too small");
    memcpy(&id, dd->nccl_id, sizeof(id));
    ncclResult_t r = ncclCommInitRank(&d.comm, nranks, id, rank);
    if (r != ncclSuccess) { s->err = std::string("ncclCommInitRank: ") + ncclGetErrorString(r); return bail(WL_ERR_CUDA); }
    if (cudaMalloc(&s->d_redg, sizeof(Red) * NRED * nranks) != cudaSuccess ||
        cudaHostAlloc(&s->h_redg, sizeof(Red) * NRED * nranks, cudaHostAllocDefault) != cudaSuccess) {
      s->err = "reduction gather buffers allocation failed";
      return bail(WL_ERR_CUDA);
    }
  }
  s->lay = make_lay(cfg->D, ng, nranks > 1 ? 2 : 0);
  const int D = s->D;
  if (nranks > 1) { if (int e2 = setup_p2p(s)) return bail(e2); }
  int e = 0;
  if ((e = alloc_field(s, s->u, s->lay, D)) || (e = alloc_field(s, s->u0, s->lay, D)) ||
      (e = alloc_field(s, s->f, s->lay, D)) || (e = alloc_field(s, s->p, s->lay, 1)) ||
      (e = alloc_field(s, s->sigma, s->lay, 1)) || (e = alloc_field(s, s->V, s->lay, D)) ||
      (e = alloc_field(s, s->mu0, s->lay, D)) || (e = alloc_field(s, s->mu1, s->lay, D * D)) ||
      (D == 3 && (e = alloc_field(s, s->w, s->lay, D))))
    return bail(e);
  if (cudaMalloc(&s->band_count, sizeof(int)) != cudaSuccess) { s->err = "band counter allocation failed"; return bail(WL_ERR_CUDA); }
  cudaMemsetAsync(s->band_count, 0, sizeof(int), s->st);
  // reduction scratch: one partial per block of the largest launch
  {
    dim3 g = grid_all(s->lay);
    const size_t ngroups = ((size_t)g.x * g.y * g.z) / RED_GROUP + 8;
    size_t nb = (size_t)g.x * g.y * g.z + ngroups + 64;     // block partials + group partials
    if (cudaMalloc(&s->redw.psum, nb * sizeof(double)) != cudaSuccess ||
        cudaMalloc(&s->redw.pmax, nb * sizeof(float)) != cudaSuccess ||
        cudaMalloc(&s->redw.ticket, (ngroups + 1) * sizeof(unsigned)) != cudaSuccess ||
        cudaMalloc(&s->d_red, sizeof(Red) * NRED) != cudaSuccess ||
        cudaHostAlloc(&s->h_red, sizeof(Red) * NRED, cudaHostAllocDefault) != cudaSuccess) {
      s->err = "reduction scratch allocation failed";
      return bail(WL_ERR_CUDA);
    }
    cudaMemsetAsync(s->redw.ticket, 0, (ngroups + 1) * sizeof(unsigned), s->st);
    cudaMemsetAsync(s->d_red, 0, sizeof(Red) * NRED, s->st);
  }
  for (int i = 0; i < 8; ++i) cudaEventCreate(&s->ev[i]);
  // Flow constructor (Flow.jl:139-145): u = uBC everywhere, BC!, exitBC!(u,u,0), u⁰ = u, μ₀ = 1 then BC!(μ₀,0)
  {
    std::vector<float> ones((size_t)s->lay.cstride * D + 64, 1.f);
    cudaMemcpyAsync(s->mu0.alloc, ones.data(), ones.size() * sizeof(float), cudaMemcpyHostToDevice, s->st);
    for (int c = 0; c < D; ++c) {
      std::fill(ones.begin(), ones.begin() + s->lay.cstride, cfg->uBC[c]);
      cudaMemcpyAsync(s->u.alloc + (size_t)s->lay.cstride * c, ones.data(), s->lay.cstride * sizeof(float),
                      cudaMemcpyHostToDevice, s->st);
      cudaStreamSynchronize(s->st);
    }
  }
  *out = s;
  if ((e = wl_finish_ic(s, cfg->uBC))) { *out = nullptr; return bail(e); }
  bc_vec(s, s->mu0, nullptr, false);
  if ((e = halo(s, s->mu0, 1))) { *out = nullptr; return bail(e); }
  if (cudaStreamSynchronize(s->st) != cudaSuccess || cudaGetLastError() != cudaSuccess) {
    s->err = "device initialisation failed";
    *out = nullptr;
    return bail(WL_ERR_CUDA);
  }
  return WL_OK;
}

int wl_create(const wl_config* cfg, wl_sim** out) { return create_impl(cfg, nullptr, out); }

// Slab decomposition: cfg describes the GLOBAL grid; every rank calls this with the
// same cfg and NCCL id and receives the sim of its slab of N[2]/nranks planes.
int wl_create_slab(const wl_config* cfg, const wl_dist* dist, wl_sim** out) {
  if (!dist) return fail(nullptr, WL_ERR_INVALID, "null argument");
  return create_impl(cfg, dist, out);
}
int wl_nccl_unique_id(char id[128]) {
  if (!id) return WL_ERR_INVALID;
  { const std::string why = nccl_load(); if (!why.empty()) return fail(nullptr, WL_ERR_CUDA, why); }
  ncclUniqueId u;
  if (ncclGetUniqueId(&u) != ncclSuccess) return fail(nullptr, WL_ERR_CUDA, "ncclGetUniqueId failed");
  memset(id, 0, 128);
  memcpy(id, &u, sizeof(u));
  return WL_OK;
}
int wl_slab_info(const wl_sim* s, int32_t* rank, int32_t* nranks, int32_t* k_offset, int32_t* n_local, int32_t* n_global) {
  if (!s) return WL_ERR_INVALID;
  if (rank) *rank = s->dist.rank;
  if (nranks) *nranks = s->dist.n;
  if (k_offset) *k_offset = s->per.kz0;
  if (n_local) *n_local = s->dist.nzl;
  if (n_global) *n_global = s->dist.nzg;
  return WL_OK;
}
int wl_halo_transport(const wl_sim* s) { return s ? (s->dist.n > 1 ? (s->p2p ? 2 : 1) : 0) : -1; }
int wl_halo_timing(wl_sim* s, double out[4]) {
  if (!s || !out) return WL_ERR_INVALID;
  for (int q = 0; q < 4; ++q) out[q] = 0.0;
  if (!s->p2p) return WL_OK;
  cudaSetDevice(s->device);
  unsigned long long st[4];
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  WL_CUDA(s, cudaMemcpy(st, s->arena + 16 * sizeof(unsigned int), sizeof(st), cudaMemcpyDeviceToHost));
  for (int q = 0; q < 4; ++q) out[q] = (double)st[q];
  return WL_OK;
}
int wl_halo_stats(const wl_sim* s, int64_t* exchanges, int64_t* bytes) {
  if (!s) return WL_ERR_INVALID;
  if (exchanges) *exchanges = s->halo_msgs;
  if (bytes) *bytes = s->halo_bytes;
  return WL_OK;
}

void wl_destroy(wl_sim* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  if (s->st) cudaStreamSynchronize(s->st);
  if (s->arena) {
    // no neighbour may still be storing into this arena: meet once before unmapping / freeing
    if (s->dist.comm && s->d_redg) { ncclAllGather(s->d_red, s->d_redg, sizeof(Red) * NRED, ncclChar, s->dist.comm, s->st); cudaStreamSynchronize(s->st); }
    bool mapped_all = false;
    for (int r = 0; r < P2P_MAXRANKS; ++r)
      if (s->peer_all[r]) { mapped_all = true; if (s->peer_all[r] != s->arena) cudaIpcCloseMemHandle(s->peer_all[r]); }
    if (!mapped_all) {
      if (s->peer_lo) cudaIpcCloseMemHandle(s->peer_lo);
      if (s->peer_hi && s->peer_hi != s->peer_lo) cudaIpcCloseMemHandle(s->peer_hi);
    }
    cudaFree(s->arena);
    s->arena = nullptr;
  }
  if (s->dist.comm) ncclCommDestroy(s->dist.comm);
  if (s->d_redg) cudaFree(s->d_redg);
  if (s->stage) cudaFree(s->stage);
  if (s->bc_tab) cudaFree(s->bc_tab);
  if (s->fw_off) cudaFree(s->fw_off);
  if (s->fw_w) cudaFree(s->fw_w);
  if (s->h_redg) cudaFreeHost(s->h_redg);
  for (size_t l = 0; l < s->lv.size(); ++l) {
    Level& v = s->lv[l];
    if (v.owns) { free_field(v.L); free_field(v.x); free_field(v.z); }
    free_field(v.r); free_field(v.eps); free_field(v.Dg); free_field(v.iD); free_field(v.tmp);
    if (v.ucore) cudaFree(v.ucore);
    if (v.uflag) cudaFree(v.uflag);
    if (v.ulist) cudaFree(v.ulist);
    if (v.ucount) cudaFree(v.ucount);
  }
  free_field(s->u); free_field(s->u0); free_field(s->f); free_field(s->p);
  free_field(s->sigma); free_field(s->V); free_field(s->mu0); free_field(s->mu1); free_field(s->w);
  free_field(s->mf_P); free_field(s->mf_U); free_field(s->mf_UU);
  free_field(s->accf[0]); free_field(s->accf[1]);
  if (s->band_list) cudaFree(s->band_list);
  if (s->band_list2) cudaFree(s->band_list2);
  if (s->band_mark) cudaFree(s->band_mark);
  if (s->band_count2) cudaFree(s->band_count2);
  if (s->band_count) cudaFree(s->band_count);
  if (s->redw.psum) cudaFree(s->redw.psum);
  if (s->redw.pmax) cudaFree(s->redw.pmax);
  if (s->redw.ticket) cudaFree(s->redw.ticket);
  if (s->d_red) cudaFree(s->d_red);
  if (s->h_red) cudaFreeHost(s->h_red);
  for (int i = 0; i < 8; ++i) if (s->ev[i]) cudaEventDestroy(s->ev[i]);
  if (s->st) cudaStreamDestroy(s->st);
  delete s;
}

int wl_nlevels(const wl_sim* s) { return s ? (int)s->lv.size() : 0; }

int wl_field_dims(const wl_sim* s, int field, int level, int32_t dims[5], int32_t* ndims) {
  if (!s) return WL_ERR_INVALID;
  Lay lay;
  DevField* f = flow_field(const_cast<wl_sim*>(s), field, level, &lay);
  if (!f) return WL_ERR_INVALID;
  int n = 0;
  dims[n++] = lay.n0; dims[n++] = lay.n1;
  if (s->D == 3) dims[n++] = lay.n2;
  if (f->ncomp == s->D * s->D && (field == WL_MU1 || field == WL_MEAN_UU)) { dims[n++] = s->D; dims[n++] = s->D; }
  else if (f->ncomp > 1) dims[n++] = f->ncomp;
  *ndims = n;
  return WL_OK;
}

static int xfer(wl_sim* s, int field, int level, float* host, size_t count, bool up) {
  if (!s || !host) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  Lay lay;
  DevField* f = flow_field(s, field, level, &lay);
  if (!f) return fail(s, WL_ERR_INVALID, "unknown field/level");
  const size_t per = (size_t)lay.n0 * lay.n1 * lay.n2;
  if (count != per * f->ncomp) return fail(s, WL_ERR_INVALID, "host buffer has the wrong number of elements");
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  const size_t bytes = count * sizeof(float);
  if (s->stage_bytes < bytes) {
    if (s->stage) cudaFree(s->stage);
    s->stage = nullptr; s->stage_bytes = 0;
    if (cudaMalloc(&s->stage, bytes) == cudaSuccess) s->stage_bytes = bytes; else cudaGetLastError();
  }
  if (s->stage) {
    // one contiguous DMA + a repack kernel (pinned host memory runs at the PCIe rate)
    PackArgs pa{lay, f->base, s->stage, f->ncomp, up ? 1 : 0};
    const unsigned nb = (unsigned)std::min<size_t>((count + 1023) / 1024, 148 * 16);
    if (up) {
      WL_CUDA(s, cudaMemcpyAsync(s->stage, host, bytes, cudaMemcpyHostToDevice, s->st));
      LAUNCH(s, k_repack, dim3(nb), dim3(256), pa);
    } else {
      LAUNCH(s, k_repack, dim3(nb), dim3(256), pa);
      WL_CUDA(s, cudaMemcpyAsync(host, s->stage, bytes, cudaMemcpyDeviceToHost, s->st));
    }
  } else {
    for (int c = 0; c < f->ncomp; ++c) {
      float* dev = f->base + lay.cstride * c;
      float* h = host + per * c;
      if (up)
        WL_CUDA(s, cudaMemcpy2DAsync(dev, lay.pitch * sizeof(float), h, lay.n0 * sizeof(float), lay.n0 * sizeof(float),
                                     (size_t)lay.n1 * lay.n2, cudaMemcpyHostToDevice, s->st));
      else
        WL_CUDA(s, cudaMemcpy2DAsync(h, lay.n0 * sizeof(float), dev, lay.pitch * sizeof(float), lay.n0 * sizeof(float),
                                     (size_t)lay.n1 * lay.n2, cudaMemcpyDeviceToHost, s->st));
    }
  }
  if (up && s->dist.n > 1 && field < WL_MEAN_P && (field < 10 || s->lv[level].dist)) {
    // the caller's slab carries one ghost plane; deeper halo planes come from the neighbours
    if (int e = halo(s, *f, (field == WL_U || field == WL_U0) ? 3 : 1)) return e;
  }
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  if (up && field == WL_MU1) { s->band_dirty = true; s->band_from_measure = false; }
  if (up && (field == WL_MU1 || field == WL_MU0 || field == WL_V)) s->free_flow = false;
  // the uniform-coefficient tile flags describe L as of the last update!: a changed L switches them off until the next one
  if (up && field == WL_MU0 && !s->lv.empty()) s->lv[0].uvalid = false;
  if (up && (field == WL_L || field == WL_DIAG || field == WL_IDIAG) && level >= 0 && level < (int)s->lv.size()) s->lv[level].uvalid = false;
  return WL_OK;
}
int wl_download(wl_sim* s, int field, int level, float* host, size_t count) { return xfer(s, field, level, host, count, false); }
int wl_upload(wl_sim* s, int field, int level, const float* host, size_t count) {
  return xfer(s, field, level, const_cast<float*>(host), count, true);
}

int wl_finish_ic(wl_sim* s, const float uBC[3]) {
  if (!s) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  bc_vec(s, s->u, uBC ? uBC : s->cfg.uBC, s->cfg.exitBC != 0);
  if (int e = exitbc(s, s->u, s->u, 0.f)) return e;  // runs even when exitBC=false (Flow.jl:141)
  if (int e = halo(s, s->u, 3)) return e;
  WL_CUDA(s, cudaMemcpyAsync(s->u0.alloc, s->u.alloc, s->u.bytes, cudaMemcpyDeviceToDevice, s->st));
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  return check_async(s, "wl_finish_ic");
}

extern "C++" {
template <class B>
static int measure_impl(wl_sim* s, const B& body, double eps, const char* what) {
  cudaSetDevice(s->device);
  MeasArgsT<B> a{s->lay, body, (float)eps, s->per.kz0, s->sigma.base, s->V.base, s->mu0.base, s->mu1.base,
                 nullptr, nullptr, 0};
  if (s->D == 3 && s->band_count) {                 // K-MOM's band list comes out of the same pass (no k_band_scan)
    WL_CUDA(s, cudaMemsetAsync(s->band_count, 0, sizeof(int), s->st));
    a.band_list = s->band_list; a.band_count = s->band_count; a.band_cap = s->band_cap;
    s->band_from_measure = true;
  }
  DISPATCH_DB(s, k_measure, B, grid_all(s->lay), blk(s->D), a);
  s->band_dirty = true; s->free_flow = false;
  if (!s->lv.empty()) s->lv[0].uvalid = false;          // μ₀ ≡ L changed: flags come back with update!
  bc_vec(s, s->mu0, nullptr, false);
  bc_vec(s, s->V, nullptr, s->cfg.exitBC != 0);
  if (int e = halo(s, s->mu0, 1)) return e;
  if (int e = halo(s, s->V, 1)) return e;
  return check_async(s, what);
}
}  // extern "C++"

int wl_bc_table_size(const wl_sim* s, size_t* count) {
  if (!s || !count) return WL_ERR_INVALID;
  const int N[3] = {s->lay.n0, s->lay.n1, s->D == 3 ? s->lay.n2 : 1};
  size_t tot = 0;
  for (int dj = 0; dj < s->D; ++dj) tot += (size_t)3 * s->D * ((size_t)N[0] * N[1] * N[2] / N[dj]);
  *count = tot;
  return WL_OK;
}

int wl_set_bc_tables(wl_sim* s, const float* tables, size_t count) {
  if (!s) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  if (!tables) {                                      // back to the constant uBC of the configuration
    WL_CUDA(s, cudaStreamSynchronize(s->st));
    if (s->bc_tab) cudaFree(s->bc_tab);
    s->bc_tab = nullptr;
    return WL_OK;
  }
  size_t want = 0;
  wl_bc_table_size(s, &want);
  if (count != want) return fail(s, WL_ERR_INVALID, "boundary tables have the wrong number of elements (wl_bc_table_size)");
  const int N[3] = {s->lay.n0, s->lay.n1, s->D == 3 ? s->lay.n2 : 1};
  int off = 0;
  for (int dj = 0; dj < 3; ++dj) s->bc_psz[dj] = dj < s->D ? N[0] * N[1] * N[2] / N[dj] : 0;
  for (int ci = 0; ci < s->D; ++ci)
    for (int dj = 0; dj < s->D; ++dj) { s->bc_off[ci * s->D + dj] = off; off += 3 * s->bc_psz[dj]; }
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  if (!s->bc_tab) WL_CUDA(s, cudaMalloc(&s->bc_tab, want * sizeof(float)));
  WL_CUDA(s, cudaMemcpyAsync(s->bc_tab, tables, want * sizeof(float), cudaMemcpyHostToDevice, s->st));
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  return WL_OK;
}

int wl_set_accel_field(wl_sim* s, int which, const float* field, size_t count) {
  if (!s || which < 0 || which > 1) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  if (!field) {
    WL_CUDA(s, cudaStreamSynchronize(s->st));
    free_field(s->accf[0]); free_field(s->accf[1]);
    s->accf_on = false;
    return WL_OK;
  }
  for (int w = 0; w < 2; ++w)
    if (!s->accf[w].alloc) {
      char* arena = s->arena; s->arena = nullptr;      // plain allocation, never peer-mapped
      const int e = alloc_field(s, s->accf[w], s->lay, s->D);
      s->arena = arena;
      if (e) return e;
    }
  if (int e = xfer(s, WL_ACCEL_T0 + which, 0, const_cast<float*>(field), count, true)) return e;
  s->accf_on = true;
  return WL_OK;
}

int wl_measure_body(wl_sim* s, const wl_body* body, double t, double eps) {
  if (!s || !body) return WL_ERR_INVALID;
  if (body->kind == 0) return WL_OK;                 // NoBody: measure! is a no-op (Body.jl:83)
  if (body->kind < 0 || body->kind > 3) return fail(s, WL_ERR_UNSUPPORTED, "unknown body kind");
  return measure_impl(s, body_at(body, s->D, (float)t), eps, "wl_measure_body");
}

int wl_measure_csg(wl_sim* s, const wl_body_node* nodes, int32_t n, double eps) {
  if (!s) return WL_ERR_INVALID;
  BodyProg p;
  if (const char* m = prog_from(nodes, n, s->D, &p)) return fail(s, WL_ERR_INVALID, m);
  if (p.n == 1 && p.nd[0].kind == 0) return WL_OK;   // NoBody
  return measure_impl(s, p, eps, "wl_measure_csg");
}

int wl_set_body_coeffs(wl_sim* s, const float* mu0, const float* mu1, const float* V, const float* sdf) {
  if (!s) return WL_ERR_INVALID;
  const size_t per = (size_t)s->lay.n0 * s->lay.n1 * s->lay.n2;
  int e = 0;
  if (mu0 && (e = wl_upload(s, WL_MU0, 0, mu0, per * s->D))) return e;
  if (mu1 && (e = wl_upload(s, WL_MU1, 0, mu1, per * s->D * s->D))) return e;
  if (V && (e = wl_upload(s, WL_V, 0, V, per * s->D))) return e;
  if (sdf && (e = wl_upload(s, WL_SIGMA, 0, sdf, per))) return e;
  s->band_dirty = true;
  bc_vec(s, s->mu0, nullptr, false);
  bc_vec(s, s->V, nullptr, s->cfg.exitBC != 0);
  if (int e = halo(s, s->mu0, 1)) return e;
  if (int e = halo(s, s->V, 1)) return e;
  return check_async(s, "wl_set_body_coeffs");
}

int wl_build_poisson(wl_sim* s) {
  if (!s) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  if (int e = build_levels(s)) return e;
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  return check_async(s, "wl_build_poisson");
}

// update!(ml) (MultiLevelPoisson.jl:79-86)
int wl_update_poisson(wl_sim* s) {
  if (!s || s->lv.empty()) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  if (int e = set_diag_h(s, s->lv[0])) return e;
  for (size_t l = 1; l < s->lv.size(); ++l) {
    if (int e = restrictL(s, s->lv[l - 1], s->lv[l])) return e;
    if (int e = set_diag_h(s, s->lv[l])) return e;
  }
  return check_async(s, "wl_update_poisson");
}

int wl_solver(wl_sim* s, int32_t* n_iter, int32_t* trace_len, float* rinf, float* r1, float* omega) {
  if (!s || s->lv.empty()) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  Trace tr;
  if (int e = mlp_solver(s, &tr)) return e;
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  if (n_iter) *n_iter = tr.n;
  if (trace_len) *trace_len = tr.n + 1;
  for (int i = 0; i <= tr.n && i < 34; ++i) {
    if (rinf) rinf[i] = tr.rinf[i];
    if (r1) r1[i] = tr.r1[i];
    if (omega) omega[i] = tr.omega[i];
  }
  return check_async(s, "wl_solver");
}

int wl_poisson_solver(wl_sim* s, int32_t itmx, int32_t* n_iter) {
  if (!s || s->lv.empty() || itmx < 0) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  int n = 0;
  if (int e = poisson_solver(s, itmx, &n)) return e;
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  if (n_iter) *n_iter = n;
  return check_async(s, "wl_poisson_solver");
}

// pattern-tile classes of a level (TileFlags, wl_common.cuh): counts[c] = tiles of class c (0 general, 1 constants,
// 2 constants with boundary faces); all zero when the level has no classes (2D, rows not a multiple of 4, invalidated)
int wl_tile_classes(wl_sim* s, int32_t level, int64_t counts[3]) {
  if (!s || !counts || level < 0 || level >= (int)s->lv.size()) return WL_ERR_INVALID;
  counts[0] = counts[1] = counts[2] = 0;
  const Level& p = s->lv[level];
  if (!p.uvalid || !p.uflag) return WL_OK;
  cudaSetDevice(s->device);
  std::vector<unsigned char> h((size_t)p.unx * p.uny * p.unz);
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  WL_CUDA(s, cudaMemcpy(h.data(), p.uflag, h.size(), cudaMemcpyDeviceToHost));
  for (unsigned char c : h) counts[c == 3 ? 2 : (c < 3 ? c : 0)]++;          // class 3 (x end only) is a boundary class
  return WL_OK;
}

int wl_get_dt(const wl_sim* s, float* dt) { if (!s || !dt) return WL_ERR_INVALID; *dt = s->dt; return WL_OK; }
int wl_set_dt(wl_sim* s, float dt) { if (!s) return WL_ERR_INVALID; s->dt = dt; return WL_OK; }

int wl_cfl(wl_sim* s, float* dt) {
  if (!s || !dt) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  if (int e = cfl(s, dt)) return e;
  return check_async(s, "wl_cfl");
}

// mom_step!(a,b) (Flow.jl:156-167) = mom_predict! (:190-196), mom_project!(w=1), mom_correct! (:205-210),
// mom_project!(w=½), push!(Δt,CFL)
int wl_mom_step(wl_sim* s, const float* uBC_t1, const float* acc0, const float* acc1, wl_step_info* info) {
  if (!s || s->lv.empty()) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  const float* U = uBC_t1 ? uBC_t1 : s->cfg.uBC;
  const bool ex = s->cfg.exitBC != 0;
  Trace tr[2];
  const bool fused = s->D == 3 && !(s->cfg.flags & WL_FLAG_NO_FUSED);
  if (fused && s->band_dirty) { if (int e = band_rebuild(s)) return e; }
  if (s->timing) cudaEventRecord(s->ev[0], s->st);
  if (fused) {
    // predictor: u⁰ .= u is a buffer rotation; K-MOM writes the new u into the work buffer
    mom3(s, s->u, s->u, s->w, 0, acc0, s->accf_on ? s->accf[0].base : nullptr);
    DevField old0 = s->u0; s->u0 = s->u; s->u = s->w; s->w = old0;
  } else {
    WL_CUDA(s, cudaMemcpyAsync(s->u0.alloc, s->u.alloc, s->u.bytes, cudaMemcpyDeviceToDevice, s->st));   // u⁰ .= u
    ++s->launches;
    mom_rhs(s, s->u0, acc0, s->accf_on ? s->accf[0].base : nullptr);
    bdim(s, 0);
  }
  bc_vec(s, s->u, U, ex);
  if (ex) { if (int e = exitbc(s, s->u, s->u0, s->dt)) return e; }
  if (int e = halo(s, s->u, 1)) return e;              // K-DIVRES reads u one plane up
  if (s->timing) cudaEventRecord(s->ev[1], s->st);
  if (int e = mom_project(s, 1.f, U, &tr[0])) return e;
  if (int e = halo(s, s->u, 3)) return e;              // K-MOM: QUICK reaches 2 planes, the band's f-neighbours 3
  if (s->timing) cudaEventRecord(s->ev[2], s->st);
  // corrector
  ghost_phi(s);
  if (fused) {
    mom3(s, s->u, s->u0, s->w, 1, acc1, s->accf_on ? s->accf[1].base : nullptr);
    DevField t = s->u; s->u = s->w; s->w = t;
  } else {
    mom_rhs(s, s->u, acc1, s->accf_on ? s->accf[1].base : nullptr);
    bdim(s, 1);
  }
  bc_vec(s, s->u, U, ex);
  if (int e = halo(s, s->u, 1)) return e;
  if (s->timing) cudaEventRecord(s->ev[3], s->st);
  if (int e = mom_project(s, 0.5f, U, &tr[1])) return e;
  if (int e = halo(s, s->u, 3)) return e;
  if (s->timing) cudaEventRecord(s->ev[4], s->st);
  float dtn = 0.f;
  if (int e = cfl(s, &dtn)) return e;
  if (s->timing) {
    cudaEventRecord(s->ev[5], s->st);
    cudaEventSynchronize(s->ev[5]);
    for (int i = 0; i < 5; ++i) { float ms = 0; cudaEventElapsedTime(&ms, s->ev[i], s->ev[i + 1]); s->phase_s[i] = ms * 1e-3; }
  }
  s->dt = dtn;
  if (info) {
    info->dt_next = dtn; info->n_pred = tr[0].n; info->n_corr = tr[1].n;
    info->flags = tr[0].flags | tr[1].flags;
    for (int q = 0; q < 2; ++q) {
      info->trace_len[q] = tr[q].n + 1;
      for (int i = 0; i <= tr[q].n && i < 34; ++i) {
        info->rinf[q][i] = tr[q].rinf[i]; info->r1[q][i] = tr[q].r1[i]; info->omega[q][i] = tr[q].omega[i];
      }
    }
  }
  return check_async(s, "wl_mom_step");
}

extern "C++" {
static int sum_df(wl_sim* s, double out[3], const char* what) {
  for (int c = 0; c < s->D; ++c) {
    SumArgs sa{s->lay, s->f.base + s->lay.cstride * c, s->redw, s->d_red + 3 + c};
    DISPATCH_D(s, k_sum_all, grid_all(s->lay), blk(s->D), sa);
  }
  if (int e = fetch_red(s)) return e;
  for (int c = 0; c < 3; ++c) out[c] = c < s->D ? s->h_red[3 + c].sum : 0.0;
  return check_async(s, what);
}
template <class B>
static int pressure_force_impl(wl_sim* s, const B& body, double out[3], const char* what) {
  cudaSetDevice(s->device);
  ForceArgsT<B> a{s->lay, body, s->p.base, s->f.base, s->per.kz0};
  DISPATCH_DB(s, k_force_df, B, grid_all(s->lay), blk(s->D), a);
  return sum_df(s, out, what);
}
template <class B>
static int viscous_force_impl(wl_sim* s, const B& body, double out[3], const char* what) {
  cudaSetDevice(s->device);
  if (int e = halo(s, s->u, 1)) return e;               // the strain rate reads u across slab interfaces
  ViscArgsT<B> a{s->lay, body, s->u.base, s->f.base, s->cfg.nu, s->per.kz0};
  DISPATCH_DB(s, k_visc_df, B, grid_all(s->lay), blk(s->D), a);
  return sum_df(s, out, what);
}
template <class B>
static int force_moment_impl(wl_sim* s, const B& body, const double x0[3], int viscous, double out[3], const char* what) {
  cudaSetDevice(s->device);
  if (viscous) { if (int e = halo(s, s->u, 1)) return e; }
  MomentArgsT<B> a{s->lay, body, s->p.base, s->u.base, s->f.base, s->cfg.nu, s->per.kz0, viscous ? 1 : 0,
                   {(float)x0[0], (float)x0[1], (float)x0[2]}};
  DISPATCH_DB(s, k_moment_df, B, grid_all(s->lay), blk(s->D), a);
  return sum_df(s, out, what);
}
}  // extern "C++"

int wl_pressure_force(wl_sim* s, const wl_body* body, double t, double out[3]) {
  if (!s || !body || !out) return WL_ERR_INVALID;
  return pressure_force_impl(s, body_at(body, s->D, (float)t), out, "wl_pressure_force");
}
// nds weights of a host-measured body (src/Metrics.jl:110-113,126-133): idx = linear index into the reference's
// (n0,n1[,n2]) scalar array (first dimension fastest, 0-based), w = D floats per entry
int wl_set_force_weights(wl_sim* s, int64_t nnz, const int64_t* idx, const float* w) {
  if (!s || nnz < 0 || (nnz > 0 && (!idx || !w))) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  if (s->fw_off) cudaFree(s->fw_off);
  if (s->fw_w) cudaFree(s->fw_w);
  s->fw_off = nullptr; s->fw_w = nullptr; s->fw_n = -1;
  if (nnz > 0x7fffffff) return fail(s, WL_ERR_INVALID, "too many force weights");
  const Lay& l = s->lay;
  std::vector<int> off((size_t)nnz);
  const long long per = (long long)l.n0 * l.n1 * l.n2;
  for (int64_t q = 0; q < nnz; ++q) {
    const long long g = idx[q];
    if (g < 0 || g >= per) return fail(s, WL_ERR_INVALID, "force weight index outside the scalar field");
    const int i = (int)(g % l.n0), j = (int)((g / l.n0) % l.n1), k = (int)(g / ((long long)l.n0 * l.n1));
    off[q] = l.off(i, j, k);
  }
  if (nnz > 0) {
    WL_CUDA(s, cudaMalloc(&s->fw_off, sizeof(int) * (size_t)nnz));
    WL_CUDA(s, cudaMalloc(&s->fw_w, sizeof(float) * (size_t)nnz * s->D));
    WL_CUDA(s, cudaMemcpy(s->fw_off, off.data(), sizeof(int) * (size_t)nnz, cudaMemcpyHostToDevice));
    WL_CUDA(s, cudaMemcpy(s->fw_w, w, sizeof(float) * (size_t)nnz * s->D, cudaMemcpyHostToDevice));
  }
  s->fw_n = (int)nnz;
  return WL_OK;
}
int wl_pressure_force_weights(wl_sim* s, double out[3]) {
  if (!s || !out) return WL_ERR_INVALID;
  if (s->fw_n < 0) return fail(s, WL_ERR_INVALID, "wl_set_force_weights has not been called");
  cudaSetDevice(s->device);
  const int nb = std::max(1, std::min((s->fw_n + 255) / 256, 592));
  for (int c = 0; c < s->D; ++c) {
    ForceWArgs a{s->p.base, s->fw_off, s->fw_w, s->fw_n, s->D, c, s->redw, s->d_red + 3 + c};
    LAUNCH(s, k_force_w, dim3(nb), dim3(256), a);
  }
  if (int e = fetch_red(s)) return e;
  for (int c = 0; c < 3; ++c) out[c] = c < s->D ? s->h_red[3 + c].sum : 0.0;
  return check_async(s, "wl_pressure_force_weights");
}
int wl_viscous_force(wl_sim* s, const wl_body* body, double t, double out[3]) {
  if (!s || !body || !out) return WL_ERR_INVALID;
  return viscous_force_impl(s, body_at(body, s->D, (float)t), out, "wl_viscous_force");
}
int wl_force_moment(wl_sim* s, const wl_body* body, double t, const double x0[3], int viscous, double out[3]) {
  if (!s || !body || !out || !x0) return WL_ERR_INVALID;
  return force_moment_impl(s, body_at(body, s->D, (float)t), x0, viscous, out, "wl_force_moment");
}
int wl_pressure_force_csg(wl_sim* s, const wl_body_node* nodes, int32_t n, double out[3]) {
  if (!s || !out) return WL_ERR_INVALID;
  BodyProg p;
  if (const char* m = prog_from(nodes, n, s->D, &p)) return fail(s, WL_ERR_INVALID, m);
  return pressure_force_impl(s, p, out, "wl_pressure_force_csg");
}
int wl_viscous_force_csg(wl_sim* s, const wl_body_node* nodes, int32_t n, double out[3]) {
  if (!s || !out) return WL_ERR_INVALID;
  BodyProg p;
  if (const char* m = prog_from(nodes, n, s->D, &p)) return fail(s, WL_ERR_INVALID, m);
  return viscous_force_impl(s, p, out, "wl_viscous_force_csg");
}
int wl_force_moment_csg(wl_sim* s, const wl_body_node* nodes, int32_t n, const double x0[3], int viscous, double out[3]) {
  if (!s || !out || !x0) return WL_ERR_INVALID;
  BodyProg p;
  if (const char* m = prog_from(nodes, n, s->D, &p)) return fail(s, WL_ERR_INVALID, m);
  return force_moment_impl(s, p, x0, viscous, out, "wl_force_moment_csg");
}

int wl_meanflow_init(wl_sim* s, int uu_stats) {
  if (!s) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  char* arena = s->arena; s->arena = nullptr;       // private to the rank: plain allocations, never peer-mapped
  int e = WL_OK;
  free_field(s->mf_P); free_field(s->mf_U); free_field(s->mf_UU);
  if (!(e = alloc_field(s, s->mf_P, s->lay, 1)) && !(e = alloc_field(s, s->mf_U, s->lay, s->D)) && uu_stats)
    e = alloc_field(s, s->mf_UU, s->lay, s->D * s->D);
  s->arena = arena;
  if (e) return e;
  return check_async(s, "wl_meanflow_init");
}
int wl_meanflow_update(wl_sim* s, float eps) {
  if (!s) return WL_ERR_INVALID;
  if (!s->mf_P.alloc) return fail(s, WL_ERR_INVALID, "wl_meanflow_init has not been called");
  cudaSetDevice(s->device);
  MeanArgs a{s->lay, s->p.base, s->u.base, s->mf_P.base, s->mf_U.base, s->mf_UU.alloc ? s->mf_UU.base : nullptr, eps};
  DISPATCH_D(s, k_mean_update, grid_all(s->lay), blk(s->D), a);
  return check_async(s, "wl_meanflow_update");
}
int wl_meanflow_copy_to_flow(wl_sim* s) {
  if (!s) return WL_ERR_INVALID;
  if (!s->mf_P.alloc) return fail(s, WL_ERR_INVALID, "wl_meanflow_init has not been called");
  cudaSetDevice(s->device);
  WL_CUDA(s, cudaMemcpyAsync(s->u.alloc, s->mf_U.alloc, s->u.bytes, cudaMemcpyDeviceToDevice, s->st));   // a.u .= b.U
  WL_CUDA(s, cudaMemcpyAsync(s->p.alloc, s->mf_P.alloc, s->p.bytes, cudaMemcpyDeviceToDevice, s->st));   // a.p .= b.P
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  return check_async(s, "wl_meanflow_copy_to_flow");
}

int wl_debug_op(wl_sim* s, int op, int level, float arg) {
  if (!s) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  if (op >= 2 && op <= 6 && (s->lv.empty() || level < 0 || level >= (int)s->lv.size())) return WL_ERR_INVALID;
  const bool ex = s->cfg.exitBC != 0;
  switch (op) {
    case 0: bc_vec(s, s->u, s->cfg.uBC, ex); break;
    case 1: mom_rhs(s, s->u0, nullptr); bdim(s, 0); break;
    case 2: { Trace tr; if (int e = mom_project(s, arg, s->cfg.uBC, &tr)) return e; break; }
    case 3: {
      Level& p = s->lv[level];
      perbc(s, p.lay, p.x.base, &p.per);
      if (p.dist) { if (int e = halo(s, p.x, 1)) return e; }
      LevArgs a = lev_args(s, p);
      a.out = s->d_red + 0;
      DISPATCH_D(s, k_residual, grid_inside(p.lay), blk(s->D), a);
      if (int e = gather_red(s)) return e;
      a.in = red_in(s, 0); a.out = s->d_red + 1;
      DISPATCH_D(s, k_shift_norms, grid_tiles(p.lay, false), blk(s->D), a);
      break;
    }
    case 4: vcycle(s, 0, arg); break;
    case 5: smooth(s, s->lv[level], arg, nullptr); break;
    case 6: jacobi(s, s->lv[level]); break;
    case 7: exitbc(s, s->u, s->u0, s->dt); break;
    case 8: ghost_phi(s); mom_rhs(s, s->u, nullptr); bdim(s, 1); break;
    case 9: cudaMemcpyAsync(s->u0.alloc, s->u.alloc, s->u.bytes, cudaMemcpyDeviceToDevice, s->st); break;
    case 10: if (s->lv.empty()) return WL_ERR_INVALID; if (int e = pcg(s, s->lv[0], (int)arg)) return e; break;
    default: return fail(s, WL_ERR_INVALID, "unknown debug op");
  }
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  return check_async(s, "wl_debug_op");
}

int wl_prof_begin(wl_sim* s, const char* filter) {
  if (!s) return WL_ERR_INVALID;
  s->prof_on = true;
  s->prof_split = filter && filter[0] == '@';                  // "@…": one line per (kernel, grid size), i.e. per multigrid level
  s->prof_filter = filter ? filter + (s->prof_split ? 1 : 0) : "";
  return WL_OK;
}
// Writes one line per kernel: "<name>\t<launches>\t<total_ms>\t<total_blocks>\n"; returns the number of kernels.
int wl_prof_end(wl_sim* s, char* buf, size_t buflen) {
  if (!s) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  cudaStreamSynchronize(s->st);
  s->prof_on = false;
  struct Acc { const char* name; long n; double ms; long long blocks; };
  std::vector<Acc> acc;
  for (auto& r : s->recs) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, r.a, r.b);
    s->evpool.push_back(r.a); s->evpool.push_back(r.b);
    size_t i = 0;
    for (; i < acc.size(); ++i)
      if ((acc[i].name == r.name || !strcmp(acc[i].name, r.name)) && (!s->prof_split || acc[i].blocks == r.blocks * acc[i].n)) break;
    if (i == acc.size()) acc.push_back({r.name, 0, 0.0, 0});
    acc[i].n++; acc[i].ms += ms; acc[i].blocks += r.blocks;
  }
  s->recs.clear();
  std::string out;
  char line[256];
  for (auto& a : acc) {
    if (s->prof_split) snprintf(line, sizeof line, "%s@%lld\t%ld\t%.6f\t%lld\n", a.name, a.blocks / a.n, a.n, a.ms, a.blocks);
    else snprintf(line, sizeof line, "%s\t%ld\t%.6f\t%lld\n", a.name, a.n, a.ms, a.blocks);
    out += line;
  }
  if (buf && buflen) { strncpy(buf, out.c_str(), buflen - 1); buf[buflen - 1] = 0; }
  return (int)acc.size();
}
// CUDA events on the sim's own stream (torch.cuda.Event only sees torch's stream)
int wl_event_record(wl_sim* s, int slot) {
  if (!s || slot < 0 || slot >= 8) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  if (!s->uev[slot]) WL_CUDA(s, cudaEventCreate(&s->uev[slot]));
  WL_CUDA(s, cudaEventRecord(s->uev[slot], s->st));
  return WL_OK;
}
int wl_event_elapsed_ms(wl_sim* s, int a, int b, float* ms) {
  if (!s || !ms || a < 0 || b < 0 || a >= 8 || b >= 8 || !s->uev[a] || !s->uev[b]) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  WL_CUDA(s, cudaEventSynchronize(s->uev[b]));
  WL_CUDA(s, cudaEventElapsedTime(ms, s->uev[a], s->uev[b]));
  return WL_OK;
}

int wl_sync(wl_sim* s) {
  if (!s) return WL_ERR_INVALID;
  cudaSetDevice(s->device);
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  if (s->p2p) {                                   // a halo kernel that gave up waiting for its neighbour
    unsigned int err = 0;
    WL_CUDA(s, cudaMemcpy(&err, s->arena + 5 * sizeof(unsigned int), sizeof(err), cudaMemcpyDeviceToHost));
    if (err) return fail(s, WL_ERR_CUDA, "halo exchange timed out waiting for a neighbouring rank");
  }
  return WL_OK;
}
int64_t wl_launch_count(const wl_sim* s) { return s ? s->launches : 0; }
int wl_enable_timing(wl_sim* s, int on) { if (!s) return WL_ERR_INVALID; s->timing = on != 0; return WL_OK; }
int wl_last_phase_times(const wl_sim* s, double out[8]) {
  if (!s || !out) return WL_ERR_INVALID;
  for (int i = 0; i < 8; ++i) out[i] = s->phase_s[i];
  return WL_OK;
}

}  // extern "C"
