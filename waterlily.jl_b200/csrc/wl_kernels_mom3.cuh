// wl_kernels_mom3.cuh -- K-MOM: the fused 3D momentum predictor/corrector.
//
// One kernel replaces, per phase, `conv_diff!` (1+9·4 launches), `accelerate!`, both loops of
// `BDIM!` and `scale_u!` (reference src/Flow.jl:38-62,69-73,176-180,211-214).  The flux scratch Φ
// and the force field f never touch HBM: the kernel reads the advecting velocity (+u⁰, V, μ₀) and
// writes the new velocity into a third rotating buffer (u⁰ .= u becomes a pointer swap).
//
// Decomposition: a block owns a 31×7 column of cells and marches KZ planes up the last
// dimension.  Every thread evaluates the 9 LOWER-face fluxes Φ[ci][dj] of its cell once; the
// upper-face fluxes come from the neighbour that owns them:
//    x: lane+1 by warp shuffle (warps overlap by one cell: 32 faces -> 31 cells),
//    y: row+1 through a double-buffered shared-memory plane (one __syncthreads per plane),
//    z: the next plane of the march (the cell is finalised one plane later, from registers).
// The sum r = Σ_j (+Φ_low −Φ_up) is accumulated in exactly the reference's order, so results are
// bit-identical to the two-launch scatter of Flow.jl:47-49.  Domain-boundary faces use the
// ϕuL/ϕuR/ϕuP variants (Flow.jl:9-11,56-62); periodic wrap values come straight from global memory.
// Blocks that touch no x/y domain boundary run a branch-free FAST body (compile-time face kinds,
// immediate x offsets, no index clamps); edge blocks run the general body.
//
// BDIM: the main kernel assumes μ₁ ≡ 0 (then μddn ≡ 0 exactly) and never reads μ₁ (36 of the 84
// algorithmic B/cell); the few cells in the body band are recomputed afterwards by k_mom_patch
// from the exact reference formula, f-neighbours on ghost rows included (Flow.jl:20-26,179).
#pragma once
#include <cuda.h>          // CUtensorMap (type only: the encoder is fetched through cudaGetDriverEntryPoint, no libcuda link)

#include <type_traits>

#include "wl_kernels_mom.cuh"

namespace wl {

// x/6 correctly rounded in 3 instructions (Markstein: q0 = x·RN(1/6), r = fma(-6,q0,x) exact,
// q = fma(r,RN(1/6),q0)); verified bit-identical to IEEE x/6.0f for EVERY float with
// 1e-30 < |x| < 1e38 (exhaustive 2^32 sweep); outside that range fall back to the IEEE division.
__device__ __forceinline__ float div6(float x) {
  const float ax = fabsf(x);
  if (ax > 1e-30f && ax < 1e38f) {
    const float y = 1.0f / 6.0f;
    const float q0 = __fmul_rn(x, y);
    const float r = __fmaf_rn(-6.0f, q0, x);
    return __fmaf_rn(r, y, q0);
  }
  if (x == 0.f) return x;
  return x / 6.f;
}
// median of three by min/max: same value as Flow.jl:27-36 for all non-NaN inputs, 4 instructions
__device__ __forceinline__ float med3(float a, float b, float c) { return fmaxf(fminf(a, b), fminf(fmaxf(a, b), c)); }
template <int S>
__device__ __forceinline__ float lamf(float u, float c, float d) {
  if (S == 0) return med3(div6(5.f * c + 2.f * d - u), c, med3(10.f * c - 9.f * u, c, d));
  if (S == 1) return (c + d) / 2.f;
  return (c <= fminf(u, d) || c >= fmaxf(u, d)) ? c : c + (d - c) * (c - u) / (d - u);
}
// kind: 0 inner (ϕu), 1 lower non-periodic (ϕuL), 2 lower periodic (ϕuP; caller passes the wrapped
// far-upstream value as um2), 3 upper non-periodic (ϕuR).
template <int S>
__device__ __forceinline__ float flux_v(float um2, float um1, float uc, float up1, float ua, float nu, int kind) {
  const bool pos = ua > 0.f;
  float v = lamf<S>(pos ? um2 : up1, pos ? um1 : uc, pos ? uc : um1);
  if ((kind == 1 && pos) || (kind == 3 && ua < 0.f)) v = (uc + um1) / 2.f;
  return ua * v - nu * (uc - um1);
}

// Branch-free x/6 for the marching kernel: the Markstein quotient of div6 for every x, while the range of
// |x| seen is recorded: with b = bits(x)<<1 (sign dropped), bmax = largest b, bmin = smallest b-1 as unsigned, so
// an exact zero wraps to 0xffffffff and never counts as "too small" (early in a run whole regions hold
// u_y = u_z = 0).  One integer add-min and one integer max per face, no predicate.  The caller re-evaluates
// the plane's fluxes with div6 when any non-zero |x| fell outside (1e-30, 1e38), so the nine face fluxes of
// a cell form one straight-line block that the scheduler can interleave.  For x = ±0 the quotient is +0
// where IEEE gives ±0: the sign of a zero flux cannot reach a non-zero value (no division by a velocity).
__device__ __forceinline__ float div6_track(float x, unsigned& bmin, unsigned& bmax) {
  const unsigned b = __float_as_uint(x) << 1;      // |x| without the sign bit, order-preserving
  bmax = max(bmax, b);
  bmin = min(bmin, b - 1u);
  const float y = 1.0f / 6.0f;
  const float q0 = __fmul_rn(x, y);
  const float r = __fmaf_rn(-6.0f, q0, x);
  return __fmaf_rn(r, y, q0);
}
// bit patterns of the bounds of div6's verified range
constexpr unsigned DIV6_LO = 0x0DA24260u << 1;   // 1e-30f
constexpr unsigned DIV6_HI = 0x7E967699u << 1;   // 1e38f
template <int S, bool EXACT>
__device__ __forceinline__ float flux_q(float um2, float um1, float uc, float up1, float ua, float nu, int kind,
                                        unsigned& amin, unsigned& amax) {
  if (S != 0 || EXACT) return flux_v<S>(um2, um1, uc, up1, ua, nu, kind);
  const bool pos = ua > 0.f;
  const float u = pos ? um2 : up1, c = pos ? um1 : uc, d = pos ? uc : um1;
  float v = med3(div6_track(5.f * c + 2.f * d - u, amin, amax), c, med3(10.f * c - 9.f * u, c, d));
  if ((kind == 1 && pos) || (kind == 3 && ua < 0.f)) v = (uc + um1) / 2.f;
  return ua * v - nu * (uc - um1);
}

struct Mom3Args {
  Lay l;
  const float* ua;    // advecting velocity: u⁰ (predictor) / projected u (corrector)
  const float* u0;    // u⁰
  const float* V;
  const float* mu0;
  float* un;          // output velocity buffer (distinct from ua and u0)
  // the same arrays per component (base + cstride·c), so every address is constant-bank base + offset
  const float* uac[3]; const float* u0c[3]; const float* Vc[3]; const float* m0c[3]; float* unc[3];
  Per per;
  float nu, dt;
  Vec3 acc;
  const float* accfc[3];  // accelerate! field g(i,x,t)+∂ₜuBC(i,x,t) per component (wl_set_accel_field): read by the ACCF variant
  int corrector;
  int KZ;
  // TMA (cp.async.bulk.tensor): 4-D maps (x+3, y, z+zpad, component) over the padded device arrays of the advecting
  // velocity, u⁰, V and μ₀; blocks whose tile lies inside the array take their planes with ONE bulk copy per operand
  int use_tma;
  alignas(64) CUtensorMap tm_ua, tm_u0, tm_V, tm_m0;
};

constexpr int M3_TY = 8;

// ---- asynchronous plane staging (cp.async / LDGSTS): the in-plane neighbourhood of the marching
// plane is served from a 4-slot shared-memory ring filled two planes ahead, so the flux
// evaluation never waits on a global load (a register-only body stalls ~6.6 cycles per issue on the
// long scoreboard with 16 warps/SM resident).
__device__ __forceinline__ void cp_async4(float* smem, const float* g) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

constexpr int M3_X0 = 4;                        // tile column of cell i0.  The stencil needs columns i0-2 .. i0+33; the tile starts two
                                                // columns earlier because a TMA box must start on a 16-byte boundary of its row (measured on
                                                // B200: an innermost coordinate that is not a multiple of 4 floats raises "illegal instruction",
                                                // scripts/r2/tma_probe.cu), and cell i0-4 = 32·bx-3 is the nearest such column (layout: cell 1
                                                // of every row is 128-byte aligned)
constexpr int M3_TW = 40;                       // tile row pitch: columns i0-4 .. i0+35
constexpr int M3_TH = M3_TY + 3;                // tile rows      j0-2 .. j0+TY
constexpr int M3_TILE = M3_TW * M3_TH;          // floats per component per plane
constexpr int M3_RING = 4;
constexpr int M3_SLOT = ((3 * M3_TILE * 4 + 127) / 128) * 32;   // floats per ring slot: [3][M3_TILE] padded to 128 bytes (TMA)
constexpr int M3_FIN = 32 * (M3_TY - 1);        // own-cell operands of one component and plane: the block's output rows

struct alignas(128) Mom3Smem {
  float ring[M3_RING][M3_SLOT];    // slot = plane & 3; component c of the tile at c * M3_TILE
  float fin[3][9][M3_FIN];         // u⁰, V, μ₀ of the thread's own cell, staged two planes ahead (slot = plane % 3)
  float shy[2][3][M3_TY][32];
  float shx[2][3][M3_TY];          // W32: x-face fluxes of the column right of the block, per row
  unsigned long long mbar[M3_RING];   // TMA: one transaction barrier per ring slot
};

// ---- TMA + mbarrier primitives (sm_90+; SASS: UTMALDG / SYNCS)
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned mb, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(mb), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(unsigned mb, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(mb), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned mb, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WL_MBAR_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WL_MBAR_DONE;\n"
      "bra WL_MBAR_WAIT;\n"
      "WL_MBAR_DONE:\n"
      "}\n" ::"r"(mb), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_4d(unsigned dst, const CUtensorMap* tm, int c0, int c1, int c2, int c3, unsigned mb) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];\n"
               ::"r"(dst), "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(mb) : "memory");
}

// NB ("no body"): the host knows V ≡ 0 and μ₀ ≡ 1 on every cell (no body was ever measured or uploaded
// and every direction is periodic, so BC!(μ₀,0) left no zero faces): then f - V = f and V + μ₀·f = f
// exactly, and the kernel skips those two fields (24 of its 84 B/cell and their address arithmetic).
// W32: the block owns 32 columns instead of 31.  Lane 31 has no lane+1 to take its upper x face from; the
// block's last row (a whole warp that only supplies y faces, six of its nine fluxes are unused) evaluates
// those 3x7 faces at column i0+32 -- the same expression on the same staged values as lane 0 of the next
// block -- and hands them over through shared memory.  Power-of-two grids then tile exactly (128 = 4x32
// instead of 5x31 columns) and every lane of the seven output rows produces a cell.
template <int S, bool FAST, bool NB, bool W32, bool ACCF>
__device__ __forceinline__ void mom3_body_async(const Mom3Args& a, Mom3Smem& sm) {
  const Lay& l = a.l;
  const int lane = threadIdx.x, ty = threadIdx.y, tid = lane + 32 * ty;
  const int N0 = l.n0, N1 = l.n1, N2 = l.n2;
  const int pitch = l.pitch, plane = l.plane;
  constexpr int BW = W32 ? 32 : 31;
  const int i0 = 1 + BW * blockIdx.x, j0 = 1 + (M3_TY - 1) * blockIdx.y;
  const int i_raw = i0 + lane, j_raw = j0 + ty;
  const int i = FAST ? i_raw : min(i_raw, N0 - 1), j = FAST ? j_raw : min(j_raw, N1 - 1);
  const int k0 = 1 + a.KZ * blockIdx.z;
  const int kend = min(k0 + a.KZ, N2 - 1);
  const bool out_ij = lane < BW && ty < M3_TY - 1 && (FAST || (i_raw <= N0 - 2 && j_raw <= N1 - 2));
  const int oij = i + pitch * j;
  const bool p0 = a.per.p[0] != 0, p1 = a.per.p[1] != 0, p2 = a.per.p[2] != 0;
  const int klo = a.per.zlo ? -1 : 0, khi = a.per.zhi ? N2 : N2 - 1;   // readable plane range (slab halos)
  // Periodic directions have no special faces here: the ghost layer holds exact images (BC! ran on the
  // advecting field), and the one value QUICK needs beyond it -- CIj(j,·,N[j]-2) of ϕuP, Flow.jl:10,61 -- is
  // fetched from its wrapped position when the tile / the column is staged.  The face on the upper ghost
  // is then the same expression on the same values as Φ of the wrapped cell (Flow.jl:57-59).
  const int kx = (FAST || p0) ? 0 : (i == 1 ? 1 : (i == N0 - 1 ? 3 : 0));
  const int ky = (FAST || p1) ? 0 : (j == 1 ? 1 : (j == N1 - 1 ? 3 : 0));
  const float* const* U = a.uac;
  const float nu = a.nu;
  auto fold = [](int g, int N, bool per) {        // array index of tile coordinate g
    if (per) g = g < 0 ? g + (N - 2) : (g > N - 1 ? g - (N - 2) : g);
    return min(max(g, 0), N - 1);
  };
  // staging duty of this thread: tile positions tid and tid+256
  const int pa = tid, pb = tid + 32 * M3_TY;
  const bool has_b = pb < M3_TILE;
  const int ga = fold(i0 - M3_X0 + pa % M3_TW, N0, p0) + pitch * fold(j0 - 2 + pa / M3_TW, N1, p1);
  const int gb = has_b ? fold(i0 - M3_X0 + pb % M3_TW, N0, p0) + pitch * fold(j0 - 2 + pb / M3_TW, N1, p1) : 0;
  // TMA: the block's tile lies inside the array (every FAST block of a non-periodic direction; periodic ones except
  // those whose halo columns/rows wrap), so one elected thread fetches a plane of the three components with ONE bulk
  // tensor copy, and the own-cell operands of the finalisation with one more per field; completion is counted on the
  // ring slot's mbarrier.  Other blocks keep the per-thread cp.async staging below.
  // (columns of the box left of cell 0 or right of the row's pitch are never used: they hold the neighbouring row's
  //  padding or the zero fill of an out-of-bounds box; the columns the stencil reads, i0-2 .. i0+33, must be real cells)
  const bool tma = FAST && a.use_tma && i0 - 2 >= 0 && i0 + 33 <= N0 - 1 && i0 + 33 + 3 <= pitch - 1 &&
                   j0 - 2 >= 0 && j0 + M3_TY <= N1 - 1;
  auto plane_of = [&](int kp) { return p2 ? (kp > N2 - 1 ? min(kp - (N2 - 2), N2 - 1) : kp) : min(kp, khi); };
  // ring plane kp (+ the finalisation operands of plane kf, or kf < 0) on the mbarrier of slot kp & 3
  auto stage_tma = [&](int kp, int kf) {
    const unsigned mb = smem_u32(&sm.mbar[kp & (M3_RING - 1)]);
    constexpr unsigned ring_bytes = 3 * M3_TILE * 4, fin_bytes = 3 * M3_FIN * 4;
    mbar_expect_tx(mb, ring_bytes + (kf >= 0 ? (NB ? 1u : 3u) * fin_bytes : 0u));
    tma_load_4d(smem_u32(sm.ring[kp & (M3_RING - 1)]), &a.tm_ua, i0 - M3_X0 + 3, j0 - 2, plane_of(kp) + l.zpad, 0, mb);
    if (kf >= 0) {
      float* dst = sm.fin[kf % 3][0];
      tma_load_4d(smem_u32(dst), &a.tm_u0, i0 + 3, j0, kf + l.zpad, 0, mb);
      if (!NB) {
        tma_load_4d(smem_u32(dst + 3 * M3_FIN), &a.tm_V, i0 + 3, j0, kf + l.zpad, 0, mb);
        tma_load_4d(smem_u32(dst + 6 * M3_FIN), &a.tm_m0, i0 + 3, j0, kf + l.zpad, 0, mb);
      }
    }
  };
  unsigned phase = 0;                             // parity bit per ring slot
  auto wait_tma = [&](int kp) {
    const int sl = kp & (M3_RING - 1);
    mbar_wait(smem_u32(&sm.mbar[sl]), (phase >> sl) & 1u);
    phase ^= 1u << sl;
  };
  auto stage = [&](int kp) {
    const int kc = plane_of(kp);
    float(*slot)[M3_TILE] = reinterpret_cast<float(*)[M3_TILE]>(sm.ring[kp & (M3_RING - 1)]);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      cp_async4(&slot[c][pa], U[c] + ga + plane * kc);
      if (has_b) cp_async4(&slot[c][pb], U[c] + gb + plane * kc);
    }
  };
  const int tc = (ty + 2) * M3_TW + lane + M3_X0;     // own position in the tile
  // Operands of the finalisation (u⁰, V, μ₀ of the own cell): ncu showed 9 % of the kernel's stall samples on
  // the first use of these loads although they were issued a whole plane of flux arithmetic earlier.  They
  // travel asynchronously too, two planes ahead and without holding registers (cp.async path: each thread reads
  // back what it staged itself, so its own wait_group is the only synchronisation).
  const bool fin_thread = out_ij && ty < M3_TY - 1;
  auto stage_fin = [&](int kp) {
    if (!fin_thread) return;
    const int o = oij + plane * kp;
    float* dst = &sm.fin[kp % 3][0][tid];
#pragma unroll
    for (int ci = 0; ci < 3; ++ci) {
      cp_async4(dst + ci * M3_FIN, a.u0c[ci] + o);
      if (!NB) {
        cp_async4(dst + (3 + ci) * M3_FIN, a.Vc[ci] + o);
        cp_async4(dst + (6 + ci) * M3_FIN, a.m0c[ci] + o);
      }
    }
  };
  // W32: lane = 7·c + r of the last row evaluates the x face of component c at column i0+32, row j0+r
  const bool xact = W32 && ty == M3_TY - 1 && lane < 3 * (M3_TY - 1);
  const int xr = lane % (M3_TY - 1), xc = min(lane / (M3_TY - 1), 2);
  const int pX = (xr + 2) * M3_TW + M3_X0 + 32;          // tile position of (i0+32, j0+xr)
  const int pA = xc == 0 ? pX - 1 : pX - M3_TW;   // the other operand of the advecting mean (xc == 2: previous plane)
  const int kxX = (FAST || p0) ? 0 : (i0 + 32 == N0 - 1 ? 3 : 0);
  float xprev = 0.f;                              // u_x(i0+32, j0+xr) one plane below

  if (tma) {
    if (tid == 0) {
#pragma unroll
      for (int q = 0; q < M3_RING; ++q) mbar_init(smem_u32(&sm.mbar[q]), 1);
      mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0) { stage_tma(k0, k0 <= kend - 1 ? k0 : -1); stage_tma(k0 + 1, -1); stage_tma(k0 + 2, -1); }
  } else {
    stage(k0); stage(k0 + 1);
    if (k0 <= kend - 1) stage_fin(k0);
    cp_async_commit();
    stage(k0 + 2); cp_async_commit();
  }
  float w[3][4];   // own column: planes k-2, k-1, k, k+1
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    w[c][0] = U[c][oij + plane * ((p2 && k0 == 1) ? N2 - 3 : max(k0 - 2, klo))];
    w[c][1] = U[c][oij + plane * (k0 - 1)];
  }
  if (xact) xprev = U[0][fold(i0 + 32, N0, p0) + pitch * fold(j0 + xr, N1, p1) + plane * (k0 - 1)];
  if (tma) { wait_tma(k0); wait_tma(k0 + 1); } else cp_async_wait<1>();
  __syncthreads();
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    w[c][2] = sm.ring[k0 & (M3_RING - 1)][c * M3_TILE + tc];
    w[c][3] = sm.ring[(k0 + 1) & (M3_RING - 1)][c * M3_TILE + tc];
  }
  float pend[3] = {0.f, 0.f, 0.f};

  // one plane of the march.  ZIN: an inner plane (ordinary z faces, not the last visited plane) -- with
  // FAST that body has no data-dependent branch at all.
  // TOP: the last row of the block only supplies the upper y faces of the row below, so of its nine
  // fluxes only the three y-face ones are evaluated (the row is one whole warp: no divergence).
  auto plane_step = [&](auto zin_tag, auto top_tag, const int k) {
    constexpr bool ZIN = decltype(zin_tag)::value;
    constexpr bool TOP = decltype(top_tag)::value;
    const bool lastp = ZIN ? false : k == kend;
    const int kz = (ZIN || p2) ? 0 : (k == 1 ? (a.per.zlo ? 0 : 1) : (k == N2 - 1 ? (a.per.zhi ? 0 : 3) : 0));
    if (!lastp) {
      if (tma) { if (tid == 0) stage_tma(k + 3, k + 1 <= kend - 1 ? k + 1 : -1); }
      else { stage(k + 3); cp_async_commit(); }
    }
    // operands of the finalisation of cell k-1: staged two planes ago, consumed after the fluxes
    float fu0[3], fV[3], fm0[3];
    const int opv = oij + plane * (k - 1);
    const bool fin = !TOP && k > k0 && out_ij;
    float(*cur)[M3_TILE] = reinterpret_cast<float(*)[M3_TILE]>(sm.ring[k & (M3_RING - 1)]);
    float nb[3][3][3];   // nb[c][d][m2,m1,p1]
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      nb[c][2][0] = w[c][0]; nb[c][2][1] = w[c][1]; nb[c][2][2] = w[c][3];
      nb[c][0][0] = cur[c][tc - 2]; nb[c][0][1] = cur[c][tc - 1]; nb[c][0][2] = cur[c][tc + 1];
      nb[c][1][0] = cur[c][tc - 2 * M3_TW]; nb[c][1][1] = cur[c][tc - M3_TW]; nb[c][1][2] = cur[c][tc + M3_TW];
    }
    float F[3][3];
    auto fluxes = [&](auto exact_tag) -> bool {
      constexpr bool EX = decltype(exact_tag)::value;
      unsigned amin = 0xffffffffu, amax = 0u;
#pragma unroll
      for (int ci = 0; ci < 3; ++ci) {
#pragma unroll
        for (int dj = 0; dj < 3; ++dj) {
          if ((lastp && dj < 2) || (TOP && dj != 1)) { F[ci][dj] = 0.f; continue; }
          const int kind = dj == 0 ? kx : (dj == 1 ? ky : kz);
          const float ua = (w[dj][2] + nb[dj][ci][1]) / 2.f;    // ϕ(i,CI(I,j),u) Flow.jl:3
          F[ci][dj] = flux_q<S, EX>(nb[ci][dj][0], nb[ci][dj][1], w[ci][2], nb[ci][dj][2], ua, nu, kind, amin, amax);
        }
      }
      return !(amin >= DIV6_LO && amax < DIV6_HI);       // bits-1 >= LO  <=>  |x| > 1e-30 (or x == 0)
    };
    if (fluxes(std::false_type{})) fluxes(std::true_type{});
    float Fxu[3], Fyu[3];
    if (!lastp) {
      const int buf = k & 1;
      if (W32 && ty == M3_TY - 1) {                          // one warp: the faces right of the block
        if (xact) {
          const float* cx = cur[xc];
          const float ux0 = cur[0][pX];
          const float ux1 = xc == 2 ? xprev : cur[0][pA];
          const float uaX = (ux0 + ux1) / 2.f;
          sm.shx[buf][xc][xr] = flux_v<S>(cx[pX - 2], cx[pX - 1], cx[pX], cx[pX + 1], uaX, nu, kxX);
          xprev = ux0;
        }
      }
#pragma unroll
      for (int ci = 0; ci < 3; ++ci) {
        Fxu[ci] = __shfl_down_sync(0xffffffffu, F[ci][0], 1);
        sm.shy[buf][ci][ty][lane] = F[ci][1];
      }
      if (tma) wait_tma(k + 2); else cp_async_wait<1>();     // plane k+2 has landed (k+3 may be in flight)
      __syncthreads();
#pragma unroll
      for (int ci = 0; ci < 3; ++ci) Fyu[ci] = sm.shy[buf][ci][min(ty + 1, M3_TY - 1)][lane];
      if (W32 && lane == 31) {
#pragma unroll
        for (int ci = 0; ci < 3; ++ci) Fxu[ci] = sm.shx[buf][ci][min(ty, M3_TY - 2)];
      }
    }
    if (lastp && !tma) cp_async_wait<0>();                   // no wait_group above on the last plane
    if (fin) {
      const float* src = &sm.fin[(k - 1) % 3][0][tid];
#pragma unroll
      for (int ci = 0; ci < 3; ++ci) {
        fu0[ci] = src[ci * M3_FIN];
        if (!NB) { fV[ci] = src[(3 + ci) * M3_FIN]; fm0[ci] = src[(6 + ci) * M3_FIN]; }
      }
    }
    if (!tma && !TOP && k + 1 <= kend - 1) stage_fin(k + 1);  // joins the group committed at the top of the next plane
    if (fin) {                                               // finalise cell (i,j,k-1)
#pragma unroll
      for (int ci = 0; ci < 3; ++ci) {
        float r = pend[ci] - F[ci][2];
        // accelerate! (Flow.jl:69-73): uniform part and sampled field are ONE term, summed first (as rhs_f does)
        if (ACCF) r += a.acc.v[ci] + a.accfc[ci][opv];
        else r += a.acc.v[ci];
        float X;
        if (NB) {
          X = fu0[ci] + a.dt * r;                            // f - 0, then 0 + 1·f
        } else {
          const float f = fu0[ci] + a.dt * r - fV[ci];       // BDIM! loop 1, Flow.jl:178
          X = fV[ci] + fm0[ci] * f;                          // loop 2 with μ₁ ≡ 0, Flow.jl:179
        }
        a.unc[ci][opv] = a.corrector ? (w[ci][1] + X) * 0.5f : X;   // scale_u!(½) Flow.jl:209
      }
    }
    if (!lastp) {
      float(*nxt)[M3_TILE] = reinterpret_cast<float(*)[M3_TILE]>(sm.ring[(k + 2) & (M3_RING - 1)]);
#pragma unroll
      for (int ci = 0; ci < 3; ++ci) {
        float r = F[ci][0];
        r -= Fxu[ci]; r += F[ci][1]; r -= Fyu[ci]; r += F[ci][2];
        pend[ci] = r;
        w[ci][0] = w[ci][1]; w[ci][1] = w[ci][2]; w[ci][2] = w[ci][3];
        w[ci][3] = nxt[ci][tc];
      }
    }
  };
  auto march = [&](auto top_tag) {
    for (int k = k0; k <= kend; ++k) {
      if ((k > 1 || p2 || a.per.zlo) && k < kend) plane_step(std::true_type{}, top_tag, k);
      else plane_step(std::false_type{}, top_tag, k);
    }
  };
  // (evaluating only the y faces in the block's last row was tried with 31-column tiles: one light warp per
  //  block does not shorten the plane, whose time is set by the SMSPs that host four full warps.  With 32-column
  //  tiles that row also evaluates the faces right of the block; dropping its six unused fluxes keeps it from
  //  becoming the warp every barrier waits for: measured +6 % without)
  if (W32 && ty == M3_TY - 1) march(std::true_type{});
  else march(std::false_type{});
  cp_async_wait<0>();
}

// ACCF: with the caller's acceleration field -- a separate instantiation, because even a uniform branch on a null
// pointer in the finalisation cost the hot variant 6 % (777 vs 730 us, profiles/r2_c17_*)
template <int S, bool NB, bool W32, bool ACCF>
__global__ void __launch_bounds__(32 * M3_TY, 2) k_mom3(const __grid_constant__ Mom3Args a) {
  pdl_entry();
  extern __shared__ __align__(128) unsigned char m3_raw[];
  Mom3Smem& sm = *reinterpret_cast<Mom3Smem*>(m3_raw);
  // FAST: every thread of the block sits on an ordinary cell with ordinary faces (in a periodic direction
  // that is every block whose last lane / row still lies inside the array)
  const int px = a.per.p[0] ? 1 : 0, py = a.per.p[1] ? 1 : 0;
  // (W32: lane 31 is an output cell, so it must be a real cell, and the face at column i0+32 an ordinary one)
  const bool inner_x = W32 ? ((int)blockIdx.x >= 1 - px && 1 + 32 * (int)blockIdx.x + 31 <= a.l.n0 - 2 &&
                              1 + 32 * (int)blockIdx.x + 32 <= a.l.n0 - 2 + px)
                           : ((int)blockIdx.x >= 1 - px && 1 + 31 * (int)blockIdx.x + 31 <= a.l.n0 - 2 + px);
  const bool inner = inner_x &&
                     (int)blockIdx.y >= 1 - py && 1 + (M3_TY - 1) * (int)blockIdx.y + (M3_TY - 1) <= a.l.n1 - 2 + py;
  if (inner) mom3_body_async<S, true, NB, W32, ACCF>(a, sm);
  else mom3_body_async<S, false, NB, W32, ACCF>(a, sm);
}

// ---- band cells (any μ₁ ≠ 0): exact BDIM with μddn, Flow.jl:20-26,176-180 (two passes, below).
struct PatchArgs {
  MomArgs m;           // ua, u0, V, per, nu, dt, acc (m.f unused)
  const float* accf;   // accelerate! field of this phase or null
  const float* mu0;
  const float* mu1;
  float* un;
  const int* list;
  const int* count;
  int cap;
  int corrector;
};
// ---- band patch in two passes.  (A one-warp-per-cell kernel evaluated every f value up to seven times -- a band
// cell and its six neighbours are mostly band cells themselves: 141 us per call on the sphere of config C, 521 us
// on the cylinder of config D; removed in round 2.)  Here the cells whose f is needed -- the band and its
// face neighbours, de-duplicated once per measure! into `list2` -- get f[I,ci] exactly once (K-PATCH-F, one thread
// per cell and component, written to the otherwise idle flow.f), then one thread per band cell and component
// applies μddn (K-PATCH-APPLY).  Same rhs_f, same summation order as the reference: bit-identical.
struct Band2Args { Lay l; const int* list; const int* count; int cap; int* mark; int epoch; int* list2; int* count2; int cap2; };
__global__ void __launch_bounds__(256) k_band_nbrs(const Band2Args a) {
  pdl_entry();
  const int n = min(*a.count, a.cap);
  const int st[3] = {1, a.l.pitch, a.l.plane};
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < 7 * n; t += gridDim.x * blockDim.x) {
    const int q = t / 7, m = t - 7 * q;
    const int o = a.list[q] + (m == 0 ? 0 : ((m & 1) ? st[(m - 1) >> 1] : -st[(m - 1) >> 1]));
    if (atomicExch(&a.mark[o], a.epoch) != a.epoch) {
      const int idx = atomicAdd(a.count2, 1);
      if (idx < a.cap2) a.list2[idx] = o;
    }
  }
}
template <int S>
__global__ void __launch_bounds__(128) k_patch_f(const PatchArgs a, const int* list2, const int* count2, int cap2) {
  pdl_entry();
  const Lay& l = a.m.l;
  const int n = min(*count2, cap2);
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < 3 * n; t += gridDim.x * blockDim.x) {
    const int ci = t / n, o = list2[t - ci * n];
    const int k = o / l.plane, rem = o - k * l.plane, j = rem / l.pitch, i = rem - j * l.pitch;
    a.m.f[l.cstride * ci + o] = a.accf ? rhs_f<3, S, true>(a.m, ci, i, j, k, a.accf) : rhs_f<3, S>(a.m, ci, i, j, k);
  }
}
__global__ void __launch_bounds__(128) k_patch_apply(const PatchArgs a) {
  pdl_entry();
  const Lay& l = a.m.l;
  const int n = min(*a.count, a.cap);
  const int st[3] = {1, l.pitch, l.plane};
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < 3 * n; t += gridDim.x * blockDim.x) {
    const int ci = t / n, o = a.list[t - ci * n];
    const float* fc = a.m.f + l.cstride * ci;
    float s = 0.f;
#pragma unroll
    for (int dj = 0; dj < 3; ++dj) s += a.mu1[l.cstride * (ci + 3 * dj) + o] * (fc[o + st[dj]] - fc[o - st[dj]]);
    const long long oc = l.cstride * ci + o;
    const float X = s / 2.f + a.m.V[oc] + a.mu0[oc] * fc[o];
    a.un[oc] = a.corrector ? (a.m.ua[oc] + X) * 0.5f : X;
  }
}

// list of interior cells with any μ₁ ≠ 0 (rebuilt whenever the coefficients change)
struct BandArgs { Lay l; const float* mu1; int* list; int* count; int cap; };
template <int D>
__global__ void __launch_bounds__(256) k_band_scan(const BandArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
  const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z + 1 : 0;
  if (i > l.n0 - 2 || j > l.n1 - 2 || (D == 3 && k > l.n2 - 2)) return;
  const int o = l.off(i, j, k);
  bool any = false;
#pragma unroll
  for (int q = 0; q < D * D; ++q) any |= a.mu1[l.cstride * q + o] != 0.f;
  if (any) {
    const int idx = atomicAdd(a.count, 1);
    if (idx < a.cap) a.list[idx] = o;
  }
}

// u[exit plane, comp 1] carried from the previous buffer (the reference updates u in place, so the
// exit plane keeps its old values until exitBC! rewrites it, core.jl:212,226-233)
struct ExitCopyArgs { Lay l; const float* src; float* dst; };
__global__ void k_copy_exit_plane(const ExitCopyArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int j = blockIdx.x * blockDim.x + threadIdx.x, k = blockIdx.y;
  if (j >= l.n1 || k >= l.n2) return;
  const int o = l.off(l.n0 - 1, j, k);
  a.dst[o] = a.src[o];
}

}  // namespace wl
