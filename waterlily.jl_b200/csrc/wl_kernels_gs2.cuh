// wl_kernels_gs2.cuh -- K-GS2: TWO red-black half sweeps of GaussSeidelRB! in one pass over the level
// (reference src/Poisson.jl:116-148: `ϵ = r·iD`, perBC!(ϵ), then four `gauss_rb` colour sweeps).
//
// The per-sweep kernels (k_gs_half_c) are DRAM-bound at ~95 % of the measured bandwidth, but each streams
// r, L(3), ϵ once per colour: 4 x 24 B/cell per smooth.  A colour only reads the other colour, so the sweep of
// colour 1 on plane p+1 and the sweep of colour 0 on plane p can run in the same march step: one pass reads
// r, iD, L once and writes ϵ once (24 B/cell with the ϵ = r·iD initialisation folded in, 28 B/cell for the
// second pair).  Two passes replace the ϵ initialisation and the four sweeps.
//
// Decomposition: a block owns 60 x 14 columns of cells and marches up the last dimension.  A warp is one ROW
// of 64 cells (a lane owns the x-pair (x, x+1): one cell of each colour on every plane), the block is 16 rows;
// the outer pair of every row and the outer rows are the halo a second sweep needs (their first sweep is
// recomputed, their second sweep is not).  In-plane neighbours travel by warp shuffle (x) and through a
// double-buffered shared-memory plane (y), z-neighbours stay in the thread's registers; one __syncthreads per
// plane.  Same operations in the same order as gauss_rb (Poisson.jl:124-128): bit-identical to the per-sweep
// kernels.  Reference semantics kept: periodic ghosts are STALE for all four sweeps (they hold the initial
// r·iD of the wrapped cell, recomputed here from r and iD, which no sweep changes), non-periodic ghosts are
// read as they are in memory, ghost cells are never swept.
#pragma once
#include "wl_kernels_mg4.cuh"

namespace wl {

constexpr int G2_XS = 60;          // output cells per row of a block (64 minus the halo pairs)
constexpr int G2_TY = 14;          // output rows per block
constexpr int G2_NR = G2_TY + 2;   // rows = warps per block
constexpr int G2_NT = 32 * G2_NR;

struct Gs2Args {
  Lay l;
  const float* L; const float* iD; const float* r;
  const float* ei;   // ϵ before the pair (second pair; the first pair forms r·iD itself and reads only non-periodic ghosts here)
  float* eo;         // ϵ after the pair: a DIFFERENT buffer than ei for the second pair (neighbouring blocks still read ei)
  Per per;
  int KZ;            // planes per z-chunk
};

struct F2 { float a, b; };
__device__ __forceinline__ F2 ld2g(const float* p) { const float2 t = __ldg(reinterpret_cast<const float2*>(p)); return F2{t.x, t.y}; }
__device__ __forceinline__ F2 ld2(const float* p) { const float2 t = *reinterpret_cast<const float2*>(p); return F2{t.x, t.y}; }

struct Gs2Smem {                   // [buffer][row][cell of the pair][lane]: conflict-free stores and loads
  float e0[2][G2_NR + 2][2][32];   // e_in of the plane loaded last / the one before; rows y0-2 .. y0+TY+1
  float e1[2][G2_NR][2][32];       // after the first sweep of the pair
};

// class of coordinate c in a direction of extent n: 0 interior, 1 lower ghost, 2 upper ghost, 3 outside
__device__ __forceinline__ int g2_class(int c, int n) { return c < 0 || c > n - 1 ? 3 : (c == 0 ? 1 : (c == n - 1 ? 2 : 0)); }

// INIT: the first pair of sweeps (input ϵ = r·iD formed here); else the second pair (input ϵ from memory).
// FAST: every cell the block touches in x/y (halo included) is an interior cell.
template <bool INIT, bool FAST>
__device__ __forceinline__ void gs_pair_body(const Gs2Args& a, Gs2Smem& sm) {
  const Lay& l = a.l;
  const int lane = threadIdx.x, w = threadIdx.y;
  const int n0 = l.n0, n1 = l.n1, n2 = l.n2, pitch = l.pitch, plane = l.plane;
  const int x = 1 + G2_XS * blockIdx.x - 2 + 2 * lane;           // cells x (a) and x+1 (b)
  const int y0 = 1 + G2_TY * blockIdx.y;
  const int y = y0 - 1 + w;
  const int kb = 1 + a.KZ * blockIdx.z, ke = min(kb + a.KZ - 1, n2 - 2);
  const float* __restrict__ R = a.r;
  const float* __restrict__ ID = a.iD;
  const float* __restrict__ Lx = a.L;
  const float* __restrict__ Ly = a.L + l.cstride;
  const float* __restrict__ Lz = a.L + 2 * l.cstride;
  const float* E = a.ei;
  float* EO = a.eo;
  const int px = a.per.p[0], py = a.per.p[1], pz = a.per.p[2];
  // fixed classification of the thread's two cells and of its row
  const int cxa = g2_class(x, n0), cxb = g2_class(x + 1, n0), cy = g2_class(y, n1);
  const bool pair_live = x + 1 >= 0 && x <= n0 - 1;                // some cell of the pair lies inside the array
  const bool row_in = cy == 0;
  const bool upd_a = cxa == 0 && row_in, upd_b = cxb == 0 && row_in;   // swept when the plane is interior too
  const int oxy = x + pitch * y;
  // extra halo rows of e_in: warp 0 also serves row y0-2, the last warp row y0+TY+1
  const bool extra = w == 0 || w == G2_NR - 1;
  const int ye = w == 0 ? y0 - 2 : y0 + G2_TY + 1;
  const int se = w == 0 ? 0 : G2_NR + 1;                           // their row in sm.e0

  // e_in of the pair at (x, yy, p): what the sweeps see there before they start
  auto ein = [&](int yy, int p) -> F2 {
    F2 v{0.f, 0.f};
    if (FAST && p >= 1 && p <= n2 - 2) {
      const int o = x + pitch * yy + plane * p;
      if (INIT) { const F2 r = ld2g(R + o), d = ld2g(ID + o); v.a = r.a * d.a; v.b = r.b * d.b; }
      else v = ld2(E + o);
      return v;
    }
    const int cyy = g2_class(yy, n1), cp = g2_class(p, n2);
    if (cyy == 3 || cp == 3 || !pair_live) return v;
    const int ng = (cyy != 0) + (cp != 0);
    if (ng == 2) return v;                                         // an edge of the array: never read by a sweep
    if (ng == 0) {
      const int o = x + pitch * yy + plane * p;
      if (cxa == 0 && cxb == 0) {
        if (INIT) { const F2 r = ld2g(R + o), d = ld2g(ID + o); v.a = r.a * d.a; v.b = r.b * d.b; }
        else v = ld2(E + o);
        return v;
      }
      // the pair straddles an x ghost
      auto cell = [&](int xx, int cx) -> float {
        if (cx == 3) return 0.f;
        const int oo = o + (xx - x);
        if (cx == 0) return INIT ? __ldg(R + oo) * __ldg(ID + oo) : E[oo];
        if (!px) return E[oo];
        const int ow = oo + (cx == 1 ? n0 - 2 : -(n0 - 2));       // stale image: r·iD of the wrapped cell
        return __ldg(R + ow) * __ldg(ID + ow);
      };
      v.a = cell(x, cxa); v.b = cell(x + 1, cxb);
      return v;
    }
    // ghost row or ghost plane (one of them): x-interior cells only
    const bool per_d = cyy != 0 ? py != 0 : pz != 0;
    const int yw = cyy == 1 ? n1 - 2 : (cyy == 2 ? 1 : yy), pw = cp == 1 ? n2 - 2 : (cp == 2 ? 1 : p);
    const int og = x + pitch * yy + plane * p, ow = x + pitch * yw + plane * pw;
    if (cxa == 0) v.a = per_d ? __ldg(R + ow) * __ldg(ID + ow) : E[og];
    if (cxb == 0) v.b = per_d ? __ldg(R + ow + 1) * __ldg(ID + ow + 1) : E[og + 1];
    return v;
  };

  // register pipeline: e_in of planes m-1, m, m+1 (z-neighbours of the first sweep), e1 of planes m-2, m-1, m
  F2 ein_m1{0.f, 0.f}, ein_0{0.f, 0.f}, ein_p1{0.f, 0.f};
  F2 e1_m2{0.f, 0.f}, e1_m1{0.f, 0.f}, e1_0{0.f, 0.f};
  // coefficients of the cell the SECOND sweep updates on plane m-1 (carried over from the step that loaded them)
  float c2_r = 0.f, c2_d = 0.f, c2_xl = 0.f, c2_xh = 0.f, c2_yl = 0.f, c2_yh = 0.f, c2_zl = 0.f, c2_zh = 0.f;

  // prologue: e_in of planes kb-2 and kb-1
  ein_0 = ein(y, kb - 2);
  ein_p1 = ein(y, kb - 1);
  {
    const int b1 = (kb - 1) & 1;
    sm.e0[b1][w + 1][0][lane] = ein_p1.a; sm.e0[b1][w + 1][1][lane] = ein_p1.b;
    if (extra) { const F2 t = ein(ye, kb - 1); sm.e0[b1][se][0][lane] = t.a; sm.e0[b1][se][1][lane] = t.b; }
  }
  // plane kb-2 only serves as the z-neighbour of plane kb-1: its first sweep is not needed, e1(kb-2) is never read
  __syncthreads();

  // step m: first sweep on plane m (m = kb-1 .. ke+1), second sweep on plane m-1 (kb .. ke)
  for (int m = kb - 1; m <= ke + 1; ++m) {
    // ---- load plane m+1's e_in (the upper z-neighbour of the first sweep) and publish it for step m+1
    ein_m1 = ein_0; ein_0 = ein_p1;
    ein_p1 = ein(y, m + 1);
    {
      const int bn = (m + 1) & 1;
      sm.e0[bn][w + 1][0][lane] = ein_p1.a; sm.e0[bn][w + 1][1][lane] = ein_p1.b;
      if (extra) { const F2 t = ein(ye, m + 1); sm.e0[bn][se][0][lane] = t.a; sm.e0[bn][se][1][lane] = t.b; }
    }
    const bool plane_in = m >= 1 && m <= n2 - 2;
    // ---- coefficients of plane m (both cells of the pair)
    F2 r{0.f, 0.f}, d{0.f, 0.f}, lx{0.f, 0.f}, ly{0.f, 0.f}, lyu{0.f, 0.f}, lz{0.f, 0.f}, lzu{0.f, 0.f};
    float lx2 = 0.f;
    const bool ld_ok = plane_in && (FAST || (row_in && x >= 0 && x + 1 <= n0 - 1));
    if (ld_ok) {
      const int o = oxy + plane * m;
      r = ld2g(R + o); d = ld2g(ID + o); lx = ld2g(Lx + o);
      ly = ld2g(Ly + o); lyu = ld2g(Ly + o + pitch); lz = ld2g(Lz + o); lzu = ld2g(Lz + o + plane);
    } else if (!FAST && plane_in && row_in && cxa == 2) {
      lx.a = __ldg(Lx + oxy + plane * m);                          // upper face of the row's last cell (periodic image or 0)
    }
    lx2 = __shfl_down_sync(0xffffffffu, lx.a, 1);                 // Lx at x+2 (lane 31: unused, its cell b is halo)
    // ---- first sweep of the pair on plane m: colour 1 (x+y+m odd) for the first pair, i.e. sweeps 1 and 3
    const int par = (x + y + m) & 1;                               // 1: cell a has colour 1, 0: cell b has it
    e1_m2 = e1_m1; e1_m1 = e1_0;
    e1_0 = ein_0;
    {
      const float (*rowm)[32] = sm.e0[m & 1][w];                   // row y-1
      const float (*rowp)[32] = sm.e0[m & 1][w + 2];               // row y+1
      // x-neighbours of the updated cell: the other cell of the pair and the adjacent lane's nearer cell
      const float up_b = __shfl_up_sync(0xffffffffu, ein_0.b, 1);      // cell x-1
      const float dn_a = __shfl_down_sync(0xffffffffu, ein_0.a, 1);    // cell x+2
      if (par) {                                                   // update a
        const float ym = rowm[0][lane], yp = rowp[0][lane];
        float s = r.a;
        s -= up_b * lx.a + ein_0.b * lx.b;
        s -= ym * ly.a + yp * lyu.a;
        s -= ein_m1.a * lz.a + ein_p1.a * lzu.a;
        if (upd_a && plane_in) e1_0.a = s * d.a;
      } else {                                                     // update b
        const float ym = rowm[1][lane], yp = rowp[1][lane];
        float s = r.b;
        s -= ein_0.a * lx.b + dn_a * lx2;
        s -= ym * ly.b + yp * lyu.b;
        s -= ein_m1.b * lz.b + ein_p1.b * lzu.b;
        if (upd_b && plane_in) e1_0.b = s * d.b;
      }
    }
    sm.e1[m & 1][w][0][lane] = e1_0.a; sm.e1[m & 1][w][1][lane] = e1_0.b;
    // ---- second sweep of the pair on plane m-1: the other colour, coefficients carried over from step m-1
    if (m - 1 >= kb) {
      const int q = m - 1;
      const int par2 = (x + y + q) & 1;                            // colour 0 cell of plane q: a if par2 == 0
      F2 out = e1_m1;
      if (w >= 1 && w <= G2_TY) {
        const float (*rowm)[32] = sm.e1[q & 1][w - 1];
        const float (*rowp)[32] = sm.e1[q & 1][w + 1];
        const float up_b = __shfl_up_sync(0xffffffffu, e1_m1.b, 1);
        const float dn_a = __shfl_down_sync(0xffffffffu, e1_m1.a, 1);
        const bool q_in = q >= 1 && q <= n2 - 2;
        if (!par2) {                                               // update a
          const float ym = rowm[0][lane], yp = rowp[0][lane];
          float s = c2_r;
          s -= up_b * c2_xl + e1_m1.b * c2_xh;
          s -= ym * c2_yl + yp * c2_yh;
          s -= e1_m2.a * c2_zl + e1_0.a * c2_zh;
          if (upd_a && q_in) out.a = s * c2_d;
        } else {
          const float ym = rowm[1][lane], yp = rowp[1][lane];
          float s = c2_r;
          s -= e1_m1.a * c2_xl + dn_a * c2_xh;
          s -= ym * c2_yl + yp * c2_yh;
          s -= e1_m2.b * c2_zl + e1_0.b * c2_zh;
          if (upd_b && q_in) out.b = s * c2_d;
        }
        // store the plane: interior cells of the block's output columns
        if (lane >= 1 && lane <= 30 && (FAST || y <= n1 - 2) && q_in) {
          const int o = oxy + plane * q;
          if (FAST || (cxa == 0 && cxb == 0)) *reinterpret_cast<float2*>(EO + o) = make_float2(out.a, out.b);
          else { if (cxa == 0) EO[o] = out.a; if (cxb == 0) EO[o + 1] = out.b; }
        }
      }
    }
    // ---- hand the coefficients of plane m's OTHER colour to the next step's second sweep
    if (par) { c2_r = r.b; c2_d = d.b; c2_xl = lx.b; c2_xh = lx2; c2_yl = ly.b; c2_yh = lyu.b; c2_zl = lz.b; c2_zh = lzu.b; }
    else { c2_r = r.a; c2_d = d.a; c2_xl = lx.a; c2_xh = lx.b; c2_yl = ly.a; c2_yh = lyu.a; c2_zl = lz.a; c2_zh = lzu.a; }
    __syncthreads();
  }
}

// ---- the same march for blocks that touch no periodic seam in x or y (every block of a non-periodic grid): all
// global data of a plane arrive by cp.async two planes ahead into a three-slot shared-memory ring (r, iD, Lx, Ly, Lz
// [, ϵ] of the block's rows), so no sweep waits on a global load and neighbouring rows' coefficients (Ly of row y+1,
// Lx of cell x+2, Lz of plane m+1) are read from the ring instead of being loaded twice.  Addresses of cells outside
// the array are clamped (their values only reach ghost cells, which are never swept); non-periodic ghosts take
// r·iD = 0 (first pair: iD is zero on those ghosts, as is the ϵ ghost the reference reads) or the ϵ in memory
// (second pair); a periodic last dimension reads the wrapped plane for its stale ghost image.
constexpr int G2_NA = 6;                                      // r, iD, Lx, Ly, Lz, ϵ
struct Gs2Ring {
  float2 v[3][G2_NA][G2_NR + 2][32];                          // [slot][array][row y0-2 .. y0+TY+1][lane]
  Gs2Smem t;
};
__device__ __forceinline__ void cp_async8(void* smem, const void* g) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(g) : "memory");
}
template <bool INIT>
__device__ __forceinline__ void gs_pair_staged(const Gs2Args& a, Gs2Ring& rg) {
  Gs2Smem& sm = rg.t;
  const Lay& l = a.l;
  const int lane = threadIdx.x, w = threadIdx.y;
  const int n0 = l.n0, n1 = l.n1, n2 = l.n2, pitch = l.pitch, plane = l.plane;
  const int x = 1 + G2_XS * blockIdx.x - 2 + 2 * lane;
  const int y0 = 1 + G2_TY * blockIdx.y;
  const int y = y0 - 1 + w;
  const int kb = 1 + a.KZ * blockIdx.z, ke = min(kb + a.KZ - 1, n2 - 2);
  const float* __restrict__ R = a.r;
  const float* __restrict__ ID = a.iD;
  const float* __restrict__ Lx = a.L;
  const float* __restrict__ Ly = a.L + l.cstride;
  const float* __restrict__ Lz = a.L + 2 * l.cstride;
  const float* E = a.ei;
  float* EO = a.eo;
  const int pz = a.per.p[2];
  const bool upd_a = x >= 1 && x <= n0 - 2 && y >= 1 && y <= n1 - 2;
  const bool upd_b = x + 1 >= 1 && x + 1 <= n0 - 2 && y >= 1 && y <= n1 - 2;
  const int xo = (x + 1 < 0 || x > n0 - 1) ? 1 : x;               // a pair wholly outside the array reads a valid address
  const int oxy = xo + pitch * min(max(y, 0), n1 - 1);
  const bool extra = w == 0 || w == G2_NR - 1;
  const int ye = w == 0 ? y0 - 2 : y0 + G2_TY + 1;
  const int se = w == 0 ? 0 : G2_NR + 1;
  const int oxe = xo + pitch * min(max(ye, 0), n1 - 1);
  const int ostore = x + pitch * y;

  auto stage = [&](int p, int slot) {
    const int pr = min(max(p, 0), n2 - 1);                         // the plane itself (L: ghosts hold the images BC! left)
    const int pw = (pz && pr == 0) ? n2 - 2 : ((pz && pr == n2 - 1) ? 1 : pr);   // r, iD: wrapped across a periodic seam
    const int ow = oxy + plane * pw, orr = oxy + plane * pr;
    cp_async8(&rg.v[slot][0][w + 1][lane], R + ow);
    cp_async8(&rg.v[slot][1][w + 1][lane], ID + ow);
    cp_async8(&rg.v[slot][2][w + 1][lane], Lx + orr);
    cp_async8(&rg.v[slot][3][w + 1][lane], Ly + orr);
    cp_async8(&rg.v[slot][4][w + 1][lane], Lz + orr);
    if (!INIT) cp_async8(&rg.v[slot][5][w + 1][lane], E + orr);
    if (extra) {
      cp_async8(&rg.v[slot][0][se][lane], R + oxe + plane * pw);   // (second pair: only a periodic ghost plane needs r·iD)
      cp_async8(&rg.v[slot][1][se][lane], ID + oxe + plane * pw);
      if (!INIT) cp_async8(&rg.v[slot][5][se][lane], E + oxe + plane * pr);
      if (w == G2_NR - 1) cp_async8(&rg.v[slot][3][se][lane], Ly + oxe + plane * pr);
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");
  };
  // e_in of the pair in ring row `row` of plane p (landed in `slot`)
  auto ein_of = [&](int p, int slot, int row) -> F2 {
    const bool wrapz = pz && (p == 0 || p == n2 - 1);
    if (INIT || wrapz) {
      const float2 r = rg.v[slot][0][row][lane], d = rg.v[slot][1][row][lane];
      return F2{r.x * d.x, r.y * d.y};
    }
    const float2 e = rg.v[slot][5][row][lane];
    return F2{e.x, e.y};
  };

  int s_m1 = 0, s_0 = 1, s_p1 = 2;                                 // slots of planes m-1 (being refilled with m+2), m, m+1
  stage(kb - 2, 0); stage(kb - 1, 1); stage(kb, 2);
  asm volatile("cp.async.wait_group 0;\n" ::: "memory");
  __syncthreads();
  F2 ein_m1{0.f, 0.f}, ein_0 = ein_of(kb - 2, 0, w + 1), ein_p1 = ein_of(kb - 1, 1, w + 1);
  {
    const int b1 = (kb - 1) & 1;
    sm.e0[b1][w + 1][0][lane] = ein_p1.a; sm.e0[b1][w + 1][1][lane] = ein_p1.b;
    if (extra) { const F2 t = ein_of(kb - 1, 1, se); sm.e0[b1][se][0][lane] = t.a; sm.e0[b1][se][1][lane] = t.b; }
  }
  F2 e1_m2{0.f, 0.f}, e1_m1{0.f, 0.f}, e1_0{0.f, 0.f};
  float c2_r = 0.f, c2_d = 0.f, c2_xl = 0.f, c2_xh = 0.f, c2_yl = 0.f, c2_yh = 0.f, c2_zl = 0.f, c2_zh = 0.f;
  __syncthreads();
  // on entry of step m: s_0 holds plane m-1 ... rotate first
  for (int m = kb - 1; m <= ke + 1; ++m) {
    // slots: plane m in s_0', plane m+1 in s_p1'; the slot of plane m-1 is refilled with plane m+2
    { const int t = s_m1; s_m1 = s_0; s_0 = s_p1; s_p1 = t; }
    // after the rotation s_m1 = plane m, s_0 = plane m+1, s_p1 = free (held plane m-1): rename for clarity
    const int sl_m = s_m1, sl_p = s_0, sl_free = s_p1;
    if (m + 2 <= ke + 2) stage(m + 2, sl_free);
    ein_m1 = ein_0; ein_0 = ein_p1;
    ein_p1 = ein_of(m + 1, sl_p, w + 1);
    {
      const int bn = (m + 1) & 1;
      sm.e0[bn][w + 1][0][lane] = ein_p1.a; sm.e0[bn][w + 1][1][lane] = ein_p1.b;
      if (extra) { const F2 t = ein_of(m + 1, sl_p, se); sm.e0[bn][se][0][lane] = t.a; sm.e0[bn][se][1][lane] = t.b; }
    }
    const bool plane_in = m >= 1 && m <= n2 - 2;
    const float2 r = rg.v[sl_m][0][w + 1][lane], d = rg.v[sl_m][1][w + 1][lane], lx = rg.v[sl_m][2][w + 1][lane],
                 ly = rg.v[sl_m][3][w + 1][lane], lyu = rg.v[sl_m][3][w + 2][lane], lz = rg.v[sl_m][4][w + 1][lane],
                 lzu = rg.v[sl_p][4][w + 1][lane];
    const float lx2 = (&rg.v[sl_m][2][w + 1][0].x)[2 * lane + 2];   // Lx at x+2 (lane 31: the next row's first value, unused)
    const int par = (x + y + m) & 1;
    e1_m2 = e1_m1; e1_m1 = e1_0;
    e1_0 = ein_0;
    {
      const float (*rowm)[32] = sm.e0[m & 1][w];
      const float (*rowp)[32] = sm.e0[m & 1][w + 2];
      const float up_b = __shfl_up_sync(0xffffffffu, ein_0.b, 1);
      const float dn_a = __shfl_down_sync(0xffffffffu, ein_0.a, 1);
      if (par) {
        const float ym = rowm[0][lane], yp = rowp[0][lane];
        float s = r.x;
        s -= up_b * lx.x + ein_0.b * lx.y;
        s -= ym * ly.x + yp * lyu.x;
        s -= ein_m1.a * lz.x + ein_p1.a * lzu.x;
        if (upd_a && plane_in) e1_0.a = s * d.x;
      } else {
        const float ym = rowm[1][lane], yp = rowp[1][lane];
        float s = r.y;
        s -= ein_0.a * lx.y + dn_a * lx2;
        s -= ym * ly.y + yp * lyu.y;
        s -= ein_m1.b * lz.y + ein_p1.b * lzu.y;
        if (upd_b && plane_in) e1_0.b = s * d.y;
      }
    }
    sm.e1[m & 1][w][0][lane] = e1_0.a; sm.e1[m & 1][w][1][lane] = e1_0.b;
    if (m - 1 >= kb) {
      const int q = m - 1;
      const int par2 = (x + y + q) & 1;
      F2 out = e1_m1;
      if (w >= 1 && w <= G2_TY) {
        const float (*rowm)[32] = sm.e1[q & 1][w - 1];
        const float (*rowp)[32] = sm.e1[q & 1][w + 1];
        const float up_b = __shfl_up_sync(0xffffffffu, e1_m1.b, 1);
        const float dn_a = __shfl_down_sync(0xffffffffu, e1_m1.a, 1);
        if (!par2) {
          const float ym = rowm[0][lane], yp = rowp[0][lane];
          float s = c2_r;
          s -= up_b * c2_xl + e1_m1.b * c2_xh;
          s -= ym * c2_yl + yp * c2_yh;
          s -= e1_m2.a * c2_zl + e1_0.a * c2_zh;
          if (upd_a) out.a = s * c2_d;
        } else {
          const float ym = rowm[1][lane], yp = rowp[1][lane];
          float s = c2_r;
          s -= e1_m1.a * c2_xl + dn_a * c2_xh;
          s -= ym * c2_yl + yp * c2_yh;
          s -= e1_m2.b * c2_zl + e1_0.b * c2_zh;
          if (upd_b) out.b = s * c2_d;
        }
        if (lane >= 1 && lane <= 30 && y <= n1 - 2) {             // q is an interior plane of the chunk
          const int o = ostore + plane * q;
          if (upd_a && upd_b) *reinterpret_cast<float2*>(EO + o) = make_float2(out.a, out.b);
          else { if (upd_a) EO[o] = out.a; if (upd_b) EO[o + 1] = out.b; }
        }
      }
    }
    if (par) { c2_r = r.y; c2_d = d.y; c2_xl = lx.y; c2_xh = lx2; c2_yl = ly.y; c2_yh = lyu.y; c2_zl = lz.y; c2_zh = lzu.y; }
    else { c2_r = r.x; c2_d = d.x; c2_xl = lx.x; c2_xh = lx.y; c2_yl = ly.x; c2_yh = lyu.x; c2_zl = lz.x; c2_zh = lzu.x; }
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncthreads();
  }
}

template <bool INIT>
__global__ void __launch_bounds__(G2_NT, 2) k_gs_pair(const Gs2Args a) {
  pdl_entry();
  extern __shared__ __align__(16) unsigned char g2_raw[];
  Gs2Ring& rg = *reinterpret_cast<Gs2Ring*>(g2_raw);
  const int xl = 1 + G2_XS * (int)blockIdx.x - 2, yl = 1 + G2_TY * (int)blockIdx.y - 2;
  const bool seam_x = a.per.p[0] && !(xl >= 1 && xl + 63 <= a.l.n0 - 2);
  const bool seam_y = a.per.p[1] && !(yl >= 1 && yl + G2_TY + 3 <= a.l.n1 - 2);
  if (!seam_x && !seam_y) gs_pair_staged<INIT>(a, rg);
  else gs_pair_body<INIT, false>(a, rg.t);
}

}  // namespace wl
