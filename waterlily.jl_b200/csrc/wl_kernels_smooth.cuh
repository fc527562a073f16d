// wl_kernels_smooth.cuh -- K-SMOOTH: GaussSeidelRB!(p;it=4,ω) + increment! (+ L₁/L∞ of the new
// residual) in ONE kernel (reference src/Poisson.jl:100-104,116-148, MultiLevelPoisson.jl:106,115).
//
// The un-fused path streams r, ϵ, L, iD through HBM six times (ϵ init, four colour half-sweeps,
// increment): ≈156 B/cell.  Here ϵ never leaves shared memory: a block owns a 32×16 column of cells
// and marches up the last dimension with a software pipeline of six stages, one plane apart:
//     stage 0  ϵ = r·iD                       on plane m
//     stage s  colour half-sweep s (s=1..4)   on plane m-s
//     stage 5  r' = r - ω·Aϵ,  x += ω·ϵ       on plane m-5
// A colour only reads the other colour, the in-plane neighbours of stage s were written one march
// step earlier and the z-neighbours belong to the same thread, so ONE __syncthreads per plane
// orders everything.  Blocks are independent: each recomputes a halo that shrinks by one cell per
// stage (5 cells around the column for stage 0 ... none for stage 5), in x/y and at the ends of its
// z-chunk.  HBM traffic: r, x, iD, L(3) read, r', x written = 36 B/cell (algorithmic 28 + iD).
//
// Reference semantics kept bit for bit:
//  * same operation order as gauss_rb / increment! / mult (Poisson.jl:70-76,101-103,124-128);
//  * STALE periodic ghosts: the reference fills ϵ's ghosts once after the init and never during the
//    sweeps, so across a periodic seam a sweep reads the neighbour's INITIAL ϵ (= r·iD, recomputed
//    from global memory), while increment! (which starts with perBC!(ϵ)) reads the final one.  The
//    halo of a tile simply continues across the seam in wrapped coordinates and evolves like any
//    other cell; only the links that cross the seam are special;
//  * non-periodic ghosts hold ϵ = 0 for the whole smoother.
// r' goes to a second buffer (the level's idle ϵ array; the host swaps the two pointers) because
// neighbouring blocks still read the old r in their halos.
#pragma once
#include "wl_kernels_mg2.cuh"

namespace wl {

constexpr int SF_TX = 32, SF_TY = 16, SF_H = 5;
constexpr int SF_W = SF_TX + 2 * SF_H, SF_HH = SF_TY + 2 * SF_H;   // 42 x 26 cells per staged plane
constexpr int SF_R = 8;                                              // ring slots (7 planes live)
constexpr int SF_NT = 512;
constexpr int SF_MAXKZ = 256;                                        // planes per z-chunk (upper bound)

struct SmoothFArgs {
  LevDev p;          // L, iD, x, r (input residual), eps unused
  float* rn;         // output residual buffer
  Per per;
  float omega;
  RedWork w; Red* out; int want_norms;
  int KZ;            // planes per z-chunk
  int zh;            // halo planes available beyond each interface ghost plane (slab decomposition)
  int seam_zlo, seam_zhi;   // slab interface that is also the periodic seam of the global grid
};

struct SfPlane { int kz; bool live, seam_lo, seam_hi; };

// class of march plane m (unwrapped index): memory plane, whether it holds evolving cells, seam flags
__device__ __forceinline__ SfPlane sf_plane(const SmoothFArgs& a, int m) {
  const int n2 = a.p.l.n2, ni = n2 - 2;
  SfPlane q;
  if (a.per.p[2]) {
    int kz = m;
    if (kz < 1) kz += ni; else if (kz > ni) kz -= ni;
    q.kz = kz; q.live = kz >= 1 && kz <= ni;
    q.seam_lo = kz == 1; q.seam_hi = kz == ni;
  } else {
    const int lo = a.per.zlo ? -a.zh : 1, hi = a.per.zhi ? ni + 1 + a.zh : ni;
    q.kz = m; q.live = m >= lo && m <= hi;
    q.seam_lo = (a.seam_zlo && m == 1) || (a.seam_zhi && m == ni + 1);
    q.seam_hi = (a.seam_zhi && m == ni) || (a.seam_zlo && m == 0);
  }
  return q;
}

struct SfSmem {
  float ring[2 * SF_R][SF_HH * SF_W];  // stages run two planes apart: 12 planes live, 16 slots
  int gxo[SF_W], gyo[SF_HH];           // memory offset of a staged column/row, -1 if it holds no evolving cell
  int sfx[SF_W], sfy[SF_HH];           // periodic-seam flags of the cell: bit0 lower link, bit1 upper link
  // march planes kb-6 .. ke+6 of this block's chunk: .x = memory offset of the plane, .y = flags
  // (bit0 live, bit4 seam below, bit5 seam above, bits 8.. distance to the chunk [kb,ke])
  int2 pl[SF_MAXKZ + 12];
};
constexpr int SF_RM = 2 * SF_R - 1;
constexpr int SF_CELLS = SF_W * SF_HH;                        // 1092
constexpr int SF_N0 = (SF_CELLS + SF_NT - 1) / SF_NT;         // stage-0 cells per thread (3)

// What a thread needs to know about ITS cell of colour half-sweep S, fixed for the whole march:
// the staged position for colour parity 0 (parity 1 is one cell to the right), the in-plane memory
// offsets for both parities (-1: no evolving cell there) and the seam flags.
struct SfCell { int si, o0, o1, fl; };

template <int S>
__device__ __forceinline__ SfCell sf_cell_init(const SfSmem& sm, int i0, int j0) {
  constexpr int PW = SF_W / 2 - S, PH = SF_HH - 2 * S, NP = PW * PH;
  static_assert(NP <= SF_NT, "one cell per thread and sweep");
  const int q = min((int)threadIdx.x, NP - 1);
  const int row = q / PW, px = q - row * PW;
  const int y = S + row, x0 = S + 2 * px;
  const bool have = (int)threadIdx.x < NP;
  const int gy = sm.gyo[y];
  SfCell c;
  c.si = y * SF_W + x0;
  c.o0 = (have && gy >= 0 && sm.gxo[x0] >= 0) ? gy + sm.gxo[x0] : -1;
  c.o1 = (have && gy >= 0 && sm.gxo[x0 + 1] >= 0) ? gy + sm.gxo[x0 + 1] : -1;
  // bits 0-3: seam flags for parity 0 (x lo, x hi, y lo, y hi), bits 4-7: parity 1, bit 8: parity of i0+j0+y
  c.fl = (sm.sfx[x0] | (sm.sfy[y] << 2)) | ((sm.sfx[x0 + 1] | (sm.sfy[y] << 2)) << 4) | (((i0 + j0 + y) & 1) << 8);
  return c;
}

// colour half-sweep S on march plane kp (gauss_rb, Poisson.jl:124-128): one cell per thread, all
// eight coefficient loads issued before the first use; a thread without a cell reads a safe address
// and skips the store.
template <int S>
__device__ __forceinline__ void sf_sweep(const SmoothFArgs& a, SfSmem& sm, const SfCell& c, int kp, int pb) {
  const int2 P = sm.pl[kp - pb];
  const Lay& l = a.p.l;
  const int pitch = l.pitch, plane = l.plane;
  const float* __restrict__ Lx = a.p.L;
  const float* __restrict__ Ly = a.p.L + l.cstride;
  const float* __restrict__ Lz = a.p.L + 2 * l.cstride;
  const int b = ((c.fl >> 8) ^ kp) & 1;                      // colour parity -> which of the thread's two columns
  const int oxy = b ? c.o1 : c.o0;
  const bool ok = (P.y & 1) && (P.y >> 8) <= SF_H - S && oxy >= 0;
  const int o = ok ? oxy + P.x : pitch + 1 + plane;          // (1,1,1) is always allocated
  const float cr = __ldg(a.p.r + o), cd = __ldg(a.p.iD + o);
  const float c0 = __ldg(Lx + o), c1 = __ldg(Lx + o + 1);
  const float c2 = __ldg(Ly + o), c3 = __ldg(Ly + o + pitch);
  const float c4 = __ldg(Lz + o), c5 = __ldg(Lz + o + plane);
  float* cur = sm.ring[kp & SF_RM] + c.si + b;
  const float* dn = sm.ring[(kp - 1) & SF_RM] + c.si + b;
  const float* up = sm.ring[(kp + 1) & SF_RM] + c.si + b;
  float em0 = cur[-1], ep0 = cur[1], em1 = cur[-SF_W], ep1 = cur[SF_W], em2 = dn[0], ep2 = up[0];
  const int fl = ((c.fl >> (4 * b)) & 15) | (P.y & 48);
  if (ok && fl) {                 // links across a periodic seam see the INITIAL ϵ of the neighbour (stale ghosts)
    const int si = c.si + b, y = si / SF_W, x = si - y * SF_W;
    const int gy = sm.gyo[y], gx = sm.gxo[x];
    if (fl & 1) { const int q = gy + sm.gxo[x - 1] + P.x; em0 = __ldg(a.p.r + q) * __ldg(a.p.iD + q); }
    if (fl & 2) { const int q = gy + sm.gxo[x + 1] + P.x; ep0 = __ldg(a.p.r + q) * __ldg(a.p.iD + q); }
    if (fl & 4) { const int q = sm.gyo[y - 1] + gx + P.x; em1 = __ldg(a.p.r + q) * __ldg(a.p.iD + q); }
    if (fl & 8) { const int q = sm.gyo[y + 1] + gx + P.x; ep1 = __ldg(a.p.r + q) * __ldg(a.p.iD + q); }
    if (fl & 16) { const int q = oxy + sm.pl[kp - 1 - pb].x; em2 = __ldg(a.p.r + q) * __ldg(a.p.iD + q); }
    if (fl & 32) { const int q = oxy + sm.pl[kp + 1 - pb].x; ep2 = __ldg(a.p.r + q) * __ldg(a.p.iD + q); }
  }
  float s_ = cr;
  s_ -= em0 * c0 + ep0 * c1;
  s_ -= em1 * c2 + ep1 * c3;
  s_ -= em2 * c4 + ep2 * c5;
  if (ok) cur[0] = s_ * cd;
}

template <int D>
__global__ void __launch_bounds__(SF_NT, 2) k_smooth_fused(const SmoothFArgs a) {
  pdl_entry();
  extern __shared__ __align__(16) unsigned char sf_raw[];
  SfSmem& sm = *reinterpret_cast<SfSmem*>(sf_raw);
  const Lay& l = a.p.l;
  const int tid = threadIdx.x;
  const int n0 = l.n0, n1 = l.n1, n2 = l.n2, pitch = l.pitch, plane = l.plane;
  const int i0 = 1 + SF_TX * blockIdx.x, j0 = 1 + SF_TY * blockIdx.y;
  const int kb = 1 + a.KZ * blockIdx.z, ke = min(kb + a.KZ - 1, n2 - 2);
  // staged columns / rows -> memory coordinates (wrapped across periodic seams) and seam flags
  for (int x = tid; x < SF_W; x += SF_NT) {
    int gi = i0 - SF_H + x;
    if (a.per.p[0]) { if (gi < 1) gi += n0 - 2; else if (gi > n0 - 2) gi -= n0 - 2; }
    const bool live = gi >= 1 && gi <= n0 - 2;
    sm.gxo[x] = live ? gi : -1;
    sm.sfx[x] = (a.per.p[0] && live) ? ((gi == 1 ? 1 : 0) | (gi == n0 - 2 ? 2 : 0)) : 0;
  }
  for (int y = tid; y < SF_HH; y += SF_NT) {
    int gj = j0 - SF_H + y;
    if (a.per.p[1]) { if (gj < 1) gj += n1 - 2; else if (gj > n1 - 2) gj -= n1 - 2; }
    const bool live = gj >= 1 && gj <= n1 - 2;
    sm.gyo[y] = live ? gj * pitch : -1;
    sm.sfy[y] = (a.per.p[1] && live) ? ((gj == 1 ? 1 : 0) | (gj == n1 - 2 ? 2 : 0)) : 0;
  }
  const int pb = kb - SF_H - 1, pe = ke + SF_H + 1;   // first / last plane of the table
  for (int t = tid; t < ke - kb + 1 + 2 * SF_H + 2; t += SF_NT) {
    const int m = pb + t;
    const SfPlane P = sf_plane(a, m);
    const int dist = max(max(kb - m, m - ke), 0);
    sm.pl[t] = make_int2(plane * P.kz, (P.live ? 1 : 0) | (P.seam_lo ? 16 : 0) | (P.seam_hi ? 32 : 0) | (dist << 8));
  }
  __syncthreads();
  // per-thread work items, fixed for the whole march
  const SfCell c1 = sf_cell_init<1>(sm, i0, j0), c2 = sf_cell_init<2>(sm, i0, j0),
               c3 = sf_cell_init<3>(sm, i0, j0), c4 = sf_cell_init<4>(sm, i0, j0);
  int o0[SF_N0];                                     // stage 0: staged cells tid, tid+NT, ...
#pragma unroll
  for (int n = 0; n < SF_N0; ++n) {
    const int q = tid + n * SF_NT;
    const int qq = min(q, SF_CELLS - 1), y = qq / SF_W, x = qq - y * SF_W;
    o0[n] = (q < SF_CELLS && sm.gyo[y] >= 0 && sm.gxo[x] >= 0) ? sm.gyo[y] + sm.gxo[x] : -1;
  }
  static_assert(SF_TX * SF_TY == SF_NT, "stage 5: one core cell per thread");
  const int gi5 = i0 + (tid & (SF_TX - 1)), gj5 = j0 + tid / SF_TX;
  const int o5 = (gi5 <= n0 - 2 && gj5 <= n1 - 2) ? gj5 * pitch + gi5 : -1;
  const int si5 = (SF_H + tid / SF_TX) * SF_W + SF_H + (tid & (SF_TX - 1));
  const int osafe = pitch + 1 + plane;
  const float* __restrict__ Lx = a.p.L;
  const float* __restrict__ Ly = a.p.L + l.cstride;
  const float* __restrict__ Lz = a.p.L + 2 * l.cstride;
  double s1 = 0.0; float mx = -INFINITY;

  // stage s works on plane m - 2s: what a stage reads was written at least one step (one barrier) ago
  for (int m = kb - SF_H; m <= ke + SF_H + 10; ++m) {
    if (m <= ke + SF_H) {                                              // stage 0: ϵ = r·iD  (Poisson.jl:142)
      const int2 P = sm.pl[m - pb];
      float* dst = sm.ring[m & SF_RM];
      const bool live = P.y & 1;
      float vr[SF_N0], vd[SF_N0];
#pragma unroll
      for (int n = 0; n < SF_N0; ++n) {
        const int o = (live && o0[n] >= 0) ? o0[n] + P.x : osafe;
        vr[n] = __ldg(a.p.r + o); vd[n] = __ldg(a.p.iD + o);
      }
#pragma unroll
      for (int n = 0; n < SF_N0; ++n)
        if (tid + n * SF_NT < SF_CELLS) dst[tid + n * SF_NT] = (live && o0[n] >= 0) ? vr[n] * vd[n] : 0.f;
    }
    // sweeps: the table's distance field switches a stage off outside its trapezoid; planes before the
    // table's first entry only occur while the pipeline fills (index clamped, entry says "far")
    sf_sweep<1>(a, sm, c1, min(max(m - 2, pb), pe), pb);
    sf_sweep<2>(a, sm, c2, min(max(m - 4, pb), pe), pb);
    sf_sweep<3>(a, sm, c3, min(max(m - 6, pb), pe), pb);
    sf_sweep<4>(a, sm, c4, min(max(m - 8, pb), pe), pb);
    {                                                                  // stage 5: increment! (Poisson.jl:100-104)
      const int kp = m - 10;
      const bool ok = kp >= kb && kp <= ke && o5 >= 0;
      const float* cur = sm.ring[kp & SF_RM] + si5;
      const float* dn = sm.ring[(kp - 1) & SF_RM] + si5;
      const float* up = sm.ring[(kp + 1) & SF_RM] + si5;
      const int o = ok ? o5 + plane * kp : osafe;
      const float lx0 = __ldg(Lx + o), lx1 = __ldg(Lx + o + 1), ly0 = __ldg(Ly + o), ly1 = __ldg(Ly + o + pitch),
                  lz0 = __ldg(Lz + o), lz1 = __ldg(Lz + o + plane);
      const float r0 = __ldg(a.p.r + o), x0 = a.p.x[o], e0 = cur[0];
      float dg = 0.f;                                                  // diag(I,L)  Poisson.jl:49-55
      dg -= lx0 + lx1; dg -= ly0 + ly1; dg -= lz0 + lz1;
      float s = e0 * dg;                                               // mult(I,L,D,ϵ)  Poisson.jl:70-76
      s += cur[-1] * lx0 + cur[1] * lx1;
      s += cur[-SF_W] * ly0 + cur[SF_W] * ly1;
      s += dn[0] * lz0 + up[0] * lz1;
      const float rr = r0 - a.omega * s;
      if (ok) {
        a.rn[o] = rr;
        a.p.x[o] = x0 + a.omega * e0;
        s1 += (double)fabsf(rr); mx = fmaxf(mx, fabsf(rr));
      }
    }
    __syncthreads();
  }
  if (a.want_norms) block_reduce_finish(s1, mx, a.w, a.out);
}

}  // namespace wl
