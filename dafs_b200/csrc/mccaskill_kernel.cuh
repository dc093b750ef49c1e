// K3/K4: McCaskill inside / outside base-pairing probabilities in scaled FP64,
// one CTA -- or one thread-block cluster when there are fewer sequences than SMs --
// per sequence (persistent grid), span-length diagonals separated by one block /
// cluster barrier each.  Replaces what RNAfold::calculate
// (src/fold.cpp:78-116) gets from ViennaRNA's pf_fold + export_bppm, including
// its read-out: float(p) > th kept, rows i < j, ascending j (:108-114).
//
// Quantities (1-based, W = n_max + 2 row stride):
//   qb(i,j)  = hairpin + sum interior + MLclose(i,j) * sum_k qm(i+1,k-1) * qm1(k,j-1)
//   qm1(i,j) = qm1(i,j-1)*mlb + qb(i,j)*MLstem(i,j)           (ViennaRNA's rolling qqm)
//   qm(i,j)  = sum_{k=i..j} (mlb^(k-i) + qm(i,k-1)) * qm1(k,j)  =  A(i,j) + R(i,j)  with
//              A(i,j) = sum_{k=i+1..j} qm(i,k-1)*qm1(k,j)   and   R(i,j) = qm1(i,j) + mlb*R(i+1,j);
//              the multiloop closing sum of qb(i,j) is then just A(i+1,j-1)
//   exterior: only q(1,j) =: q5[j] and q(i,n) =: q3[i] are ever used (Z = q5[n] and the
//              first term of O), so instead of ViennaRNA's O(n^3) q(i,j) table
//              q5[j] = q5[j-1]*sc + sum_k q5[k-1] * qb(k,j)*Ext(k,j)     (j unpaired | j pairs with k)
//              q3[i] = q3[i+1]*sc + sum_k qb(i,k)*Ext(i,k) * q3[k+1]     (i unpaired | i pairs with k)
//              -- the same sums over structures, O(n^2) in total, evaluated by the cell (1,j) /
//              (i,n) of the diagonal that completes column j / row i of qb
//   O(i,j)   = q5[i-1]*q3[j+1]*Ext(i,j) + sum_{interior (p,q)} O(p,q)*wI*sc^(..)
//            + MLstem(i,j) * [ sum_{p<i-1} PL(p,j)*qm(p+1,i-1) + tB(i,j) ],  PL = prm_l + prml,
//              tB(i,j) = prml(i-1,j) + mlb*tB(i-1,j)
//   prm_l(p,j) = prm_l(p,j+1)*mlb + OM(p,j+1),  prml(p,j) = sum_{q>j+1} OM(p,q)*qm(j+1,q-1),
//   OM = O*MLclose*sc^2,   P(i,j) = qb*O / Z.
//
// Work layout: a cell of the diagonal belongs to a GROUP of GL lanes (GL = 8 for sequences up to
// 480 nt: four cells per warp; 16 up to 1100 nt; whole warps beyond): the sums of a cell run over
// one index with the group striding it, the serial parts (hairpin, the seven small interior loops,
// the epilogue with its table look-ups) occupy one lane of eight instead of one of thirty-two, and
// four times as many cells are in flight per SM.  Cells are handed out from a per-diagonal list
// that holds the cells whose ends can pair first, so the groups of a warp mostly do the same work.
// Arrays that are read along their first index are kept transposed (qm1T, qmT, E5T, GT, PLT, HT)
// so that every stream is contiguous; quantities that are only read one or two diagonals later
// (A, R, prm_l, prml, tB, the V rows of the interior-loop recurrence) live in rolling per-diagonal
// rows instead of W x W arrays.  Eight W x W FP64 arrays per team remain (was thirteen): qb qm qmT
// O, and qm1T / E5T / G / GT whose storage the outside pass reuses for OM / PLT / H / HT.
// Launch forms (mccaskill.cu): at least one sequence per SM -> two CTAs of 384 threads per SM, each
// with its own sequence (one sequence's diagonal barrier is covered by the other's work); fewer ->
// one CTA of 768 threads, in a thread-block cluster per sequence.
#pragma once
#include <cooperative_groups.h>

#include "mccaskill_model.cuh"

namespace dafs {

struct McArgs {
  const uint8_t* seq;          // nucleotide codes 0..4, concatenated
  const int64_t* seq_off;      // [nseq]
  const int32_t* seq_len;      // [nseq]
  const int32_t* order;        // queue position -> sequence (longest first)
  int nseq;
  int cluster;                 // CTAs per sequence (= the launch's cluster size)
  int W;                       // row stride of the DP arrays = max n + 2
  float th;
  const uint16_t* cons_cls;    // hard constraints (NULL: none), per nucleotide like seq: bit 15 = must stay
                               // unpaired, bit 14 = has a given partner, low 14 bits = id of the loop of the
                               // given pairs the position lies in; see mc_allowed
  const uint16_t* cons_prt;    // 1-based partner of a position with a given pair, else 0
  const McTables* tab;         // in HBM; staged to shared memory by every CTA
  double* ws;                  // per-team workspace: MC_NARRAYS arrays of W*W doubles + MC_NROWS rows of W
  size_t ws_stride;            // doubles per team
  // outputs
  const int64_t* row_off;      // [nseq+1] start of each sequence's row_ptr block (n+1 entries)
  int32_t* row_ptr;
  const int64_t* pool_off;     // [nseq+1] start of each sequence's region in pool_col / pool_val
  uint32_t* pool_col;
  float* pool_val;
  int32_t* seq_nnz;            // [nseq]
  double* seq_lnz;             // [nseq] ln Z (diagnostic, = -G_ensemble/kT)
};

#ifndef MC_THREADS
#define MC_THREADS 768      // 80 registers per thread: one CTA of 24 warps per SM
#endif
#ifndef MC_UNR
#define MC_UNR 4            // pairs of stream loads a lane has in flight in the O(span) sums
#endif
#define MC_WARPS (MC_THREADS / 32)   // most warps a CTA can have; the launch may use fewer (blockDim.x)

__device__ __forceinline__ double mc_warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
template <int GL>
__device__ __forceinline__ double mc_group_sum(double v) {
#pragma unroll
  for (int o = GL / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// The CTAs of one thread-block cluster share a sequence: entry e of a diagonal's cell list
// belongs to CTA e % nc, the per-diagonal barrier is the cluster barrier (release /
// acquire at cluster scope, so the global-memory DP arrays written by one CTA are
// visible to the others), a lone CTA uses the block barrier.
struct McTeam {
  int rank, nc;
};
__device__ __forceinline__ void mc_sync(const McTeam& tm) {
  if (tm.nc == 1) __syncthreads();
  else cooperative_groups::this_cluster().sync();
}

struct McWs {
  double *qb, *qm, *qmT, *O;
  double *qm1T, *E5T, *G, *GT; // inside only
  double *OM, *PLT, *H, *HT;   // outside only: they reuse the storage of qm1T, E5T, G, GT
  double *Vroll;               // V rows of the last three diagonals: [3][W][32]
  // rows of W doubles
  double *q5, *q3;             // exterior partition functions of prefixes / suffixes
  double *Aroll, *Rroll;       // inside: A of the last three diagonals, R of the last two (indexed by i)
  double *plroll, *prroll, *tbroll;   // outside: prm_l, prml, tB of the last two diagonals (indexed by i)
  double *far0, *far1;         // far parts of the O(span) sums of the current band of diagonals: [MC_NB][W] each
};
#define MC_NARRAYS 8
#define MC_NB 16             // diagonals per band (see mc_far_tile)
#define MC_TI 32             // cells per diagonal in a tile of the far pass
#define MC_KC 32             // terms staged per round of the far pass
#define MC_FAR_ROWS (MC_TI + MC_TI + MC_NB - 1)                         // staged rows: P | D
#define MC_FAR_STAGE (MC_KC * (MC_TI + 1) + MC_KC * (MC_TI + MC_NB))   // doubles of shared memory (also holds 5 x 512 partial sums)
#define MC_NROWS (14 + 96 + 2 * MC_NB)   // q5 q3 | Aroll x3 Rroll x2 | plroll x2 prroll x2 tbroll x2 (+1 spare) | Vroll 3 x 32 | far0, far1

// Hard constraints (RNAfold::calculate(seq, str, bp), src/fold.cpp:118-166): ViennaRNA applies the
// dot-bracket constraint by removing every pair that conflicts with it.  Pair (i,j) survives when it
// is a given pair, or when both ends are unconstrained and lie in the same loop of the given pairs
// (then no given pair separates them, i.e. none is crossed).
struct McCons {
  const uint16_t* cls;         // shared memory, 1-based; NULL = unconstrained
  const uint16_t* prt;
};
__device__ __forceinline__ bool mc_allowed(const McCons& c, int i, int j) {
  if (!c.cls) return true;
  const unsigned ci = c.cls[i], cj = c.cls[j];
  if (c.prt[i] == (uint16_t)j) return true;
  return ((ci | cj) >> 14) == 0 && ci == cj;
}

// Per-CTA scheduling state in static shared memory.
struct McSched {
  int next[2];                 // hand-out counter of the diagonal (double-buffered)
  int wcnt[MC_WARPS];          // per-warp counts of the cell list being built
};

// Interior loops.  A loop with u1, u2 unpaired bases is "generic" when ViennaRNA's
// exp_E_IntLoop takes its last branch, internal[u1+u2] * mismatchI(outer) *
// mismatchI(inner) * ninio[|u1-u2|]: every (u1,u2) with min >= 1 except 1x1, 1x2,
// 2x1, 2x2.  With
//   G(k,l) = qb(k,l) * mismatchI[rtype(k,l)][S[l+1]][S[k-1]]      (inside)
//   H(p,q) = O(p,q)  * mismatchI[type(p,q)][S[p+1]][S[q-1]]       (outside)
// the generic part of a cell is mismatchI(outer) * sum_s internal[s]*scale[s+2] * V(i,j,s),
//   V(i,j,s) = sum_{u1+u2=s, u1,u2>=1} ninio[|u1-u2|] * G(i+1+u1, j-1-u2),
// and growing a loop by one unpaired base on EACH side keeps |u1-u2|, so
//   V(i,j,s) = V(i+1,j-1,s-2) + ninio[s-2] * ( G(i+2, j-s) + G(i+s, j-2) )      (one term for s = 2)
// -- O(30) per cell from the row kept two diagonals ago instead of a 435-term sum per cell (the
// idea of the O(n^3) interior-loop evaluation of Lyngso et al.).  Every cell maintains its V row
// (an enclosing pair needs it whether or not the cell's own ends pair).  The four small loops the
// generic formula does not cover sit at s = 2 (1x1), s = 3 (1x2, 2x1) and in V(i+1,j-1,2) of s = 4
// (2x2), so the combination takes nothing from s = 2, 3 and only the two new terms at s = 4.
// Outside: Vo(i,j,s) = Vo(i-1,j+1,s-2) + ninio[s-2] * ( H(i-2, j+s) + H(i-s, j+2) ) likewise.
// Bulges of 2..30 bases are bulge[u]*scale[u+2] * tAU(outer) * tAU(inner) along one row and one
// column of qb (O outside); the remaining seven loops (stack, 1-bulges, 1x1, 1x2, 2x1, 2x2) go
// through mc_interior on seven lanes.
struct McCoef {              // per-scale coefficient tables in shared memory
  double fgen[32];           // internal[s] * scale[s+2]
  double gnin[32];           // ninio[k]
  double cbul[32];           // bulge[u] * scale[u+2]
};

// The cell list of diagonal d (cells i = 1 .. n-d, j = i+d): cells for which `front(i)` holds are
// packed at the head in ascending i, the others at the tail in descending i.  Built by the whole
// CTA before the barrier that ends the previous diagonal; the order is a pure function of the
// predicate, so every CTA of a cluster builds the same list and takes entries e = rank (mod nc).
// promote_last: the front cell with the largest i moves to the second place (inside pass: cells i = 1 and
// j = n are front cells and carry the exterior sums, so they start first).
template <class F>
__device__ __forceinline__ void mc_build_list(uint16_t* list, int ncell, McSched& sc, F front, bool promote_last = false) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NT = blockDim.x;
  int base = 0;                                  // front cells among those of earlier passes
  for (int b = 0; b < ncell; b += NT) {
    const int c = b + tid;                       // cell i = c + 1
    const bool in = c < ncell;
    const bool f = in && front(c + 1);
    const unsigned mf = __ballot_sync(0xffffffffu, f);
    if (lane == 0) sc.wcnt[warp] = __popc(mf);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int k = 0; k < MC_WARPS; ++k) { const int x = k < (NT >> 5) ? sc.wcnt[k] : 0; total += x; if (k < warp) before += x; }
    const int nf = base + before + __popc(mf & ((1u << lane) - 1u));    // front cells with a smaller index
    if (f) list[nf] = (uint16_t)(c + 1);
    else if (in) list[ncell - 1 - (c - nf)] = (uint16_t)(c + 1);
    base += total;
    if (b + NT < ncell) __syncthreads();          // wcnt is rewritten by the next pass
  }
  if (promote_last && base >= 3) {
    __syncthreads();
    if (tid == 0) { const uint16_t t = list[1]; list[1] = list[base - 1]; list[base - 1] = t; }
  }
}

// ---- the O(span) sums, banded -------------------------------------------------------------------
// A(i,j) = sum_k qm(i,k-1) qm1(k,j) and the two outside sums are dot products of a row that belongs to
// one end of the cell with a row that belongs to the other end.  Evaluated cell by cell they are chains of
// dependent HBM round trips (span / lanes / MC_UNR of them per cell) that stream 2 x span doubles per cell.
// Here the diagonals are taken in bands of MC_NB: the terms whose two factors were complete before the
// band began -- all but fewer than MC_NB at either end of the range -- are summed for the whole band in one
// pass over tiles of MC_TI cells x MC_NB diagonals.  In a tile one factor depends on a (the cell's position
// in the diagonal: MC_TI rows, "P") and the other on a +- b (b = diagonal within the band: MC_TI + MC_NB - 1
// rows, "D"); both are staged in shared memory MC_KC terms at a time, k fastest (coalesced), the next round's
// values being fetched into registers while this round's are used, and every staged value feeds up to MC_NB
// (P) or MC_TI (D) products.  The bounds of a cell's range are a property of ITS P row on one side and of ITS
// D row on the other, so entries outside a row's bounds are staged as zeros and the inner loop has no tests.
// The near terms are left to the cell-by-cell pass (mc_inside / mc_outside), which adds far + near.
//   P entry: MP[(rp0 + a) * W + k + cp]  for k in prange(a),  D entry: MD[r * W + k + cd] for k in drange(r),
//   r = rd0 + a + sgn * b.
template <class FP, class FD, class FS>
__device__ __forceinline__ void mc_far_tile(int W, int nmax, double* stage, const double* __restrict__ MP, int rp0, int cp,
                                            const double* __restrict__ MD, int rd0, int sgn, int cd, int k0, int k1,
                                            FP prange, FD drange, FS store) {
  constexpr int PP = MC_TI + 1, PD = MC_TI + MC_NB, MB = 4;
  constexpr int QM = 10;                               // staged rows per thread: warp, warp + nwp, ... (CTAs of >= 256 threads)
  static_assert(MC_KC == 32, "a lane stages term kk = lane of its rows");
  double* sP = stage;
  double* sD = stage + MC_KC * PP;
  const int tid = threadIdx.x, NT = blockDim.x, nwp = NT >> 5;
  const int a = tid & 31, bw = tid >> 5, uoff = sgn < 0 ? MC_NB - 1 : 0;
  // products: warp (bw & 3) owns diagonals 4 (bw & 3) .. + 3 of the band for all 32 cells of the tile (one P value,
  // four D values, four FMAs per term); the nwp / 4 groups of four warps take the terms of a round in turn and
  // their partial sums meet in shared memory at the end
  static_assert(MC_NB == 16, "four warps x four diagonals");
  const int G = max(nwp >> 2, 1), g = bw >> 2, b4 = (bw & 3) * 4;
  const bool works = g < G;
  double acc[MB];
#pragma unroll
  for (int m = 0; m < MB; ++m) acc[m] = 0.0;
  double nxt[QM];
  auto fetch = [&](int kc) {
    const int k = kc + a;
#pragma unroll
    for (int q = 0; q < QM; ++q) {
      const int row = bw + q * nwp;
      double v = 0.0;
      if (row < MC_FAR_ROWS && k <= k1) {
        int lo, hi;
        if (row < MC_TI) {
          const int r = rp0 + row;
          prange(row, lo, hi);
          if (r >= 0 && r <= nmax + 1 && k >= lo && k <= hi) v = MP[(size_t)r * W + k + cp];
        } else {
          const int r = rd0 + (row - MC_TI) - uoff;
          drange(r, lo, hi);
          if (r >= 0 && r <= nmax + 1 && k >= lo && k <= hi) v = MD[(size_t)r * W + k + cd];
        }
      }
      nxt[q] = v;
    }
  };
  fetch(k0);
  const double* dcol = sD + a + sgn * b4 + uoff;       // + sgn * m per diagonal, + kk * PD per term
  for (int kc = k0; kc <= k1; kc += MC_KC) {
#pragma unroll
    for (int q = 0; q < QM; ++q) {
      const int row = bw + q * nwp;
      if (row < MC_TI) sP[a * PP + row] = nxt[q];
      else if (row < MC_FAR_ROWS) sD[a * PD + (row - MC_TI)] = nxt[q];
    }
    __syncthreads();
    if (kc + MC_KC <= k1) fetch(kc + MC_KC);
    if (works) {
      for (int kk = g; kk < MC_KC; kk += G) {
        const double pv = sP[kk * PP + a];
#pragma unroll
        for (int m = 0; m < MB; ++m) acc[m] += pv * dcol[kk * PD + sgn * m];
      }
    }
    __syncthreads();
  }
  if (G > 1) {                                         // (the last barrier above freed the staging tiles)
    if (works && g > 0) {
#pragma unroll
      for (int m = 0; m < MB; ++m) stage[(g - 1) * (MC_TI * MC_NB) + (b4 + m) * MC_TI + a] = acc[m];
    }
    __syncthreads();
    if (g == 0) {
      for (int h = 1; h < G; ++h) {
#pragma unroll
        for (int m = 0; m < MB; ++m) acc[m] += stage[(h - 1) * (MC_TI * MC_NB) + (b4 + m) * MC_TI + a];
      }
    }
  }
  if (g == 0) {
#pragma unroll
    for (int m = 0; m < MB; ++m) store(a, b4 + m, acc[m]);
  }
  if (G > 1) __syncthreads();                          // the next tile stages into the same memory
}

// near terms of a banded sum: up to 2 * (MC_NB - 1) products x[t] * y[t], t = gl, gl + GL, ..., all loads first
template <int GL, class FX>
__device__ __forceinline__ double mc_near(int nterms, int gl, FX term) {
  constexpr int NI = (2 * MC_NB + GL - 1) / GL;
  double x[NI], y[NI];
#pragma unroll
  for (int u = 0; u < NI; ++u) {
    const int t = gl + GL * u;
    x[u] = 0.0; y[u] = 0.0;
    if (t < nterms) term(t, x[u], y[u]);
  }
  double s = 0.0;
#pragma unroll
  for (int u = 0; u < NI; ++u) s += x[u] * y[u];
  return s;
}

// two such sums in one round of loads: terms t < n1 go to s1, the others to s2
template <int GL, class FX>
__device__ __forceinline__ void mc_near2(int n1, int nterms, int gl, FX term, double& s1, double& s2) {
  constexpr int NI = (2 * MC_NB + GL - 1) / GL;
  double x[NI], y[NI];
#pragma unroll
  for (int u = 0; u < NI; ++u) {
    const int t = gl + GL * u;
    x[u] = 0.0; y[u] = 0.0;
    if (t < nterms) term(t, x[u], y[u]);
  }
  s1 = 0.0; s2 = 0.0;
#pragma unroll
  for (int u = 0; u < NI; ++u) { const double pr = x[u] * y[u]; if (gl + GL * u < n1) s1 += pr; else s2 += pr; }
}

// Inside recursion over all spans for the current scale.  Z = q5[n] afterwards.
template <int GL, bool BANDED>
__device__ void mc_inside(const McTables& T, const uint8_t* S, const McCons& cons, int n, int W, const double* scale,
                          const double* mlb, const McCoef& cf, McSched& sc, uint16_t* lists, const McWs& w,
                          double* stage, const McTeam& tm, int lane) {
#define IX(i, j) ((size_t)(i) * W + (j))
  constexpr int WG = 32 / GL;         // cells per warp
  const int gl = lane & (GL - 1), grp = lane / GL, NT = blockDim.x;
  // The banded evaluation (mc_far_tile) pays when one CTA owns the sequence; a cluster that shares one has few
  // tiles per CTA and band, and its whole-warp groups already keep many loads in flight per cell.
  constexpr bool banded = BANDED;                            // chosen by the host: one CTA of >= 256 threads per sequence
  if (threadIdx.x < 2) sc.next[threadIdx.x] = 0;
  if (tm.rank == 0 && threadIdx.x == 0) { w.q5[0] = 1.0; w.q3[n + 1] = 1.0; }
  __syncthreads();
  mc_build_list(lists, n, sc, [&](int) { return true; });
  mc_sync(tm);
  for (int d = 0; d < n; ++d) {
    const int ncell = n - d;
    const int d0 = banded ? (d & ~(MC_NB - 1)) : 0;             // first diagonal of the band (0: no far part)
    if (banded && d == d0 && d0 > 0) {
      // far part of A for the cells of diagonals d0 .. d0 + MC_NB - 1: k in [j - d0 + 1, i + d0]
      // (both qm(i,k-1) and qm1(k,j) span less than d0)
      for (int t = tm.rank; t * MC_TI < n - d0; t += tm.nc) {
        const int i0 = 1 + t * MC_TI;
        mc_far_tile(W, n, stage, w.qm, i0, -1, w.qm1T, i0 + d0, +1, 0, i0 + 1, min(i0 + MC_TI - 1 + d0, n),
                    [&](int a, int& lo, int& hi) { lo = i0 + a + 1; hi = i0 + a + d0; },      // never the unwritten lower triangle: 0 x garbage
                    [&](int r, int& lo, int& hi) { lo = r - d0 + 1; hi = r; },                 // k <= j
                    [&](int a, int b, double v) { if (i0 + a + d0 + b <= n) w.far0[b * W + i0 + a] = v; });
      }
      mc_sync(tm);
    }
    const uint16_t* list = lists + (d & 1) * W;
    const int usum = min(MC_MAXLOOP, d - MC_TURN - 3);          // u1 + u2 <= usum keeps l - k > TURN
    double* Ad = w.Aroll + (d % 3) * W;
    const double* Ad2 = w.Aroll + ((d + 1) % 3) * W;            // diagonal d - 2
    double* Rd = w.Rroll + (d & 1) * W;
    const double* Rd1 = w.Rroll + ((d + 1) & 1) * W;            // diagonal d - 1
    double* Vd = w.Vroll + (size_t)(d % 3) * W * 32;
    const double* Vd2 = w.Vroll + (size_t)((d + 1) % 3) * W * 32;   // diagonal d - 2
    for (;;) {
      int h = 0;
      if (lane == 0) h = atomicAdd(&sc.next[d & 1], WG);
      h = __shfl_sync(0xffffffffu, h, 0);
      if (h * tm.nc + tm.rank >= ncell) break;
      const int e = (h + grp) * tm.nc + tm.rank;
      const bool valid = e < ncell;
      const int i = valid ? (int)list[e] : 1;
      const int j = i + d;
      const int type = (valid && mc_allowed(cons, i, j)) ? mc_pair(S[i], S[j]) : 0;
      const bool closes = (d > MC_TURN) && type != 0;
      const bool lead = valid && gl == 0;
      // operands of the group's epilogue, requested now so that their latency hides behind the sums
      double pre_A = 0.0, pre_m1 = 0.0, pre_R = 0.0, pre_far = 0.0;
      if (lead && d0 > 0) pre_far = w.far0[(d - d0) * W + i];
      if (lead && d > 0) {
        pre_m1 = w.qm1T[IX(j - 1, i)];
        pre_R = Rd1[i + 1];
        if (closes && d > 2) pre_A = Ad2[i + 1];
      }
      // V row of this cell (every cell), and the generic interior-loop part of a closing one
      double vb = 0.0;          // hairpin + interior part of qb, lane-partial
      if (usum >= 2) {
        constexpr int NC = 32 / GL;
        const double* Vold = Vd2 + (size_t)(i + 1) * 32;
        double* Vnew = Vd + (size_t)i * 32;
        const double* Grow = w.G + IX(i + 2, j);          // G(i+2, j-s) = Grow[-s]
        const double* GTrow = w.GT + IX(j - 2, i);        // G(i+s, j-2) = GT[j-2][i+s] = GTrow[s]
        double gen = 0.0;
        double g1[NC], g2[NC], vo[NC];           // every load before the first store (the rows may alias for the compiler)
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          const int sz = 2 + gl + GL * c;
          const bool on = valid && sz <= usum;
          g1[c] = on ? Grow[-sz] : 0.0;
          g2[c] = (on && sz > 2) ? GTrow[sz] : 0.0;
          vo[c] = (on && sz >= 4) ? Vold[sz - 2] : 0.0;
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          const int sz = 2 + gl + GL * c;
          if (valid && sz <= usum) {
            const double bnd = (g1[c] + g2[c]) * cf.gnin[sz - 2];
            const double v = bnd + vo[c];
            Vnew[sz] = v;
            gen += cf.fgen[sz] * (sz >= 5 ? v : (sz == 4 ? bnd : 0.0));
          }
        }
        if (closes) vb = (double)T.mmI[type - 1][S[i + 1]][S[j - 1]] * gen;
      }
      if (__any_sync(0xffffffffu, closes)) {
        if (closes) {
          const int si1 = S[i + 1], sj1 = S[j - 1];
          if (gl == GL - 1) {
            const int u = d - 1;
            int s6 = 0;
            if (u == 4) { for (int c = 0; c < 6; ++c) s6 = s6 * 5 + S[i + c]; }
            vb += mc_hairpin(T, u, type, si1, sj1, s6) * scale[u + 2];
          }
          if (usum >= 0) {
            // bulges of 2 .. usum bases: row i+1 and column j-1 of qb
            double bs = 0.0;
            const double* qb_r = w.qb + IX(i + 1, j - 1);
            const double tA = (double)T.termAU;
            for (int u = 2 + gl; u <= usum; u += GL) {
              double r = qb_r[-u];
              if (r != 0.0 && (int)S[i + 1] * (int)S[j - 1 - u] != 6) r *= tA;
              double c = qb_r[(ptrdiff_t)u * W];
              if (c != 0.0 && (int)S[i + 1 + u] * (int)S[j - 1] != 6) c *= tA;
              bs += cf.cbul[u] * (r + c);
            }
            vb += (type > 2 ? tA : 1.0) * bs;
            // stack, 1-bulges, 1x1, 1x2, 2x1, 2x2 on lanes 0..6
            if (gl < 7) {
              const int u1 = (0x2211100 >> (4 * gl)) & 15, u2 = (0x2121010 >> (4 * gl)) & 15;
              if (u1 + u2 <= usum) {
                const int k = i + 1 + u1, l = j - 1 - u2;
                const double v = w.qb[IX(k, l)];
                if (v != 0.0)
                  vb += v * mc_interior(T, u1, u2, type, mc_rtype(mc_pair(S[k], S[l])), si1, sj1, S[k - 1], S[l + 1]) *
                        scale[u1 + u2 + 2];
              }
            }
          }
        }
      }
      // A(i,j): in the first band every k = i+1 .. j; later the d - d0 terms at either end of the range
      // (the far part comes from mc_far_tile)
      double sm = 0.0;
      {
        const double* qm_i = w.qm + IX(i, 0);
        const double* qm1T_j = w.qm1T + IX(j, 0);
        if (d0 == 0) {
          for (int kb = i + 1 + gl; kb <= j; kb += MC_UNR * GL) {
            double a0[MC_UNR], a1[MC_UNR];
#pragma unroll
            for (int u = 0; u < MC_UNR; ++u) {
              const int k = kb + GL * u;
              const bool ok = valid && k <= j;
              a0[u] = ok ? qm_i[k - 1] : 0.0;  a1[u] = ok ? qm1T_j[k] : 0.0;
            }
#pragma unroll
            for (int u = 0; u < MC_UNR; ++u) sm += a0[u] * a1[u];
          }
        } else {
          const int nb = d - d0;                                 // k = i+1 .. j-d0  and  k = i+d0+1 .. j
          sm = mc_near<GL>(valid ? 2 * nb : 0, gl, [&](int t, double& x, double& y) {
            const int k = t < nb ? i + 1 + t : i + d0 + 1 + (t - nb);
            x = qm_i[k - 1]; y = qm1T_j[k];
          });
        }
      }
      vb = mc_group_sum<GL>(vb); sm = mc_group_sum<GL>(sm);
      double e_cur = 0.0;        // qb(i,j) * Ext(i,j)
      if (lead) {
        if (closes)
          vb += pre_A * ((double)T.ml_closing * mc_stem(T, mc_rtype(type), S[j - 1], S[i + 1], true)) * scale[2];
        w.qb[IX(i, j)] = vb;
        const double gij = (vb != 0.0) ? vb * (double)T.mmI[mc_rtype(type) - 1][S[j + 1]][S[i - 1]] : 0.0;
        w.G[IX(i, j)] = gij;
        w.GT[IX(j, i)] = gij;
        const int sl = i > 1 ? (int)S[i - 1] : -1, sr = j < n ? (int)S[j + 1] : -1;
        double m1 = pre_m1 * mlb[1];
        if (vb != 0.0) { m1 += vb * mc_stem(T, type, sl, sr, true); e_cur = vb * mc_stem(T, type, sl, sr, false); }
        w.qm1T[IX(j, i)] = m1;
        w.E5T[IX(j, i)] = e_cur;
        const double r = m1 + pre_R * mlb[1];
        sm += pre_far;
        Ad[i] = sm;
        Rd[i] = r;
        const double vqm = sm + r;
        w.qm[IX(i, j)] = vqm;
        w.qmT[IX(j, i)] = vqm;
      }
      // exterior: the cell that completes column j of qb extends q5, the one that completes row i extends q3.
      // These two cells of every diagonal carry an O(n) sum each; the whole warp takes it (32 lanes, MC_UNR pairs
      // of loads in flight per lane) instead of the cell's own group, or they would be what the barrier waits for.
      const bool ext5 = valid && i == 1, ext3 = valid && j == n;
      const unsigned m5 = __ballot_sync(0xffffffffu, ext5 && gl == 0), m3 = __ballot_sync(0xffffffffu, ext3 && gl == 0);
      if (m5 | m3) {
        double s5 = 0.0, s3 = 0.0;
        if (m5) {                                      // k = 2 .. j - TURN - 1 (k = 1 is this cell)
          const int jj = __shfl_sync(0xffffffffu, j, __ffs(m5) - 1);
          const double* e5 = w.E5T + IX(jj, 0);
          for (int kb = 2 + lane; kb < jj - MC_TURN; kb += 32 * MC_UNR) {
            double x[MC_UNR], y[MC_UNR];
#pragma unroll
            for (int u = 0; u < MC_UNR; ++u) {
              const int k = kb + 32 * u;
              const bool ok = k < jj - MC_TURN;
              x[u] = ok ? w.q5[k - 1] : 0.0; y[u] = ok ? e5[k] : 0.0;
            }
#pragma unroll
            for (int u = 0; u < MC_UNR; ++u) s5 += x[u] * y[u];
          }
          s5 = mc_warp_sum(s5);
        }
        if (m3) {                                      // k = i + TURN + 1 .. n - 1 (k = n is this cell)
          const int ii = __shfl_sync(0xffffffffu, i, __ffs(m3) - 1);
          const double* qb_i = w.qb + IX(ii, 0);
          const int sl = ii > 1 ? (int)S[ii - 1] : -1;
          const int si = S[ii];
          for (int kb = ii + MC_TURN + 1 + lane; kb < n; kb += 32 * MC_UNR) {
            double x[MC_UNR], y[MC_UNR];
#pragma unroll
            for (int u = 0; u < MC_UNR; ++u) {
              const int k = kb + 32 * u;
              const bool ok = k < n;
              x[u] = ok ? qb_i[k] : 0.0; y[u] = ok ? w.q3[k + 1] : 0.0;
            }
#pragma unroll
            for (int u = 0; u < MC_UNR; ++u) {
              const int k = kb + 32 * u;
              if (x[u] != 0.0) s3 += x[u] * mc_stem(T, mc_pair(si, S[k]), sl, (int)S[k + 1], false) * y[u];
            }
          }
          s3 = mc_warp_sum(s3);
        }
        if (lead && ext5) w.q5[j] = w.q5[j - 1] * scale[1] + e_cur + s5;
        if (lead && ext3) w.q3[i] = w.q3[i + 1] * scale[1] + e_cur + s3;
      }
    }
    // next diagonal: reset its counters, build its list (cells that close first)
    if (threadIdx.x == 0) { sc.next[(d + 1) & 1] = 0; }
    if (d + 1 < n) {
      const int dn = d + 1;
      mc_build_list(lists + (dn & 1) * W, n - dn, sc,
                    [&](int i) { return dn <= MC_TURN || i == 1 || i == n - dn || (mc_pair(S[i], S[i + dn]) != 0 && mc_allowed(cons, i, i + dn)); },
                    dn > MC_TURN);
    }
    mc_sync(tm);
  }
#undef IX
}

// Outside recursion; needs the inside arrays of the same scale.
template <int GL, bool BANDED>
__device__ void mc_outside(const McTables& T, const uint8_t* S, int n, int W, const double* scale,
                           const double* mlb, const McCoef& cf, McSched& sc, uint16_t* lists, const McWs& w,
                           double* stage, const McTeam& tm, int lane) {
#define IX(i, j) ((size_t)(i) * W + (j))
  constexpr int WG = 32 / GL;         // cells per warp
  const int gl = lane & (GL - 1), grp = lane / GL, NT = blockDim.x;
  // O, H, OM, PLT start at zero: entries that are never written (pairs that cannot form, q > n) are
  // read as 0; so do the rolling rows.
  for (size_t c = (size_t)tm.rank * NT + threadIdx.x; c < (size_t)(n + 2) * W; c += (size_t)tm.nc * NT) {
    w.O[c] = 0.0; w.H[c] = 0.0; w.HT[c] = 0.0; w.OM[c] = 0.0; w.PLT[c] = 0.0;
  }
  for (int c = tm.rank * NT + threadIdx.x; c < 6 * W; c += tm.nc * NT) w.plroll[c] = 0.0;   // plroll | prroll | tbroll are contiguous
  if (threadIdx.x < 2) sc.next[threadIdx.x] = 0;
  __syncthreads();
  {
    const int d0 = n - 1;
    if (d0 > MC_TURN)
      mc_build_list(lists + (d0 & 1) * W, n - d0, sc,
                    [&](int i) { return mc_pair(S[i], S[i + d0]) != 0 && w.qb[IX(i, i + d0)] != 0.0; });
  }
  mc_sync(tm);
  constexpr bool banded = BANDED;                               // see mc_inside
  int dtop = n - 1;                                             // top diagonal of the current band
  for (int d = n - 1; d > MC_TURN; --d) {
    const int ncell = n - d;
    if (banded && (d == n - 1 || (d & (MC_NB - 1)) == MC_NB - 1)) {
      // Band d .. d - (d mod MC_NB): the terms of the two O(span) sums whose outside factor belongs to a cell of
      // span > dtop (complete):  (a) q >= i + dtop + 1,  (c) p <= j - dtop - 1 (and p <= i - 2).
      dtop = d;
      const int dbot = max(d & ~(MC_NB - 1), MC_TURN + 1);
      for (int t = tm.rank; t * MC_TI < n - dbot; t += tm.nc) {
        const int i0 = 1 + t * MC_TI;
        // (a): P = OM(i, q), D = qm(j+1, q-1), j + 1 = i0 + dtop + 1 + a - b
        mc_far_tile(W, n, stage, w.OM, i0, 0, w.qm, i0 + dtop + 1, -1, -1, i0 + dtop + 1, n,
                    [&](int a, int& lo, int& hi) { lo = i0 + a + dtop + 1; hi = n; },
                    [&](int r, int& lo, int& hi) { lo = r + 1; hi = r <= n ? n : -1; },         // q - 1 >= j + 1
                    [&](int a, int b, double v) { if (dtop - b > MC_TURN && i0 + a + dtop - b <= n) w.far0[b * W + i0 + a] = v; });
        // (c): P = qm(p+1, i-1) = qmT[i-1][p+1], D = PL(p, j) = PLT[j][p], j = i0 + dtop + a - b
        mc_far_tile(W, n, stage, w.qmT, i0 - 1, +1, w.PLT, i0 + dtop, -1, 0, 1, min(i0 + MC_TI - 3, n),
                    [&](int a, int& lo, int& hi) { lo = 1; hi = i0 + a - 2; },
                    [&](int r, int& lo, int& hi) { lo = 1; hi = r <= n ? r - dtop - 1 : -1; },
                    [&](int a, int b, double v) { if (dtop - b > MC_TURN && i0 + a + dtop - b <= n) w.far1[b * W + i0 + a] = v; });
      }
      mc_sync(tm);
    }
    const int bnd_b = dtop - d;                                 // diagonal within the band
    const uint16_t* list = lists + (d & 1) * W;
    double* pl_d = w.plroll + (d & 1) * W;   const double* pl_d1 = w.plroll + ((d + 1) & 1) * W;   // diagonal d + 1
    double* pr_d = w.prroll + (d & 1) * W;   const double* pr_d1 = w.prroll + ((d + 1) & 1) * W;
    double* tb_d = w.tbroll + (d & 1) * W;   const double* tb_d1 = w.tbroll + ((d + 1) & 1) * W;
    double* Vd = w.Vroll + (size_t)(d % 3) * W * 32;
    const double* Vd2 = w.Vroll + (size_t)((d + 2) % 3) * W * 32;   // diagonal d + 2
    for (;;) {
      int h = 0;
      if (lane == 0) h = atomicAdd(&sc.next[d & 1], WG);
      h = __shfl_sync(0xffffffffu, h, 0);
      if (h * tm.nc + tm.rank >= ncell) break;
      const int e = (h + grp) * tm.nc + tm.rank;
      const bool valid = e < ncell;
      const int i = valid ? (int)list[e] : 1;
      const int j = i + d;
      const bool lead = valid && gl == 0;
      const int type = valid ? mc_pair(S[i], S[j]) : 0;
      const double qbij = type ? w.qb[IX(i, j)] : 0.0;
      const bool pairs = type != 0 && qbij != 0.0;
      double pre_pl = 0.0, pre_om = 0.0, pre_prml = 0.0, pre_tB = 0.0, pre_q1 = 0.0, pre_q2 = 0.0, pre_fa = 0.0, pre_fc = 0.0;
      if (lead) {
        if (banded) {
          pre_fa = w.far0[bnd_b * W + i];
          if (pairs) pre_fc = w.far1[bnd_b * W + i];
        }
        if (j + 1 <= n) { pre_pl = pl_d1[i]; pre_om = w.OM[IX(i, j + 1)]; }
        if (i > 1) { pre_prml = pr_d1[i - 1]; pre_tB = tb_d1[i - 1]; }
        if (pairs) { pre_q1 = w.q5[i - 1]; pre_q2 = w.q3[j + 1]; }
      }
      // (a) prml(i,j) = sum_{q = j+2 .. n} OM(i,q) * qm(j+1,q-1)            [every cell]
      // (c) sum_{p = 1 .. i-2} PL(p,j) * qm(p+1,i-1)                        [cells that pair]
      // here only the terms next to the cell (fewer than MC_NB each, one round of loads); the rest is in far0 / far1
      double ta = 0.0, tc = 0.0;
      if (banded) {
        const int na = valid ? max(0, min(i + dtop, n) - (j + 2) + 1) : 0;
        const int p0 = max(1, i - bnd_b);
        const int ncc = pairs ? max(0, (i - 2) - p0 + 1) : 0;
        const double* OM_i = w.OM + IX(i, 0) + (j + 2);          // OM(i, j+2+t)
        const double* qm_j1 = w.qm + IX(min(j + 1, n), 0) + (j + 1);   // qm(j+1, j+1+t)
        const double* pl_j = w.PLT + IX(j, 0) + p0;              // PL(p0+t, j)
        const double* qmT_i1 = w.qmT + IX(i - 1, 0) + p0 + 1;    // qm(p0+1+t, i-1) = qmT[i-1][p0+1+t]
        mc_near2<GL>(na, na + ncc, gl, [&](int t, double& x, double& y) {
          if (t < na) { x = OM_i[t]; y = qm_j1[t]; }
          else { x = pl_j[t - na]; y = qmT_i1[t - na]; }
        }, ta, tc);
      } else {
        // the whole of both sums, fused into one loop of na + nc steps: both operands of a step are contiguous
        const int na = valid ? max(0, n - j - 1) : 0, ncc = pairs ? max(0, i - 2) : 0;
        const double* OM_i = w.OM + IX(i, 0) + (j + 2);          // OM(i, j+2+t)
        const double* qm_j1 = w.qm + IX(min(j + 1, n), 0) + (j + 1);   // qm(j+1, j+1+t)
        const double* pl_j = w.PLT + IX(j, 0) + 1;               // PL(1+t, j)
        const double* qmT_i1 = w.qmT + IX(i - 1, 0) + 2;         // qm(2+t, i-1) = qmT[i-1][2+t]
        const int tot = na + ncc;
        const int wtot = __reduce_max_sync(0xffffffffu, tot);
        for (int tb0 = gl; tb0 < wtot; tb0 += MC_UNR * GL) {
          double x0[MC_UNR], x1[MC_UNR];
          bool first[MC_UNR];
#pragma unroll
          for (int u = 0; u < MC_UNR; ++u) {
            const int t = tb0 + GL * u;
            first[u] = t < na;
            const bool ok = t < tot;
            const int t2 = t - na;
            x0[u] = ok ? (first[u] ? OM_i[t] : pl_j[t2]) : 0.0;
            x1[u] = ok ? (first[u] ? qm_j1[t] : qmT_i1[t2]) : 0.0;
          }
#pragma unroll
          for (int u = 0; u < MC_UNR; ++u) { const double pr = x0[u] * x1[u]; if (first[u]) ta += pr; else tc += pr; }
        }
      }
      // (b) O(i,j): the Vo row of this cell (every cell) and the generic enclosing loops of a pairing one
      double o = 0.0;
      {
        constexpr int NC = 32 / GL;
        const bool has_old = valid && i >= 2 && j + 1 <= n;        // cell (i-1, j+1) of diagonal d + 2 exists
        const double* Vold = Vd2 + (size_t)(i - 1) * 32;
        double* Vnew = Vd + (size_t)i * 32;
        const double* Hrow = w.H + IX(i - 2, j);                   // H(i-2, j+s) = Hrow[s]
        const double* HTrow = w.HT + IX(j + 2, i);                 // H(i-s, j+2) = HT[j+2][i-s] = HTrow[-s]
        double gen = 0.0;
        double g1[NC], g2[NC], vo[NC];           // every load before the first store
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          const int sz = 2 + gl + GL * c;
          const bool on = valid && sz <= MC_MAXLOOP;
          g1[c] = (on && i >= 3 && j + sz <= n) ? Hrow[sz] : 0.0;
          g2[c] = (on && sz > 2 && i - sz >= 1 && j + 2 <= n) ? HTrow[-sz] : 0.0;
          vo[c] = (on && sz >= 4 && has_old) ? Vold[sz - 2] : 0.0;
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          const int sz = 2 + gl + GL * c;
          if (valid && sz <= MC_MAXLOOP) {
            const double bnd = (g1[c] + g2[c]) * cf.gnin[sz - 2];
            const double v = bnd + vo[c];
            Vnew[sz] = v;
            gen += cf.fgen[sz] * (sz >= 5 ? v : (sz == 4 ? bnd : 0.0));
          }
        }
        if (pairs) o = (double)T.mmI[mc_rtype(type) - 1][S[j + 1]][S[i - 1]] * gen;
      }
      if (__any_sync(0xffffffffu, pairs)) {
        if (pairs && i > 1 && j < n) {
          // (p,q) = (i-1-u1, j+1+u2) encloses (i,j)
          const int t2 = mc_rtype(type);
          const int sp1 = S[i - 1], sq1 = S[j + 1];
          const int u1max = min(MC_MAXLOOP, i - 2), u2max = min(MC_MAXLOOP, n - j - 1);
          // bulges of 2 .. 30 bases: row i-1 and column j+1 of O
          double bs = 0.0;
          const double* O_r = w.O + IX(i - 1, j + 1);
          const double tA = (double)T.termAU;
          for (int u = 2 + gl; u <= MC_MAXLOOP; u += GL) {
            double r = u <= u2max ? O_r[u] : 0.0;
            if (r != 0.0 && sp1 * (int)S[j + 1 + u] != 6) r *= tA;
            double c = u <= u1max ? O_r[-(ptrdiff_t)u * W] : 0.0;
            if (c != 0.0 && (int)S[i - 1 - u] * sq1 != 6) c *= tA;
            bs += cf.cbul[u] * (r + c);
          }
          o += (type > 2 ? tA : 1.0) * bs;
          if (gl < 7) {
            const int u1 = (0x2211100 >> (4 * gl)) & 15, u2 = (0x2121010 >> (4 * gl)) & 15;
            if (u1 <= u1max && u2 <= u2max) {
              const int p = i - 1 - u1, qx = j + 1 + u2;
              const double v = w.O[IX(p, qx)];
              if (v != 0.0)
                o += v * mc_interior(T, u1, u2, mc_pair(S[p], S[qx]), t2, S[p + 1], S[qx - 1], sp1, sq1) *
                     scale[u1 + u2 + 2];
            }
          }
        }
      }
      ta = mc_group_sum<GL>(ta); tc = mc_group_sum<GL>(tc); o = mc_group_sum<GL>(o);
      if (lead) {
        if (j + 1 <= n) {
          const double pl = pre_pl * mlb[1] + pre_om;
          ta += pre_fa;
          pr_d[i] = ta;
          pl_d[i] = pl;
          w.PLT[IX(j, i)] = ta + pl;
        }
        const double tb = pre_prml + mlb[1] * pre_tB;
        if (i > 1) tb_d[i] = tb;
        if (pairs) {
          const int sl = i > 1 ? (int)S[i - 1] : -1, sr = j < n ? (int)S[j + 1] : -1;
          o += pre_q1 * pre_q2 * mc_stem(T, type, sl, sr, false);
          o += (tc + pre_fc + tb) * mc_stem(T, type, sl, sr, true);
          w.O[IX(i, j)] = o;
          const double hij = o * (double)T.mmI[type - 1][S[i + 1]][S[j - 1]];
          w.H[IX(i, j)] = hij;
          w.HT[IX(j, i)] = hij;
          w.OM[IX(i, j)] = o * ((double)T.ml_closing * mc_stem(T, mc_rtype(type), S[j - 1], S[i + 1], true)) * scale[2];
        }
      }
    }
    if (threadIdx.x == 0) { sc.next[(d - 1) & 1] = 0; }
    if (d - 1 > MC_TURN) {
      const int dn = d - 1;
      mc_build_list(lists + (dn & 1) * W, n - dn, sc,
                    [&](int i) { return mc_pair(S[i], S[i + dn]) != 0 && w.qb[IX(i, i + dn)] != 0.0; });
    }
    mc_sync(tm);
  }
#undef IX
}

}  // namespace dafs

namespace dafs {

// Dynamic shared memory: McTables | McCoef | staging tiles of the far pass | scale[W+2] | mlb[W+2] |
// lists[2][W] (uint16) | codes[W+8] | constraint classes[W+2], partners[W+2] (uint16)
template <int GL, bool BANDED>
__global__ void __launch_bounds__(MC_THREADS) mccaskill_kernel(const McArgs a) {
  extern __shared__ __align__(16) char smem[];
  __shared__ int s_total;
  __shared__ McSched s_sched;
  McTables& T = *reinterpret_cast<McTables*>(smem);
  const size_t tab_bytes = (sizeof(McTables) + 15) & ~size_t(15);
  McCoef& cf = *reinterpret_cast<McCoef*>(smem + tab_bytes);
  double* stage = reinterpret_cast<double*>(smem + tab_bytes + sizeof(McCoef));
  double* scale = stage + MC_FAR_STAGE;
  double* mlb = scale + (a.W + 2);
  uint16_t* lists = reinterpret_cast<uint16_t*>(mlb + (a.W + 2));
  uint8_t* S = reinterpret_cast<uint8_t*>(lists + 2 * a.W);
  uint16_t* s_cls = reinterpret_cast<uint16_t*>(S + ((a.W + 8 + 1) & ~1));
  uint16_t* s_prt = s_cls + (a.W + 2);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NT = blockDim.x;
  McTeam tm;
  tm.nc = a.cluster;
  tm.rank = tm.nc > 1 ? (int)cooperative_groups::this_cluster().block_rank() : 0;
  const int team = blockIdx.x / tm.nc, nteams = gridDim.x / tm.nc;
  const int gwarp = tm.rank * (NT >> 5) + warp, nwarps = tm.nc * (NT >> 5);
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(a.tab);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&T);
    for (int k = tid; k < (int)(sizeof(McTables) / 4); k += NT) dst[k] = src[k];
  }
  const int W = a.W;
  McWs w;
  {
    double* base = a.ws + (size_t)team * a.ws_stride;
    const size_t A = (size_t)W * W;
    w.qb = base; w.qm = base + A; w.qmT = base + 2 * A; w.O = base + 3 * A;
    w.qm1T = base + 4 * A; w.E5T = base + 5 * A; w.G = base + 6 * A; w.GT = base + 7 * A;
    w.OM = w.qm1T; w.PLT = w.E5T; w.H = w.G; w.HT = w.GT;
    double* rows = base + (size_t)MC_NARRAYS * A;
    w.q5 = rows; w.q3 = rows + W; w.Aroll = rows + 2 * W; w.Rroll = rows + 5 * W;
    w.plroll = rows + 7 * W; w.prroll = rows + 9 * W; w.tbroll = rows + 11 * W;
    w.Vroll = rows + 14 * W;
    w.far0 = rows + (14 + 96) * W; w.far1 = w.far0 + MC_NB * W;
  }
#define IX(i, j) ((size_t)(i) * W + (j))
  // sequences are dealt to the teams round-robin in length order (longest first)
  for (int sq = team; sq < a.nseq; sq += nteams) {
    mc_sync(tm);
    const int s = a.order[sq];
    const int n = a.seq_len[s];
    const uint8_t* gs = a.seq + a.seq_off[s];
    for (int k = tid; k <= n + 1; k += NT) S[k] = (k >= 1 && k <= n) ? gs[k - 1] : (uint8_t)0;
    McCons cons{nullptr, nullptr};
    if (a.cons_cls) {
      const uint16_t* gc = a.cons_cls + a.seq_off[s];
      const uint16_t* gp = a.cons_prt + a.seq_off[s];
      for (int k = tid; k <= n + 1; k += NT) {
        const bool in = k >= 1 && k <= n;
        s_cls[k] = in ? gc[k - 1] : (uint16_t)0; s_prt[k] = in ? gp[k - 1] : (uint16_t)0;
      }
      cons.cls = s_cls; cons.prt = s_prt;
    }
    // ---- scale: start from ViennaRNA's default per-nucleotide guess and move to
    // exp(lnZ/n) when the first inside pass ends far from 1 (the reference takes
    // its pf_scale from an MFE run instead, fold.cpp:86-95; p does not depend on it)
    double lns = 0.30010459245;                        // ln 1.35
    const double ln_mlbase = log((double)T.ml_base);
    double Z = 0.0, lnZ = 0.0;
    for (int attempt = 0; attempt < 8; ++attempt) {
      __syncthreads();
      for (int k = tid; k <= n + 2; k += NT) { scale[k] = exp(-(double)k * lns); mlb[k] = exp((double)k * (ln_mlbase - lns)); }
      if (tid < 32) {
        const int k = tid;
        cf.fgen[k] = k <= MC_MAXLOOP ? (double)T.internal[k] * exp(-(double)(k + 2) * lns) : 0.0;
        cf.gnin[k] = k <= MC_MAXLOOP ? (double)T.ninio[k] : 0.0;
        cf.cbul[k] = k <= MC_MAXLOOP ? (double)T.bulge[k] * exp(-(double)(k + 2) * lns) : 0.0;
      }
      __syncthreads();
      mc_inside<GL, BANDED>(T, S, cons, n, W, scale, mlb, cf, s_sched, lists, w, stage, tm, lane);
      Z = w.q5[n];
      if (isnan(Z) || isinf(Z)) { lns += 0.35; continue; }
      if (!(Z > 0.0)) { lns -= 0.25; continue; }
      lnZ = log(Z) + (double)n * lns;
      if (fabs(log(Z)) <= 150.0) break;
      lns = lnZ / (double)n;                            // rerun with the ensemble's own scale
    }
    mc_outside<GL, BANDED>(T, S, n, W, scale, mlb, cf, s_sched, lists, w, stage, tm, lane);
    // ---- read-out: p = float(qb*O/Z) > th, rows ascending, columns ascending ----
    int32_t* row_ptr = a.row_ptr + a.row_off[s];
    const float th = a.th;
    const double invZ = 1.0 / Z;
    for (int i = 1 + gwarp; i <= n; i += nwarps) {
      int cnt = 0;
      for (int jb = i + 1; jb <= n; jb += 32) {
        const int j = jb + lane;
        bool keep = false;
        if (j <= n) { const float p = (float)(w.qb[IX(i, j)] * w.O[IX(i, j)] * invZ); keep = p > th; }
        cnt += __popc(__ballot_sync(0xffffffffu, keep));
      }
      if (lane == 0) row_ptr[i] = cnt;                  // count of row i-1 (0-based), scanned below
    }
    mc_sync(tm);
    if (gwarp == 0) {                                   // exclusive scan over rows by one warp
      int run = 0;
      for (int b = 1; b <= n; b += 32) {
        const int i = b + lane;
        const int c = i <= n ? row_ptr[i] : 0;
        int x = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (i <= n) row_ptr[i] = run + x;               // inclusive -> row_ptr[i] = end of row i-1
        run += __shfl_sync(0xffffffffu, x, 31);
      }
      if (lane == 0) { row_ptr[0] = 0; s_total = run; }
    }
    mc_sync(tm);
    {
      uint32_t* col = a.pool_col + a.pool_off[s];
      float* val = a.pool_val + a.pool_off[s];
      const int64_t cap = a.pool_off[s + 1] - a.pool_off[s];
      const unsigned lt = (1u << lane) - 1u;
      for (int i = 1 + gwarp; i <= n; i += nwarps) {
        int run = row_ptr[i - 1];
        for (int jb = i + 1; jb <= n; jb += 32) {
          const int j = jb + lane;
          float p = 0.f;
          bool keep = false;
          if (j <= n) { p = (float)(w.qb[IX(i, j)] * w.O[IX(i, j)] * invZ); keep = p > th; }
          const unsigned m = __ballot_sync(0xffffffffu, keep);
          if (keep) { const int k = run + __popc(m & lt); if (k < cap) { col[k] = (uint32_t)(j - 1); val[k] = p; } }
          run += __popc(m);
        }
      }
      if (tid == 0 && tm.rank == 0) { a.seq_nnz[s] = s_total; a.seq_lnz[s] = lnZ; }
    }
  }
#undef IX
}

}  // namespace dafs
