// K5: the dual-decomposition loop of DAFS::solve_by_dd (src/dafs.cpp:1006-1295)
// entirely on the device: up to t_max iterations of
//   SparseNussinov::decode (x), SparseNussinov::decode (y)   src/nussinov.cpp:207-298
//   SparseNeedlemanWunsch::decode (z)                        src/needleman_wunsch.cpp:255-338
//   multiplier tally / update / step size                    src/dafs.cpp:1099-1291
// with no host round trip.  Bit-exact with the reference's x86-64 build: every
// FP32 operation is rounded on its own, compares are strict in the reference's
// order, the Lagrangian value is summed in CBP order by one thread.
//
// Mapping.  One CTA of three warps per DD problem (persistent grid, atomic work
// queue).  The problem's whole sparse description (a few KB) is staged into
// shared memory once; multipliers, DP rows and traceback codes live there too.
// Warp 0 decodes x, warp 1 decodes y, warp 2 decodes z, concurrently; then all
// threads tally and update the multipliers.  Neither decoder uses a block
// barrier inside its DP:
//   * Nussinov runs on the COMPRESSED grid of live rows x live columns (rows /
//     columns that carry at least one decoder candidate).  A cell in a dead
//     column equals its left neighbour and a cell in a dead row equals the one
//     below (both exactly: the maxima range over identical float sums), so only
//     live x live cells are computed; the traceback re-derives the reference's
//     choices in dead rows / columns from one extra bit per live cell.  Lanes own
//     32 consecutive live rows (a wave, bottom first) and walk the live columns
//     left to right, one column behind the row below, whose values arrive by
//     warp shuffle; bifurcations read the lane's own row and the column's
//     candidate list in the reference's push order (bp[j], descending i).
//   * NW: only across the alignment envelope; p_z and q_z are
//     kept as dense BANDS (row r holds columns ef(r)..el(r) of the envelope, the
//     only cells the recursion ever evaluates), so a step is two indexed loads,
//     three adds and two compares.  There is no gap cost, so a row is a prefix
//     maximum over its band: one warp does a whole row per step (see nw_warp).
// Multipliers are sparse: for x and y the decoder's candidates plus the CBP
// projections c_x / c_y, for z the projection c_z -- nothing else can ever
// become non-zero (see dd.cu).
#pragma once
#include <cfloat>
#include <cstdint>
#include <cuda_runtime.h>

namespace dafs {

#define DD_NOPAIR 0xffffffffu

// Offsets (bytes from the start of the problem's constant block) of one
// Nussinov structure.  Entries [0, nd) are the decoder candidates (span > 2) in
// column order -- ascending j, then descending i, the order in which the
// reference pushes to bp[j] (nussinov.cpp:236-245); entries [nd, n) are CBP
// projections with span <= 2, which only carry multipliers.
struct DDNusHdr {
  int L, n, nd, nI, nJ, pad;
  float w;                     // w_*2*N1/(N1+N2)  (dafs.cpp:1091)
  uint32_t col_ptr;            // int[nJ+1]   candidates of live column b
  uint32_t row_ptr;            // int[nI+1]
  uint32_t row_ent;            // int[nd]     entry ids of live row a, ascending b
  uint32_t ent_ab;             // u32[nd]     a | b << 16 (live row / live column index)
  uint32_t ent_c;              // i32[nd]     a << 16 | (1 + largest live column index with J <= i-1, 0 if none)
  uint32_t ent_ij;             // u32[n]      i | j << 16
  uint32_t ent_p;              // f32[n]
  uint32_t ent_flag;           // u8[n]       bit0: in c_x (dafs.cpp:1037)
  uint32_t rowI;               // u16[nI]     live rows, ascending
  uint32_t colJ;               // u16[nJ]     live columns, ascending
  uint32_t bstart;             // u16[nI]     first live column index with J > I[a]
  uint32_t rowmap;             // u16[L+1]    i -> index of the first live row >= i (nI if none)
  uint32_t colmap;             // u16[L+1]    j -> 1 + index of the last live column <= j (0 if none)
};

struct DDHdr {
  int L1, L2, ncbp, t_max, bytes, stage;   // stage: bytes staged to shared memory = everything before cbp
  float th_s, th_a, eta0;
  int pad0;
  DDNusHdr x, y;
  // z: the envelope, the band of p_z inside it, and the cells of c_z (dafs.cpp:1039-1040)
  int nb, ncz;                 // band cells, |c_z|
  uint32_t z_env;              // uint2[L1+1] {first, second}   (needleman_wunsch.cpp:198-253)
  uint32_t z_boff;             // int[L1+2]   band index of (row r, column k) = boff[r] + k - max(1, first(r))
  uint32_t z_pb;               // f32[nb]     p_z over the band
  uint32_t cz_row_ptr;         // int[L1+1]   c_z cells of row i (0-based), ascending k
  uint32_t cz_col;             // u16[ncz]
  uint32_t cz_bidx;            // i32[ncz]    band index of the cell, -1 outside the envelope
  uint32_t cbp;                // int4 {entry in x, entry in y, c_z id of (i,k), c_z id of (j,l)}; last in the
                               // block and read from HBM: only the tally touches it, coalesced, once per iteration
};

struct DDProbRef {
  const char* cst;             // constant block in HBM
  uint32_t* x_out;             // [L1]
  uint32_t* y_out;             // [L2]
  uint32_t* z_out;             // [L1]
  float* result;               // {score, iters, violated, status} as 4 x 32 bit, then 8 cycle counters (float),
                               // then the completion flag (word 12)
  float* trace;                // optional [t_max*3] {L, eta, violated} per iteration (-v 2, dafs.cpp:1273)
};

// Per-CTA scratch arrays; each lives in shared memory when it fits (host decides,
// in this order) and in a per-CTA global workspace otherwise.
enum {
  DD_CST, DD_Q_X, DD_Q_Y, DD_Q_Z, DD_Q_CZ, DD_ES_X, DD_ES_Y, DD_ROW_X, DD_ROW_Y, DD_BROW_X, DD_BROW_Y, DD_BROW_Z,
  DD_TRZ, DD_TR_X, DD_TR_Y, DD_SS_X, DD_SS_Y, DD_ZRES, DD_STK_X, DD_STK_Y, DD_TAL_X, DD_TAL_Y, DD_TAL_Z,
  DD_SW, DD_NITEMS
};
struct DDLayout {
  uint32_t off[DD_NITEMS];     // byte offset in smem or in the CTA's workspace
  uint32_t in_smem;            // bit per item
  uint32_t tr16;               // Nussinov traceback codes need 16 bits (a column with > 120 candidates)
  int lp1, lp2;                // row stride of the Nussinov row buffers (max live columns, odd)
  int trs1, trs2;              // row stride of the Nussinov traceback arrays (max live columns)
  int ts;                      // NW traceback words (uint2) per row: ceil(widest band / 32)
  int nw_chunked;              // see NwScratch::chunked
  char* ws;                    // global workspace
  size_t ws_stride;            // per CTA
};

struct DDArgs {
  const DDProbRef* prob;
  const int32_t* order;        // queue position -> problem (largest first)
  int nprob;
  int* queue;
  DDLayout lay;
};

// One Nussinov structure as the decoder sees it (pointers into the staged block).
struct DDNus {
  int L, n, nd, nI, nJ;
  float w;
  const int* col_ptr; const int* row_ptr; const int* row_ent;
  const uint32_t* ent_ab; const int* ent_c; const uint32_t* ent_ij; const float* ent_p; const uint8_t* ent_flag;
  const uint16_t* rowI; const uint16_t* colJ; const uint16_t* bstart; const uint16_t* rowmap; const uint16_t* colmap;
  float* ent_q;                // [n] multipliers q_x[i][j]
};

__device__ __forceinline__ DDNus dd_make_nus(const char* blk, const DDNusHdr& h, float* q) {
  DDNus S;
  S.L = h.L; S.n = h.n; S.nd = h.nd; S.nI = h.nI; S.nJ = h.nJ; S.w = h.w; S.ent_q = q;
  S.col_ptr = reinterpret_cast<const int*>(blk + h.col_ptr);
  S.row_ptr = reinterpret_cast<const int*>(blk + h.row_ptr);
  S.row_ent = reinterpret_cast<const int*>(blk + h.row_ent);
  S.ent_ab = reinterpret_cast<const uint32_t*>(blk + h.ent_ab);
  S.ent_c = reinterpret_cast<const int*>(blk + h.ent_c);
  S.ent_ij = reinterpret_cast<const uint32_t*>(blk + h.ent_ij);
  S.ent_p = reinterpret_cast<const float*>(blk + h.ent_p);
  S.ent_flag = reinterpret_cast<const uint8_t*>(blk + h.ent_flag);
  S.rowI = reinterpret_cast<const uint16_t*>(blk + h.rowI);
  S.colJ = reinterpret_cast<const uint16_t*>(blk + h.colJ);
  S.bstart = reinterpret_cast<const uint16_t*>(blk + h.bstart);
  S.rowmap = reinterpret_cast<const uint16_t*>(blk + h.rowmap);
  S.colmap = reinterpret_cast<const uint16_t*>(blk + h.colmap);
  return S;
}

struct NusScratch {
  float* es;        // [nd] value a candidate contributes: dp[i+1][j-1] + s, or -inf when inactive
  float* row;       // [T][32*lp] dp rows (over live columns) of the wave each team warp works on
  float* brow;      // [T][lp] top row of that wave = the row below the next wave's lane 0
  volatile int* prog;   // [T] progress of each team warp: wave << 16 | columns its top row has finished
  float* score_slot;    // dp[0][L-1], written by the warp that owns the top wave
  void* tr;         // [nI*trs] traceback codes: t | down_ge << 7 (or << 15)
  uint32_t* ss;     // [L] result
  uint32_t* stk;    // traceback stack
  int lp, trs;
  bool tr16;
};

// The alignment part: the envelope and the dense bands of p_z / q_z inside it.
struct DDAln {
  int L1, L2;
  float th;
  const uint2* env;            // [L1+1] {first, second}   (needleman_wunsch.cpp:198-253)
  const int* boff;             // [L1+2] band offsets (row r starts at boff[r], column max(1, first(r)))
  const float* pb;             // [nb] p_z over the band
  const float* qb;             // [nb] multipliers q_z over the band (non-zero only on c_z, see dd.cu)
};

struct NwScratch {
  float* rowbuf;    // 2 x [L2+2] dp rows (even / odd row)
  uint2* trz;       // [(L1+1) * ts] traceback codes (0 ' ', 1 'M', 2 'X', 3 'Y') as bit planes {low, high}
                    //               per 32 band columns
  uint32_t* zres;   // [L1] result
  int ts;           // words per row = ceil(widest band / 32)
  int chunked;      // 1: 32 columns per scan even where one scan per row would do (DAFS_DD_NW_CHUNKED, A/B runs)
};

}  // namespace dafs

namespace dafs {

__device__ __forceinline__ void nus_tr_store(const NusScratch& sc, int idx, int t, bool flag) {
  if (sc.tr16) reinterpret_cast<uint16_t*>(sc.tr)[idx] = (uint16_t)(t | (flag ? 0x8000 : 0));
  else reinterpret_cast<uint8_t*>(sc.tr)[idx] = (uint8_t)(t | (flag ? 0x80 : 0));
}
// returns t in the low bits and the down_ge flag in bit 16
__device__ __forceinline__ int nus_tr_load(const NusScratch& sc, int idx) {
  if (sc.tr16) { const int v = reinterpret_cast<const uint16_t*>(sc.tr)[idx]; return (v & 0x7fff) | ((v & 0x8000) << 1); }
  const int v = reinterpret_cast<const uint8_t*>(sc.tr)[idx];
  return (v & 0x7f) | ((v & 0x80) << 9);
}

// SparseNussinov::decode(w, p, q, ss): fill + traceback by one warp on the
// compressed grid.  With f(i,j) the reference's dp[i][j], I[] / J[] the live rows
// / columns and D[a][b] = f(I[a], J[b]) (0 where I[a] >= J[b]):
//   f(i+1, J[b])   = D[a+1][b]        ("below":  value of the lane one row down)
//   f(i, J[b]-1)   = D[a][b-1]        ("left":   this lane's previous value)
//   f(i+1, J[b]-1) = D[a+1][b-1]      ("diag")
//   f(i, k-1)      = D[a][c(k)]       (c(k) = last live column <= k-1, ent_c)
// Returns dp[0][L-1] in every lane.
template <int T>
__device__ __forceinline__ float nussinov_warp(const DDNus& S, const NusScratch& sc, float th, int lane, int m, int bar_id,
                                               unsigned long long* prof) {
  // T > 1: a TEAM of T warps fills the table.  Waves (32 live rows, bottom first) are dealt
  // round-robin; wave v may work on live column b once wave v-1 has finished b in its top
  // row (then every row below has: its own lanes are ahead of lane 31, and it waited for
  // the wave below it in turn).  The top row of a wave is mirrored into brow[m] as it is
  // produced; progress goes through prog[m] (shared memory, fence before the flag).  The
  // waves overlap as a pipeline: nJ + 32 * waves steps instead of waves * (nJ + 31).
  const long long t0 = clock64();
  const int nI = S.nI, nJ = S.nJ, lp = sc.lp, trs = sc.trs;
  float* myrow = sc.row + (size_t)m * 32 * lp + lane * lp;
  float* mybrow = sc.brow + (size_t)m * lp;
  const float* below_row = sc.brow + (size_t)((m + T - 1) % T) * lp;
  volatile int* prog_mine = sc.prog + m;
  volatile int* prog_prev = sc.prog + (m + T - 1) % T;
  int seen = 0;
  int wv = 0;
  if (m == 0 && lane == 0) *sc.score_slot = 0.0f;            // before wave 0, which every later wave waits for
  for (int ahi = nI - 1; ahi >= 0; ahi -= 32, ++wv) {
    if (T > 1 && (wv % T) != m) continue;
    const int a = ahi - lane;
    const bool live = a >= 0;
    int i = 0, bs = 0x3fffffff, rcur = 0, rend = 0;
    if (live) { i = S.rowI[a]; bs = S.bstart[a]; rcur = S.row_ptr[a]; rend = S.row_ptr[a + 1]; }
    int next_b = (rcur < rend) ? (int)(S.ent_ab[S.row_ent[rcur]] >> 16) : 0x7fffffff;
    // first column of the live row just below lane 0 (nothing below the bottom wave)
    const int bsb = (wv > 0) ? (int)S.bstart[ahi + 1] : 0x3fffffff;
    // lane t works on live column B0 + s - t at step s; every lane must reach its
    // first valid column bs at some s >= 0
    int B0 = live ? bs + lane : 0x3fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) B0 = min(B0, __shfl_xor_sync(0xffffffffu, B0, o));
    const int nsteps = (nJ - B0) + 31;
    const int prev_base = (wv - 1) << 16, my_base = wv << 16;
    if (T > 1) {
      if (lane == 0) *prog_mine = my_base;
      if (wv > 0) {                                  // column B0 of the row below
        const int tgt = prev_base + min(B0 + 1, nJ);
        if (seen < tgt) { do { seen = *prog_prev; } while (seen < tgt); __threadfence_block(); }
      }
    }
    float cur = 0.0f;          // D[a][b-1] entering a step, D[a][b] leaving it
    float below_prev = 0.0f;   // D[a+1][b-1]
    float brow_next = (lane == 0 && B0 >= bsb && B0 < nJ) ? below_row[B0] : 0.0f;
    for (int s = 0; s < nsteps; ++s) {
      const int b = B0 + s - lane;
      if (T > 1 && wv > 0) {                         // lane 0 fetches column B0 + s + 1 of the row below now
        const int tgt = prev_base + min(B0 + s + 2, nJ);
        if (seen < tgt) { do { seen = *prog_prev; } while (seen < tgt); __threadfence_block(); }
      }
      float below = __shfl_up_sync(0xffffffffu, cur, 1);
      if (lane == 0) { below = brow_next; brow_next = (b + 1 >= bsb && b + 1 < nJ) ? below_row[b + 1] : 0.0f; }
      if (live && b >= bs && b < nJ) {
        const int j = S.colJ[b];
        const int span = j - i;
        float v = 0.0f;
        int t = 0;
        if (span >= 2) { v = below; t = 1; }                 // nussinov.cpp:224-228
        if (span >= 2 && v < cur) { v = cur; t = 2; }        // :229-233
        if (b == next_b) {                                   // (i, j) is a candidate (:234)
          const int e = S.row_ent[rcur];
          ++rcur;
          next_b = (rcur < rend) ? (int)(S.ent_ab[S.row_ent[rcur]] >> 16) : 0x7fffffff;
          const float sv = __fsub_rn(__fmul_rn(S.w, __fsub_rn(S.ent_p[e], th)), S.ent_q[e]);   // :236
          float c = -INFINITY;
          if (sv > 0.0f) {
            c = __fadd_rn(below_prev, sv);                   // :239
            if (v < c) { v = c; t = 3; }                     // :240-244
          }
          sc.es[e] = c;
        }
        // bifurcations: candidates of column j with k > i, in push order (:247-259).
        // The list is in descending row order, so the candidates with k > i are a prefix.
        // Four at a time: the packed (row, column) words and the candidate values are
        // fetched first, the row lookups next, and only the compare chain is serial.
        const int cb = S.col_ptr[b], ce = S.col_ptr[b + 1];
        for (int e = cb; e < ce; e += 4) {
          int w[4]; float sv[4], lv[4]; bool ok[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const bool in = e + u < ce;
            w[u] = in ? S.ent_c[e + u] : 0;                  // row 0 is never below a live row
            sv[u] = in ? sc.es[e + u] : 0.0f;
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            ok[u] = (w[u] >> 16) > a;
            const int cc = (w[u] & 0xffff) - 1;
            lv[u] = (ok[u] && cc >= bs) ? myrow[cc] : 0.0f;  // dp[i][k-1]
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const float c = __fadd_rn(lv[u], sv[u]);
            if (ok[u] && v < c) { v = c; t = 4 + (e + u - cb); }
          }
          if (!ok[3]) break;
        }
        myrow[b] = v;
        if (lane == 31) mybrow[b] = v;                       // the row below the next wave's lane 0
        nus_tr_store(sc, a * trs + b, t, !(below < v));
        cur = v;
        if (a == 0 && b == nJ - 1) *sc.score_slot = v;
      }
      below_prev = below;
      __syncwarp();
      if (T > 1 && ((s & 3) == 3)) {                         // lane 31 has finished every column <= B0 + s - 31
        __threadfence_block();
        if (lane == 0) *prog_mine = my_base + max(0, min(B0 + s - 30, nJ));
      }
    }
    if (T > 1) {
      __threadfence_block();
      if (lane == 0) *prog_mine = my_base + 0xffff;          // the whole wave is done
    }
    __syncwarp();
  }
  if (T > 1) {                                               // team barrier: 1 = x team, 2 = y team
    if (bar_id == 1) asm volatile("bar.sync 1, %0;" :: "r"(32 * T) : "memory");
    else asm volatile("bar.sync 2, %0;" :: "r"(32 * T) : "memory");
  } else __syncwarp();
  const float score = (nI > 0) ? *sc.score_slot : 0.0f;
  const long long t1 = clock64();
  if (T > 1 && m != 0) return score;                         // the team's first warp walks back

  // ---- traceback (:265-295) on full coordinates, all lanes in lock step.  The
  // order in which the reference's stack is served does not change ss.
  for (int k = lane; k < S.L; k += 32) sc.ss[k] = DD_NOPAIR;
  __syncwarp();
  int sp = 0;
  int i = 0, j = S.L - 1;
  for (;;) {
    bool pop = false;
    if (i >= j) pop = true;
    else {
      const int a = S.rowmap[i];
      const int bj = (int)S.colmap[j] - 1;
      if (a >= nI || bj < 0) pop = true;                    // no candidate inside [i, j]
      else {
        const int ia = S.rowI[a], jb = S.colJ[bj];
        if (ia >= j || jb <= ia) pop = true;                // dead rows: t = 1 down to span 1; or nothing to pair
        else {
          i = ia;                                           // t = 1 through the dead rows above I[a]
          const int code = nus_tr_load(sc, a * trs + bj);
          if (jb != j) {
            // dead column: t = 2 iff f(i+1, j-1) < f(i, j-1), both equal their values at J[bj]
            if (code & 0x10000) i = i + 1; else j = jb;
          } else {
            const int t = code & 0xffff;
            if (t == 0) pop = true;
            else if (t == 1) i = i + 1;
            else if (t == 2) j = j - 1;
            else if (t == 3) { if (lane == 0) sc.ss[i] = (uint32_t)j; ++i; --j; }
            else {
              const int e = S.col_ptr[bj] + (t - 4);
              const int k = (int)(S.ent_ij[e] & 0xffffu);
              if (lane == 0) { sc.ss[k] = (uint32_t)j; sc.stk[sp] = (uint32_t)(k + 1) | ((uint32_t)(j - 1) << 16); }
              ++sp;
              j = k - 1;
            }
          }
        }
      }
    }
    if (pop) {
      if (sp == 0) break;
      --sp;
      __syncwarp();
      const uint32_t w = sc.stk[sp];
      __syncwarp();
      i = (int)(w & 0xffffu); j = (int)(w >> 16);
    }
  }
  __syncwarp();
  if (prof && lane == 0) { const long long t2 = clock64(); prof[0] += t1 - t0; prof[1] += t2 - t1; }
  return score;
}

}  // namespace dafs

namespace dafs {

// NW fill when every band of the problem is at most 32 columns wide (alignments of close
// homologs): lane l IS band column ef(r) + l, and the row above stays in registers -- its
// cells dp[r-1][k] and dp[r-1][k-1] come by two independent shuffles from the lanes that
// held them (the band moved right by ef(r) - ef(r-1) columns), not through a row buffer in
// shared memory; the band data of row r+1 is requested before row r's scan.  Same
// operands, compares and bits as the chunked row in nw_warp.
__device__ __forceinline__ float nw_fill_narrow(const DDAln& A, const NwScratch& sc, int lane) {
  const int L1 = A.L1, L2 = A.L2;
  const float th = A.th;
  uint2* trow = sc.trz;
  float vprev = 0.0f, cprev = 0.0f, v = 0.0f;              // row above: dp[r-1][efp + lane], dp[r-1][efp - 1]
  int efp = 1, wprev = 31;                                 // its first column and width - 1
  uint2 ev = A.env[1];
  int ef = max(1, (int)ev.x), el = (int)ev.y;
  int bidx = A.boff[1] + min(lane, max(el - ef, 0));       // band index of (r, ef + lane), clamped into the row
  float pv = A.pb[bidx], qv = A.qb[bidx];
  for (int r = 1; r <= L1; ++r) {
    // row r + 1: envelope and band data, nothing of it depends on the recursion
    const int rn = min(r + 1, L1);
    const uint2 evn = A.env[rn];
    const int efn = max(1, (int)evn.x), eln = (int)evn.y;
    const int bn = A.boff[rn] + min(lane, max(eln - efn, 0));
    const float pvn = A.pb[bn], qvn = A.qb[bn];
    const bool in = ef + lane <= el;
    float up = 0.0f, dg = 0.0f;                            // row 0 is 0 everywhere (:270-274)
    if (r > 1) {
      const int ju = lane + (ef - efp), jd = ju - 1;       // lanes of the row above that hold columns k and k - 1
      const float su = __shfl_sync(0xffffffffu, vprev, ju & 31);
      const float sd = __shfl_sync(0xffffffffu, vprev, jd & 31);
      up = (ju <= wprev) ? su : -FLT_MAX;                  // right of the band above: never filled
      dg = (jd < 0) ? cprev : ((jd <= wprev) ? sd : -FLT_MAX);
    }
    const float m = __fadd_rn(__fsub_rn(__fadd_rn(dg, pv), th), qv);               // :282
    const bool c2 = m < up;                                                        // :284
    float x = in ? (c2 ? up : m) : -FLT_MAX;
    const float a = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) x = fmaxf(x, __shfl_up_sync(0xffffffffu, x, d));
    const float carry = (ef == 1) ? 0.0f : -FLT_MAX;       // dp[r][ef-1]
    v = fmaxf(x, carry);                                                           // dp[r][k]
    float left = __shfl_up_sync(0xffffffffu, v, 1);                                // dp[r][k-1]
    if (lane == 0) left = carry;
    const bool c3 = a < left;                                                      // :289
    const unsigned lo = __ballot_sync(0xffffffffu, in && (c3 || !c2));             // codes 1 and 3
    const unsigned hi = __ballot_sync(0xffffffffu, in && (c3 || c2));              // codes 2 and 3
    if (lane == 0) trow[r] = make_uint2(lo, hi);           // ts == 1
    vprev = v; cprev = carry; efp = ef; wprev = el - ef;
    ef = efn; el = eln; pv = pvn; qv = qvn;
  }
  __syncwarp();
  return __shfl_sync(0xffffffffu, v, (L2 - efp) & 31);     // el(L1) = L2: dp[L1][L2] sits in lane L2 - ef(L1)
}

// One row of the NW fill when every band of the problem is at most 32 CN columns wide:
// lane l takes the CN consecutive band columns ef + CN l .. ef + CN l + CN - 1, so the
// row's prefix maximum is ONE five-shuffle scan over the lanes' own maxima instead of one
// scan per 32 columns (the scans of a row are a dependent chain through the carry; at
// 33..64 columns this halves the row's latency).  Same operands, same compares, same
// bits as the chunked row in nw_warp; the code bits are regrouped into the band-relative
// words the traceback reads by a second round of ballots.
template <int CN>
__device__ __forceinline__ float nw_fill_blocked(const DDAln& A, const NwScratch& sc, int lane) {
  const int L1 = A.L1, L2 = A.L2;
  const float th = A.th;
  for (int k = lane; k < 2 * (L2 + 2); k += 32) sc.rowbuf[k] = -FLT_MAX;
  __syncwarp();
  float* prev = sc.rowbuf + (L2 + 2);
  float* mine = sc.rowbuf;
  uint2* trow = sc.trz;
  for (int r = 1; r <= L1; ++r) {
    { float* t = prev; prev = mine; mine = t; }
    trow += CN;                                            // ts == CN
    const uint2 ev = A.env[r];
    const int ef = max(1, (int)ev.x), el = (int)ev.y;
    const float* pb = A.pb + (A.boff[r] - ef);             // band index of (r, k) = boff[r] - ef + k
    const float* qb = A.qb + (A.boff[r] - ef);
    const float carry = (ef == 1) ? 0.0f : -FLT_MAX;      // dp[r][ef-1]
    if (lane == 0) mine[ef - 1] = carry;
    const int kb = ef + CN * lane;
    float pv[CN], qv[CN], upv[CN + 1], a[CN];
#pragma unroll
    for (int j = 0; j < CN; ++j) {                         // clamped into the band: no branch on a load
      const int kc = min(kb + j, el);
      pv[j] = pb[kc]; qv[j] = qb[kc];
    }
#pragma unroll
    for (int j = 0; j <= CN; ++j) upv[j] = (r > 1) ? prev[min(kb + j, el + 1) - 1] : 0.0f;   // dp[r-1][kb-1+j]
    unsigned c2m = 0u, inm = 0u;
    float x = -FLT_MAX;
#pragma unroll
    for (int j = 0; j < CN; ++j) {
      const bool in = kb + j <= el;
      const float m = __fadd_rn(__fsub_rn(__fadd_rn(upv[j], pv[j]), th), qv[j]);   // :282
      const bool c2 = m < upv[j + 1];                                             // :284
      a[j] = in ? (c2 ? upv[j + 1] : m) : -FLT_MAX;
      c2m |= (c2 ? 1u : 0u) << j;
      inm |= (in ? 1u : 0u) << j;
      x = fmaxf(x, a[j]);
    }
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) x = fmaxf(x, __shfl_up_sync(0xffffffffu, x, d));
    float left = __shfl_up_sync(0xffffffffu, x, 1);        // maximum over the lanes to the left
    left = (lane == 0) ? carry : fmaxf(left, carry);       // dp[r][kb-1]
    unsigned c3m = 0u;
#pragma unroll
    for (int j = 0; j < CN; ++j) {
      const bool c3 = a[j] < left;                                                // :289
      const float v = fmaxf(a[j], left);                                          // dp[r][kb+j]
      c3m |= (c3 ? 1u : 0u) << j;
      if ((inm >> j) & 1u) mine[kb + j] = v;
      left = v;
    }
    const unsigned lom = inm & (c3m | ~c2m);               // codes 1 and 3
    const unsigned him = inm & (c3m | c2m);                // codes 2 and 3
    unsigned bl[CN], bh[CN];
#pragma unroll
    for (int j = 0; j < CN; ++j) {
      bl[j] = __ballot_sync(0xffffffffu, (lom >> j) & 1u);
      bh[j] = __ballot_sync(0xffffffffu, (him >> j) & 1u);
    }
    // band column c = CN l + j  ->  word c / 32, bit c % 32
#pragma unroll
    for (int w = 0; w < CN; ++w) {
      const int c = 32 * w + lane, lsrc = c / CN, j = c - lsrc * CN;
      unsigned sl = bl[0], sh = bh[0];
#pragma unroll
      for (int t = 1; t < CN; ++t) if (j == t) { sl = bl[t]; sh = bh[t]; }
      const unsigned lo = __ballot_sync(0xffffffffu, (sl >> lsrc) & 1u);
      const unsigned hi = __ballot_sync(0xffffffffu, (sh >> lsrc) & 1u);
      if (lane == 0) trow[w] = make_uint2(lo, hi);
    }
    __syncwarp();
  }
  const float score = mine[L2];                            // el(L1) = L2: dp[L1][L2] was written
  __syncwarp();
  return score;
}

// SparseNeedlemanWunsch::decode(p, q, al) by one warp, one ROW per step.
// The reference's cell update (needleman_wunsch.cpp:282-293)
//   v = dp[i-1][k-1] + p - th + q;  if (v < dp[i-1][k]) v = dp[i-1][k];  if (v < dp[i][k-1]) v = dp[i][k-1];
// has no gap cost, so along a row dp[i][k] = max(a[k], dp[i][k-1]) with
// a[k] = max(dp[i-1][k-1] + p - th + q, dp[i-1][k]) depending on the row above only:
// the row is an (exact, order-independent) prefix maximum of a.  Lanes take 32
// consecutive columns of the row's band [ef, el] (needleman_wunsch.cpp:262-264: nothing
// outside it is ever evaluated), compute a, scan it with five shuffles, and derive the
// reference's traceback code from the strict compares in its order: 3 if a < dp[i][k-1],
// else 2 if the diagonal term < dp[i-1][k], else 1.  Codes are stored band-relative as
// two bit planes per 32 columns.  The row above comes from a double-buffered row in
// shared memory; cells of it outside ITS band are -FLT_MAX, column 0 and row 0 are 0
// (:265-274).  Returns dp[L1][L2] in every lane; *bad is set if the traceback meets an
// unfilled cell (the reference would spin there, :309).
__device__ __forceinline__ float nw_warp(const DDAln& A, const NwScratch& sc, int lane, int* bad,
                                         unsigned long long* prof) {
  const long long t0 = clock64();
  const int L1 = A.L1, L2 = A.L2, ts = sc.ts;
  const float th = A.th;
  float score;
  if (ts == 1 && !sc.chunked && L1 > 0 && L2 > 0) score = nw_fill_narrow(A, sc, lane);
  else if (ts == 2 && !sc.chunked) score = nw_fill_blocked<2>(A, sc, lane);
  else if (ts == 3 && !sc.chunked) score = nw_fill_blocked<3>(A, sc, lane);
  else if (ts == 4 && !sc.chunked) score = nw_fill_blocked<4>(A, sc, lane);
  else {
  // Both row buffers start at -FLT_MAX ("outside every band").  A row writes its band
  // and the cell just left of it (0 in column 0, :265-268, else -FLT_MAX, :262-264); the
  // bands only move right, so whatever the next row reads outside this row's band is
  // either that cell or was never written since the fill: no band tests on the reads.
  for (int k = lane; k < 2 * (L2 + 2); k += 32) sc.rowbuf[k] = -FLT_MAX;
  __syncwarp();
  float* prev = sc.rowbuf + (L2 + 2);
  float* mine = sc.rowbuf;
  uint2* trow = sc.trz;
  for (int r = 1; r <= L1; ++r) {
    { float* t = prev; prev = mine; mine = t; }
    trow += ts;
    const uint2 ev = A.env[r];
    const int ef = max(1, (int)ev.x), el = (int)ev.y;
    const float* pb = A.pb + (A.boff[r] - ef);             // band index of (r, k) = boff[r] - ef + k
    const float* qb = A.qb + (A.boff[r] - ef);
    float carry = (ef == 1) ? 0.0f : -FLT_MAX;            // dp[r][ef-1]
    if (lane == 0) mine[ef - 1] = carry;
    int chunk = 0;
    for (int k0 = ef; k0 <= el; k0 += 32, ++chunk) {
      const int k = k0 + lane;
      const bool in = k <= el;
      float up = 0.0f, dg = 0.0f, pv = 0.0f, qv = 0.0f;   // row 0 is 0 everywhere (:270-274)
      if (in) {
        pv = pb[k]; qv = qb[k];
        if (r > 1) { up = prev[k]; dg = prev[k - 1]; }     // dp[r-1][k], dp[r-1][k-1]
      }
      const float m = __fadd_rn(__fsub_rn(__fadd_rn(dg, pv), th), qv);             // :282
      const bool c2 = m < up;                                                      // :284
      float x = in ? (c2 ? up : m) : -FLT_MAX;
      const float a = x;
      // inclusive prefix maximum; a lane below the shuffle distance gets its own value
      // back, which a maximum ignores, so no lane test is needed
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) x = fmaxf(x, __shfl_up_sync(0xffffffffu, x, d));
      const float v = fmaxf(x, carry);                                             // dp[r][k]
      float left = __shfl_up_sync(0xffffffffu, v, 1);                              // dp[r][k-1]
      if (lane == 0) left = carry;
      const bool c3 = a < left;                                                    // :289
      const unsigned lo = __ballot_sync(0xffffffffu, in && (c3 || !c2));           // codes 1 and 3
      const unsigned hi = __ballot_sync(0xffffffffu, in && (c3 || c2));            // codes 2 and 3
      if (lane == 0) trow[chunk] = make_uint2(lo, hi);
      if (in) mine[k] = v;
      carry = __shfl_sync(0xffffffffu, v, 31);
    }
    __syncwarp();
  }
  score = mine[L2];                                        // el(L1) = L2: dp[L1][L2] was written
  __syncwarp();
  }
  const long long t1 = clock64();
  // ---- traceback (:300-312) + alignment read-out (:317-335) ------------------
  // A row is left through the cell the walk reaches after a run of 'Y' codes (3, leftwards,
  // nothing recorded): code 1 records the pair and enters the row above one column to the
  // left, code 2 enters it in the same column.  Alignments are mostly diagonal, so lane l
  // decodes row i - l on the guess that the walk enters it at column k - l; the guesses are
  // right up to the first lane whose predecessor left its row somewhere else, that prefix
  // is committed (up to 32 rows per round) and the walk resumes from the first wrong guess.
  for (int i = lane; i < L1; i += 32) sc.zres[i] = DD_NOPAIR;
  __syncwarp();
  {
    int i = L1, k = L2;
    while (i > 0 && k > 0) {       // row 0 is all 'Y', column 0 all 'X': nothing more to record
      const int ri = i - lane, kg = k - lane;
      int status = 3;              // 0 left the row upwards, 1 reached column 0, 2 unfilled cell, 3 no guess
      int kx = 0, rec = -1;
      if (ri >= 1 && kg >= 1) {
        const uint2 ev = A.env[ri];
        const int ef = max(1, (int)ev.x);
        int kk = kg, c = kk - ef;
        if (c < 0 || kk > (int)ev.y) status = 2;                     // a cell that was never filled
        else {
          const uint2* trow = sc.trz + (size_t)ri * ts;
          for (;;) {
            const uint2 w = trow[c >> 5];
            // jump to the nearest column at or left of kk in this word whose code is not 3
            const int b = c & 31;
            const uint32_t stop = ~(w.x & w.y) & (0xffffffffu >> (31 - b));
            if (stop == 0u) {                                        // the run leaves the word
              kk -= b + 1; c -= b + 1;
              if (kk <= 0) { status = 1; break; }
              if (c < 0) { status = 2; break; }
              continue;
            }
            const int b2 = 31 - __clz(stop);
            kk -= b - b2;
            const uint32_t code = ((w.x >> b2) & 1u) | (((w.y >> b2) & 1u) << 1);
            if (code == 1u) { rec = kk - 1; kx = kk - 1; status = 0; }
            else if (code == 2u) { kx = kk; status = 0; }
            else status = 2;
            break;
          }
        }
      }
      const int pkx = __shfl_up_sync(0xffffffffu, kx, 1);
      const int pst = __shfl_up_sync(0xffffffffu, status, 1);
      const bool link = lane == 0 || (status != 3 && pst == 0 && pkx == kg);
      const unsigned lm = __ballot_sync(0xffffffffu, link);
      const int nacc = (lm == 0xffffffffu) ? 32 : __ffs(~lm) - 1;    // >= 1: lane 0 holds the true state
      if (lane < nacc && rec >= 0) sc.zres[ri - 1] = (uint32_t)rec;
      const int lst = __shfl_sync(0xffffffffu, status, nacc - 1);
      const int lkx = __shfl_sync(0xffffffffu, kx, nacc - 1);
      if (lst == 2) { if (lane == 0) *bad = 1; break; }
      if (lst == 1) break;
      i -= nacc; k = lkx;
    }
  }
  __syncwarp();
  if (prof && lane == 0) { const long long t2 = clock64(); prof[0] += t1 - t0; prof[1] += t2 - t1; }
  return score;
}

// max(0.0f, v) exactly as std::max(0.0f, v): (0.0f < v) ? v : 0.0f  (NaN -> 0)
__device__ __forceinline__ float dd_max0(float v) { return (0.0f < v) ? v : 0.0f; }

struct DDShared {
  float part[3];
  float s, s_prev, c, eta;
  unsigned violated;
  int done, bad, q;
  int prog[8];                 // Nussinov team progress flags: x team [0..3], y team [4..7]
  float nus_score[2];
  // cycles: 0 nussinov x fill, 1 x traceback, 2 y fill, 3 y traceback, 4 nw fill, 5 nw traceback,
  // 6 decode phase (CTA), 7 tally + update phase (CTA)
  unsigned long long prof[8];
};

__device__ __forceinline__ void dd_update_nus(const DDNus& S, const uint32_t* ss, int* tal, float eta,
                                              unsigned& viol, int tid, int nthreads) {
  // dafs.cpp:1121-1150 (x) and :1172-1201 (y), one thread per entry
  for (int e = tid; e < S.n; e += nthreads) {
    const uint32_t ij = S.ent_ij[e];
    const int i = (int)(ij & 0xffffu), j = (int)(ij >> 16);
    const int t = tal[e];
    tal[e] = 0;
    if (ss[i] == (uint32_t)j) {
      if (t != 1) { ++viol; S.ent_q[e] = __fsub_rn(S.ent_q[e], __fmul_rn(eta, (float)(t - 1))); }
    } else if ((S.ent_flag[e] & 1) && t != 0) {
      ++viol;
      S.ent_q[e] = __fsub_rn(S.ent_q[e], __fmul_rn(eta, (float)t));
    }
  }
}

#define DD_THREADS 96          // T = 1: one warp per decoder
#define DD_MAX_TEAM 4

// SMEM: every scratch item and the constant block are in shared memory (the host
// says so), which lets the compiler address them as shared (LDS/STS, 32-bit
// addresses); otherwise pointers are generic and items may sit in HBM.
// T: warps per Nussinov structure.  T = 1 is the throughput configuration (three warps per
// problem, several CTAs per SM); T > 1 is used when a launch has fewer problems than the GPU
// has SMs (the upper waves of a guide tree), where only the latency of one problem counts.
template <bool SMEM, int T>
__global__ void __launch_bounds__(32 * (2 * T + 1)) dd_kernel(const DDArgs a) {
  constexpr int NT = 32 * (2 * T + 1);
  extern __shared__ __align__(16) char smem[];
  __shared__ DDShared sh;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const DDLayout& lay = a.lay;
  char* const wsb = SMEM ? smem : lay.ws + (size_t)blockIdx.x * lay.ws_stride;
#define DD_ITEM(T, item) reinterpret_cast<T*>((SMEM || ((lay.in_smem >> (item)) & 1u) ? smem : wsb) + lay.off[item])
  // this warp's Nussinov structure: x for warps [0, T), y for warps [T, 2T); warp 2T aligns
  const int wy = (warp >= T && warp < 2 * T) ? 1 : 0;
  const int team_m = warp < T ? warp : warp - T;
  NusScratch ns;
  ns.es = DD_ITEM(float, DD_ES_X + wy); ns.row = DD_ITEM(float, DD_ROW_X + wy);
  ns.brow = DD_ITEM(float, DD_BROW_X + wy); ns.tr = DD_ITEM(void, DD_TR_X + wy);
  ns.ss = DD_ITEM(uint32_t, DD_SS_X + wy); ns.stk = DD_ITEM(uint32_t, DD_STK_X + wy);
  ns.lp = wy ? lay.lp2 : lay.lp1; ns.trs = wy ? lay.trs2 : lay.trs1; ns.tr16 = lay.tr16 != 0;
  ns.prog = sh.prog + 4 * wy; ns.score_slot = sh.nus_score + wy;
  NwScratch nz;
  nz.rowbuf = DD_ITEM(float, DD_BROW_Z); nz.trz = DD_ITEM(uint2, DD_TRZ); nz.zres = DD_ITEM(uint32_t, DD_ZRES);
  nz.ts = lay.ts;
  nz.chunked = lay.nw_chunked;
  uint32_t* ss_x = DD_ITEM(uint32_t, DD_SS_X);
  uint32_t* ss_y = DD_ITEM(uint32_t, DD_SS_Y);
  float* es_x = DD_ITEM(float, DD_ES_X);
  float* es_y = DD_ITEM(float, DD_ES_Y);
  int* tal_x = DD_ITEM(int, DD_TAL_X);
  int* tal_y = DD_ITEM(int, DD_TAL_Y);
  int* tal_z = DD_ITEM(int, DD_TAL_Z);
  float* sw = DD_ITEM(float, DD_SW);
  float* q_x = DD_ITEM(float, DD_Q_X);
  float* q_y = DD_ITEM(float, DD_Q_Y);
  float* q_z = DD_ITEM(float, DD_Q_Z);       // band
  float* q_cz = DD_ITEM(float, DD_Q_CZ);     // one per c_z cell
  char* cst_s = DD_ITEM(char, DD_CST);
  const bool cst_in_smem = SMEM || ((lay.in_smem >> DD_CST) & 1u);

  for (;;) {
    __syncthreads();
    if (tid == 0) sh.q = atomicAdd(a.queue, 1);
    __syncthreads();
    if (sh.q >= a.nprob) break;
    const DDProbRef pr = a.prob[a.order[sh.q]];
    const char* blk = pr.cst;
    if (cst_in_smem) {            // stage the constant block: coalesced 16-byte copies
      const int n16 = reinterpret_cast<const DDHdr*>(pr.cst)->stage >> 4;
      const uint4* src = reinterpret_cast<const uint4*>(pr.cst);
      uint4* dst = reinterpret_cast<uint4*>(cst_s);
      for (int k = tid; k < n16; k += NT) dst[k] = src[k];
      blk = cst_s;
    }
    __syncthreads();
    const DDHdr& h = *reinterpret_cast<const DDHdr*>(blk);
    const DDNus X = dd_make_nus(blk, h.x, q_x), Y = dd_make_nus(blk, h.y, q_y);
    const DDNus& mine = wy ? Y : X;
    DDAln Z;
    Z.L1 = h.L1; Z.L2 = h.L2; Z.th = h.th_a; Z.qb = q_z;
    Z.env = reinterpret_cast<const uint2*>(blk + h.z_env);
    Z.boff = reinterpret_cast<const int*>(blk + h.z_boff);
    Z.pb = reinterpret_cast<const float*>(blk + h.z_pb);
    const int* cz_row_ptr = reinterpret_cast<const int*>(blk + h.cz_row_ptr);
    const uint16_t* cz_col = reinterpret_cast<const uint16_t*>(blk + h.cz_col);
    const int* cz_bidx = reinterpret_cast<const int*>(blk + h.cz_bidx);
    const int nb = h.nb, ncz = h.ncz;
    const int4* cbp = reinterpret_cast<const int4*>(pr.cst + h.cbp);
    const int L1 = h.L1, L2 = h.L2, ncbp = h.ncbp, t_max = h.t_max;
    const float th_s = h.th_s, eta0 = h.eta0;

    // ---- q = 0 (dafs.cpp:1067-1069), scratch reset ----------------------------
    for (int e = tid; e < X.n; e += NT) { q_x[e] = 0.0f; tal_x[e] = 0; if (e < X.nd) es_x[e] = -INFINITY; }
    for (int e = tid; e < Y.n; e += NT) { q_y[e] = 0.0f; tal_y[e] = 0; if (e < Y.nd) es_y[e] = -INFINITY; }
    for (int e = tid; e < nb; e += NT) q_z[e] = 0.0f;
    for (int e = tid; e < ncz; e += NT) { q_cz[e] = 0.0f; tal_z[e] = 0; }
    for (int k = tid; k < L1; k += NT) { ss_x[k] = DD_NOPAIR; nz.zres[k] = DD_NOPAIR; }
    for (int k = tid; k < L2; k += NT) ss_y[k] = DD_NOPAIR;
    if (tid == 0) {
      sh.c = 0.0f; sh.eta = eta0; sh.s_prev = 0.0f; sh.violated = 0; sh.done = 0; sh.bad = 0;
      for (int k = 0; k < 8; ++k) sh.prog[k] = 0;
      for (int k = 0; k < 8; ++k) sh.prof[k] = 0;
    }
    __syncthreads();

    int t = 0;
    for (; t < t_max; ++t) {
      // ---- the three subproblems (dafs.cpp:1090-1093) -------------------------
      const long long tp0 = clock64();
      if (warp < 2 * T) {
        const float v = nussinov_warp<T>(mine, ns, th_s, lane, team_m, 1 + wy, team_m == 0 ? sh.prof + 2 * wy : nullptr);
        if (lane == 0 && team_m == 0) sh.part[wy] = v;
      } else { const float v = nw_warp(Z, nz, lane, &sh.bad, sh.prof + 4); if (lane == 0) sh.part[2] = v; }
      if (tid == 0) sh.violated = 0;
      __syncthreads();
      if (T > 1 && tid < 8) sh.prog[tid] = 0;                 // wave numbers restart with every decode
      const long long tp1 = clock64();
      // ---- tally (:1103-1117) ---------------------------------------------------
      for (int u = tid; u < ncbp; u += NT) {
        const int4 cb = cbp[u];
        const float s_w = __fsub_rn(__fsub_rn(__fadd_rn(q_x[cb.x], q_y[cb.y]), q_cz[cb.z]), q_cz[cb.w]);
        sw[u] = s_w;
        if (s_w > 0.0f) {
          atomicAdd(&tal_x[cb.x], 1); atomicAdd(&tal_y[cb.y], 1);
          atomicAdd(&tal_z[cb.z], 1); atomicAdd(&tal_z[cb.w], 1);
        }
      }
      __syncthreads();
      // ---- Lagrangian value, summed in CBP order (:1090-1111) ----
      // The reference adds only the positive s_w, one after the other.  The last warp walks the list 32
      // entries at a time and chains an addition per POSITIVE entry (in order, broadcast by shuffle):
      // the serial part is as long as the number of violated candidates, not as the list (2 112 CBPs at
      // the root of RF00017 cost 100 k cycles per iteration with one thread reading every entry).
      if (tid >= NT - 32) {
        const int ln = tid & 31;
        float s = __fadd_rn(0.0f, sh.part[0]);
        s = __fadd_rn(s, sh.part[1]);
        s = __fadd_rn(s, sh.part[2]);
        for (int base = 0; base < ncbp; base += 32) {
          const int u = base + ln;
          const float v = u < ncbp ? sw[u] : 0.0f;
          unsigned m = __ballot_sync(0xffffffffu, v > 0.0f);
          while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            s = __fadd_rn(s, __shfl_sync(0xffffffffu, v, b));
          }
        }
        if (ln == 31) sh.s = s;
      }
      // ---- multiplier updates (:1121-1254) -----------------------------------------
      const float eta = sh.eta;
      unsigned viol = 0;
      dd_update_nus(X, ss_x, tal_x, eta, viol, tid, NT);
      dd_update_nus(Y, ss_y, tal_y, eta, viol, tid, NT);
      // z (:1223-1254), one thread per row over its c_z cells.  Cells outside c_z
      // have t_z = 0 and q_z = 0, and stay there: max(0, 0 - eta*(1-0)) = 0 for
      // eta >= 0 or NaN (the host rejects eta0 < 0).
      for (int i = tid; i < L1; i += NT) {
        const uint32_t zi = nz.zres[i];
        for (int e = cz_row_ptr[i]; e < cz_row_ptr[i + 1]; ++e) {
          const int tz = tal_z[e];
          tal_z[e] = 0;
          float qn;
          if (zi == (uint32_t)cz_col[e]) {                      // z_ik == 1  (:1226-1238)
            if (tz > 1) ++viol;
            qn = dd_max0(__fsub_rn(q_cz[e], __fmul_rn(eta, (float)(1 - tz))));
          } else {                                               // z_ik == 0  (:1239-1253)
            if (tz > 0) ++viol;
            qn = dd_max0(__fadd_rn(q_cz[e], __fmul_rn(eta, (float)tz)));
          }
          q_cz[e] = qn;
          const int bi = cz_bidx[e];
          if (bi >= 0) q_z[bi] = qn;                             // the copy the aligner reads
        }
      }
      if (viol) atomicAdd(&sh.violated, viol);
      __syncthreads();
      if (tid == 0) {
        const unsigned violated = sh.violated;
        const float s = sh.s;
        if (pr.trace) { pr.trace[3 * t] = s; pr.trace[3 * t + 1] = eta; pr.trace[3 * t + 2] = (float)violated; }
        if (violated == 0) sh.done = 1;                     // :1278
        else {
          if (s > sh.s_prev || t == 0) {                    // :1283-1288
            const float n4 = __fmul_rn(4.0f, (float)(unsigned long long)ncbp);
            const float av = dd_max0(__fsub_rn(n4, (float)violated));
            sh.c = (float)((double)sh.c + (double)av / (4.0 * (double)ncbp));
            sh.eta = (float)((double)eta0 / (1.0 + (double)sh.c));
          }
          sh.s_prev = s;
        }

      }
      __syncthreads();
      if (tid == 0) { const long long tp2 = clock64(); sh.prof[6] += tp1 - tp0; sh.prof[7] += tp2 - tp1; }
      if (sh.done) break;
    }
    // ---- results -------------------------------------------------------------------
    for (int k = tid; k < L1; k += NT) { pr.x_out[k] = ss_x[k]; pr.z_out[k] = nz.zres[k]; }
    for (int k = tid; k < L2; k += NT) pr.y_out[k] = ss_y[k];
    if (tid == 0) {
      pr.result[0] = sh.s_prev;
      reinterpret_cast<int*>(pr.result)[1] = t;
      reinterpret_cast<unsigned*>(pr.result)[2] = sh.violated;
      reinterpret_cast<int*>(pr.result)[3] = sh.bad;
      for (int k = 0; k < 8; ++k) pr.result[4 + k] = (float)sh.prof[k];
    }
    // completion flag for dafs_cuda_dd_plan_poll: set after everything above is visible
    __threadfence();
    __syncthreads();
    if (tid == 0) { __threadfence(); reinterpret_cast<volatile int*>(pr.result)[12] = 1; }
  }
#undef DD_ITEM
}

}  // namespace dafs
