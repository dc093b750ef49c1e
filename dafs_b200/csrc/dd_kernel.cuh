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
// queue).  Warp 0 decodes x, warp 1 decodes y, warp 2 decodes z, concurrently;
// then all threads tally and update the multipliers.  Neither decoder uses a
// block barrier inside its DP:
//   * Nussinov: lanes own 32 consecutive rows (a "wave", bottom rows first) and
//     walk their row left to right, one column behind the row below, whose
//     values arrive by warp shuffle; bifurcations read the lane's own row from
//     shared memory and the candidate values of the column (reference: bp[j]).
//   * NW: lanes own C consecutive columns, rows run through the lanes as a
//     skewed pipeline (like the pair-HMM kernel); traceback codes are packed 2
//     bits per cell.
// Multipliers are sparse for x and y (only entries that can ever be non-zero:
// the decoder's candidates and the CBP projections c_x / c_y) and dense for z.
#pragma once
#include <cfloat>
#include <cstdint>
#include <cuda_runtime.h>

namespace dafs {

#define DD_NOPAIR 0xffffffffu

// One Nussinov structure (x or y) of one problem.  Entries are kept in "column
// order": ascending j, then descending i -- the order in which the reference
// pushes candidates to bp[j] (nussinov.cpp:236-245).
struct DDNus {
  int L, n;
  const int* col_ptr;          // [L+1] entries of column j
  const int* row_ptr;          // [L+1]
  const int* row_ent;          // [n] entry ids of row i, ascending j
  const uint32_t* ent_ij;      // [n] i | j << 16
  const float* ent_p;          // [n]
  const uint8_t* ent_flag;     // [n] bit0: in c_x (dafs.cpp:1037), bit1: j-i > 2 (decoder candidate)
  float* ent_q;                // [n] multipliers q_x[i][j]
  float w;                     // w_*2*N1/(N1+N2)  (dafs.cpp:1091)
  uint32_t* out;               // [L] result in the arena
};

struct DDProblem {
  DDNus sx, sy;
  int L1, L2, nz, ncbp, t_max;
  float th_s, th_a, eta0;
  const float* pz;             // [L1*L2]
  float* qz;                   // [L1*L2] multipliers q_z
  const uint2* env;            // [L1+1] {first, second}   (needleman_wunsch.cpp:198-253)
  const int* cz_ptr;           // [L1+1] c_z rows (dafs.cpp:1039-1040, sorted unique)
  const int* cz_col;           // [nz]
  const int* cz_cell;          // [nz] i*L2 + k
  const int4* cbp;             // [ncbp] {entry in x, entry in y, c_z id of (i,k), c_z id of (j,l)}
  uint32_t* z_out;             // [L1]
  float* result;               // {score, iters, violated, status} as 4 x 32 bit
  float* trace;                // optional [t_max*3] {s, eta, violated} per iteration (-v 2, dafs.cpp:1273)
};

// Per-CTA scratch arrays; each lives in shared memory when it fits (host decides,
// most latency-critical first) and in a per-CTA global workspace otherwise.
enum {
  DD_ES_X, DD_ES_Y, DD_ROW_X, DD_ROW_Y, DD_BROW_X, DD_BROW_Y, DD_TRZ, DD_TR_X, DD_TR_Y,
  DD_SS_X, DD_SS_Y, DD_ZRES, DD_STK_X, DD_STK_Y, DD_TAL_X, DD_TAL_Y, DD_TAL_Z, DD_SW, DD_BND,
  DD_NITEMS
};
struct DDLayout {
  uint32_t off[DD_NITEMS];     // byte offset in smem or in the CTA's workspace
  uint32_t in_smem;            // bit per item
  uint32_t tr16;               // traceback codes of Nussinov need 16 bits (a column with > 250 entries)
  int lp1, lp2;                // padded row length of the Nussinov row buffers
  int nstripes;                // NW column stripes of 32*C
  char* ws;                    // global workspace
  size_t ws_stride;            // per CTA
};

struct DDArgs {
  const DDProblem* prob;
  const int32_t* order;        // queue position -> problem (largest first)
  int nprob;
  int* queue;
  DDLayout lay;
};

template <class T>
__device__ __forceinline__ T* dd_item(const DDLayout& lay, char* smem, int item) {
  char* base = (lay.in_smem >> item) & 1u ? smem : lay.ws + (size_t)blockIdx.x * lay.ws_stride;
  return reinterpret_cast<T*>(base + lay.off[item]);
}

__device__ __forceinline__ int dd_tri(int L, int i, int j) {   // i < j, row-major upper triangle
  return i * (2 * L - i - 1) / 2 + (j - i - 1);
}

struct NusScratch {
  float* es;        // [n] value a candidate contributes: dp[i+1][j-1] + s, or -inf when inactive
  float* row;       // [32*lp] dp rows of the current wave
  float* brow;      // [lp] dp row just above the next wave (row of the wave's last lane)
  void* tr;         // traceback codes, triangular
  uint32_t* ss;     // [L] result
  uint32_t* stk;    // traceback stack
  int lp;
  bool tr16;
};

__device__ __forceinline__ void nus_tr_store(const NusScratch& sc, int idx, int t) {
  if (sc.tr16) reinterpret_cast<uint16_t*>(sc.tr)[idx] = (uint16_t)t;
  else reinterpret_cast<uint8_t*>(sc.tr)[idx] = (uint8_t)t;
}
__device__ __forceinline__ int nus_tr_load(const NusScratch& sc, int idx) {
  return sc.tr16 ? (int)reinterpret_cast<const uint16_t*>(sc.tr)[idx]
                 : (int)reinterpret_cast<const uint8_t*>(sc.tr)[idx];
}

// SparseNussinov::decode(w, p, q, ss): fill + traceback by one warp.
// Returns dp[0][L-1] in every lane.
__device__ __noinline__ float nussinov_warp(const DDNus& S, const NusScratch& sc, float th, int lane) {
  const int L = S.L, lp = sc.lp;
  float score = 0.0f;
  float* myrow = sc.row + lane * lp;
  for (int hi = L - 1; hi >= 0; hi -= 32) {
    const int row = hi - lane;
    const bool live = row >= 0;
    int rcur = 0, rend = 0;
    if (live) { rcur = S.row_ptr[row]; rend = S.row_ptr[row + 1]; myrow[row] = 0.0f; }
    int next_j = (rcur < rend) ? (int)(S.ent_ij[S.row_ent[rcur]] >> 16) : 0x7fffffff;
    float cur = 0.0f;          // dp[row][j-1] entering a step, dp[row][j] leaving it
    float below_prev = 0.0f;   // dp[row+1][j-1]
    const bool have_brow = (hi + 1 < L);
    const int nsteps = (L - 1 - hi) + 32;
    for (int s = 0; s < nsteps; ++s) {
      const int j = row + 1 + s;
      float below = __shfl_up_sync(0xffffffffu, cur, 1);
      if (lane == 0) below = (have_brow && j < L) ? sc.brow[j] : 0.0f;
      if (live && j < L) {
        const int l = s + 1;                               // j - row
        float v = 0.0f;
        int t = 0;
        if (l >= 2) { v = below; t = 1; }                  // nussinov.cpp:224-228
        if (l >= 2 && v < cur) { v = cur; t = 2; }         // :229-233
        if (j == next_j) {                                 // (row, j) is an entry
          const int e = S.row_ent[rcur];
          ++rcur;
          next_j = (rcur < rend) ? (int)(S.ent_ij[S.row_ent[rcur]] >> 16) : 0x7fffffff;
          if (S.ent_flag[e] & 2) {                         // :234  i+1 < j-1
            const float sv = __fsub_rn(__fmul_rn(S.w, __fsub_rn(S.ent_p[e], th)), S.ent_q[e]);   // :236
            float c = -INFINITY;
            if (sv > 0.0f) {
              c = __fadd_rn(below_prev, sv);               // :239
              if (v < c) { v = c; t = 3; }                 // :240-244
            }
            sc.es[e] = c;
          }
        }
        // bifurcations: entries of column j with k > row, in push order (:247-259)
        const int cb = S.col_ptr[j], ce = S.col_ptr[j + 1];
        for (int e = cb; e < ce; ++e) {
          const int k = (int)(S.ent_ij[e] & 0xffffu);
          if (k <= row) break;
          const float c = __fadd_rn(myrow[k - 1], sc.es[e]);
          if (v < c) { v = c; t = 4 + (e - cb); }
        }
        myrow[j] = v;
        nus_tr_store(sc, dd_tri(L, row, j), t);
        below_prev = below;
        cur = v;
        if (row == 0 && j == L - 1) score = v;
      }
      __syncwarp();
    }
    // the row of the last lane becomes the "row below" of the next wave
    if (hi - 32 >= 0) {
      const int r31 = hi - 31;
      for (int j = r31 + lane; j < L; j += 32) sc.brow[j] = sc.row[31 * lp + j];
    }
    __syncwarp();
  }
  {
    const int owner = (L - 1) & 31;     // lane of row 0 in the last wave: hi = (L-1)%32 there
    score = __shfl_sync(0xffffffffu, score, owner);
  }

  // ---- traceback (:265-295), all lanes in lock step; runs of code 1 / 2 are
  // skipped 32 cells at a time.  The order in which the reference's stack is
  // served does not change ss.
  for (int i = lane; i < L; i += 32) sc.ss[i] = DD_NOPAIR;
  __syncwarp();
  int sp = 0;
  int i = 0, j = L - 1;
  for (;;) {
    bool pop = false;
    if (i >= j) pop = true;
    else {
      const int c = nus_tr_load(sc, dd_tri(L, i, j));
      if (c == 0) pop = true;
      else if (c == 1) {
        const int ii = i + 1 + lane;
        const int cc = (ii < j) ? nus_tr_load(sc, dd_tri(L, ii, j)) : 0;
        const unsigned m = __ballot_sync(0xffffffffu, cc != 1);
        i += 1 + (m ? (__ffs(m) - 1) : 32);
      } else if (c == 2) {
        const int jj = j - 1 - lane;
        const int cc = (jj > i) ? nus_tr_load(sc, dd_tri(L, i, jj)) : 0;
        const unsigned m = __ballot_sync(0xffffffffu, cc != 2);
        j -= 1 + (m ? (__ffs(m) - 1) : 32);
      } else if (c == 3) {
        if (lane == 0) sc.ss[i] = (uint32_t)j;
        ++i; --j;
      } else {
        const int e = S.col_ptr[j] + (c - 4);
        const int k = (int)(S.ent_ij[e] & 0xffffu);
        if (lane == 0) { sc.ss[k] = (uint32_t)j; sc.stk[sp] = (uint32_t)(k + 1) | ((uint32_t)(j - 1) << 16); }
        ++sp;
        j = k - 1;
      }
    }
    if (pop) {
      if (sp == 0) break;
      --sp;
      __syncwarp();
      const uint32_t w = sc.stk[sp];
      i = (int)(w & 0xffffu); j = (int)(w >> 16);
    }
  }
  __syncwarp();
  return score;
}

}  // namespace dafs

namespace dafs {

// SparseNeedlemanWunsch::decode(p, q, al) by one warp.  trz: packed traceback
// codes (2 bits per cell: 0 ' ', 1 'M', 2 'X', 3 'Y'), one word per (row,
// stripe, lane).  Returns dp[L1][L2] in every lane; *bad is set if the
// traceback meets an unfilled cell (the reference would spin there).
template <int C>
__device__ __noinline__ float nw_warp(const DDProblem& P, uint32_t* trz, float* bnd, uint32_t* zres,
                                      int lane, int* bad) {
  const int L1 = P.L1, L2 = P.L2;
  const int SW = 32 * C;
  const int nstripes = (L2 + 1 + SW - 1) / SW;
  const float th = P.th_a;
  const float* __restrict__ pz = P.pz;
  const float* qz = P.qz;
  float score = 0.0f;
  for (int stripe = 0; stripe < nstripes; ++stripe) {
    const int k0 = stripe * SW + lane * C;
    const float* bnd_in = bnd + ((stripe & 1) ? 0 : (L1 + 1));
    float* bnd_out = bnd + ((stripe & 1) ? (L1 + 1) : 0);
    float pd[C];                 // dp[i-1][k0+c]
#pragma unroll
    for (int c = 0; c < C; ++c) pd[c] = 0.0f;
    float dgl = 0.0f;            // dp[i-1][k0-1]
    float o = 0.0f;              // my last column of the row done in the previous step
    float pn[C], qn[C];          // p/q of the row this lane does next
#pragma unroll
    for (int c = 0; c < C; ++c) { pn[c] = 0.0f; qn[c] = 0.0f; }
    const int nsteps = L1 + 1 + 31;
    for (int step = 0; step < nsteps; ++step) {
      float left = __shfl_up_sync(0xffffffffu, o, 1);
      const int i = step - lane;
      float pc[C], qc[C];
#pragma unroll
      for (int c = 0; c < C; ++c) { pc[c] = pn[c]; qc[c] = qn[c]; }
      if (i + 1 >= 1 && i + 1 <= L1) {             // prefetch array row i for DP row i+1
        const size_t base = (size_t)i * L2;
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const int k = k0 + c;
          if (k >= 1 && k <= L2) { pn[c] = __ldg(pz + base + k - 1); qn[c] = qz[base + k - 1]; }
        }
      }
      if (i < 0 || i > L1) continue;
      if (lane == 0) left = (stripe > 0) ? bnd_in[i] : 0.0f;
      uint32_t word = 0;
      if (i == 0) {
        // dp[0][k] = 0, tr 'Y' (needleman_wunsch.cpp:270-274); (0,0) is never traced
#pragma unroll
        for (int c = 0; c < C; ++c) { pd[c] = 0.0f; if (k0 + c >= 1) word |= 3u << (2 * c); }
      } else {
        const uint2 ev = P.env[i];
        const int ef = max(1, (int)ev.x), el = (int)ev.y;
        float diag = dgl, lf = left;
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const int k = k0 + c;
          const float up = pd[c];
          float v;
          uint32_t code;
          if (k == 0) { v = 0.0f; code = 2u; }                 // :265-268
          else if (k >= ef && k <= el) {
            v = __fadd_rn(__fsub_rn(__fadd_rn(diag, pc[c]), th), qc[c]);   // :282
            code = 1u;
            if (v < up) { v = up; code = 2u; }                 // :284-288
            if (v < lf) { v = lf; code = 3u; }                 // :289-293
          } else { v = -FLT_MAX; code = 0u; }                  // :262-264
          diag = up;
          pd[c] = v;
          lf = v;
          word |= code << (2 * c);
        }
      }
      trz[(i * nstripes + stripe) * 32 + lane] = word;
      dgl = left;
      o = pd[C - 1];
      if (lane == 31 && stripe + 1 < nstripes) bnd_out[i] = o;
      if (i == L1) {
#pragma unroll
        for (int c = 0; c < C; ++c)
          if (k0 + c == L2) score = pd[c];
      }
    }
    __syncwarp();
  }
  {
    const int owner = (L2 % SW) / C;
    score = __shfl_sync(0xffffffffu, score, owner);
  }
  // ---- traceback (:300-312) + alignment read-out (:317-335) ------------------
  for (int i = lane; i < L1; i += 32) zres[i] = DD_NOPAIR;
  __syncwarp();
  if (lane == 0) {
    int i = L1, k = L2;
    while (i > 0 || k > 0) {
      const int st = k / SW, r = k - st * SW, ln = r / C, c = r - ln * C;
      const uint32_t code = (trz[(i * nstripes + st) * 32 + ln] >> (2 * c)) & 3u;
      if (code == 1u) { --i; --k; zres[i] = (uint32_t)k; }
      else if (code == 2u) --i;
      else if (code == 3u) --k;
      else { *bad = 1; break; }
    }
  }
  __syncwarp();
  return score;
}

// max(0.0f, v) exactly as std::max(0.0f, v): (0.0f < v) ? v : 0.0f  (NaN -> 0)
__device__ __forceinline__ float dd_max0(float v) { return (0.0f < v) ? v : 0.0f; }

struct DDShared {
  DDProblem P;
  float part[3];
  float s, s_prev, c, eta;
  unsigned violated;
  int done, bad, q;
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

template <int C>
__global__ void __launch_bounds__(96) dd_kernel(const DDArgs a) {
  extern __shared__ __align__(16) char smem[];
  __shared__ DDShared sh;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const DDLayout& lay = a.lay;
  for (;;) {
    __syncthreads();
    if (tid == 0) sh.q = atomicAdd(a.queue, 1);
    __syncthreads();
    if (sh.q >= a.nprob) break;
    {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(a.prob + a.order[sh.q]);
      uint32_t* dst = reinterpret_cast<uint32_t*>(&sh.P);
      for (int k = tid; k < (int)(sizeof(DDProblem) / 4); k += 96) dst[k] = src[k];
    }
    __syncthreads();
    const DDProblem& P = sh.P;
    NusScratch nx, ny;
    nx.es = dd_item<float>(lay, smem, DD_ES_X); ny.es = dd_item<float>(lay, smem, DD_ES_Y);
    nx.row = dd_item<float>(lay, smem, DD_ROW_X); ny.row = dd_item<float>(lay, smem, DD_ROW_Y);
    nx.brow = dd_item<float>(lay, smem, DD_BROW_X); ny.brow = dd_item<float>(lay, smem, DD_BROW_Y);
    nx.tr = dd_item<void>(lay, smem, DD_TR_X); ny.tr = dd_item<void>(lay, smem, DD_TR_Y);
    nx.ss = dd_item<uint32_t>(lay, smem, DD_SS_X); ny.ss = dd_item<uint32_t>(lay, smem, DD_SS_Y);
    nx.stk = dd_item<uint32_t>(lay, smem, DD_STK_X); ny.stk = dd_item<uint32_t>(lay, smem, DD_STK_Y);
    nx.lp = lay.lp1; ny.lp = lay.lp2; nx.tr16 = ny.tr16 = lay.tr16 != 0;
    uint32_t* trz = dd_item<uint32_t>(lay, smem, DD_TRZ);
    uint32_t* zres = dd_item<uint32_t>(lay, smem, DD_ZRES);
    float* bnd = dd_item<float>(lay, smem, DD_BND);
    int* tal_x = dd_item<int>(lay, smem, DD_TAL_X);
    int* tal_y = dd_item<int>(lay, smem, DD_TAL_Y);
    int* tal_z = dd_item<int>(lay, smem, DD_TAL_Z);
    float* sw = dd_item<float>(lay, smem, DD_SW);

    // ---- q = 0 (dafs.cpp:1067-1069), scratch reset ----------------------------
    for (int e = tid; e < P.sx.n; e += 96) { P.sx.ent_q[e] = 0.0f; nx.es[e] = -INFINITY; tal_x[e] = 0; }
    for (int e = tid; e < P.sy.n; e += 96) { P.sy.ent_q[e] = 0.0f; ny.es[e] = -INFINITY; tal_y[e] = 0; }
    for (int e = tid; e < P.nz; e += 96) tal_z[e] = 0;
    for (int k = tid; k < P.L1 * P.L2; k += 96) P.qz[k] = 0.0f;
    for (int k = tid; k < P.L1; k += 96) { nx.ss[k] = DD_NOPAIR; zres[k] = DD_NOPAIR; }
    for (int k = tid; k < P.L2; k += 96) ny.ss[k] = DD_NOPAIR;
    if (tid == 0) { sh.c = 0.0f; sh.eta = P.eta0; sh.s_prev = 0.0f; sh.violated = 0; sh.done = 0; sh.bad = 0; }
    __syncthreads();

    int t = 0;
    for (; t < P.t_max; ++t) {
      // ---- the three subproblems (dafs.cpp:1090-1093) -------------------------
      if (warp == 0) { const float v = nussinov_warp(P.sx, nx, P.th_s, lane); if (lane == 0) sh.part[0] = v; }
      else if (warp == 1) { const float v = nussinov_warp(P.sy, ny, P.th_s, lane); if (lane == 0) sh.part[1] = v; }
      else { const float v = nw_warp<C>(P, trz, bnd, zres, lane, &sh.bad); if (lane == 0) sh.part[2] = v; }
      if (tid == 0) sh.violated = 0;
      __syncthreads();
      // ---- tally (:1103-1117) ---------------------------------------------------
      for (int u = tid; u < P.ncbp; u += 96) {
        const int4 cb = P.cbp[u];
        const float s_w = __fsub_rn(__fsub_rn(__fadd_rn(P.sx.ent_q[cb.x], P.sy.ent_q[cb.y]),
                                              P.qz[P.cz_cell[cb.z]]), P.qz[P.cz_cell[cb.w]]);
        sw[u] = s_w;
        if (s_w > 0.0f) {
          atomicAdd(&tal_x[cb.x], 1); atomicAdd(&tal_y[cb.y], 1);
          atomicAdd(&tal_z[cb.z], 1); atomicAdd(&tal_z[cb.w], 1);
        }
      }
      __syncthreads();
      // ---- Lagrangian value, summed in CBP order by one thread (:1090-1111) ----
      if (tid == 95) {
        float s = __fadd_rn(0.0f, sh.part[0]);
        s = __fadd_rn(s, sh.part[1]);
        s = __fadd_rn(s, sh.part[2]);
        for (int u = 0; u < P.ncbp; ++u) { const float v = sw[u]; if (v > 0.0f) s = __fadd_rn(s, v); }
        sh.s = s;
      }
      // ---- multiplier updates (:1121-1254) -----------------------------------------
      const float eta = sh.eta;
      unsigned viol = 0;
      dd_update_nus(P.sx, nx.ss, tal_x, eta, viol, tid, 96);
      dd_update_nus(P.sy, ny.ss, tal_y, eta, viol, tid, 96);
      for (int i = tid; i < P.L1; i += 96) {               // z_ik == 1  (:1226-1238)
        const uint32_t k = zres[i];
        if (k != DD_NOPAIR) {
          int tz = 0;
          for (int e = P.cz_ptr[i]; e < P.cz_ptr[i + 1]; ++e)
            if (P.cz_col[e] == (int)k) { tz = tal_z[e]; break; }
          if (tz > 1) ++viol;
          float* qc = P.qz + (size_t)i * P.L2 + k;
          *qc = dd_max0(__fsub_rn(*qc, __fmul_rn(eta, (float)(1 - tz))));
        }
      }
      __syncthreads();     // tal_z of matched cells read above before it is cleared below
      for (int e = tid; e < P.nz; e += 96) {                // z_ik == 0  (:1239-1253)
        const int tz = tal_z[e];
        tal_z[e] = 0;
        const int cell = P.cz_cell[e];
        const int i = cell / P.L2, k = cell - i * P.L2;
        if (zres[i] != (uint32_t)k) {
          if (tz > 0) ++viol;
          P.qz[cell] = dd_max0(__fadd_rn(P.qz[cell], __fmul_rn(eta, (float)tz)));
        }
      }
      if (viol) atomicAdd(&sh.violated, viol);
      __syncthreads();
      if (tid == 0) {
        const unsigned violated = sh.violated;
        const float s = sh.s;
        if (P.trace) { P.trace[3 * t] = s; P.trace[3 * t + 1] = eta; P.trace[3 * t + 2] = (float)violated; }
        if (violated == 0) sh.done = 1;                     // :1278
        else {
          if (s > sh.s_prev || t == 0) {                    // :1283-1288
            const float n4 = __fmul_rn(4.0f, (float)(unsigned long long)P.ncbp);
            const float av = dd_max0(__fsub_rn(n4, (float)violated));
            sh.c = (float)((double)sh.c + (double)av / (4.0 * (double)P.ncbp));
            sh.eta = (float)((double)P.eta0 / (1.0 + (double)sh.c));
          }
          sh.s_prev = s;
        }
      }
      __syncthreads();
      if (sh.done) break;
    }
    // ---- results -------------------------------------------------------------------
    for (int k = tid; k < P.L1; k += 96) { P.sx.out[k] = nx.ss[k]; P.z_out[k] = zres[k]; }
    for (int k = tid; k < P.L2; k += 96) P.sy.out[k] = ny.ss[k];
    if (tid == 0) {
      P.result[0] = sh.s_prev;
      reinterpret_cast<int*>(P.result)[1] = t;
      reinterpret_cast<unsigned*>(P.result)[2] = sh.violated;
      reinterpret_cast<int*>(P.result)[3] = sh.bad;
    }
  }
}

}  // namespace dafs
