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
//   * Nussinov: lanes own 32 consecutive rows (a "wave", bottom rows first) and
//     walk their row left to right, one column behind the row below, whose
//     values arrive by warp shuffle; bifurcations read the lane's own row from
//     shared memory and the candidate values of the column (reference: bp[j]).
//   * NW: the mirror image (32 rows per wave, top rows first, one column behind
//     the row above); p_z and q_z come from sparse row lists, traceback codes
//     are packed 2 bits per cell.
// Multipliers are sparse: for x and y the decoder's candidates plus the CBP
// projections c_x / c_y, for z the projection c_z -- nothing else can ever
// become non-zero (see dd.cu).
#pragma once
#include <cfloat>
#include <cstdint>
#include <cuda_runtime.h>

namespace dafs {

#define DD_NOPAIR 0xffffffffu

// Header of one problem's constant block (built on the host, dd.cu).  The block
// is contiguous and position independent: the kernel copies it to shared memory
// when it fits and works on it in place in HBM otherwise.  Offsets are bytes from
// the start of the block, every section 16-byte aligned.
struct DDHdr {
  int L1, L2, nx, ny, nz, ncbp, t_max, bytes;
  float th_s, th_a, eta0, wx, wy;          // wx = w_*2*N1/(N1+N2), wy likewise (dafs.cpp:1091-1092)
  int pad0, pad1, pad2;
  // x / y: Nussinov entries in column order (ascending j, descending i: the push
  // order of the reference's bp[j], nussinov.cpp:236-245)
  uint32_t x_col_ptr, x_row_ptr, x_row_ent, x_ij, x_p, x_flag;
  uint32_t y_col_ptr, y_row_ptr, y_row_ent, y_ij, y_p, y_flag;
  // z: per row i the cells (i,k) with p_z != 0 or in c_z, ascending k
  uint32_t z_row_ptr, z_col, z_p, z_flag, z_env;
  uint32_t cbp;                            // int4 {entry in x, entry in y, z entry of (i,k), z entry of (j,l)}
  uint32_t pad3, pad4;
};

struct DDProbRef {
  const char* cst;             // constant block in HBM
  uint32_t* x_out;             // [L1]
  uint32_t* y_out;             // [L2]
  uint32_t* z_out;             // [L1]
  float* result;               // {score, iters, violated, status} as 4 x 32 bit, then 8 cycle counters (float)
  float* trace;                // optional [t_max*3] {L, eta, violated} per iteration (-v 2, dafs.cpp:1273)
};

// One Nussinov structure (x or y) of the problem a CTA works on.
struct DDNus {
  int L, n;
  const int* col_ptr;          // [L+1] entries of column j
  const int* row_ptr;          // [L+1]
  const int* row_ent;          // [n] entry ids of row i, ascending j
  const uint32_t* ent_ij;      // [n] i | j << 16
  const float* ent_p;          // [n]
  const uint8_t* ent_flag;     // [n] bit0: in c_x (dafs.cpp:1037), bit1: j-i > 2 (decoder candidate)
  float* ent_q;                // [n] multipliers q_x[i][j]
  float w;
  uint32_t* out;               // [L] result in HBM
};

// The alignment part: sparse p_z / q_z rows and the envelope.
struct DDAln {
  int L1, L2, n;
  float th;
  const int* row_ptr;          // [L1+1]
  const uint16_t* col;         // [n] 0-based column k
  const float* p;              // [n] p_z[i][k]
  const uint8_t* flag;         // [n] bit0: (i,k) in c_z (dafs.cpp:1039-1040)
  float* q;                    // [n] multipliers q_z[i][k] (zero everywhere else, see dd.cu)
  const uint2* env;            // [L1+1] {first, second}   (needleman_wunsch.cpp:198-253)
};

// Per-CTA scratch arrays; each lives in shared memory when it fits (host decides,
// in this order) and in a per-CTA global workspace otherwise.
enum {
  DD_CST, DD_Q_X, DD_Q_Y, DD_Q_Z, DD_ES_X, DD_ES_Y, DD_ROW_X, DD_ROW_Y, DD_BROW_X, DD_BROW_Y, DD_BROW_Z,
  DD_TRZ, DD_TR_X, DD_TR_Y, DD_SS_X, DD_SS_Y, DD_ZRES, DD_STK_X, DD_STK_Y, DD_TAL_X, DD_TAL_Y, DD_TAL_Z,
  DD_SW, DD_NITEMS
};
struct DDLayout {
  uint32_t off[DD_NITEMS];     // byte offset in smem or in the CTA's workspace
  uint32_t in_smem;            // bit per item
  uint32_t tr16;               // traceback codes of Nussinov need 16 bits (a column with > 250 entries)
  int lp1, lp2;                // padded row length of the Nussinov row buffers
  int w16;                     // NW traceback words per row: ceil(max L2 / 16)
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
__device__ __noinline__ float nussinov_warp(const DDNus S, const NusScratch sc, float th, int lane,
                                            unsigned long long* prof) {
  const long long t0 = clock64();
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

  const long long t1 = clock64();
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
  if (prof && lane == 0) { const long long t2 = clock64(); prof[0] += t1 - t0; prof[1] += t2 - t1; }
  return score;
}

}  // namespace dafs

namespace dafs {

struct NwScratch {
  float* brow;      // [L2+1] dp row above the current wave (row of the previous wave's last lane)
  uint32_t* trz;    // traceback codes, 2 bits per cell (0 ' ', 1 'M', 2 'X', 3 'Y'), w16 words per row
  uint32_t* zres;   // [L1] result
  int w16;
};

// SparseNeedlemanWunsch::decode(p, q, al) by one warp: lanes own 32 consecutive
// rows (a wave, top rows first) and walk their row left to right, one column
// behind the row above, whose values arrive by warp shuffle.  p_z and q_z are
// read from the sparse row lists (every other cell has p = q = 0).  Returns
// dp[L1][L2] in every lane; *bad is set if the traceback meets an unfilled cell
// (the reference would spin there, needleman_wunsch.cpp:309).
__device__ __noinline__ float nw_warp(const DDAln A, const NwScratch sc, int lane, int* bad,
                                      unsigned long long* prof) {
  const long long t0 = clock64();
  const int L1 = A.L1, L2 = A.L2, w16 = sc.w16;
  const float th = A.th;
  float score = 0.0f;
  for (int lo = 1; lo <= L1; lo += 32) {
    const int row = lo + lane;
    const bool live = row <= L1;
    int ecur = 0, eend = 0, ef = 1, el = 0;
    if (live) {
      ecur = A.row_ptr[row - 1]; eend = A.row_ptr[row];
      const uint2 ev = A.env[row];
      ef = max(1, (int)ev.x); el = (int)ev.y;
    }
    int next_k = (ecur < eend) ? (int)A.col[ecur] + 1 : 0x7fffffff;
    float cur = 0.0f;      // dp[row][k-1] entering a step (dp[row][0] = 0, :265-268)
    float diag = 0.0f;     // dp[row-1][k-1]        (dp[.][0] = 0)
    uint32_t word = 0;
    uint32_t* trow = sc.trz + row * w16;
    const int nsteps = L2 + 31;
    for (int s = 0; s < nsteps; ++s) {
      const int k = s - lane + 1;
      float up = __shfl_up_sync(0xffffffffu, cur, 1);        // dp[row-1][k]
      if (lane == 0) up = (lo > 1 && k <= L2) ? sc.brow[k] : 0.0f;   // row 0: dp[0][k] = 0 (:270-274)
      if (live && k >= 1 && k <= L2) {
        float pv = 0.0f, qv = 0.0f;
        if (k == next_k) {
          pv = A.p[ecur]; qv = A.q[ecur];
          ++ecur;
          next_k = (ecur < eend) ? (int)A.col[ecur] + 1 : 0x7fffffff;
        }
        float v;
        uint32_t code;
        if (k >= ef && k <= el) {
          v = __fadd_rn(__fsub_rn(__fadd_rn(diag, pv), th), qv);     // :282
          code = 1u;
          if (v < up) { v = up; code = 2u; }                         // :284-288
          if (v < cur) { v = cur; code = 3u; }                       // :289-293
        } else { v = -FLT_MAX; code = 0u; }                          // :262-264
        word |= code << (2 * ((k - 1) & 15));
        if (((k & 15) == 0) || k == L2) { trow[(k - 1) >> 4] = word; word = 0; }
        diag = up;
        cur = v;
        if (lane == 31) sc.brow[k] = v;
        if (row == L1 && k == L2) score = v;
      }
      __syncwarp();
    }
  }
  score = __shfl_sync(0xffffffffu, score, (L1 - 1) & 31);
  const long long t1 = clock64();
  // ---- traceback (:300-312) + alignment read-out (:317-335) ------------------
  for (int i = lane; i < L1; i += 32) sc.zres[i] = DD_NOPAIR;
  __syncwarp();
  if (lane == 0) {
    int i = L1, k = L2;
    while (i > 0 && k > 0) {       // row 0 is all 'Y', column 0 all 'X': nothing more to record
      const uint32_t code = (sc.trz[i * w16 + ((k - 1) >> 4)] >> (2 * ((k - 1) & 15))) & 3u;
      if (code == 1u) { --i; --k; sc.zres[i] = (uint32_t)k; }
      else if (code == 2u) --i;
      else if (code == 3u) --k;
      else { *bad = 1; break; }
    }
  }
  __syncwarp();
  if (prof && lane == 0) { const long long t2 = clock64(); prof[0] += t1 - t0; prof[1] += t2 - t1; }
  return score;
}

// max(0.0f, v) exactly as std::max(0.0f, v): (0.0f < v) ? v : 0.0f  (NaN -> 0)
__device__ __forceinline__ float dd_max0(float v) { return (0.0f < v) ? v : 0.0f; }

struct DDShared {
  DDNus sx, sy;
  DDAln az;
  const int4* cbp;
  int ncbp, t_max;
  float th_s, eta0;
  float* trace;
  float part[3];
  float s, s_prev, c, eta;
  unsigned violated;
  int done, bad, q;
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

#define DD_THREADS 96

__global__ void __launch_bounds__(DD_THREADS) dd_kernel(const DDArgs a) {
  extern __shared__ __align__(16) char smem[];
  __shared__ DDShared sh;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const DDLayout& lay = a.lay;
  NusScratch nx, ny;
  nx.es = dd_item<float>(lay, smem, DD_ES_X); ny.es = dd_item<float>(lay, smem, DD_ES_Y);
  nx.row = dd_item<float>(lay, smem, DD_ROW_X); ny.row = dd_item<float>(lay, smem, DD_ROW_Y);
  nx.brow = dd_item<float>(lay, smem, DD_BROW_X); ny.brow = dd_item<float>(lay, smem, DD_BROW_Y);
  nx.tr = dd_item<void>(lay, smem, DD_TR_X); ny.tr = dd_item<void>(lay, smem, DD_TR_Y);
  nx.ss = dd_item<uint32_t>(lay, smem, DD_SS_X); ny.ss = dd_item<uint32_t>(lay, smem, DD_SS_Y);
  nx.stk = dd_item<uint32_t>(lay, smem, DD_STK_X); ny.stk = dd_item<uint32_t>(lay, smem, DD_STK_Y);
  nx.lp = lay.lp1; ny.lp = lay.lp2; nx.tr16 = ny.tr16 = lay.tr16 != 0;
  NwScratch nz;
  nz.brow = dd_item<float>(lay, smem, DD_BROW_Z);
  nz.trz = dd_item<uint32_t>(lay, smem, DD_TRZ);
  nz.zres = dd_item<uint32_t>(lay, smem, DD_ZRES);
  nz.w16 = lay.w16;
  int* tal_x = dd_item<int>(lay, smem, DD_TAL_X);
  int* tal_y = dd_item<int>(lay, smem, DD_TAL_Y);
  int* tal_z = dd_item<int>(lay, smem, DD_TAL_Z);
  float* sw = dd_item<float>(lay, smem, DD_SW);
  float* q_x = dd_item<float>(lay, smem, DD_Q_X);
  float* q_y = dd_item<float>(lay, smem, DD_Q_Y);
  float* q_z = dd_item<float>(lay, smem, DD_Q_Z);
  char* cst_s = dd_item<char>(lay, smem, DD_CST);
  const bool cst_in_smem = (lay.in_smem >> DD_CST) & 1u;

  for (;;) {
    __syncthreads();
    if (tid == 0) sh.q = atomicAdd(a.queue, 1);
    __syncthreads();
    if (sh.q >= a.nprob) break;
    const DDProbRef pr = a.prob[a.order[sh.q]];
    const DDHdr* gh = reinterpret_cast<const DDHdr*>(pr.cst);
    const char* blk = pr.cst;
    if (cst_in_smem) {            // stage the constant block: coalesced 16-byte copies
      const int n16 = gh->bytes >> 4;
      const uint4* src = reinterpret_cast<const uint4*>(pr.cst);
      uint4* dst = reinterpret_cast<uint4*>(cst_s);
      for (int k = tid; k < n16; k += DD_THREADS) dst[k] = src[k];
      blk = cst_s;
    }
    __syncthreads();
    if (tid == 0) {
      const DDHdr& h = *reinterpret_cast<const DDHdr*>(blk);
      DDNus& X = sh.sx; DDNus& Y = sh.sy; DDAln& Z = sh.az;
      X.L = h.L1; X.n = h.nx; X.w = h.wx; X.out = pr.x_out; X.ent_q = q_x;
      X.col_ptr = reinterpret_cast<const int*>(blk + h.x_col_ptr);
      X.row_ptr = reinterpret_cast<const int*>(blk + h.x_row_ptr);
      X.row_ent = reinterpret_cast<const int*>(blk + h.x_row_ent);
      X.ent_ij = reinterpret_cast<const uint32_t*>(blk + h.x_ij);
      X.ent_p = reinterpret_cast<const float*>(blk + h.x_p);
      X.ent_flag = reinterpret_cast<const uint8_t*>(blk + h.x_flag);
      Y.L = h.L2; Y.n = h.ny; Y.w = h.wy; Y.out = pr.y_out; Y.ent_q = q_y;
      Y.col_ptr = reinterpret_cast<const int*>(blk + h.y_col_ptr);
      Y.row_ptr = reinterpret_cast<const int*>(blk + h.y_row_ptr);
      Y.row_ent = reinterpret_cast<const int*>(blk + h.y_row_ent);
      Y.ent_ij = reinterpret_cast<const uint32_t*>(blk + h.y_ij);
      Y.ent_p = reinterpret_cast<const float*>(blk + h.y_p);
      Y.ent_flag = reinterpret_cast<const uint8_t*>(blk + h.y_flag);
      Z.L1 = h.L1; Z.L2 = h.L2; Z.n = h.nz; Z.th = h.th_a; Z.q = q_z;
      Z.row_ptr = reinterpret_cast<const int*>(blk + h.z_row_ptr);
      Z.col = reinterpret_cast<const uint16_t*>(blk + h.z_col);
      Z.p = reinterpret_cast<const float*>(blk + h.z_p);
      Z.flag = reinterpret_cast<const uint8_t*>(blk + h.z_flag);
      Z.env = reinterpret_cast<const uint2*>(blk + h.z_env);
      sh.cbp = reinterpret_cast<const int4*>(blk + h.cbp);
      sh.ncbp = h.ncbp; sh.t_max = h.t_max; sh.th_s = h.th_s; sh.eta0 = h.eta0; sh.trace = pr.trace;
      sh.c = 0.0f; sh.eta = h.eta0; sh.s_prev = 0.0f; sh.violated = 0; sh.done = 0; sh.bad = 0;
      for (int k = 0; k < 8; ++k) sh.prof[k] = 0;
    }
    __syncthreads();
    const int L1 = sh.az.L1, L2 = sh.az.L2, ncbp = sh.ncbp, t_max = sh.t_max;

    // ---- q = 0 (dafs.cpp:1067-1069), scratch reset ----------------------------
    for (int e = tid; e < sh.sx.n; e += DD_THREADS) { q_x[e] = 0.0f; nx.es[e] = -INFINITY; tal_x[e] = 0; }
    for (int e = tid; e < sh.sy.n; e += DD_THREADS) { q_y[e] = 0.0f; ny.es[e] = -INFINITY; tal_y[e] = 0; }
    for (int e = tid; e < sh.az.n; e += DD_THREADS) { q_z[e] = 0.0f; tal_z[e] = 0; }
    for (int k = tid; k < L1; k += DD_THREADS) { nx.ss[k] = DD_NOPAIR; nz.zres[k] = DD_NOPAIR; }
    for (int k = tid; k < L2; k += DD_THREADS) ny.ss[k] = DD_NOPAIR;
    __syncthreads();

    int t = 0;
    for (; t < t_max; ++t) {
      // ---- the three subproblems (dafs.cpp:1090-1093) -------------------------
      const long long tp0 = clock64();
      if (warp == 0) { const float v = nussinov_warp(sh.sx, nx, sh.th_s, lane, sh.prof + 0); if (lane == 0) sh.part[0] = v; }
      else if (warp == 1) { const float v = nussinov_warp(sh.sy, ny, sh.th_s, lane, sh.prof + 2); if (lane == 0) sh.part[1] = v; }
      else { const float v = nw_warp(sh.az, nz, lane, &sh.bad, sh.prof + 4); if (lane == 0) sh.part[2] = v; }
      if (tid == 0) sh.violated = 0;
      __syncthreads();
      const long long tp1 = clock64();
      // ---- tally (:1103-1117) ---------------------------------------------------
      for (int u = tid; u < ncbp; u += DD_THREADS) {
        const int4 cb = sh.cbp[u];
        const float s_w = __fsub_rn(__fsub_rn(__fadd_rn(q_x[cb.x], q_y[cb.y]), q_z[cb.z]), q_z[cb.w]);
        sw[u] = s_w;
        if (s_w > 0.0f) {
          atomicAdd(&tal_x[cb.x], 1); atomicAdd(&tal_y[cb.y], 1);
          atomicAdd(&tal_z[cb.z], 1); atomicAdd(&tal_z[cb.w], 1);
        }
      }
      __syncthreads();
      // ---- Lagrangian value, summed in CBP order by one thread (:1090-1111) ----
      if (tid == DD_THREADS - 1) {
        float s = __fadd_rn(0.0f, sh.part[0]);
        s = __fadd_rn(s, sh.part[1]);
        s = __fadd_rn(s, sh.part[2]);
        for (int u = 0; u < ncbp; ++u) { const float v = sw[u]; if (v > 0.0f) s = __fadd_rn(s, v); }
        sh.s = s;
      }
      // ---- multiplier updates (:1121-1254) -----------------------------------------
      const float eta = sh.eta;
      unsigned viol = 0;
      dd_update_nus(sh.sx, nx.ss, tal_x, eta, viol, tid, DD_THREADS);
      dd_update_nus(sh.sy, ny.ss, tal_y, eta, viol, tid, DD_THREADS);
      // z (:1223-1254), one thread per c_z entry.  Cells outside c_z have t_z = 0
      // and q_z = 0, and stay there: max(0, 0 - eta*(1-0)) = 0 for eta >= 0 or NaN
      // (the host rejects eta0 < 0).
      {
        const DDAln& Z = sh.az;
        for (int i = tid; i < L1; i += DD_THREADS) {
          const uint32_t zi = nz.zres[i];
          for (int e = Z.row_ptr[i]; e < Z.row_ptr[i + 1]; ++e) {
            if (!(Z.flag[e] & 1)) continue;
            const int tz = tal_z[e];
            tal_z[e] = 0;
            if (zi == (uint32_t)Z.col[e]) {                       // z_ik == 1  (:1226-1238)
              if (tz > 1) ++viol;
              q_z[e] = dd_max0(__fsub_rn(q_z[e], __fmul_rn(eta, (float)(1 - tz))));
            } else {                                               // z_ik == 0  (:1239-1253)
              if (tz > 0) ++viol;
              q_z[e] = dd_max0(__fadd_rn(q_z[e], __fmul_rn(eta, (float)tz)));
            }
          }
        }
      }
      if (viol) atomicAdd(&sh.violated, viol);
      __syncthreads();
      if (tid == 0) {
        const unsigned violated = sh.violated;
        const float s = sh.s;
        if (sh.trace) { sh.trace[3 * t] = s; sh.trace[3 * t + 1] = eta; sh.trace[3 * t + 2] = (float)violated; }
        if (violated == 0) sh.done = 1;                     // :1278
        else {
          if (s > sh.s_prev || t == 0) {                    // :1283-1288
            const float n4 = __fmul_rn(4.0f, (float)(unsigned long long)ncbp);
            const float av = dd_max0(__fsub_rn(n4, (float)violated));
            sh.c = (float)((double)sh.c + (double)av / (4.0 * (double)ncbp));
            sh.eta = (float)((double)sh.eta0 / (1.0 + (double)sh.c));
          }
          sh.s_prev = s;
        }
      }
      __syncthreads();
      if (tid == 0) { sh.prof[6] += tp1 - tp0; sh.prof[7] += clock64() - tp1; }
      if (sh.done) break;
    }
    // ---- results -------------------------------------------------------------------
    for (int k = tid; k < L1; k += DD_THREADS) { pr.x_out[k] = nx.ss[k]; pr.z_out[k] = nz.zres[k]; }
    for (int k = tid; k < L2; k += DD_THREADS) pr.y_out[k] = ny.ss[k];
    if (tid == 0) {
      pr.result[0] = sh.s_prev;
      reinterpret_cast<int*>(pr.result)[1] = t;
      reinterpret_cast<unsigned*>(pr.result)[2] = sh.violated;
      reinterpret_cast<int*>(pr.result)[3] = sh.bad;
      for (int k = 0; k < 8; ++k) pr.result[4 + k] = (float)sh.prof[k];
    }
  }
}

}  // namespace dafs
