// K6-K9: the probabilistic consistency transformation of DAFS::run()
// (src/dafs.cpp:1796-1827) on device-resident posteriors:
//   K6 transpose_mp                     src/dafs.cpp:155-167
//   K7 calculate_similarity_score       src/dafs.cpp:713-764
//   K8 relax_basepairing_probability    src/dafs.cpp:326-375
//   K9 relax_matching_probability       src/dafs.cpp:258-324
// Bit-exact with the reference's x86-64 build: every output cell is accumulated in
// the reference's loop order with unfused FP32 multiplies / adds, the weights follow
// its mixed float / double arithmetic, the similarity DP reproduces its strict-'<'
// tie-breaking.
//
// Matrices.  mp_[a][b] for a < b is the caller's CSR (U), for a > b its transpose
// (T, built by K6, same entry count), for a == b the identity (align.cpp:41-44,
// never stored).  Rows of a relaxed matrix are independent, so K8 / K9 give one
// warp a ROW of one output matrix: the row is accumulated in shared memory in the
// reference's order -- the reference's loops over (z, k) [(y, k, l)] visit a cell
// (i, j) in ascending z, then ascending k, and that is also the order in which
// row i of mp_[x][z] lists its k -- then thresholded and appended to a pool; a
// scan + gather pass puts rows in CSR order.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace dafs {

#ifndef PCT_MP_MINB
#define PCT_MP_MINB 3        // resident CTAs of 256 threads per SM for K9 (register cap 80)
#endif
#define PCT_CUTOFF 0.01   // src/dafs.cpp:47, a double constant

struct PctMat {             // one mp_[a][b]: rows of a, columns of b
  const int* row_ptr;       // nullptr: identity
  const uint32_t* col;
  const float* val;
};

// Where mp_[a][b] lives: offsets into the combined arrays: U first, then T, then the n
// identity matrices mp_[a][a] (align.cpp:41-44) written out as CSR like the others.
struct PctDesc { int64_t rp, nz; };

struct PctArgs {
  int n;                    // sequences
  const int* len;           // [n]
  // mp: U = pairs a < b in row-major order, T = their transposes; both in one set of
  // arrays (rp_all = [U row_ptr blocks | T row_ptr blocks], col_all / val_all likewise)
  const int64_t* u_row_off; // [npairs+1] into u_row_ptr (len[a]+1 per pair)
  const int64_t* t_row_off; // [npairs+1] into t_row_ptr (len[b]+1 per pair)
  const int64_t* nz_off;    // [npairs+1] entries of pair u in both U and T
  const int* u_row_ptr; const uint32_t* u_col; const float* u_val;
  int* t_row_ptr; uint32_t* t_col; float* t_val;
  const int* rp_all; const uint32_t* col_all; const float* val_all;
  uint2* ent_all;           // the same entries interleaved, {column, value bits}: one 8-byte load per entry
  int64_t ru, nnz;          // sizes of the U parts of rp_all / col_all
  PctDesc* desc;            // [n*n]  desc[a*n+b]  = mp_[a][b]
  PctDesc* descT;           // [n*n]  descT[b*n+a] = mp_[a][b]
  // per row of rp_all, then per row of the n identity blocks (len+1 each, at ru + rt +
  // b_row_off[a]): {entries, k0 | k1 << 16, p0, p1}.  A missing k is the number of columns,
  // which in the matrix that k indexes next is the spare last row: its head is empty.
  uint4* head;
  int64_t rt;
  // bp: one CSR per sequence
  const int64_t* b_row_off; // [n+1] (len+1 per sequence)
  const int64_t* b_nz_off;  // [n+1]
  const int* b_row_ptr; const uint32_t* b_col; const float* b_val;
  float* sim;               // [n*n]
};

__device__ __forceinline__ int64_t pct_pair_index(int n, int a, int b) {   // a < b
  return (int64_t)a * n - (int64_t)a * (a + 1) / 2 + (b - a - 1);
}

__global__ void pct_desc_kernel(PctArgs A) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= A.n * A.n) return;
  const int a = t / A.n, b = t % A.n;
  PctDesc d;
  if (a == b) { d.rp = A.ru + A.rt + A.b_row_off[a]; d.nz = 2 * A.nnz + A.b_row_off[a] - a; }
  else if (a < b) { const int64_t u = pct_pair_index(A.n, a, b); d.rp = A.u_row_off[u]; d.nz = A.nz_off[u]; }
  else { const int64_t u = pct_pair_index(A.n, b, a); d.rp = A.ru + A.t_row_off[u]; d.nz = A.nnz + A.nz_off[u]; }
  A.desc[(int64_t)a * A.n + b] = d;
  A.descT[(int64_t)b * A.n + a] = d;
}

__device__ __forceinline__ PctMat pct_mat(const PctArgs& A, int a, int b) {
  const PctDesc d = A.desc[(int64_t)a * A.n + b];
  PctMat m;
  m.row_ptr = A.rp_all + d.rp; m.col = A.col_all + d.nz; m.val = A.val_all + d.nz;
  return m;
}

__device__ __forceinline__ int pct_warp_incl_scan(int v, int lane) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += y; }
  return v;
}

// ---- K6: transpose_mp ------------------------------------------------------------
// One warp per pair.  Column counts by atomics, a warp scan, then the rows of U are
// walked in ascending order so that every row of T lists its columns ascending (the
// reference sorts (column, value) pairs whose columns are distinct).  cursor:
// scratch, same size as t_row_ptr.
__global__ void __launch_bounds__(128) pct_transpose_kernel(PctArgs A, int64_t npairs, const int2* __restrict__ pairs,
                                                            int* __restrict__ cursor) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t u = wid; u < npairs; u += nw) {
    const int2 pr = pairs[u];
    const int L1 = A.len[pr.x], L2 = A.len[pr.y];
    const int* rp = A.u_row_ptr + A.u_row_off[u];
    const uint32_t* col = A.u_col + A.nz_off[u];
    const float* val = A.u_val + A.nz_off[u];
    int* trp = A.t_row_ptr + A.t_row_off[u];
    int* cur = cursor + A.t_row_off[u];
    uint32_t* tcol = A.t_col + A.nz_off[u];
    float* tval = A.t_val + A.nz_off[u];
    const int nnz = rp[L1];
    uint2* uent = A.ent_all + A.nz_off[u];
    uint2* tent = A.ent_all + A.nnz + A.nz_off[u];
    for (int k = lane; k < nnz; k += 32) uent[k] = make_uint2(col[k], __float_as_uint(val[k]));
    for (int j = lane; j <= L2; j += 32) trp[j] = 0;
    __syncwarp();
    for (int k = lane; k < nnz; k += 32) atomicAdd(&trp[col[k] + 1], 1);
    __syncwarp();
    int carry = 0;
    for (int j0 = 0; j0 <= L2; j0 += 32) {
      const int j = j0 + lane;
      const int c = j <= L2 ? __ldcg(trp + j) : 0;
      const int incl = pct_warp_incl_scan(c, lane) + carry;
      if (j <= L2) { trp[j] = incl; cur[j] = incl; }       // cur[j] = start of row j (trp[j] counts columns < j)
      carry = __shfl_sync(0xffffffffu, incl, 31);
    }
    __syncwarp();
    for (int i = 0; i < L1; ++i) {
      const int s = rp[i], e = rp[i + 1];
      for (int k = s + lane; k < e; k += 32) {
        const uint32_t j = col[k];
        const int pos = __ldcg(cur + j);
        tcol[pos] = (uint32_t)i; tval[pos] = val[k];
        tent[pos] = make_uint2((unsigned)i, __float_as_uint(val[k]));
        __stcg(cur + j, pos + 1);
      }
      __syncwarp();
    }
    // row heads of U and T: entry count and the first two entries of every row
    uint4* hu = A.head + A.u_row_off[u];
    for (int i = lane; i <= L1; i += 32) {
      const int b = rp[min(i, L1 - 1)], c = i < L1 ? rp[i + 1] - b : 0;
      uint4 h = make_uint4((unsigned)c, (unsigned)L2 | ((unsigned)L2 << 16), 0u, 0u);
      if (c > 0) { h.y = col[b] | ((unsigned)L2 << 16); h.z = __float_as_uint(val[b]); }
      if (c > 1) { h.y = col[b] | (col[b + 1] << 16); h.w = __float_as_uint(val[b + 1]); }
      hu[i] = h;
    }
    uint4* ht = A.head + A.ru + A.t_row_off[u];
    for (int j = lane; j <= L2; j += 32) {
      const int b = __ldcg(trp + min(j, L2 - 1)), c = j < L2 ? __ldcg(trp + j + 1) - b : 0;
      uint4 h = make_uint4((unsigned)c, (unsigned)L1 | ((unsigned)L1 << 16), 0u, 0u);
      if (c > 0) { h.y = __ldcg(tcol + b) | ((unsigned)L1 << 16); h.z = __float_as_uint(__ldcg(tval + b)); }
      if (c > 1) { h.y = __ldcg(tcol + b) | (__ldcg(tcol + b + 1) << 16); h.w = __float_as_uint(__ldcg(tval + b + 1)); }
      ht[j] = h;
    }
  }
}

// the identity matrices mp_[a][a]: CSR rows {(r, 1.0f)} and heads {1, r, 1.0f}, spare row empty
__global__ void pct_identity_kernel(PctArgs A, int* __restrict__ rp_all, uint32_t* __restrict__ col_all,
                                    float* __restrict__ val_all) {
  const int a = blockIdx.x;
  const int L = A.len[a];
  const int64_t rb = A.ru + A.rt + A.b_row_off[a], zb = 2 * A.nnz + A.b_row_off[a] - a;
  for (int r = threadIdx.x; r <= L; r += blockDim.x) {
    A.head[rb + r] = r < L ? make_uint4(1u, (unsigned)r | ((unsigned)L << 16), 0x3f800000u, 0u)
                           : make_uint4(0u, (unsigned)L | ((unsigned)L << 16), 0u, 0u);
    rp_all[rb + r] = r;
    if (r < L) { col_all[zb + r] = (uint32_t)r; val_all[zb + r] = 1.0f; A.ent_all[zb + r] = make_uint2((unsigned)r, 0x3f800000u); }
  }
}

// ---- K7: calculate_similarity_score ------------------------------------------------
// dp[i][j] = first maximum, in the order {diagonal + p (entries only), left, up}, of
// its predecessors, and tr[i][j] = tr[chosen] + 1 (:721-760).  "left" beats "up" and
// "diagonal" beats "left" on ties, so along a row the running (dp, tr) is replaced
// by column j's own candidate c_j = (max(diag + p, up) with diag winning ties)
// exactly when c_j.v > running (candidate from "up") or c_j.v >= running (from
// "diagonal"), and otherwise tr grows by one.  The running dp is a prefix maximum,
// and the tr of column j is tr(c_m) + (j - m) for the last accepted column m <= j:
// two warp scans per 32 columns.  One warp per pair; rows of dp / tr and the row's
// match values in per-warp scratch: 2*(W) floats + 2*(W) ints + (W) floats, W = maxL+2.
__global__ void __launch_bounds__(128) pct_similarity_kernel(PctArgs A, int64_t npairs, const int2* __restrict__ pairs,
                                                             float* __restrict__ scratch, int W) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
  float* base = scratch + wid * (int64_t)W * 5;
  float* dpa = base; float* dpb = base + W;
  int* tra = reinterpret_cast<int*>(base + 2 * W); int* trb = reinterpret_cast<int*>(base + 3 * W);
  float* pv = base + 4 * W;
  for (int64_t u = wid; u < npairs; u += nw) {
    const int2 pr = pairs[u];
    const int L1 = A.len[pr.x], L2 = A.len[pr.y];
    const int* rp = A.u_row_ptr + A.u_row_off[u];
    const uint32_t* col = A.u_col + A.nz_off[u];
    const float* val = A.u_val + A.nz_off[u];
    for (int j = lane; j <= L2; j += 32) { dpa[j] = 0.0f; tra[j] = 0; pv[j] = -1.0f; }
    __syncwarp();
    float* dpp = dpa; float* dpn = dpb; int* trp = tra; int* trn = trb;
    for (int i = 1; i <= L1; ++i) {
      const int s = rp[i - 1], e = rp[i];
      for (int k = s + lane; k < e; k += 32) pv[col[k] + 1] = val[k];      // match value of column j = col + 1
      __syncwarp();
      float cV = 0.0f; int cT = 0;                                          // column 0: dp = 0, tr = 0
      if (lane == 0) { dpn[0] = 0.0f; trn[0] = 0; }
      for (int j0 = 1; j0 <= L2; j0 += 32) {
        const int j = j0 + lane;
        const bool in = j <= L2;
        float v = -1.0f; int t = 0; bool strict = true;
        if (in) {
          const float up = dpp[j]; const int tu = trp[j];
          const float p = pv[j];
          v = up; t = tu + 1;
          if (p >= 0.0f) {
            const float dg = __fadd_rn(dpp[j - 1], p);
            if (!(dg < up)) { v = dg; t = trp[j - 1] + 1; strict = false; }
          }
        }
        // exclusive prefix maximum of the candidates, seeded with the running value
        float incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const float y = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl = fmaxf(incl, y); }
        float excl = __shfl_up_sync(0xffffffffu, incl, 1);
        excl = lane == 0 ? cV : fmaxf(excl, cV);
        const bool acc = in && (strict ? (excl < v) : !(v < excl));
        int m = acc ? lane : -1;                                            // last accepted lane <= me
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, m, d); if (lane >= d) m = max(m, y); }
        const float vm = __shfl_sync(0xffffffffu, v, m < 0 ? 0 : m);
        const int tm = __shfl_sync(0xffffffffu, t, m < 0 ? 0 : m);
        const float V = m < 0 ? cV : vm;
        const int T = m < 0 ? cT + lane + 1 : tm + (lane - m);
        if (in) { dpn[j] = V; trn[j] = T; }
        const int last = min(31, L2 - j0);
        cV = __shfl_sync(0xffffffffu, V, last); cT = __shfl_sync(0xffffffffu, T, last);
      }
      for (int k = s + lane; k < e; k += 32) pv[col[k] + 1] = -1.0f;
      __syncwarp();
      { float* x = dpp; dpp = dpn; dpn = x; int* y = trp; trp = trn; trn = y; }
    }
    if (lane == 0) {
      const float sc = __fdiv_rn(dpp[L2], (float)trp[L2]);
      A.sim[(int64_t)pr.x * A.n + pr.y] = sc;
      A.sim[(int64_t)pr.y * A.n + pr.x] = sc;
    }
    __syncwarp();
  }
}

// unit of a global row: off[u] <= row < off[u+1]; hint = the previous answer
__device__ __forceinline__ int64_t pct_find_unit(const int64_t* __restrict__ off, int64_t n, int64_t row, int64_t hint) {
  if (off[hint] <= row && row < off[hint + 1]) return hint;
  int64_t lo = 0, hi = n - 1;
  while (lo < hi) { const int64_t mid = (lo + hi + 1) >> 1; if (off[mid] <= row) lo = mid; else hi = mid - 1; }
  return lo;
}

// Input check (one warp per unit): row_ptr starts at 0, never decreases, ends at the
// unit's entry count; columns ascend strictly inside a row and stay below ncols.
__global__ void pct_validate_kernel(int64_t nunits, const int64_t* __restrict__ row_off, const int64_t* __restrict__ nz_off,
                                    const int* __restrict__ row_ptr, const uint32_t* __restrict__ col,
                                    const int* __restrict__ ncols, int* __restrict__ bad) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t u = wid; u < nunits; u += nw) {
    const int* rp = row_ptr + row_off[u];
    const int64_t rows = row_off[u + 1] - row_off[u] - 1;
    const int64_t nnz = nz_off[u + 1] - nz_off[u];
    const uint32_t* c = col + nz_off[u];
    const uint32_t nc = (uint32_t)ncols[u];
    bool ok = rows >= 0 && nnz >= 0 && rp[0] == 0 && rp[rows] == nnz;
    for (int64_t r = lane; ok && r < rows; r += 32) {
      const int s = rp[r], e = rp[r + 1];
      if (s > e || s < 0 || e > nnz) { ok = false; break; }
      for (int k = s; k < e; ++k)
        if (c[k] >= nc || (k > s && c[k] <= c[k - 1])) ok = false;
    }
    if (!ok) atomicExch(bad, 1);
  }
}

__global__ void pct_sim_diag_kernel(int n, float* __restrict__ sim) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) sim[(int64_t)i * n + i] = 1.0f;                 // :1816
}

// ---- weights ----------------------------------------------------------------------
// relax_basepairing_probability, :341-348: w(x, y) and the float sum over y in order.
__global__ void pct_weights_bp_kernel(int n, const float* __restrict__ sim, float w_pct_s, float* __restrict__ w,
                                      float* __restrict__ sum_w) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n) return;
  float sw = 0.0f;
  for (int y = 0; y < n; ++y) {
    float v = sim[(int64_t)y * n + x];
    if (w_pct_s < 0.0) v = (float)((double)v * (1.0 / (double)(unsigned)n));
    else if (y == x) v = (float)((double)v * (1.0 - (double)w_pct_s));
    else v = __fmul_rn(v, __fdiv_rn(w_pct_s, (float)(unsigned)(n - 1)));
    w[(int64_t)x * n + y] = v;
    sw = __fadd_rn(sw, v);
  }
  sum_w[x] = sw;
}

// relax_matching_probability, :280-288: w(x, y, z) and the float sum over z in order.
__global__ void pct_weights_mp_kernel(int n, int64_t npairs, const int2* __restrict__ pairs,
                                      const float* __restrict__ sim, float w_pct_a, float* __restrict__ w,
                                      float* __restrict__ sum_w) {
  const int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= npairs) return;
  const int x = pairs[u].x, y = pairs[u].y;
  float sw = 0.0f;
  for (int z = 0; z < n; ++z) {
    float v = __fmul_rn(sim[(int64_t)z * n + x], sim[(int64_t)z * n + y]);
    if (w_pct_a < 0.0) v = (float)((double)v * (1.0 / (double)(unsigned)n));
    else if (z == x || z == y) v = (float)((double)v * ((1.0 - (double)w_pct_a) / 2));
    else v = __fmul_rn(v, __fdiv_rn(w_pct_a, (float)(unsigned)(n - 2)));
    w[u * n + z] = v;
    sw = __fadd_rn(sw, v);
  }
  sum_w[u] = sw;
}

// ---- row emission: acc / sum_w > CUTOFF, ascending columns, into the pool -----------
struct PctOut {
  int64_t* row_pool;        // [total rows] where the row's entries start in the pool
  int* row_cnt;             // [total rows]
  unsigned long long* pool_count;
  uint32_t* pool_col; float* pool_val; int64_t pool_cap;
};

__device__ __forceinline__ void pct_emit_row(const PctOut& O, int64_t row, float* acc, int jlo, int jhi, float sum_w,
                                             int lane) {
  int cnt = 0;
  for (int j0 = jlo; j0 < jhi; j0 += 32) {
    const int j = j0 + lane;
    const bool keep = j < jhi && (double)__fdiv_rn(acc[j], sum_w) > PCT_CUTOFF;
    cnt += __popc(__ballot_sync(0xffffffffu, keep));
  }
  unsigned long long b = 0;
  if (lane == 0) { b = atomicAdd(O.pool_count, (unsigned long long)cnt); O.row_pool[row] = (int64_t)b; O.row_cnt[row] = cnt; }
  const int64_t base = (int64_t)__shfl_sync(0xffffffffu, b, 0);
  const bool fits = base + cnt <= O.pool_cap;
  int run = 0;
  const unsigned lt = (1u << lane) - 1u;
  for (int j0 = jlo; j0 < jhi; j0 += 32) {
    const int j = j0 + lane;
    float v = 0.0f;
    bool keep = false;
    if (j < jhi) { v = __fdiv_rn(acc[j], sum_w); keep = (double)v > PCT_CUTOFF; }
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (keep && fits) { const int64_t k = base + run + __popc(m & lt); O.pool_col[k] = (uint32_t)j; O.pool_val[k] = v; }
    run += __popc(m);
  }
}

// ---- K9: relax_matching_probability ---------------------------------------------------
// One THREAD per row i of one pair (x < y); a warp holds 32 consecutive rows of the same
// pair (task = (pair, block of 32 rows), listed by the host in tiles of pairs so that
// the matrices a tile reads stay in L2).  For z = 0..n-1 in order: the k's of row i of
// mp_[x][z] in ascending order, for each the row k of mp_[z][y],
// acc[j] += (p_ik * p_jk) * w -- the reference's order for every cell.
// Rows are read through their 16-byte heads (count + first two entries), so a z step is
// a chain of two loads; the chain is software-pipelined over z (head of row i for z+2,
// heads of its k's for z+1, arithmetic for z).  Rows with more than two entries take
// the CSR for the rest.
// The accumulator row of a thread lives in a per-warp scratch of W x 32 floats in
// global memory, skewed so that the 32 rows of a warp meet on one 128-byte line when
// they follow the same diagonal (cell (lane l, column j) at ((j - l) mod W) * 32 + l);
// it is zeroed lazily over the range of columns the row has touched, which is also the
// range the read-out scans.
struct PctAcc {
  float* a; int W, lane, jmin, jw;                          // touched columns: [jmin, jmin + jw)
  __device__ __forceinline__ int idx(int j) const { int c = j - lane; if (c < 0) c += W; return c * 32 + lane; }
  __device__ __forceinline__ void add(int j, float t);
};

// first touch of a row, or a column outside the touched range: zero the new cells.
// Out of line and by value (the hot loop keeps PctAcc in registers); returns (jmin, jw).
__device__ __noinline__ int2 pct_acc_extend(float* a, int W, int lane, int jmin, int jw, int j) {
  PctAcc t; t.a = a; t.W = W; t.lane = lane;
  if (jw == 0) { a[t.idx(j)] = 0.0f; return make_int2(j, 1); }
  if (j < jmin) { for (int q = j; q < jmin; ++q) a[t.idx(q)] = 0.0f; return make_int2(j, jw + jmin - j); }
  for (int q = jmin + jw; q <= j; ++q) a[t.idx(q)] = 0.0f;
  return make_int2(jmin, j - jmin + 1);
}

__device__ __forceinline__ void PctAcc::add(int j, float t) {
  if ((unsigned)(j - jmin) >= (unsigned)jw) { const int2 r = pct_acc_extend(a, W, lane, jmin, jw, j); jmin = r.x; jw = r.y; }
  const int k = idx(j);
  a[k] = __fadd_rn(a[k], t);                                // 0 + t = t on the first touch
}

struct PctPtrs { const int* rp; const uint32_t* col; const float* val; const uint4* head; };

// entries 3.. of row k of mp_[z][y] (rare: sharp posteriors keep one or two per row)
__device__ __noinline__ int2 pct_mp_row_tail(PctPtrs P, PctAcc acc, int64_t rp, int64_t nz, int k, float p_ik, int cnt, float wz) {
  const int bs = P.rp[rp + k];
  for (int c = bs + 2; c < bs + cnt; ++c)
    acc.add((int)P.col[nz + c], __fmul_rn(__fmul_rn(p_ik, P.val[nz + c]), wz));
  return make_int2(acc.jmin, acc.jw);
}

// all products of one k: row k of mp_[z][y] through its head hb
__device__ __forceinline__ void pct_mp_row(const PctPtrs& P, PctAcc& acc, const PctDesc& dB, int k, float p_ik,
                                           const uint4& hb, float wz) {
  if (hb.x >= 1u) acc.add((int)(hb.y & 0xffffu), __fmul_rn(__fmul_rn(p_ik, __uint_as_float(hb.z)), wz));
  if (hb.x >= 2u) acc.add((int)(hb.y >> 16), __fmul_rn(__fmul_rn(p_ik, __uint_as_float(hb.w)), wz));
  if (hb.x > 2u) { const int2 r = pct_mp_row_tail(P, acc, dB.rp, dB.nz, k, p_ik, (int)hb.x, wz); acc.jmin = r.x; acc.jw = r.y; }
}

// entries 3.. of row i of mp_[x][z]
__device__ __noinline__ int2 pct_mp_tail(PctPtrs P, PctAcc acc, int64_t arp, int64_t anz, int64_t brp, int64_t bnz, int i,
                                         int cnt, float wz) {
  const int s = P.rp[arp + i];
  PctDesc db; db.rp = brp; db.nz = bnz;
  for (int ke = s + 2; ke < s + cnt; ++ke) {
    const int k = (int)P.col[anz + ke];
    const uint4 hb = P.head[brp + k];
    pct_mp_row(P, acc, db, k, P.val[anz + ke], hb, wz);
  }
  return make_int2(acc.jmin, acc.jw);
}

__global__ void __launch_bounds__(256, PCT_MP_MINB) pct_relax_mp_kernel(PctArgs A, int64_t ntasks, const int2* __restrict__ tasks,
                                                           const int2* __restrict__ pairs,
                                                           const int64_t* __restrict__ out_row_off,
                                                           const float* __restrict__ w, const float* __restrict__ sum_w,
                                                           PctOut O, float* __restrict__ scratch, int W) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int n = A.n;
  PctAcc acc;
  acc.a = scratch + wid * (int64_t)W * 32; acc.W = W; acc.lane = lane;
  const uint4* __restrict__ head = A.head;
  PctPtrs P; P.rp = A.rp_all; P.col = A.col_all; P.val = A.val_all; P.head = A.head;
  for (int64_t task = wid; task < ntasks; task += nw) {
    const int2 tk = tasks[task];
    const int64_t u = tk.x;
    const int x = pairs[u].x, y = pairs[u].y;
    const int L1 = A.len[x];
    const int i = tk.y * 32 + lane;
    const bool live = i < L1;
    const int ii = live ? i : L1;                            // the spare row: empty head
    acc.jmin = 0; acc.jw = 0;
    const float* wu = w + u * n;
    const PctDesc* dA = A.desc + (int64_t)x * n;             // mp_[x][z]
    const PctDesc* dB = A.descT + (int64_t)y * n;            // mp_[z][y]
    // pipeline: ha2 = head of row i for z+2, ha1 for z+1 (heads of its k's in flight), ha0 / hb0x for z
    uint4 ha0 = __ldg(head + dA[0].rp + ii);
    uint4 ha1 = __ldg(head + dA[n > 1 ? 1 : 0].rp + ii);
    PctDesc db0 = dB[0];
    uint4 hb00 = __ldg(head + db0.rp + (ha0.y & 0xffffu));
    uint4 hb01 = __ldg(head + db0.rp + (ha0.y >> 16));
    for (int z = 0; z < n; ++z) {
      const float wz = wu[z];
      // loads for the steps ahead (past the end: repeated, never used)
      const uint4 ha2 = __ldg(head + dA[min(z + 2, n - 1)].rp + ii);
      const PctDesc db1 = dB[min(z + 1, n - 1)];
      const uint4 hb10 = __ldg(head + db1.rp + (ha1.y & 0xffffu));
      const uint4 hb11 = __ldg(head + db1.rp + (ha1.y >> 16));
      // step z
      if (ha0.x >= 1u) pct_mp_row(P, acc, db0, (int)(ha0.y & 0xffffu), __uint_as_float(ha0.z), hb00, wz);
      if (ha0.x >= 2u) pct_mp_row(P, acc, db0, (int)(ha0.y >> 16), __uint_as_float(ha0.w), hb01, wz);
      if (ha0.x > 2u) {
        const PctDesc da = dA[z];
        const int2 r = pct_mp_tail(P, acc, da.rp, da.nz, db0.rp, db0.nz, i, (int)ha0.x, wz);
        acc.jmin = r.x; acc.jw = r.y;
      }
      ha0 = ha1; ha1 = ha2; db0 = db1; hb00 = hb10; hb01 = hb11;
    }
    // read-out: v = acc / sum_w > CUTOFF, ascending j (:302-311)
    const float sw = sum_w[u];
    int cnt = 0;
    for (int j = acc.jmin; j < acc.jmin + acc.jw; ++j) cnt += (double)__fdiv_rn(acc.a[acc.idx(j)], sw) > PCT_CUTOFF;
    const int incl = pct_warp_incl_scan(cnt, lane);
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    unsigned long long b = 0;
    if (lane == 0 && total) b = atomicAdd(O.pool_count, (unsigned long long)total);
    const int64_t base = (int64_t)__shfl_sync(0xffffffffu, b, 0) + (incl - cnt);
    if (live) {
      const int64_t row = out_row_off[u] + i;
      O.row_pool[row] = base; O.row_cnt[row] = cnt;
      if (base + cnt <= O.pool_cap) {
        int64_t k = base;
        for (int j = acc.jmin; j < acc.jmin + acc.jw; ++j) {
          const float v = __fdiv_rn(acc.a[acc.idx(j)], sw);
          if ((double)v > PCT_CUTOFF) { O.pool_col[k] = (uint32_t)j; O.pool_val[k] = v; ++k; }
        }
      }
    }
    __syncwarp();
  }
}

// ---- K9, rows with many entries --------------------------------------------------------
// Unrelated sequences give rows of 5-10 entries, i.e. ~50 products per (row, z): the
// accumulator rows are then dense and must stay on chip.  G lanes share a row (32 / G
// rows per warp), the accumulator row is in shared memory (SMEM) or, for sequences too
// long for that, in global scratch.  Per z: the k's of row i are taken one per step in
// ascending order, the G lanes take the entries of row k of mp_[z][y] (distinct
// columns, so no two lanes meet on a cell inside a step; __syncwarp between steps).
// The three dependent loads that lead to a step (row bounds of i, its k's, their row
// bounds) are software-pipelined over z.
// Every warp takes exactly one unit (32 / G rows of one pair) and walks z = 0..n-1 once.
// Units are numbered in the host's tile order and CTAs start in that order, so the
// pairs in flight at any time come from one or two tiles, whose matrices fit in L2
// (unit0: first unit of this launch -- the global-scratch variant is launched in waves
// of resident warps because its scratch is sized for one wave).
template <int G, bool SMEM>
__global__ void __launch_bounds__(256) pct_relax_mp_dense_kernel(PctArgs A, int64_t unit0, int64_t ntasks,
                                                                 const int2* __restrict__ tasks,
                                                                 const int2* __restrict__ pairs,
                                                                 const int64_t* __restrict__ out_row_off,
                                                                 const float* __restrict__ w, const float* __restrict__ sum_w,
                                                                 PctOut O, float* __restrict__ gacc, int Wp) {
  extern __shared__ float pct_smem[];
  constexpr int RW = 32 / G;
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int g = lane / G, q = lane % G;
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + wib;
  const int64_t unit = unit0 + wid;                          // (task, block of RW rows inside its 32)
  if (unit >= ntasks * (32 / RW)) return;
  float* acc = (SMEM ? pct_smem + (size_t)wib * RW * Wp : gacc + wid * (int64_t)(RW * Wp)) + (size_t)g * Wp;
  const int n = A.n;
  const int* __restrict__ rp = A.rp_all;
  const uint2* __restrict__ ent = A.ent_all;
  const int2 tk = tasks[unit / (32 / RW)];
  const int64_t u = tk.x;
  const int x = pairs[u].x, y = pairs[u].y;
  const int L1 = A.len[x], L2 = A.len[y];
  const int i0 = tk.y * 32 + (int)(unit % (32 / RW)) * RW;
  if (i0 >= L1) return;
  const float* wu = w + u * n;
  const PctDesc* dA = A.desc + (int64_t)x * n;               // mp_[x][z]
  const PctDesc* dB = A.descT + (int64_t)y * n;              // mp_[z][y]
  const float sw = sum_w[u];
  const int i = i0 + g;
  const bool live = i < L1;
  for (int j = q; j < L2; j += G) acc[j] = 0.0f;
  __syncwarp();
  // pipeline registers: row bounds for z .. z+3, k / p for z .. z+2, bounds of row k for z, z+1
  int s0 = 0, e0 = 0, s1 = 0, e1 = 0, s2 = 0, e2 = 0, k1 = 0, b0 = 0, c0 = 0;
  float p0 = 0.0f, p1 = 0.0f;
  if (live) {
    PctDesc d = dA[0]; s0 = rp[d.rp + i]; e0 = rp[d.rp + i + 1];
    if (s0 + q < e0) { const uint2 t = ent[d.nz + s0 + q]; p0 = __uint_as_float(t.y); const PctDesc b = dB[0]; b0 = rp[b.rp + t.x]; c0 = rp[b.rp + t.x + 1]; }
    d = dA[min(1, n - 1)]; s1 = rp[d.rp + i]; e1 = rp[d.rp + i + 1];
    if (s1 + q < e1) { const uint2 t = ent[d.nz + s1 + q]; k1 = (int)t.x; p1 = __uint_as_float(t.y); }
    d = dA[min(2, n - 1)]; s2 = rp[d.rp + i]; e2 = rp[d.rp + i + 1];
  }
  for (int z = 0; z < n; ++z) {
    const float wz = wu[z];
    const PctDesc dbz = dB[z];
    const uint2* __restrict__ eb = ent + dbz.nz;             // entries of mp_[z][y]
    // loads for the steps ahead (past the end: repeated, never used)
    int s3 = 0, e3 = 0, k2 = 0, b1 = 0, c1 = 0;
    float p2 = 0.0f;
    if (live) {
      const PctDesc d3 = dA[min(z + 3, n - 1)];
      s3 = rp[d3.rp + i]; e3 = rp[d3.rp + i + 1];
      if (s2 + q < e2) { const uint2 t = ent[dA[min(z + 2, n - 1)].nz + s2 + q]; k2 = (int)t.x; p2 = __uint_as_float(t.y); }
      if (s1 + q < e1) { const int* rb = rp + dB[min(z + 1, n - 1)].rp; b1 = rb[k1]; c1 = rb[k1 + 1]; }
    }
    // step z: chunks of G k's
    const int cnt = e0 - s0;
    for (int ch = 0; __any_sync(FULL, ch < cnt); ch += G) {
      int kb = b0, kc = c0; float pp = p0;
      if (ch > 0) {                                          // more than G entries in row i: fetch now
        kb = 0; kc = 0; pp = 0.0f;
        if (s0 + ch + q < e0) {
          const uint2 t = ent[dA[z].nz + s0 + ch + q];
          pp = __uint_as_float(t.y);
          kb = rp[dbz.rp + t.x]; kc = rp[dbz.rp + t.x + 1];
        }
      }
      const int left = cnt - ch;
      for (int t = 0; __any_sync(FULL, t < min(left, G)); ++t) {
        const int src = (lane & ~(G - 1)) + t;
        const float pt = __shfl_sync(FULL, pp, src);
        const int bs = __shfl_sync(FULL, kb, src), be = __shfl_sync(FULL, kc, src);
        const bool on = t < left;
        for (int f = bs + q; __any_sync(FULL, on && f < be); f += G) {
          if (on && f < be) {
            const uint2 en = eb[f];
            acc[en.x] = __fadd_rn(acc[en.x], __fmul_rn(__fmul_rn(pt, __uint_as_float(en.y)), wz));
          }
        }
        __syncwarp();
      }
    }
    s0 = s1; e0 = e1; s1 = s2; e1 = e2; s2 = s3; e2 = e3;
    p0 = p1; k1 = k2; p1 = p2; b0 = b1; c0 = c1;
  }
  // read-out: v = acc / sum_w > CUTOFF, ascending j (:302-311); one pool reservation per warp
  const unsigned gmask = (G == 32 ? FULL : ((1u << G) - 1u)) << (g * G);
  int cntk = 0;
  for (int j0 = 0; j0 < L2; j0 += G) {
    const int j = j0 + q;
    const bool keep = live && j < L2 && (double)__fdiv_rn(acc[j], sw) > PCT_CUTOFF;
    cntk += __popc(__ballot_sync(FULL, keep) & gmask);
  }
  int off = 0, total = 0;
#pragma unroll
  for (int r = 0; r < RW; ++r) { const int c = __shfl_sync(FULL, cntk, r * G); if (r < g) off += c; total += c; }
  unsigned long long b = 0;
  if (lane == 0 && total) b = atomicAdd(O.pool_count, (unsigned long long)total);
  const int64_t base = (int64_t)__shfl_sync(FULL, b, 0) + off;
  if (live && q == 0) { const int64_t row = out_row_off[u] + i; O.row_pool[row] = base; O.row_cnt[row] = cntk; }
  const bool fits = base + cntk <= O.pool_cap;
  int run = 0;
  for (int j0 = 0; j0 < L2; j0 += G) {
    const int j = j0 + q;
    float v = 0.0f; bool keep = false;
    if (live && j < L2) { v = __fdiv_rn(acc[j], sw); keep = (double)v > PCT_CUTOFF; }
    const unsigned m = __ballot_sync(FULL, keep) & gmask;
    if (keep && fits) { const int64_t k = base + run + __popc(m & ((1u << lane) - 1u)); O.pool_col[k] = (uint32_t)j; O.pool_val[k] = v; }
    run += __popc(m);
  }
}

// ---- K8: relax_basepairing_probability ------------------------------------------------
// One warp per row i of one sequence x.  For y = 0..n-1 in order: the k's of row i of
// mp_[x][y] (ascending), for each the pairs (k, l) of bp_[y] (ascending l), for each the
// row l of mp_[y][x]; lanes take its columns j > i.  p_kl * p_ik * p_jl * w (:357).
__global__ void __launch_bounds__(128) pct_relax_bp_kernel(PctArgs A, const int64_t* __restrict__ out_row_off,
                                                           const float* __restrict__ w, const float* __restrict__ sum_w,
                                                           PctOut O, int accW) {
  extern __shared__ float pct_smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float* acc = pct_smem + (size_t)wib * accW;
  const int n = A.n;
  const int64_t total_rows = out_row_off[n];
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + wib;
  const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
  int x = 0;
  for (int64_t row = wid; row < total_rows; row += nw) {
    x = (int)pct_find_unit(out_row_off, n, row, x);
    const int i = (int)(row - out_row_off[x]);
    const int L1 = A.len[x];
    for (int j = lane; j < L1; j += 32) acc[j] = 0.0f;
    __syncwarp();
    for (int y = 0; y < n; ++y) {
      const float wy = w[(int64_t)x * n + y];
      const PctMat MA = pct_mat(A, x, y), MB = pct_mat(A, y, x);
      const int* brp = A.b_row_ptr + A.b_row_off[y];
      const uint32_t* bcol = A.b_col + A.b_nz_off[y];
      const float* bval = A.b_val + A.b_nz_off[y];
      int s = i, e = i + 1;
      if (MA.row_ptr) { s = MA.row_ptr[i]; e = MA.row_ptr[i + 1]; }
      for (int ke = s; ke < e; ++ke) {
        int k = i; float p_ik = 1.0f;
        if (MA.row_ptr) { k = (int)MA.col[ke]; p_ik = MA.val[ke]; }
        const int ls = brp[k], le = brp[k + 1];
        for (int a = ls; a < le; ++a) {
          const int l = (int)bcol[a];
          const float t1 = __fmul_rn(bval[a], p_ik);                        // p_kl * p_ik
          int js = l, je = l + 1;
          if (MB.row_ptr) { js = MB.row_ptr[l]; je = MB.row_ptr[l + 1]; }
          for (int c = js + lane; c < je; c += 32) {
            int j = l; float p_jl = 1.0f;
            if (MB.row_ptr) { j = (int)MB.col[c]; p_jl = MB.val[c]; }
            if (i < j) acc[j] = __fadd_rn(acc[j], __fmul_rn(__fmul_rn(t1, p_jl), wy));
          }
          __syncwarp();
        }
      }
    }
    __syncwarp();
    // rows i = 0 .. L1-2, columns j = i+1 .. L1-1 (:364-370); the last row stays empty
    pct_emit_row(O, row, acc, i + 1, (i < L1 - 1) ? L1 : i + 1, sum_w[x], lane);
    __syncwarp();
  }
}

// ---- pool -> CSR ------------------------------------------------------------------------
// entries per unit
__global__ void pct_unit_nnz_kernel(int64_t nunits, const int64_t* __restrict__ out_row_off,
                                    const int* __restrict__ row_cnt, int* __restrict__ unit_nnz) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t u = wid; u < nunits; u += nw) {
    int s = 0;
    for (int64_t r = out_row_off[u] + lane; r < out_row_off[u + 1]; r += 32) s += row_cnt[r];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) unit_nnz[u] = s;
  }
}

// exclusive scan of unit_nnz -> nz_off[nunits+1]; one block of 1024
__global__ void pct_scan_kernel(const int* __restrict__ nnz, int64_t n, int64_t* __restrict__ off) {
  __shared__ long long part[1024];
  const int t = threadIdx.x, T = blockDim.x;
  const int64_t per = (n + T - 1) / T;
  const int64_t b = min(n, (int64_t)t * per), e = min(n, b + per);
  long long s = 0;
  for (int64_t k = b; k < e; ++k) s += nnz[k];
  part[t] = s;
  __syncthreads();
  if (t == 0) { long long run = 0; for (int k = 0; k < T; ++k) { const long long v = part[k]; part[k] = run; run += v; } off[n] = run; }
  __syncthreads();
  long long run = part[t];
  for (int64_t k = b; k < e; ++k) { off[k] = run; run += nnz[k]; }
}

// one warp per unit: row_ptr block (rows+1 entries, relative to the unit) and the entries
__global__ void pct_gather_kernel(int64_t nunits, const int64_t* __restrict__ out_row_off, PctOut O,
                                  const int64_t* __restrict__ nz_off, int* __restrict__ row_ptr,
                                  uint32_t* __restrict__ col, float* __restrict__ val) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t u = wid; u < nunits; u += nw) {
    const int64_t r0 = out_row_off[u], rows = out_row_off[u + 1] - r0;
    int* rp = row_ptr + r0 + u;                              // the CSR block has rows + 1 entries per unit
    const int64_t dst0 = nz_off[u];
    int carry = 0;
    if (lane == 0) rp[0] = 0;
    for (int64_t q0 = 0; q0 < rows; q0 += 32) {
      const int64_t q = q0 + lane;
      const int c = q < rows ? O.row_cnt[r0 + q] : 0;
      const int incl = pct_warp_incl_scan(c, lane) + carry;
      if (q < rows) {
        rp[q + 1] = incl;
        const int64_t src = O.row_pool[r0 + q];
        if (src + c <= O.pool_cap)
          for (int k = 0; k < c; ++k) { col[dst0 + incl - c + k] = O.pool_col[src + k]; val[dst0 + incl - c + k] = O.pool_val[src + k]; }
      }
      carry = __shfl_sync(0xffffffffu, incl, 31);
    }
  }
}

}  // namespace dafs
