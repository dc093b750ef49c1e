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

#define PCT_CUTOFF 0.01   // src/dafs.cpp:47, a double constant

struct PctMat {             // one mp_[a][b]: rows of a, columns of b
  const int* row_ptr;       // nullptr: identity
  const uint32_t* col;
  const float* val;
};

struct PctArgs {
  int n;                    // sequences
  const int* len;           // [n]
  // mp: U = pairs a < b in row-major order, T = their transposes
  const int64_t* u_row_off; // [npairs+1] into u_row_ptr (len[a]+1 per pair)
  const int64_t* t_row_off; // [npairs+1] into t_row_ptr (len[b]+1 per pair)
  const int64_t* nz_off;    // [npairs+1] entries of pair u in both U and T
  const int* u_row_ptr; const uint32_t* u_col; const float* u_val;
  int* t_row_ptr; uint32_t* t_col; float* t_val;
  // bp: one CSR per sequence
  const int64_t* b_row_off; // [n+1] (len+1 per sequence)
  const int64_t* b_nz_off;  // [n+1]
  const int* b_row_ptr; const uint32_t* b_col; const float* b_val;
  float* sim;               // [n*n]
};

__device__ __forceinline__ int64_t pct_pair_index(int n, int a, int b) {   // a < b
  return (int64_t)a * n - (int64_t)a * (a + 1) / 2 + (b - a - 1);
}

__device__ __forceinline__ PctMat pct_mat(const PctArgs& A, int a, int b) {
  PctMat m;
  if (a == b) { m.row_ptr = nullptr; m.col = nullptr; m.val = nullptr; return m; }
  if (a < b) {
    const int64_t u = pct_pair_index(A.n, a, b);
    m.row_ptr = A.u_row_ptr + A.u_row_off[u]; m.col = A.u_col + A.nz_off[u]; m.val = A.u_val + A.nz_off[u];
  } else {
    const int64_t u = pct_pair_index(A.n, b, a);
    m.row_ptr = A.t_row_ptr + A.t_row_off[u]; m.col = A.t_col + A.nz_off[u]; m.val = A.t_val + A.nz_off[u];
  }
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
        __stcg(cur + j, pos + 1);
      }
      __syncwarp();
    }
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
// One warp per row i of one pair (x < y).  For z = 0..n-1 in order: the k's of row i
// of mp_[x][z] in ascending order, for each the row k of mp_[z][y]; lanes take the
// flattened (k, j) products, which are loaded in two rounds (all k's of the row at
// once, then all their j's) and added group by group in k order -- within a group the
// columns j are distinct.  Accumulator: one float per column of y in shared memory.
__global__ void __launch_bounds__(128) pct_relax_mp_kernel(PctArgs A, int64_t npairs, const int2* __restrict__ pairs,
                                                           const int64_t* __restrict__ out_row_off,
                                                           const float* __restrict__ w, const float* __restrict__ sum_w,
                                                           PctOut O, int accW) {
  extern __shared__ float pct_smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float* acc = pct_smem + (size_t)wib * accW;
  const int64_t total_rows = out_row_off[npairs];
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + wib;
  const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
  int64_t u = 0;
  for (int64_t row = wid; row < total_rows; row += nw) {
    // unit of this row: rows of pair u are [out_row_off[u], out_row_off[u+1]) (len[x] rows each, no +1)
    u = pct_find_unit(out_row_off, npairs, row, u);
    const int i = (int)(row - out_row_off[u]);
    const int x = pairs[u].x, y = pairs[u].y;
    const int L2 = A.len[y];
    for (int j = lane; j < L2; j += 32) acc[j] = 0.0f;
    __syncwarp();
    const float* wu = w + u * A.n;
    for (int z = 0; z < A.n; ++z) {
      const float wz = wu[z];
      const PctMat MA = pct_mat(A, x, z), MB = pct_mat(A, z, y);
      int s = i, e = i + 1;                                 // identity: the single entry (i, 1.0f)
      if (MA.row_ptr) { s = MA.row_ptr[i]; e = MA.row_ptr[i + 1]; }
      for (int k0 = s; k0 < e; k0 += 32) {
        // round 1: up to 32 k's of the row, and where their rows of mp_[z][y] live
        const int ke = k0 + lane;
        int ak = 0, bs = 0, cnt = 0; float ap = 0.0f;
        if (ke < e) {
          if (MA.row_ptr) { ak = (int)MA.col[ke]; ap = MA.val[ke]; } else { ak = i; ap = 1.0f; }
          if (MB.row_ptr) { bs = MB.row_ptr[ak]; cnt = MB.row_ptr[ak + 1] - bs; } else { bs = ak; cnt = 1; }
        }
        const int incl = pct_warp_incl_scan(cnt, lane);
        const int off = incl - cnt;
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        const int nk = min(32, e - k0);
        for (int f0 = 0; f0 < total; f0 += 32) {
          // round 2: 32 flattened products; g = the k group of mine
          const int f = f0 + lane;
          int g = 0;
          for (int t = 1; t < nk; ++t) { const int o = __shfl_sync(0xffffffffu, off, t); if (f >= o) g = t; }
          const int go = __shfl_sync(0xffffffffu, off, g), gb = __shfl_sync(0xffffffffu, bs, g);
          const float gp = __shfl_sync(0xffffffffu, ap, g);
          int j = 0; float t2 = 0.0f;
          const bool live = f < total;
          if (live) {
            float pj;
            if (MB.row_ptr) { j = (int)MB.col[gb + (f - go)]; pj = MB.val[gb + (f - go)]; } else { j = gb; pj = 1.0f; }
            t2 = __fmul_rn(__fmul_rn(gp, pj), wz);                          // p_ik * p_jk * w  (:297)
          }
          // add group by group (ascending k); columns inside a group are distinct
          const int gfirst = __shfl_sync(0xffffffffu, g, 0);
          const int glast = __shfl_sync(0xffffffffu, live ? g : -1, min(31, total - f0 - 1));
          for (int cur = gfirst; cur <= glast; ++cur) {
            if (live && g == cur) acc[j] = __fadd_rn(acc[j], t2);
            __syncwarp();
          }
        }
      }
    }
    __syncwarp();
    pct_emit_row(O, row, acc, 0, L2, sum_w[u], lane);
    __syncwarp();
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
