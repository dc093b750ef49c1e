// K10-K12: the per-node inputs of the progressive stage (DAFS::align_alignments,
// src/dafs.cpp:913-951) on device-resident posteriors, for a whole WAVE of guide-tree
// nodes at once (all nodes whose children are finished):
//   K10 transpose_mp                        src/dafs.cpp:155-167 (run(), :1797-1799)
//   K11 average_basepairing_probability     src/dafs.cpp:557-607 (use_alifold == false)
//   K12 average_matching_probability        src/dafs.cpp:513-555
// Bit-exact with the reference's x86-64 build.  A cell of an averaged matrix receives
// at most one contribution per member sequence (K11) or member pair (K12), added in
// the reference's member order; one warp owns one output row, keeps it in shared
// memory, and walks the members in that order.  Inside one member (pair) the entries of
// a sparse row map to distinct cells, so the lanes take them side by side.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace dafs {

#define PROG_CUTOFF 0.01   // src/dafs.cpp:65, a double constant

struct ProgDesc { int64_t rp, nz; };   // where mp_[a][b] lives in rp_all / col_all / val_all

// One side of a node = a sub-alignment: `rows` member sequences over L columns.
// Member m (global index mem0 + m): sequence mem_seq[], residue -> column map at
// maps + mem_col[], column -> residue (or -1) map at maps + mem_inv[].
struct ProgSide { int L, rows, mem0, pad; int64_t out; };   // out: offset of the dense L x L matrix (floats)
struct ProgNode { int side1, side2, pad0, pad1; int64_t out; };   // out: offset of the dense L1 x L2 matrix

struct ProgArgs {
  int n;                                  // sequences
  const int* len;
  // mp: U (pairs a < b, the caller's) then T (their transposes), one set of arrays
  const int* rp_all; const uint32_t* col_all; const float* val_all;
  const ProgDesc* desc;                   // [n*n]
  // bp: one CSR per sequence
  const int64_t* b_row_off; const int64_t* b_nz_off;
  const int* b_row_ptr; const uint32_t* b_col; const float* b_val;
  // the wave
  const ProgSide* sides; const ProgNode* nodes;
  const int* mem_seq; const int64_t* mem_col; const int64_t* mem_inv;
  const int* maps;
  float* dense;                           // all p_x / p_y / p_z of the wave
};

__device__ __forceinline__ int prog_warp_incl_scan(int v, int lane) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += y; }
  return v;
}

// ---- input check: the CSR arrays index device memory --------------------------------
__global__ void prog_validate_kernel(int64_t nunits, const int64_t* __restrict__ row_off, const int64_t* __restrict__ nz_off,
                                     const int* __restrict__ row_ptr, const uint32_t* __restrict__ col,
                                     const int* __restrict__ unit_cols, int* __restrict__ bad) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t u = wid; u < nunits; u += nw) {
    const int rows = (int)(row_off[u + 1] - row_off[u]) - 1;
    const int64_t nnz = nz_off[u + 1] - nz_off[u];
    const int* rp = row_ptr + row_off[u];
    const uint32_t* c = col + nz_off[u];
    const int ncols = unit_cols[u];
    bool ok = rows >= 0 && nnz >= 0 && rp[0] == 0 && (int64_t)rp[rows < 0 ? 0 : rows] == nnz;
    if (ok) {
      for (int r = lane; r < rows; r += 32) {
        const int b = rp[r], e = rp[r + 1];
        if (b < 0 || e < b || (int64_t)e > nnz) { ok = false; break; }
        for (int k = b; k < e; ++k)
          if (c[k] >= (uint32_t)ncols || (k > b && c[k] <= c[k - 1])) ok = false;
      }
    }
    if (!ok) atomicExch(bad, 1);
  }
}

// ---- K10: transposes of all pairs ------------------------------------------------------
// Warp per pair.  Column histogram -> exclusive scan = row_ptr of the transpose; rows of U are
// then walked in ascending order, so inside a row of T the columns come out ascending, which
// is what the reference's std::sort over (row, value) pairs with distinct rows yields.
__global__ void __launch_bounds__(128) prog_transpose_kernel(int n, const int* __restrict__ len, int64_t npairs,
                                                             const int2* __restrict__ pairs,
                                                             const int64_t* __restrict__ u_row_off,
                                                             const int64_t* __restrict__ t_row_off,
                                                             const int64_t* __restrict__ nz_off,
                                                             const int* __restrict__ u_rp, const uint32_t* __restrict__ u_col,
                                                             const float* __restrict__ u_val, int* __restrict__ t_rp,
                                                             uint32_t* __restrict__ t_col, float* __restrict__ t_val,
                                                             int* __restrict__ fill) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t u = wid; u < npairs; u += nw) {
    const int La = len[pairs[u].x], Lb = len[pairs[u].y];
    const int* rp = u_rp + u_row_off[u];
    const uint32_t* col = u_col + nz_off[u];
    const float* val = u_val + nz_off[u];
    int* trp = t_rp + t_row_off[u];
    int* cur = fill + t_row_off[u];
    uint32_t* tc = t_col + nz_off[u];
    float* tv = t_val + nz_off[u];
    const int nnz = rp[La];
    for (int j = lane; j <= Lb; j += 32) cur[j] = 0;
    __syncwarp();
    for (int k = lane; k < nnz; k += 32) atomicAdd(&cur[col[k]], 1);
    __syncwarp();
    int carry = 0;
    for (int j0 = 0; j0 <= Lb; j0 += 32) {
      const int j = j0 + lane;
      const int c = j < Lb ? __ldcg(cur + j) : 0;
      const int incl = prog_warp_incl_scan(c, lane) + carry;
      if (j <= Lb) { trp[j] = incl - c; __stcg(cur + j, incl - c); }
      carry = __shfl_sync(0xffffffffu, incl, 31);
    }
    __syncwarp();
    for (int i = 0; i < La; ++i) {
      const int b = rp[i], e = rp[i + 1];
      for (int k = b + lane; k < e; k += 32) {      // distinct columns inside a row: no two lanes share a cursor
        const uint32_t j = col[k];
        const int pos = __ldcg(cur + j);
        tc[pos] = (uint32_t)i; tv[pos] = val[k];
        __stcg(cur + j, pos + 1);
      }
      __syncwarp();
    }
  }
}

__global__ void prog_desc_kernel(int n, const int64_t* __restrict__ u_row_off, const int64_t* __restrict__ t_row_off,
                                 const int64_t* __restrict__ nz_off, int64_t ru, int64_t nnz, ProgDesc* __restrict__ desc) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * n) return;
  const int a = t / n, b = t % n;
  ProgDesc d{0, 0};
  if (a < b) { const int64_t u = (int64_t)a * n - (int64_t)a * (a + 1) / 2 + (b - a - 1); d.rp = u_row_off[u]; d.nz = nz_off[u]; }
  else if (a > b) { const int64_t u = (int64_t)b * n - (int64_t)b * (b + 1) / 2 + (a - b - 1); d.rp = ru + t_row_off[u]; d.nz = nnz + nz_off[u]; }
  desc[t] = d;
}

// ---- K11: average_basepairing_probability -----------------------------------------------
// task = (side, alignment column a).  p[idx[i]][idx[j]] += p_ij / N for the members in
// alignment order (:566-578); then p <= CUTOFF -> 0 on the upper triangle (:598-604).
__global__ void __launch_bounds__(256) prog_avg_bp_kernel(ProgArgs A, int64_t ntasks, const int2* __restrict__ tasks, int accW) {
  extern __shared__ float prog_smem[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  float* acc = prog_smem + (size_t)w * accW;
  for (int64_t t = (int64_t)blockIdx.x * wpb + w; t < ntasks; t += (int64_t)gridDim.x * wpb) {
    const int2 tk = tasks[t];
    const ProgSide S = A.sides[tk.x];
    const int a = tk.y, L = S.L;
    for (int j = lane; j < L; j += 32) acc[j] = 0.0f;
    __syncwarp();
    const float fN = (float)(unsigned)S.rows;
    for (int m = 0; m < S.rows; ++m) {
      const int ii = A.maps[A.mem_inv[S.mem0 + m] + a];
      if (ii < 0) continue;
      const int s = A.mem_seq[S.mem0 + m];
      const int* cmap = A.maps + A.mem_col[S.mem0 + m];
      const int* rp = A.b_row_ptr + A.b_row_off[s];
      const int b = rp[ii], e = rp[ii + 1];
      const uint32_t* col = A.b_col + A.b_nz_off[s];
      const float* val = A.b_val + A.b_nz_off[s];
      for (int k = b + lane; k < e; k += 32) {
        const int j = cmap[col[k]];
        acc[j] = __fadd_rn(acc[j], __fdiv_rn(val[k], fN));
      }
      __syncwarp();
    }
    float* out = A.dense + S.out + (int64_t)a * L;
    for (int j = lane; j < L; j += 32) {
      float v = acc[j];
      if (j <= a || (double)v <= PROG_CUTOFF) v = 0.0f;
      out[j] = v;
    }
    __syncwarp();
  }
}

// ---- K12: average_matching_probability ----------------------------------------------------
// task = (node, alignment column i of aln1).  For the members r1 of aln1 that have a residue
// in column i, and all members r2 of aln2, in that order: the entries (jj, p) of row ii of
// mp_[s1][s2] add p / (N1 * N2) to column colmap2[r2][jj] (:520-544).  The row bounds (and the
// first entry) of 32 member pairs are fetched at once, one per lane, and handed round with
// shuffles, so the serial walk over the pairs does not wait for them.  Then p <= CUTOFF -> 0,
// p > 1 -> 1 (:546-553).
__global__ void __launch_bounds__(256) prog_avg_mp_kernel(ProgArgs A, int64_t ntasks, const int2* __restrict__ tasks, int accW) {
  extern __shared__ float prog_smem[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  float* acc = prog_smem + (size_t)w * accW;
  for (int64_t t = (int64_t)blockIdx.x * wpb + w; t < ntasks; t += (int64_t)gridDim.x * wpb) {
    const int2 tk = tasks[t];
    const ProgNode nd = A.nodes[tk.x];
    const ProgSide S1 = A.sides[nd.side1], S2 = A.sides[nd.side2];
    const int i = tk.y, L2 = S2.L;
    for (int j = lane; j < L2; j += 32) acc[j] = 0.0f;
    __syncwarp();
    const float fNN = (float)((unsigned)S1.rows * (unsigned)S2.rows);
    for (int m1 = 0; m1 < S1.rows; ++m1) {
      const int ii = A.maps[A.mem_inv[S1.mem0 + m1] + i];
      if (ii < 0) continue;
      const int s1 = A.mem_seq[S1.mem0 + m1];
      for (int m20 = 0; m20 < S2.rows; m20 += 32) {
        const int m2 = m20 + lane;
        int64_t base = 0; int cnt = 0; const int* cmap = A.maps;
        int j0 = 0; float v0 = 0.0f;
        if (m2 < S2.rows) {
          const int s2 = A.mem_seq[S2.mem0 + m2];
          const ProgDesc d = A.desc[(int64_t)s1 * A.n + s2];
          const int* rp = A.rp_all + d.rp;
          const int b = rp[ii];
          cnt = rp[ii + 1] - b;
          base = d.nz + b;
          cmap = A.maps + A.mem_col[S2.mem0 + m2];
          if (cnt > 0) { j0 = cmap[A.col_all[base]]; v0 = A.val_all[base]; }
        }
        const int lim = min(32, S2.rows - m20);
        for (int q = 0; q < lim; ++q) {
          const int c = __shfl_sync(0xffffffffu, cnt, q);
          if (c == 0) continue;
          const int jq = __shfl_sync(0xffffffffu, j0, q);
          const float vq = __shfl_sync(0xffffffffu, v0, q);
          if (c == 1) {
            if (lane == 0) acc[jq] = __fadd_rn(acc[jq], __fdiv_rn(vq, fNN));
          } else {
            const int64_t bq = __shfl_sync(0xffffffffu, base, q);
            const int* cm = reinterpret_cast<const int*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(cmap), q));
            if (lane == 0) acc[jq] = __fadd_rn(acc[jq], __fdiv_rn(vq, fNN));
            for (int k = 1 + lane; k < c; k += 32) {
              const int j = cm[A.col_all[bq + k]];
              acc[j] = __fadd_rn(acc[j], __fdiv_rn(A.val_all[bq + k], fNN));
            }
          }
          __syncwarp();
        }
      }
    }
    float* out = A.dense + nd.out + (int64_t)i * L2;
    for (int j = lane; j < L2; j += 32) {
      float v = acc[j];
      if ((double)v <= PROG_CUTOFF) v = 0.0f;
      if (v > 1.0f) v = 1.0f;
      out[j] = v;
    }
    __syncwarp();
  }
}

}  // namespace dafs
