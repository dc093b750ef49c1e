// Device-side set-up of DD problems ("p_x, p_y, p_z never visit the host"):
// what build_problem() in dd.cu does on the host cores -- consensus-base-pair enumeration
// (src/dafs.cpp:1016-1061), the sorted-unique projections c_x / c_y / c_z, the Nussinov entry lists
// with their live-row / live-column maps, SparseNeedlemanWunsch::initialize
// (src/needleman_wunsch.cpp:198-253), the dense band of p_z and the CBP index quadruples -- from the
// dense matrices where they already are: in HBM, written by the averaging kernels of the progressive
// stage.  Seven small kernels; the row-parallel ones run one warp per (problem, row) so that even a
// batch of one problem spreads over the SMs (a single CTA per problem was bound by the latency of its
// own dependent loads: as slow as the host cores):
//   dd_pack_gates     clear the flags, gate lists of p_x / p_y / p_z                (warp per row)
//   dd_pack_cbp       CBP enumeration: count per x entry, projection flags          (warp per x row)
//   dd_pack_classify  entry counts per row / column, envelope row scans, c_z counts (warp per row)
//   dd_pack_scan      per problem: the reference's monotone envelope passes, band offsets, every
//                     prefix sum, and the line of totals the host turns into the block layout
//   dd_pack_head      per problem: header, index maps, row / column pointers        (after the layout)
//   dd_pack_entries   entry arrays by column, band of p_z, c_z cells                (warp per column / row)
//   dd_pack_rows      row_ent lists and the CBP quadruples                          (warp per row)
// The block is BYTE-IDENTICAL to the one the host packer builds (DAFS_DD_PACK_VERIFY=1 compares
// them), so everything downstream (kernel, tie-breaks, results) is unchanged.
#pragma once
#include <cstdint>

#include "dd_kernel.cuh"

namespace dafs {

struct PackCounts {            // totals of one problem, read back by the host
  int n_x, nd_x, nI_x, nJ_x, maxcol_x;
  int n_y, nd_y, nI_y, nJ_y, maxcol_y;
  int nb, bw, ncz, ncbp;
  int pad0, pad1;
};

struct PackProb {
  int L1, L2;
  unsigned N1, N2;
  const float* p_x;            // device, [L1*L1]
  const float* p_y;            // device, [L2*L2]
  const float* p_z;            // device, [L1*L2]
  float w, wx, wy, th_s, min_th_s, th_a;
  char* scratch;               // device, pack_scratch_bytes(L1, L2)
  char* block;                 // device, the constant block (dd_pack_fill)
  const DDHdr* hdr;            // device, the block's header as the host laid it out from the counts
};

// ---- scratch layout (per problem) ---------------------------------------------------
struct PackNusScratch {        // one Nussinov structure of side L
  int* gate_cnt;               // [L]    entries of row i with p > CUTOFF (j > i)
  int* gate_col;               // [L*L]  their columns, segment of row i at i*L
  uint8_t* proj;               // [L*L]  (i,j) is the projection of a CBP
  int* idx;                    // [L*L]  (i,j) -> entry id
  int* ccnt;                   // [L]    entries with span > 2 in column j
  int* ncnt;                   // [L]    entries with span <= 2 in column j
  int* rcnt;                   // [L]    entries with span > 2 in row i
  int* cbase;                  // [L+1]  exclusive prefix of ccnt
  int* nbase;                  // [L+1]  nd + exclusive prefix of ncnt
  int* rowmap;                 // [L+1]  live rows below i
  int* colmap;                 // [L+1]  live columns before j (exclusive prefix)
  int* rpre;                   // [L+1]  exclusive prefix of rcnt
};
struct PackScratch {
  PackNusScratch x, y;
  int* cbpcnt;                 // [L1*L1] CBPs of x-gate entry (i, t), same indexing as x.gate_col
  int* cbprow;                 // [L1+1]  exclusive prefix over rows of the CBP counts
  int* zgate_cnt;              // [L1]
  int* zgate_col;              // [L1*L2]
  uint8_t* czflag;             // [L1*L2] cell is in c_z
  int* zmap;                   // [L1*L2] cell -> c_z id
  int* czcnt;                  // [L1+1]  c_z cells per row, then their exclusive prefix
  uint2* env;                  // [L1+1]
  int* boff;                   // [L1+2]
};

__host__ __device__ inline size_t pack_up(size_t x) { return (x + 255) & ~size_t(255); }
__host__ __device__ inline size_t pack_nus_bytes(int L) {
  const size_t LL = (size_t)L * L;
  return pack_up(4 * (size_t)L) + pack_up(4 * LL) + pack_up(LL) + pack_up(4 * LL) + 3 * pack_up(4 * (size_t)L) +
         5 * pack_up(4 * (size_t)(L + 1));
}
__host__ __device__ inline size_t pack_scratch_bytes(int L1, int L2) {
  const size_t Z = (size_t)L1 * L2;
  return pack_nus_bytes(L1) + pack_nus_bytes(L2) + pack_up(4 * (size_t)L1 * L1) + pack_up(4 * (size_t)(L1 + 1)) +
         pack_up(4 * (size_t)L1) + pack_up(4 * Z) + pack_up(Z) + pack_up(4 * Z) + pack_up(4 * (size_t)(L1 + 1)) +
         pack_up(8 * (size_t)(L1 + 1)) + pack_up(4 * (size_t)(L1 + 2));
}
__host__ __device__ inline char* pack_carve_nus(char* p, int L, PackNusScratch& s) {
  const size_t LL = (size_t)L * L;
  s.gate_cnt = (int*)p; p += pack_up(4 * (size_t)L);
  s.gate_col = (int*)p; p += pack_up(4 * LL);
  s.proj = (uint8_t*)p; p += pack_up(LL);
  s.idx = (int*)p; p += pack_up(4 * LL);
  s.ccnt = (int*)p; p += pack_up(4 * (size_t)L);
  s.ncnt = (int*)p; p += pack_up(4 * (size_t)L);
  s.rcnt = (int*)p; p += pack_up(4 * (size_t)L);
  s.cbase = (int*)p; p += pack_up(4 * (size_t)(L + 1));
  s.nbase = (int*)p; p += pack_up(4 * (size_t)(L + 1));
  s.rowmap = (int*)p; p += pack_up(4 * (size_t)(L + 1));
  s.colmap = (int*)p; p += pack_up(4 * (size_t)(L + 1));
  s.rpre = (int*)p; p += pack_up(4 * (size_t)(L + 1));
  return p;
}
__host__ __device__ inline void pack_carve(char* p, int L1, int L2, PackScratch& s) {
  const size_t Z = (size_t)L1 * L2;
  p = pack_carve_nus(p, L1, s.x);
  p = pack_carve_nus(p, L2, s.y);
  s.cbpcnt = (int*)p; p += pack_up(4 * (size_t)L1 * L1);
  s.cbprow = (int*)p; p += pack_up(4 * (size_t)(L1 + 1));
  s.zgate_cnt = (int*)p; p += pack_up(4 * (size_t)L1);
  s.zgate_col = (int*)p; p += pack_up(4 * Z);
  s.czflag = (uint8_t*)p; p += pack_up(Z);
  s.zmap = (int*)p; p += pack_up(4 * Z);
  s.czcnt = (int*)p; p += pack_up(4 * (size_t)(L1 + 1));
  s.env = (uint2*)p; p += pack_up(8 * (size_t)(L1 + 1));
  s.boff = (int*)p; p += pack_up(4 * (size_t)(L1 + 2));
}

#define DD_PACK_THREADS 256
#define DD_PACK_CUTOFF 0.01      // CUTOFF, src/dafs.cpp:65 (a double literal: float > double compares)

// Exclusive prefix sum of in[0..n) into out[0..n], out[n] = total; in and out may alias.  Whole CTA.
__device__ inline int pack_block_scan(const int* in, int* out, int n, int* s_warp /* [9] */) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int carry = 0;
  for (int b = 0; b < n; b += DD_PACK_THREADS) {
    const int k = b + tid;
    const int v = k < n ? in[k] : 0;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    int wbase = 0, tot = 0;
#pragma unroll
    for (int q = 0; q < DD_PACK_THREADS / 32; ++q) { const int c = s_warp[q]; if (q < warp) wbase += c; tot += c; }
    if (k < n) out[k] = carry + wbase + x - v;
    carry += tot;
    __syncthreads();
  }
  if (tid == 0) out[n] = carry;
  __syncthreads();
  return carry;
}
__device__ inline int pack_block_sum(int v, int* s_warp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) s_warp[warp] = v;
  __syncthreads();
  int t = 0;
#pragma unroll
  for (int q = 0; q < DD_PACK_THREADS / 32; ++q) t += s_warp[q];
  return t;
}
__device__ inline int pack_block_max(int v, int* s_warp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
  __syncthreads();
  if (lane == 0) s_warp[warp] = v;
  __syncthreads();
  int t = 0;
#pragma unroll
  for (int q = 0; q < DD_PACK_THREADS / 32; ++q) t = max(t, s_warp[q]);
  return t;
}

// Is (i,j), i < j, an entry of the structure, and does it count as a decoder candidate (span > 2)?
// build_nus(), dd.cu: candidates are w*(p - th) > 0 with span > 2 (src/nussinov.cpp:234-237), plus
// every projection of a CBP (src/dafs.cpp:1037).
__device__ __forceinline__ bool pack_is_entry(float p, bool proj, int span, float w, float th) {
  const float s0 = __fmul_rn(w, __fsub_rn(p, th));
  return (span > 2 && s0 > 0.0f) || proj;
}

// The CBPs of x-gate entry (i,j) whose (i,k) is the b-th z-gate entry of row i, in l order
// (src/dafs.cpp:1022-1045: the reference's loops run i, j, k, l with exactly these tests).
template <class F>
__device__ __forceinline__ int pack_enum_b(const PackProb& P, const PackScratch& S, int i, int j, float pxv, int b, F emit) {
  const int L2 = P.L2;
  const int k = S.zgate_col[(size_t)i * L2 + b];
  if (k + 1 >= L2) return 0;
  const float fN1 = (float)P.N1, fN2 = (float)P.N2, fN = (float)(P.N1 + P.N2);
  const float pz1 = P.p_z[(size_t)i * L2 + k];
  const int ny = S.y.gate_cnt[k];
  int cnt = 0;
  for (int c = 0; c < ny; ++c) {
    const int l = S.y.gate_col[(size_t)k * L2 + c];
    const float pz2 = P.p_z[(size_t)j * L2 + l];
    if (!((double)pz2 > DD_PACK_CUTOFF)) continue;
    const float p = __fdiv_rn(__fadd_rn(__fmul_rn(fN1, pxv), __fmul_rn(fN2, P.p_y[(size_t)k * L2 + l])), fN);   // :1030
    const float qq = __fdiv_rn(__fadd_rn(pz1, pz2), 2.0f);                                                    // :1031
    const float dp = __fsub_rn(p, P.min_th_s);
    if (dp > 0.0f && __fadd_rn(__fmul_rn(P.w, dp), __fsub_rn(qq, P.th_a)) > 0.0f) { emit(k, l); ++cnt; }    // :1032
  }
  return cnt;
}

#define DD_PACK_ROWS (DD_PACK_THREADS / 32)     // rows (warps) per CTA of the row-parallel kernels

// ---- 1. clear + gates: warp per row r of every matrix ------------------------------------
__device__ __forceinline__ void pack_gate_row(const float* p, int i, int cols, bool upper, int* cnt, int* col, int lane) {
  int run = 0;
  for (int jb = upper ? i + 1 : 0; jb < cols; jb += 32) {
    const int j = jb + lane;
    const bool g = j < cols && (double)p[(size_t)i * cols + j] > DD_PACK_CUTOFF;
    const unsigned m = __ballot_sync(0xffffffffu, g);
    if (g) col[(size_t)i * cols + run + __popc(m & ((1u << lane) - 1u))] = j;
    run += __popc(m);
  }
  if (lane == 0) cnt[i] = run;
}

__global__ void __launch_bounds__(DD_PACK_THREADS) dd_pack_gates(const PackProb* probs) {
  const PackProb P = probs[blockIdx.y];
  const int lane = threadIdx.x & 31, r = blockIdx.x * DD_PACK_ROWS + (threadIdx.x >> 5);
  const int L1 = P.L1, L2 = P.L2;
  if (r >= max(L1, L2)) return;
  PackScratch S;
  pack_carve(P.scratch, L1, L2, S);
  if (r < L1) {
    for (int c = lane; c < L1; c += 32) { S.x.proj[(size_t)r * L1 + c] = 0; S.cbpcnt[(size_t)r * L1 + c] = 0; }
    for (int c = lane; c < L2; c += 32) S.czflag[(size_t)r * L2 + c] = 0;
    if (lane == 0) { S.x.ccnt[r] = 0; S.x.ncnt[r] = 0; S.cbprow[r] = 0; }
    pack_gate_row(P.p_x, r, L1, true, S.x.gate_cnt, S.x.gate_col, lane);
    pack_gate_row(P.p_z, r, L2, false, S.zgate_cnt, S.zgate_col, lane);
  }
  if (r < L2) {
    for (int c = lane; c < L2; c += 32) S.y.proj[(size_t)r * L2 + c] = 0;
    if (lane == 0) { S.y.ccnt[r] = 0; S.y.ncnt[r] = 0; }
    pack_gate_row(P.p_y, r, L2, true, S.y.gate_cnt, S.y.gate_col, lane);
  }
}

// ---- 2. CBP enumeration, counting: warp per row i of p_x; lanes over the z-gate entries of row i ----
__global__ void __launch_bounds__(DD_PACK_THREADS) dd_pack_cbp(const PackProb* probs) {
  const PackProb P = probs[blockIdx.y];
  const int lane = threadIdx.x & 31, i = blockIdx.x * DD_PACK_ROWS + (threadIdx.x >> 5);
  const int L1 = P.L1, L2 = P.L2;
  if (i + 1 >= L1) return;
  PackScratch S;
  pack_carve(P.scratch, L1, L2, S);
  const int nx = S.x.gate_cnt[i], nz = S.zgate_cnt[i];
  int row_total = 0;
  for (int t = 0; t < nx; ++t) {
    const int j = S.x.gate_col[(size_t)i * L1 + t];
    const float pxv = P.p_x[(size_t)i * L1 + j];
    int c = 0;
    for (int b = lane; b < nz; b += 32)
      c += pack_enum_b(P, S, i, j, pxv, b, [&](int k, int l) {
        S.y.proj[(size_t)k * L2 + l] = 1;
        S.czflag[(size_t)i * L2 + k] = 1;
        S.czflag[(size_t)j * L2 + l] = 1;
      });
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) { S.cbpcnt[(size_t)i * L1 + t] = c; if (c) S.x.proj[(size_t)i * L1 + j] = 1; }
    row_total += c;
  }
  if (lane == 0) S.cbprow[i] = row_total;
}

// ---- 3. entry counts, envelope row scans, c_z counts: warp per row ------------------------------
__device__ __forceinline__ void pack_classify_row(const float* p, int L, int i, float w, float th, const PackNusScratch& s, int lane) {
  int rc = 0;
  for (int jb = i + 1; jb < L; jb += 32) {
    const int j = jb + lane;
    bool cand = false;
    if (j < L) {
      const bool ent = pack_is_entry(p[(size_t)i * L + j], s.proj[(size_t)i * L + j] != 0, j - i, w, th);
      cand = ent && (j - i > 2);
      if (ent) atomicAdd(cand ? &s.ccnt[j] : &s.ncnt[j], 1);
    }
    rc += __popc(__ballot_sync(0xffffffffu, cand));
  }
  if (lane == 0) s.rcnt[i] = rc;
}

__global__ void __launch_bounds__(DD_PACK_THREADS) dd_pack_classify(const PackProb* probs) {
  const PackProb P = probs[blockIdx.y];
  const int lane = threadIdx.x & 31, r = blockIdx.x * DD_PACK_ROWS + (threadIdx.x >> 5);
  const int L1 = P.L1, L2 = P.L2;
  if (r >= max(L1, L2)) return;
  PackScratch S;
  pack_carve(P.scratch, L1, L2, S);
  if (r < L1) {
    pack_classify_row(P.p_x, L1, r, P.wx, P.th_s, S.x, lane);
    // envelope of row r + 1 (1-based): first / last column with p - th >= 0 (needleman_wunsch.cpp:201-222)
    int first = 0, last = 0, cz = 0;
    for (int kb = 0; kb < L2; kb += 32) {
      const int k = kb + lane;
      const bool g = k < L2 && __fsub_rn(P.p_z[(size_t)r * L2 + k], P.th_a) >= 0.0f;
      const unsigned m = __ballot_sync(0xffffffffu, g);
      if (m) { if (!first) first = kb + __ffs(m); last = kb + 32 - __clz(m); }
      cz += __popc(__ballot_sync(0xffffffffu, k < L2 && S.czflag[(size_t)r * L2 + k] != 0));
    }
    if (lane == 0) { S.env[r + 1] = make_uint2((unsigned)first, (unsigned)last); S.czcnt[r] = cz; }
  }
  if (r < L2) pack_classify_row(P.p_y, L2, r, P.wy, P.th_s, S.y, lane);
}

// ---- 4. per problem: envelope passes, band offsets, prefix sums, totals ------------------------
__device__ inline void pack_scan_nus(const PackNusScratch& s, int L, int* s_warp, int& n, int& nd, int& nI, int& nJ, int& mc) {
  const int tid = threadIdx.x;
  int a_mc = 0;
  for (int c = tid; c < L; c += DD_PACK_THREADS) a_mc = max(a_mc, s.ccnt[c]);
  mc = pack_block_max(a_mc, s_warp);
  nd = pack_block_scan(s.ccnt, s.cbase, L, s_warp);
  const int nn = pack_block_scan(s.ncnt, s.nbase, L, s_warp);
  n = nd + nn;
  for (int k = tid; k < L; k += DD_PACK_THREADS) { s.rowmap[k] = s.rcnt[k] > 0; s.colmap[k] = s.ccnt[k] > 0; }
  __syncthreads();
  nI = pack_block_scan(s.rowmap, s.rowmap, L, s_warp);
  nJ = pack_block_scan(s.colmap, s.colmap, L, s_warp);
  pack_block_scan(s.rcnt, s.rpre, L, s_warp);
}

__global__ void __launch_bounds__(DD_PACK_THREADS) dd_pack_scan(const PackProb* probs, PackCounts* counts) {
  extern __shared__ uint2 s_env[];             // [L1 + 1] when it fits (the monotone passes are serial)
  __shared__ int s_warp[DD_PACK_THREADS / 32 + 1];
  __shared__ int s_nb, s_bw;
  const PackProb P = probs[blockIdx.x];
  PackScratch S;
  pack_carve(P.scratch, P.L1, P.L2, S);
  const int L1 = P.L1, L2 = P.L2, tid = threadIdx.x;
  unsigned dyn = 0;
  asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn));
  const bool in_smem = (size_t)(L1 + 1) * sizeof(uint2) <= (size_t)dyn;
  uint2* env = in_smem ? s_env : S.env;
  if (in_smem) for (int k = 1 + tid; k <= L1; k += DD_PACK_THREADS) env[k] = S.env[k];
  __syncthreads();
  if (tid == 0) {
    // src/needleman_wunsch.cpp:198-253 on the per-row first / last columns found by dd_pack_classify
    env[0] = make_uint2(0, 0);
    for (int i = 1; i <= L1; ++i) {
      const uint2 fl = env[i];
      if (fl.x == 0) env[i] = env[i - 1];
      else { uint2 pv = env[i - 1]; pv.x = min(pv.x, fl.x - 1); pv.y = max(pv.y, fl.y - 1); env[i - 1] = pv; }
    }
    env[L1].y = (unsigned)L2;
    { unsigned v = (unsigned)L2; for (int i = L1; i != 0; --i) env[i].x = v = min(v, env[i].x); }
    { unsigned v = 0; for (int i = 0; i != L1 + 1; ++i) env[i].y = v = max(v, env[i].y); }
    for (int i = 1; i != L1 + 1; ++i) if (env[i - 1].y < env[i].x) env[i].x = env[i - 1].y;
    int nb = 0, bw = 0;
    S.boff[0] = 0;
    for (int r = 1; r <= L1; ++r) {
      S.boff[r] = nb;
      const int ef = max(1, (int)env[r].x), el = (int)env[r].y;
      if (el >= ef) { nb += el - ef + 1; bw = max(bw, el - ef + 1); }
    }
    S.boff[L1 + 1] = nb;
    s_nb = nb; s_bw = bw;
  }
  __syncthreads();
  if (in_smem) for (int k = tid; k <= L1; k += DD_PACK_THREADS) S.env[k] = env[k];
  PackCounts C;
  pack_scan_nus(S.x, L1, s_warp, C.n_x, C.nd_x, C.nI_x, C.nJ_x, C.maxcol_x);
  pack_scan_nus(S.y, L2, s_warp, C.n_y, C.nd_y, C.nI_y, C.nJ_y, C.maxcol_y);
  C.ncz = pack_block_scan(S.czcnt, S.czcnt, L1, s_warp);
  C.ncbp = pack_block_scan(S.cbprow, S.cbprow, L1, s_warp);
  if (tid == 0) { C.nb = s_nb; C.bw = s_bw; C.pad0 = C.pad1 = 0; counts[blockIdx.x] = C; }
}

// ---- 5. per problem: header, maps, pointers (the block's layout is known now) ---------------------
__device__ inline void pack_head_nus(const PackNusScratch& s, int L, char* blk, const DDNusHdr& h) {
  const int tid = threadIdx.x;
  int* col_ptr = reinterpret_cast<int*>(blk + h.col_ptr);
  int* row_ptr = reinterpret_cast<int*>(blk + h.row_ptr);
  uint16_t* rowI = reinterpret_cast<uint16_t*>(blk + h.rowI);
  uint16_t* colJ = reinterpret_cast<uint16_t*>(blk + h.colJ);
  uint16_t* bstart = reinterpret_cast<uint16_t*>(blk + h.bstart);
  uint16_t* rowmap = reinterpret_cast<uint16_t*>(blk + h.rowmap);
  uint16_t* colmap = reinterpret_cast<uint16_t*>(blk + h.colmap);
  for (int k = tid; k <= L; k += DD_PACK_THREADS) {
    rowmap[k] = (uint16_t)s.rowmap[k];                                              // live rows below k
    colmap[k] = (uint16_t)(k < L ? s.colmap[k] + (s.ccnt[k] > 0) : s.colmap[L]);    // live columns up to and including k
  }
  for (int k = tid; k < L; k += DD_PACK_THREADS) {
    if (s.rcnt[k] > 0) { const int a = s.rowmap[k]; rowI[a] = (uint16_t)k; row_ptr[a] = s.rpre[k]; bstart[a] = (uint16_t)(s.colmap[k] + (s.ccnt[k] > 0)); }
    if (s.ccnt[k] > 0) { const int b = s.colmap[k]; colJ[b] = (uint16_t)k; col_ptr[b] = s.cbase[k]; }
  }
  if (tid == 0) { col_ptr[s.colmap[L]] = s.cbase[L]; row_ptr[s.rowmap[L]] = s.cbase[L]; }
}

__global__ void __launch_bounds__(DD_PACK_THREADS) dd_pack_head(const PackProb* probs) {
  const PackProb P = probs[blockIdx.x];
  PackScratch S;
  pack_carve(P.scratch, P.L1, P.L2, S);
  const int L1 = P.L1, tid = threadIdx.x;
  char* blk = P.block;
  {
    const int* src = reinterpret_cast<const int*>(P.hdr);
    int* dst = reinterpret_cast<int*>(blk);
    for (int k = tid; k < (int)(sizeof(DDHdr) / 4); k += DD_PACK_THREADS) dst[k] = src[k];
  }
  const DDHdr H = *P.hdr;
  pack_head_nus(S.x, L1, blk, H.x);
  pack_head_nus(S.y, P.L2, blk, H.y);
  uint2* env = reinterpret_cast<uint2*>(blk + H.z_env);
  int* boff = reinterpret_cast<int*>(blk + H.z_boff);
  int* cz_row_ptr = reinterpret_cast<int*>(blk + H.cz_row_ptr);
  for (int k = tid; k <= L1; k += DD_PACK_THREADS) { env[k] = S.env[k]; cz_row_ptr[k] = S.czcnt[k]; }
  for (int k = tid; k <= L1 + 1; k += DD_PACK_THREADS) boff[k] = S.boff[k];
}

// ---- 6. entries by column, band of p_z, c_z cells: warp per column / row ---------------------------
// Entries with span > 2 first, in column order (ascending j, descending i: the order of the reference's
// bp[j] lists), then the others likewise (pack_nus() of dd.cu).
__device__ __forceinline__ void pack_entries_col(const float* p, int L, int j, float w, float th, const PackNusScratch& s,
                                                 char* blk, const DDNusHdr& h, int lane) {
  uint32_t* ent_ab = reinterpret_cast<uint32_t*>(blk + h.ent_ab);
  int* ent_c = reinterpret_cast<int*>(blk + h.ent_c);
  uint32_t* ent_ij = reinterpret_cast<uint32_t*>(blk + h.ent_ij);
  float* ent_p = reinterpret_cast<float*>(blk + h.ent_p);
  uint8_t* ent_flag = reinterpret_cast<uint8_t*>(blk + h.ent_flag);
  int run_c = s.cbase[j], run_n = s.cbase[L] + s.nbase[j];
  if (s.ccnt[j] + s.ncnt[j] == 0) return;
  const int bcol = s.colmap[j];                       // live column index of j (when it is live)
  const unsigned lt = (1u << lane) - 1u;
  for (int ib = j - 1; ib >= 0; ib -= 32) {
    const int i = ib - lane;
    bool ent = false, cand = false, pr = false;
    float pv = 0.0f;
    if (i >= 0) {
      pv = p[(size_t)i * L + j];
      pr = s.proj[(size_t)i * L + j] != 0;
      ent = pack_is_entry(pv, pr, j - i, w, th);
      cand = ent && (j - i > 2);
    }
    const unsigned mc = __ballot_sync(0xffffffffu, cand), mn = __ballot_sync(0xffffffffu, ent && !cand);
    if (ent) {
      const int e = cand ? run_c + __popc(mc & lt) : run_n + __popc(mn & lt);
      ent_ij[e] = (uint32_t)i | ((uint32_t)j << 16);
      ent_p[e] = pv;
      ent_flag[e] = pr ? 1 : 0;
      s.idx[(size_t)i * L + j] = e;
      if (cand) {
        const int a = s.rowmap[i];
        ent_ab[e] = (uint32_t)a | ((uint32_t)bcol << 16);
        // 1 + (last live column <= i - 1) = live columns up to and including i - 1
        ent_c[e] = (a << 16) | (i >= 1 ? s.colmap[i - 1] + (s.ccnt[i - 1] > 0) : 0);
      }
    }
    run_c += __popc(mc); run_n += __popc(mn);
  }
}

__global__ void __launch_bounds__(DD_PACK_THREADS) dd_pack_entries(const PackProb* probs) {
  const PackProb P = probs[blockIdx.y];
  const int lane = threadIdx.x & 31, r = blockIdx.x * DD_PACK_ROWS + (threadIdx.x >> 5);
  const int L1 = P.L1, L2 = P.L2;
  if (r >= max(L1, L2)) return;
  PackScratch S;
  pack_carve(P.scratch, L1, L2, S);
  char* blk = P.block;
  const DDHdr H = *P.hdr;
  if (r < L2) pack_entries_col(P.p_y, L2, r, P.wy, P.th_s, S.y, blk, H.y, lane);
  if (r < L1) {
    pack_entries_col(P.p_x, L1, r, P.wx, P.th_s, S.x, blk, H.x, lane);
    // band of row r + 1, c_z cells of row r
    float* pb = reinterpret_cast<float*>(blk + H.z_pb);
    uint16_t* cz_col = reinterpret_cast<uint16_t*>(blk + H.cz_col);
    int* cz_bidx = reinterpret_cast<int*>(blk + H.cz_bidx);
    const uint2 e = S.env[r + 1];
    const int ef = max(1, (int)e.x), el = (int)e.y, b0 = S.boff[r + 1];
    for (int k = ef + lane; k <= el; k += 32) pb[b0 + k - ef] = P.p_z[(size_t)r * L2 + (k - 1)];
    int run = S.czcnt[r];
    for (int kb = 0; kb < L2; kb += 32) {
      const int k = kb + lane;
      const bool f = k < L2 && S.czflag[(size_t)r * L2 + k] != 0;
      const unsigned m = __ballot_sync(0xffffffffu, f);
      if (f) {
        const int c = run + __popc(m & ((1u << lane) - 1u));
        cz_col[c] = (uint16_t)k;
        cz_bidx[c] = (k + 1 >= ef && k + 1 <= el) ? b0 + (k + 1 - ef) : -1;
        S.zmap[(size_t)r * L2 + k] = c;
      }
      run += __popc(m);
    }
  }
}

// ---- 7. row_ent lists and CBP quadruples: warp per row ---------------------------------------------
__device__ __forceinline__ void pack_row_ent(const float* p, int L, int i, float w, float th, const PackNusScratch& s,
                                             char* blk, const DDNusHdr& h, int lane) {
  if (s.rcnt[i] == 0) return;
  int* row_ent = reinterpret_cast<int*>(blk + h.row_ent);
  int run = s.rpre[i];
  for (int jb = i + 3; jb < L; jb += 32) {
    const int j = jb + lane;
    const bool cand = j < L && pack_is_entry(p[(size_t)i * L + j], s.proj[(size_t)i * L + j] != 0, j - i, w, th);
    const unsigned m = __ballot_sync(0xffffffffu, cand);
    if (cand) row_ent[run + __popc(m & ((1u << lane) - 1u))] = s.idx[(size_t)i * L + j];
    run += __popc(m);
  }
}

__global__ void __launch_bounds__(DD_PACK_THREADS) dd_pack_rows(const PackProb* probs) {
  const PackProb P = probs[blockIdx.y];
  const int lane = threadIdx.x & 31, r = blockIdx.x * DD_PACK_ROWS + (threadIdx.x >> 5);
  const int L1 = P.L1, L2 = P.L2;
  if (r >= max(L1, L2)) return;
  PackScratch S;
  pack_carve(P.scratch, L1, L2, S);
  char* blk = P.block;
  const DDHdr H = *P.hdr;
  if (r < L2) pack_row_ent(P.p_y, L2, r, P.wy, P.th_s, S.y, blk, H.y, lane);
  if (r >= L1) return;
  pack_row_ent(P.p_x, L1, r, P.wx, P.th_s, S.x, blk, H.x, lane);
  if (r + 1 >= L1) return;
  // CBP quadruples of row i = r in the reference's (i, j, k, l) order
  int4* cb = reinterpret_cast<int4*>(blk + H.cbp);
  const int i = r, nx = S.x.gate_cnt[i], nz = S.zgate_cnt[i];
  int base = S.cbprow[i];
  for (int t = 0; t < nx; ++t) {
    const int total = S.cbpcnt[(size_t)i * L1 + t];
    if (total == 0) continue;
    const int j = S.x.gate_col[(size_t)i * L1 + t];
    const float pxv = P.p_x[(size_t)i * L1 + j];
    const int ex = S.x.idx[(size_t)i * L1 + j];
    int run = base;
    for (int bb = 0; bb < nz; bb += 32) {
      const int b = bb + lane;
      const int c = b < nz ? pack_enum_b(P, S, i, j, pxv, b, [](int, int) {}) : 0;
      int x = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
      if (c) {
        int out = run + x - c;
        pack_enum_b(P, S, i, j, pxv, b, [&](int k, int l) {
          cb[out++] = make_int4(ex, S.y.idx[(size_t)k * L2 + l], S.zmap[(size_t)i * L2 + k], S.zmap[(size_t)j * L2 + l]);
        });
      }
      run += __shfl_sync(0xffffffffu, x, 31);
    }
    base += total;
  }
}

}  // namespace dafs
