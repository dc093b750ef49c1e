// Per-pair driver: forward stripes, backward stripes, total probability,
// posterior + threshold + CSR emission; and the persistent kernel around it.
#pragma once
#include "pairhmm_backward.cuh"

namespace dafs {

__device__ __forceinline__ int col_class(const uint8_t* __restrict__ s2, int L2, int j) {
  // column 0 carries no symbol (finite class so its cells equal the reference's
  // LOG_ZERO exactly); columns past L2 are padding.
  return j == 0 ? (int)kClsOther : (j <= L2 ? (int)s2[j - 1] : (int)kClsPad);
}

// One warp, one pair.  Returns nnz (same value in every lane).
template <int C>
__device__ __forceinline__ void pairhmm_pair(const PairHmmArgs& a, const PairHmmSmem& sm,
                                             int pair, int slot, int lane) {
  const int2 pr = a.pairs[pair];
  const uint8_t* s1 = a.seq + a.seq_off[pr.x];
  const uint8_t* s2 = a.seq + a.seq_off[pr.y];
  const int L1 = a.seq_len[pr.x], L2 = a.seq_len[pr.y];
  const int SW = 32 * C;
  const int nstripes = (L2 + 1 + SW - 1) / SW;
  const int PW = nstripes * SW;
  float* planeF = a.planes + (int64_t)slot * 2 * a.plane_stride;
  float* planeB = planeF + a.plane_stride;
  float4* bnd0 = a.bnd + (int64_t)slot * 2 * a.bnd_stride;
  float4* bnd1 = bnd0 + a.bnd_stride;

  // ---- forward ------------------------------------------------------------
  float fend[3] = {DAFS_LOG_ZERO, DAFS_LOG_ZERO, DAFS_LOG_ZERO};
  for (int s = 0; s < nstripes; ++s) {
    StripeCols<C> sc;
    const int j0 = s * SW + lane * C;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int cls = col_class(s2, L2, j0 + c);
      sc.mofs[c] = cls * 32 + lane;
      sc.ins[c] = sm.ins[cls * 32 + lane];
    }
    forward_stripe<C>(sm, s1, L1, L2, s, nstripes, lane, sc, planeF, PW,
                      (s & 1) ? bnd0 : bnd1, (s & 1) ? bnd1 : bnd0, fend);
    __syncwarp();
  }
  // F_k(L1, L2) lives in the lane that owns column L2.
  {
    const int owner = ((L2 % SW) / C);
#pragma unroll
    for (int k = 0; k < 3; ++k) fend[k] = __shfl_sync(0xffffffffu, fend[k], owner);
  }

  // ---- backward -----------------------------------------------------------
  float bcap[3] = {DAFS_LOG_ZERO, DAFS_LOG_ZERO, DAFS_LOG_ZERO};
  for (int s = nstripes - 1; s >= 0; --s) {
    StripeColsB<C> sc;
    const int j0 = s * SW + lane * C;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int cls = col_class(s2, L2, j0 + c + 1);
      sc.mofs[c] = cls * 32 + lane;
      sc.ins[c] = sm.ins[cls * 32 + lane];
    }
    backward_stripe<C>(sm, s1, L1, L2, s, nstripes, lane, sc, planeB, PW,
                       (s & 1) ? bnd0 : bnd1, (s & 1) ? bnd1 : bnd0, bcap);
    __syncwarp();
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) bcap[k] = __shfl_sync(0xffffffffu, bcap[k], 0);

  // ---- total probability: ProbabilisticModel.h:337-365 ----------------------
  float tf = DAFS_LOG_ZERO;
#pragma unroll
  for (int k = 0; k < 3; ++k) tf = pairhmm_lse(sm.lut, tf, __fadd_rn(fend[k], c_tab.init[k]));
  const int c11 = s1[0], c21 = s2[0];
  const float f11 = __fadd_rn(c_tab.init[0], c_tab.match[c11 * 8 + c21]);   // :123-131 seeds
  const float f10 = __fadd_rn(c_tab.init[1], c_tab.ins[c11]);
  const float f01 = __fadd_rn(c_tab.init[2], c_tab.ins[c21]);
  float tb = __fadd_rn(f11, bcap[0]);
  tb = pairhmm_lse(sm.lut, tb, __fadd_rn(f10, bcap[1]));
  tb = pairhmm_lse(sm.lut, tb, __fadd_rn(f01, bcap[2]));
  const float total = __fmul_rn(__fadd_rn(tf, tb), 0.5f);                  // (tf+tb)/2 exactly
  if (lane == 0) a.pair_total[pair] = total;

  // ---- posterior + sparsify: :393-398, wrapper.cpp:125-128, align.cpp:69-78 --
  // One coalesced sweep over rows 1..L1 of both planes.  Kept entries are
  // compacted, already in CSR order, into the part of the planes the sweep has
  // consumed (entry k lands at index k < PW <= every index still to be read:
  // row 0 is never read), columns over B_M and values over F_M.  Then the warp
  // reserves its range of the output pool and copies the compact run there.
  int32_t* row_ptr = a.row_ptr + a.row_off[pair];
  const unsigned lt_mask = (1u << lane) - 1u;
  uint32_t* ccol = reinterpret_cast<uint32_t*>(planeB);
  float* cval = planeF;
  int run = 0;
  if (lane == 0) row_ptr[0] = 0;
  for (int i = 1; i <= L1; ++i) {
    const float* frow = planeF + (int64_t)i * PW;
    const float* brow = planeB + (int64_t)i * PW;
    for (int jb = 1; jb <= L2; jb += 128) {
      float f[4], bb[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = jb + 32 * u + lane;
        const bool in = j <= L2;
        f[u] = in ? frow[j] : DAFS_LOG_ZERO;
        bb[u] = in ? brow[j] : DAFS_LOG_ZERO;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (jb + 32 * u > L2) break;                      // warp-uniform
        const int j = jb + 32 * u + lane;
        const float x = __fsub_rn(__fadd_rn(f[u], bb[u]), total);   // forward + backward - total
        float p = 0.f;
        bool keep = false;
        if (j <= L2 && x > a.xcut) {
          p = pairhmm_exp(fminf(x, 0.0f));
          keep = (p >= a.th) && (p > a.th);
        }
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (keep) {
          const int k = run + __popc(m & lt_mask);
          ccol[k] = (uint32_t)(j - 1);
          cval[k] = p;
        }
        run += __popc(m);
      }
    }
    if (lane == 0) row_ptr[i] = run;
  }
  unsigned long long b = 0;
  if (lane == 0) {
    b = atomicAdd(a.pool_count, (unsigned long long)run);
    a.pair_pool_off[pair] = (int64_t)b;
    a.pair_nnz[pair] = run;
  }
  const int64_t base = (int64_t)__shfl_sync(0xffffffffu, b, 0);
  __syncwarp();
  if (base + run <= a.pool_cap) {
    for (int k = lane; k < run; k += 32) {
      a.pool_col[base + k] = ccol[k];
      a.pool_val[base + k] = cval[k];
    }
  }
}

template <int C, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB) pairhmm_kernel(const PairHmmArgs a) {
  __shared__ PairHmmSmem sm;
  if (threadIdx.x < 4) {
    const float4 rows[4] = {DAFS_LOOKUP_ROWS};
    sm.lut[threadIdx.x] = rows[threadIdx.x];
  }
  for (int k = threadIdx.x; k < 64 * 32; k += blockDim.x) sm.match[k] = c_tab.match[k >> 5];
  for (int k = threadIdx.x; k < 8 * 32; k += blockDim.x) sm.ins[k] = c_tab.ins[k >> 5];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int slot = blockIdx.x * WARPS + (threadIdx.x >> 5);
  for (;;) {
    int q = 0;
    if (lane == 0) q = atomicAdd(a.queue, 1);
    q = __shfl_sync(0xffffffffu, q, 0);
    if (q >= a.npairs) break;
    pairhmm_pair<C>(a, sm, a.order[q], slot, lane);
    __syncwarp();
  }
}

}  // namespace dafs
