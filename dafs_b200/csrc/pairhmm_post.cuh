// Per-pair driver: forward stripes, backward stripes, total probability,
// posterior + threshold + CSR emission; and the persistent kernel around it.
#pragma once
#include "pairhmm_backward.cuh"

namespace dafs {

#define DAFS_PAIRHMM_CNT_ROWS 2048

// Posterior + sparsify (ProbabilisticModel.h:393-398, wrapper.cpp:125-128, align.cpp:69-78) of one pair's
// candidate list.  Pass 1 applies the reference's test to every candidate (the posterior replaces
// the sum, a dropped candidate gets key ~0) and counts the survivors per row;
// a scan turns the counts into row_ptr; pass 2 walks
// the list in order and places each survivor from the END of its row backwards,
// which turns the descending column order of the list into ascending CSR order.
// Both passes are chains of memory round trips: pass 1 keeps four blocks of 32 candidates in flight, pass 2
// fetches the next block while it places the current one.
template <bool SM>
__device__ __forceinline__ void pairhmm_emit(const PairHmmArgs& a, int* cnt, int pair, int L1, int lane, int run,
                                             float total, uint32_t* __restrict__ ckey, float* __restrict__ csum,
                                             int32_t* __restrict__ row_ptr) {
  auto cnt_ld = [&](int i) -> int { return SM ? cnt[i] : __ldcg(cnt + i); };
  auto cnt_st = [&](int i, int v) { if (SM) cnt[i] = v; else __stcg(cnt + i, v); };
  const unsigned lt_mask = (1u << lane) - 1u;
  for (int i = lane; i <= L1; i += 32) cnt_st(i, 0);
  __syncwarp();
  for (int k0 = 0; k0 < run; k0 += 128) {
    uint32_t key[4];
    float sv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = k0 + u * 32 + lane;
      key[u] = 0; sv[u] = 0.f;
      if (k < run) { key[u] = ckey[k]; sv[u] = csum[k]; }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = k0 + u * 32 + lane;
      if (k < run) {
        const float x = __fsub_rn(sv[u], total);                           // forward + backward - total
        float p = 0.f;
        bool keep = false;
        if (x > a.xcut) {
          p = pairhmm_exp(fminf(x, 0.0f));
          keep = (p >= a.th) && (p > a.th);
        }
        if (keep) { csum[k] = p; atomicAdd(cnt + (key[u] >> 16), 1); }
        else ckey[k] = 0xffffffffu;
      }
    }
  }
  __syncwarp();
  int kept = 0;
  if (lane == 0) row_ptr[0] = 0;
  for (int i0 = 1; i0 <= L1; i0 += 32) {
    const int i = i0 + lane;
    int incl = i <= L1 ? cnt_ld(i) : 0;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (i <= L1) { row_ptr[i] = kept + incl; cnt_st(i, kept + incl); }     // cnt[i] <- end of row i
    kept += __shfl_sync(0xffffffffu, incl, 31);
  }
  unsigned long long b = 0;
  if (lane == 0) {
    b = atomicAdd(a.pool_count, (unsigned long long)kept);
    a.pair_pool_off[pair] = (int64_t)b;
    a.pair_nnz[pair] = kept;
  }
  const int64_t base = (int64_t)__shfl_sync(0xffffffffu, b, 0);
  __syncwarp();
  if (base + kept <= a.pool_cap) {
    uint32_t keyn = lane < run ? ckey[lane] : 0xffffffffu;
    float pn = lane < run ? csum[lane] : 0.f;
    for (int k0 = 0; k0 < run; k0 += 32) {
      const uint32_t key = keyn;
      const float pv = pn;
      {
        const int kn = k0 + 32 + lane;
        keyn = 0xffffffffu;
        if (kn < run) { keyn = ckey[kn]; pn = csum[kn]; }
      }
      const bool keep = key != 0xffffffffu;
      const int row = (int)(key >> 16);
      const unsigned same = __match_any_sync(0xffffffffu, keep ? row : 0x10000 + lane);
      const int end = keep ? cnt_ld(row) : 0;
      __syncwarp();                                                             // every read before any update
      if (keep) {
        const int pos = end - 1 - __popc(same & lt_mask);
        a.pool_col[base + pos] = (key & 0xffffu) - 1u;
        a.pool_val[base + pos] = pv;
        if ((same >> lane) == 1u) cnt_st(row, end - __popc(same));              // highest lane of the group
      }
      __syncwarp();
    }
  }
}

// One warp, one pair.  Returns nnz (same value in every lane).
template <int C>
__device__ __forceinline__ void pairhmm_pair(const PairHmmArgs& a, const PairHmmSmem& sm, PairHmmRing& ring, int* cnt_sm,
                                             int pair, int slot, int lane) {
  const int2 pr = a.pairs[pair];
  const uint8_t* s1 = a.seq + a.seq_off[pr.x];
  const uint8_t* s2 = a.seq + a.seq_off[pr.y];
  const int L1 = a.seq_len[pr.x], L2 = a.seq_len[pr.y];
  const uint32_t lut = *(volatile const uint32_t*)&sm.lut_addr[lane];
  const int SW = 32 * C;
  const int nstripes = (L2 + 1 + SW - 1) / SW;
  const int PW = nstripes * SW;
  // slot = forward plane | candidate keys | candidate sums (plane_stride words each)
  float* planeF = a.planes + (int64_t)slot * 3 * a.plane_stride;
  uint32_t* ckey = reinterpret_cast<uint32_t*>(planeF + a.plane_stride);
  float* csum = planeF + 2 * a.plane_stride;
  float4* bnd0 = a.bnd + (int64_t)slot * 2 * a.bnd_stride;
  float4* bnd1 = bnd0 + a.bnd_stride;

  // ---- forward ------------------------------------------------------------
  float fend[3] = {DAFS_LOG_ZERO, DAFS_LOG_ZERO, DAFS_LOG_ZERO};
  forward_sweep<C>(sm, ring, s1, s2, L1, L2, nstripes, lane, planeF, bnd0, bnd1, fend);
  __syncwarp();
  // F_k(L1, L2) lives in the lane that owns column L2.
  {
    const int owner = ((L2 % SW) / C);
#pragma unroll
    for (int k = 0; k < 3; ++k) fend[k] = __shfl_sync(0xffffffffu, fend[k], owner);
  }

  // ---- total probability, forward half: ProbabilisticModel.h:337-349 -----------
  PairHmmCoef KT;
  float tf = DAFS_LOG_ZERO;
#pragma unroll
  for (int k = 0; k < 3; ++k) tf = pairhmm_lse(lut, KT, tf, __fadd_rn(fend[k], c_tab.init[k]));
  const int c11 = s1[0], c21 = s2[0];
  const float f11 = __fadd_rn(c_tab.init[0], c_tab.match[c11 * 8 + c21]);   // :123-131 seeds
  const float f10 = __fadd_rn(c_tab.init[1], c_tab.ins[c11]);
  const float f01 = __fadd_rn(c_tab.init[2], c_tab.ins[c21]);

  // ---- backward + candidate cells ---------------------------------------------
  // The posterior of a cell is EXP(F_M + B_M - total) with total = (tf + tb) / 2,
  // and tb is only known when the backward sweep ends.  The sweep therefore keeps
  // every cell with F_M + B_M > tf + xcut - kSlack; as long as |total - tf| <
  // kSlack / 2 that is a superset of the cells the exact test below keeps.  If the
  // two totals ever disagree by more, the sweep is repeated keeping every cell.
  const float kSlack = 0.25f;
  float lthr = __fsub_rn(__fadd_rn(tf, a.xcut), kSlack);
  float total;
  int run;
  for (;;) {
    float bcap[3] = {DAFS_LOG_ZERO, DAFS_LOG_ZERO, DAFS_LOG_ZERO};
    run = 0;
    backward_sweep<C>(sm, ring, s1, s2, L1, L2, nstripes, lane, planeF, bnd0, bnd1, bcap,
                      lthr, lthr == -INFINITY, ckey, csum, run);
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 3; ++k) bcap[k] = __shfl_sync(0xffffffffu, bcap[k], 0);
    // backward half of the total: ProbabilisticModel.h:350-361
    float tb = __fadd_rn(f11, bcap[0]);
    tb = pairhmm_lse(lut, KT, tb, __fadd_rn(f10, bcap[1]));
    tb = pairhmm_lse(lut, KT, tb, __fadd_rn(f01, bcap[2]));
    total = __fmul_rn(__fadd_rn(tf, tb), 0.5f);                            // (tf+tb)/2 exactly
    if (lthr == -INFINITY || fabsf(__fsub_rn(total, tf)) < 0.5f * kSlack) break;
    lthr = -INFINITY;
  }
  if (lane == 0) a.pair_total[pair] = total;

  // ---- posterior + sparsify (pairhmm_emit) --------------------------------------
  // The per-row counters live in shared memory when the pair has at most DAFS_PAIRHMM_CNT_ROWS rows (both
  // passes wait for them), else in the dead forward plane.
  int32_t* row_ptr = a.row_ptr + a.row_off[pair];
  if (L1 < DAFS_PAIRHMM_CNT_ROWS) pairhmm_emit<true>(a, cnt_sm, pair, L1, lane, run, total, ckey, csum, row_ptr);
  else pairhmm_emit<false>(a, reinterpret_cast<int*>(planeF), pair, L1, lane, run, total, ckey, csum, row_ptr);
}

template <int C, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB) pairhmm_kernel(const PairHmmArgs a) {
  __shared__ PairHmmSmem sm;
  __shared__ PairHmmRing rings[WARPS];
  __shared__ int cnt_sm[WARPS][DAFS_PAIRHMM_CNT_ROWS];
  for (int k = threadIdx.x; k < DAFS_LOOKUP_SLOTS * DAFS_LOOKUP_COPIES; k += blockDim.x) {
    const float4 rows[4] = {DAFS_LOOKUP_ROWS};
    sm.lut[k] = rows[pairhmm_lookup_piece(k / DAFS_LOOKUP_COPIES)];
  }
  if (threadIdx.x < 32) sm.lut_addr[threadIdx.x] = pairhmm_lut_addr(sm.lut, threadIdx.x);
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
    pairhmm_pair<C>(a, sm, rings[threadIdx.x >> 5], cnt_sm[threadIdx.x >> 5], a.order[q], slot, lane);
    __syncwarp();
  }
}

}  // namespace dafs
