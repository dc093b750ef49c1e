// Backward sweep of one column stripe (ProbabilisticModel.h:197-259), fused with
// the first half of the posterior: B_M never goes to memory.  The forward plane is
// stored in pipeline order (pairhmm_kernel.cuh), so the C values a lane needs at
// step s' are the ones it wrote at forward step L1 + 31 - s': C coalesced loads,
// issued one whole step before they are used.  Every cell whose F_M + B_M clears a
// loose bound
// (candidate test, see pairhmm_pair) is appended to the pair's candidate list as
// (row << 16 | column, F_M + B_M).  Rows of one stripe leave the pipeline from the
// highest column to the lowest and stripes run right to left, so inside the list
// the entries of any one row are in strictly descending column order.
// Mirror image of forward_stripe: lane 31 starts, rows run L1..0, each lane walks
// its columns right to left and hands its first column to lane t-1.
#pragma once
#include "pairhmm_kernel.cuh"

namespace dafs {

template <int C>
struct StripeColsB {      // per-lane constants: column j+1 of each owned column j
  int mofs[C];            // class(column j+1)*32 + lane
  float ins[C];           // log insert emission of column j+1
};

// bcap[0..2] receive B_M(1,1), B_X(1,0), B_Y(0,1) (needed by
// ComputeTotalProbability, :350-361) when this is stripe 0.
template <int C>
__device__ __forceinline__ void backward_stripe(
    const PairHmmSmem& sm, const uint8_t* __restrict__ s1, int L1, int L2, int stripe,
    int nstripes, int lane, const StripeColsB<C>& sc, float* __restrict__ planeF,
    const float4* __restrict__ bnd_in, float4* __restrict__ bnd_out, float (&bcap)[3],
    float lthr, uint32_t* __restrict__ ckey, float* __restrict__ csum, int& run) {
  const float T00 = c_tab.trans[0], T01 = c_tab.trans[1], T02 = c_tab.trans[2];
  const float T10 = c_tab.trans[3], T11 = c_tab.trans[4];
  const float T20 = c_tab.trans[6], T22 = c_tab.trans[8];
  const int j0 = stripe * 32 * C + lane * C;
  const int rl = 31 - lane;                          // pipeline position
  const bool has_right = stripe + 1 < nstripes;
  const float* pin = stripe_plane<C>(planeF, stripe, L1) + lane;

  float pM[C], pX[C], pY[C];                         // B(i+1, j0+c)
#pragma unroll
  for (int c = 0; c < C; ++c) pM[c] = pX[c] = pY[c] = DAFS_LOG_ZERO;
  float dM = DAFS_LOG_ZERO;                          // B_M(i+1, j0+C)
  float oM = DAFS_LOG_ZERO, oY = DAFS_LOG_ZERO;      // my first column, last step
  // Row i needs the symbol of row i+1 (row L1 has none: any finite class, it only
  // meets LOG_ZERO terms).  Rows enter at lane 31 and move down one lane per step:
  // lane j of the window of block b holds what lane 31 needs at step 32 b + j, the
  // symbol of row L1 - (32 b + j) + 1 and the boundary column of the stripe to the right.
  auto load_sym = [&](int b) -> int { const int r = L1 - (32 * b + lane) + 1; return (r >= 1 && r <= L1) ? (int)s1[r - 1] : (int)kClsOther; };
  auto load_bnd = [&](int b) -> float4 {
    const int r = L1 - (32 * b + lane);
    return (has_right && r >= 0) ? bnd_in[r] : make_float4(DAFS_LOG_ZERO, 0.f, DAFS_LOG_ZERO, 0.f);
  };
  int win = load_sym(0), win_next = load_sym(1);
  float4 bw = load_bnd(0), bw_next = load_bnd(1);
  int cls1 = kClsOther;
  const int nsteps = L1 + 1 + 31;
  // forward values of the cells of step s': written at forward step nsteps - 1 - s'
  float fnext[C];
#pragma unroll
  for (int c = 0; c < C; ++c) fnext[c] = __ldcs(pin + ((int64_t)(nsteps - 1) * C + c) * 32);
  for (int step = 0; step < nsteps; ++step) {
    const int wl = step & 31;
    float rM = __shfl_down_sync(0xffffffffu, oM, 1);
    float rY = __shfl_down_sync(0xffffffffu, oY, 1);
    {
      const int c_in = __shfl_sync(0xffffffffu, win, wl);
      const int c_dn = __shfl_down_sync(0xffffffffu, cls1, 1);
      cls1 = lane == 31 ? c_in : c_dn;
    }
    if (has_right) {
      const float bM = __shfl_sync(0xffffffffu, bw.x, wl);
      const float bY = __shfl_sync(0xffffffffu, bw.z, wl);
      if (lane == 31) { rM = bM; rY = bY; }
    } else if (lane == 31) {
      rM = rY = DAFS_LOG_ZERO;
    }
    if (wl == 31) {
      win = win_next; bw = bw_next;
      const int b = (step >> 5) + 2;
      win_next = load_sym(b); bw_next = load_bnd(b);
    }
    float fv[C];
#pragma unroll
    for (int c = 0; c < C; ++c) fv[c] = fnext[c];
    if (step + 1 < nsteps) {
#pragma unroll
      for (int c = 0; c < C; ++c) fnext[c] = __ldcs(pin + ((int64_t)(nsteps - 2 - step) * C + c) * 32);
    }
    const int i = L1 - (step - rl);
    unsigned cm = 0;                                 // candidate cells of this lane's row
    if (i >= 0 && i <= L1) {
    const float insRow = sm.ins[cls1 * 32 + lane];
    const float* mrow = sm.match + cls1 * 8 * 32;
    float rightY = rY;                               // B_Y(i, j+1)
    float gM = dM;                                   // B_M(i+1, j+1)
    const int seed = (i == L1) ? (L2 - j0) : -1;     // column slot of cell (L1, L2)
#pragma unroll
    for (int c = C - 1; c >= 0; --c) {
      // ProbabilisticModel.h:233-237
      const float pxy = __fadd_rn(gM, mrow[sc.mofs[c]]);
      // :238-243   (bx + ins) + trans, left to right like the reference
      const float bxi = __fadd_rn(pX[c], insRow);
      // :244-249
      const float byi = __fadd_rn(rightY, sc.ins[c]);
      float nM = pairhmm_lse(sm.lut, pairhmm_lse(sm.lut, __fadd_rn(pxy, T00), __fadd_rn(bxi, T01)), __fadd_rn(byi, T02));
      float nX = pairhmm_lse(sm.lut, __fadd_rn(pxy, T10), __fadd_rn(bxi, T11));
      float nY = pairhmm_lse(sm.lut, __fadd_rn(pxy, T20), __fadd_rn(byi, T22));
      if (c == seed) { nM = c_tab.init[0]; nX = c_tab.init[1]; nY = c_tab.init[2]; }   // :213-214
      gM = pM[c];
      pM[c] = nM; pX[c] = nX; pY[c] = nY;
      rightY = nY;
      fv[c] = __fadd_rn(fv[c], nM);                  // forward + backward (:393-397)
      if (fv[c] > lthr) cm |= 1u << c;
    }
    {
      // rows 1..L1 and columns 1..L2 only (ComputePosteriorMatrix, :386-399)
      const int lo = max(0, 1 - j0), hi = min(C, L2 + 1 - j0);
      cm = (i >= 1 && lo < hi) ? (cm >> lo << lo) & ((1u << hi) - 1u) : 0u;
    }
    dM = rM;
    oM = pM[0]; oY = pY[0];
    if (lane == 0 && stripe > 0) bnd_out[i] = make_float4(oM, 0.f, oY, 0.f);
    if (stripe == 0 && lane == 0) {
      if (i == 1) { bcap[0] = pM[1]; bcap[1] = pX[0]; }
      if (i == 0) bcap[2] = pY[1];
    }
    }  // active row
    // ---- append this step's candidates (warp-uniform control flow) ----------
    const int n = __popc(cm);
    if (__any_sync(0xffffffffu, n != 0)) {
      int incl = n;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
      }
      int pos = run + incl - n;
#pragma unroll
      for (int c = C - 1; c >= 0; --c)
        if (cm >> c & 1u) {
          ckey[pos] = ((uint32_t)i << 16) | (uint32_t)(j0 + c);
          csum[pos] = fv[c];
          ++pos;
        }
      run += __shfl_sync(0xffffffffu, incl, 31);
    }
  }
}

}  // namespace dafs
