// Backward sweep of one column stripe (ProbabilisticModel.h:197-259); B_M goes to
// its own plane so the DP loop issues no global loads.
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
    int nstripes, int lane, const StripeColsB<C>& sc, float* __restrict__ plane, int PW,
    const float4* __restrict__ bnd_in, float4* __restrict__ bnd_out, float (&bcap)[3]) {
  const float T00 = c_tab.trans[0], T01 = c_tab.trans[1], T02 = c_tab.trans[2];
  const float T10 = c_tab.trans[3], T11 = c_tab.trans[4];
  const float T20 = c_tab.trans[6], T22 = c_tab.trans[8];
  const int j0 = stripe * 32 * C + lane * C;
  const int rl = 31 - lane;                          // pipeline position

  float pM[C], pX[C], pY[C];                         // B(i+1, j0+c)
#pragma unroll
  for (int c = 0; c < C; ++c) pM[c] = pX[c] = pY[c] = DAFS_LOG_ZERO;
  float dM = DAFS_LOG_ZERO;                          // B_M(i+1, j0+C)
  float oM = DAFS_LOG_ZERO, oY = DAFS_LOG_ZERO;      // my first column, last step
  // Row i needs the symbol of row i+1; row L1 has none (any finite class: it only
  // meets LOG_ZERO terms).
  int cls_next = kClsOther;
  const int nsteps = L1 + 1 + 31;
  for (int step = 0; step < nsteps; ++step) {
    float rM = __shfl_down_sync(0xffffffffu, oM, 1);
    float rY = __shfl_down_sync(0xffffffffu, oY, 1);
    const int i = L1 - (step - rl);
    const int cls1 = cls_next;                       // class of seq1[i+1] (1-based)
    // next step works on row i-1 and needs the symbol of row i: s1[i-1]
    if (i >= 1 && i <= L1) cls_next = s1[i - 1];
    if (i < 0 || i > L1) continue;
    if (lane == 31) {
      if (stripe + 1 < nstripes) { const float4 b = bnd_in[i]; rM = b.x; rY = b.z; }
      else { rM = rY = DAFS_LOG_ZERO; }
    }
    const float insRow = sm.ins[cls1 * 32 + lane];
    const float* mrow = sm.match + cls1 * 8 * 32;
    float rightY = rY;                               // B_Y(i, j+1)
    float gM = dM;                                   // B_M(i+1, j+1)
    float* prow = plane + (int64_t)i * PW + j0;            // B_M plane
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
      prow[c] = nM;
    }
    dM = rM;
    oM = pM[0]; oY = pY[0];
    if (lane == 0 && stripe > 0) bnd_out[i] = make_float4(oM, 0.f, oY, 0.f);
    if (stripe == 0 && lane == 0) {
      if (i == 1) { bcap[0] = pM[1]; bcap[1] = pX[0]; }
      if (i == 0) bcap[2] = pY[1];
    }
  }
}

}  // namespace dafs
