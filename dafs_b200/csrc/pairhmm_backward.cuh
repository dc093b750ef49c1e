// Backward sweep over all stripes of one pair (ProbabilisticModel.h:197-259), fused with
// the first half of the posterior: B_M never goes to memory.  The forward plane is
// stored in pipeline order (pairhmm_kernel.cuh), so the C values a lane needs for a row
// are the ones it wrote for that row in the forward sweep: C coalesced loads,
// issued one whole step before they are used.  Every cell whose F_M + B_M clears a
// loose bound (candidate test, see pairhmm_pair) is appended to the pair's candidate list as
// (row << 16 | column, F_M + B_M).  Rows of one stripe leave the pipeline from the
// highest column to the lowest and stripes run right to left (a row of the next stripe enters
// lane 31 at least one period after the same row of this stripe), so inside the list
// the entries of any one row are in strictly descending column order.
// Mirror image of forward_sweep: lane 31 starts, rows run L1..0, each lane walks
// its columns right to left and hands its first column to lane t-1; stripes follow each
// other without a gap.
#pragma once
#include "pairhmm_kernel.cuh"

namespace dafs {

// bcap[0..2] receive B_M(1,1), B_X(1,0), B_Y(0,1) (needed by ComputeTotalProbability, :350-361) in lane 0.
// keep_all: no candidate test (every cell of rows 1..L1, columns 1..L2 is appended).
template <int C>
__device__ __forceinline__ void backward_sweep(
    const PairHmmSmem& sm, PairHmmRing& ring, const uint8_t* __restrict__ s1, const uint8_t* __restrict__ s2,
    int L1, int L2, int nstripes, int lane, float* __restrict__ planeF,
    float4* __restrict__ bnd0, float4* __restrict__ bnd1, float (&bcap)[3],
    float lthr, bool keep_all, uint32_t* __restrict__ ckey, float* __restrict__ csum, int& run) {
  const uint32_t lut = *(volatile const uint32_t*)&sm.lut_addr[lane];
  const uint32_t match0 = (uint32_t)__cvta_generic_to_shared(sm.match);
  const float T00 = c_tab.trans[0], T01 = c_tab.trans[1], T02 = c_tab.trans[2];
  const float T10 = c_tab.trans[3], T11 = c_tab.trans[4];
  const float T20 = c_tab.trans[6], T22 = c_tab.trans[8];
  constexpr int SW = 32 * C;
  const int P = max(L1 + 1, DAFS_PAIRHMM_PERIOD_MIN);
  const int G = (nstripes - 1) * P + L1 + 1 + 31;
  const int kNever = -0x40000000;
  const int rl = 31 - lane;                          // pipeline position

  // my current row: the q-th stripe from the right (stripe nstripes - 1 - q), row L1 - r
  int q = 0, r = -rl;
  uint32_t mofs[C];                                  // shared address of match[0][class(column j+1)][lane]
  float insc[C];                                     // log insert emission of column j+1
  int j0 = 0;
  auto load_cols = [&]() {
    j0 = (nstripes - 1 - q) * SW + lane * C;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int cls = col_class(s2, L2, j0 + c + 1);
      mofs[c] = match0 + (uint32_t)(cls * 32 + lane) * 4u;
      insc[c] = sm.ins[cls * 32 + lane];
    }
  };
  load_cols();
  float pM[C], pX[C], pY[C];                         // B(i+1, j0+c)
#pragma unroll
  for (int c = 0; c < C; ++c) pM[c] = pX[c] = pY[c] = DAFS_LOG_ZERO;
  float dM = DAFS_LOG_ZERO;                          // B_M(i+1, j0+C)
  // forward values of row L1 - r of my stripe: stripe_plane + (L1 - r + lane) * SW + lane  (+ c * 32)
  const float* pin = stripe_plane<C>(planeF, nstripes - 1, L1) + (int64_t)(L1 - r + lane) * SW + lane;
  float fnext[C];
#pragma unroll
  for (int c = 0; c < C; ++c) fnext[c] = 0.f;
  if (r == 0) {
#pragma unroll
    for (int c = 0; c < C; ++c) fnext[c] = __ldcs(pin + c * 32);
  }

  // ring entry of global step g = kq * P + rq: row L1 - rq of the kq-th stripe from the right needs the symbol of
  // the row below it and the boundary column written by lane 0 of the stripe to its right
  int kq = 0, rq = lane;
  auto fetch = [&]() -> float4 {
    float4 v = make_float4(DAFS_LOG_ZERO, 0.f, DAFS_LOG_ZERO, __int_as_float((int)kClsOther));
    if (rq <= L1 && kq < nstripes) {
      const int row = L1 - rq;
      if (kq >= 1) {
        const float4 t = __ldcg(((kq - 1) & 1 ? bnd1 : bnd0) + row);
        v.x = t.x; v.z = t.z;
      }
      if (row < L1) v.w = __int_as_float((int)s1[row]);
    }
    rq += 32;
    if (rq >= P) { rq -= P; ++kq; }
    return v;
  };
  float4 pre = fetch();
  int cls1 = kClsOther;
  PairHmmCoef KM[DAFS_PAIRHMM_MSETS], KX, KY;                            // coefficient sets: calls that share one run one after the other

  auto step = [&](int g) {
    float rM = __shfl_down_sync(0xffffffffu, pM[0], 1);
    float rY = __shfl_down_sync(0xffffffffu, pY[0], 1);
    int cd = __shfl_down_sync(0xffffffffu, cls1, 1);
    if (r == P) {                                    // on to row L1 of the next stripe to the left -- only now: the
      ++q;                                           // shuffles above still had to hand row 0 to the next lane
      if (q < nstripes) {
        r = 0;
        load_cols();
#pragma unroll
        for (int c = 0; c < C; ++c) pM[c] = pX[c] = pY[c] = DAFS_LOG_ZERO;
        dM = DAFS_LOG_ZERO;
        pin = stripe_plane<C>(planeF, nstripes - 1 - q, L1) + (int64_t)(L1 + lane) * SW + lane;
      } else {
        r = kNever;
      }
    }
    if (lane == 31) {
      const float4 b = ring.e[(g >> 5) & 1][g & 31];
      rM = b.x; rY = b.z; cd = __float_as_int(b.w);
    }
    cls1 = cd;
    float fv[C];
#pragma unroll
    for (int c = 0; c < C; ++c) fv[c] = fnext[c];
    {
      // fetch the forward values of the next step's row
      int rn = r + 1;
      const float* pn = pin - SW;
      if (rn == P) {
        rn = kNever;
        if (q + 1 < nstripes) { rn = 0; pn = stripe_plane<C>(planeF, nstripes - 2 - q, L1) + (int64_t)(L1 + lane) * SW + lane; }
      }
      if ((unsigned)rn <= (unsigned)L1) {
#pragma unroll
        for (int c = 0; c < C; ++c) fnext[c] = __ldcs(pn + c * 32);
      }
    }
    unsigned cm = 0;                                 // candidate cells of this lane's row
    const int i = L1 - r;
    if ((unsigned)r <= (unsigned)L1) {
      const float insRow = sm.ins[cls1 * 32 + lane];
      const uint32_t mrow = (uint32_t)cls1 * (8 * 32 * 4);
      float rightY = rY;                             // B_Y(i, j+1)
      float gM = dM;                                 // B_M(i+1, j+1)
#pragma unroll
      for (int c = C - 1; c >= 0; --c) {
        // ProbabilisticModel.h:233-237
        const float pxy = __fadd_rn(gM, pairhmm_lds(mofs[c] + mrow));
        // :238-243   (bx + ins) + trans, left to right like the reference
        const float bxi = __fadd_rn(pX[c], insRow);
        // :244-249
        const float byi = __fadd_rn(rightY, insc[c]);
        float nM = pairhmm_lse(lut, KM[c % DAFS_PAIRHMM_MSETS], pairhmm_lse(lut, KM[c % DAFS_PAIRHMM_MSETS], __fadd_rn(pxy, T00), __fadd_rn(bxi, T01)), __fadd_rn(byi, T02));
        float nX = pairhmm_lse(lut, KX, __fadd_rn(pxy, T10), __fadd_rn(bxi, T11));
        float nY = pairhmm_lse(lut, KY, __fadd_rn(pxy, T20), __fadd_rn(byi, T22));
        gM = pM[c];
        pM[c] = nM; pX[c] = nX; pY[c] = nY;
        rightY = nY;
        fv[c] = __fadd_rn(fv[c], nM);                // forward + backward (:393-397)
        if (fv[c] > lthr) cm |= 1u << c;
      }
      if (r == 0 && q == 0 && (unsigned)(L2 - j0) < (unsigned)C) {
        // Row L1 of the lane that owns column L2 (once per pair, one lane): cell (L1, L2) holds the initial
        // distribution (:213-214).  The cells to its left in this lane depend on it, and in row L1 -- nothing but
        // LOG_ZERO below -- the recurrence collapses to its Y term: with byi = B_Y(L1, j+1) + ins(j+1),
        // B_M = byi + T02, B_Y = byi + T22 (every LOG_ADD returns its larger argument, the other one being
        // LOG_ZERO) and B_X stays as computed.  Their forward values are fetched again.
        asm volatile("");
        const int cs = L2 - j0;
        float ry = DAFS_LOG_ZERO;
#pragma unroll
        for (int c = C - 1; c >= 0; --c) {
          if (c == cs) { pM[c] = c_tab.init[0]; pX[c] = c_tab.init[1]; pY[c] = c_tab.init[2]; }
          else if (c < cs) {
            const float byi = __fadd_rn(ry, insc[c]);
            pM[c] = __fadd_rn(byi, T02);
            pY[c] = __fadd_rn(byi, T22);
          }
          if (c <= cs) {
            fv[c] = __fadd_rn(__ldcs(pin + c * 32), pM[c]);
            cm = (cm & ~(1u << c)) | ((fv[c] > lthr ? 1u : 0u) << c);
          }
          ry = pY[c];
        }
      }
      if (keep_all) {
        // rows 1..L1 and columns 1..L2 only (ComputePosteriorMatrix, :386-399); with a finite bound the cells
        // outside carry a forward value of LOG_ZERO and fail the test by themselves
        const int lo = max(0, 1 - j0), hi = min(C, L2 + 1 - j0);
        cm = (i >= 1 && lo < hi) ? (cm >> lo << lo) & ((1u << hi) - 1u) : 0u;
      }
      dM = rM;
      if (lane == 0 && q + 1 < nstripes) ((q & 1) ? bnd1 : bnd0)[i] = make_float4(pM[0], 0.f, pY[0], 0.f);
      if (q == nstripes - 1 && lane == 0 && i <= 1) {     // the last two steps of the sweep
        asm volatile("");
        if (i == 1) { bcap[0] = pM[1]; bcap[1] = pX[0]; }
        if (i == 0) bcap[2] = pY[1];
      }
    }
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
    ++r; pin -= SW;
  };

  for (int g0 = 0; g0 < G; g0 += 32) {
    __syncwarp();
    ring.e[(g0 >> 5) & 1][lane] = pre;
    __syncwarp();
    pre = fetch();
    const int gend = min(g0 + 32, G);
#pragma unroll 1
    for (int g = g0; g < gend; ++g) step(g);
  }
}

}  // namespace dafs
