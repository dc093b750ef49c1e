// K1/K2: batched ProbCons pair-HMM forward / backward / posterior for sm_100a.
//
// Reference: ProbabilisticModel::ComputeForwardMatrix / ComputeBackwardMatrix /
// ComputeTotalProbability / ComputePosteriorMatrix
// (src/probconsRNA/ProbabilisticModel.h:105-179, :197-259, :337-365, :374-403),
// the >= th / > th sparsification of wrapper.cpp:125-128 and align.cpp:69-78.
//
// Mapping (no tensor cores: nothing here is a contraction).  One warp owns one
// sequence pair at a time and pulls pairs from an atomic queue (persistent
// grid).  The DP matrix is cut into column stripes of 32*C columns; inside a
// stripe lane t owns C consecutive columns whose three state values of the
// previous row stay in registers, and the rows run through the lanes as a
// skewed pipeline: at step s lane t works on row s-t and hands its last column
// to lane t+1 with warp shuffles.  There is no block-level barrier anywhere in
// the DP.  The forward match plane F_M (the only O(L1*L2) intermediate) goes to
// a per-warp slot in HBM in PIPELINE order -- element (step, column slot c, lane),
// i.e. what the warp produces in one step is contiguous -- so the stores and, in
// the backward sweep, the loads are fully coalesced 128-byte lines (lane t at
// backward step s' needs exactly what it wrote at forward step L1 + 31 - s').
// The backward sweep fetches those lines one step ahead, never stores B_M, and
// emits the few cells that can pass the threshold as a candidate list, from which
// the CSR rows are built (pairhmm_post.cuh).  Row symbols and stripe-boundary
// columns reach the pipeline through 32-step register windows (one coalesced load
// per 32 steps, fetched a window ahead) and shuffles, so the inner loop has no
// load whose latency it has to wait for.
#pragma once
#include <cstdint>

#include "pairhmm_math.cuh"

namespace dafs {

struct PairHmmTables {   // log-space, symbol classes A C G U/T N other pad
  float init[3];         // ProbabilisticModel.h:76
  float trans[9];        // [from*3+to], :77-78
  float ins[8];          // :83-84   (class 6 = pad = LOG_ZERO)
  float match[64];       // [c1*8+c2], :85-86
};

struct PairHmmArgs {
  const uint8_t* seq;        // concatenated symbol classes
  const int64_t* seq_off;    // [nseq]
  const int32_t* seq_len;    // [nseq]
  const int2* pairs;         // [npairs] (row sequence, column sequence)
  const int32_t* order;      // [npairs] queue position -> pair index (longest first)
  int npairs;
  float th;                  // keep p > th
  float xcut;                // cells with F+B-total <= xcut cannot exceed th
  float* planes;             // per-warp-slot: F_M plane, candidate keys, candidate sums
  int64_t plane_stride;      // words per plane (a slot holds three)
  float4* bnd;               // per-warp-slot stripe-boundary columns (2 buffers)
  int64_t bnd_stride;        // float4 per buffer
  int* queue;                // work counter
  // outputs
  const int64_t* row_off;    // [npairs+1]
  int32_t* row_ptr;          // concatenated per-pair row_ptr blocks
  unsigned long long* pool_count;
  uint32_t* pool_col;
  float* pool_val;
  int64_t pool_cap;
  int64_t* pair_pool_off;    // [npairs] where the pair's entries start in the pool
  int32_t* pair_nnz;         // [npairs]
  float* pair_total;         // [npairs] total log-probability (diagnostic)
};

__constant__ PairHmmTables c_tab;

// Per-block shared tables, replicated once per lane (index*32 + lane) so that a
// warp whose lanes look up different symbols never has a bank conflict.
struct PairHmmSmem {
  float4 lut[DAFS_LOOKUP_SLOTS];   // LOOKUP coefficient rows {c3,c2,c1,c0} by ceil(2 d)
  float match[64 * 32];
  float ins[8 * 32];
};

template <int C>
struct StripeCols {       // per-lane constants of the current stripe
  int mofs[C];            // class(column j)*32 + lane : offset into smem.match row
  float ins[C];           // log insert emission of column j
};

// ---------------------------------------------------------------------------
// Forward sweep over one stripe.  Returns (through fend) F_k(L1, L2) when the
// stripe contains column L2.
// plane_row(plane, stripe, L1, C): start of the stripe's region, (L1 + 32) steps of 32*C floats
template <int C>
__device__ __forceinline__ float* stripe_plane(float* plane, int stripe, int L1) {
  return plane + (int64_t)stripe * (L1 + 32) * (32 * C);
}

template <int C>
__device__ __forceinline__ void forward_stripe(
    const PairHmmSmem& sm, const uint8_t* __restrict__ s1, int L1, int L2, int stripe,
    int nstripes, int lane, const StripeCols<C>& sc, float* __restrict__ plane,
    const float4* __restrict__ bnd_in, float4* __restrict__ bnd_out, float (&fend)[3]) {
  const float T00 = c_tab.trans[0], T01 = c_tab.trans[1], T02 = c_tab.trans[2];
  const float T10 = c_tab.trans[3], T11 = c_tab.trans[4];
  const float T20 = c_tab.trans[6], T22 = c_tab.trans[8];
  const int j0 = stripe * 32 * C + lane * C;
  const bool origin = (stripe == 0 && lane == 0);   // owns cells (0,0),(0,1),(1,0),(1,1)
  const bool has_left = stripe > 0;
  float* pout = stripe_plane<C>(plane, stripe, L1) + lane;

  float pM[C], pX[C], pY[C];                        // F(i-1, j0+c)
#pragma unroll
  for (int c = 0; c < C; ++c) pM[c] = pX[c] = pY[c] = DAFS_LOG_ZERO;
  float dM = DAFS_LOG_ZERO, dX = DAFS_LOG_ZERO, dY = DAFS_LOG_ZERO;   // F(i-1, j0-1)
  float oM = DAFS_LOG_ZERO, oX = DAFS_LOG_ZERO, oY = DAFS_LOG_ZERO;   // my last column, last step
  // Lane t works on row s-t at step s, so a row's symbol enters at lane 0 and moves
  // up one lane per step.  Lane j of the window of block b holds the symbol lane 0
  // needs at step 32 b + j (row 0 has no symbol; any finite class works there
  // because every term it feeds is LOG_ZERO-absorbed); the stripe-boundary column
  // of the stripe to the left enters the same way.
  auto load_sym = [&](int b) -> int { const int r = 32 * b + lane; return (r >= 1 && r <= L1) ? (int)s1[r - 1] : (int)kClsOther; };
  auto load_bnd = [&](int b) -> float4 {
    const int r = 32 * b + lane;
    return (has_left && r <= L1) ? bnd_in[r] : make_float4(DAFS_LOG_ZERO, DAFS_LOG_ZERO, DAFS_LOG_ZERO, 0.f);
  };
  int win = load_sym(0), win_next = load_sym(1);
  float4 bw = load_bnd(0), bw_next = load_bnd(1);
  int cls1 = kClsOther;
  const int nsteps = L1 + 1 + 31;
  for (int step = 0; step < nsteps; ++step) {
    const int wl = step & 31;
    float lM = __shfl_up_sync(0xffffffffu, oM, 1);
    float lX = __shfl_up_sync(0xffffffffu, oX, 1);
    float lY = __shfl_up_sync(0xffffffffu, oY, 1);
    {
      const int c_in = __shfl_sync(0xffffffffu, win, wl);
      const int c_up = __shfl_up_sync(0xffffffffu, cls1, 1);
      cls1 = lane == 0 ? c_in : c_up;
    }
    if (has_left) {
      const float bM = __shfl_sync(0xffffffffu, bw.x, wl);
      const float bX = __shfl_sync(0xffffffffu, bw.y, wl);
      const float bY = __shfl_sync(0xffffffffu, bw.z, wl);
      if (lane == 0) { lM = bM; lX = bX; lY = bY; }
    } else if (lane == 0) {
      lM = lX = lY = DAFS_LOG_ZERO;
    }
    if (wl == 31) {
      win = win_next; bw = bw_next;
      const int b = (step >> 5) + 2;
      win_next = load_sym(b); bw_next = load_bnd(b);
    }
    const int i = step - lane;
    if (i < 0 || i > L1) continue;
    const float insRow = sm.ins[cls1 * 32 + lane];
    const float* mrow = sm.match + cls1 * 8 * 32;
    float leftM = lM, leftY = lY;
    float gM = dM, gX = dX, gY = dY;                           // diagonal F(i-1, j-1)
    float* prow = pout + (int64_t)step * (32 * C);
#pragma unroll
    for (int c = 0; c < C; ++c) {
      // ProbabilisticModel.h:151-156
      float nM = pairhmm_lse(sm.lut, pairhmm_lse(sm.lut, __fadd_rn(gM, T00), __fadd_rn(gX, T10)), __fadd_rn(gY, T20));
      nM = __fadd_rn(nM, mrow[sc.mofs[c]]);
      // :157-162
      float nX = __fadd_rn(insRow, pairhmm_lse(sm.lut, __fadd_rn(pM[c], T01), __fadd_rn(pX[c], T11)));
      // :163-168
      float nY = __fadd_rn(sc.ins[c], pairhmm_lse(sm.lut, __fadd_rn(leftM, T02), __fadd_rn(leftY, T22)));
      if (c < 2 && origin && i <= 1) {
        // seeds, ProbabilisticModel.h:123-131; (0,0) and the other states stay LOG_ZERO (:150)
        nM = (i == 1 && c == 1) ? __fadd_rn(c_tab.init[0], mrow[sc.mofs[1]]) : DAFS_LOG_ZERO;
        nX = (i == 1 && c == 0) ? __fadd_rn(c_tab.init[1], insRow) : DAFS_LOG_ZERO;
        nY = (i == 0 && c == 1) ? __fadd_rn(c_tab.init[2], sc.ins[1]) : DAFS_LOG_ZERO;
      }
      gM = pM[c]; gX = pX[c]; gY = pY[c];
      pM[c] = nM; pX[c] = nX; pY[c] = nY;
      leftM = nM; leftY = nY;
      prow[c * 32] = nM;
    }
    dM = lM; dX = lX; dY = lY;
    oM = pM[C - 1]; oX = pX[C - 1]; oY = pY[C - 1];
    if (lane == 31 && stripe + 1 < nstripes) bnd_out[i] = make_float4(oM, oX, oY, 0.f);
    if (i == L1) {
#pragma unroll
      for (int c = 0; c < C; ++c)
        if (j0 + c == L2) { fend[0] = pM[c]; fend[1] = pX[c]; fend[2] = pY[c]; }
    }
  }
}

}  // namespace dafs
