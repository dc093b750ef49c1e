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
#include <type_traits>

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
  float4 lut[DAFS_LOOKUP_SLOTS * DAFS_LOOKUP_COPIES];   // LOOKUP coefficient rows {c3,c2,c1,c0} by ceil(2 d), 8 copies each
  float match[64 * 32];
  float ins[8 * 32];
  uint32_t lut_addr[32];           // pairhmm_lut_addr(lut, lane), written by one warp, read by all
};

// Per-warp ring in shared memory: what lane 0 (forward) / lane 31 (backward) takes in at each step --
// the boundary column handed over by the neighbouring stripe and the symbol class of the row (.w, as
// bits).  Two blocks of 32 steps; a block is fetched from HBM one block ahead (one coalesced load per
// lane), parked in registers, and stored here when its turn comes.
struct PairHmmRing { float4 e[2][32]; };

// plane_row(plane, stripe, L1, C): start of the stripe's region, (L1 + 32) steps of 32*C floats
template <int C>
__device__ __forceinline__ float* stripe_plane(float* plane, int stripe, int L1) {
  return plane + (int64_t)stripe * (L1 + 32) * (32 * C);
}

__device__ __forceinline__ int col_class(const uint8_t* __restrict__ s2, int L2, int j) {
  // column 0 carries no symbol (finite class so its cells equal the reference's
  // LOG_ZERO exactly); columns past L2 are padding.
  return j == 0 ? (int)kClsOther : (j <= L2 ? (int)s2[j - 1] : (int)kClsPad);
}

// Rows of consecutive stripes follow each other through the lanes without a gap: row 0 of stripe
// k + 1 enters lane 0 one step after row L1 of stripe k (kPeriodMin keeps the hand-over of the
// boundary column, written by lane 31 and read two blocks ahead by the ring, race-free for short
// sequences: their stripes are spaced by idle rows).
#define DAFS_PAIRHMM_PERIOD_MIN 100

// ---------------------------------------------------------------------------
// Forward sweep over all stripes of one pair (ProbabilisticModel.h:105-179).
// fend receives F_k(L1, L2) in the lane that owns column L2.
template <int C>
__device__ __forceinline__ void forward_sweep(
    const PairHmmSmem& sm, PairHmmRing& ring, const uint8_t* __restrict__ s1, const uint8_t* __restrict__ s2,
    int L1, int L2, int nstripes, int lane, float* __restrict__ plane,
    float4* __restrict__ bnd0, float4* __restrict__ bnd1, float (&fend)[3]) {
  const uint32_t lut = *(volatile const uint32_t*)&sm.lut_addr[lane];
  const uint32_t match0 = (uint32_t)__cvta_generic_to_shared(sm.match);
  const float T00 = c_tab.trans[0], T01 = c_tab.trans[1], T02 = c_tab.trans[2];
  const float T10 = c_tab.trans[3], T11 = c_tab.trans[4];
  const float T20 = c_tab.trans[6], T22 = c_tab.trans[8];
  constexpr int SW = 32 * C;
  const int P = max(L1 + 1, DAFS_PAIRHMM_PERIOD_MIN);          // steps from one stripe's row 0 to the next's
  const int G = (nstripes - 1) * P + L1 + 1 + 31;              // steps of the whole sweep
  const int kNever = -0x40000000;

  // my current row: stripe k, row i (negative: not there yet)
  int k = 0, i = -lane;
  uint32_t mofs[C];                                            // shared address of match[0][class(column)][lane]
  float insc[C];                                               // log insert emission of the column
  int j0 = lane * C;
  auto load_cols = [&]() {
    j0 = k * SW + lane * C;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int cls = col_class(s2, L2, j0 + c);
      mofs[c] = match0 + (uint32_t)(cls * 32 + lane) * 4u;
      insc[c] = sm.ins[cls * 32 + lane];
    }
  };
  load_cols();
  float pM[C], pX[C], pY[C];                                   // F(i-1, j0+c)
#pragma unroll
  for (int c = 0; c < C; ++c) pM[c] = pX[c] = pY[c] = DAFS_LOG_ZERO;
  float dM = DAFS_LOG_ZERO, dX = DAFS_LOG_ZERO, dY = DAFS_LOG_ZERO;   // F(i-1, j0-1)
  float* prow = plane + lane;                                  // stripe_plane(k) + (i + lane) * SW + lane

  // ring entry of global step g = kq * P + rq
  int kq = 0, rq = lane;
  auto fetch = [&]() -> float4 {
    float4 v = make_float4(DAFS_LOG_ZERO, DAFS_LOG_ZERO, DAFS_LOG_ZERO, __int_as_float((int)kClsOther));
    if (rq <= L1 && kq < nstripes) {
      if (kq >= 1) {
        const float4 t = __ldcg(((kq - 1) & 1 ? bnd1 : bnd0) + rq);
        v.x = t.x; v.y = t.y; v.z = t.z;
      }
      if (rq >= 1) v.w = __int_as_float((int)s1[rq - 1]);
    }
    rq += 32;
    if (rq >= P) { rq -= P; ++kq; }
    return v;
  };
  float4 pre = fetch();
  int cls1 = kClsOther;
  PairHmmCoef KM[DAFS_PAIRHMM_MSETS], KX, KY;                                      // coefficient sets: calls that share one run one after the other

  auto step = [&](int g) {
    float lM = __shfl_up_sync(0xffffffffu, pM[C - 1], 1);
    float lX = __shfl_up_sync(0xffffffffu, pX[C - 1], 1);
    float lY = __shfl_up_sync(0xffffffffu, pY[C - 1], 1);
    int cu = __shfl_up_sync(0xffffffffu, cls1, 1);
    if (i == P) {                                              // on to row 0 of the next stripe -- only now: the shuffles
      ++k;                                                     // above still had to hand row L1 to the next lane
      if (k < nstripes) {
        i = 0;
        load_cols();
#pragma unroll
        for (int c = 0; c < C; ++c) pM[c] = pX[c] = pY[c] = DAFS_LOG_ZERO;
        dM = dX = dY = DAFS_LOG_ZERO;
        prow = stripe_plane<C>(plane, k, L1) + (int64_t)lane * SW + lane;
      } else {
        i = kNever;
      }
    }
    if (lane == 0) {
      const float4 b = ring.e[(g >> 5) & 1][g & 31];
      lM = b.x; lX = b.y; lY = b.z; cu = __float_as_int(b.w);
    }
    cls1 = cu;
    if ((unsigned)i <= (unsigned)L1) {
      const float insRow = sm.ins[cls1 * 32 + lane];
      const uint32_t mrow = (uint32_t)cls1 * (8 * 32 * 4);
      float leftM = lM, leftY = lY;
      float gM = dM, gX = dX, gY = dY;                         // diagonal F(i-1, j-1)
      const bool seeded = g < 2 && lane == 0 && k == 0;        // this lane's row holds cells (0,0),(0,1) or (1,0),(1,1)
#pragma unroll
      for (int c = 0; c < C; ++c) {
        // ProbabilisticModel.h:151-156
        float nM = pairhmm_lse(lut, KM[c % DAFS_PAIRHMM_MSETS], pairhmm_lse(lut, KM[c % DAFS_PAIRHMM_MSETS], __fadd_rn(gM, T00), __fadd_rn(gX, T10)), __fadd_rn(gY, T20));
        nM = __fadd_rn(nM, pairhmm_lds(mofs[c] + mrow));
        // :157-162
        float nX = __fadd_rn(insRow, pairhmm_lse(lut, KX, __fadd_rn(pM[c], T01), __fadd_rn(pX[c], T11)));
        // :163-168
        float nY = __fadd_rn(insc[c], pairhmm_lse(lut, KY, __fadd_rn(leftM, T02), __fadd_rn(leftY, T22)));
        if (c < 2 && seeded) {
          // seeds, ProbabilisticModel.h:123-131; (0,0) and the other states stay LOG_ZERO (:150)
          // (a branch taken by one lane in the first two steps of a pair; the asm keeps it a branch)
          asm volatile("");
          nM = (i == 1 && c == 1) ? __fadd_rn(c_tab.init[0], pairhmm_lds(mofs[1] + mrow)) : DAFS_LOG_ZERO;
          nX = (i == 1 && c == 0) ? __fadd_rn(c_tab.init[1], insRow) : DAFS_LOG_ZERO;
          nY = (i == 0 && c == 1) ? __fadd_rn(c_tab.init[2], insc[1]) : DAFS_LOG_ZERO;
        }
        gM = pM[c]; gX = pX[c]; gY = pY[c];
        pM[c] = nM; pX[c] = nX; pY[c] = nY;
        leftM = nM; leftY = nY;
        prow[c * 32] = nM;
      }
      dM = lM; dX = lX; dY = lY;
      if (lane == 31 && k + 1 < nstripes) ((k & 1) ? bnd1 : bnd0)[i] = make_float4(pM[C - 1], pX[C - 1], pY[C - 1], 0.f);
      if (i == L1 && k == nstripes - 1) {                      // once per lane and pair
        asm volatile("");
#pragma unroll
        for (int c = 0; c < C; ++c)
          if (j0 + c == L2) { fend[0] = pM[c]; fend[1] = pX[c]; fend[2] = pY[c]; }
      }
    }
    ++i; prow += SW;
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
