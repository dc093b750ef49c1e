// Device arithmetic of the ProbCons pair-HMM, bit-compatible with the
// reference's x86-64 build: every FP32 multiply and add is rounded on its own
// (__fmul_rn/__fadd_rn are never contracted into FMA), polynomial pieces and
// evaluation order follow probconsRNA/ScoreType.h exactly.
#pragma once
#include <cuda_runtime.h>

namespace dafs {

// probconsRNA/ScoreType.h:18 -- "log 0".  ulp(2e20) ~ 1.8e13, so LOG_ZERO plus
// any log-probability is LOG_ZERO again.
#define DAFS_LOG_ZERO (-2e20f)

// LOOKUP(d) = log(1 + e^-d)-ish on [0, 7.5): four cubic pieces, Horner form,
// float literals (ScoreType.h:187-198).  The piece boundaries 1.0 / 2.5 / 4.5 are
// multiples of one half and the reference's tests are strict (d > 1.0f ...), so
// the piece is a function of n = ceil(2 d): 0..2 -> piece 0, 3..5 -> 1, 6..9 -> 2,
// 10.. -> 3.  n comes out of ONE instruction: fma(d, 2, 1.5 * 2^23) rounded
// towards +inf leaves ceil(2 d) in the low mantissa bits (2 d is exact, the sum is
// rounded once).  The coefficient rows {c3,c2,c1,c0} live in shared memory as a
// 17-entry float4 table indexed by n, fetched with one LDS.128 -- this replaces
// three compares and three selects per LOOKUP and moves the work from the ALU
// pipe to the FMA pipe.  d is clamped to 8 first (n <= 16); for d >= 7.5 the
// caller discards the value anyway.
#define DAFS_LOOKUP_ROWS                                                                   \
  {-0.009350833524763f, 0.130659527668286f, 0.498799810682272f, 0.693203116424741f},      \
  {-0.014532321752540f, 0.139942324101744f, 0.495635523139337f, 0.692140569840976f},      \
  {-0.004605031767994f, 0.063427417320019f, 0.695956496475118f, 0.514272634594009f},      \
  {-0.000458661602210f, 0.009695946122598f, 0.930734667215156f, 0.168037164329057f}
#define DAFS_LOOKUP_SLOTS 17
__host__ __device__ __forceinline__ int pairhmm_lookup_piece(int n) { return n >= 10 ? 3 : n >= 6 ? 2 : n >= 3 ? 1 : 0; }

__device__ __forceinline__ float pairhmm_lookup(const float4* __restrict__ tab, float d) {
  const float t = __fmaf_ru(fminf(d, 8.0f), 2.0f, 12582912.0f);
  const float4 k = tab[__float_as_int(t) - 0x4b400000];
  float r = __fadd_rn(__fmul_rn(k.x, d), k.y);
  r = __fadd_rn(__fmul_rn(r, d), k.z);
  r = __fadd_rn(__fmul_rn(r, d), k.w);
  return r;
}

// LOG_ADD / LOG_PLUS_EQUALS (ScoreType.h:233-238, :259-262), branch-free.
// The reference's "== LOG_ZERO" test is numerically redundant: with lo ==
// LOG_ZERO either hi is finite (then hi-lo >= 7.5 picks hi, the same answer)
// or hi == LOG_ZERO too (then LOOKUP(0)+LOG_ZERO rounds to LOG_ZERO, the same
// answer).  The same holds for the even more negative values that padded DP
// cells carry, so "zero" never leaks into a finite cell.
__device__ __forceinline__ float pairhmm_lse(const float4* __restrict__ tab, float x, float y) {
  const float hi = fmaxf(x, y), lo = fminf(x, y);
  const float d = __fsub_rn(hi, lo);
  const float r = __fadd_rn(pairhmm_lookup(tab, d), lo);
  return d >= 7.5f ? hi : r;
}

// EXP (ScoreType.h:37-57) for x <= 0: quartic pieces with double literals, so
// the reference evaluates in double and rounds once to float on return.
__device__ __forceinline__ float pairhmm_exp(float xf) {
  const double x = (double)xf;
  double a, b, c, d, e;
  if (x > -2.0) {
    if (x > -0.5) { a = 0.03254409303190190000; b = 0.16280432765779600000; c = 0.49929760485974900000; d = 0.99995149601363700000; e = 0.99999925508501600000; }
    else if (x > -1.0) { a = 0.01973899026052090000; b = 0.13822379685007000000; c = 0.48056651562365000000; d = 0.99326940370383500000; e = 0.99906756856399500000; }
    else { a = 0.00940528203591384000; b = 0.09414963667859410000; c = 0.40825793595877300000; d = 0.93933625499130400000; e = 0.98369508190545300000; }
  } else if (x > -8.0) {
    if (x > -4.0) { a = 0.00217245711583303000; b = 0.03484829428350620000; c = 0.22118199801337800000; d = 0.67049462206469500000; e = 0.83556950223398500000; }
    else { a = 0.00012398771025456900; b = 0.00349155785951272000; c = 0.03727721426017900000; d = 0.17974997741536900000; e = 0.33249299994217400000; }
  } else if (x > -16.0) {
    a = 0.00000051741713416603; b = 0.00002721456879608080; c = 0.00053418601865636800; d = 0.00464101989351936000; e = 0.01507447981459420000;
  } else {
    return 0.0f;
  }
  double r = __dadd_rn(__dmul_rn(a, x), b);
  r = __dadd_rn(__dmul_rn(r, x), c);
  r = __dadd_rn(__dmul_rn(r, x), d);
  r = __dadd_rn(__dmul_rn(r, x), e);
  return __double2float_rn(r);
}

}  // namespace dafs
