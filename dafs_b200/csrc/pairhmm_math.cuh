// Device arithmetic of the ProbCons pair-HMM, bit-compatible with the
// reference's x86-64 build: every FP32 multiply and add is rounded on its own
// (__fmul_rn/__fadd_rn are never contracted into FMA), polynomial pieces and
// evaluation order follow probconsRNA/ScoreType.h exactly.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

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
// pipe to the FMA pipe.  For d >= 7.5 the value is not needed and the fetch is skipped.
// The shared-memory data pipe is the unit this kernel loads most (an LDS.128 is four
// quarter-warp wavefronts of 128 bytes): every row is stored EIGHT times, copy (lane & 7) of row n
// at float4 index 8 n + (lane & 7), so the eight lanes of a wavefront always read eight different
// 16-byte bank groups whatever rows they want -- no bank conflict is possible.
#define DAFS_LOOKUP_ROWS                                                                   \
  {-0.009350833524763f, 0.130659527668286f, 0.498799810682272f, 0.693203116424741f},      \
  {-0.014532321752540f, 0.139942324101744f, 0.495635523139337f, 0.692140569840976f},      \
  {-0.004605031767994f, 0.063427417320019f, 0.695956496475118f, 0.514272634594009f},      \
  {-0.000458661602210f, 0.009695946122598f, 0.930734667215156f, 0.168037164329057f}
#define DAFS_LOOKUP_SLOTS 17
__host__ __device__ __forceinline__ int pairhmm_lookup_piece(int n) { return n >= 10 ? 3 : n >= 6 ? 2 : n >= 3 ? 1 : 0; }

#define DAFS_LOOKUP_COPIES 8
// Shared-window address of this lane's copy of row 0 plus the constant that cancels the exponent bits
// of the magic sum below: bits(t) = 0x4b400000 + n, (bits << 7) mod 2^32 = 0xa0000000 + 128 n.
// The sum is kept in shared memory and read back (PairHmmSmem::lut_addr[lane]): ptxas folds a value it can
// trace to the cvta into "base register + immediate" again and spends a second integer add per call.
__device__ __forceinline__ uint32_t pairhmm_lut_addr(const float4* tab, int lane) {
  return (uint32_t)__cvta_generic_to_shared(tab) + (uint32_t)(lane & 7) * 16u + 0x60000000u;
}

__device__ __forceinline__ float pairhmm_lds(uint32_t addr) {
  float v;
  asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}

// LOG_ADD / LOG_PLUS_EQUALS (ScoreType.h:233-238, :259-262), branch-free.
// The reference's "== LOG_ZERO" test is numerically redundant: with lo ==
// LOG_ZERO either hi is finite (then hi-lo >= 7.5 picks hi, the same answer)
// or hi == LOG_ZERO too (then LOOKUP(0)+LOG_ZERO rounds to LOG_ZERO, the same
// answer).  The same holds for the even more negative values that padded DP
// cells carry, so "zero" never leaks into a finite cell.
// 14 instructions: max, min, sub, setp, fma.rp, lea, lds (skipped, like the final add, when
// d >= 7.5: no clamp of the index is needed), 3 x (mul, add), add.  Every mul / add carries .rn,
// which PTX never contracts.  The four coefficient registers are the caller's (PairHmmCoef, set
// to anything once): a predicated load leaves them partly defined, and registers that every call
// site declared for itself would be live across the whole DP loop.  Calls that share a set run
// one after the other, so a caller keeps one set per independent chain.
#ifndef DAFS_PAIRHMM_MSETS
#define DAFS_PAIRHMM_MSETS 2     // sets for the M state (two calls per cell): cells alternate
#endif
struct PairHmmCoef { float k0 = 0.f, k1 = 0.f, k2 = 0.f, k3 = 0.f; };

__device__ __forceinline__ float pairhmm_lse(uint32_t lut, PairHmmCoef& K, float x, float y) {
  float r;
  asm("{\n\t"
      ".reg .pred p;\n\t"
      ".reg .f32 hi, lo, d, t, q;\n\t"
      ".reg .b32 a;\n\t"
      "max.f32 hi, %5, %6;\n\t"
      "min.f32 lo, %5, %6;\n\t"
      "sub.rn.f32 d, hi, lo;\n\t"
      "setp.ge.f32 p, d, 0f40F00000;\n\t"             // 7.5
      "fma.rp.f32 t, d, 0f40000000, 0f4B400000;\n\t"  // ceil(2 d) in the low mantissa bits
      "mov.b32 a, t;\n\t"
      "shl.b32 a, a, 7;\n\t"
      "add.u32 a, a, %7;\n\t"
      "@!p ld.shared.v4.f32 {%1, %2, %3, %4}, [a];\n\t"
#ifdef DAFS_PAIRHMM_FMA   /* A/B build only (tools/ab_pairhmm_fma.sh): fused polynomial, NOT bit-exact with the reference */
      "fma.rn.f32 q, %1, d, %2;\n\t"
      "fma.rn.f32 q, q, d, %3;\n\t"
      "fma.rn.f32 q, q, d, %4;\n\t"
#else
      "mul.rn.f32 q, %1, d;\n\t"
      "add.rn.f32 q, q, %2;\n\t"
      "mul.rn.f32 q, q, d;\n\t"
      "add.rn.f32 q, q, %3;\n\t"
      "mul.rn.f32 q, q, d;\n\t"
      "add.rn.f32 q, q, %4;\n\t"
#endif
      "mov.f32 %0, hi;\n\t"
      "@!p add.rn.f32 %0, q, lo;\n\t"
      "}"
      : "=&f"(r), "+f"(K.k0), "+f"(K.k1), "+f"(K.k2), "+f"(K.k3)
      : "f"(x), "f"(y), "r"(lut));
  return r;
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
