// Prepared kNN operand of warp-resident rows: fp16 rounding, half squared norm of the rounded row,
// rounded-up norm of the rounding error, range flag.  ONE implementation shared by knn_prep_kernel
// (knn_screen.cu, R = 1) and the rep producers (rep_producer.cu, R rows per warp), so that a
// producer's operand is bit-identical to exb_knn_prep_f16 of the rep it wrote.
//
// Arithmetic (DESIGN.md §4 relies on it): hf = fp16_rn(f) ; hf*hf is exact in fp32 (11-bit
// significands) ; the norm sum is accumulated round-to-nearest in two interleaved chains per lane
// (packed FFMA2), <= 8 + 1 + 5 roundings ; f - hf is exact ; the error sum is rounded UP at every
// step (FFMA2.RU, FADD.RU) so round_err stays a rigorous upper bound of ||x - fp16(x)||_2.
#pragma once
#include "common.cuh"
#include <cuda_fp16.h>
#include <math.h>

namespace exb {

struct OpOut {
  __half* f16;        // [n, dpad] or null
  float* hn;          // [n]
  float* er;          // [n]
  PrepScalars* sc;    // max-accumulated over rows (and calls); may be null
  int dpad;
  // The rows are s * fl(x - mu) of an affine-prepared operand (PrepAffine): fl(x - mu) carries a
  // relative rounding error <= 2^-24 per element against the exact difference, so
  // || s (x - mu) - v ||_2 <= 2^-24 ||v|| <= 2^-24 (||v^|| + round_err); extra_rel = 2^-22 (slack
  // for the round-to-nearest norm sum) adds that to round_err.  0 = plain rows (no term added:
  // the producers' operands stay bit-identical to exb_knn_prep_f16).
  float extra_rel = 0.f;
};

// round_err of an affine-prepared row (see OpOut::extra_rel); nn = ||v^||^2, every step rounded up
__device__ __forceinline__ float affine_round_err(float er, float nn, float extra_rel) {
  return __fadd_ru(er, __fmul_ru(extra_rel, __fadd_ru(__fsqrt_ru(nn), er)));
}

// v[r][i] = float4 chunk c = lane + 32 i of row row0 + r (zero for c >= dim/4).  Rows r >= rows
// are computed but not stored.  FULL: dim == dpad == NCH * 128 (no column guards).
template <int R, int NCH, bool FULL>
__device__ __forceinline__ void emit_rows(const float4 (&v)[R][NCH], int lane, int dim,
                                          int64_t row0, int rows, float* __restrict__ out_rep,
                                          const OpOut& op) {
  const int nv = dim >> 2;
  if (out_rep != nullptr) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float4* o4 = reinterpret_cast<float4*>(out_rep + (row0 + r) * dim);
#pragma unroll
      for (int i = 0; i < NCH; ++i) {
        const int c = lane + 32 * i;
        if (r < rows && (FULL || c < nv)) o4[c] = v[r][i];
      }
    }
  }
  if (op.f16 == nullptr) return;
  const int nvp = op.dpad >> 2;
  float nn[R], ee[R];
  uint32_t bits[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    uint2* o2 = reinterpret_cast<uint2*>(op.f16 + (row0 + r) * op.dpad);
    float2 n2 = make_float2(0.f, 0.f), e2 = make_float2(0.f, 0.f);
    uint32_t b = 0;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const int c = lane + 32 * i;
      if (FULL || c < nvp) {
        const float4 f = v[r][i];
        const __half2 h01 = __floats2half2_rn(f.x, f.y);
        const __half2 h23 = __floats2half2_rn(f.z, f.w);
        const float2 g01 = __half22float2(h01), g23 = __half22float2(h23);
        n2 = __ffma2_rn(g01, g01, n2);
        n2 = __ffma2_rn(g23, g23, n2);
        const float2 d01 = __fadd2_rn(make_float2(f.x, f.y), make_float2(-g01.x, -g01.y));
        const float2 d23 = __fadd2_rn(make_float2(f.z, f.w), make_float2(-g23.x, -g23.y));
        e2 = __ffma2_ru(d01, d01, e2);
        e2 = __ffma2_ru(d23, d23, e2);
        uint2 pk;
        pk.x = *reinterpret_cast<const uint32_t*>(&h01);
        pk.y = *reinterpret_cast<const uint32_t*>(&h23);
        // bit 15 of a half-word is set iff its magnitude is >= 0x7c00 (inf / NaN after rounding)
        b |= ((pk.x & 0x7fff7fffu) + 0x04000400u) | ((pk.y & 0x7fff7fffu) + 0x04000400u);
        if (r < rows) o2[c] = pk;
      }
    }
    nn[r] = n2.x + n2.y;
    ee[r] = __fadd_ru(e2.x, e2.y);
    bits[r] = b;
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      nn[r] += __shfl_xor_sync(0xffffffffu, nn[r], off);
      ee[r] = __fadd_ru(ee[r], __shfl_xor_sync(0xffffffffu, ee[r], off));
    }
  }
  // lane r finishes row r
  float hn = 0.f, er = 0.f;
  bool any_bad = false;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const bool bad = __any_sync(0xffffffffu, (bits[r] & 0x80008000u) != 0);
    if (lane == r) {
      hn = bad ? 0.f : 0.5f * nn[r];
      er = bad ? INFINITY : __fsqrt_ru(ee[r]);   // rounded up: the bound stays rigorous
      if (op.extra_rel != 0.f && !bad) er = affine_round_err(er, nn[r], op.extra_rel);
    }
    any_bad |= bad && r < rows;
  }
  const bool mine = lane < rows && lane < R;
  if (mine) {
    op.hn[row0 + lane] = hn;
    op.er[row0 + lane] = er;
  }
  if (op.sc) {
    // non-negative floats order like their bit patterns; rows not stored contribute 0
    const int er_b = __reduce_max_sync(0xffffffffu, mine ? __float_as_int(er) : 0);
    const int hn_b = __reduce_max_sync(0xffffffffu, mine ? __float_as_int(hn) : 0);
    if (lane == 0) {
      // the maxima settle after a few rows: look first, RED only when this warp raises them
      // (10^6 same-address REDs per call were the kernel's bottleneck)
      int* pe = reinterpret_cast<int*>(&op.sc->err_max);
      int* ph = reinterpret_cast<int*>(&op.sc->hn_max);
      if (er_b > *reinterpret_cast<volatile int*>(pe)) atomicMax(pe, er_b);
      if (hn_b > *reinterpret_cast<volatile int*>(ph)) atomicMax(ph, hn_b);
      if (any_bad) atomicExch(&op.sc->range_flag, 1);
    }
  }
}

}  // namespace exb
