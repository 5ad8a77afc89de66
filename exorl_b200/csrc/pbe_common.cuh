// RMS update + PBE per-row epilogue shared by pbe_epilogue.cu and the fused per-update chain
// (knn_rerank.cuh).  Follows /root/reference/utils/utils.py:264-276 and :301-318.
#pragma once
#include "common.cuh"
#include <math.h>

namespace exb {

constexpr int PBE_THREADS = 256;

// The running-moment state of the reference's RMS object, resident on the device and updated in
// place: M and S are the fp32 (1,) tensors RMS.M / RMS.S, n is RMS.n as one fp64 word.
struct RmsParams {
  float* m;
  float* s;
  double* n;
};

// values entering the moments: avg -> all nq*k, else column k-1 of each row
__device__ __forceinline__ int64_t value_index(int64_t i, int k, int avg) {
  return avg ? i : i * k + (k - 1);
}

__device__ __forceinline__ double2 warp_sum2(double2 v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    v.x += __shfl_xor_sync(0xffffffffu, v.x, off);
    v.y += __shfl_xor_sync(0xffffffffu, v.y, off);
  }
  return v;
}

// Block-wide (sum, sumsq); result valid in thread 0.
template <int THREADS>
__device__ __forceinline__ double2 block_sum2(double2 v, double2* smem) {
  v = warp_sum2(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  if (warp == 0) {
    double2 t = (lane < THREADS / 32) ? smem[lane] : make_double2(0.0, 0.0);
    t = warp_sum2(t);
    if (lane == 0) smem[0] = t;
  }
  __syncthreads();
  return smem[0];
}

// utils/utils.py:264-276 on scalars.  sums = {count, sum, sumsq}.
__device__ __forceinline__ void rms_apply(const double* sums, const RmsParams& p) {
  const double bs = sums[0], sum = sums[1], sumsq = sums[2];
  const double M = (double)*p.m, S = (double)*p.s, n = *p.n;
  const double mean = sum / bs;
  // torch.var(unbiased): NaN for a single element, like the reference.
  const double var = (bs > 1.0) ? fmax(sumsq - sum * mean, 0.0) / (bs - 1.0) : (double)NAN;
  const double tot = n + bs;
  const double delta = mean - M;
  const double new_m = M + delta * bs / tot;
  const double new_s = (S * n + var * bs + delta * delta * n * bs / tot) / tot;
  *p.m = (float)new_m;
  *p.s = (float)new_s;
  *p.n = tot;       // utils/utils.py:274
}

// torch.maximum(t, 0) propagates NaN (0/0 when every distance is 0, e.g. k = 1 on source =
// target); fmaxf would not.
__device__ __forceinline__ float clip_nan(float t) { return (t < 0.f) ? 0.f : t; }

// r / M (true division, utils.py:304/311), max(r - clip, 0) (:305-308/:312-315), mean over k
// (:317), log(r + 1) (:318) on k values in registers (k <= 32).
__device__ __forceinline__ float pbe_row_regs(const float (&d)[32], int k, int avg, float inv_on,
                                              float m, int use_clip, float clip) {
  float acc = 0.f, last = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    if (j < k) {
      float v = d[j];
      if (inv_on != 0.f) v = v / m;
      if (use_clip) v = clip_nan(v - clip);
      acc += v;
      last = v;
    }
  }
  return logf((avg ? acc / (float)k : last) + 1.0f);
}

// Row in memory: all k loads are issued together (L2 reads: the fused chain reads rows that other
// SMs just wrote).
__device__ __forceinline__ float pbe_row(const float* __restrict__ d, int k, int avg, float inv_on,
                                         float m, int use_clip, float clip) {
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = (j < k) ? __ldcg(d + j) : 0.f;
  return pbe_row_regs(v, k, avg, inv_on, m, use_clip, clip);
}

}  // namespace exb
