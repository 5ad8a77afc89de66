// Rerank: exact fp32 direct-difference distances of the screen's survivors, one warp per query
// row (included by knn_screen.cu).  Restores the reference's formula for the rows that matter:
// sqrt(sum_c (a_c - b_c)^2), utils/utils.py:293-296, then the k smallest with (distance, index)
// order, :297-300.
//
//   LAT = false  dataset-scale path: scalars (EB, HBmax, range flags) come from PrepScalars,
//                undecided rows are appended to a device list for the exact kernel.
//   LAT = true   per-update path: scalars come from the two floats the latency screen kernel
//                wrote, undecided rows (rare) are recomputed inline against the whole bank, so the
//                chain has no fallback launches; optional fused PBE epilogue by the last block.
//   VEC          dim % 4 == 0 and 16-byte aligned rows: float4 loads, 8 survivors per batch.
#pragma once

namespace exb {

struct RerankStatus {   // device, zeroed per call
  int n_fallback;
  int invalid;
  unsigned long long pairs;
};

constexpr int RR_PER_LANE = 12;   // 32 * 12 = 384 >= 16 splits * 24 slots
constexpr int RR_SURV = 64;       // survivors recomputed exactly per row (else undecided)
constexpr int RR_WARPS = 4;

struct RerankParams {
  const float* Q;
  const float* B;
  int64_t nq, nb;
  int dim, k, idx_base, splits;
  const float* cand_val;
  const int* cand_idx;
  const int* cand_cnt;
  const float* ha;
  const float* ea;
  const PrepScalars* qsc;    // !LAT
  const PrepScalars* bsc;    // !LAT
  const float* bank_max;     // LAT: {EB, HBmax}
  float* out_dist;
  int* out_idx;
  int* fallback_rows;        // !LAT
  int fallback_cap;
  RerankStatus* status;
  long long* trace;
};

// Warp-resident ascending top-k of (value, index) pairs: lane r holds the r-th smallest.
struct WarpTopK {
  float v;
  int i;
  __device__ __forceinline__ void init() { v = FLT_MAX; i = INT_MAX; }
  __device__ __forceinline__ void insert(float nv, int ni, int k, int lane) {
    const unsigned full = 0xffffffffu;
    const float thr_v = __shfl_sync(full, v, k - 1);
    const int thr_i = __shfl_sync(full, i, k - 1);
    if (pair_less(nv, ni, thr_v, thr_i)) {
      const int pos = __popc(__ballot_sync(full, pair_less(v, i, nv, ni)));
      const float up_v = __shfl_up_sync(full, v, 1);
      const int up_i = __shfl_up_sync(full, i, 1);
      if (lane == pos) { v = nv; i = ni; }
      else if (lane > pos) { v = up_v; i = up_i; }
    }
  }
};

// Exact squared distances of `count` bank rows (ids from idx_of(c)) to the query held in
// registers, inserted into `top`.  NB rows per batch so that NB x (dim/32) loads are in flight.
template <bool VEC, typename IdxOf>
__device__ __forceinline__ void exact_scan(const float (&q)[16], const float* __restrict__ B,
                                           int dim, int k, int lane, int count, IdxOf idx_of,
                                           WarpTopK& top) {
  const unsigned full = 0xffffffffu;
  constexpr int NB = VEC ? 8 : 4;
  for (int c0 = 0; c0 < count; c0 += NB) {
    int j[NB];
    float a[NB];
#pragma unroll
    for (int u = 0; u < NB; ++u) {
      j[u] = idx_of(min(c0 + u, count - 1));
      a[u] = 0.f;
    }
    if constexpr (VEC) {
      const int nv = dim >> 2;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int c = lane + 32 * i;
        if (c < nv) {
          float4 b[NB];
#pragma unroll
          for (int u = 0; u < NB; ++u)
            b[u] = __ldg(reinterpret_cast<const float4*>(B + (int64_t)j[u] * dim) + c);
#pragma unroll
          for (int u = 0; u < NB; ++u) {
            float d = q[4 * i] - b[u].x;
            a[u] = fmaf(d, d, a[u]);
            d = q[4 * i + 1] - b[u].y;
            a[u] = fmaf(d, d, a[u]);
            d = q[4 * i + 2] - b[u].z;
            a[u] = fmaf(d, d, a[u]);
            d = q[4 * i + 3] - b[u].w;
            a[u] = fmaf(d, d, a[u]);
          }
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int c = lane + 32 * i;
        if (c < dim) {
#pragma unroll
          for (int u = 0; u < NB; ++u) {
            const float d = q[i] - __ldg(B + (int64_t)j[u] * dim + c);
            a[u] = fmaf(d, d, a[u]);
          }
        }
      }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
#pragma unroll
      for (int u = 0; u < NB; ++u) a[u] += __shfl_xor_sync(full, a[u], off);
#pragma unroll
    for (int u = 0; u < NB; ++u)
      if (c0 + u < count) top.insert(a[u], j[u], k, lane);
  }
}

template <bool LAT, bool VEC>
__global__ void __launch_bounds__(RR_WARPS * 32)
knn_rerank_kernel(const RerankParams p) {
  __shared__ int s_surv[RR_WARPS][RR_SURV];
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t row = (int64_t)blockIdx.x * RR_WARPS + wib;
  if (row >= p.nq) return;
  const bool tr = p.trace != nullptr && row == 0;
  if (tr && lane == 0) p.trace[8] = clock_now();
  const int k = p.k, splits = p.splits, dim = p.dim;

  // query row into registers first: its latency hides behind the list gathering
  float q[16];
  if constexpr (VEC) {
    const float4* q4 = reinterpret_cast<const float4*>(p.Q + row * dim);
    const int nv = dim >> 2;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = lane + 32 * i;
      const float4 v = (c < nv) ? __ldg(q4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      q[4 * i] = v.x; q[4 * i + 1] = v.y; q[4 * i + 2] = v.z; q[4 * i + 3] = v.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int c = lane + 32 * i;
      q[i] = (c < dim) ? __ldg(p.Q + row * dim + c) : 0.f;
    }
  }

  float eb_max, hb_max;
  bool undecided;
  if constexpr (LAT) {
    eb_max = p.bank_max[0];
    hb_max = p.bank_max[1];
    undecided = !(eb_max < INFINITY);     // a bank row outside the fp16 range
  } else {
    eb_max = p.bsc->err_max;
    hb_max = p.bsc->hn_max;
    undecided = (p.qsc->range_flag | p.bsc->range_flag) != 0;
  }

  // ---- per-split counts: lane sp holds the count of split sp (splits <= 16)
  const int my_cnt = (lane < splits) ? p.cand_cnt[row * splits + lane] : 0;
  if (__any_sync(full, my_cnt < 0)) undecided = true;
  int total = max(my_cnt, 0);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) total += __shfl_xor_sync(full, total, off);
  if (total < k) undecided = true;

  // ---- gather the per-split lists of this row into registers (entry e = lane + 32 i)
  float ev[RR_PER_LANE];
  int ei[RR_PER_LANE];
#pragma unroll
  for (int i = 0; i < RR_PER_LANE; ++i) {
    const int e = lane + 32 * i;
    const int sp = e / SCR_CAP, c = e - sp * SCR_CAP;
    const int cnt = __shfl_sync(full, my_cnt, sp & 31);
    ev[i] = -FLT_MAX;
    ei[i] = -1;
    if (sp < splits && c < cnt) {
      ev[i] = p.cand_val[(row * splits + sp) * SCR_CAP + c];
      ei[i] = p.cand_idx[(row * splits + sp) * SCR_CAP + c];
    }
  }
  if (tr && lane == 0) p.trace[9] = clock_now();

  int n_surv = 0;
  const float ea_row = p.ea[row];
  if (!undecided && !(ea_row < INFINITY)) undecided = true;   // query row outside the fp16 range
  if (!undecided) {
    // ---- k-th best screened score over the union of splits (k rounds of warp arg-max)
    float wv[RR_PER_LANE];
#pragma unroll
    for (int i = 0; i < RR_PER_LANE; ++i) wv[i] = ev[i];
    float kth = -FLT_MAX;
    for (int round = 0; round < k; ++round) {
      float lm = wv[0];
      int ls = 0;
#pragma unroll
      for (int i = 1; i < RR_PER_LANE; ++i)
        if (wv[i] > lm) { lm = wv[i]; ls = i; }
      float wm = lm;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) wm = fmaxf(wm, __shfl_xor_sync(full, wm, off));
      const int owner = __ffs(__ballot_sync(full, lm == wm)) - 1;
      if (lane == owner) {
#pragma unroll
        for (int i = 0; i < RR_PER_LANE; ++i)
          if (i == ls) wv[i] = -FLT_MAX;
      }
      kth = wm;
    }
    // same keep rule as the screening kernel, now with the global k-th best
    ListState st;
    st.ha = p.ha[row];
    st.e2 = 2.f * (ea_row + eb_max);
    st.dl = (st.ha + hb_max) * SCR_ACC_EPS;
    st.tk = kth;
    const float tkeep = keep_threshold(st);
#pragma unroll
    for (int i = 0; i < RR_PER_LANE; ++i) {
      const bool keep = ei[i] >= 0 && ev[i] >= tkeep;
      const unsigned m = __ballot_sync(full, keep);
      const int pos = n_surv + __popc(m & ((1u << lane) - 1u));
      if (keep && pos < RR_SURV) s_surv[wib][pos] = ei[i];
      n_surv += __popc(m);
    }
    if (n_surv > RR_SURV) undecided = true;
    __syncwarp();
  }
  if (tr && lane == 0) p.trace[10] = clock_now();

  WarpTopK top;
  top.init();
  if (undecided) {
    if constexpr (LAT) {
      // rare: recompute this row against the whole bank here (no fallback launches in the chain)
      if (lane == 0) atomicAdd(&p.status->n_fallback, 1);
      exact_scan<VEC>(q, p.B, dim, k, lane, (int)p.nb, [](int c) { return c; }, top);
    } else {
      if (lane == 0) {
        const int pos = atomicAdd(&p.status->n_fallback, 1);
        if (pos < p.fallback_cap) p.fallback_rows[pos] = (int)row;
        else p.status->invalid = 1;
      }
      return;
    }
  } else {
    const int* surv = s_surv[wib];
    exact_scan<VEC>(q, p.B, dim, k, lane, n_surv, [surv](int c) { return surv[c]; }, top);
    if (lane == 0) atomicAdd(&p.status->pairs, (unsigned long long)n_surv);
  }
  if (lane < k) {
    p.out_dist[row * k + lane] = sqrtf(top.v);
    p.out_idx[row * k + lane] = top.i + p.idx_base;
  }
  if (tr && lane == 0) p.trace[11] = clock_now();
}

}  // namespace exb
