// Symmetric screen (source == bank, dataset scale): kernels around knn_screen_kernel<CG, 4>
// (included by knn_screen.cu after knn_rerank.cuh).
//
// PBE.__call__ and the relabel score a set against ITSELF (utils/utils.py:290: source = target =
// rep): the tensor-core product a^_i . a^_j serves both the pair (query i, bank j) and the pair
// (query j, bank i), so only the tile pairs I <= J need an MMA — half the flops of the full
// screen, which already runs at the tensor roofline.  What makes that possible is a keep
// threshold per row that is FIXED before the pass, so that the kernel can test an element for a
// row it does not own:
//   1. every row's exact distance to its k-th nearest neighbour inside a strided SAMPLE of the
//      bank (an ordinary screened call, n x S) bounds its true k-th-neighbour distance from above;
//   2. sym_threshold_kernel turns it into a score threshold: any bank row j among the true k
//      nearest of i has ||a - b_j|| <= d_S, hence ||a^ - b^_j|| <= R' = s d_S + ea + EB in the
//      operand space (scale s of the affine record), hence a computed score
//      >= ha - R'^2 / 2 - dl  (DESIGN.md §4 with the distance bound in place of the chain);
//   3. knn_screen_kernel<CG, 4> logs, per row, the direct hits of its own scan (tiles J >= I) and
//      the transposed hits it finds for the column rows;
//   4. sym_scatter_kernel regroups the transposed hits by the row they belong to;
//   5. knn_rerank_sym_kernel: one warp per row merges its direct logs and its received hits,
//      finds the k-th best score, applies the usual keep rule and recomputes the survivors in
//      fp32 (exact_scan) — from there on identical to knn_rerank_kernel.
// A direct log or a received list that overflows makes its row undecided (exact kernel); the
// transposed hits go through one append buffer sized from the expected total (hub rows are
// candidates of thousands of rows, so no per-row log could be sized) — if even that overflows,
// entries of unknown rows were lost and the pass is invalid (status->invalid |= 2: the host reruns
// the call with the full screen).
#pragma once

namespace exb {

constexpr int SYM_INVALID = 2;

struct SymThresholdParams {
  const float* d_sample;   // [n] exact distance to the k-th nearest SAMPLE row (stride floats apart)
  int64_t stride;
  const float* ha;         // [n]
  const float* ea;         // [n]
  const PrepScalars* bsc;
  const float* aff;        // affine record (scale) or null
  float* tkeep;            // [n] out
  int64_t n;
};

__global__ void sym_threshold_kernel(const SymThresholdParams p) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.n) return;
  const float scale = (p.aff != nullptr && p.aff[AFF_ON] != 0.f) ? p.aff[AFF_SCALE] : 1.f;
  const float ha = p.ha[i], ea = p.ea[i];
  const float eb_max = p.bsc->err_max, hb_max = p.bsc->hn_max;
  // the sample distance is an fp32 result (<= 1e-6 relative from the exact value): inflate it
  const float r = __fmul_ru(__fmul_ru(p.d_sample[i * p.stride], scale), 1.00001f);
  const float bound = __fadd_ru(r, __fadd_ru(ea, eb_max));
  const float dl = (ha + hb_max) * SCR_ACC_EPS;
  float t = ha - 0.5f * bound * bound * 1.000001f - dl;
  t -= 4e-6f * (fabsf(ha) + 0.5f * bound * bound);
  // rows the operand cannot represent (round_err = inf) or a failed sample: keep everything
  p.tkeep[i] = (t == t && bound < INFINITY) ? t : -FLT_MAX;
}

struct SymLists {
  const float* dlog_val;   // [n, 2, log_cap]
  const int* dlog_idx;
  const int* dlog_cnt;     // [n, 2]
  int log_cap;
  const int* tq_target;    // append buffer of transposed hits (query row j, score, bank row i)
  const float* tq_val;
  const int* tq_src;
  const unsigned long long* tq_count;
  long long tq_cap;
  float* rval;             // [n, rcap] received (transposed) hits per row
  int* ridx;
  int* rcnt;               // [n], zeroed per call
  int rcap;
};

// One thread per transposed hit: entry -> slot atomicAdd(rcnt[target]) of its row.  n_hits < 0:
// the entries are the workspace's own append buffer (count on the device); else `n_hits` entries
// of the given arrays (hits received from other ranks).
__global__ void __launch_bounds__(256)
sym_scatter_kernel(const SymLists l, const int* __restrict__ target, const float* __restrict__ val,
                   const int* __restrict__ src, long long n_hits, RerankStatus* status) {
  long long n_ent = n_hits;
  if (n_hits < 0) {
    const unsigned long long total = *l.tq_count;
    if (blockIdx.x == 0 && threadIdx.x == 0 && total > (unsigned long long)l.tq_cap)
      atomicOr(&status->invalid, SYM_INVALID);    // hits were dropped: nobody can be certified
    n_ent = (long long)min(total, (unsigned long long)l.tq_cap);
  }
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n_ent;
       e += (long long)gridDim.x * blockDim.x) {
    const int t = target[e];
    const int pos = atomicAdd(l.rcnt + t, 1);
    if (pos < l.rcap) {
      l.rval[(int64_t)t * l.rcap + pos] = val[e];
      l.ridx[(int64_t)t * l.rcap + pos] = src[e];
    }
  }
}

// Debug (EXB_SYM_STATS): sums and maxima of the log / list counts into the trace words
// [0] sum direct, [1] max direct (per half), [2] transposed hits appended, [3] append capacity,
// [4] sum received, [5] max received, [6] overflowed direct logs
__global__ void sym_stats_kernel(const SymLists l, int64_t n, long long* out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 2 * n) return;
  const int d = l.dlog_cnt[i];
  unsigned long long* o = reinterpret_cast<unsigned long long*>(out);
  atomicAdd(o + 0, (unsigned long long)max(d, 0));
  atomicMax(o + 1, (unsigned long long)max(d, 0));
  if (d < 0) atomicAdd(o + 6, 1ull);
  if (i == 0) { o[2] = *l.tq_count; o[3] = (unsigned long long)l.tq_cap; }
  if (i < n) {
    atomicAdd(o + 4, (unsigned long long)l.rcnt[i]);
    atomicMax(o + 5, (unsigned long long)l.rcnt[i]);
  }
}

// Warp-resident DESCENDING top-k of scores: lane r holds the r-th largest.
struct WarpTopScores {
  float v;
  __device__ __forceinline__ void init() { v = -FLT_MAX; }
  __device__ __forceinline__ void insert(float nv, int k, int lane) {
    const unsigned full = 0xffffffffu;
    const int pos = __popc(__ballot_sync(full, v >= nv));     // entries that stay in front
    const float up = __shfl_up_sync(full, v, 1);
    if (lane == pos) v = nv;
    else if (lane > pos) v = up;
    (void)k;
  }
};

// One warp per row.  Same outputs / fallback protocol as knn_rerank_kernel<false, NI>.
template <int NI>
__global__ void __launch_bounds__(RR_WARPS * 32)
knn_rerank_sym_kernel(const RerankParams p, const SymLists l, const int64_t row_lo) {
  __shared__ int s_surv[RR_WARPS][RR_SURV];
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t row = row_lo + (int64_t)blockIdx.x * RR_WARPS + wib;   // rows [row_lo, p.nq)
  if (row >= p.nq) return;
  const int k = p.k, dim = p.dim;
  float q[16];
  if constexpr (NI > 0) {
    const float4* q4 = reinterpret_cast<const float4*>(p.Q + row * dim);
    const int nv = dim >> 2;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
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
  const float eb_max = p.bsc->err_max, hb_max = p.bsc->hn_max;
  const float ea_row = p.ea[row], ha_row = p.ha[row];
  const int c0 = l.dlog_cnt[row * 2], c1 = l.dlog_cnt[row * 2 + 1], cr = l.rcnt[row];
  bool undecided = (p.bsc->range_flag != 0) || !(ea_row < INFINITY) || c0 < 0 || c1 < 0 ||
                   cr > l.rcap || (c0 + c1 + cr) < k;
  // the three lists as one index space [0, n0) [n0, n0 + n1) [n0 + n1, total)
  const int n0 = max(c0, 0), n1 = max(c1, 0), nr = min(max(cr, 0), l.rcap);
  const int total = n0 + n1 + nr;
  const float* v0 = l.dlog_val + (row * 2) * l.log_cap;
  const float* v1 = l.dlog_val + (row * 2 + 1) * l.log_cap;
  const float* vr = l.rval + row * l.rcap;
  const int* i0 = l.dlog_idx + (row * 2) * l.log_cap;
  const int* i1 = l.dlog_idx + (row * 2 + 1) * l.log_cap;
  const int* ir = l.ridx + row * l.rcap;
  auto score_at = [&](int e) -> float {
    return e < n0 ? v0[e] : (e < n0 + n1 ? v1[e - n0] : vr[e - n0 - n1]);
  };
  auto index_at = [&](int e) -> int {
    return e < n0 ? i0[e] : (e < n0 + n1 ? i1[e - n0] : ir[e - n0 - n1]);
  };
  int n_surv = 0;
  if (!undecided) {
    // ---- pass 1: k-th best score of the union
    WarpTopScores ts;
    ts.init();
    float kth = -FLT_MAX;
    for (int e0 = 0; e0 < total; e0 += 32) {
      const int e = e0 + lane;
      const float v = (e < total) ? score_at(e) : -FLT_MAX;
      unsigned m = __ballot_sync(full, v > kth);
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1;
        const float nv = __shfl_sync(full, v, src);
        if (nv > kth) {                              // warp-uniform
          ts.insert(nv, k, lane);
          kth = __shfl_sync(full, ts.v, k - 1);
        }
      }
    }
    // ---- same keep rule as the screening kernels, with the global k-th best
    ListState st;
    st.ha = ha_row;
    st.e2 = 2.f * (ea_row + eb_max);
    st.dl = (st.ha + hb_max) * SCR_ACC_EPS;
    st.tk = kth;
    const float tkeep = keep_threshold(st);
    // ---- pass 2: survivors
    for (int e0 = 0; e0 < total; e0 += 32) {
      const int e = e0 + lane;
      const bool keep = e < total && score_at(e) >= tkeep;
      const unsigned m = __ballot_sync(full, keep);
      const int pos = n_surv + __popc(m & ((1u << lane) - 1u));
      if (keep && pos < RR_SURV) s_surv[wib][pos] = index_at(e);
      n_surv += __popc(m);
    }
    if (n_surv > RR_SURV) undecided = true;
    __syncwarp();
  }
  if (undecided) {
    if (lane == 0) {
      const int pos = atomicAdd(&p.status->n_fallback, 1);
      if (pos < p.fallback_cap) {
        p.fallback_rows[pos] = (int)row;
      } else {
        atomicOr(&p.status->invalid, 1);
        p.out_idx[(int64_t)row * p.k] = -1;
      }
    }
    return;
  }
  WarpTopK top;
  top.init();
  float sv[2] = {FLT_MAX, FLT_MAX};
  int si[2] = {INT_MAX, INT_MAX};
  const int* surv = s_surv[wib];
  exact_scan<NI, true>(q, p.B, dim, k, lane, n_surv, [surv](int c) { return surv[c]; }, top, sv,
                       si);
  if (n_surv <= 32) bitonic_pairs_asc<32>(sv, si, lane);
  else bitonic_pairs_asc<64>(sv, si, lane);
  top.v = sv[0];
  top.i = si[0];
  if (lane == 0) atomicAdd(&p.status->pairs, (unsigned long long)n_surv);
  rerank_store(p, row, lane, top);
}

}  // namespace exb
