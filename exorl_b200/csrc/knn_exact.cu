// Exact fp32 kNN on CUDA cores: direct-difference squared distances, tiled, with a warp-resident
// sorted top-k per query row.  No (nq, nb) matrix is written anywhere.
//
// Replaces /root/reference/utils/utils.py:293-300 and agents/unsupervised_learning/proto.py:115-116
// (the reference materialises a (b1, b2, c) temporary and calls torch.topk).
//
// Layout: block = 256 threads = 8 warps handles TQ=32 query rows against one bank split, TB=128
// bank rows per step.  Warp w owns query rows 4w..4w+3; lane l owns bank columns l, l+32, l+64,
// l+96 of the step, so after the K loop every lane of a warp holds four candidates for each of
// the warp's four rows and the per-row top-k lives across the 32 lanes of that warp (lane r holds
// the r-th smallest).  Inserts are a ballot + shuffle, no shared memory.
#include "common.cuh"
#include <float.h>
#include <limits.h>

namespace exb {

constexpr int TQ = 32;
constexpr int TB = 128;
constexpr int KC = 32;
constexpr int EXACT_THREADS = 256;
constexpr int MERGE_PER_LANE = 16;  // merge handles up to 32*16 = 512 entries per row

__global__ void __launch_bounds__(EXACT_THREADS, 2)
knn_exact_kernel(const float* __restrict__ Q, int64_t nq, const float* __restrict__ B, int64_t nb,
                 int dim, int k, int64_t split_len, int splits,
                 const int* __restrict__ row_list, const int* __restrict__ row_count,
                 float* __restrict__ part_d2, int* __restrict__ part_idx) {
  __shared__ __align__(16) float sQ[KC][TQ + 4];
  __shared__ float sB[KC][TB + 1];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned full = 0xffffffffu;
  const int64_t nrows = row_count ? (int64_t)min((int64_t)*row_count, nq) : nq;
  const int64_t q0 = (int64_t)blockIdx.x * TQ;
  if (q0 >= nrows) return;
  const int split = blockIdx.y;
  const int64_t b_begin = (int64_t)split * split_len;
  const int64_t b_end = min(nb, b_begin + split_len);

  float lv[4];
  int li[4];
  float thr[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) { lv[r] = FLT_MAX; li[r] = INT_MAX; thr[r] = FLT_MAX; }
  // FLT_MAX (not +inf) as the sentinel so that a genuine +inf distance never ties with it.

  // global query row of each smem slot handled by this thread's loads
  for (int64_t b0 = b_begin; b0 < b_end; b0 += TB) {
    float acc[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[r][j] = 0.f;

    for (int k0 = 0; k0 < dim; k0 += KC) {
      __syncthreads();
      const int kk = lane;
      const bool kin = (k0 + kk) < dim;
#pragma unroll
      for (int i = 0; i < (TQ * KC) / EXACT_THREADS; ++i) {
        const int row = warp + 8 * i;
        const int64_t slot = q0 + row;
        float v = 0.f;
        if (kin && slot < nrows) {
          const int64_t qrow = row_list ? (int64_t)row_list[slot] : slot;
          v = Q[qrow * dim + k0 + kk];
        }
        sQ[kk][row] = v;
      }
#pragma unroll
      for (int i = 0; i < (TB * KC) / EXACT_THREADS; ++i) {
        const int row = warp + 8 * i;
        const int64_t brow = b0 + row;
        sB[kk][row] = (kin && brow < b_end) ? B[brow * dim + k0 + kk] : 0.f;
      }
      __syncthreads();
#pragma unroll
      for (int c = 0; c < KC; ++c) {
        const float4 q4 = *reinterpret_cast<const float4*>(&sQ[c][warp * 4]);
        const float qv[4] = {q4.x, q4.y, q4.z, q4.w};
        float bv[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) bv[j] = sB[c][lane + 32 * j];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float df = qv[r] - bv[j];
            acc[r][j] = fmaf(df, df, acc[r][j]);
          }
      }
    }

    // running top-k update, candidates visited in increasing bank index
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t col = b0 + lane + 32 * j;
        const float v = acc[r][j];
        const bool cand = (col < b_end) && (v < thr[r]);
        unsigned m = __ballot_sync(full, cand);
        while (m) {
          const int src = __ffs(m) - 1;
          m &= m - 1;
          const float cv = __shfl_sync(full, v, src);
          const int ci = (int)(b0 + src + 32 * j);
          if (cv < thr[r]) {
            const int pos = __popc(__ballot_sync(full, lv[r] <= cv));
            const float up_v = __shfl_up_sync(full, lv[r], 1);
            const int up_i = __shfl_up_sync(full, li[r], 1);
            if (lane == pos) { lv[r] = cv; li[r] = ci; }
            else if (lane > pos) { lv[r] = up_v; li[r] = up_i; }
            thr[r] = __shfl_sync(full, lv[r], k - 1);
          }
        }
      }
    }
  }

#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int64_t slot = q0 + warp * 4 + r;
    if (slot < nrows && lane < k) {
      const int64_t o = (slot * splits + split) * k + lane;
      part_d2[o] = lv[r];
      part_idx[o] = (li[r] == INT_MAX) ? -1 : li[r];
    }
  }
}

// One warp per output row merges n_lists ascending lists of k (value, index) pairs.
//   entry e of list l of slot s lives at  l*list_stride + s*row_stride + e.
// Values of FLT_MAX / index -1 mark unfilled entries.  take_sqrt converts squared distances.
__global__ void __launch_bounds__(256)
topk_merge_kernel(const float* __restrict__ vals, const int* __restrict__ idxs, int n_lists,
                  int64_t list_stride, int64_t row_stride, int64_t nslots, int k, int take_sqrt,
                  int idx_base, const int* __restrict__ row_list,
                  const int* __restrict__ row_count, float* __restrict__ out_dist,
                  int* __restrict__ out_idx) {
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int64_t slot = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t limit = row_count ? (int64_t)min((int64_t)*row_count, nslots) : nslots;
  if (slot >= limit) return;
  const int total = n_lists * k;

  float v[MERGE_PER_LANE];
  int ix[MERGE_PER_LANE];
#pragma unroll
  for (int i = 0; i < MERGE_PER_LANE; ++i) {
    const int e = lane + 32 * i;
    v[i] = FLT_MAX;
    ix[i] = INT_MAX;
    if (e < total) {
      const int l = e / k, r = e - l * k;
      const int64_t o = (int64_t)l * list_stride + slot * row_stride + r;
      const int id = idxs[o];
      if (id >= 0) { v[i] = vals[o]; ix[i] = id; }
    }
  }
  float my_v = FLT_MAX;
  int my_i = -1;
  for (int round = 0; round < k; ++round) {
    float bv = v[0];
    int bi = ix[0], bs = 0;
#pragma unroll
    for (int i = 1; i < MERGE_PER_LANE; ++i)
      if (pair_less(v[i], ix[i], bv, bi)) { bv = v[i]; bi = ix[i]; bs = i; }
    float wv = bv;
    int wi = bi;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const float ov = __shfl_xor_sync(full, wv, off);
      const int oi = __shfl_xor_sync(full, wi, off);
      if (pair_less(ov, oi, wv, wi)) { wv = ov; wi = oi; }
    }
    if (wi == bi && wv == bv && bi != INT_MAX) {  // this lane owns the winner (indices unique)
#pragma unroll
      for (int i = 0; i < MERGE_PER_LANE; ++i)
        if (i == bs) { v[i] = FLT_MAX; ix[i] = INT_MAX; }
    }
    if (lane == round) { my_v = wv; my_i = (wi == INT_MAX) ? -1 : wi; }
  }
  if (lane < k) {
    const int64_t orow = row_list ? (int64_t)row_list[slot] : slot;
    float d = my_v;
    if (take_sqrt && my_i >= 0) d = sqrtf(d);
    out_dist[orow * k + lane] = d;
    out_idx[orow * k + lane] = (my_i >= 0) ? my_i + idx_base : -1;
  }
}

int exact_splits(int64_t nq, int64_t nb, int k) {
  const int64_t q_tiles = (nq + TQ - 1) / TQ;
  const int64_t b_tiles = (nb + TB - 1) / TB;
  int64_t want = (4LL * sm_count() + q_tiles - 1) / q_tiles;
  int64_t cap = (32 * MERGE_PER_LANE) / k;
  if (cap > 16) cap = 16;
  if (want > cap) want = cap;
  if (want > b_tiles) want = b_tiles;
  if (want < 1) want = 1;
  return (int)want;
}

// Shared launcher: direct mode (row_list == nullptr) or indirect fallback mode.
int launch_exact(const float* src, int64_t nq, const float* bank, int64_t nb, int dim, int k,
                 int idx_base, float* out_dist, int* out_idx, float* part_d2, int* part_idx,
                 int splits, const int* row_list, const int* row_count, int64_t max_slots,
                 cudaStream_t st) {
  const int64_t b_tiles = (nb + TB - 1) / TB;
  const int64_t split_len = ((b_tiles + splits - 1) / splits) * TB;
  const int64_t q_tiles = (max_slots + TQ - 1) / TQ;
  if (q_tiles == 0) return EXB_OK;
  dim3 grid((unsigned)q_tiles, (unsigned)splits);
  knn_exact_kernel<<<grid, EXACT_THREADS, 0, st>>>(src, max_slots, bank, nb, dim, k, split_len, splits,
                                                   row_list, row_count, part_d2, part_idx);
  EXB_LAUNCH_CHECK();
  const int wpb = 8;
  topk_merge_kernel<<<(unsigned)((max_slots + wpb - 1) / wpb), wpb * 32, 0, st>>>(
      part_d2, part_idx, splits, (int64_t)k, (int64_t)splits * k, max_slots, k, 1, idx_base,
      row_list, row_count, out_dist, out_idx);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

}  // namespace exb

using namespace exb;

extern "C" size_t exb_knn_exact_workspace_bytes(int64_t nq, int64_t nb, int dim, int k) {
  (void)dim;
  if (nq <= 0 || nb <= 0 || k <= 0 || k > EXB_MAX_K) return 0;
  const int splits = exact_splits(nq, nb, k);
  return 2 * align_up((size_t)nq * splits * k * 4, 256) + 512;
}

static int check_knn_args(const float* src, int64_t nq, const float* bank, int64_t nb, int dim,
                          int k, const float* out_dist, const int32_t* out_idx) {
  EXB_CHECK_ARG(src && bank && out_dist && out_idx, "null pointer");
  EXB_CHECK_ARG(nq > 0 && nb > 0 && dim > 0, "empty input: nq=%lld nb=%lld dim=%d",
                (long long)nq, (long long)nb, dim);
  EXB_CHECK_ARG(nb < (1LL << 31) && nq < (1LL << 31), "more than 2^31 rows");
  EXB_CHECK_ARG(k >= 1 && k <= EXB_MAX_K, "k=%d outside [1,%d]", k, EXB_MAX_K);
  if (k > nb) {
    set_error("selected index k out of range (k=%d > bank rows=%lld)", k, (long long)nb);
    return EXB_EKRANGE;
  }
  return EXB_OK;
}

extern "C" int exb_knn_exact_f32(const float* src, int64_t nq, const float* bank, int64_t nb,
                                 int dim, int k, int32_t idx_base, float* out_dist,
                                 int32_t* out_idx, void* workspace, size_t workspace_bytes,
                                 exb_stream_t stream) {
  int rc = check_knn_args(src, nq, bank, nb, dim, k, out_dist, out_idx);
  if (rc) return rc;
  const size_t need = exb_knn_exact_workspace_bytes(nq, nb, dim, k);
  EXB_CHECK_ARG(workspace && workspace_bytes >= need, "workspace too small: %zu < %zu",
                workspace_bytes, need);
  const int splits = exact_splits(nq, nb, k);
  Carver cv(workspace);
  float* part_d2 = cv.take<float>((size_t)nq * splits * k);
  int* part_idx = cv.take<int>((size_t)nq * splits * k);
  return launch_exact(src, nq, bank, nb, dim, k, idx_base, out_dist, out_idx, part_d2, part_idx,
                      splits, nullptr, nullptr, nq, (cudaStream_t)stream);
}

extern "C" int exb_topk_merge(const float* dist_lists, const int32_t* idx_lists, int n_lists,
                              int64_t nq, int k, float* out_dist, int32_t* out_idx,
                              exb_stream_t stream) {
  EXB_CHECK_ARG(dist_lists && idx_lists && out_dist && out_idx, "null pointer");
  EXB_CHECK_ARG(nq > 0 && n_lists >= 1 && k >= 1 && k <= EXB_MAX_K, "bad sizes");
  EXB_CHECK_ARG(n_lists * k <= 32 * MERGE_PER_LANE, "n_lists*k=%d exceeds merge capacity %d",
                n_lists * k, 32 * MERGE_PER_LANE);
  const int wpb = 8;
  topk_merge_kernel<<<(unsigned)((nq + wpb - 1) / wpb), wpb * 32, 0, (cudaStream_t)stream>>>(
      dist_lists, idx_lists, n_lists, nq * (int64_t)k, (int64_t)k, nq, k, 0, 0, nullptr, nullptr,
      out_dist, out_idx);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}
