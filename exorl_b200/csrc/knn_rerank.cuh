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
//   NI > 0       dim % 4 == 0, dim <= 128 NI, 16-byte aligned rows: float4 loads, 8 survivors per
//                batch, all loads of a batch in flight; NI = 0: scalar loads, any dim <= 512.
#pragma once
#include "pbe_common.cuh"

namespace exb {

struct RerankStatus {   // device, zeroed per call (24 bytes: the prep kernel clears 3 words)
  int n_fallback;
  int invalid;
  unsigned long long pairs;
  unsigned ticket;      // fused PBE epilogue: blocks finished
  unsigned pad;
};

// Fused PBE epilogue of the per-update chain (LAT only): utils/utils.py:301-318 + RMS :264-276
// the moments are accumulated by the rerank blocks and the last block to finish applies the RMS
// update; without RMS every warp finishes its own row.  (With RMS the per-row finish needs the
// updated mean of ALL rows, so it stays a separate, fully parallel launch.)
struct FusedPbe {
  int on, avg, use_rms, use_clip;
  int raw_kth;           // no RMS, no clip, no log: reward = the k-th distance (proto.py:117-118)
  float clip;
  RmsParams rms;
  float* out_reward;     // [nq]
  double2* partials;     // [gridDim.x]
};

constexpr int RR_PER_LANE = 12;   // 32 * 12 = 384 >= 16 splits * 24 slots
constexpr int RR_SURV = 64;       // survivors recomputed exactly per row (else undecided)
constexpr int RR_WARPS = 4;

struct RerankParams {
  const float* Q;
  const float* B;
  int64_t nq, nb;
  int dim, k, idx_base, splits;
  // Lists gathered from several bank shards (exb_knn_rerank_lists_f32): list s of a row belongs
  // to shard s / lists_per_shard, whose arrays start shard_stride rows further on and whose
  // indices are local to the shard (global = local + shard_lo).  One shard: lists_per_shard =
  // splits, shard_rows = 0.
  int lists_per_shard;
  int64_t shard_rows;        // rows per shard block in cand_* ([n_shards, shard_rows, lps, CAP])
  int shard_lo[16];
  const float* cand_val;
  const int* cand_idx;
  const int* cand_cnt;
  const float* cand_tk;      // LAT: per-split lower bound of the k-th best score
  const float* cand_tr;      // LAT: per-split rsel-th best pair maximum, rsel = ceil(k / splits)
  const float* ha;
  const float* ea;
  const PrepScalars* qsc;    // !LAT
  const PrepScalars* bsc;    // !LAT
  const float* bank_max;     // LAT: {EB, HBmax}
  int split_cols;            // LAT: bank columns per split (tiles_per_split * 64)
  const int* coop_flags;     // LAT: [m_tiles, splits, 4] written by the latency screen (LatParams)
  float* out_dist;
  int* out_idx;
  int* fallback_rows;        // !LAT
  int fallback_cap;
  RerankStatus* status;
  long long* trace;
  FusedPbe pbe;
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

// Bitonic sorts across a warp, two elements per lane (element e = lane + 32 * slot).
// Scores: descending.  Partner distance 32 is the only in-register stage.
__device__ __forceinline__ void bitonic64_desc(float& x0, float& x1, int lane) {
  const unsigned full = 0xffffffffu;
#pragma unroll
  for (int ksz = 2; ksz <= 64; ksz <<= 1) {
#pragma unroll
    for (int j = ksz >> 1; j > 0; j >>= 1) {
      if (j == 32) {
        const float lo = fminf(x0, x1), hi = fmaxf(x0, x1);
        x0 = hi;   // (e & 64) == 0 for every element: one descending run
        x1 = lo;
      } else {
        const float y0 = __shfl_xor_sync(full, x0, j), y1 = __shfl_xor_sync(full, x1, j);
        const bool lower = (lane & j) == 0;
        const bool d0 = ((lane & ksz) == 0), d1 = (((lane + 32) & ksz) == 0);
        // descending run: the lower position keeps the max
        x0 = (lower == d0) ? fmaxf(x0, y0) : fminf(x0, y0);
        x1 = (lower == d1) ? fmaxf(x1, y1) : fminf(x1, y1);
      }
    }
  }
}

// (value, index) pairs ascending by pair_less; N = 32 (slot 0 only) or 64.
template <int N>
__device__ __forceinline__ void bitonic_pairs_asc(float (&v)[2], int (&ix)[2], int lane) {
  const unsigned full = 0xffffffffu;
#pragma unroll
  for (int ksz = 2; ksz <= N; ksz <<= 1) {
#pragma unroll
    for (int j = ksz >> 1; j > 0; j >>= 1) {
      if (j == 32) {
        if (pair_less(v[1], ix[1], v[0], ix[0])) {
          const float tv = v[0]; v[0] = v[1]; v[1] = tv;
          const int ti = ix[0]; ix[0] = ix[1]; ix[1] = ti;
        }
      } else {
#pragma unroll
        for (int s = 0; s < N / 32; ++s) {
          const float yv = __shfl_xor_sync(full, v[s], j);
          const int yi = __shfl_xor_sync(full, ix[s], j);
          const bool lower = (lane & j) == 0;
          const bool asc = (((lane + 32 * s) & ksz) == 0);
          const bool y_less = pair_less(yv, yi, v[s], ix[s]);
          // ascending run: the lower position keeps the smaller pair
          const bool take = (lower == asc) ? y_less : !y_less;
          if (take) { v[s] = yv; ix[s] = yi; }
        }
      }
    }
  }
}

// Exact squared distances of `count` bank rows (ids from idx_of(c)) to the query held in
// registers.  NI = float4 loads per lane and row (dim <= 128 * NI, dim % 4 == 0, aligned rows);
// NI = 0: scalar loads for any dim <= 512.  NB rows per batch and ALL of a batch's loads are
// issued before the first use, so a batch costs one memory round trip.
// STORE: lane (c & 31) keeps candidate c in slot c >> 5 (count <= 64); else insert into `top`.
template <int NI, bool STORE, typename IdxOf>
__device__ __forceinline__ void exact_scan(const float (&q)[16], const float* __restrict__ B,
                                           int dim, int k, int lane, int count, IdxOf idx_of,
                                           WarpTopK& top, float (&sv)[2], int (&si)[2]) {
  const unsigned full = 0xffffffffu;
  constexpr int NB = NI > 0 ? 8 : 4;
  for (int c0 = 0; c0 < count; c0 += NB) {
    int j[NB];
    float a[NB];
#pragma unroll
    for (int u = 0; u < NB; ++u) {
      j[u] = idx_of(min(c0 + u, count - 1));
      a[u] = 0.f;
    }
    if constexpr (NI > 0) {
      const int nv = dim >> 2;
      float4 b[NI][NB];
#pragma unroll
      for (int i = 0; i < NI; ++i) {
        const int c = lane + 32 * i;
        const int cc = (c < nv) ? c : 0;       // clamped, masked below: loads stay unconditional
#pragma unroll
        for (int u = 0; u < NB; ++u)
          b[i][u] = __ldg(reinterpret_cast<const float4*>(B + (int64_t)j[u] * dim) + cc);
      }
#pragma unroll
      for (int i = 0; i < NI; ++i) {
        const bool ok = (lane + 32 * i) < nv;
#pragma unroll
        for (int u = 0; u < NB; ++u) {
          float d = q[4 * i] - b[i][u].x;
          float t = d * d;
          d = q[4 * i + 1] - b[i][u].y;
          t = fmaf(d, d, t);
          d = q[4 * i + 2] - b[i][u].z;
          t = fmaf(d, d, t);
          d = q[4 * i + 3] - b[i][u].w;
          t = fmaf(d, d, t);
          a[u] += ok ? t : 0.f;
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
    for (int u = 0; u < NB; ++u) {
      if constexpr (STORE) {
        const int c = c0 + u;
        if (c < count && (c & 31) == lane) {
          if (c < 32) { sv[0] = a[u]; si[0] = j[u]; }
          else { sv[1] = a[u]; si[1] = j[u]; }
        }
      } else {
        if (c0 + u < count) top.insert(a[u], j[u], k, lane);
      }
    }
  }
}

// Rare path of the per-update chain, kept out of line so that it does not weigh on the register
// allocation of the common one: exact distances of `count` bank rows (list[c], or base + c when
// list is null) inserted into `top`.
template <int NI>
__device__ __noinline__ void scan_insert(const float (&q)[16], const float* __restrict__ B,
                                         int dim, int k, int lane, int count, const int* list,
                                         int base, WarpTopK& top) {
  float sv[2];
  int si[2];
  exact_scan<NI, false>(q, B, dim, k, lane, count,
                        [list, base](int c) { return list ? list[c] : base + c; }, top, sv, si);
}

// One warp, one query row: leaves the row's k smallest SQUARED distances (ascending, lane r = the
// r-th) in `top` and the query row in `q`.  Returns false when the row has no result here (!LAT:
// it went to the fallback list).  LAT: `need` receives the bank ranges that still have to be
// scanned for this row — bit sp = the columns of split sp, RR_NEED_ALL = the whole bank — which
// the block then scans cooperatively (coop_scan).
constexpr unsigned RR_NEED_ALL = 0xffffffffu;
template <bool LAT, int NI, bool SH = false, bool COOP = false>
__device__ __forceinline__ bool rerank_row(const RerankParams& p, int64_t row, int lane,
                                           int* surv_w, float* val_w, float (&q)[16],
                                           WarpTopK& top, unsigned& need) {
  // surv_w / val_w: this warp's RR_SURV-entry scratch rows in shared memory
  const unsigned full = 0xffffffffu;
  const bool tr = p.trace != nullptr && row == 0;
  if (tr && lane == 0) p.trace[8] = clock_now();
  const int k = p.k, splits = p.splits, dim = p.dim;

  // query row into registers first: its latency hides behind the list gathering
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

  // ---- everything this row needs from the screen is requested at once (one round trip): counts
  // (lane sp holds split sp, splits <= 16), per-split thresholds, the row's scalars, and all list
  // slots (entry e = lane + 32 i; slots beyond a split's count are read and masked afterwards).
  // SH (lists gathered from several shards): lane L < splits owns list L of the row and works out
  // once where it lives (its shard's block, its slot) and the shard's index offset; the gather
  // below fetches them by shuffle.  !SH: list sp of the row is simply row * splits + sp.
  int64_t my_base = row * splits + lane;
  int my_lo = 0;
  if constexpr (SH) {
    const int lps = p.lists_per_shard;
    const int sh = lane / lps;
    my_base = ((int64_t)sh * p.shard_rows + row) * lps + (lane - sh * lps);
#pragma unroll
    for (int i = 0; i < 16; ++i) my_lo = (sh == i) ? p.shard_lo[i] : my_lo;
  }
  const int my_cnt = (lane < splits) ? p.cand_cnt[my_base] : 0;
  float my_tk = -FLT_MAX, my_tr = -FLT_MAX;
  if constexpr (LAT) {
    if (lane < splits) {
      my_tk = p.cand_tk[my_base];
      my_tr = p.cand_tr[my_base];
    }
  }
  const float ea_row = p.ea[row];
  const float ha_row = p.ha[row];
  float ev[RR_PER_LANE];
  int ei[RR_PER_LANE];
#pragma unroll
  for (int i = 0; i < RR_PER_LANE; ++i) {
    const int e = lane + 32 * i;
    const int sp = e / SCR_CAP;
    ev[i] = -FLT_MAX;
    ei[i] = -1;
    if constexpr (SH) {
      const int64_t b = __shfl_sync(full, my_base, sp & 31);
      const int lo = __shfl_sync(full, my_lo, sp & 31);
      if (sp < splits) {
        const int64_t a = b * SCR_CAP + (e - sp * SCR_CAP);
        ev[i] = p.cand_val[a];
        ei[i] = p.cand_idx[a] + lo;
      }
    } else if (sp < splits) {
      ev[i] = p.cand_val[row * splits * SCR_CAP + e];
      ei[i] = p.cand_idx[row * splits * SCR_CAP + e];
    }
  }
  int why = undecided ? 4 : 0;           // debug only (trace[12..15] of an undecided row)
  if (__any_sync(full, my_cnt < 0)) { undecided = true; why |= 1; }
  int total = max(my_cnt, 0);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) total += __shfl_xor_sync(full, total, off);
  if (total < k) { undecided = true; why |= 2; }
#pragma unroll
  for (int i = 0; i < RR_PER_LANE; ++i) {
    const int e = lane + 32 * i;
    const int sp = e / SCR_CAP, c = e - sp * SCR_CAP;
    const int cnt = __shfl_sync(full, my_cnt, sp & 31);
    if (!(sp < splits && c < cnt)) {
      ev[i] = -FLT_MAX;
      ei[i] = -1;
    }
  }
  if (tr && lane == 0) p.trace[9] = clock_now();

  int n_surv = 0;
  if (!undecided && !(ea_row < INFINITY)) undecided = true;   // query row outside the fp16 range
  if (!undecided) {
    // ---- k-th best screened score over the union of splits.
    // T0 = the best per-split lower bound of the k-th best: the global k-th best is >= T0 and is
    // the k-th largest of the entries >= T0 (the split that attains T0 alone holds k of them).
    // L1: with r = ceil(k / splits), every split holds r scores >= its r-th best, so the m-th
    // largest of those per-split values, m = ceil(k / r), has m * r >= k scores at or above it.
    float t0 = -FLT_MAX;
    if constexpr (LAT) {
      if (lane < splits && my_cnt >= k) t0 = my_tk;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) t0 = fmaxf(t0, __shfl_xor_sync(full, t0, off));
      const int rsel = (k + splits - 1) / splits;
      const int msel = (k + rsel - 1) / rsel;
      const float mine = (lane < splits) ? my_tr : -FLT_MAX;
      int above = 0;       // how many splits' values beat this lane's (ties broken by lane id)
#pragma unroll
      for (int s2 = 0; s2 < 16; ++s2) {    // splits <= 16; unrolled so that the shuffles pipeline
        const float o = __shfl_sync(full, mine, s2);
        above += (s2 < splits && (o > mine || (o == mine && s2 < lane))) ? 1 : 0;
      }
      const unsigned who = __ballot_sync(full, lane < splits && above == msel - 1);
      if (who != 0u) t0 = fmaxf(t0, __shfl_sync(full, mine, __ffs(who) - 1));
    }
    int n0 = 0;
#pragma unroll
    for (int i = 0; i < RR_PER_LANE; ++i) {
      const bool keep = ei[i] >= 0 && ev[i] >= t0;
      const unsigned m = __ballot_sync(full, keep);
      const int pos = n0 + __popc(m & ((1u << lane) - 1u));
      if (keep && pos < RR_SURV) val_w[pos] = ev[i];
      n0 += __popc(m);
    }
    __syncwarp();
    float kth = -FLT_MAX;
    if (n0 <= 32) {
      float xv[2] = {(lane < n0) ? -val_w[lane] : FLT_MAX, FLT_MAX};
      int xi[2] = {lane, lane};
      bitonic_pairs_asc<32>(xv, xi, lane);     // ascending on the negated scores
      kth = -__shfl_sync(full, xv[0], k - 1);
    } else if (n0 <= RR_SURV) {
      // one 64-element bitonic sort instead of k dependent arg-max rounds
      float x0 = (lane < n0) ? val_w[lane] : -FLT_MAX;
      float x1 = (lane + 32 < n0) ? val_w[lane + 32] : -FLT_MAX;
      bitonic64_desc(x0, x1, lane);
      kth = __shfl_sync(full, x0, k - 1);      // k <= 16: always in slot 0
    } else {
      float wv[RR_PER_LANE];
#pragma unroll
      for (int i = 0; i < RR_PER_LANE; ++i) wv[i] = ev[i];
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
    }
    // same keep rule as the screening kernel, now with the global k-th best
    ListState st;
    st.ha = ha_row;
    st.e2 = 2.f * (ea_row + eb_max);
    st.dl = (st.ha + hb_max) * SCR_ACC_EPS;
    st.tk = kth;
    const float tkeep = keep_threshold(st);
#pragma unroll
    for (int i = 0; i < RR_PER_LANE; ++i) {
      const bool keep = ei[i] >= 0 && ev[i] >= tkeep;
      const unsigned m = __ballot_sync(full, keep);
      const int pos = n_surv + __popc(m & ((1u << lane) - 1u));
      if (keep && pos < RR_SURV) surv_w[pos] = ei[i];
      n_surv += __popc(m);
    }
    if (n_surv > RR_SURV) { undecided = true; why |= 8; }
    __syncwarp();
  }
  if (!LAT && undecided && p.trace != nullptr && lane == 0) {
    p.trace[12] = row;
    p.trace[13] = why;
    p.trace[14] = total;
    p.trace[15] = n_surv;
    // histogram of the reasons (EXB_SCREEN_TRACE=1): [0] undecided rows, [1] a screen list
    // overflowed, [2] fewer than k candidates, [3] fp16 range, [4] more than RR_SURV survivors
    unsigned long long* h = reinterpret_cast<unsigned long long*>(p.trace);
    atomicAdd(h + 0, 1ull);
    if (why & 1) atomicAdd(h + 1, 1ull);
    if (why & 2) atomicAdd(h + 2, 1ull);
    if (why & 4) atomicAdd(h + 3, 1ull);
    if (why & 8) atomicAdd(h + 4, 1ull);
  }
  if (tr && lane == 0) p.trace[10] = clock_now();

  top.init();
  float sv[2] = {FLT_MAX, FLT_MAX};
  int si[2] = {INT_MAX, INT_MAX};
  if (undecided) {
    if constexpr (LAT) {
      // rare: recompute this row inside the chain (no fallback launches).  The bank ranges it
      // lost are recorded in `need` and scanned by the whole block together (coop_scan): with
      // every row of a batch undecided, per-warp scans re-read the bank once per ROW from L2
      // (1024 x 2 MB: 369 us); the block-wide scan reads it once per 4 rows.
      if (lane == 0) atomicAdd(&p.status->n_fallback, 1);
      // copies for the out-of-line calls: q and top themselves stay in registers
      float qq[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) qq[i] = q[i];
      WarpTopK tt;
      tt.init();
      if ((why & 4) || !(ea_row < INFINITY)) {
        // outside the fp16 range: no screened score means anything, scan the whole bank
        if constexpr (COOP) need = RR_NEED_ALL;
        else scan_insert<NI>(qq, p.B, dim, k, lane, (int)p.nb, nullptr, 0, tt);
      } else {
        // Only the splits whose list overflowed lost entries.  Every split's k-th chain entry is
        // still a lower bound of the k-th best score, so the best of them gives a valid (looser)
        // keep threshold: the complete lists are filtered with it, the overflowed splits' column
        // ranges are scanned whole.
        float t0 = (lane < splits) ? my_tk : -FLT_MAX;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) t0 = fmaxf(t0, __shfl_xor_sync(full, t0, off));
        ListState st;
        st.ha = ha_row;
        st.e2 = 2.f * (ea_row + eb_max);
        st.dl = (st.ha + hb_max) * SCR_ACC_EPS;
        st.tk = t0;
        const float tkeep0 = (t0 > -FLT_MAX) ? keep_threshold(st) : -FLT_MAX;
        unsigned ovf = __ballot_sync(full, lane < splits && my_cnt < 0);   // splits <= 16
        if constexpr (COOP) {
          need = ovf;                              // scanned by the block together (coop_scan)
        } else {
          while (ovf) {
            const int sp = __ffs(ovf) - 1;
            ovf &= ovf - 1;
            const int lo = sp * p.split_cols;
            const int cnt = min((int)p.nb - lo, p.split_cols);
            if (cnt > 0) scan_insert<NI>(qq, p.B, dim, k, lane, cnt, nullptr, lo, tt);
          }
        }
        int n = 0;
        __syncwarp();
#pragma unroll
        for (int i = 0; i < RR_PER_LANE; ++i) {
          const bool keep = ei[i] >= 0 && ev[i] >= tkeep0;
          const unsigned m = __ballot_sync(full, keep);
          if (n + __popc(m) > RR_SURV) {
            __syncwarp();
            scan_insert<NI>(qq, p.B, dim, k, lane, n, surv_w, 0, tt);
            __syncwarp();
            n = 0;
          }
          if (keep) surv_w[n + __popc(m & ((1u << lane) - 1u))] = ei[i];
          n += __popc(m);
        }
        __syncwarp();
        if (n > 0) scan_insert<NI>(qq, p.B, dim, k, lane, n, surv_w, 0, tt);
      }
      top = tt;
    } else {
      if (lane == 0) {
        const int pos = atomicAdd(&p.status->n_fallback, 1);
        if (pos < p.fallback_cap) {
          p.fallback_rows[pos] = (int)row;
        } else {
          // more undecided rows than the device-side fallback holds (heavily duplicated data, or
          // values outside the fp16 range): flag the row so that the host can recompute it with
          // the exact kernel (exorl_b200/knn.py:_repair_overflow_rows) — never a silent result
          p.status->invalid = 1;
          p.out_idx[(int64_t)row * p.k] = -1;
        }
      }
      return false;
    }
  } else {
    // survivors' exact distances stay distributed over the lanes, then one bitonic sort
    const int* surv = surv_w;
    exact_scan<NI, true>(q, p.B, dim, k, lane, n_surv, [surv](int c) { return surv[c]; }, top,
                          sv, si);
    if (n_surv <= 32) bitonic_pairs_asc<32>(sv, si, lane);
    else bitonic_pairs_asc<64>(sv, si, lane);
    top.v = sv[0];
    top.i = si[0];
    if (lane == 0) atomicAdd(&p.status->pairs, (unsigned long long)n_surv);
  }
  if (tr && lane == 0) p.trace[11] = clock_now();
  return true;
}

// sqrt + store of a finished row; returns the row's distances (lane r = the r-th smallest).
__device__ __forceinline__ float rerank_store(const RerankParams& p, int64_t row, int lane,
                                              const WarpTopK& top) {
  const float dist = sqrtf(top.v);
  if (lane < p.k) {
    p.out_dist[row * p.k + lane] = dist;
    p.out_idx[row * p.k + lane] = top.i + p.idx_base;
  }
  return dist;
}

// Block-cooperative exact scan of the bank ranges the block's rows still need (per-update chain).
// Every warp passes its row's `need` mask (0 = nothing); the union of the ranges is streamed
// through shared memory in tiles of RR_COOP_ROWS bank rows (consecutive rows are one contiguous
// piece of memory) with cp.async into two buffers — tile t+1 is in flight while tile t is
// scored — and each warp whose mask covers the tile's range takes the tile's distances into its
// top-k with the arithmetic of exact_scan.  Must be called by all threads of the block.
constexpr int RR_COOP_ROWS = 8;
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <int NI, int NW = RR_WARPS>
__device__ __noinline__ void coop_scan(const RerankParams& p, const float (&q)[16], WarpTopK& top,
                                       unsigned need, int lane, int wib, float* s_tile,
                                       unsigned* s_need) {
  const unsigned full = 0xffffffffu;
  if (lane == 0) s_need[wib] = need;
  __syncthreads();
  unsigned uni = 0;
#pragma unroll
  for (int w = 0; w < NW; ++w) uni |= s_need[w];
  if (uni == 0) return;                               // block-uniform
  const int dim = p.dim, k = p.k, nb = (int)p.nb;
  const bool all = uni == RR_NEED_ALL;                // some row needs the whole bank
  const int split_cols = max(p.split_cols, RR_COOP_ROWS);   // a multiple of 64 in the chain
  constexpr int TILE_FLOATS = RR_COOP_ROWS * EXB_SCREEN_MAX_DIM;
  // a tile lies inside one split's column range (split_cols % RR_COOP_ROWS == 0)
  auto wanted = [&](int r0) { return all || ((uni >> min(r0 / split_cols, 31)) & 1u); };
  auto fetch = [&](int r0, float* dst) {              // rows [r0, r0 + RR_COOP_ROWS) ∩ bank
    const int n_el = min(RR_COOP_ROWS, nb - r0) * dim;
    const float* src = p.B + (int64_t)r0 * dim;
    if constexpr (NI > 0) {
      for (int i = threadIdx.x; i < (n_el >> 2); i += NW * 32)
        cp_async16(dst + 4 * i, src + 4 * i);
    } else {
      for (int i = threadIdx.x; i < n_el; i += NW * 32) dst[i] = __ldg(src + i);
    }
    cp_async_commit();
  };
  auto next_tile = [&](int r0) {                      // first wanted tile at or after r0, else nb
    while (r0 < nb && !wanted(r0)) r0 += RR_COOP_ROWS;
    return r0;
  };
  float thr = __shfl_sync(full, top.v, k - 1);       // warp-uniform: current k-th smallest
  int buf = 0;
  int r0 = next_tile(0);
  if (r0 < nb) fetch(r0, s_tile);
  while (r0 < nb) {
    cp_async_wait_all();
    __syncthreads();                                  // tile r0 landed; everyone left tile r0 - 1
    const int r1 = next_tile(r0 + RR_COOP_ROWS);
    if (r1 < nb) fetch(r1, s_tile + (buf ^ 1) * TILE_FLOATS);
    const float* tile = s_tile + buf * TILE_FLOATS;
    const int rows = min(RR_COOP_ROWS, nb - r0);
    const bool mine = need == RR_NEED_ALL || ((need >> min(r0 / split_cols, 31)) & 1u);
    if (mine) {
      float a[RR_COOP_ROWS];
#pragma unroll
      for (int u = 0; u < RR_COOP_ROWS; ++u) a[u] = 0.f;
      if constexpr (NI > 0) {
        const int nv = dim >> 2;
        const float4* t4 = reinterpret_cast<const float4*>(tile);
#pragma unroll
        for (int i = 0; i < NI; ++i) {
          const int c = lane + 32 * i;
          const bool ok = c < nv;
          const int cc = ok ? c : 0;
#pragma unroll
          for (int u = 0; u < RR_COOP_ROWS; ++u) {
            const float4 bv = t4[min(u, rows - 1) * nv + cc];
            float d = q[4 * i] - bv.x;
            float t = d * d;
            d = q[4 * i + 1] - bv.y;
            t = fmaf(d, d, t);
            d = q[4 * i + 2] - bv.z;
            t = fmaf(d, d, t);
            d = q[4 * i + 3] - bv.w;
            t = fmaf(d, d, t);
            a[u] += ok ? t : 0.f;
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int c = lane + 32 * i;
          if (c < dim) {
#pragma unroll
            for (int u = 0; u < RR_COOP_ROWS; ++u) {
              const float d = q[i] - tile[min(u, rows - 1) * dim + c];
              a[u] = fmaf(d, d, a[u]);
            }
          }
        }
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1)
#pragma unroll
        for (int u = 0; u < RR_COOP_ROWS; ++u) a[u] += __shfl_xor_sync(full, a[u], off);
      // a[u] and thr are warp-uniform: most candidates are rejected without a shuffle
      bool changed = false;
#pragma unroll
      for (int u = 0; u < RR_COOP_ROWS; ++u)
        if (u < rows && a[u] <= thr) {
          top.insert(a[u], r0 + u, k, lane);
          changed = true;
        }
      if (changed) thr = __shfl_sync(full, top.v, k - 1);
    }
    buf ^= 1;
    r0 = r1;
  }
  __syncthreads();
}

// COOP (LAT only): rows that lost bank ranges are scanned by the block together (coop_scan)
// instead of by their own warp.  A separate instance because its presence alone costs the common
// path 1.3 us (146 instead of 104 registers, 35 KB of static shared memory, a jump over the scan
// code: 15.9 vs 14.6 us for the C1 rerank, ncu) — the chain launches it only on request.
template <bool LAT, int NI, bool SH = false, bool COOP = false>
__global__ void __launch_bounds__(RR_WARPS * 32)
knn_rerank_kernel(const RerankParams p) {
  __shared__ int s_surv[RR_WARPS][RR_SURV];
  __shared__ float s_val[RR_WARPS][RR_SURV];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t row = (int64_t)blockIdx.x * RR_WARPS + wib;
  const bool active = row < p.nq;
  if constexpr (LAT) {   // per-update chain (dependent launch): lists come from the screen
    pdl_wait();
    pdl_launch_dependents();
  }
  float dist = 0.f;
  float q[16];
  WarpTopK top;
  unsigned need = 0;
  bool done = false;
  // COOP: whether ANY row of this query tile lost a bank range is known from the flags the latency
  // screen wrote (the 4 rows of a block lie in one 128-row tile, so every warp reads the same
  // words: block-uniform without a barrier).  Loaded first: the round trip hides behind the row.
  int f0 = 1, f1 = 0;
  if constexpr (LAT && COOP) {
    if (p.coop_flags != nullptr) {
      const int m_tile = (int)(((int64_t)blockIdx.x * RR_WARPS) / SCR_M);
      const int nflag = p.splits * 4;                         // <= 64
      const int* fl = p.coop_flags + (int64_t)m_tile * nflag;
      f0 = lane < nflag ? __ldcg(fl + lane) : 0;
      f1 = lane + 32 < nflag ? __ldcg(fl + lane + 32) : 0;
    }
  }
  if (active)
    done = rerank_row<LAT, NI, SH, COOP>(p, row, lane, s_surv[wib], s_val[wib], q, top, need);
  if constexpr (LAT && COOP) {
    // the common case skips the scan and its barriers
    const bool any_lost = __any_sync(0xffffffffu, (f0 | f1) != 0);
    if (any_lost) {
      __shared__ __align__(16) float s_tile[2 * RR_COOP_ROWS * EXB_SCREEN_MAX_DIM];
      __shared__ unsigned s_need[RR_WARPS];
      // copies for the out-of-line call: q and top themselves stay in registers on the common
      // path (a reference to them would put them in local memory for the whole kernel)
      float qq[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) qq[i] = q[i];
      WarpTopK tt = top;
      coop_scan<NI>(p, qq, tt, active ? need : 0u, lane, wib, s_tile, s_need);
      top = tt;
    }
  }
  if (active && done) dist = rerank_store(p, row, lane, top);
  if constexpr (LAT) {
    if (!p.pbe.on) return;
    const FusedPbe& f = p.pbe;
    const int k = p.k;
    if (!f.use_rms) {
      // no running moments: every warp finishes its own row (same order of operations as pbe_row)
      if (active) {
        float acc = 0.f, r = 0.f;
        for (int j = 0; j < k; ++j) {
          float v = __shfl_sync(0xffffffffu, dist, j);
          if (f.use_clip) v = clip_nan(v - f.clip);
          acc += v;
          r = v;
        }
        if (lane == 0)
          f.out_reward[row] = f.raw_kth ? r : logf((f.avg ? acc / (float)k : r) + 1.0f);
      }
      return;
    }
    __shared__ double2 s_part[RR_WARPS];
    __shared__ double2 s_red[RR_WARPS];
    __shared__ bool s_last;
    double2 acc = make_double2(0.0, 0.0);
    if (active && (f.avg ? lane < k : lane == k - 1)) {
      acc.x = (double)dist;
      acc.y = (double)dist * (double)dist;
    }
    acc = warp_sum2(acc);
    if (lane == 0) s_part[wib] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      double2 t = s_part[0];
      for (int w = 1; w < RR_WARPS; ++w) { t.x += s_part[w].x; t.y += s_part[w].y; }
      f.partials[blockIdx.x] = t;
      __threadfence();
      s_last = atomicAdd(&p.status->ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double2 a = make_double2(0.0, 0.0);
    for (int i = threadIdx.x; i < (int)gridDim.x; i += RR_WARPS * 32) {
      const double2 t = __ldcg(f.partials + i);
      a.x += t.x;
      a.y += t.y;
    }
    const double2 all = block_sum2<RR_WARPS * 32>(a, s_red);
    if (threadIdx.x == 0) {
      const double sums[3] = {(double)(f.avg ? p.nq * k : p.nq), all.x, all.y};
      rms_apply(sums, f.rms);   // M', S' for the per-row finish kernel that follows
    }
  }
}

}  // namespace exb
