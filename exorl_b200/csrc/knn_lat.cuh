// Latency-mode screening kernel (included by knn_screen.cu after the list helpers).
//
// The per-update sizes of the reference (B = 1024 reps against themselves, utils/utils.py:290-300;
// 1024 z against a 2048-row queue, proto.py:115-116) are pure latency: ~1 GFLOP.  The steady-state
// kernel would give each CTA pair one 256-row bank tile and leave 116 SMs idle with a single
// epilogue warp per scheduler grinding through 256 columns.  Here instead:
//   grid = (query tiles of 128) x (bank splits), bank tiles of LAT_N = 64 rows, so that a
//   1024 x 1024 problem fills 128 SMs;
//   the whole K extent of the query tile AND of a bank tile is resident in shared memory, so every
//   TMA load of the CTA is in flight at once (no ring; the MMA of chunk kc starts when chunk kc of
//   both operands has landed);
//   cta_group::1 MMA, M = 128, N = 64, K = 16; two 64-column TMEM accumulators;
//   epilogue thread = query row: its 64 scores go through a branch-free sorted register chain of
//   KS stages (exact k-th best so far), then every score >= tkeep is appended to the row's list.
// Output format = the steady-state kernel's (cand_val / cand_idx / cand_cnt per (row, split)), so
// the rerank kernel is shared.  EB / HBmax (maxima over the bank of the rounding error and of the
// half norm) are reduced by the epilogue warps from the per-row arrays while the loads and MMAs are
// in flight: no atomics, no memset in front of the prep kernel.
#pragma once

namespace exb {

constexpr int LAT_N = 64;
constexpr int LAT_B_CHUNK_BYTES = LAT_N * SCR_KC * 2;   // 8 KB
constexpr int LAT_THREADS = 192;
constexpr int LAT_TMEM_COLS = 128;
constexpr int LAT_LS = SCR_M + 1;   // list row stride in shared memory: slot-major reads and
                                    // row-major writes are both bank-conflict free

struct LatParams {
  const float* ha;   // [nq]
  const float* ea;   // [nq]
  const float* hb;   // [nb]
  const float* eb;   // [nb]
  int64_t nq, nb;
  int nkc, n_tiles, tiles_per_split, splits, k;
  float* cand_val;
  int* cand_idx;
  int* cand_cnt;
  float* cand_tk;    // [nq, splits] lower bound of the split's k-th best score (rerank prefilter)
  float* cand_tr;    // [nq, splits] the split's rsel-th best pair maximum (rerank prefilter)
  int rsel;          // ceil(k / splits)
  float* bank_max;   // [2] out: EB, HBmax (written by block (0,0) for the rerank kernel)
  long long* trace;  // optional phase timestamps (debug)
  unsigned long long* zero_status;   // caller-prepared operands: nobody ran prep (which zeroes the
                                     // 24-byte rerank status), so block (0,0) does it here
  int* coop_flags;   // [m_tiles, splits, 4]: every epilogue warp writes (unconditionally, so that
                     // nothing has to be zeroed) whether one of its rows lost a bank range — list
                     // overflow, operand outside the fp16 range; the rerank blocks of the query
                     // tile then scan cooperatively (knn_rerank.cuh:coop_scan), else they skip it
};

template <int KS>
__device__ __forceinline__ void chain_push_ks(float (&tv)[KS], float x) {
#pragma unroll
  for (int i = 0; i < KS; ++i) {
    const float hi = fmaxf(tv[i], x);
    x = fminf(tv[i], x);
    tv[i] = hi;
  }
}

// The kernel body as a function of (query tile, bank split): knn_screen_lat_kernel runs it on a
// (m_tiles, splits) grid of LAT_THREADS threads; the fused per-update chain (knn_chain.cuh) runs
// it as its second phase on 256-thread blocks (warps >= 6 only take part in the block barriers).
// CHAINED: no programmatic-dependent-launch calls, a proxy fence in front of the TMA loads (the
// operands were written by other blocks of the SAME kernel before a grid barrier).
template <int KS, bool CHAINED>
__device__ __forceinline__ void lat_screen_body(const CUtensorMap& tmap_q, const CUtensorMap& tmap_b,
                                                const LatParams& p, uint8_t* smem_raw,
                                                const int m_tile, const int split) {
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (base - raw);
  const int nkc = p.nkc;
  const uint32_t a_off = 0;
  const uint32_t b_off = a_off + (uint32_t)nkc * SCR_A_CHUNK_BYTES;
  const uint32_t lv_off = b_off + (uint32_t)nkc * LAT_B_CHUNK_BYTES;
  const uint32_t li_off = lv_off + (SCR_CAP + 1) * LAT_LS * 4;
  const uint32_t hn_off = (li_off + (SCR_CAP + 1) * LAT_LS * 4 + 15u) & ~15u;
  const uint32_t red_off = hn_off + 2 * LAT_N * 4;
  const uint32_t bar_off = red_off + 16 * 4;
  float* s_lv = reinterpret_cast<float*>(sm + lv_off);
  int* s_li = reinterpret_cast<int*>(sm + li_off);
  float* s_hn = reinterpret_cast<float*>(sm + hn_off);
  float* s_red = reinterpret_cast<float*>(sm + red_off);
  // barrier slots: [0,8) full[kc], 8 b_empty, 9-10 tfull, 11-12 tempty, then the TMEM address
  const uint32_t bar_base = base + bar_off;
  auto bar_full = [&](int kc) { return bar_base + 8u * kc; };
  const uint32_t bar_bempty = bar_base + 8u * 8;
  auto bar_tfull = [&](int s) { return bar_base + 8u * (9 + s); };
  auto bar_tempty = [&](int s) { return bar_base + 8u * (11 + s); };
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(sm + bar_off + 8 * 13);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t_begin = split * p.tiles_per_split;
  const int t_end = min(p.n_tiles, t_begin + p.tiles_per_split);
  const int n_local = max(t_end - t_begin, 0);
  const bool tr = p.trace != nullptr && m_tile == 0 && split == 0;
  if (tr && threadIdx.x == 0) p.trace[0] = clock_now();

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_q);
    tma_prefetch_desc(&tmap_b);
    for (int kc = 0; kc < nkc; ++kc) mbar_init(bar_full(kc), 1);
    mbar_init(bar_bempty, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_tfull(s), 1);
      mbar_init(bar_tempty(s), 4);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_cols(smem_u32(s_tmem), LAT_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  // barriers, TMEM and descriptors are ready: let the rerank's blocks be scheduled, then wait for
  // the producer of the operands (prep / rep producer) to finish before the first global read
  if constexpr (!CHAINED) {
    pdl_launch_dependents();
    pdl_wait();
  }

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0 && n_local > 0) {
      if constexpr (CHAINED) fence_proxy_async_all();   // generic-proxy writes -> TMA reads
      for (int j = 0; j < n_local; ++j) {
        const int brow = (t_begin + j) * LAT_N;
        if (j > 0) mbar_wait(bar_bempty, (uint32_t)(j - 1) & 1u);   // MMAs of tile j-1 done
        for (int kc = 0; kc < nkc; ++kc) {
          mbar_expect_tx(bar_full(kc), LAT_B_CHUNK_BYTES + (j == 0 ? SCR_A_CHUNK_BYTES : 0));
          if (j == 0)
            tma_load_2d(base + a_off + kc * SCR_A_CHUNK_BYTES, &tmap_q, kc * SCR_KC,
                        m_tile * SCR_M, bar_full(kc));
          tma_load_2d(base + b_off + kc * LAT_B_CHUNK_BYTES, &tmap_b, kc * SCR_KC, brow,
                      bar_full(kc));
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0 && n_local > 0) {
      for (int j = 0; j < n_local; ++j) {
        const int as = j & 1;
        if (j >= 2) {
          mbar_wait(bar_tempty(as), (((uint32_t)j >> 1) & 1u) ^ 1u);
          tc_fence_after();
        }
        const uint32_t d_tmem = tmem_base + (uint32_t)as * LAT_N;
        for (int kc = 0; kc < nkc; ++kc) {
          mbar_wait(bar_full(kc), (uint32_t)j & 1u);
          tc_fence_after();
          if (tr && j == 0 && kc == 0) p.trace[1] = clock_now();
          const uint64_t adesc = make_sw128_desc(base + a_off + kc * SCR_A_CHUNK_BYTES);
          const uint64_t bdesc = make_sw128_desc(base + b_off + kc * LAT_B_CHUNK_BYTES);
#pragma unroll
          for (int ki = 0; ki < SCR_KC / 16; ++ki)
            umma_f16(d_tmem, adesc + (uint64_t)(2 * ki), bdesc + (uint64_t)(2 * ki),
                     umma_idesc_f16(SCR_M, LAT_N), (kc | ki) != 0 ? 1u : 0u);
        }
        umma_commit(bar_bempty);
        umma_commit(bar_tfull(as));
        if (tr && j == 0) p.trace[2] = clock_now();
      }
    }
  } else if (warp < 6) {
    // ===================== epilogue: exact running k-th best + candidate lists =====================
    const int quad = warp & 3;                    // TMEM lane quadrant this warp may access
    const int e = quad * 32 + lane;               // row inside the tile
    const int et = (warp - 2) * 32 + lane;        // 0..127 linear epilogue thread id
    const int64_t row = (int64_t)m_tile * SCR_M + e;
    const bool row_ok = row < p.nq;
    float* lv = s_lv + e;
    int* li = s_li + e;

    // maxima over the whole bank of the rounding error and the half norm (while loads fly)
    float eb_max = 0.f, hb_max = 0.f;
    for (int64_t i = et; i < p.nb; i += 128) {
      eb_max = fmaxf(eb_max, __ldg(p.eb + i));
      hb_max = fmaxf(hb_max, __ldg(p.hb + i));
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      eb_max = fmaxf(eb_max, __shfl_xor_sync(0xffffffffu, eb_max, off));
      hb_max = fmaxf(hb_max, __shfl_xor_sync(0xffffffffu, hb_max, off));
    }
    if (lane == 0) {
      s_red[warp - 2] = eb_max;
      s_red[4 + warp - 2] = hb_max;
    }
    // first tile's bank half norms: thread et < 64 owns column et
    if (n_local > 0 && et < LAT_N) {
      const int64_t c0 = (int64_t)t_begin * LAT_N + et;
      s_hn[et] = (c0 < p.nb) ? __ldg(p.hb + c0) : INFINITY;
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    eb_max = fmaxf(fmaxf(s_red[0], s_red[1]), fmaxf(s_red[2], s_red[3]));
    hb_max = fmaxf(fmaxf(s_red[4], s_red[5]), fmaxf(s_red[6], s_red[7]));
    if (et == 0 && m_tile == 0 && split == 0) {
      p.bank_max[0] = eb_max;
      p.bank_max[1] = hb_max;
      if (p.zero_status != nullptr) {
        p.zero_status[0] = 0ull; p.zero_status[1] = 0ull; p.zero_status[2] = 0ull;
      }
    }

    ListState st;   // tv[] of ListState unused here: the chain is the KS-stage array below
    st.n = 0;
    st.dropped = -FLT_MAX;
    st.tk = -FLT_MAX;
    st.tkeep = -FLT_MAX;
    st.ha = row_ok ? p.ha[row] : 0.f;
    st.e2 = 2.f * ((row_ok ? p.ea[row] : 0.f) + eb_max);
    st.dl = (st.ha + hb_max) * SCR_ACC_EPS;
    float tv[KS];
#pragma unroll
    for (int i = 0; i < KS; ++i) tv[i] = (i < KS - p.k) ? FLT_MAX : -FLT_MAX;

    for (int j = 0; j < n_local; ++j) {
      const int t = t_begin + j;
      const int as = j & 1;
      const float* hn = s_hn + as * LAT_N;
      if (j + 1 < n_local && et < LAT_N) {   // next tile's half norms into the other buffer
        const int64_t c0 = (int64_t)(t + 1) * LAT_N + et;
        s_hn[(as ^ 1) * LAT_N + et] = (c0 < p.nb) ? __ldg(p.hb + c0) : INFINITY;
      }
      mbar_wait(bar_tfull(as), ((uint32_t)j >> 1) & 1u);
      tc_fence_after();
      if (tr && j == 0 && et == 0) p.trace[3] = clock_now();
      const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)as * LAT_N;
      uint32_t ra[32], rb[32];
      tmem_ld32(t_row, ra);
      tmem_ld32(t_row + 32, rb);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty(as));   // accumulator is in registers: free it

      float sc[2 * 32];
      const float4* hn4 = reinterpret_cast<const float4*>(hn);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float4 h = hn4[q], g = hn4[8 + q];
        sc[4 * q] = __uint_as_float(ra[4 * q]) - h.x;
        sc[4 * q + 1] = __uint_as_float(ra[4 * q + 1]) - h.y;
        sc[4 * q + 2] = __uint_as_float(ra[4 * q + 2]) - h.z;
        sc[4 * q + 3] = __uint_as_float(ra[4 * q + 3]) - h.w;
        sc[32 + 4 * q] = __uint_as_float(rb[4 * q]) - g.x;
        sc[32 + 4 * q + 1] = __uint_as_float(rb[4 * q + 1]) - g.y;
        sc[32 + 4 * q + 2] = __uint_as_float(rb[4 * q + 2]) - g.z;
        sc[32 + 4 * q + 3] = __uint_as_float(rb[4 * q + 3]) - g.w;
      }
      // Chain fed with maxima of column pairs: its k-th entry is the k-th best of 32 real scores,
      // a lower bound of the true k-th best (conservative threshold) at half the min/max work.
      // Pairs are (c, c + 32), not neighbours: in a batch of consecutive trajectory states a row's
      // k nearest are adjacent columns, and adjacent pairs would hide half of them (the chain's
      // k-th entry drops to about the 2k-th best and the list overflows its capacity).
#pragma unroll
      for (int c = 0; c < 32; ++c) chain_push_ks<KS>(tv, fmaxf(sc[c], sc[32 + c]));
      st.tk = tv[KS - 1];
      if (st.tk > -FLT_MAX) st.tkeep = fmaxf(st.tkeep, keep_threshold(st));
      if (tr && j == 0 && et == 0) p.trace[4] = clock_now();
      if (j > 0 && st.n > SCR_CAP - 14) list_compact<LAT_LS>(st, lv, li);
      // branch-free appends: slot SCR_CAP is a trash row for hits beyond the capacity
      const float tkeep = st.tkeep;
      int n = st.n;
      float dropped = st.dropped;
#pragma unroll
      for (int c = 0; c < 2 * 32; ++c) {
        const bool hit = sc[c] >= tkeep;
        const int slot = min(n, SCR_CAP);
        if (hit) {
          lv[slot * LAT_LS] = sc[c];
          li[slot * LAT_LS] = t * LAT_N + c;
        }
        dropped = (hit && n >= SCR_CAP) ? fmaxf(dropped, sc[c]) : dropped;   // vs the final tkeep
        n += hit ? 1 : 0;
      }
      st.n = min(n, SCR_CAP);
      st.dropped = dropped;
      if (tr && j == 0 && et == 0) p.trace[5] = clock_now();
      // the next iteration rewrites s_hn[as ^ 1 ^ 1] two tiles later: order the 64 writers/readers
      asm volatile("bar.sync 1, 128;" ::: "memory");
    }

    if (n_local > 1) list_compact<LAT_LS>(st, lv, li);
    const bool overflow = st.dropped > -FLT_MAX && st.dropped >= st.tkeep;
    if (p.coop_flags != nullptr) {
      const bool lost = __any_sync(0xffffffffu, row_ok && (overflow || !(st.e2 < INFINITY)));
      if (lane == 0) p.coop_flags[(m_tile * p.splits + split) * 4 + (warp - 2)] = lost ? 1 : 0;
    }
    if (row_ok) {
      const int64_t o = row * p.splits + split;
      p.cand_cnt[o] = overflow ? -1 : st.n;
      p.cand_tk[o] = st.tk;
      // right-aligned chain: slot KS - k holds the best, KS - k + r the (r+1)-th best
      float tr = -FLT_MAX;
#pragma unroll
      for (int i = 0; i < KS; ++i)
        if (i == KS - p.k + p.rsel - 1) tr = tv[i];
      p.cand_tr[o] = tr;
    }
    // coalesced list output: the warp writes one row's slots (96 contiguous bytes) per step
    __syncwarp();
    if (tr && et == 0) p.trace[7] = clock_now();
    // (32-bit element offsets — the per-update sizes are far below 2^32 list slots — and the
    // addresses formed outside the predicate: with 64-bit index math inside a divergent region
    // per row this loop was 1000 instructions per warp, 2.6 us of the kernel's 6.9)
    const int64_t row0 = (int64_t)m_tile * SCR_M + quad * 32;
    const int rows_here = (int)max((int64_t)0, min((int64_t)32, p.nq - row0));
    const uint32_t rstride = (uint32_t)(p.splits * SCR_CAP);
    const uint32_t o0 = ((uint32_t)row0 * (uint32_t)p.splits + (uint32_t)split) * SCR_CAP + lane;
    const int sl = min(lane, SCR_CAP) * LAT_LS + quad * 32;
#pragma unroll
    for (int r = 0; r < 32; ++r) {
      const int n_r = __shfl_sync(0xffffffffu, st.n, r);
      const float v = s_lv[sl + r];
      const int ix = s_li[sl + r];
      float* pv = p.cand_val + (o0 + r * rstride);
      int* pi = p.cand_idx + (o0 + r * rstride);
      if (r < rows_here && lane < n_r) {
        *pv = v;
        *pi = ix;
      }
    }
    if (tr && et == 0) p.trace[6] = clock_now();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_cols(tmem_base, LAT_TMEM_COLS);
  }
}

template <int KS>
__global__ void __launch_bounds__(LAT_THREADS, 1)
knn_screen_lat_kernel(const __grid_constant__ CUtensorMap tmap_q,
                      const __grid_constant__ CUtensorMap tmap_b, const LatParams p) {
  extern __shared__ uint8_t smem_raw[];
  lat_screen_body<KS, false>(tmap_q, tmap_b, p, smem_raw, (int)blockIdx.x, (int)blockIdx.y);
}

}  // namespace exb
