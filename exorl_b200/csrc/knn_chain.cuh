// The per-update reward as ONE kernel (included by knn_screen.cu after knn_lat.cuh / knn_rerank.cuh).
//
// PBE.__call__ on a 1024 x 512 batch (utils/utils.py:289-319) is ~1 GFLOP: the four kernels of the
// chain (prep -> latency screen -> rerank + moments -> per-row finish) run for ~17 us in total,
// but even replayed from a CUDA graph the call took 30.7 us — three kernel boundaries at 4-5 us
// each (grid drain, dependency resolution, ramp-up of the next grid).  Here the four stages are
// phases of one cooperative kernel with one block per SM, separated by a grid-wide barrier in
// global memory (~1 us): every block
//   0. prepares its share of the fp16 operand rows (prep_one_row; skipped for caller-prepared
//      operands),                                                                     | barrier
//   1. runs the latency screen of its (query tile, bank split) (lat_screen_body),     | barrier
//   2. reranks its share of the query rows, one warp per row, block-cooperative scan of lost
//      bank ranges, fp64 moment partials,                                             | barrier
//   3. (knn_rms) reduces the partials of all blocks in a fixed order — every block gets the same
//      sums and derives the same updated mean, block 0 stores the RMS state — and finishes its
//      own rows (utils/utils.py:301-318).
// The barrier is a counter + generation word pair: the last block to arrive resets the counter
// and bumps the generation the others spin on (ld.acquire), so the pair cleans itself up and a
// launch needs no memset.  Co-residency of the grid comes from cudaLaunchCooperativeKernel.
#pragma once

namespace exb {

constexpr int CH_WARPS = 8;
constexpr int CH_THREADS = CH_WARPS * 32;
// phase-2 shared memory (aliases the screen's operand tiles, dead by then)
constexpr int CH_P2_SMEM = CH_WARPS * RR_SURV * 8 + 2 * RR_COOP_ROWS * EXB_SCREEN_MAX_DIM * 4 +
                           CH_WARPS * 4 + 2 * CH_WARPS * 16 + 64;

struct ChainParams {
  // phase 0
  const float* bank;
  const float* src;
  int64_t nb, nq;
  int dim, dpad, same, do_prep, vec;
  __half* b16;
  float *hb, *eb;
  __half* q16;
  float *ha, *ea;
  // phase 1 / 2
  LatParams lp;
  RerankParams rp;
  int m_tiles, splits;
  unsigned* gbar;          // {arrivals, generation}
  long long* trace;        // debug (EXB_CHAIN_TRACE): 8 phase stamps of block 0, 8 of the last block
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// All blocks of the (co-resident) grid.  `gen` is meaningful in thread 0 only.
__device__ __forceinline__ void grid_barrier(unsigned* gbar, unsigned n_cta, unsigned& gen) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();                                    // this block's writes before the arrival
    const unsigned old = atomicAdd(gbar, 1u);
    if (old == n_cta - 1) {
      atomicExch(gbar, 0u);                             // clean for the next barrier / launch
      __threadfence();
      atomicAdd(gbar + 1, 1u);                          // release the others
    } else {
      while (ld_acquire_u32(gbar + 1) == gen) {
      }
    }
    ++gen;
    __threadfence();
  }
  __syncthreads();
}

// utils/utils.py:264-276 on scalars, values only (rms_apply stores them).
__device__ __forceinline__ void rms_values(double bs, double sum, double sumsq, double M, double S,
                                           double n, double* new_m, double* new_s, double* tot) {
  const double mean = sum / bs;
  const double var = (bs > 1.0) ? fmax(sumsq - sum * mean, 0.0) / (bs - 1.0) : (double)NAN;
  *tot = n + bs;
  const double delta = mean - M;
  *new_m = M + delta * bs / *tot;
  *new_s = (S * n + var * bs + delta * delta * n * bs / *tot) / *tot;
}

template <int KS, int NI>
__global__ void __launch_bounds__(CH_THREADS, 1)
apt_chain_kernel(const __grid_constant__ CUtensorMap tmap_q,
                 const __grid_constant__ CUtensorMap tmap_b, const ChainParams c) {
  extern __shared__ uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cta = blockIdx.x, n_cta = gridDim.x;
  unsigned gen = 0;
  if (threadIdx.x == 0) gen = ld_acquire_u32(c.gbar + 1);
  long long* tr = nullptr;
  if (c.trace != nullptr && threadIdx.x == 0 && (cta == 0 || cta == n_cta - 1))
    tr = c.trace + (cta == 0 ? 0 : 8);
  if (tr) tr[0] = clock_now();
  if (cta == 0 && threadIdx.x < 3)    // rerank status {n_fallback, invalid | pairs | ticket, pad}
    reinterpret_cast<unsigned long long*>(c.rp.status)[threadIdx.x] = 0ull;

  // ---------------- phase 0: fp16 operands ----------------
  if (c.do_prep) {
    for (int64_t row = (int64_t)cta * CH_WARPS + warp; row < c.nb; row += (int64_t)n_cta * CH_WARPS) {
      if (c.vec) prep_one_row<true>(c.bank, row, c.dim, c.dpad, c.b16, c.hb, c.eb, nullptr, nullptr, lane);
      else prep_one_row<false>(c.bank, row, c.dim, c.dpad, c.b16, c.hb, c.eb, nullptr, nullptr, lane);
    }
    if (!c.same) {
      for (int64_t row = (int64_t)cta * CH_WARPS + warp; row < c.nq;
           row += (int64_t)n_cta * CH_WARPS) {
        if (c.vec) prep_one_row<true>(c.src, row, c.dim, c.dpad, c.q16, c.ha, c.ea, nullptr, nullptr, lane);
        else prep_one_row<false>(c.src, row, c.dim, c.dpad, c.q16, c.ha, c.ea, nullptr, nullptr, lane);
      }
    }
    if (tr) tr[1] = clock_now();
    grid_barrier(c.gbar, n_cta, gen);
  }
  if (tr) tr[2] = clock_now();

  // ---------------- phase 1: latency screen ----------------
  if (cta < c.m_tiles * c.splits)
    lat_screen_body<KS, true>(tmap_q, tmap_b, c.lp, smem_raw, cta % c.m_tiles, cta / c.m_tiles);
  if (tr) tr[3] = clock_now();
  grid_barrier(c.gbar, n_cta, gen);
  if (tr) tr[4] = clock_now();

  // ---------------- phase 2: rerank, one warp per row ----------------
  const RerankParams& p = c.rp;
  uint8_t* sm = smem_raw + ((16u - (smem_u32(smem_raw) & 15u)) & 15u);
  int* s_surv = reinterpret_cast<int*>(sm);                                   // [W][RR_SURV]
  float* s_val = reinterpret_cast<float*>(s_surv + CH_WARPS * RR_SURV);       // [W][RR_SURV]
  float* s_tile = s_val + CH_WARPS * RR_SURV;                                 // 2 coop tiles
  unsigned* s_need = reinterpret_cast<unsigned*>(s_tile + 2 * RR_COOP_ROWS * EXB_SCREEN_MAX_DIM);
  double2* s_part = reinterpret_cast<double2*>(s_need + CH_WARPS);
  double2* s_red = s_part + CH_WARPS;
  float* s_mean = reinterpret_cast<float*>(s_red + CH_WARPS);
  const FusedPbe& f = p.pbe;
  const int k = p.k;
  const bool moments = f.on && f.use_rms;
  double old_m = 0.0, old_s = 1.0, old_n = 0.0;
  if (moments && threadIdx.x == 0) {   // before anyone can store the updated state (barrier 3)
    old_m = (double)*f.rms.m;
    old_s = (double)*f.rms.s;
    old_n = *f.rms.n;
  }
  double2 acc = make_double2(0.0, 0.0);
  for (int64_t row0 = (int64_t)cta * CH_WARPS; row0 < p.nq; row0 += (int64_t)n_cta * CH_WARPS) {
    const int64_t row = row0 + warp;
    const bool active = row < p.nq;
    float q[16];
    WarpTopK top;
    unsigned need = 0;
    if (active)
      rerank_row<true, NI, false, true>(p, row, lane, s_surv + warp * RR_SURV,
                                        s_val + warp * RR_SURV, q,
                                  top, need);
    {
      float qq[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) qq[i] = q[i];
      WarpTopK tt = top;
      coop_scan<NI, CH_WARPS>(p, qq, tt, active ? need : 0u, lane, warp, s_tile, s_need);
      top = tt;
    }
    float dist = 0.f;
    if (active) dist = rerank_store(p, row, lane, top);
    if (!f.on) continue;
    if (!f.use_rms) {
      // no running moments: every warp finishes its own row (same order of operations as pbe_row)
      if (active) {
        float a = 0.f, r = 0.f;
        for (int j = 0; j < k; ++j) {
          float v = __shfl_sync(0xffffffffu, dist, j);
          if (f.use_clip) v = clip_nan(v - f.clip);
          a += v;
          r = v;
        }
        if (lane == 0)
          f.out_reward[row] = f.raw_kth ? r : logf((f.avg ? a / (float)k : r) + 1.0f);
      }
    } else if (active && (f.avg ? lane < k : lane == k - 1)) {
      acc.x += (double)dist;
      acc.y += (double)dist * (double)dist;
    }
  }
  if (tr) tr[5] = clock_now();
  if (!moments) return;

  // ---------------- phase 3: global moments, RMS update, per-row finish ----------------
  acc = warp_sum2(acc);
  if (lane == 0) s_part[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double2 t = s_part[0];
    for (int w = 1; w < CH_WARPS; ++w) { t.x += s_part[w].x; t.y += s_part[w].y; }
    f.partials[cta] = t;
  }
  grid_barrier(c.gbar, n_cta, gen);
  if (tr) tr[6] = clock_now();
  double2 a = make_double2(0.0, 0.0);
  for (int i = threadIdx.x; i < n_cta; i += CH_THREADS) {
    const double2 t = __ldcg(f.partials + i);
    a.x += t.x;
    a.y += t.y;
  }
  const double2 all = block_sum2<CH_THREADS>(a, s_red);   // same order in every block
  if (threadIdx.x == 0) {
    double new_m, new_s, tot;
    rms_values((double)(f.avg ? p.nq * k : p.nq), all.x, all.y, old_m, old_s, old_n, &new_m,
               &new_s, &tot);
    *s_mean = (float)new_m;
    if (cta == 0) {
      *f.rms.m = (float)new_m;
      *f.rms.s = (float)new_s;
      *f.rms.n = tot;                       // utils/utils.py:274
    }
  }
  __syncthreads();
  const float m = *s_mean;
  for (int64_t row = (int64_t)cta * CH_WARPS + warp; row < p.nq; row += (int64_t)n_cta * CH_WARPS)
    if (lane == 0)
      f.out_reward[row] = pbe_row(p.out_dist + row * k, k, f.avg, 1.f, m, f.use_clip, f.clip);
  if (tr) tr[7] = clock_now();
}

}  // namespace exb
