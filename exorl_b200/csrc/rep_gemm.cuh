// Wide ICM trunk Linear on the tensor cores: pre[n, D] = obs[n, K] . W[D, K]^T for input widths
// whose W^T does not fit shared memory (pixel APT: K = 39 200 encoder features -> D = 512,
// agents/unsupervised_learning/icm_apt.py:17-18 with ddpg.py:23-24's encoder in front).
// Included by rep_producer.cu; LayerNorm + Tanh + operand emit follow in rep_sum_ln_tanh_kernel.
//
// fp32 parity (rep within 1e-5 of the reference's fp32 Linear) on tf32 tensor cores — "3xTF32":
//   x = hi + lo,  hi = x with the low 13 mantissa bits cleared (what kind::tf32 reads from a 32-bit
//   container), lo = x - hi (exact in fp32);  a.b ~= a_hi b_hi + a_hi b_lo + a_lo b_hi, the dropped
//   a_lo b_lo term and the truncation of lo to tf32 are <= 2^-20 |a b| each.
// Accumulation: the tensor core adds into fp32 TMEM accumulators; over K = 39 200 its per-add
// rounding would build up to ~1e-4 relative, so an accumulator only ever holds a WINDOW of
// TG_WIN_CHUNKS x 32 input columns (16 MMA k-steps): the main term and the two correction terms
// go to separate TMEM accumulators (the corrections are 2^-11 of the main term and would be lost
// in its roundings), drain warps read a finished window back with tcgen05.ld while the next one
// is being accumulated in the second TMEM buffer, and add it to per-thread fp32 registers with
// round-to-nearest adds.  Error model and measurement: DESIGN.md K6b.
//
// One CTA = one 128 x 128 output tile over one K split (split-K so that a 1024-row batch fills
// the GPU: 8 x 4 tiles x 4 splits = 128 CTAs); partial tiles go to part[split][n][D] and are
// summed in split order by the LayerNorm kernel (deterministic).
//   warp 0      TMA producer: per 32-column chunk obs tile, W tile, W_lo tile (128 rows x 128 B
//               each, SWIZZLE_128B) into a TG_STAGES-deep ring
//   warp 1      tcgen05.mma kind::tf32 issuer (one lane): 4 k-steps x 2 MMAs (N = 256, N = 128) per chunk
//   warps 2-5   converters: lo tile of the obs chunk, element-wise at the swizzled offsets
//               (shared -> shared, then fence.proxy.async)
//   warps 6-13  drain: TMEM -> registers per window, final store of the partial tile
#pragma once
#include "tc_ptx.cuh"
#include "tmap_host.cuh"

namespace exb {

constexpr int TG_M = 128;               // rows per tile
constexpr int TG_N = 128;               // output columns per tile
constexpr int TG_KC = 32;               // fp32 columns per chunk = one 128-byte swizzle row
constexpr int TG_TILE_BYTES = TG_M * TG_KC * 4;     // 16 KB
constexpr int TG_STAGE_BYTES = 4 * TG_TILE_BYTES;   // obs, obs_lo, W, W_lo
constexpr int TG_STAGES = 3;
constexpr int TG_WIN_CHUNKS = 4;        // chunks per accumulation window (16 k-steps of 8)
constexpr int TG_THREADS = 14 * 32;
constexpr int TG_CONV_THREADS = 128;
constexpr int TG_DRAIN_THREADS = 256;
constexpr int TG_SMEM_BYTES = TG_STAGES * TG_STAGE_BYTES + 1024 + 256;

// kind::tf32 instruction descriptor: D = F32 (bit 4), A = B = TF32 (2 at [7,10) and [10,13)),
// both K-major, N >> 3 at [17,23), M >> 4 at [24,29).
constexpr uint32_t umma_idesc_tf32(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// x - (x with the low 13 mantissa bits cleared): the part of x a tf32 operand does not carry.
__device__ __forceinline__ float tf32_lo(float x) {
  return x - __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
}

// W_lo for the whole weight (one pass per call; W itself is the hi operand as it lies).
__global__ void tf32_lo_kernel(const float4* __restrict__ w, int64_t n4, float4* __restrict__ lo) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (int64_t)gridDim.x * blockDim.x) {
    const float4 v = __ldg(w + i);
    lo[i] = make_float4(tf32_lo(v.x), tf32_lo(v.y), tf32_lo(v.z), tf32_lo(v.w));
  }
}

struct TrunkGemmParams {
  float* part;            // [splits][n][dim]
  int64_t n;
  int dim;                // output columns (rep_dim)
  int chunks;             // ceil(in_dim / 32)
  int chunks_per_split;
};

__global__ void __launch_bounds__(TG_THREADS, 1)
trunk_gemm_kernel(const __grid_constant__ CUtensorMap t_obs, const __grid_constant__ CUtensorMap t_w,
                  const __grid_constant__ CUtensorMap t_wlo, const TrunkGemmParams p) {
  extern __shared__ unsigned char tg_smem_raw[];
  const uint32_t raw = smem_u32(tg_smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;            // SWIZZLE_128B tiles: 1 KB aligned
  unsigned char* gen = tg_smem_raw + (base - raw);
  const uint32_t bars = base + TG_STAGES * TG_STAGE_BYTES;
  // barriers (8 bytes each): full[S], conv[S], empty[S], accfull[2], accempty[2], then tmem slot
  auto bar_full = [&](int s) { return bars + 8u * s; };
  auto bar_conv = [&](int s) { return bars + 8u * (TG_STAGES + s); };
  auto bar_empty = [&](int s) { return bars + 8u * (2 * TG_STAGES + s); };
  auto bar_accfull = [&](int b) { return bars + 8u * (3 * TG_STAGES + b); };
  auto bar_accempty = [&](int b) { return bars + 8u * (3 * TG_STAGES + 2 + b); };
  const uint32_t tmem_slot = bars + 8u * (3 * TG_STAGES + 4);
  volatile uint32_t* tmem_slot_gen =
      reinterpret_cast<volatile uint32_t*>(gen + TG_STAGES * TG_STAGE_BYTES + 8 * (3 * TG_STAGES + 4));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * TG_M, n0 = blockIdx.y * TG_N, split = blockIdx.z;
  const int c_begin = split * p.chunks_per_split;
  const int c_end = min(p.chunks, c_begin + p.chunks_per_split);
  const int n_chunks = c_end - c_begin;                    // >= 1 by construction of the grid
  const int n_win = (n_chunks + TG_WIN_CHUNKS - 1) / TG_WIN_CHUNKS;

  if (threadIdx.x == 0) {
    for (int s = 0; s < TG_STAGES; ++s) {
      mbar_init(bar_full(s), 1);
      mbar_init(bar_conv(s), TG_CONV_THREADS);
      mbar_init(bar_empty(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(bar_accfull(b), 1);
      mbar_init(bar_accempty(b), TG_DRAIN_THREADS);
    }
    fence_barrier_init();
    tma_prefetch_desc(&t_obs);
    tma_prefetch_desc(&t_w);
    tma_prefetch_desc(&t_wlo);
  }
  if (warp == 1) tmem_alloc_512(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_gen;

  if (warp == 0) {
    if (lane == 0) {
      for (int i = 0; i < n_chunks; ++i) {
        const int s = i % TG_STAGES;
        if (i >= TG_STAGES) mbar_wait(bar_empty(s), ((i / TG_STAGES) - 1) & 1);
        const uint32_t st = base + s * TG_STAGE_BYTES;
        mbar_expect_tx(bar_full(s), 3 * TG_TILE_BYTES);
        const int c = (c_begin + i) * TG_KC;
        tma_load_2d(st, &t_obs, c, m0, bar_full(s));
        tma_load_2d(st + 2 * TG_TILE_BYTES, &t_w, c, n0, bar_full(s));
        tma_load_2d(st + 3 * TG_TILE_BYTES, &t_wlo, c, n0, bar_full(s));
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // W and W_lo tiles are adjacent in the stage (256 rows x 128 B, 1 KB between 8-row
      // groups), and so are the main and correction accumulators in TMEM: ONE N = 256 MMA gives
      // a_hi . [w_hi ; w_lo]^T = {main, first correction} and reads the obs tile once; the
      // second correction a_lo . w_hi^T is an N = 128 MMA into the correction columns.
      constexpr uint32_t idesc2 = umma_idesc_tf32(TG_M, 2 * TG_N);
      constexpr uint32_t idesc = umma_idesc_tf32(TG_M, TG_N);
      for (int i = 0; i < n_chunks; ++i) {
        const int s = i % TG_STAGES;
        const int win = i / TG_WIN_CHUNKS, j = i - win * TG_WIN_CHUNKS;
        const int buf = win & 1;
        if (j == 0 && win >= 2) mbar_wait(bar_accempty(buf), ((win >> 1) - 1) & 1);
        mbar_wait(bar_full(s), (i / TG_STAGES) & 1);
        mbar_wait(bar_conv(s), (i / TG_STAGES) & 1);
        tc_fence_after();
        const uint32_t st = base + s * TG_STAGE_BYTES;
        const uint32_t d_main = tmem + buf * 256, d_corr = d_main + 128;
#pragma unroll
        for (int ks = 0; ks < TG_KC / 8; ++ks) {
          const uint64_t a_hi = make_sw128_desc(st + ks * 32);
          const uint64_t a_lo = make_sw128_desc(st + TG_TILE_BYTES + ks * 32);
          const uint64_t w_hi = make_sw128_desc(st + 2 * TG_TILE_BYTES + ks * 32);
          const uint32_t first = (j == 0 && ks == 0) ? 0u : 1u;
          umma_tf32(d_main, a_hi, w_hi, idesc2, first);
          umma_tf32(d_corr, a_lo, w_hi, idesc, 1u);
        }
        umma_commit(bar_empty(s));
        if (j == TG_WIN_CHUNKS - 1 || i == n_chunks - 1) umma_commit(bar_accfull(buf));
      }
    }
  } else if (warp < 6) {
    // converters: obs_lo tile = tf32_lo(obs tile), same (swizzled) offsets
    const int t = threadIdx.x - 64;
    for (int i = 0; i < n_chunks; ++i) {
      const int s = i % TG_STAGES;
      mbar_wait(bar_full(s), (i / TG_STAGES) & 1);
      const float4* src = reinterpret_cast<const float4*>(gen + s * TG_STAGE_BYTES);
      float4* dst = reinterpret_cast<float4*>(gen + s * TG_STAGE_BYTES + TG_TILE_BYTES);
#pragma unroll
      for (int u = 0; u < TG_TILE_BYTES / 16 / TG_CONV_THREADS; ++u) {
        const float4 v = src[u * TG_CONV_THREADS + t];
        dst[u * TG_CONV_THREADS + t] =
            make_float4(tf32_lo(v.x), tf32_lo(v.y), tf32_lo(v.z), tf32_lo(v.w));
      }
      fence_proxy_async();
      mbar_arrive(bar_conv(s));
    }
  } else {
    // drain: TMEM lanes of this warp's quadrant, 64 of the tile's 128 columns
    const int dw = warp - 6;                       // 0..7
    const int quad = warp & 3;                     // TMEM lane quadrant this warp may read
    const int col0 = (dw >> 2) * 64;               // warps 6-9: columns 0-63, 10-13: 64-127
    float acc[64];
#pragma unroll
    for (int c = 0; c < 64; ++c) acc[c] = 0.f;
    for (int w = 0; w < n_win; ++w) {
      const int buf = w & 1;
      mbar_wait(bar_accfull(buf), (w >> 1) & 1);
      tc_fence_after();
      const uint32_t t_main = tmem + ((uint32_t)(quad * 32) << 16) + buf * 256 + col0;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t a[32], b[32];
        tmem_ld32(t_main + h * 32, a);
        tmem_ld32(t_main + 128 + h * 32, b);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 32; ++c)
          acc[h * 32 + c] += __uint_as_float(a[c]) + __uint_as_float(b[c]);
      }
      tc_fence_before();
      mbar_arrive(bar_accempty(buf));
    }
    const int64_t row = (int64_t)m0 + quad * 32 + lane;
    const int col = n0 + col0;
    if (row < p.n) {
      float* o = p.part + ((int64_t)split * p.n + row) * p.dim + col;
#pragma unroll
      for (int c = 0; c < 64; c += 4)
        if (col + c < p.dim)       // dim % 4 == 0
          *reinterpret_cast<float4*>(o + c) = make_float4(acc[c], acc[c + 1], acc[c + 2], acc[c + 3]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc_512(tmem);
}

}  // namespace exb
