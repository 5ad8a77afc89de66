// n-step replay gather over a device-resident flat buffer + counter-based index draw.
//
// Replaces the per-sample python of /root/reference/utils/replay_buffer.py:223-235
// (ReplayBuffer._sample), :87-96 (OfflineReplayBuffer._sample) and the collate / pin / H2D of
// :246-277.  HBM-bound byte work: every (sample, column) row is copied with the widest aligned
// vector the layout allows (16 B for walker obs = 96 B and pixel frames = 63 504 B), one thread
// group per row, groups sized so that small rows still fill warps.
#include "common.cuh"
#include <stdlib.h>

namespace exb {

constexpr int GATHER_MAX_COLS = 8;
constexpr int GATHER_THREADS = 256;

struct GatherCols {
  exb_gather_col c[GATHER_MAX_COLS];
  int vec[GATHER_MAX_COLS];    // bytes per vector access: 16, 8, 4 or 1
  int group[GATHER_MAX_COLS];  // threads cooperating on one row: 32 or 256
  int n;
};

template <typename V>
__device__ __forceinline__ void copy_row(const char* __restrict__ s, char* __restrict__ d,
                                         int64_t bytes, int t, int group) {
  const int64_t nv = bytes / (int64_t)sizeof(V);
  const V* sv = reinterpret_cast<const V*>(s);
  V* dv = reinterpret_cast<V*>(d);
  int64_t i = t;
  // 4 independent loads in flight per thread before the stores (latency hiding on long rows)
  for (; i + 3 * (int64_t)group < nv; i += 4 * (int64_t)group) {
    const V a = __ldg(sv + i), b = __ldg(sv + i + group), c = __ldg(sv + i + 2 * group),
            e = __ldg(sv + i + 3 * group);
    dv[i] = a;
    dv[i + group] = b;
    dv[i + 2 * group] = c;
    dv[i + 3 * group] = e;
  }
  for (; i < nv; i += group) dv[i] = __ldg(sv + i);
}

// grid.y = column id (n == reward/discount pseudo-column), grid.x covers the batch.
__global__ void __launch_bounds__(GATHER_THREADS)
replay_gather_kernel(const int64_t* __restrict__ rows, int64_t batch, GatherCols cols,
                     const float* __restrict__ reward_src, const float* __restrict__ discount_src,
                     int nstep, int nstep_mode, float gamma, float* __restrict__ out_reward,
                     float* __restrict__ out_discount) {
  const int col = blockIdx.y;
  if (col == cols.n) {
    // utils/replay_buffer.py:229-234 — fp32, multiply then add (no fused multiply-add), in order.
    const int64_t b = (int64_t)blockIdx.x * GATHER_THREADS + threadIdx.x;
    if (b >= batch) return;
    const int64_t row = rows[b];
    float r, g;
    if (nstep_mode) {
      r = 0.f;
      g = 1.f;
      for (int i = 0; i < nstep; ++i) {
        r = __fadd_rn(r, __fmul_rn(g, reward_src[row + i]));
        g = __fmul_rn(g, __fmul_rn(discount_src[row + i], gamma));
      }
    } else {  // :94-95
      r = reward_src[row];
      g = __fmul_rn(discount_src[row], gamma);
    }
    out_reward[b] = r;
    out_discount[b] = g;
    return;
  }
  const int group = cols.group[col];
  const int per_block = GATHER_THREADS / group;
  const int64_t b = (int64_t)blockIdx.x * per_block + threadIdx.x / group;
  if (b >= batch) return;
  const int t = threadIdx.x % group;
  const exb_gather_col c = cols.c[col];
  const int64_t srow = rows[b] + c.row_offset;
  const char* s = static_cast<const char*>(c.src) + srow * c.row_bytes;
  char* d = static_cast<char*>(c.dst) + b * c.row_bytes;
  switch (cols.vec[col]) {
    case 16: copy_row<uint4>(s, d, c.row_bytes, t, group); break;
    case 8: copy_row<uint2>(s, d, c.row_bytes, t, group); break;
    case 4: copy_row<uint32_t>(s, d, c.row_bytes, t, group); break;
    default: copy_row<uint8_t>(s, d, c.row_bytes, t, group); break;
  }
}

// ------------------------------------------------------------------------------------------
// Long rows (pixel frames: 63 504 B): DMA-style copy with the bulk async-copy engine (TMA 1-D).
// One thread per CTA drives a ring of 16 KB shared-memory stages: global -> smem (mbarrier
// complete_tx) -> global (bulk group).  No register traffic, full-line reads and writes,
// (STAGES - 1) x 16 KB of loads in flight per CTA, two CTAs per SM, persistent over the chunks.
// ------------------------------------------------------------------------------------------
constexpr int BULK_CHUNK = 16384;
constexpr int BULK_STAGES = 6;
constexpr int BULK_MAX_COLS = 4;

struct BulkCols {
  exb_gather_col c[BULK_MAX_COLS];
  int n;
  int chunks_per_row[BULK_MAX_COLS];
  int chunk_base[BULK_MAX_COLS + 1];   // prefix sum of chunks per sample over the columns
};

__device__ __forceinline__ uint32_t g_smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

__global__ void __launch_bounds__(32)
replay_gather_bulk_kernel(const int64_t* __restrict__ rows, int64_t batch, BulkCols cols) {
  extern __shared__ __align__(128) uint8_t bulk_smem[];
  __shared__ __align__(8) uint64_t full[BULK_STAGES];
  if (threadIdx.x != 0) return;
  const int per_sample = cols.chunk_base[cols.n];
  const int64_t total = batch * per_sample;
  const int64_t first = blockIdx.x, stride = gridDim.x;
  const int64_t n_mine = first < total ? (total - first + stride - 1) / stride : 0;
  if (n_mine == 0) return;
  for (int s = 0; s < BULK_STAGES; ++s)
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(g_smem_u32(&full[s])));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");

  auto locate = [&](int64_t i, const char*& src, char*& dst, uint32_t& bytes) {
    const int64_t id = first + i * stride;
    const int64_t b = id / per_sample;
    int r = (int)(id - b * per_sample);
    int col = 0;
    while (r >= cols.chunk_base[col + 1]) ++col;
    const int piece = r - cols.chunk_base[col];
    const exb_gather_col c = cols.c[col];
    const int64_t off = (int64_t)piece * BULK_CHUNK;
    const int64_t left = c.row_bytes - off;
    bytes = (uint32_t)(left < BULK_CHUNK ? left : BULK_CHUNK);
    src = static_cast<const char*>(c.src) + (rows[b] + c.row_offset) * c.row_bytes + off;
    dst = static_cast<char*>(c.dst) + b * c.row_bytes + off;
  };
  auto load = [&](int64_t i) {
    const char* src;
    char* dst;
    uint32_t bytes;
    locate(i, src, dst, bytes);
    const int s = (int)(i % BULK_STAGES);
    const uint32_t bar = g_smem_u32(&full[s]);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(g_smem_u32(bulk_smem + (size_t)s * BULK_CHUNK)), "l"(src), "r"(bytes), "r"(bar)
        : "memory");
  };
  for (int64_t i = 0; i < BULK_STAGES - 1 && i < n_mine; ++i) load(i);
  for (int64_t i = 0; i < n_mine; ++i) {
    const int s = (int)(i % BULK_STAGES);
    const uint32_t bar = g_smem_u32(&full[s]);
    const uint32_t parity = (uint32_t)((i / BULK_STAGES) & 1);
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "BULK_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra BULK_DONE;\n"
        "bra BULK_WAIT;\n"
        "BULK_DONE:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
    const char* src;
    char* dst;
    uint32_t bytes;
    locate(i, src, dst, bytes);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst),
                 "r"(g_smem_u32(bulk_smem + (size_t)s * BULK_CHUNK)), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    const int64_t j = i + BULK_STAGES - 1;
    if (j < n_mine) {
      // stage of chunk j was last used by chunk i - 1: its store must have finished READING smem
      asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      load(j);
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
  c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}

__global__ void replay_draw_kernel(uint64_t seed, uint64_t call,
                                   const int64_t* __restrict__ eps_start,
                                   const int32_t* __restrict__ eps_len, int64_t n_episodes,
                                   int nstep, int64_t batch, int64_t* __restrict__ out_rows,
                                   int32_t* __restrict__ out_episode) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= batch) return;
  uint32_t c[4] = {(uint32_t)s, (uint32_t)call, (uint32_t)(call >> 32), 0u};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  const int64_t ep = (int64_t)(((uint64_t)c[0] * (uint64_t)n_episodes) >> 32);
  const uint64_t span = (uint64_t)(eps_len[ep] - nstep + 1);
  const int64_t idx = (int64_t)(((uint64_t)c[1] * span) >> 32) + 1;
  out_rows[s] = eps_start[ep] + idx;
  if (out_episode) out_episode[s] = (int32_t)ep;
}

}  // namespace exb

using namespace exb;

extern "C" int exb_replay_gather(const int64_t* rows, int64_t batch,
                                 const exb_gather_col* host_cols, int n_cols,
                                 const float* reward_src, const float* discount_src, int nstep,
                                 int nstep_mode, float gamma, float* out_reward,
                                 float* out_discount, exb_stream_t stream) {
  EXB_CHECK_ARG(rows && batch > 0, "bad rows/batch");
  EXB_CHECK_ARG(n_cols >= 0 && n_cols <= GATHER_MAX_COLS && (n_cols == 0 || host_cols),
                "n_cols=%d outside [0,%d]", n_cols, GATHER_MAX_COLS);
  EXB_CHECK_ARG(reward_src && discount_src && out_reward && out_discount, "null reward pointers");
  EXB_CHECK_ARG(nstep >= 1, "nstep must be >= 1");
  static const int bulk_enabled = [] {
    // Measured on B200 (C4, B = 1024 frames of 63 504 B, device time incl. the draw kernel):
    // 128-bit SIMT copies 54.3 us (74 % of the HBM copy peak), bulk-copy ring 60.2 us (67 %).
    // The SIMT path is therefore the default; EXB_GATHER_BULK=1 selects the bulk engine.
    const char* e = getenv("EXB_GATHER_BULK");
    return e ? atoi(e) : 0;
  }();
  GatherCols g;
  BulkCols bk;
  g.n = 0;
  bk.n = 0;
  bk.chunk_base[0] = 0;
  for (int i = 0; i < n_cols; ++i) {
    const exb_gather_col& c = host_cols[i];
    EXB_CHECK_ARG(c.src && c.dst && c.row_bytes > 0, "column %d: bad pointer or row_bytes", i);
    const uintptr_t a = (uintptr_t)c.src | (uintptr_t)c.dst | (uintptr_t)c.row_bytes;
    if (bulk_enabled && a % 16 == 0 && c.row_bytes >= 4096 && bk.n < BULK_MAX_COLS) {
      bk.c[bk.n] = c;
      bk.chunks_per_row[bk.n] = (int)((c.row_bytes + BULK_CHUNK - 1) / BULK_CHUNK);
      bk.chunk_base[bk.n + 1] = bk.chunk_base[bk.n] + bk.chunks_per_row[bk.n];
      ++bk.n;
      continue;
    }
    const int j = g.n++;
    g.c[j] = c;
    g.vec[j] = (a % 16 == 0) ? 16 : (a % 8 == 0) ? 8 : (a % 4 == 0) ? 4 : 1;
    g.group[j] = (c.row_bytes / g.vec[j] > 128) ? GATHER_THREADS : 32;
  }
  if (bk.n > 0) {
    static bool attr_set = false;
    const int smem = BULK_STAGES * BULK_CHUNK;
    if (!attr_set) {
      EXB_CUDA(cudaFuncSetAttribute(replay_gather_bulk_kernel,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      attr_set = true;
    }
    const int64_t total = batch * bk.chunk_base[bk.n];
    int64_t ctas = 2LL * sm_count();
    if (ctas > total) ctas = total;
    replay_gather_bulk_kernel<<<(unsigned)ctas, 32, smem, (cudaStream_t)stream>>>(rows, batch, bk);
    EXB_LAUNCH_CHECK();
  }
  n_cols = g.n;
  // grid.x must cover the column with the fewest rows per block and the reward pseudo-column
  const int64_t per_block_min = 1;  // group == GATHER_THREADS -> one row per block
  (void)per_block_min;
  int64_t gx = 0;
  for (int i = 0; i < n_cols; ++i) {
    const int64_t per_block = GATHER_THREADS / g.group[i];
    const int64_t need = (batch + per_block - 1) / per_block;
    if (need > gx) gx = need;
  }
  const int64_t need_r = (batch + GATHER_THREADS - 1) / GATHER_THREADS;
  if (need_r > gx) gx = need_r;
  dim3 grid((unsigned)gx, (unsigned)(n_cols + 1));
  replay_gather_kernel<<<grid, GATHER_THREADS, 0, (cudaStream_t)stream>>>(
      rows, batch, g, reward_src, discount_src, nstep, nstep_mode, gamma, out_reward,
      out_discount);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

extern "C" int exb_replay_draw(uint64_t seed, uint64_t call, const int64_t* eps_start,
                               const int32_t* eps_len, int64_t n_episodes, int nstep,
                               int64_t batch, int64_t* out_rows, int32_t* out_episode,
                               exb_stream_t stream) {
  EXB_CHECK_ARG(eps_start && eps_len && out_rows && n_episodes > 0 && batch > 0 && nstep >= 1,
                "bad arguments");
  replay_draw_kernel<<<(unsigned)((batch + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      seed, call, eps_start, eps_len, n_episodes, nstep, batch, out_rows, out_episode);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}
