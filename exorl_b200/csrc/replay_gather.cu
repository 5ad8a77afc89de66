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
constexpr int GATHER_UNROLL = 4;                                  // vectors in flight per thread
constexpr int GATHER_PIECE_VECS = GATHER_THREADS * GATHER_UNROLL;  // vectors per long-row piece

// Work list of one launch.  Columns are ordered long rows first; block ids are a flat prefix sum
// over the columns (+ the reward/discount pseudo-column last), so no block is launched idle.
//   long row  (more than 128 vectors): one block per PIECE of 1024 vectors (16 KB at 16 B/vector);
//                                      a 63 504-B frame is 4 pieces, a 1024-batch 8192 blocks.
//   short row: one warp per row, 8 rows per block.
struct GatherPlan {
  exb_gather_col c[GATHER_MAX_COLS];
  int vec[GATHER_MAX_COLS];      // bytes per vector access: 16, 8, 4 or 1
  int pieces[GATHER_MAX_COLS];   // 0 = short row (warp per row), else pieces per row
  unsigned block_base[GATHER_MAX_COLS + 2];
  int n;
};

// Optional fused index draw (replay_draw_kernel's arithmetic, below): rows == nullptr.
struct GatherDraw {
  uint64_t seed, call;
  const int64_t* eps_start;
  const int32_t* eps_len;
  int64_t n_episodes;
  int nstep;
  int64_t* out_rows;   // may be null
};

__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
  c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}

// Philox4x32-10, counter (sample, call_lo, call_hi, 0), key = seed.  episode ~ U{0..n_episodes-1}
// (random.choice, utils/replay_buffer.py:168-170), idx ~ U{1..eps_len-nstep+1} (:222; nstep = 1
// is the offline range of :90); returns the global row = episode start + idx.
__device__ __forceinline__ int64_t draw_row(int64_t s, uint64_t seed, uint64_t call,
                                            const int64_t* __restrict__ eps_start,
                                            const int32_t* __restrict__ eps_len,
                                            int64_t n_episodes, int nstep, int32_t* episode) {
  uint32_t c[4] = {(uint32_t)s, (uint32_t)call, (uint32_t)(call >> 32), 0u};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  const int64_t ep = (int64_t)(((uint64_t)c[0] * (uint64_t)n_episodes) >> 32);
  // An episode shorter than nstep has no valid start (the reference raises ValueError from
  // np.random.randint, :222); the host mirror raises before launching (replay.py:_check_spans).
  // The clamp only keeps a direct C caller's reads inside the episode's own rows.
  const int32_t span_i = __ldg(eps_len + ep) - nstep + 1;
  const uint64_t span = (uint64_t)(span_i > 1 ? span_i : 1);
  const int64_t idx = (int64_t)(((uint64_t)c[1] * span) >> 32) + 1;
  if (episode) *episode = (int32_t)ep;
  return __ldg(eps_start + ep) + idx;
}

template <typename V>
__device__ __forceinline__ V ld_stream(const V* p) { return __ldcs(p); }
template <>
__device__ __forceinline__ uint8_t ld_stream<uint8_t>(const uint8_t* p) { return __ldg(p); }

// One piece of a long row: GATHER_UNROLL independent vector loads per thread in flight (source
// rows are read once: streaming / evict-first), then the stores.
template <typename V>
__device__ __forceinline__ void copy_piece(const char* __restrict__ s, char* __restrict__ d,
                                           int64_t nv, int64_t v0, int t) {
  const V* sv = reinterpret_cast<const V*>(s);
  V* dv = reinterpret_cast<V*>(d);
  V r[GATHER_UNROLL];
#pragma unroll
  for (int u = 0; u < GATHER_UNROLL; ++u) {
    const int64_t i = v0 + t + u * GATHER_THREADS;
    if (i < nv) r[u] = ld_stream(sv + i);
  }
#pragma unroll
  for (int u = 0; u < GATHER_UNROLL; ++u) {
    const int64_t i = v0 + t + u * GATHER_THREADS;
    if (i < nv) dv[i] = r[u];
  }
}

// Short row: one warp, at most 128 vectors -> at most 4 per lane, all in flight.
template <typename V>
__device__ __forceinline__ void copy_short(const char* __restrict__ s, char* __restrict__ d,
                                           int nv, int lane) {
  const V* sv = reinterpret_cast<const V*>(s);
  V* dv = reinterpret_cast<V*>(d);
  V r[4];
#pragma unroll
  for (int u = 0; u < 4; ++u)
    if (lane + 32 * u < nv) r[u] = __ldg(sv + lane + 32 * u);
#pragma unroll
  for (int u = 0; u < 4; ++u)
    if (lane + 32 * u < nv) dv[lane + 32 * u] = r[u];
}

__global__ void __launch_bounds__(GATHER_THREADS)
replay_gather_kernel(const int64_t* __restrict__ rows, int64_t batch, GatherPlan plan,
                     GatherDraw draw, const float* __restrict__ reward_src,
                     const float* __restrict__ discount_src, int nstep, int nstep_mode,
                     float gamma, float* __restrict__ out_reward,
                     float* __restrict__ out_discount) {
  int col = 0;
  while (blockIdx.x >= plan.block_base[col + 1]) ++col;
  const unsigned local = blockIdx.x - plan.block_base[col];
  auto row_of = [&](int64_t b) -> int64_t {
    return rows ? rows[b]
                : draw_row(b, draw.seed, draw.call, draw.eps_start, draw.eps_len,
                           draw.n_episodes, draw.nstep, nullptr);
  };
  if (col == plan.n) {
    // utils/replay_buffer.py:229-234 — fp32, multiply then add (no fused multiply-add), in order.
    const int64_t b = (int64_t)local * GATHER_THREADS + threadIdx.x;
    if (b >= batch) return;
    const int64_t row = row_of(b);
    if (!rows && draw.out_rows) draw.out_rows[b] = row;
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
  const exb_gather_col c = plan.c[col];
  const int vec = plan.vec[col];
  const int pieces = plan.pieces[col];
  if (pieces == 0) {
    const int64_t b = (int64_t)local * (GATHER_THREADS / 32) + (threadIdx.x >> 5);
    if (b >= batch) return;
    const int lane = threadIdx.x & 31;
    const char* s = static_cast<const char*>(c.src) + (row_of(b) + c.row_offset) * c.row_bytes;
    char* d = static_cast<char*>(c.dst) + b * c.row_bytes;
    const int nv = (int)(c.row_bytes / vec);
    switch (vec) {
      case 16: copy_short<uint4>(s, d, nv, lane); break;
      case 8: copy_short<uint2>(s, d, nv, lane); break;
      case 4: copy_short<uint32_t>(s, d, nv, lane); break;
      default: copy_short<uint8_t>(s, d, nv, lane); break;
    }
    return;
  }
  const int64_t b = local / (unsigned)pieces;
  const int64_t v0 = (int64_t)(local % (unsigned)pieces) * GATHER_PIECE_VECS;
  const char* s = static_cast<const char*>(c.src) + (row_of(b) + c.row_offset) * c.row_bytes;
  char* d = static_cast<char*>(c.dst) + b * c.row_bytes;
  const int64_t nv = c.row_bytes / vec;
  switch (vec) {
    case 16: copy_piece<uint4>(s, d, nv, v0, threadIdx.x); break;
    case 8: copy_piece<uint2>(s, d, nv, v0, threadIdx.x); break;
    case 4: copy_piece<uint32_t>(s, d, nv, v0, threadIdx.x); break;
    default: copy_piece<uint8_t>(s, d, nv, v0, threadIdx.x); break;
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

__global__ void replay_draw_kernel(uint64_t seed, uint64_t call,
                                   const int64_t* __restrict__ eps_start,
                                   const int32_t* __restrict__ eps_len, int64_t n_episodes,
                                   int nstep, int64_t batch, int64_t* __restrict__ out_rows,
                                   int32_t* __restrict__ out_episode) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= batch) return;
  int32_t ep;
  out_rows[s] = draw_row(s, seed, call, eps_start, eps_len, n_episodes, nstep, &ep);
  if (out_episode) out_episode[s] = ep;
}

}  // namespace exb

using namespace exb;

static int launch_gather(const int64_t* rows, const GatherDraw& draw, int64_t batch,
                         const exb_gather_col* host_cols, int n_cols, const float* reward_src,
                         const float* discount_src, int nstep, int nstep_mode, float gamma,
                         float* out_reward, float* out_discount, exb_stream_t stream) {
  EXB_CHECK_ARG(batch > 0, "bad batch");
  EXB_CHECK_ARG(n_cols >= 0 && n_cols <= GATHER_MAX_COLS && (n_cols == 0 || host_cols),
                "n_cols=%d outside [0,%d]", n_cols, GATHER_MAX_COLS);
  EXB_CHECK_ARG(reward_src && discount_src && out_reward && out_discount, "null reward pointers");
  EXB_CHECK_ARG(nstep >= 1, "nstep must be >= 1");
  static const int bulk_enabled = [] {
    // Measured on B200 (C4, B = 1024 frames of 63 504 B, device time incl. the draw kernel):
    // 128-bit SIMT copies 54.3 us (74 % of the HBM copy peak), bulk-copy ring 60.2 us (67 %).
    // The SIMT path is therefore the default; EXB_GATHER_BULK=1 selects the bulk engine.
    return knob("EXB_GATHER_BULK", 0);
  }();
  GatherPlan g;
  BulkCols bk;
  g.n = 0;
  bk.n = 0;
  bk.chunk_base[0] = 0;
  // long rows first, so the short columns and the reward pseudo-column fill the tail
  for (int pass = 0; pass < 2; ++pass) {
    for (int i = 0; i < n_cols; ++i) {
      const exb_gather_col& c = host_cols[i];
      EXB_CHECK_ARG(c.src && c.dst && c.row_bytes > 0, "column %d: bad pointer or row_bytes", i);
      const uintptr_t a = (uintptr_t)c.src | (uintptr_t)c.dst | (uintptr_t)c.row_bytes;
      const int vec = (a % 16 == 0) ? 16 : (a % 8 == 0) ? 8 : (a % 4 == 0) ? 4 : 1;
      const int64_t nv = c.row_bytes / vec;
      const bool is_long = nv > 128;
      if (is_long != (pass == 0)) continue;
      if (bulk_enabled && rows && a % 16 == 0 && c.row_bytes >= 4096 && bk.n < BULK_MAX_COLS) {
        bk.c[bk.n] = c;
        bk.chunks_per_row[bk.n] = (int)((c.row_bytes + BULK_CHUNK - 1) / BULK_CHUNK);
        bk.chunk_base[bk.n + 1] = bk.chunk_base[bk.n] + bk.chunks_per_row[bk.n];
        ++bk.n;
        continue;
      }
      const int j = g.n++;
      g.c[j] = c;
      g.vec[j] = vec;
      g.pieces[j] = is_long ? (int)((nv + GATHER_PIECE_VECS - 1) / GATHER_PIECE_VECS) : 0;
    }
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
  int64_t blocks = 0;
  for (int i = 0; i < g.n; ++i) {
    g.block_base[i] = (unsigned)blocks;
    blocks += g.pieces[i] ? batch * g.pieces[i]
                          : (batch + GATHER_THREADS / 32 - 1) / (GATHER_THREADS / 32);
  }
  g.block_base[g.n] = (unsigned)blocks;
  blocks += (batch + GATHER_THREADS - 1) / GATHER_THREADS;
  g.block_base[g.n + 1] = (unsigned)blocks;
  EXB_CHECK_ARG(blocks < (int64_t)0x7fffffff, "batch too large for one launch");
  replay_gather_kernel<<<(unsigned)blocks, GATHER_THREADS, 0, (cudaStream_t)stream>>>(
      rows, batch, g, draw, reward_src, discount_src, nstep, nstep_mode, gamma, out_reward,
      out_discount);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

extern "C" int exb_replay_gather(const int64_t* rows, int64_t batch,
                                 const exb_gather_col* host_cols, int n_cols,
                                 const float* reward_src, const float* discount_src, int nstep,
                                 int nstep_mode, float gamma, float* out_reward,
                                 float* out_discount, exb_stream_t stream) {
  EXB_CHECK_ARG(rows, "null rows");
  GatherDraw none = {};
  return launch_gather(rows, none, batch, host_cols, n_cols, reward_src, discount_src, nstep,
                       nstep_mode, gamma, out_reward, out_discount, stream);
}

extern "C" int exb_replay_sample(uint64_t seed, uint64_t call, const int64_t* eps_start,
                                 const int32_t* eps_len, int64_t n_episodes, int64_t batch,
                                 const exb_gather_col* host_cols, int n_cols,
                                 const float* reward_src, const float* discount_src, int nstep,
                                 int nstep_mode, float gamma, float* out_reward,
                                 float* out_discount, int64_t* out_rows, exb_stream_t stream) {
  EXB_CHECK_ARG(eps_start && eps_len && n_episodes > 0, "bad episode tables");
  GatherDraw d;
  d.seed = seed;
  d.call = call;
  d.eps_start = eps_start;
  d.eps_len = eps_len;
  d.n_episodes = n_episodes;
  d.nstep = nstep_mode ? nstep : 1;
  d.out_rows = out_rows;
  return launch_gather(nullptr, d, batch, host_cols, n_cols, reward_src, discount_src, nstep,
                       nstep_mode, gamma, out_reward, out_discount, stream);
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
