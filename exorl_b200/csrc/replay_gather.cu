// n-step replay gather over a device-resident flat buffer + counter-based index draw.
//
// Replaces the per-sample python of /root/reference/utils/replay_buffer.py:223-235
// (ReplayBuffer._sample), :87-96 (OfflineReplayBuffer._sample) and the collate / pin / H2D of
// :246-277.  HBM-bound byte work: every (sample, column) row is copied with the widest aligned
// vector the layout allows (16 B for walker obs = 96 B and pixel frames = 63 504 B), one thread
// group per row, groups sized so that small rows still fill warps.
#include "common.cuh"

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
  GatherCols g;
  g.n = n_cols;
  int min_group = GATHER_THREADS;
  for (int i = 0; i < n_cols; ++i) {
    const exb_gather_col& c = host_cols[i];
    EXB_CHECK_ARG(c.src && c.dst && c.row_bytes > 0, "column %d: bad pointer or row_bytes", i);
    g.c[i] = c;
    const uintptr_t a = (uintptr_t)c.src | (uintptr_t)c.dst | (uintptr_t)c.row_bytes;
    g.vec[i] = (a % 16 == 0) ? 16 : (a % 8 == 0) ? 8 : (a % 4 == 0) ? 4 : 1;
    g.group[i] = (c.row_bytes / g.vec[i] > 128) ? GATHER_THREADS : 32;
    if (g.group[i] < min_group) min_group = g.group[i];
  }
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
