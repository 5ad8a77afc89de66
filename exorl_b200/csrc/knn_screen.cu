// Screened kNN for sm_100a: TMA-fed tcgen05 GEMM with the accumulator filtered straight out of
// TMEM against a per-row running threshold, then an exact fp32 recompute of the survivors.
//
// Replaces /root/reference/utils/utils.py:293-300 (and proto.py:115-116) with the same results
// as the exact kernel (knn_exact.cu) at tensor-core speed.  Nothing of size nq x nb reaches HBM.
//
// Math.  With a^ = fp16(a), b^ = fp16(b):  ||a^ - b^||^2 = 2 (ha - s),  s = a^.b^ - hb,
// ha = ||a^||^2/2, hb = ||b^||^2/2.  The GEMM gives a^.b^ (fp32 accumulate), so "nearest" =
// "largest s".  The true distance obeys | ||a-b|| - ||a^-b^|| | <= ea + eb with
// ea = ||a - a^||, eb = ||b - b^|| (triangle inequality), both computed exactly per row by the
// prep kernel.  While scanning, each query row keeps a descending list of scores; with tk its
// k-th largest score, every bank row that can still belong to the true top-k satisfies
//     s >= tkeep = ha - (sqrt(2(ha - tk) + 2 dl) + 2(ea + EB))^2 / 2 - dl          
// (EB = max eb, dl bounds the fp32 accumulation error).  The list keeps everything >= tkeep; if
// it would need more than SCREEN_CAP slots the row is marked undecided and recomputed by the
// exact kernel.  The survivors' distances are then recomputed in fp32 direct-difference form,
// so the returned distances/indices are those of the exact path.
//
// Kernel structure (one CTA per 128 query rows x a range of 256-row bank tiles; CTA pairs with
// cta_group::2 share each bank tile):
//   warp 0      TMA producer: query tile once (resident in smem), bank tiles through a ring
//   warp 1      TMEM allocator + single-thread tcgen05.mma issuer (M=128|256, N=256, K=16)
//   warps 2-5   epilogue (FILT 1, 2): tcgen05.ld 32 columns at a time (thread = TMEM lane =
//   (2-9)       query row), subtract hb, compare with tkeep; hits are rare and handled by a
//               per-thread loop (k-th-best chain in registers, candidate list in shared memory).
//               FILT 3 runs eight epilogue warps — two trackers per row, one per 128-column half
//               of the tile, lists in global memory — for scans short enough that hit handling,
//               not the tensor pipe, would otherwise set the pace.
//   two 256-column TMEM accumulators so the MMA of tile t+1 overlaps the filter of tile t.
// Scan order: every CTA walks its tile range cyclically, starting where the CTAs already running
// are (a shared stream head in global memory), so the co-resident CTAs pull the same window of
// the bank through L2: DRAM reads fell from 146 GB to 37 GB per 1M x 500k launch (L2 hit 90 ->
// 98 %), which under the 1 kW power cap is clock, not just bandwidth.  The first few tiles are
// scored twice (FILT 2, 3): once only to seed the k-th-best chain from their group maxima.
#include "common.cuh"
#include "prep_row.cuh"
#include "tc_ptx.cuh"
#include "tmap_host.cuh"
#include <cuda.h>
#include <algorithm>
#include <atomic>
#include <cuda_fp16.h>
#include <float.h>
#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>

namespace exb {

int exact_splits(int64_t nq, int64_t nb, int k);
int launch_pbe_finish(const float* dist, int64_t nq, int k, int avg, int use_rms, int use_clip,
                      float clip, const float* m_ptr, float* out, cudaStream_t st);
int launch_exact(const float* src, int64_t nq, const float* bank, int64_t nb, int dim, int k,
                 int idx_base, float* out_dist, int* out_idx, float* part_d2, int* part_idx,
                 int splits, const int* row_list, const int* row_count, int64_t max_slots,
                 cudaStream_t st);

constexpr int SCR_M = 128;          // query rows per CTA (UMMA M)
constexpr int SCR_N = 256;          // bank rows per tile (UMMA N)
constexpr int SCR_KC = 64;          // fp16 elements per K chunk = one 128-byte swizzle row
constexpr int SCR_CAP = 24;         // list slots per (row, split)
constexpr int SCR_MAX_K = 16;       // k supported by the screened path
// Cycle-counter instrumentation of the streaming kernel (exb_debug_trace_sync) costs registers in
// the hot loop, so it is compiled in only for diagnostic builds:
//   EXB_EXTRA_NVCC_FLAGS=-DEXB_SCREEN_STATS=1 python -m exorl_b200._build --force
#ifndef EXB_SCREEN_STATS
#define EXB_SCREEN_STATS 0
#endif
constexpr int SCR_THREADS = 192;
constexpr int SCR_THREADS_V3 = 320;
constexpr int SCR_SCAN_CAP = 64;      // FILT 3: in-scan scratch list slots per tracker (global memory)   // FILT 3: TMA warp, MMA warp, eight epilogue warps
constexpr int SCR_A_CHUNK_BYTES = SCR_M * SCR_KC * 2;   // 16 KB
constexpr int SCR_B_STAGE_BYTES = SCR_N * SCR_KC * 2;   // 32 KB
constexpr int SCR_MAX_STAGES = 8;
constexpr int SCR_SMEM_LIMIT = 232448;                  // 227 KB
constexpr int FALLBACK_CAP = 4096;
constexpr int FALLBACK_SPLITS = 42;   // x k <= 512 entries, the merge kernel's capacity
// fp32 accumulation error of a <= 512-term fp16 dot product, relative to ||a^|| ||b^|| <= ha + hb:
// 512 additions x 2^-23 per (truncating) add = 2^-14, plus 2^-18 for the fp32
// rounding of the two half norms themselves (21 roundings x 2^-24 each, for ha and for hb).
constexpr float SCR_ACC_EPS = 1.0f / 16384.0f + 1.0f / 262144.0f;

// ------------------------------------------------------------------------------------------
// prep: fp32 rows -> padded fp16 operand, half squared norm of the rounded row, rounding error
// ------------------------------------------------------------------------------------------
// PrepScalars {err_max, hn_max, range_flag, pad}: common.cuh (shared with rep_producer.cu)

// Affine record of a range-proof call (device, AFF_FLOATS floats): [0, dim) the column shift mu,
// [AFF_SCALE] the power-of-two scale s, [AFF_ON] != 0 when the transform applies, [AFF_SPREAD]
// max |x - mu| it was derived from.  The screen sees s (x - mu): distances are translation
// invariant and scale by s, the certificate of DESIGN.md §4 is scale free, and the rerank
// recomputes the survivors from the caller's fp32 rows, so ANY (mu, s) gives the exact answer —
// the choice only decides how much of the fp16 format the rows use.
constexpr int AFF_SCALE = EXB_SCREEN_MAX_DIM;
constexpr int AFF_ON = EXB_SCREEN_MAX_DIM + 1;
constexpr int AFF_SPREAD = EXB_SCREEN_MAX_DIM + 2;
constexpr int AFF_FLOATS = EXB_AFFINE_FLOATS;
constexpr float AFF_EXTRA_REL = 1.0f / 4194304.0f;   // 2^-22, see OpOut::extra_rel

// One warp per row.  VEC: dim % 4 == 0 and 16-byte aligned rows -> every lane issues all of its
// float4 loads up front (<= 4 for dim <= 512) instead of a dependent scalar loop.  A row with a
// value outside the fp16 range gets round_err = +inf (the screen then keeps nothing certain for
// it and every row falls to the exact kernel) in addition to the range flag.
template <bool VEC>
__device__ __forceinline__ void prep_one_row(const float* __restrict__ x, int64_t row, int dim,
                                             int dpad, __half* __restrict__ out,
                                             float* __restrict__ half_norm,
                                             float* __restrict__ round_err,
                                             PrepScalars* __restrict__ sc,
                                             const float* __restrict__ aff, int lane) {
  const float* xr = x + row * dim;
  // affine-prepared operand (range-proof screen, see colstats below): rows become
  // s * fl(x - mu) with the call's column shift mu and power-of-two scale s
  const bool affine = aff != nullptr && aff[AFF_ON] != 0.f;
  const float scale = affine ? aff[AFF_SCALE] : 1.f;
  if constexpr (VEC) {
    // shared with the rep producers (prep_row.cuh): bit-identical operands
    const float4* x4 = reinterpret_cast<const float4*>(xr);
    const int nv = dim >> 2;
    float4 v[1][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = lane + 32 * i;
      v[0][i] = (c < nv) ? __ldg(x4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    OpOut op{out, half_norm, round_err, sc, dpad};
    if (affine) {
      const float4* m4 = reinterpret_cast<const float4*>(aff);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int c = lane + 32 * i;
        if (c < nv) {
          const float4 m = m4[c];
          v[0][i].x = __fmul_rn(__fsub_rn(v[0][i].x, m.x), scale);
          v[0][i].y = __fmul_rn(__fsub_rn(v[0][i].y, m.y), scale);
          v[0][i].z = __fmul_rn(__fsub_rn(v[0][i].z, m.z), scale);
          v[0][i].w = __fmul_rn(__fsub_rn(v[0][i].w, m.w), scale);
        }
      }
      op.extra_rel = AFF_EXTRA_REL;
    }
    emit_rows<1, 4, false>(v, lane, dim, row, 1, nullptr, op);
  } else {
    // scalar path (dim % 4 != 0): fp32 throughout, norm sum rounded to nearest, error sum rounded
    // UP at every step so that round_err stays a rigorous upper bound (f - hf is exact)
    __half* orow = out + row * dpad;
    float nn = 0.f, ee = 0.f;
    bool bad = false;
    for (int c = lane; c < dpad; c += 32) {
      float v = (c < dim) ? xr[c] : 0.f;
      if (affine && c < dim) v = __fmul_rn(__fsub_rn(v, aff[c]), scale);
      bad |= !(fabsf(v) <= 65504.f);
      const __half h = __float2half_rn(v);
      const float hf = __half2float(h);
      orow[c] = h;
      nn = fmaf(hf, hf, nn);
      const float e = v - hf;
      ee = __fmaf_ru(e, e, ee);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      nn += __shfl_xor_sync(0xffffffffu, nn, off);
      ee = __fadd_ru(ee, __shfl_xor_sync(0xffffffffu, ee, off));
    }
    bad = __any_sync(0xffffffffu, bad);
    if (lane == 0) {
      const float hn = bad ? 0.f : 0.5f * nn;
      float er = bad ? INFINITY : __fsqrt_ru(ee);   // rounded up: the bound stays rigorous
      if (affine && !bad) er = affine_round_err(er, nn, AFF_EXTRA_REL);
      half_norm[row] = hn;
      round_err[row] = er;
      if (sc) {
        int* pe = reinterpret_cast<int*>(&sc->err_max);
        int* ph = reinterpret_cast<int*>(&sc->hn_max);
        if (__float_as_int(er) > *reinterpret_cast<volatile int*>(pe))
          atomicMax(pe, __float_as_int(er));
        if (__float_as_int(hn) > *reinterpret_cast<volatile int*>(ph))
          atomicMax(ph, __float_as_int(hn));
        if (bad) atomicExch(&sc->range_flag, 1);
      }
    }
  }
}

template <bool VEC>
__global__ void __launch_bounds__(256)
knn_prep_kernel(const float* __restrict__ x, int64_t n, int dim, int dpad,
                __half* __restrict__ out, float* __restrict__ half_norm,
                float* __restrict__ round_err, PrepScalars* __restrict__ sc,
                unsigned long long* __restrict__ zero16, const float* __restrict__ aff) {
  pdl_launch_dependents();   // per-update chain: the screen's blocks may set up while this runs
  if (zero16 != nullptr && blockIdx.x == 0 && threadIdx.x < 3) zero16[threadIdx.x] = 0ull;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= n) return;
  prep_one_row<VEC>(x, row, dim, dpad, out, half_norm, round_err, sc, aff, threadIdx.x & 31);
}

// ------------------------------------------------------------------------------------------
// Column statistics -> affine record (range-proof screen).
//
// One streaming pass over the rows: per-column sum, min and max.  colstats_kernel<E> (E = float4
// for 16-byte rows, float otherwise) lets thread t own element-column t % ncol of every
// (blockDim / ncol)-th row of its block's stripe; the row groups of a block are combined in a
// fixed order and the block writes one partial [3][dim].  colstats_reduce_kernel adds the
// partials in block order (fp64 sum), so the statistics — and with them the operand — do not
// depend on scheduling.  affine_kernel turns {sum, min, max} into the record: mu = sum / n,
// A = max_c max(max_c - mu_c, mu_c - min_c) rounded up, s = 2^(6 - floor(log2 A)) so that
// s * A lies in [64, 128) (fp16 normal range 2^-14 .. 65504: 9 binades of headroom above, 20
// below).  Non-finite statistics (a NaN / inf anywhere) leave the transform off; the range flag
// of the prep then routes the call as before.
// ------------------------------------------------------------------------------------------
constexpr int CS_MAX_BLOCKS = 592;      // 4 x 148
constexpr int CS_ROWS_PER_BLOCK = 256;

// fminf / fmaxf drop a NaN operand; a NaN stays visible through the sum (NaN + x = NaN)
__device__ __forceinline__ void cs_merge(float& s, float& lo, float& hi, float s2, float lo2,
                                         float hi2) {
  s += s2;
  lo = fminf(lo, lo2);
  hi = fmaxf(hi, hi2);
}
__device__ __forceinline__ void cs_merge(float4& s, float4& lo, float4& hi, const float4 s2,
                                         const float4 lo2, const float4 hi2) {
  cs_merge(s.x, lo.x, hi.x, s2.x, lo2.x, hi2.x);
  cs_merge(s.y, lo.y, hi.y, s2.y, lo2.y, hi2.y);
  cs_merge(s.z, lo.z, hi.z, s2.z, lo2.z, hi2.z);
  cs_merge(s.w, lo.w, hi.w, s2.w, lo2.w, hi2.w);
}
__device__ __forceinline__ void cs_init(float4& s, float4& lo, float4& hi) {
  s = make_float4(0.f, 0.f, 0.f, 0.f);
  lo = make_float4(INFINITY, INFINITY, INFINITY, INFINITY);
  hi = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
}
__device__ __forceinline__ void cs_init(float& s, float& lo, float& hi) {
  s = 0.f; lo = INFINITY; hi = -INFINITY;
}

// part: [gridDim.x][3][ncol] elements of E (sum, min, max); dynamic smem = 3 * ncol * sizeof(E)
template <typename E>
__global__ void __launch_bounds__(512)
colstats_kernel(const E* __restrict__ x, int64_t n, int ncol, int64_t rows_per_block,
                E* __restrict__ part) {
  extern __shared__ __align__(16) unsigned char cs_smem[];
  E* s_sum = reinterpret_cast<E*>(cs_smem);
  E* s_lo = s_sum + ncol;
  E* s_hi = s_lo + ncol;
  const int groups = blockDim.x / ncol;          // row groups of this block (>= 1)
  const int g = threadIdx.x / ncol, c = threadIdx.x - g * ncol;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = min(n, r0 + rows_per_block);
  E sm, lo, hi;
  cs_init(sm, lo, hi);
  if (g < groups) {
    const int64_t st = groups;
    int64_t r = r0 + g;
    for (; r + 3 * st < r1; r += 4 * st) {       // four independent loads in flight per thread
      const E v0 = __ldg(x + r * ncol + c);
      const E v1 = __ldg(x + (r + st) * ncol + c);
      const E v2 = __ldg(x + (r + 2 * st) * ncol + c);
      const E v3 = __ldg(x + (r + 3 * st) * ncol + c);
      cs_merge(sm, lo, hi, v0, v0, v0);
      cs_merge(sm, lo, hi, v1, v1, v1);
      cs_merge(sm, lo, hi, v2, v2, v2);
      cs_merge(sm, lo, hi, v3, v3, v3);
    }
    for (; r < r1; r += st) {
      const E v = __ldg(x + r * ncol + c);
      cs_merge(sm, lo, hi, v, v, v);
    }
  }
  for (int gg = 0; gg < groups; ++gg) {          // fixed combination order
    if (g == gg) {
      if (gg == 0) {
        s_sum[c] = sm; s_lo[c] = lo; s_hi[c] = hi;
      } else {
        E a = s_sum[c], b = s_lo[c], d = s_hi[c];
        cs_merge(a, b, d, sm, lo, hi);
        s_sum[c] = a; s_lo[c] = b; s_hi[c] = d;
      }
    }
    __syncthreads();
  }
  if (threadIdx.x < ncol) {
    E* o = part + (int64_t)blockIdx.x * 3 * ncol;
    o[c] = s_sum[c];
    o[ncol + c] = s_lo[c];
    o[2 * ncol + c] = s_hi[c];
  }
}

// stats: [3][dim] floats (sum, min, max) over the partials of up to two row sets; one thread per
// column.  A partial block is [3][dim] floats whatever E was (float4 columns are 4 float columns).
__global__ void colstats_reduce_kernel(const float* __restrict__ part_a, int blocks_a,
                                       const float* __restrict__ part_b, int blocks_b, int dim,
                                       float* __restrict__ stats) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= dim) return;
  double sm = 0.0;
  float lo = INFINITY, hi = -INFINITY;
  for (int side = 0; side < 2; ++side) {
    const float* p = side ? part_b : part_a;
    const int nblk = side ? blocks_b : blocks_a;
    for (int b = 0; b < nblk; ++b) {
      const float* q = p + (int64_t)b * 3 * dim;
      sm += (double)q[c];
      lo = fminf(lo, q[dim + c]);
      hi = fmaxf(hi, q[2 * dim + c]);
    }
  }
  stats[c] = (float)sm;
  stats[dim + c] = lo;
  stats[2 * dim + c] = hi;
}

// One block of EXB_SCREEN_MAX_DIM threads.
__global__ void __launch_bounds__(EXB_SCREEN_MAX_DIM)
affine_kernel(const float* __restrict__ stats, double n_rows, int dim, float* __restrict__ aff) {
  __shared__ float s_max[EXB_SCREEN_MAX_DIM / 32];
  __shared__ int s_bad[EXB_SCREEN_MAX_DIM / 32];
  const int c = threadIdx.x;
  float mu = 0.f, spread = 0.f;
  int bad = 0;
  if (c < dim) {
    const float sm = stats[c], lo = stats[dim + c], hi = stats[2 * dim + c];
    mu = (float)((double)sm / n_rows);
    // mu may fall outside [lo, hi] by rounding: take the larger one-sided distance, rounded up
    spread = fmaxf(__fsub_ru(hi, mu), __fsub_ru(mu, lo));
    bad = !(fabsf(sm) < INFINITY) || !(fabsf(lo) < INFINITY) || !(fabsf(hi) < INFINITY) ||
          !(spread < INFINITY);
  }
  float wm = spread;
  int wb = bad;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    wm = fmaxf(wm, __shfl_xor_sync(0xffffffffu, wm, off));
    wb |= __shfl_xor_sync(0xffffffffu, wb, off);
  }
  if ((c & 31) == 0) { s_max[c >> 5] = wm; s_bad[c >> 5] = wb; }
  __syncthreads();
  float a = 0.f;
  int any_bad = 0;
  for (int w = 0; w < EXB_SCREEN_MAX_DIM / 32; ++w) { a = fmaxf(a, s_max[w]); any_bad |= s_bad[w]; }
  const bool on = !any_bad && a > 0.f;
  float scale = 1.f;
  if (on) {
    int e = 6 - ilogbf(a);                       // s * a in [64, 128)
    e = e < -100 ? -100 : (e > 100 ? 100 : e);
    scale = ldexpf(1.f, e);
  }
  aff[c] = (on && c < dim) ? mu : 0.f;
  if (c == 0) {
    aff[AFF_SCALE] = scale;
    aff[AFF_ON] = on ? 1.f : 0.f;
    aff[AFF_SPREAD] = a;
    for (int i = AFF_SPREAD + 1; i < AFF_FLOATS; ++i) aff[i] = 0.f;
  }
}

// ------------------------------------------------------------------------------------------
// screening kernel
// ------------------------------------------------------------------------------------------
struct SymPeerTable;   // multi-GPU symmetric screen, below
struct ScreenParams {
  const float* ha;        // [nq] query half norms
  const float* ea;        // [nq] query rounding errors
  const float* hb;        // [nb] bank half norms
  const PrepScalars* bsc; // bank scalars (EB, HBmax)
  int64_t nq, nb;
  int nkc;                // K chunks of 64
  int n_tiles;            // bank tiles of 256
  int tiles_per_split;
  int splits;
  int k;
  int stages;
  int prefetch;           // L2 prefetch distance in bank tiles (0 = off)
  int debug;              // measurement only (EXB_SCREEN_DEBUG): 1 = skip filter, 2 = load only
  float* cand_val;        // [nq, splits, CAP] screened scores s = a^.b^ - hb
  int* cand_idx;          // [nq, splits, CAP]
  int* cand_cnt;          // [nq, splits]; -1 = undecided
  long long* stats;       // debug (EXB_SCREEN_TRACE): 8 accumulated cycle counters, else null
  int lists;              // candidate lists per query row = splits x trackers per (row, split)
  int seed_tiles;         // FILT 3: tiles scored twice — first only to seed the k-th-best chain
  int scan_cap;           // FILT 3: slots of the in-scan scratch list per tracker
  float* scan_val;        // FILT 3: [nq, lists, scan_cap] scratch lists (global)
  int* scan_idx;
  int* head;              // [splits] shared stream head (virtual tile position), zeroed per call;
                          // null = every CTA scans its tile range from the first tile
  // Hybrid tail (1-D grid): the first full_units CTA groups scan the whole bank for one query
  // tile each; the query tiles of the partly filled last wave are cut into tail_splits bank
  // ranges of tail_tps tiles so that the idle SMs of that wave share their scans.  0 = off.
  int tail_splits, full_units, tail_tps;
  int chain_reset;        // 1 = drop the seed chain after the seed phase (default); 0 = A/B only
  // FILT 4 (symmetric screen, source == bank): every query tile pair scans the bank tiles from its
  // own upwards; an accumulator element g_ij is tested in both directions against the rows' fixed
  // thresholds and logged per (row, column half): direct hits (bank row j for query i) and
  // transposed hits (bank row i for query j).  log_cap slots per log; count -1 = overflowed.
  // Query tile pairs [sym_pair_lo, sym_pair_lo + gridDim.x / CG) of this launch; pair u scans the
  // bank tiles u, u + 1, ... (cyclic, modulo n_tiles), (sym_span_end - u) of them, plus
  // sym_extra more for the pairs u >= sym_mid_pair.  One GPU: pair_lo 0, span_end n_tiles, no
  // extra.  Several GPUs (relabel.py): rank r owns a block of pairs and the block pairs
  // (r, r), (r, r + 1), ..., cyclically half way round, the antipodal block split in halves.
  int sym_pair_lo, sym_span_end, sym_mid_pair, sym_extra;
  const float* sym_tkeep; // [n] stateless keep thresholds
  float* dlog_val;        // [n, 2, log_cap] direct hits: score g - hb_j
  int* dlog_idx;          //                  bank row j
  int* dlog_cnt;          // [n, 2]
  int log_cap;
  // transposed hits go to ONE append buffer (hub rows are candidates of thousands of rows: a
  // per-row log cannot be sized): entries (query row j, score g - hb_i, bank row i), staged per
  // warp in shared memory and appended 64+ at a time with one atomicAdd on tq_count
  int* tq_target;
  float* tq_val;
  int* tq_src;
  unsigned long long* tq_count;   // entries appended (may exceed tq_cap: the excess is dropped)
  long long tq_cap;
  // Several GPUs with peer-mapped memory (null = the one buffer above): the hits found for the
  // rows of bank tile t are appended to the buffer of the rank that OWNS those rows, straight
  // over NVLink — the exchange step of the multi-GPU relabel happens inside this kernel.
  const SymPeerTable* sym_peers;
};

// Device-resident routing table of the multi-GPU symmetric screen: rank r owns the bank tiles
// [tile_lo[r], tile_lo[r + 1]); target/val/src/count/cap describe ITS append buffer as mapped in
// this process (CUDA IPC, capi_common.cu: exb_peer_buffer_*; the own rank's entry is local).
struct SymPeerTable {
  int ranks, pad;
  int tile_lo[EXB_SYM_MAX_RANKS + 2];
  int* target[EXB_SYM_MAX_RANKS];
  float* val[EXB_SYM_MAX_RANKS];
  int* src[EXB_SYM_MAX_RANKS];
  unsigned long long* count[EXB_SYM_MAX_RANKS];
  long long cap[EXB_SYM_MAX_RANKS];
};
__global__ void sym_peer_table_kernel(const SymPeerTable t, SymPeerTable* out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *out = t;
}

// Per-row candidate state of one epilogue thread.  The k best scores seen so far are kept sorted
// (descending) in registers by a branch-free min/max chain, so the running k-th best `tk` costs no
// memory traffic; the candidates themselves (score, bank index) are appended to an unsorted
// shared-memory buffer and pruned lazily (entries that fell below the rising `tkeep`).
struct ListState {
  int n;                     // entries in the buffer
  float dropped;             // largest score ever dropped for lack of room (-FLT_MAX if none);
                             // the row is undecided iff dropped >= the FINAL tkeep
  float tk, tkeep;
  float ha, e2, dl;
  // k best scores, descending, RIGHT-aligned in a fixed 16-slot chain: slots [0, 16-k) hold
  // +FLT_MAX sentinels (values pass through them unchanged), so the k-th best is always the
  // static last slot and nothing is ever indexed dynamically (keeps the chain in registers).
  float tv[SCR_MAX_K];
};

__device__ __forceinline__ float keep_threshold(const ListState& s) {
  // DESIGN.md §4; evaluated conservatively (round the kept region outwards).  sqrt(x) as
  // x * rsqrt(x) (2 ulp) inflated by 1e-6: a hit pays for this on its critical path.
  const float x = fmaxf(2.f * (s.ha - s.tk), 0.f) + 2.f * s.dl + 1e-30f;
  const float dk = x * rsqrtf(x) * 1.000001f;
  const float bound = dk + s.e2;
  float t = s.ha - 0.5f * bound * bound - s.dl;
  t -= 4e-6f * (fabsf(s.ha) + fabsf(s.tk));
  return t;
}

// bv/bi: this thread's column of the [CAP][STRIDE] arrays
template <int STRIDE = SCR_M>
__device__ __forceinline__ void list_compact(ListState& s, float* bv, int* bi) {
  int w = 0;
  for (int r = 0; r < s.n; ++r) {
    const float v = bv[r * STRIDE];
    const int j = bi[r * STRIDE];
    if (v >= s.tkeep) {
      bv[w * STRIDE] = v;
      bi[w * STRIDE] = j;
      ++w;
    }
  }
  s.n = w;
}

// Insert one score into the sorted (descending) register chain, branch-free and with every slot
// independent of the others: tv'[i] = max(tv[i], min(tv[i-1], x)) — x lands where it belongs,
// smaller entries shift down one slot, the old last entry falls off.  (Depth 2 instead of a
// 2*16-deep dependent min/max chain: an epilogue warp runs alone on its scheduler, so dependent
// latency, not issue slots, is what a hit costs.)
__device__ __forceinline__ void chain_push(ListState& s, float x) {
#pragma unroll
  for (int i = SCR_MAX_K - 1; i > 0; --i) s.tv[i] = fmaxf(s.tv[i], fminf(s.tv[i - 1], x));
  s.tv[0] = fmaxf(s.tv[0], x);
}

// After the seed phase the chain holds GROUP MAXIMA (one score per 4 columns) of the sample
// tiles.  That is a valid lower bound of the k-th best, but when a row's true neighbours sit in
// consecutive columns of those tiles (trajectory data: consecutive states are near-duplicates) it
// is far too loose, and because the sample tiles' scores could not be pushed a second time it
// stayed loose: whole query tiles overflowed their lists (18 % of the rows of a 200-episode
// random-walk set, 93 % with 10x smaller steps; profiles/r01i_trajectory_*.jsonl).  So the seed
// chain only fixes the initial tkeep; the main phase starts an EMPTY chain and pushes every
// inserted score exactly once (each column is visited once), which makes tk the exact k-th best.
__device__ __forceinline__ void chain_reset(ListState& s, int k) {
#pragma unroll
  for (int i = 0; i < SCR_MAX_K; ++i) s.tv[i] = (i < SCR_MAX_K - k) ? FLT_MAX : -FLT_MAX;
  s.tk = -FLT_MAX;
}

__device__ __forceinline__ void refresh_threshold(ListState& s) {
  s.tk = s.tv[SCR_MAX_K - 1];
  if (s.tk > -FLT_MAX) s.tkeep = fmaxf(s.tkeep, keep_threshold(s));   // k scores seen
}

// Scores of one 32-column chunk (thread = query row) and their maxima over groups of 4 columns.
__device__ __forceinline__ float chunk_scores(const uint32_t (&r)[32], const float* hn_chunk,
                                              float (&sc)[32], float (&g)[8]) {
  const float4* hn4 = reinterpret_cast<const float4*>(hn_chunk);
  float m = -INFINITY;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const float4 h = hn4[q];
    sc[4 * q] = __uint_as_float(r[4 * q]) - h.x;
    sc[4 * q + 1] = __uint_as_float(r[4 * q + 1]) - h.y;
    sc[4 * q + 2] = __uint_as_float(r[4 * q + 2]) - h.z;
    sc[4 * q + 3] = __uint_as_float(r[4 * q + 3]) - h.w;
    g[q] = fmaxf(fmaxf(sc[4 * q], sc[4 * q + 1]), fmaxf(sc[4 * q + 2], sc[4 * q + 3]));
    m = fmaxf(m, g[q]);
  }
  return m;
}

// Seed pass (first tile of a CTA): only the running k-th best is updated, from group maxima.
// Every group maximum is a real score, so the chain's k-th entry never exceeds the true k-th
// best score: thresholds derived from it are conservative.
__device__ __forceinline__ void seed_chunk(ListState& st, const uint32_t (&r)[32],
                                           const float* hn_chunk) {
  float sc[32], g[8];
  chunk_scores(r, hn_chunk, sc, g);
#pragma unroll
  for (int q = 0; q < 8; ++q)
    if (g[q] > st.tv[SCR_MAX_K - 1]) chain_push(st, g[q]);
}

// Filter one 32-column chunk: append every column whose score reaches tkeep (tiny predicated
// stores), optionally feed the chain with the group maxima, refresh the thresholds once.
__device__ __forceinline__ void filter_chunk(ListState& st, const uint32_t (&r)[32], int col0,
                                             const float* hn_chunk, float* bv, int* bi,
                                             bool feed_chain) {
  float sc[32], g[8];
  const float m = chunk_scores(r, hn_chunk, sc, g);
  if (__any_sync(0xffffffffu, m >= st.tkeep)) {
    // prune stale entries only when the buffer is nearly full (steady state holds ~k+2 live ones)
    if (st.n >= SCR_CAP - 4) list_compact(st, bv, bi);
    const float tkeep = st.tkeep;
#pragma unroll
    for (int c = 0; c < 32; ++c) {
      if (sc[c] >= tkeep) {
        if (st.n < SCR_CAP) {
          bv[st.n * SCR_M] = sc[c];
          bi[st.n * SCR_M] = col0 + c;
          ++st.n;
        } else {
          st.dropped = fmaxf(st.dropped, sc[c]);   // judged against the final tkeep
        }
      }
    }
    if (feed_chain && m > st.tv[SCR_MAX_K - 1]) {
#pragma unroll
      for (int q = 0; q < 8; ++q)
        if (g[q] > st.tv[SCR_MAX_K - 1]) chain_push(st, g[q]);
      refresh_threshold(st);
    }
  }
}

// ---- steady-state path (every tile after a CTA's first): hits are rare, so the hit columns of
// the warp are revisited one by one with uniform control flow and inserted individually; every
// inserted score also goes through the chain, so tk is the exact k-th best of what was kept.
__device__ __forceinline__ void list_insert(ListState& s, float* bv, int* bi, float v, int j,
                                            int k, bool push = true) {
  // precondition: v >= s.tkeep
  if (s.n == SCR_CAP) list_compact(s, bv, bi);
  if (s.n < SCR_CAP) {
    bv[s.n * SCR_M] = v;
    bi[s.n * SCR_M] = j;
    ++s.n;
  } else {
    // still full (only while tkeep is loose, early in a scan): evict the smallest score and
    // remember it; whether that lost a needed candidate is decided against the final tkeep
    int pm = 0;
    float vm = bv[0];
    for (int r = 1; r < SCR_CAP; ++r) {
      const float x = bv[r * SCR_M];
      if (x < vm) { vm = x; pm = r; }
    }
    if (v > vm) {
      bv[pm * SCR_M] = v;
      bi[pm * SCR_M] = j;
      s.dropped = fmaxf(s.dropped, vm);
    } else {
      s.dropped = fmaxf(s.dropped, v);
    }
  }
  if (push && v > s.tk) {    // enters the k best: push through the sorted register chain
    chain_push(s, v);
    refresh_threshold(s);
  }
}

// Select the 4 accumulators of column group q (runtime) without dynamic register indexing.
__device__ __forceinline__ void select_group(const uint32_t (&r)[32], int q, uint32_t (&o)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t a0 = (q & 1) ? r[4 + i] : r[i], a1 = (q & 1) ? r[12 + i] : r[8 + i];
    const uint32_t a2 = (q & 1) ? r[20 + i] : r[16 + i], a3 = (q & 1) ? r[28 + i] : r[24 + i];
    const uint32_t b0 = (q & 2) ? a1 : a0, b1 = (q & 2) ? a3 : a2;
    o[i] = (q & 4) ? b1 : b0;
  }
}

// Per-thread hit loop: every thread walks ITS OWN hit groups (4 columns each, found from the
// group maxima the fast path already has) and inserts from its registers, so a warp pays
// max-over-rows (not sum-over-rows) insertions per chunk and no TMEM re-read.  `push` = false on
// the first tile after a seed pass (its group maxima are already in the chain; pushing them again
// would count a bank row twice).
__device__ __forceinline__ void filter_chunk_bits(ListState& st, const uint32_t (&r)[32], int ch,
                                                  const float* hn, float* lv, int* li, int t,
                                                  int k, bool push, long long* slow = nullptr) {
  const float4* hn4 = reinterpret_cast<const float4*>(hn + ch * 32);
  float g[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const float4 h = hn4[q];
    g[q] = fmaxf(fmaxf(__uint_as_float(r[4 * q]) - h.x, __uint_as_float(r[4 * q + 1]) - h.y),
                 fmaxf(__uint_as_float(r[4 * q + 2]) - h.z, __uint_as_float(r[4 * q + 3]) - h.w));
  }
  const float m = fmaxf(fmaxf(fmaxf(g[0], g[1]), fmaxf(g[2], g[3])),
                        fmaxf(fmaxf(g[4], g[5]), fmaxf(g[6], g[7])));
  if (__any_sync(0xffffffffu, m >= st.tkeep)) {
    const long long c0 = slow ? clock64() : 0;
    uint32_t gm = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) gm |= (g[q] >= st.tkeep ? 1u : 0u) << q;
    while (gm) {
      const int q = __ffs(gm) - 1;
      gm &= gm - 1;
      uint32_t u[4];
      select_group(r, q, u);
      const float4 h = hn4[q];
      const float sv[4] = {__uint_as_float(u[0]) - h.x, __uint_as_float(u[1]) - h.y,
                           __uint_as_float(u[2]) - h.z, __uint_as_float(u[3]) - h.w};
      const int col = t * SCR_N + ch * 32 + 4 * q;
      uint32_t hm = 0;
#pragma unroll
      for (int i = 0; i < 4; ++i) hm |= (sv[i] >= st.tkeep ? 1u : 0u) << i;
      // one (non-unrolled) insertion site: the slow path must stay small in the instruction cache
#pragma unroll 1
      while (hm) {
        const int i = __ffs(hm) - 1;
        hm &= hm - 1;
        const float lo = (i & 1) ? sv[1] : sv[0], hi = (i & 1) ? sv[3] : sv[2];
        const float v = (i & 2) ? hi : lo;
        if (v >= st.tkeep) list_insert(st, lv, li, v, col + i, k, push);
      }
    }
    __syncwarp();
    if (slow) {
      slow[0] += 1;
      slow[1] += clock64() - c0;
    }
  }
}

// ---- FILT 3: candidate lists live in GLOBAL memory (the row's final [CAP] slots of cand_val /
// cand_idx), which frees the shared memory for eight epilogue warps.  Appends are two stores;
// when the CAP slots are full the whole list makes one register round trip: entries below the
// risen tkeep are dropped, and if it is still full the smallest entry is evicted (remembered in
// `dropped`, judged against the final tkeep like everywhere else).
// Compact `n` (score, index) entries of a global list: entries below tkeep are dropped, the rest
// written to dst (in place when dst == src: writes trail the reads), at most cap_dst of them.
// 16 loads in flight per round trip.  Returns the number of live entries found.
__device__ __forceinline__ int list_compact_g(const float* sv, const int* si, int n, float tkeep,
                                              float* dv, int* di, int cap_dst) {
  int w = 0;
  for (int base = 0; base < n; base += 16) {
    float v[16];
    int j[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const bool ok = base + r < n;
      v[r] = ok ? sv[base + r] : -FLT_MAX;
      j[r] = ok ? si[base + r] : 0;
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      if (base + r < n && v[r] >= tkeep) {
        if (w < cap_dst) {
          dv[w] = v[r];
          di[w] = j[r];
        }
        ++w;
      }
    }
  }
  return w;
}

__device__ __forceinline__ void list_insert_g(ListState& s, float* gv, int* gi, int cap, float v,
                                              int j, bool push, long long* slow = nullptr) {
  // precondition: v >= s.tkeep
  const long long c0 = slow ? clock64() : 0;
  if (s.n == cap) {
    if (slow) slow[4] += 1;
    s.n = min(list_compact_g(gv, gi, cap, s.tkeep, gv, gi, cap), cap);
  }
  if (s.n < cap) {
    gv[s.n] = v;
    gi[s.n] = j;
    ++s.n;
  } else {
    s.dropped = fmaxf(s.dropped, v);   // every slot still live: judged against the final tkeep
  }
  const long long c1 = slow ? clock64() : 0;
  if (push && v > s.tk) {
    chain_push(s, v);
    refresh_threshold(s);
  }
  if (slow) {
    slow[2] += 1;
    slow[3] += c1 - c0;
    slow[5] += clock64() - c1;
  }
}

// One 32-column chunk of a tracker (thread = query row, columns of this warp's half tile).
__device__ __forceinline__ void filter_chunk_g(ListState& st, const uint32_t (&r)[32],
                                               const float* hn_chunk, float* gv, int* gi, int cap,
                                               int col0, bool push, long long* slow) {
  const float4* hn4 = reinterpret_cast<const float4*>(hn_chunk);
  float g[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const float4 h = hn4[q];
    g[q] = fmaxf(fmaxf(__uint_as_float(r[4 * q]) - h.x, __uint_as_float(r[4 * q + 1]) - h.y),
                 fmaxf(__uint_as_float(r[4 * q + 2]) - h.z, __uint_as_float(r[4 * q + 3]) - h.w));
  }
  const float m = fmaxf(fmaxf(fmaxf(g[0], g[1]), fmaxf(g[2], g[3])),
                        fmaxf(fmaxf(g[4], g[5]), fmaxf(g[6], g[7])));
  if (__any_sync(0xffffffffu, m >= st.tkeep)) {
    const long long c0 = slow ? clock64() : 0;
    uint32_t gm = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) gm |= (g[q] >= st.tkeep ? 1u : 0u) << q;
    while (gm) {
      const int q = __ffs(gm) - 1;
      gm &= gm - 1;
      uint32_t u[4];
      select_group(r, q, u);
      const float4 h = hn4[q];
      const float sv[4] = {__uint_as_float(u[0]) - h.x, __uint_as_float(u[1]) - h.y,
                           __uint_as_float(u[2]) - h.z, __uint_as_float(u[3]) - h.w};
      uint32_t hm = 0;
#pragma unroll
      for (int i = 0; i < 4; ++i) hm |= (sv[i] >= st.tkeep ? 1u : 0u) << i;
#pragma unroll 1
      while (hm) {
        const int i = __ffs(hm) - 1;
        hm &= hm - 1;
        const float lo = (i & 1) ? sv[1] : sv[0], hi = (i & 1) ? sv[3] : sv[2];
        const float v = (i & 2) ? hi : lo;
        if (v >= st.tkeep) list_insert_g(st, gv, gi, cap, v, col0 + 4 * q + i, push, slow);
      }
    }
    __syncwarp();
    if (slow) {
      slow[0] += 1;
      slow[1] += clock64() - c0;
    }
  }
}

// FILT 4: per-warp staging queue of transposed hits in shared memory (see ScreenParams); the entry
// count lives in a warp-uniform register of the owning warp.
constexpr int SYM_QCAP = 128;
struct SymQueue {
  int* target; float* val; int* src;                    // shared memory, this warp's
  float* scratch;                                       // 32 floats: one lane's chunk, see sym_chunk
  int* g_target; float* g_val; int* g_src;              // the global append buffer
  unsigned long long* g_count;
  long long g_cap;
  bool sys;                                             // the buffer may be another GPU's memory
};
__device__ __forceinline__ void sym_queue_flush(const SymQueue& q, int& qn, int lane) {
  __syncwarp();
  if (qn > 0) {
    unsigned long long base = 0;
    if (lane == 0)
      base = q.sys ? atomicAdd_system(q.g_count, (unsigned long long)qn)
                   : atomicAdd(q.g_count, (unsigned long long)qn);
    base = __shfl_sync(0xffffffffu, base, 0);
    for (int e = lane; e < qn; e += 32) {
      const long long at = (long long)base + e;
      if (at < q.g_cap) {
        q.g_target[at] = q.target[e];
        q.g_val[at] = q.val[e];
        q.g_src[at] = q.src[e];
      }
    }
  }
  __syncwarp();
  qn = 0;
}

// FILT 4: one 32-column chunk tested in both directions against fixed thresholds (thread = row i
// of the query tile = bank row i; column c = bank row j = query row j):
//   direct      s_ij = g - hb_j >= tkeep_i          -> (s_ij, j) into row i's direct log
//   transposed  s_ji = g - hb_i >= tkeep_j, tested as g - tkeep_j >= hb_i (the 4e-6 relative slack
//               of the thresholds covers the different rounding) -> (j, s_ji, i) into the queue:
//               "bank row i is a candidate of query row j"
// Fast path: group maxima of both differences, one vote.  Hit path, WARP-COOPERATIVE: a lane whose
// row has a hit parks its 32 accumulators in shared memory and the whole warp — lane c = column c
// — redoes that row's two tests in parallel, so a hit costs ~60 warp instructions instead of a
// ~200-instruction serial walk by one lane with 31 idle (measured at 1M rows: the serial hit path
// was 97 ms of the kernel's 445).  nd counts every direct hit; entries beyond cap are dropped and
// the log reads as overflowed.
struct SymRow {          // per-lane constants of the lane's query row
  float tkeep_i, hb_t, hb_i;
  int row;
};
__device__ __forceinline__ void sym_chunk(const uint32_t (&r)[32], const float* hn_chunk,
                                          const float* tk_chunk, const SymRow& me, int col0,
                                          int cap, float* dlog_val, int* dlog_idx, int half,
                                          int& nd, const SymQueue& sq, int& qn, int lane,
                                          bool record = true) {
  const unsigned full = 0xffffffffu;
  const float4* hn4 = reinterpret_cast<const float4*>(hn_chunk);
  const float4* tk4 = reinterpret_cast<const float4*>(tk_chunk);
  float gd[8], gt[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const float4 h = hn4[q], t = tk4[q];
    const float a0 = __uint_as_float(r[4 * q]), a1 = __uint_as_float(r[4 * q + 1]);
    const float a2 = __uint_as_float(r[4 * q + 2]), a3 = __uint_as_float(r[4 * q + 3]);
    gd[q] = fmaxf(fmaxf(a0 - h.x, a1 - h.y), fmaxf(a2 - h.z, a3 - h.w));
    gt[q] = fmaxf(fmaxf(a0 - t.x, a1 - t.y), fmaxf(a2 - t.z, a3 - t.w));
  }
  const float md = fmaxf(fmaxf(fmaxf(gd[0], gd[1]), fmaxf(gd[2], gd[3])),
                         fmaxf(fmaxf(gd[4], gd[5]), fmaxf(gd[6], gd[7])));
  const float mt = fmaxf(fmaxf(fmaxf(gt[0], gt[1]), fmaxf(gt[2], gt[3])),
                         fmaxf(fmaxf(gt[4], gt[5]), fmaxf(gt[6], gt[7])));
  unsigned hl = __ballot_sync(full, md >= me.tkeep_i || mt >= me.hb_t);
  if (hl == 0u || !record) return;
  const unsigned lt = (1u << lane) - 1u;
  const float h_c = hn_chunk[lane], t_c = tk_chunk[lane];   // this lane's column
#pragma unroll 1
  while (hl) {                                   // warp-uniform: the lanes (rows) with a hit
    const int L = __ffs(hl) - 1;
    hl &= hl - 1;
    if (qn > SYM_QCAP - 32) sym_queue_flush(sq, qn, lane);
    if (lane == L) {
      uint4* s4 = reinterpret_cast<uint4*>(sq.scratch);
#pragma unroll
      for (int q = 0; q < 8; ++q) s4[q] = make_uint4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]);
    }
    __syncwarp();
    const float g = sq.scratch[lane];
    const float tkeep_L = __shfl_sync(full, me.tkeep_i, L);
    const float hbt_L = __shfl_sync(full, me.hb_t, L);
    const int row_L = __shfl_sync(full, me.row, L);
    const float sd = g - h_c;
    const bool hit_d = sd >= tkeep_L;
    const bool hit_t = (g - t_c) >= hbt_L;
    const unsigned m_d = __ballot_sync(full, hit_d), m_t = __ballot_sync(full, hit_t);
    if (m_d) {
      const int nd_L = __shfl_sync(full, nd, L);
      const int pos = nd_L + __popc(m_d & lt);
      if (hit_d && pos < cap) {
        const int64_t o = ((int64_t)row_L * 2 + half) * cap + pos;
        dlog_val[o] = sd;
        dlog_idx[o] = col0 + lane;
      }
      if (lane == L) nd += __popc(m_d);
    }
    if (m_t) {
      const float hbi_L = __shfl_sync(full, me.hb_i, L);
      const int slot = qn + __popc(m_t & lt);      // < SYM_QCAP by the flush above
      if (hit_t) {
        sq.target[slot] = col0 + lane;
        sq.val[slot] = g - hbi_L;
        sq.src[slot] = row_L;
      }
      qn += __popc(m_t);
    }
    __syncwarp();                                // the scratch row is free again
  }
  if (qn >= 64) sym_queue_flush(sq, qn, lane);
}

// FILT 1: seed + inline appends (short scans); 2: sample-tile seed + per-thread hit loop;
// 3: as 2 with EIGHT epilogue warps — two trackers per query row, one per 128-column half of
// the bank tile, lists in global memory, no barrier between the epilogue warps.
template <int CG, int FILT>
__global__ void __launch_bounds__(FILT >= 3 ? SCR_THREADS_V3 : SCR_THREADS, 1)
knn_screen_kernel(const __grid_constant__ CUtensorMap tmap_q,
                  const __grid_constant__ CUtensorMap tmap_b, const ScreenParams p) {
  extern __shared__ uint8_t smem_raw[];
  // carve (1024-byte aligned for SWIZZLE_128B)
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (base - raw);
  constexpr uint32_t B_STAGE = SCR_B_STAGE_BYTES / CG;   // this CTA's share of a bank tile chunk
  constexpr int B_ROWS = SCR_N / CG;
  const uint32_t cta_rank = (CG == 2) ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0;
  const int nkc = p.nkc, stages = p.stages;
  const uint32_t a_off = 0;
  const uint32_t b_off = a_off + (uint32_t)nkc * SCR_A_CHUNK_BYTES;
  constexpr bool SYM = FILT == 4;
  constexpr bool V3 = FILT == 3 || SYM;          // eight epilogue warps, no shared-memory lists
  const uint32_t lv_off = b_off + (uint32_t)stages * B_STAGE;
  const uint32_t li_off = lv_off + (V3 ? 0 : SCR_CAP * SCR_M * 4);
  const uint32_t hn_off = li_off + (V3 ? 0 : SCR_CAP * SCR_M * 4);
  const uint32_t sq_off = hn_off + 16 * (SCR_N / 2) * 4;   // SYM: per-warp hit queues
  const uint32_t bar_off = SYM ? sq_off + 8 * (SYM_QCAP * 12 + 128)
                               : hn_off + (V3 ? 8 * (SCR_N / 2) * 4 : 2 * SCR_N * 4);
  float* s_lv = reinterpret_cast<float*>(sm + lv_off);
  int* s_li = reinterpret_cast<int*>(sm + li_off);
  float* s_hn = reinterpret_cast<float*>(sm + hn_off);
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + bar_off);
  // barrier slots: [0, S) full, [S, 2S) empty, 2S a_full, 2S+1.. tmem_full[2], tmem_empty[2]
  const uint32_t bar_base = base + bar_off;
  auto bar_full = [&](int s) { return bar_base + 8u * s; };
  auto bar_empty = [&](int s) { return bar_base + 8u * (SCR_MAX_STAGES + s); };
  const uint32_t bar_a = bar_base + 8u * (2 * SCR_MAX_STAGES);
  auto bar_tfull = [&](int s) { return bar_base + 8u * (2 * SCR_MAX_STAGES + 1 + s); };
  auto bar_tempty = [&](int s) { return bar_base + 8u * (2 * SCR_MAX_STAGES + 3 + s); };
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 2 * SCR_MAX_STAGES + 5);
  uint32_t* s_start = s_tmem + 1;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int m_tile = blockIdx.x, split = blockIdx.y;
  int t_begin = split * p.tiles_per_split;
  int t_end = min(p.n_tiles, t_begin + p.tiles_per_split);
  int head_slot = split;
  if (p.tail_splits > 0) {   // hybrid tail: 1-D grid of CTA groups (see ScreenParams)
    const int unit = (int)blockIdx.x / CG, r = (int)blockIdx.x % CG;
    if (unit < p.full_units) {
      m_tile = unit * CG + r; split = 0; t_begin = 0; t_end = p.n_tiles; head_slot = 0;
    } else {
      const int t = unit - p.full_units;
      m_tile = (p.full_units + t / p.tail_splits) * CG + r;
      split = t % p.tail_splits;
      t_begin = split * p.tail_tps;
      t_end = min(p.n_tiles, t_begin + p.tail_tps);
      head_slot = 1 + split;
    }
  }
  if constexpr (SYM) {
    // own bank tile (SCR_N = CG * SCR_M rows: the pair's query rows ARE that tile) and onwards
    // Launch order = longest scan first: the scan length falls with the pair index, except that
    // the pairs from sym_mid_pair on carry sym_extra more tiles (ranks of the upper half of an
    // even world) — those go first, i.e. the local pair index is rotated by (mid - lo).
    if (p.sym_extra > 0) {
      const int n_pairs = (int)gridDim.x / CG;
      int lp = m_tile / CG + (p.sym_mid_pair - p.sym_pair_lo);
      if (lp >= n_pairs) lp -= n_pairs;
      m_tile = lp * CG + m_tile % CG;
    }
    m_tile += p.sym_pair_lo * CG;
    t_begin = m_tile / CG;
    t_end = p.sym_span_end + (t_begin >= p.sym_mid_pair ? p.sym_extra : 0);   // may pass n_tiles
  }
  const int n_local = max(t_end - t_begin, 0);

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_q);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < stages; ++s) {
      mbar_init(bar_full(s), 1);
      mbar_init(bar_empty(s), 1);
    }
    mbar_init(bar_a, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_tfull(s), 1);
      mbar_init(bar_tempty(s), (V3 ? 8 : 4) * CG);
    }
    fence_barrier_init();
    // Shared stream head: CTAs that start later join the bank scan where the running CTAs are
    // (cyclic tile order), so that all co-resident CTAs stream the same window of the bank
    // through L2 instead of each pulling its own copy from HBM.
    *s_start = (p.head != nullptr && leader)
                   ? (uint32_t)*reinterpret_cast<volatile int*>(p.head + head_slot) : 0u;
  }
  if (warp == 1) {
    if constexpr (CG == 2) tmem_alloc_512_pair(smem_u32(s_tmem));
    else tmem_alloc_512(smem_u32(s_tmem));
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (CG == 2) cluster_sync_all();   // peer barriers initialised before any remote signal
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  uint32_t start_v = *s_start;
  if constexpr (CG == 2) {
    if (!leader) start_v = ld_cluster_u32(smem_u32(s_start), 0);   // the pair scans in step
  }
  const int start = n_local > 0 ? (int)(start_v % (uint32_t)n_local) : 0;
  auto tile_of = [&](int j) {   // j-th tile of this CTA's scan
    int tt = start + j;
    if (tt >= n_local) tt -= n_local;
    tt += t_begin;
    if constexpr (SYM) { if (tt >= p.n_tiles) tt -= p.n_tiles; }   // cyclic over the bank
    return tt;
  };
  // FILT 3: the first n_seed tiles are scored twice (seed phase, then again in the main phase)
  const int n_seed = (FILT >= 2 && !SYM) ? min(p.seed_tiles, n_local) : 0;
  const int n_total = n_local + n_seed;
  auto tile_seq = [&](int jj) { return tile_of(jj < n_seed ? jj : jj - n_seed); };

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0 && n_local > 0) {
      // In pair mode both CTAs' loads complete on the LEADER's barriers; only the leader arms them
      // (with the bytes of both CTAs).
      if (leader) mbar_expect_tx(bar_a, (uint32_t)nkc * SCR_A_CHUNK_BYTES * CG);
      for (int kc = 0; kc < nkc; ++kc) {
        if constexpr (CG == 2)
          tma_load_2d_pair(base + a_off + kc * SCR_A_CHUNK_BYTES, &tmap_q, kc * SCR_KC,
                           m_tile * SCR_M, bar_a);
        else
          tma_load_2d(base + a_off + kc * SCR_A_CHUNK_BYTES, &tmap_q, kc * SCR_KC,
                      m_tile * SCR_M, bar_a);
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int j = 0; j < n_total; ++j) {
        const int t = tile_seq(j);
        if (p.head != nullptr && leader && (j & 7) == 7 && j >= n_seed)
          atomicMax(p.head + head_slot, (int)(start_v + (uint32_t)(j - n_seed)));
        const int brow = t * SCR_N + (int)cta_rank * B_ROWS;
        for (int kc = 0; kc < nkc; ++kc) {
          // pull this CTA's share of a later tile into L2 so the ring only sees L2 latency
          if (p.prefetch > 0 && t + p.prefetch < t_end)
            tma_prefetch_l2_2d(&tmap_b, kc * SCR_KC, brow + p.prefetch * SCR_N);
          mbar_wait(bar_empty(stage), phase ^ 1u);
          if (leader) mbar_expect_tx(bar_full(stage), SCR_B_STAGE_BYTES);
          if constexpr (CG == 2)
            tma_load_2d_pair(base + b_off + stage * B_STAGE, &tmap_b, kc * SCR_KC, brow,
                             bar_full(stage));
          else
            tma_load_2d(base + b_off + stage * B_STAGE, &tmap_b, kc * SCR_KC, brow,
                        bar_full(stage));
          if (++stage == stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0 && n_local > 0 && leader) {
      mbar_wait(bar_a, 0);
      tc_fence_after();
      int stage = 0;
      uint32_t phase = 0;
      long long w_empty = 0, w_full = 0;
      const long long m0 = (EXB_SCREEN_STATS && p.stats) ? clock64() : 0;
      for (int j = 0; j < n_total; ++j) {
        const int as = j & 1;
        const long long c0 = (EXB_SCREEN_STATS && p.stats) ? clock64() : 0;
        mbar_wait(bar_tempty(as), (((uint32_t)j >> 1) & 1u) ^ 1u);
        if (EXB_SCREEN_STATS && p.stats) w_empty += clock64() - c0;
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)as * SCR_N;
        for (int kc = 0; kc < nkc; ++kc) {
          const long long c1 = (EXB_SCREEN_STATS && p.stats) ? clock64() : 0;
          mbar_wait(bar_full(stage), phase);
          if (EXB_SCREEN_STATS && p.stats) w_full += clock64() - c1;
          tc_fence_after();
          const uint64_t adesc = make_sw128_desc(base + a_off + kc * SCR_A_CHUNK_BYTES);
          const uint64_t bdesc = make_sw128_desc(base + b_off + stage * B_STAGE);
#pragma unroll
          for (int ki = 0; ki < SCR_KC / 16; ++ki) {
            // advance 16 fp16 = 32 bytes inside the swizzle atom: +2 in the (addr >> 4) field
            if constexpr (CG == 2)
              umma_f16_pair(d_tmem, adesc + (uint64_t)(2 * ki), bdesc + (uint64_t)(2 * ki),
                            umma_idesc_f16(2 * SCR_M, SCR_N), (kc | ki) != 0 ? 1u : 0u);
            else
              umma_f16(d_tmem, adesc + (uint64_t)(2 * ki), bdesc + (uint64_t)(2 * ki),
                       umma_idesc_f16(SCR_M, SCR_N), (kc | ki) != 0 ? 1u : 0u);
          }
          if constexpr (CG == 2) umma_commit_pair(bar_empty(stage));
          else umma_commit(bar_empty(stage));
          if (++stage == stages) { stage = 0; phase ^= 1u; }
        }
        if constexpr (CG == 2) umma_commit_pair(bar_tfull(as));
        else umma_commit(bar_tfull(as));
      }
      if (EXB_SCREEN_STATS && p.stats) {
        atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 4), (unsigned long long)w_empty);
        atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 5), (unsigned long long)w_full);
        atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 6),
                  (unsigned long long)(clock64() - m0));
        atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 7), 1ull);
      }
    }
  } else if constexpr (SYM) {
    // ===================== symmetric epilogue, eight warps: (lane quadrant, column half) ========
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    const int e = quad * 32 + lane;
    const int64_t row = (int64_t)m_tile * SCR_M + e;
    const bool row_ok = row < p.nq;
    const int cap = p.log_cap;
    const int64_t o = (row_ok ? row : 0) * 2 + half;
    SymQueue sq;
    {
      uint8_t* qb = sm + sq_off + (warp - 2) * (SYM_QCAP * 12 + 128);
      sq.target = reinterpret_cast<int*>(qb);
      sq.val = reinterpret_cast<float*>(qb + SYM_QCAP * 4);
      sq.src = reinterpret_cast<int*>(qb + SYM_QCAP * 8);
      sq.scratch = reinterpret_cast<float*>(qb + SYM_QCAP * 12);
      sq.g_target = p.tq_target; sq.g_val = p.tq_val; sq.g_src = p.tq_src;
      sq.g_count = p.tq_count; sq.g_cap = p.tq_cap;
      sq.sys = false;
    }
    // peer routing: [dest_lo, dest_hi) = bank tiles of the rank the queue currently drains to
    const SymPeerTable* pt = p.sym_peers;
    int dest_lo = 0, dest_hi = pt != nullptr ? 0 : 0x7fffffff;
    int qn = 0;                                   // queued transposed hits (warp-uniform)
    // rows past nq: unreachable thresholds in both directions
    SymRow me;
    me.tkeep_i = row_ok ? __ldg(p.sym_tkeep + row) : FLT_MAX;
    me.hb_i = row_ok ? __ldg(p.ha + row) : INFINITY;              // this row as a BANK row
    me.hb_t = me.hb_i;
    me.row = (int)(row_ok ? row : 0);
    int nd = 0;
    float* hn = s_hn + (warp - 2) * (SCR_N / 2);                  // 128 bank half norms
    float* tk = s_hn + (8 + warp - 2) * (SCR_N / 2);              // 128 column-row thresholds
    auto load4 = [&](const float* arr, int t, float fill) -> float4 {
      const int64_t c0 = (int64_t)t * SCR_N + half * (SCR_N / 2) + 4 * lane;
      float4 h;
      if (c0 + 3 < p.nb) {
        h = __ldg(reinterpret_cast<const float4*>(arr + c0));
      } else {
        h.x = (c0 < p.nb) ? arr[c0] : fill;
        h.y = (c0 + 1 < p.nb) ? arr[c0 + 1] : fill;
        h.z = (c0 + 2 < p.nb) ? arr[c0 + 2] : fill;
        h.w = (c0 + 3 < p.nb) ? arr[c0 + 3] : fill;
      }
      return h;
    };
    float4 hnext = make_float4(0.f, 0.f, 0.f, 0.f), tnext = hnext;
    if (n_total > 0) {
      hnext = load4(p.hb, tile_seq(0), INFINITY);
      tnext = load4(p.sym_tkeep, tile_seq(0), INFINITY);
    }
    constexpr int NCH = SCR_N / 2 / 32;
    for (int jj = 0; jj < n_total; ++jj) {
      const int t = tile_seq(jj);
      const int as = jj & 1;
      const bool diag = jj == 0;                 // own tile: the direct test covers both orders
      if (t < dest_lo || t >= dest_hi) {         // (peer table only) rows of another rank from here
        sym_queue_flush(sq, qn, lane);
        int r = 0;
        while (r + 1 < pt->ranks && t >= pt->tile_lo[r + 1]) ++r;
        dest_lo = pt->tile_lo[r]; dest_hi = pt->tile_lo[r + 1];
        sq.g_target = pt->target[r]; sq.g_val = pt->val[r]; sq.g_src = pt->src[r];
        sq.g_count = pt->count[r]; sq.g_cap = pt->cap[r];
        sq.sys = true;
      }
      __syncwarp();
      reinterpret_cast<float4*>(hn)[lane] = hnext;
      reinterpret_cast<float4*>(tk)[lane] = tnext;
      __syncwarp();
      if (jj + 1 < n_total) {
        hnext = load4(p.hb, tile_seq(jj + 1), INFINITY);
        tnext = load4(p.sym_tkeep, tile_seq(jj + 1), INFINITY);
      }
      mbar_wait(bar_tfull(as), ((uint32_t)jj >> 1) & 1u);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)as * SCR_N +
                             (uint32_t)half * (SCR_N / 2);
      const int col_base = t * SCR_N + half * (SCR_N / 2);
      me.hb_t = diag ? INFINITY : me.hb_i;        // transposed test off on the diagonal tile
      uint32_t ra[32], rb[32];
      tmem_ld32(t_row, ra);
#pragma unroll 1
      for (int ch = 0; ch < NCH; ch += 2) {
        tmem_ld_wait();
        tmem_ld32(t_row + (ch + 1) * 32, rb);
        sym_chunk(ra, hn + ch * 32, tk + ch * 32, me, col_base + ch * 32, cap, p.dlog_val,
                  p.dlog_idx, half, nd, sq, qn, lane, p.debug != 3);
        tmem_ld_wait();
        if (ch + 2 < NCH) tmem_ld32(t_row + (ch + 2) * 32, ra);
        sym_chunk(rb, hn + (ch + 1) * 32, tk + (ch + 1) * 32, me, col_base + (ch + 1) * 32, cap,
                  p.dlog_val, p.dlog_idx, half, nd, sq, qn, lane, p.debug != 3);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG == 2) mbar_arrive_leader(bar_tempty(as));
        else mbar_arrive(bar_tempty(as));
      }
    }
    sym_queue_flush(sq, qn, lane);
    if (row_ok) p.dlog_cnt[o] = nd > cap ? -1 : nd;
  } else if constexpr (V3) {
    // ===================== epilogue, eight warps: (lane quadrant, column half) ================
    const int quad = warp & 3;                    // TMEM lane quadrant this warp may access
    const int half = (warp - 2) >> 2;             // columns [128 half, 128 half + 128) of a tile
    const int e = quad * 32 + lane;               // row inside the tile
    const int64_t row = (int64_t)m_tile * SCR_M + e;
    const bool row_ok = row < p.nq;
    const int64_t o = (row_ok ? row : 0) * p.lists + split * 2 + half;
    const int cap = p.scan_cap;
    float* gv = p.scan_val + o * SCR_SCAN_CAP;    // in-scan scratch list of this tracker
    int* gi = p.scan_idx + o * SCR_SCAN_CAP;
    float* hn = s_hn + (warp - 2) * (SCR_N / 2);  // this warp's private copy of its 128 half norms
    ListState st;
    st.n = 0;
    st.dropped = -FLT_MAX;
    st.tk = -FLT_MAX;
    // padding rows past nq own no list: an unreachable threshold keeps them out of the slow path
    st.tkeep = row_ok ? -FLT_MAX : FLT_MAX;
#pragma unroll
    for (int i = 0; i < SCR_MAX_K; ++i) st.tv[i] = (i < SCR_MAX_K - p.k) ? FLT_MAX : -FLT_MAX;
    st.ha = row_ok ? p.ha[row] : 0.f;
    const float eb_max = p.bsc->err_max, hb_max = p.bsc->hn_max;
    st.e2 = 2.f * ((row_ok ? p.ea[row] : 0.f) + eb_max);
    st.dl = (st.ha + hb_max) * SCR_ACC_EPS;

    auto load_hn4 = [&](int t) -> float4 {
      const int64_t c0 = (int64_t)t * SCR_N + half * (SCR_N / 2) + 4 * lane;
      float4 h;
      if (c0 + 3 < p.nb) {
        h = __ldg(reinterpret_cast<const float4*>(p.hb + c0));
      } else {
        h.x = (c0 < p.nb) ? p.hb[c0] : INFINITY;
        h.y = (c0 + 1 < p.nb) ? p.hb[c0 + 1] : INFINITY;
        h.z = (c0 + 2 < p.nb) ? p.hb[c0 + 2] : INFINITY;
        h.w = (c0 + 3 < p.nb) ? p.hb[c0 + 3] : INFINITY;
      }
      return h;
    };
    float4 hnext = make_float4(0.f, 0.f, 0.f, 0.f);
    if (n_total > 0) hnext = load_hn4(tile_seq(0));
    long long e_wait = 0, e_work = 0, e_slow[6] = {0, 0, 0, 0, 0, 0};
    long long* slow = (EXB_SCREEN_STATS && p.stats) ? e_slow : nullptr;
    constexpr int NCH = SCR_N / 2 / 32;            // 4 chunks of 32 columns per tile and tracker

    for (int jj = 0; jj < n_total; ++jj) {
      const int t = tile_seq(jj);
      const int as = jj & 1;
      __syncwarp();                                // every lane is done with the previous tile's hn
      reinterpret_cast<float4*>(hn)[lane] = hnext;
      __syncwarp();
      if (jj + 1 < n_total) hnext = load_hn4(tile_seq(jj + 1));
      const long long c0 = (EXB_SCREEN_STATS && p.stats) ? clock64() : 0;
      mbar_wait(bar_tfull(as), ((uint32_t)jj >> 1) & 1u);
      const long long c1 = (EXB_SCREEN_STATS && p.stats) ? clock64() : 0;
      e_wait += c1 - c0;
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)as * SCR_N +
                             (uint32_t)half * (SCR_N / 2);
      const int col_base = t * SCR_N + half * (SCR_N / 2);
      uint32_t ra[32], rb[32];
      tmem_ld32(t_row, ra);
      if (jj < n_seed) {
        // seed phase: group maxima of the sample tiles feed the chain only (nothing is recorded;
        // the same tiles come round again in the main phase)
#pragma unroll 1
        for (int ch = 0; ch < NCH; ch += 2) {
          tmem_ld_wait();
          tmem_ld32(t_row + (ch + 1) * 32, rb);
          seed_chunk(st, ra, hn + ch * 32);
          tmem_ld_wait();
          if (ch + 2 < NCH) tmem_ld32(t_row + (ch + 2) * 32, ra);
          seed_chunk(st, rb, hn + (ch + 1) * 32);
        }
        if (jj == n_seed - 1) {
          refresh_threshold(st);
          if (p.chain_reset) chain_reset(st, p.k);
        }
      } else {
        // main phase: every inserted score goes through the (fresh) chain exactly once
        const bool push = p.chain_reset ? true : jj >= 2 * n_seed;
#pragma unroll 1
        for (int ch = 0; ch < NCH; ch += 2) {
          tmem_ld_wait();
          tmem_ld32(t_row + (ch + 1) * 32, rb);
          filter_chunk_g(st, ra, hn + ch * 32, gv, gi, cap, col_base + ch * 32, push, slow);
          tmem_ld_wait();
          if (ch + 2 < NCH) tmem_ld32(t_row + (ch + 2) * 32, ra);
          filter_chunk_g(st, rb, hn + (ch + 1) * 32, gv, gi, cap, col_base + (ch + 1) * 32, push,
                         slow);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG == 2) mbar_arrive_leader(bar_tempty(as));
        else mbar_arrive(bar_tempty(as));
      }
      if (EXB_SCREEN_STATS && p.stats) e_work += clock64() - c1;
    }
    if (EXB_SCREEN_STATS && p.stats) {
      if (lane == 0) {
        atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 0), (unsigned long long)e_wait);
        atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 1), (unsigned long long)e_work);
        atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 2), (unsigned long long)e_slow[0]);
        atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 3), (unsigned long long)e_slow[1]);
      }
      // per-lane: inserts, cycles storing, repacks, cycles in chain + threshold refresh
      for (int q = 2; q < 6; ++q)
        atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 10 + q),
                  (unsigned long long)e_slow[q]);
    }
    if (row_ok) {
      // final prune: scratch list -> the row's SCR_CAP result slots, entries below the final tkeep go
      const int w = list_compact_g(gv, gi, st.n, st.tkeep, p.cand_val + o * SCR_CAP,
                                   p.cand_idx + o * SCR_CAP, SCR_CAP);
      const bool overflow = w > SCR_CAP || (st.dropped > -FLT_MAX && st.dropped >= st.tkeep);
      p.cand_cnt[o] = overflow ? -1 : w;
    }
  } else {
    // ===================== epilogue: filter accumulators =====================
    const int quad = warp & 3;                    // TMEM lane quadrant this warp may access
    const int e = quad * 32 + lane;               // row inside the tile
    const int et = (warp - 2) * 32 + lane;        // 0..127 linear epilogue thread id
    const int64_t row = (int64_t)m_tile * SCR_M + e;
    const bool row_ok = row < p.nq;
    float* lv = s_lv + e;
    int* li = s_li + e;
    ListState st;
    st.n = 0;
    st.dropped = -FLT_MAX;
    st.tk = -FLT_MAX;
    st.tkeep = -FLT_MAX;
#pragma unroll
    for (int i = 0; i < SCR_MAX_K; ++i) st.tv[i] = (i < SCR_MAX_K - p.k) ? FLT_MAX : -FLT_MAX;
    st.ha = row_ok ? p.ha[row] : 0.f;
    const float eb_max = p.bsc->err_max, hb_max = p.bsc->hn_max;
    st.e2 = 2.f * ((row_ok ? p.ea[row] : 0.f) + eb_max);
    st.dl = (st.ha + hb_max) * SCR_ACC_EPS;

    auto load_hn = [&](int t, float& h0, float& h1) {
      const int64_t c0 = (int64_t)t * SCR_N + 2 * et;
      h0 = (c0 < p.nb) ? p.hb[c0] : INFINITY;
      h1 = (c0 + 1 < p.nb) ? p.hb[c0 + 1] : INFINITY;
    };
    float h0 = 0.f, h1 = 0.f;
    if (n_total > 0) load_hn(tile_seq(0), h0, h1);
    long long e_wait = 0, e_work = 0, e_slow[2] = {0, 0};
    long long* slow = (EXB_SCREEN_STATS && p.stats) ? e_slow : nullptr;

    for (int jj = 0; jj < n_total; ++jj) {
      const int j = jj < n_seed ? jj : jj - n_seed;   // tile of the scan (sample tiles come twice)
      const int t = tile_of(j);
      const int as = jj & 1;
      float* hn = s_hn + as * SCR_N;
      hn[2 * et] = h0;
      hn[2 * et + 1] = h1;
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (jj + 1 < n_total) load_hn(tile_seq(jj + 1), h0, h1);
      const long long c0 = (EXB_SCREEN_STATS && p.stats) ? clock64() : 0;
      mbar_wait(bar_tfull(as), ((uint32_t)jj >> 1) & 1u);
      const long long c1 = (EXB_SCREEN_STATS && p.stats) ? clock64() : 0;
      e_wait += c1 - c0;
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)as * SCR_N;
      if (p.debug != 1) {
        uint32_t ra[32], rb[32];
        // the seed chain is dropped after the seed phase (chain_reset), so every score of the main
        // phase is pushed exactly once
        const bool first = p.chain_reset ? false : (n_seed > 0 ? (jj < 2 * n_seed) : (j == 0));
        const bool seed_now = n_seed > 0 ? (jj < n_seed) : (j == 0);
        if (seed_now && p.debug == 0) {
          // seed pass: group maxima -> chain only.  Without sample tiles (n_seed == 0) this is a
          // first pass over tile 0, whose accumulator stays in TMEM for the real pass below.
          tmem_ld32(t_row, ra);
#pragma unroll 1
          for (int ch = 0; ch < SCR_N / 32; ch += 2) {
            tmem_ld_wait();
            tmem_ld32(t_row + (ch + 1) * 32, rb);
            seed_chunk(st, ra, hn + ch * 32);
            tmem_ld_wait();
            if (ch + 2 < SCR_N / 32) tmem_ld32(t_row + (ch + 2) * 32, ra);
            seed_chunk(st, rb, hn + (ch + 1) * 32);
          }
          if (n_seed == 0 || jj == n_seed - 1) {
            refresh_threshold(st);
            if (p.chain_reset) chain_reset(st, p.k);
          }
        }
        if (jj >= n_seed) {
        // Software-pipelined drain: the TMEM load of chunk c+1 is in flight while chunk c is
        // filtered (tcgen05.wait::ld waits for every outstanding load of the thread, so the order
        // is wait -> issue next -> process current, alternating two register buffers).
        tmem_ld32(t_row, ra);
#pragma unroll 1
        for (int ch = 0; ch < SCR_N / 32; ch += 2) {
          tmem_ld_wait();
          tmem_ld32(t_row + (ch + 1) * 32, rb);
          if (p.debug == 0) {
            if constexpr (FILT == 1)
              filter_chunk(st, ra, t * SCR_N + ch * 32, hn + ch * 32, lv, li, !first);
            else
              filter_chunk_bits(st, ra, ch, hn, lv, li, t, p.k, !first, slow);
          }
          tmem_ld_wait();
          if (ch + 2 < SCR_N / 32) tmem_ld32(t_row + (ch + 2) * 32, ra);
          if (p.debug == 0) {
            if constexpr (FILT == 1)
              filter_chunk(st, rb, t * SCR_N + (ch + 1) * 32, hn + (ch + 1) * 32, lv, li, !first);
            else
              filter_chunk_bits(st, rb, ch + 1, hn, lv, li, t, p.k, !first, slow);
          }
        }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG == 2) mbar_arrive_leader(bar_tempty(as));
        else mbar_arrive(bar_tempty(as));
      }
      if (EXB_SCREEN_STATS && p.stats) e_work += clock64() - c1;
    }
    if (EXB_SCREEN_STATS && p.stats && lane == 0) {
      atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 0), (unsigned long long)e_wait);
      atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 1), (unsigned long long)e_work);
      atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 2), (unsigned long long)e_slow[0]);
      atomicAdd(reinterpret_cast<unsigned long long*>(p.stats + 3), (unsigned long long)e_slow[1]);
    }

    if (row_ok) {
      const int64_t o = row * p.splits + split;
      list_compact(st, lv, li);
      const bool overflow = st.dropped > -FLT_MAX && st.dropped >= st.tkeep;
      p.cand_cnt[o] = overflow ? -1 : st.n;
      for (int c = 0; c < st.n; ++c) {
        p.cand_idx[o * SCR_CAP + c] = li[c * SCR_M];
        p.cand_val[o * SCR_CAP + c] = lv[c * SCR_M];
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if constexpr (CG == 2) cluster_sync_all();   // the pair shares TMEM/smem until both are done
  if (warp == 1) {
    tc_fence_after();
    if constexpr (CG == 2) tmem_dealloc_512_pair(tmem_base);
    else tmem_dealloc_512(tmem_base);
  }
}

}  // namespace exb

#include "knn_lat.cuh"
#include "knn_rerank.cuh"
#include "knn_chain.cuh"
#include "knn_sym.cuh"

namespace exb {

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
// fp16 [rows, dpad] row-major -> tiles of box_rows x 64 columns, 128-byte swizzle, zero OOB fill
static int make_tmap(CUtensorMap* map, const void* ptr, int64_t rows, int dpad, int box_rows) {
  return tmap_encode_2d(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, ptr, rows, dpad,
                        (int64_t)dpad * 2, SCR_KC, box_rows);
}

// Optional CUDA-event bracket around every screening-kernel launch (bench.py's roofline leg).
struct ScreenProfile {
  bool on = false;
  int n = 0;
  cudaEvent_t e0[64], e1[64];
  bool created = false;
};
static ScreenProfile g_prof;

static inline int dim_pad(int dim) { return (dim + SCR_KC - 1) / SCR_KC * SCR_KC; }

struct ScreenPlan {
  bool lat; // per-update sizes: latency kernel (64-row bank tiles, everything resident, knn_lat.cuh)
  int cg;   // CTAs cooperating on one MMA (1, or 2 = cta_group::2 pair)
  int dpad, nkc, n_tiles, m_tiles, splits, tiles_per_split, stages;
  int filt;      // streaming kernel variant (knn_screen_kernel<CG, FILT>)
  int trackers;  // candidate lists per (row, split): 2 for FILT 3, else 1
  size_t smem;
  int fb_cap, fb_splits;
  int tail_splits, full_units, tail_tps;   // hybrid tail (ScreenParams); tail_splits 0 = off
};

static bool make_plan(int64_t nq, int64_t nb, int dim, int k, ScreenPlan* pl) {
  if (dim < 1 || dim > EXB_SCREEN_MAX_DIM || k < 1 || k > SCR_MAX_K || nq < 1 || nb < 1)
    return false;
  pl->dpad = dim_pad(dim);
  pl->nkc = pl->dpad / SCR_KC;
  pl->fb_cap = (int)(nq < FALLBACK_CAP ? nq : FALLBACK_CAP);
  int fbs = (int)((nb + 127) / 128);
  if (fbs > FALLBACK_SPLITS) fbs = FALLBACK_SPLITS;
  if (fbs * k > 512) fbs = 512 / k;
  pl->fb_splits = fbs < 1 ? 1 : fbs;
  // ---- latency mode: few query tiles, a bank that one wave of 64-row tiles (x <= 8) covers
  static const int lat_enabled = knob("EXB_SCREEN_LAT", 1);
  pl->lat = false;
  {
    const int mt = (int)((nq + SCR_M - 1) / SCR_M);
    const int64_t nt = (nb + LAT_N - 1) / LAT_N;
    int sp = sm_count() / mt;
    if (sp > 16) sp = 16;
    if (sp > nt) sp = (int)nt;
    if (lat_enabled && mt <= sm_count() && sp >= 1 && nt <= (int64_t)sp * 8) {
      pl->lat = true;
      pl->cg = 1;
      pl->m_tiles = mt;
      pl->n_tiles = (int)nt;
      pl->tiles_per_split = (int)((nt + sp - 1) / sp);
      pl->splits = (int)((nt + pl->tiles_per_split - 1) / pl->tiles_per_split);
      pl->stages = 0;
      pl->filt = 1;
      pl->trackers = 1;
      pl->smem = (size_t)pl->nkc * (SCR_A_CHUNK_BYTES + LAT_B_CHUNK_BYTES) +
                 2 * (SCR_CAP + 1) * LAT_LS * 4 + 32 + 2 * LAT_N * 4 + 64 + 13 * 8 + 16 + 1024;
      return true;
    }
  }
  pl->n_tiles = (int)((nb + SCR_N - 1) / SCR_N);
  pl->m_tiles = (int)((nq + SCR_M - 1) / SCR_M);
  // CTA pairs (cta_group::2) halve each CTA's share of the bank-tile traffic and leave room for a
  // 4-deep TMA ring; EXB_SCREEN_CG=1 forces the single-CTA kernel (A/B measurements).
  static const int forced_cg = knob("EXB_SCREEN_CG", 0);
  pl->cg = (forced_cg == 1 || forced_cg == 2) ? forced_cg : (pl->m_tiles >= 2 ? 2 : 1);
  pl->m_tiles = (pl->m_tiles + pl->cg - 1) / pl->cg * pl->cg;
  int splits = sm_count() / pl->m_tiles;
  if (splits > pl->n_tiles) splits = pl->n_tiles;
  if (splits > 16) splits = 16;
  if (splits < 1) {
    // More query tiles than SMs: one bank range per CTA.  Cutting the bank into s <= 4 ranges
    // (shorter work units, a cheaper partial last wave) was measured on 125k queries x 1M rows
    // (978 CTAs = 6.6 waves): 93.4 ms (s = 1) vs 95.6 / 96.2 / 97.8 ms (s = 2 / 3 / 4) — the CTAs
    // do not finish in lock-step, so the wave model does not hold, and every extra range restarts
    // the k-th-best chain.  Kept as an A/B knob only (EXB_SCREEN_WAVESPLITS=1..4).
    splits = 1;
    static const int wave_splits = knob("EXB_SCREEN_WAVESPLITS", 0);
    if (wave_splits > 0) splits = wave_splits > 4 ? 4 : wave_splits;
    if (splits > pl->n_tiles) splits = pl->n_tiles;
  }
  pl->tiles_per_split = (pl->n_tiles + splits - 1) / splits;
  pl->splits = (pl->n_tiles + pl->tiles_per_split - 1) / pl->tiles_per_split;
  // variant: short scans -> 1 (the list fill phase dominates); mid-length scans (a bank shard of
  // <= 262k rows per CTA) -> 3 (eight epilogue warps: the hit handling, which stalls the MMA on
  // short scans, runs on twice the warps; measured 103 vs 114 ms on a 125k-row shard); long scans
  // -> 2 (four epilogue warps: fewest instructions, the kernel is power-bound there; measured 762
  // vs 769 ms on 1M rows).  EXB_SCREEN_FILTER=2|3 forces one of them (A/B).
  static const int long_filt = [] {
    const int v = knob("EXB_SCREEN_FILTER", 0);
    return (v == 2 || v == 3) ? v : 0;
  }();
  pl->filt = pl->tiles_per_split <= 16 ? 1
             : long_filt ? long_filt : (pl->tiles_per_split <= 1024 ? 3 : 2);
  if (pl->filt == 3 && pl->splits > 8) {   // the rerank gathers at most 16 lists per row
    pl->tiles_per_split = (pl->n_tiles + 7) / 8;
    pl->splits = (pl->n_tiles + pl->tiles_per_split - 1) / pl->tiles_per_split;
  }
  pl->trackers = pl->filt == 3 ? 2 : 1;
  // Hybrid tail.  More query tiles than SMs run in waves of sm_count CTAs; when the last wave is
  // partly empty (125k queries: 489 CTA pairs on 74 pair slots = 6 full waves + 45 pairs; 250k
  // queries: 13 waves + 15 pairs) its query tiles are cut into 2-4 bank ranges so that the idle
  // SMs take part.  Only those rows pay for restarted k-th-best chains and extra lists.
  pl->tail_splits = 0; pl->full_units = 0; pl->tail_tps = 0;
  static const int tail_env = knob("EXB_SCREEN_TAIL", 1);   // A/B: 0 = off
  if (tail_env && pl->filt == 2 && pl->splits == 1) {
    const int slots = sm_count() / pl->cg, units = pl->m_tiles / pl->cg;
    const int full = units / slots * slots, rem = units - full;
    // ts ranges per tail tile -> the tail takes ceil(rem * ts / slots) rounds of 1/ts wave
    int ts = 0;
    double best = 0.8;   // worth it only if the tail shrinks to < 0.8 of a wave
    for (int c = 2; c <= 4 && rem > 0; ++c) {
      const double cost = (double)((rem * c + slots - 1) / slots) / c;
      if (cost < best - 1e-9) {
        best = cost;
        ts = c;
      }
    }
    if (full > 0 && ts >= 2 && pl->n_tiles / ts >= 256) {
      pl->tail_splits = ts;
      pl->full_units = full;
      pl->tail_tps = (pl->n_tiles + ts - 1) / ts;
      pl->splits = ts;          // candidate lists per row (rows of full units use list 0 only)
    }
  }
  const size_t fixed = (size_t)pl->nkc * SCR_A_CHUNK_BYTES +
                       (pl->filt == 3 ? 8 * (SCR_N / 2) * 4 : 2 * SCR_CAP * SCR_M * 4 + 2 * SCR_N * 4) +
                       (2 * SCR_MAX_STAGES + 5) * 8 + 16 + 1024;
  const int b_stage = SCR_B_STAGE_BYTES / pl->cg;
  int stages = (int)((SCR_SMEM_LIMIT - fixed) / b_stage);
  if (stages > SCR_MAX_STAGES) stages = SCR_MAX_STAGES;
  if (stages < 2) return false;
  pl->stages = stages;
  pl->smem = fixed + (size_t)stages * b_stage;
  return true;
}

template <bool LAT, bool SH = false, bool COOP = false>
static void launch_rerank(const RerankParams& rp, unsigned grid, int ni, cudaStream_t st) {
  if constexpr (LAT) {   // per-update chain: dependent launch behind the latency screen
    switch (ni) {
      case 1: launch_pdl(knn_rerank_kernel<LAT, 1, SH, COOP>, dim3(grid), dim3(RR_WARPS * 32), 0, st, rp); break;
      case 2: launch_pdl(knn_rerank_kernel<LAT, 2, SH, COOP>, dim3(grid), dim3(RR_WARPS * 32), 0, st, rp); break;
      case 3: launch_pdl(knn_rerank_kernel<LAT, 3, SH, COOP>, dim3(grid), dim3(RR_WARPS * 32), 0, st, rp); break;
      case 4: launch_pdl(knn_rerank_kernel<LAT, 4, SH, COOP>, dim3(grid), dim3(RR_WARPS * 32), 0, st, rp); break;
      default: launch_pdl(knn_rerank_kernel<LAT, 0, SH, COOP>, dim3(grid), dim3(RR_WARPS * 32), 0, st, rp); break;
    }
    return;
  }
  switch (ni) {
    case 1: knn_rerank_kernel<LAT, 1, SH><<<grid, RR_WARPS * 32, 0, st>>>(rp); break;
    case 2: knn_rerank_kernel<LAT, 2, SH><<<grid, RR_WARPS * 32, 0, st>>>(rp); break;
    case 3: knn_rerank_kernel<LAT, 3, SH><<<grid, RR_WARPS * 32, 0, st>>>(rp); break;
    case 4: knn_rerank_kernel<LAT, 4, SH><<<grid, RR_WARPS * 32, 0, st>>>(rp); break;
    default: knn_rerank_kernel<LAT, 0, SH><<<grid, RR_WARPS * 32, 0, st>>>(rp); break;
  }
}

// Per-update chain: how rows that lost bank ranges are recomputed inside the rerank — 0 (default)
// every row's own warp scans its ranges, 1 the rerank block scans the union together
// (knn_rerank.cuh:coop_scan).  Process-wide, set through exb_knn_set_per_update_scan.
static std::atomic<int> g_coop_scan{0};

// Launch with programmatic stream serialization (see common.cuh) when EXB_PDL=1.  Measured on
// B200 (profiles/r01g_secondary_pdl.jsonl vs _nopdl): C1 39.9 us either way (stream), 36.9 vs
// 34.8 us as a CUDA graph, C2 33.8 vs 35.8 us — inside the 2 us resolution of the event timer, so
// the ordinary launch stays the default and the dependent-launch chain an A/B knob.
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t st, Args... args) {
  static const int pdl_on = knob("EXB_PDL", 0);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_on ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

struct ScreenWs {
  long long* trace;   // 16 phase timestamps (debug, EXB_SCREEN_TRACE)
  int* head;          // 16 shared stream heads (one per bank split), zeroed per call
  RerankStatus* status;
  PrepScalars* qsc;
  PrepScalars* bsc;
  __half* q16;
  __half* b16;
  float *ha, *ea, *hb, *eb;
  float* cand_val;
  int *cand_idx, *cand_cnt, *fb_rows;
  float* cand_tk;
  float* cand_tr;
  float* scan_val;    // FILT 3 scratch lists [nq, lists, SCR_SCAN_CAP]
  int* scan_idx;
  double2* pbe_partials;   // fused per-update epilogue: one per rerank block
  float* fb_d2;
  int* fb_idx;
  int* coop_flags;    // per-update chain: [m_tiles, splits, 4] (LatParams::coop_flags)
  float* aff;         // streaming: affine record of the call (AFF_FLOATS)
  float* cs_stats;    // streaming: [3][dim] column statistics
  float* cs_part_b;   // streaming: [CS_MAX_BLOCKS][3][dim] partials of the bank / of the queries
  float* cs_part_q;
};

static size_t carve_screen(void* ws, int64_t nq, int64_t nb, const ScreenPlan& pl, int k,
                           bool same, ScreenWs* o) {
  Carver cv(ws ? ws : reinterpret_cast<void*>(256));  // dry run computes the size only
  o->status = cv.take<RerankStatus>(1);
  o->qsc = cv.take<PrepScalars>(1);
  o->bsc = cv.take<PrepScalars>(1);
  o->trace = cv.take<long long>(16);
  o->head = cv.take<int>(16);
  o->b16 = cv.take<__half>((size_t)nb * pl.dpad);
  o->q16 = same ? o->b16 : cv.take<__half>((size_t)nq * pl.dpad);
  o->hb = cv.take<float>(nb);
  o->eb = cv.take<float>(nb);
  o->ha = same ? o->hb : cv.take<float>(nq);
  o->ea = same ? o->eb : cv.take<float>(nq);
  const size_t lists = (size_t)pl.splits * pl.trackers;
  o->cand_val = cv.take<float>((size_t)nq * lists * SCR_CAP);
  o->cand_idx = cv.take<int>((size_t)nq * lists * SCR_CAP);
  o->cand_cnt = cv.take<int>((size_t)nq * lists);
  o->cand_tk = cv.take<float>((size_t)nq * lists);
  o->cand_tr = cv.take<float>((size_t)nq * lists);
  o->scan_val = cv.take<float>(pl.filt == 3 ? (size_t)nq * lists * SCR_SCAN_CAP : 0);
  o->scan_idx = cv.take<int>(pl.filt == 3 ? (size_t)nq * lists * SCR_SCAN_CAP : 0);
  // one per rerank block of the launch chain, or one per block of the fused chain kernel
  o->pbe_partials = cv.take<double2>(
      pl.lat ? std::max<size_t>((size_t)((nq + RR_WARPS - 1) / RR_WARPS), 256) : 1);
  o->fb_rows = cv.take<int>(pl.fb_cap);
  o->fb_d2 = cv.take<float>((size_t)pl.fb_cap * pl.fb_splits * k);
  o->fb_idx = cv.take<int>((size_t)pl.fb_cap * pl.fb_splits * k);
  o->coop_flags = cv.take<int>(pl.lat ? (size_t)pl.m_tiles * pl.splits * 4 : 0);
  const size_t part = pl.lat ? 0 : (size_t)CS_MAX_BLOCKS * 3 * pl.dpad;
  o->aff = cv.take<float>(pl.lat ? 0 : AFF_FLOATS);
  o->cs_stats = cv.take<float>(pl.lat ? 0 : 3 * (size_t)pl.dpad);
  o->cs_part_b = cv.take<float>(part);
  o->cs_part_q = cv.take<float>(same ? 0 : part);
  return cv.off + 256;
}

static int launch_prep(const float* x, int64_t n, int dim, int dpad, __half* out, float* hn,
                       float* er, PrepScalars* sc, cudaStream_t st,
                       unsigned long long* zero16 = nullptr, const float* aff = nullptr) {
  const bool vec = (dim % 4 == 0) && (((uintptr_t)x & 15) == 0) && dim <= 512;
  if (vec)
    knn_prep_kernel<true><<<(unsigned)((n + 7) / 8), 256, 0, st>>>(x, n, dim, dpad, out, hn,
                                                                   er, sc, zero16, aff);
  else
    knn_prep_kernel<false><<<(unsigned)((n + 7) / 8), 256, 0, st>>>(x, n, dim, dpad, out, hn,
                                                                    er, sc, zero16, aff);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

// Blocks launch_colstats uses for n rows (partials it writes: that many x [3][dim] floats).
static int colstats_blocks(int64_t n) {
  const int64_t b = (n + CS_ROWS_PER_BLOCK - 1) / CS_ROWS_PER_BLOCK;
  return (int)(b < 1 ? 1 : (b > CS_MAX_BLOCKS ? CS_MAX_BLOCKS : b));
}

static int launch_colstats(const float* x, int64_t n, int dim, float* part, cudaStream_t st) {
  const int blocks = colstats_blocks(n);
  const int64_t rpb = (n + blocks - 1) / blocks;
  if ((dim % 4 == 0) && (((uintptr_t)x & 15) == 0)) {
    const int ncol = dim / 4;                   // <= 128 float4 columns, 256 threads
    colstats_kernel<float4><<<blocks, 256, 3 * ncol * sizeof(float4), st>>>(
        reinterpret_cast<const float4*>(x), n, ncol, rpb, reinterpret_cast<float4*>(part));
  } else {
    colstats_kernel<float><<<blocks, 512, 3 * dim * sizeof(float), st>>>(x, n, dim, rpb, part);
  }
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

// Affine record of a call from its row sets (queries == nullptr: one set).
static int launch_affine(const float* bank, int64_t nb, const float* queries, int64_t nq, int dim,
                         float* part_b, float* part_q, float* stats, float* aff,
                         cudaStream_t st) {
  int rc = launch_colstats(bank, nb, dim, part_b, st);
  if (rc) return rc;
  if (queries) {
    rc = launch_colstats(queries, nq, dim, part_q, st);
    if (rc) return rc;
  }
  colstats_reduce_kernel<<<(dim + 127) / 128, 128, 0, st>>>(
      part_b, colstats_blocks(nb), part_q, queries ? colstats_blocks(nq) : 0, dim, stats);
  EXB_LAUNCH_CHECK();
  affine_kernel<<<1, EXB_SCREEN_MAX_DIM, 0, st>>>(stats, (double)nb + (queries ? (double)nq : 0.0),
                                                  dim, aff);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

}  // namespace exb

using namespace exb;

extern "C" int exb_knn_prep_f16(const float* x, int64_t n, int dim, void* out_f16,
                                float* half_norm, float* round_err, float* scalars,
                                exb_stream_t stream) {
  EXB_CHECK_ARG(x && out_f16 && half_norm && round_err && scalars && n > 0 && dim > 0,
                "bad arguments");
  return launch_prep(x, n, dim, dim_pad(dim), static_cast<__half*>(out_f16), half_norm,
                     round_err, reinterpret_cast<PrepScalars*>(scalars), (cudaStream_t)stream);
}

extern "C" size_t exb_knn_colstats_workspace_bytes(int64_t n, int dim) {
  if (n < 1 || dim < 1 || dim > EXB_SCREEN_MAX_DIM) return 0;
  return (size_t)colstats_blocks(n) * 3 * dim * sizeof(float) + 256;
}

extern "C" int exb_knn_colstats_f32(const float* x, int64_t n, int dim, float* stats,
                                    void* workspace, size_t workspace_bytes,
                                    exb_stream_t stream) {
  EXB_CHECK_ARG(x && stats && workspace && n > 0, "bad arguments");
  EXB_CHECK_ARG(dim >= 1 && dim <= EXB_SCREEN_MAX_DIM, "dim %d outside [1, %d]", dim,
                EXB_SCREEN_MAX_DIM);
  EXB_CHECK_ARG(workspace_bytes >= exb_knn_colstats_workspace_bytes(n, dim) &&
                ((uintptr_t)workspace & 15) == 0, "workspace too small or unaligned");
  cudaStream_t st = (cudaStream_t)stream;
  float* part = static_cast<float*>(workspace);
  const int rc = launch_colstats(x, n, dim, part, st);
  if (rc) return rc;
  colstats_reduce_kernel<<<(dim + 127) / 128, 128, 0, st>>>(part, colstats_blocks(n), nullptr, 0,
                                                            dim, stats);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

extern "C" int exb_knn_affine_from_stats(const float* stats, double n_rows, int dim,
                                         float* affine, exb_stream_t stream) {
  EXB_CHECK_ARG(stats && affine && n_rows >= 1.0, "bad arguments");
  EXB_CHECK_ARG(dim >= 1 && dim <= EXB_SCREEN_MAX_DIM, "dim %d outside [1, %d]", dim,
                EXB_SCREEN_MAX_DIM);
  EXB_CHECK_ARG(((uintptr_t)affine & 15) == 0, "affine record must be 16-byte aligned");
  affine_kernel<<<1, EXB_SCREEN_MAX_DIM, 0, (cudaStream_t)stream>>>(stats, n_rows, dim, affine);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

extern "C" int exb_knn_prep_affine_f16(const float* x, int64_t n, int dim, const float* affine,
                                       void* out_f16, float* half_norm, float* round_err,
                                       float* scalars, exb_stream_t stream) {
  EXB_CHECK_ARG(x && affine && out_f16 && half_norm && round_err && scalars && n > 0 && dim > 0,
                "bad arguments");
  EXB_CHECK_ARG(dim <= EXB_SCREEN_MAX_DIM, "dim %d > %d", dim, EXB_SCREEN_MAX_DIM);
  EXB_CHECK_ARG(((uintptr_t)affine & 15) == 0, "affine record must be 16-byte aligned");
  return launch_prep(x, n, dim, dim_pad(dim), static_cast<__half*>(out_f16), half_norm,
                     round_err, reinterpret_cast<PrepScalars*>(scalars), (cudaStream_t)stream,
                     nullptr, affine);
}

extern "C" size_t exb_knn_screened_workspace_bytes(int64_t nq, int64_t nb, int dim, int k) {
  ScreenPlan pl;
  if (!make_plan(nq, nb, dim, k, &pl)) return 0;
  ScreenWs w;
  // worst case: src and bank distinct
  return carve_screen(nullptr, nq, nb, pl, k, false, &w);
}

// pbe: optional fused epilogue request; *fused tells the caller whether it was honoured (only the
// per-update latency chain fuses it).
__global__ void fold_scalars_kernel(const PrepScalars* bsc, const PrepScalars* qsc,
                                    PrepScalars* out) {
  PrepScalars o = *bsc;
  if (qsc) o.range_flag |= qsc->range_flag;   // a query row outside the fp16 range
  o.pad = 0;
  *out = o;
}

// Screen-only mode (exb_knn_screen_lists_f32): the candidate lists, the query rows' half norms /
// rounding errors and the bank scalars go to caller buffers; no rerank, no fallback.
// ------------------------------------------------------------------------------------------
// Instantiated CUDA graphs of the per-update chain, cached inside the library.
//
// A call is identified by everything its kernels were launched with (pointers, shapes, modes):
// nothing else changes from one update to the next now that the RMS count lives on the device.
// Miss: the chain is enqueued once on the library's capture stream (stream capture records, it
// does not execute), instantiated, cached (8 entries, least recently used out) and launched on the
// caller's stream.  Hit: one cudaGraphLaunch.  Callers whose stream is itself being captured, or
// whose pointers never repeat (32 misses in a row), get the plain launch chain.
// ------------------------------------------------------------------------------------------
struct ChainKey {
  const void *src, *bank, *out_dist, *out_idx, *ws, *pre_q, *pre_b;
  const void *rms_m, *rms_s, *rms_n, *out_reward;
  int64_t nq, nb;
  int dim, k, idx_base, same, device;
  int pbe_on, avg, use_rms, use_clip, raw_kth, coop;
  float clip;
};

class ChainGraphs {
 public:
  // true: the call was served (or failed) here and *rc holds its status; false: not applicable,
  // the caller launches the chain itself.
  template <typename Run>
  bool launch(ChainKey& key, cudaStream_t st, Run&& run, int* rc) {
    std::lock_guard<std::mutex> lock(mu_);
    if (bypass_ > 0) { --bypass_; return false; }
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) {
      cudaGetLastError();
      return false;                 // inside a caller's capture: its graph gets the kernels
    }
    if (cudaGetDevice(&key.device) != cudaSuccess) return false;
    ++tick_;
    for (Entry& e : entries_) {
      if (e.exec != nullptr && memcmp(&e.key, &key, sizeof(key)) == 0) {
        e.used = tick_;
        misses_ = 0;
        *rc = EXB_OK;
        if (cudaGraphLaunch(e.exec, st) != cudaSuccess) {
          set_error("cudaGraphLaunch of the cached per-update chain failed: %s",
                    cudaGetErrorString(cudaGetLastError()));
          *rc = EXB_ECUDA;
        }
        return true;
      }
    }
    if (++misses_ > 32) {           // pointers that never repeat: graphs would only cost
      misses_ = 0;
      bypass_ = 4096;
      return false;
    }
    cudaStream_t cs = capture_stream(key.device);
    if (cs == nullptr) return false;
    if (cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
      cudaGetLastError();
      return false;
    }
    const int run_rc = run(cs);
    cudaGraph_t graph = nullptr;
    const cudaError_t end = cudaStreamEndCapture(cs, &graph);
    if (run_rc != EXB_OK || end != cudaSuccess || graph == nullptr) {
      cudaGetLastError();
      if (graph) cudaGraphDestroy(graph);
      if (run_rc != EXB_OK) { *rc = run_rc; return true; }   // bad arguments: report them
      return false;
    }
    cudaGraphExec_t exec = nullptr;
    if (cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) {
      cudaGetLastError();
      cudaGraphDestroy(graph);
      return false;
    }
    cudaGraphDestroy(graph);
    Entry* slot = &entries_[0];
    for (Entry& e : entries_)
      if (e.exec == nullptr || e.used < slot->used) { slot = &e; if (e.exec == nullptr) break; }
    if (slot->exec) cudaGraphExecDestroy(slot->exec);
    slot->key = key;
    slot->exec = exec;
    slot->used = tick_;
    *rc = EXB_OK;
    if (cudaGraphLaunch(exec, st) != cudaSuccess) {
      set_error("cudaGraphLaunch of the per-update chain failed: %s",
                cudaGetErrorString(cudaGetLastError()));
      *rc = EXB_ECUDA;
    }
    return true;
  }

 private:
  struct Entry {
    ChainKey key;
    cudaGraphExec_t exec = nullptr;
    uint64_t used = 0;
  };
  cudaStream_t capture_stream(int dev) {
    if (dev < 0 || dev >= 64) return nullptr;
    if (streams_[dev] == nullptr &&
        cudaStreamCreateWithFlags(&streams_[dev], cudaStreamNonBlocking) != cudaSuccess) {
      cudaGetLastError();
      streams_[dev] = nullptr;
    }
    return streams_[dev];
  }
  std::mutex mu_;
  Entry entries_[8];
  cudaStream_t streams_[64] = {};
  uint64_t tick_ = 0;
  int misses_ = 0, bypass_ = 0;
};
static ChainGraphs g_chain_graphs;

struct ListsOut {
  float* val;      // [nq, lists, SCR_CAP]
  int* idx;        // [nq, lists, SCR_CAP], indices local to `bank`
  int* cnt;        // [nq, lists]
  float* ha;       // [nq]
  float* ea;       // [nq]
  void* scalars;   // 16 bytes: PrepScalars of the bank (query-side range flag folded in)
};

// Grid-barrier words of the fused chain kernel: a per-device pool of 64 {arrivals, generation}
// pairs in device memory, zeroed once (a barrier leaves its pair clean), handed out round-robin so
// that chains running concurrently on different streams do not share a pair.  nullptr: the
// device cannot launch cooperatively / allocation failed / switched off -> the launch chain.
static unsigned* chain_gbar_slot() {
  // Off by default — measured on B200 (profiles/r02f_chain_trace.txt): in-kernel 31.0 us in a hot
  // loop (prep 0.7, barrier 2.5, screen 6.5, barrier 3.5, rerank 10.7, barrier 2.9, finish 4.2)
  // and 47.5 us after an L2 flush, against 30.7 us for the graph-replayed launch chain under the
  // same flush: a grid barrier costs about what a kernel boundary inside a graph does, and the
  // phases run slower than the standalone kernels (one 254-register block per SM).
  static const int fused_on = knob("EXB_CHAIN_FUSED", 0);
  if (!fused_on) return nullptr;
  static std::mutex mu;
  static unsigned* pool[64] = {};
  static int state[64] = {};          // 0 = untried, 1 = ready, -1 = unavailable
  static unsigned next[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  if (state[dev] == 0) {
    int coop = 0;
    state[dev] = -1;
    if (cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev) == cudaSuccess && coop &&
        cudaMalloc(&pool[dev], 64 * 2 * sizeof(unsigned)) == cudaSuccess) {
      if (cudaMemset(pool[dev], 0, 64 * 2 * sizeof(unsigned)) == cudaSuccess) state[dev] = 1;
    }
    cudaGetLastError();
  }
  if (state[dev] != 1) return nullptr;
  return pool[dev] + 2 * (next[dev]++ % 64);
}

template <int KS, int NI>
static int launch_chain_one(const CUtensorMap& tq, const CUtensorMap& tb, const ChainParams& cp,
                            int grid, size_t smem, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    EXB_CUDA(cudaFuncSetAttribute(apt_chain_kernel<KS, NI>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, SCR_SMEM_LIMIT));
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(CH_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  EXB_CUDA(cudaLaunchKernelEx(&cfg, apt_chain_kernel<KS, NI>, tq, tb, cp));
  return EXB_OK;
}

// Instantiated for the two per-update shapes of the benchmark configs only (C1: k = 12, rep 512;
// C2: k = 3, dim 128); other shapes report "not available" and take the launch chain.
static int launch_chain_kernel(const CUtensorMap& tq, const CUtensorMap& tb, const ChainParams& cp,
                               int grid, size_t smem, int k, int ni, cudaStream_t st) {
  if (k > 4 && k <= 12 && ni == 4) return launch_chain_one<12, 4>(tq, tb, cp, grid, smem, st);
  if (k <= 4 && ni == 1) return launch_chain_one<4, 1>(tq, tb, cp, grid, smem, st);
  return EXB_EUNSUPPORTED;
}

static int screened_impl(const float* src, int64_t nq, const float* bank, int64_t nb, int dim,
                         int k, int32_t idx_base, int src_is_bank, float* out_dist,
                         int32_t* out_idx, void* workspace, size_t workspace_bytes,
                         exb_stream_t stream, const FusedPbe* pbe, bool* fused,
                         const ListsOut* lo = nullptr, const exb_operand* pre_q = nullptr,
                         const exb_operand* pre_b = nullptr) {
  if (fused) *fused = false;
  EXB_CHECK_ARG(src && bank && (lo || (out_dist && out_idx)), "null pointer");
  EXB_CHECK_ARG(nq > 0 && nb > 0 && nq < (1LL << 31) && nb < (1LL << 31), "bad row counts");
  if (k > nb) {
    set_error("selected index k out of range (k=%d > bank rows=%lld)", k, (long long)nb);
    return EXB_EKRANGE;
  }
  ScreenPlan pl;
  if (!make_plan(nq, nb, dim, k, &pl)) {
    set_error("screened kNN supports 1 <= dim <= %d and 1 <= k <= %d (got dim=%d k=%d)",
              EXB_SCREEN_MAX_DIM, SCR_MAX_K, dim, k);
    return EXB_EUNSUPPORTED;
  }
  const bool same = src_is_bank && src == bank && nq == nb;
  ScreenWs w;
  const size_t need = carve_screen(nullptr, nq, nb, pl, k, same, &w);
  EXB_CHECK_ARG(workspace && workspace_bytes >= need, "workspace too small: %zu < %zu",
                workspace_bytes, need);
  EXB_CHECK_ARG(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  carve_screen(workspace, nq, nb, pl, k, same, &w);
  cudaStream_t st = (cudaStream_t)stream;
  // caller-prepared operands (rep producers): use them in place of the workspace copies and skip
  // the prep launches.  The per-update chain derives the bank maxima itself (lat kernel), so its
  // scalars stay in the workspace; the streaming path reads the producer's scalars.
  const bool pre = pre_b != nullptr;
  if (pre) {
    w.b16 = static_cast<__half*>(pre_b->f16); w.hb = pre_b->half_norm; w.eb = pre_b->round_err;
    const exb_operand* q = same ? pre_b : pre_q;
    w.q16 = static_cast<__half*>(q->f16); w.ha = q->half_norm; w.ea = q->round_err;
    if (!pl.lat) {
      w.bsc = static_cast<PrepScalars*>(pre_b->scalars);
      w.qsc = static_cast<PrepScalars*>(q->scalars);
    }
  }

  static const int trace_on = knob("EXB_SCREEN_TRACE", 0);
  const bool vec = (dim % 4 == 0) && (((uintptr_t)src | (uintptr_t)bank) & 15) == 0;
  const int ni = vec ? (dim + 127) / 128 : 0;   // float4 loads per lane and row (0 = scalar path)
  RerankParams rp;
  rp.Q = src; rp.B = bank; rp.nq = nq; rp.nb = nb; rp.dim = dim; rp.k = k; rp.idx_base = idx_base;
  rp.splits = pl.splits * pl.trackers; rp.cand_val = w.cand_val; rp.cand_idx = w.cand_idx;
  rp.cand_cnt = w.cand_cnt; rp.cand_tk = w.cand_tk; rp.cand_tr = w.cand_tr; rp.ha = w.ha; rp.ea = w.ea; rp.qsc = same ? w.bsc : w.qsc;
  rp.bsc = w.bsc; rp.bank_max = reinterpret_cast<const float*>(w.bsc); rp.out_dist = out_dist;
  rp.out_idx = out_idx; rp.fallback_rows = w.fb_rows; rp.fallback_cap = pl.fb_cap;
  rp.status = w.status; rp.trace = trace_on ? w.trace : nullptr;
  rp.pbe.on = 0;
  rp.split_cols = pl.lat ? pl.tiles_per_split * LAT_N : 0;
  rp.coop_flags = pl.lat ? w.coop_flags : nullptr;
  rp.lists_per_shard = rp.splits;
  rp.shard_rows = 0;
  for (int i = 0; i < 16; ++i) rp.shard_lo[i] = 0;
  const unsigned rr_grid = (unsigned)((nq + RR_WARPS - 1) / RR_WARPS);

  if (pl.lat) {
    if (lo) {
      set_error("screen-only lists need the streaming kernel (input too small: %lld x %lld)",
                (long long)nq, (long long)nb);
      return EXB_EUNSUPPORTED;
    }
    // per-update chain: prep -> latency screen -> rerank (inline fallback) [-> finish]; no memset,
    // no atomics.  `run_chain` enqueues it on a stream; normally that stream is the library's
    // capture stream, once per distinct call, and the call itself is ONE cudaGraphLaunch of the
    // instantiated chain (ChainGraphs below): kernels of a graph follow each other with less
    // dead time than kernels of a stream (30.7 vs 37.9 us for PBE 1024 x 512).
    if (pbe) {
      rp.pbe = *pbe;
      rp.pbe.on = 1;
      rp.pbe.partials = w.pbe_partials;
      if (fused) *fused = true;
    }
    static bool lat_attr_set = false;
    if (!lat_attr_set) {
      EXB_CUDA(cudaFuncSetAttribute(knn_screen_lat_kernel<4>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, SCR_SMEM_LIMIT));
      EXB_CUDA(cudaFuncSetAttribute(knn_screen_lat_kernel<12>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, SCR_SMEM_LIMIT));
      EXB_CUDA(cudaFuncSetAttribute(knn_screen_lat_kernel<16>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, SCR_SMEM_LIMIT));
      lat_attr_set = true;
    }
    // ONE cooperative kernel (knn_chain.cuh) when the device can hold its grid; else (or with
    // EXB_CHAIN_FUSED=0 in a knob build) the launch chain below
    const bool coop = g_coop_scan.load(std::memory_order_relaxed) != 0;
    unsigned* gbar = trace_on ? nullptr : chain_gbar_slot();
    auto run_chain = [&](cudaStream_t st) -> int {
    int rc = EXB_OK;
    if (gbar != nullptr) {
      CUtensorMap tq, tb;
      rc = make_tmap(&tq, w.q16, nq, pl.dpad, SCR_M);
      if (rc) return rc;
      rc = make_tmap(&tb, w.b16, nb, pl.dpad, LAT_N);
      if (rc) return rc;
      ChainParams cp;
      cp.bank = bank; cp.src = src; cp.nb = nb; cp.nq = nq; cp.dim = dim; cp.dpad = pl.dpad;
      cp.same = same ? 1 : 0; cp.do_prep = pre ? 0 : 1;
      cp.vec = ((dim % 4 == 0) && (((uintptr_t)src | (uintptr_t)bank) & 15) == 0 && dim <= 512) ? 1 : 0;
      cp.b16 = w.b16; cp.hb = w.hb; cp.eb = w.eb; cp.q16 = w.q16; cp.ha = w.ha; cp.ea = w.ea;
      LatParams& lp = cp.lp;
      lp.ha = w.ha; lp.ea = w.ea; lp.hb = w.hb; lp.eb = w.eb; lp.nq = nq; lp.nb = nb;
      lp.nkc = pl.nkc; lp.n_tiles = pl.n_tiles; lp.tiles_per_split = pl.tiles_per_split;
      lp.splits = pl.splits; lp.k = k; lp.cand_val = w.cand_val; lp.cand_idx = w.cand_idx;
      lp.cand_cnt = w.cand_cnt; lp.cand_tk = w.cand_tk; lp.cand_tr = w.cand_tr;
      lp.rsel = (k + pl.splits - 1) / pl.splits; lp.bank_max = reinterpret_cast<float*>(w.bsc);
      lp.trace = nullptr; lp.zero_status = nullptr; lp.coop_flags = nullptr;
      cp.rp = rp;
      cp.rp.coop_flags = nullptr;                  // the fused kernel always runs the barriers
      cp.m_tiles = pl.m_tiles; cp.splits = pl.splits; cp.gbar = gbar;
      static const int chain_trace = knob("EXB_CHAIN_TRACE", 0);
      cp.trace = chain_trace ? w.trace : nullptr;
      const int64_t row_blocks = (std::max(nq, nb) + CH_WARPS - 1) / CH_WARPS;
      const int grid = std::max<int>(pl.m_tiles * pl.splits,
                                     (int)std::min<int64_t>(row_blocks, sm_count()));
      const size_t smem = std::max<size_t>(pl.smem, CH_P2_SMEM + 1024);
      rc = launch_chain_kernel(tq, tb, cp, grid, smem, k, ni, st);
      if (rc != EXB_EUNSUPPORTED) return rc;
      rc = EXB_OK;                                   // shape without a fused instance
    }
    if (!pre) {
      rc = launch_prep(bank, nb, dim, pl.dpad, w.b16, w.hb, w.eb, nullptr, st,
                       reinterpret_cast<unsigned long long*>(w.status));
      if (rc) return rc;
      if (!same) {
        rc = launch_prep(src, nq, dim, pl.dpad, w.q16, w.ha, w.ea, nullptr, st);
        if (rc) return rc;
      }
    }
    CUtensorMap tq, tb;
    rc = make_tmap(&tq, w.q16, nq, pl.dpad, SCR_M);
    if (rc) return rc;
    rc = make_tmap(&tb, w.b16, nb, pl.dpad, LAT_N);
    if (rc) return rc;
    LatParams lp;
    lp.ha = w.ha; lp.ea = w.ea; lp.hb = w.hb; lp.eb = w.eb; lp.nq = nq; lp.nb = nb;
    lp.nkc = pl.nkc; lp.n_tiles = pl.n_tiles; lp.tiles_per_split = pl.tiles_per_split;
    lp.splits = pl.splits; lp.k = k; lp.cand_val = w.cand_val; lp.cand_idx = w.cand_idx;
    lp.cand_cnt = w.cand_cnt; lp.cand_tk = w.cand_tk; lp.cand_tr = w.cand_tr;
    lp.rsel = (k + pl.splits - 1) / pl.splits; lp.bank_max = reinterpret_cast<float*>(w.bsc);
    lp.trace = trace_on ? w.trace : nullptr;
    lp.zero_status = pre ? reinterpret_cast<unsigned long long*>(w.status) : nullptr;
    lp.coop_flags = coop ? w.coop_flags : nullptr;
    dim3 grid((unsigned)pl.m_tiles, (unsigned)pl.splits);
    const bool prof = g_prof.on && g_prof.n < 64;
    if (prof) EXB_CUDA(cudaEventRecord(g_prof.e0[g_prof.n], st));
    if (k <= 4) EXB_CUDA(launch_pdl(knn_screen_lat_kernel<4>, grid, dim3(LAT_THREADS), pl.smem, st, tq, tb, lp));
    else if (k <= 12) EXB_CUDA(launch_pdl(knn_screen_lat_kernel<12>, grid, dim3(LAT_THREADS), pl.smem, st, tq, tb, lp));
    else EXB_CUDA(launch_pdl(knn_screen_lat_kernel<16>, grid, dim3(LAT_THREADS), pl.smem, st, tq, tb, lp));
    EXB_LAUNCH_CHECK();
    if (prof) EXB_CUDA(cudaEventRecord(g_prof.e1[g_prof.n++], st));
    if (coop) launch_rerank<true, false, true>(rp, rr_grid, ni, st);
    else launch_rerank<true>(rp, rr_grid, ni, st);
    EXB_LAUNCH_CHECK();
    if (pbe && pbe->use_rms)
      return launch_pbe_finish(out_dist, nq, k, pbe->avg, 1, pbe->use_clip, pbe->clip,
                               pbe->rms.m, pbe->out_reward, st);
    return EXB_OK;
    };   // run_chain
    if (!trace_on && !g_prof.on) {
      ChainKey key;
      memset(&key, 0, sizeof(key));
      key.src = src; key.bank = bank; key.nq = nq; key.nb = nb; key.dim = dim; key.k = k;
      key.idx_base = idx_base; key.same = same; key.out_dist = out_dist; key.out_idx = out_idx;
      key.ws = workspace; key.coop = coop ? 1 : 0;
      if (pre) { key.pre_b = pre_b->f16; key.pre_q = (same ? pre_b : pre_q)->f16; }
      if (pbe) {
        key.pbe_on = 1; key.avg = pbe->avg; key.use_rms = pbe->use_rms;
        key.use_clip = pbe->use_clip; key.raw_kth = pbe->raw_kth; key.clip = pbe->clip;
        key.rms_m = pbe->rms.m; key.rms_s = pbe->rms.s; key.rms_n = pbe->rms.n;
        key.out_reward = pbe->out_reward;
      }
      int rc = EXB_OK;
      if (g_chain_graphs.launch(key, st, run_chain, &rc)) return rc;
    }
    return run_chain(st);
  }

  // header (status + scalars) is the first 3 x 256 bytes (+ the 16 debug counters when tracing)
  EXB_CUDA(cudaMemsetAsync(workspace, 0, 1024 + 64, st));
  int rc = EXB_OK;
  if (!pre) {
    // range-proof screen: the operands are s (x - mu) with the call's own column means and a
    // power-of-two scale (one extra streaming pass over the rows; any (mu, s) is exact, this
    // one keeps shifted / large / tiny reps inside the fp16 format).  Not for the screen-only
    // form: the lists of several shards are compared by score and must share one transform —
    // exb_knn_colstats_f32 / exb_knn_affine_from_stats / exb_knn_prep_affine_f16 give a caller
    // that shards the bank the pieces to agree on one.
    const float* aff = nullptr;
    if (!lo) {
      rc = launch_affine(bank, nb, same ? nullptr : src, nq, dim, w.cs_part_b, w.cs_part_q,
                         w.cs_stats, w.aff, st);
      if (rc) return rc;
      aff = w.aff;
    }
    rc = launch_prep(bank, nb, dim, pl.dpad, w.b16, w.hb, w.eb, w.bsc, st, nullptr, aff);
    if (rc) return rc;
    if (!same) {
      rc = launch_prep(src, nq, dim, pl.dpad, w.q16, w.ha, w.ea, w.qsc, st, nullptr, aff);
      if (rc) return rc;
    }
  }

  CUtensorMap tq, tb;
  rc = make_tmap(&tq, w.q16, nq, pl.dpad, SCR_M);
  if (rc) return rc;
  rc = make_tmap(&tb, w.b16, nb, pl.dpad, SCR_N / pl.cg);
  if (rc) return rc;

  static bool attr_set = false;
  if (!attr_set) {
    EXB_CUDA(cudaFuncSetAttribute(knn_screen_kernel<1, 1>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, SCR_SMEM_LIMIT));
    EXB_CUDA(cudaFuncSetAttribute(knn_screen_kernel<2, 1>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, SCR_SMEM_LIMIT));
    EXB_CUDA(cudaFuncSetAttribute(knn_screen_kernel<1, 2>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, SCR_SMEM_LIMIT));
    EXB_CUDA(cudaFuncSetAttribute(knn_screen_kernel<2, 2>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, SCR_SMEM_LIMIT));
    EXB_CUDA(cudaFuncSetAttribute(knn_screen_kernel<1, 3>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, SCR_SMEM_LIMIT));
    EXB_CUDA(cudaFuncSetAttribute(knn_screen_kernel<2, 3>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, SCR_SMEM_LIMIT));
    attr_set = true;
  }
  ScreenParams sp;
  sp.ha = w.ha; sp.ea = w.ea; sp.hb = w.hb; sp.bsc = w.bsc;
  sp.nq = nq; sp.nb = nb; sp.nkc = pl.nkc; sp.n_tiles = pl.n_tiles;
  sp.tiles_per_split = pl.tiles_per_split; sp.splits = pl.splits; sp.k = k;
  static const int dbg = knob("EXB_SCREEN_DEBUG", 0);
  sp.debug = dbg;
  static const int pf = knob("EXB_SCREEN_PREFETCH", 0);   // measured: hurts (L2->SM bandwidth), so off
  sp.prefetch = pf;
  sp.stats = trace_on ? w.trace : nullptr;
  sp.stages = pl.stages; sp.cand_val = w.cand_val; sp.cand_idx = w.cand_idx; sp.cand_cnt = w.cand_cnt;
  if (lo) {
    sp.cand_val = lo->val;
    sp.cand_idx = lo->idx;
    sp.cand_cnt = lo->cnt;
  }
  dim3 grid((unsigned)pl.m_tiles, (unsigned)pl.splits);
  sp.tail_splits = pl.tail_splits; sp.full_units = pl.full_units; sp.tail_tps = pl.tail_tps;
  static const int chain_reset_env = knob("EXB_SCREEN_CHAINRESET", 1);   // A/B: 0 = keep the seed chain (old scheme)
  sp.chain_reset = chain_reset_env;
  if (pl.tail_splits > 0) {
    const int units = pl.m_tiles / pl.cg;
    grid = dim3((unsigned)((pl.full_units + (units - pl.full_units) * pl.tail_splits) * pl.cg), 1);
    sp.tiles_per_split = pl.n_tiles;
    // rows scanned by a full unit never touch lists 1..: their counts must read as empty
    EXB_CUDA(cudaMemsetAsync(sp.cand_cnt, 0, (size_t)nq * pl.splits * sizeof(int), st));
  }
  sp.lists = pl.splits * pl.trackers;
  // sample tiles that seed the k-th-best chain before the real scan (FILT 3): ~1.5 % extra MMA
  // work buys roughly half the list insertions
  static const int seed_env = knob("EXB_SCREEN_SEED", 0);
  int seed_tiles = seed_env != 0 ? seed_env : pl.tiles_per_split / 16;
  if (seed_env == 0) seed_tiles = seed_tiles < 2 ? 2 : (seed_tiles > 32 ? 32 : seed_tiles);
  if (seed_tiles < 0) seed_tiles = 0;      // EXB_SCREEN_SEED=-1: no sample tiles
  sp.seed_tiles = pl.filt >= 2 ? seed_tiles : 0;
  static const int scan_cap_env = [] {
    const int v = knob("EXB_SCREEN_SCANCAP", SCR_SCAN_CAP);   // A/B: repack period of the scratch lists
    return v < 16 ? 16 : (v > SCR_SCAN_CAP ? SCR_SCAN_CAP : v);
  }();
  sp.scan_cap = scan_cap_env;
  sp.scan_val = w.scan_val;
  sp.scan_idx = w.scan_idx;
  static const int share_head = knob("EXB_SCREEN_HEAD", 1);   // A/B: 0 = every CTA scans from its first tile
  sp.head = share_head ? w.head : nullptr;
  const int filt = pl.filt;
  const unsigned threads = filt == 3 ? SCR_THREADS_V3 : SCR_THREADS;
  const bool prof = g_prof.on && g_prof.n < 64;
  if (prof) EXB_CUDA(cudaEventRecord(g_prof.e0[g_prof.n], st));
  if (pl.cg == 2) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = pl.smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (filt == 1) EXB_CUDA(cudaLaunchKernelEx(&cfg, knn_screen_kernel<2, 1>, tq, tb, sp));
    else if (filt == 3) EXB_CUDA(cudaLaunchKernelEx(&cfg, knn_screen_kernel<2, 3>, tq, tb, sp));
    else EXB_CUDA(cudaLaunchKernelEx(&cfg, knn_screen_kernel<2, 2>, tq, tb, sp));
  } else {
    if (filt == 1) knn_screen_kernel<1, 1><<<grid, SCR_THREADS, pl.smem, st>>>(tq, tb, sp);
    else if (filt == 3) knn_screen_kernel<1, 3><<<grid, SCR_THREADS_V3, pl.smem, st>>>(tq, tb, sp);
    else knn_screen_kernel<1, 2><<<grid, SCR_THREADS, pl.smem, st>>>(tq, tb, sp);
  }
  EXB_LAUNCH_CHECK();
  if (prof) EXB_CUDA(cudaEventRecord(g_prof.e1[g_prof.n++], st));

  if (lo) {
    EXB_CUDA(cudaMemcpyAsync(lo->ha, w.ha, (size_t)nq * 4, cudaMemcpyDeviceToDevice, st));
    EXB_CUDA(cudaMemcpyAsync(lo->ea, w.ea, (size_t)nq * 4, cudaMemcpyDeviceToDevice, st));
    fold_scalars_kernel<<<1, 1, 0, st>>>(w.bsc, same ? nullptr : w.qsc,
                                         static_cast<PrepScalars*>(lo->scalars));
    EXB_LAUNCH_CHECK();
    return EXB_OK;
  }

  launch_rerank<false>(rp, rr_grid, ni, st);
  EXB_LAUNCH_CHECK();

  // undecided rows -> exact kernel, indirect through the device-side row list / count
  return launch_exact(src, nq, bank, nb, dim, k, idx_base, out_dist, out_idx, w.fb_d2, w.fb_idx,
                      pl.fb_splits, w.fb_rows, &w.status->n_fallback, pl.fb_cap, st);
}

// ------------------------------------------------------------------------------------------
// Symmetric screen (knn_sym.cuh): source == bank at dataset scale.
// ------------------------------------------------------------------------------------------
struct SymPlan {
  int dpad, nkc, m_tiles, n_tiles, stages;
  size_t smem;
  int fb_cap, fb_splits;
};

static bool sym_plan(int64_t n, int dim, int k, SymPlan* pl) {
  if (dim < 1 || dim > EXB_SCREEN_MAX_DIM || k < 1 || k > SCR_MAX_K || n < 4 * SCR_N ||
      n >= (1LL << 31) || k > n)
    return false;
  pl->dpad = dim_pad(dim);
  pl->nkc = pl->dpad / SCR_KC;
  pl->n_tiles = (int)((n + SCR_N - 1) / SCR_N);
  pl->m_tiles = 2 * pl->n_tiles;               // CTA pairs: query rows of pair u = bank tile u
  pl->fb_cap = (int)(n < FALLBACK_CAP ? n : FALLBACK_CAP);
  int fbs = (int)((n + 127) / 128);
  if (fbs > FALLBACK_SPLITS) fbs = FALLBACK_SPLITS;
  if (fbs * k > 512) fbs = 512 / k;
  pl->fb_splits = fbs < 1 ? 1 : fbs;
  const size_t fixed = (size_t)pl->nkc * SCR_A_CHUNK_BYTES + 16 * (SCR_N / 2) * 4 +
                       8 * (SYM_QCAP * 12 + 128) + (2 * SCR_MAX_STAGES + 5) * 8 + 16 + 1024;
  const int b_stage = SCR_B_STAGE_BYTES / 2;
  int stages = (int)((SCR_SMEM_LIMIT - fixed) / b_stage);
  if (stages > SCR_MAX_STAGES) stages = SCR_MAX_STAGES;
  if (stages < 2) return false;
  pl->stages = stages;
  pl->smem = fixed + (size_t)stages * b_stage;
  return true;
}

struct SymWs {
  RerankStatus* status;
  PrepScalars* bsc;
  __half* b16;
  float *hb, *eb, *aff, *cs_stats, *cs_part, *tkeep;
  float *dlog_val, *tq_val, *rval, *fb_d2;
  int *dlog_idx, *tq_target, *tq_src, *ridx, *dlog_cnt, *rcnt, *fb_rows, *fb_idx;
  unsigned long long* tq_count;
  long long tq_cap;
  SymPeerTable* peers;
};

static size_t carve_sym(void* ws, int64_t n, const SymPlan& pl, int k, int log_cap, int rcap,
                        SymWs* o) {
  Carver cv(ws ? ws : reinterpret_cast<void*>(256));
  o->status = cv.take<RerankStatus>(1);
  cv.take<PrepScalars>(1);                     // (slot of the query scalars in carve_screen)
  o->bsc = cv.take<PrepScalars>(1);
  cv.take<long long>(16);
  o->b16 = cv.take<__half>((size_t)n * pl.dpad);
  o->hb = cv.take<float>(n);
  o->eb = cv.take<float>(n);
  o->aff = cv.take<float>(AFF_FLOATS);
  o->cs_stats = cv.take<float>(3 * (size_t)pl.dpad);
  o->cs_part = cv.take<float>((size_t)CS_MAX_BLOCKS * 3 * pl.dpad);
  o->tkeep = cv.take<float>(n);
  o->dlog_val = cv.take<float>((size_t)n * 2 * log_cap);
  o->dlog_idx = cv.take<int>((size_t)n * 2 * log_cap);
  // transposed hits: every row receives rcap / 3 on average by the caller's sizing (rcap = 3x
  // the expectation of a row that receives ALL its candidates); half of all candidates travel
  o->tq_cap = (long long)n * (rcap / 3 + 16);
  o->tq_target = cv.take<int>((size_t)o->tq_cap);
  o->tq_val = cv.take<float>((size_t)o->tq_cap);
  o->tq_src = cv.take<int>((size_t)o->tq_cap);
  o->tq_count = cv.take<unsigned long long>(2);
  o->dlog_cnt = cv.take<int>((size_t)n * 2);
  o->rval = cv.take<float>((size_t)n * rcap);
  o->ridx = cv.take<int>((size_t)n * rcap);
  o->rcnt = cv.take<int>(n);
  o->fb_rows = cv.take<int>(pl.fb_cap);
  o->fb_d2 = cv.take<float>((size_t)pl.fb_cap * pl.fb_splits * k);
  o->fb_idx = cv.take<int>((size_t)pl.fb_cap * pl.fb_splits * k);
  o->peers = cv.take<SymPeerTable>(1);
  return cv.off + 256;
}

extern "C" size_t exb_knn_sym_workspace_bytes(int64_t n, int dim, int k, int log_cap,
                                              int recv_cap) {
  SymPlan pl;
  if (!sym_plan(n, dim, k, &pl) || log_cap < 32 || recv_cap < 32) return 0;
  SymWs w;
  return carve_sym(nullptr, n, pl, k, log_cap, recv_cap, &w);
}

struct SymSpan {            // which tile pairs a launch takes (ScreenParams::sym_*)
  int pair_lo, pair_hi, span_end, mid_pair, extra;
  const SymPeerTable* peers;   // host copy of the routing table, or null (one append buffer)
  bool operand_ready;          // the caller filled the operand slots (exb_knn_sym_operand_slots)
};

// Phase 1: operand, thresholds, symmetric screen of the pairs of `sp_`; leaves the direct logs
// and the append buffer of transposed hits in the workspace.
static int sym_screen_phase(const float* rep, int64_t n, int dim, int k, const float* d_sample,
                            int64_t d_stride, int log_cap, int recv_cap, const SymSpan& span,
                            void* workspace, size_t workspace_bytes, cudaStream_t st,
                            SymPlan* pl_out, SymWs* w_out) {
  SymPlan pl;
  if (!sym_plan(n, dim, k, &pl) || log_cap < 32 || recv_cap < 32) {
    set_error("symmetric screen needs n >= %d, dim <= %d, k <= %d (got n=%lld dim=%d k=%d)",
              4 * SCR_N, EXB_SCREEN_MAX_DIM, SCR_MAX_K, (long long)n, dim, k);
    return EXB_EUNSUPPORTED;
  }
  EXB_CHECK_ARG(span.pair_lo >= 0 && span.pair_lo < span.pair_hi && span.pair_hi <= pl.n_tiles &&
                span.span_end >= span.pair_hi && span.extra >= 0 &&
                span.span_end - span.pair_lo + span.extra <= pl.n_tiles,
                "bad tile span");
  SymWs w;
  const size_t need = carve_sym(nullptr, n, pl, k, log_cap, recv_cap, &w);
  EXB_CHECK_ARG(workspace_bytes >= need, "workspace too small: %zu < %zu", workspace_bytes, need);
  EXB_CHECK_ARG(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  carve_sym(workspace, n, pl, k, log_cap, recv_cap, &w);
  int rc = EXB_OK;
  if (span.operand_ready) {
    // operand, norms, scalars and affine record are in their slots already (each rank prepared
    // its shard, the shards were all-gathered): clear the status and trace words only
    EXB_CUDA(cudaMemsetAsync(w.status, 0, 256, st));
    EXB_CUDA(cudaMemsetAsync(static_cast<char*>(workspace) + 768, 0, 256, st));   // b16 at 1024
  } else {
    EXB_CUDA(cudaMemsetAsync(workspace, 0, 1024 + 64, st));
  }
  EXB_CUDA(cudaMemsetAsync(w.rcnt, 0, (size_t)n * sizeof(int), st));
  EXB_CUDA(cudaMemsetAsync(w.tq_count, 0, 16, st));
  if (!span.operand_ready) {
    rc = launch_affine(rep, n, nullptr, 0, dim, w.cs_part, nullptr, w.cs_stats, w.aff, st);
    if (rc) return rc;
    rc = launch_prep(rep, n, dim, pl.dpad, w.b16, w.hb, w.eb, w.bsc, st, nullptr, w.aff);
    if (rc) return rc;
  }
  SymThresholdParams tp;
  tp.d_sample = d_sample; tp.stride = d_stride; tp.ha = w.hb; tp.ea = w.eb; tp.bsc = w.bsc;
  tp.aff = w.aff; tp.tkeep = w.tkeep; tp.n = n;
  sym_threshold_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(tp);
  EXB_LAUNCH_CHECK();

  CUtensorMap tq, tb;
  rc = make_tmap(&tq, w.b16, n, pl.dpad, SCR_M);
  if (rc) return rc;
  rc = make_tmap(&tb, w.b16, n, pl.dpad, SCR_N / 2);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    EXB_CUDA(cudaFuncSetAttribute(knn_screen_kernel<2, 4>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, SCR_SMEM_LIMIT));
    attr_set = true;
  }
  ScreenParams sp;
  memset(&sp, 0, sizeof(sp));
  sp.ha = w.hb; sp.ea = w.eb; sp.hb = w.hb; sp.bsc = w.bsc;
  sp.nq = n; sp.nb = n; sp.nkc = pl.nkc; sp.n_tiles = pl.n_tiles;
  sp.tiles_per_split = pl.n_tiles; sp.splits = 1; sp.k = k; sp.stages = pl.stages;
  sp.lists = 2; sp.head = nullptr;
  static const int sym_debug = knob("EXB_SYM_DEBUG", 0);   // 3 = measure without the hit path
  sp.debug = sym_debug;
  sp.sym_pair_lo = span.pair_lo; sp.sym_span_end = span.span_end;
  sp.sym_mid_pair = span.mid_pair; sp.sym_extra = span.extra;
  sp.sym_tkeep = w.tkeep; sp.dlog_val = w.dlog_val; sp.dlog_idx = w.dlog_idx;
  sp.dlog_cnt = w.dlog_cnt; sp.log_cap = log_cap;
  sp.tq_target = w.tq_target; sp.tq_val = w.tq_val; sp.tq_src = w.tq_src;
  sp.tq_count = w.tq_count; sp.tq_cap = w.tq_cap;
  if (span.peers != nullptr) {
    sym_peer_table_kernel<<<1, 32, 0, st>>>(*span.peers, w.peers);
    EXB_LAUNCH_CHECK();
    sp.sym_peers = w.peers;
  }
  {
    const bool prof = g_prof.on && g_prof.n < 64;
    if (prof) EXB_CUDA(cudaEventRecord(g_prof.e0[g_prof.n], st));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * (span.pair_hi - span.pair_lo)), 1);
    cfg.blockDim = dim3(SCR_THREADS_V3);
    cfg.dynamicSmemBytes = pl.smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    EXB_CUDA(cudaLaunchKernelEx(&cfg, knn_screen_kernel<2, 4>, tq, tb, sp));
    if (prof) EXB_CUDA(cudaEventRecord(g_prof.e1[g_prof.n++], st));
  }
  *pl_out = pl;
  *w_out = w;
  return EXB_OK;
}

static SymLists sym_lists(const SymWs& w, int log_cap, int recv_cap) {
  SymLists sl;
  sl.dlog_val = w.dlog_val; sl.dlog_idx = w.dlog_idx; sl.dlog_cnt = w.dlog_cnt;
  sl.log_cap = log_cap; sl.tq_target = w.tq_target; sl.tq_val = w.tq_val; sl.tq_src = w.tq_src;
  sl.tq_count = w.tq_count; sl.tq_cap = w.tq_cap;
  sl.rval = w.rval; sl.ridx = w.ridx; sl.rcnt = w.rcnt; sl.rcap = recv_cap;
  return sl;
}

// Phase 3: rerank rows [row_lo, row_hi) from their direct logs + received lists.
static int sym_rerank_phase(const float* rep, int64_t n, int dim, int k, int32_t idx_base,
                            int64_t row_lo, int64_t row_hi, float* out_dist, int32_t* out_idx,
                            const SymPlan& pl, const SymWs& w, const SymLists& sl,
                            cudaStream_t st) {
  const bool vec = (dim % 4 == 0) && (((uintptr_t)rep) & 15) == 0;
  const int ni = vec ? (dim + 127) / 128 : 0;
  RerankParams rp;
  memset(&rp, 0, sizeof(rp));
  rp.Q = rep; rp.B = rep; rp.nq = row_hi; rp.nb = n; rp.dim = dim; rp.k = k;
  rp.idx_base = idx_base;
  rp.splits = 2; rp.ha = w.hb; rp.ea = w.eb; rp.qsc = w.bsc; rp.bsc = w.bsc;
  // outputs are [row_hi - row_lo, k] arrays of the caller: address them by global row
  rp.out_dist = out_dist - row_lo * k; rp.out_idx = out_idx - row_lo * k;
  rp.fallback_rows = w.fb_rows;
  rp.fallback_cap = pl.fb_cap; rp.status = w.status;
  const int64_t rows = row_hi - row_lo;
  const unsigned rr_grid = (unsigned)((rows + RR_WARPS - 1) / RR_WARPS);
  switch (ni) {
    case 1: knn_rerank_sym_kernel<1><<<rr_grid, RR_WARPS * 32, 0, st>>>(rp, sl, row_lo); break;
    case 2: knn_rerank_sym_kernel<2><<<rr_grid, RR_WARPS * 32, 0, st>>>(rp, sl, row_lo); break;
    case 3: knn_rerank_sym_kernel<3><<<rr_grid, RR_WARPS * 32, 0, st>>>(rp, sl, row_lo); break;
    case 4: knn_rerank_sym_kernel<4><<<rr_grid, RR_WARPS * 32, 0, st>>>(rp, sl, row_lo); break;
    default: knn_rerank_sym_kernel<0><<<rr_grid, RR_WARPS * 32, 0, st>>>(rp, sl, row_lo); break;
  }
  EXB_LAUNCH_CHECK();
  // undecided rows (global row ids in the list) -> exact kernel, written through the same
  // row-offset output views
  return launch_exact(rep, row_hi, rep, n, dim, k, idx_base, rp.out_dist, rp.out_idx, w.fb_d2,
                      w.fb_idx, pl.fb_splits, w.fb_rows, &w.status->n_fallback, pl.fb_cap, st);
}

extern "C" int exb_knn_sym_f32(const float* rep, int64_t n, int dim, int k, int32_t idx_base,
                               const float* d_sample, int64_t d_stride, int log_cap, int recv_cap,
                               float* out_dist, int32_t* out_idx, void* workspace,
                               size_t workspace_bytes, exb_stream_t stream) {
  EXB_CHECK_ARG(rep && d_sample && out_dist && out_idx && workspace, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  SymPlan pl;
  SymWs w;
  const int tiles = (int)((n + SCR_N - 1) / SCR_N);
  const SymSpan span = {0, tiles, tiles, tiles, 0, nullptr, false};
  int rc = sym_screen_phase(rep, n, dim, k, d_sample, d_stride, log_cap, recv_cap, span,
                            workspace, workspace_bytes, st, &pl, &w);
  if (rc) return rc;
  const SymLists sl = sym_lists(w, log_cap, recv_cap);
  sym_scatter_kernel<<<148 * 8, 256, 0, st>>>(sl, sl.tq_target, sl.tq_val, sl.tq_src, -1, w.status);
  EXB_LAUNCH_CHECK();
  static const int sym_stats = knob("EXB_SYM_STATS", 0);
  if (sym_stats) {   // counts into the 16 trace words (exb_debug_trace_sync), zeroed by the memset
    sym_stats_kernel<<<(unsigned)((2 * n + 255) / 256), 256, 0, st>>>(
        sl, n, reinterpret_cast<long long*>(static_cast<char*>(workspace) + 768));
    EXB_LAUNCH_CHECK();
  }
  return sym_rerank_phase(rep, n, dim, k, idx_base, 0, n, out_dist, out_idx, pl, w, sl, st);
}

// ---- the same in three calls, for ranks that each own a block of the tile pairs (relabel.py):
// screen -> [exchange of the transposed hits between the ranks] -> scatter -> rerank.
extern "C" int exb_knn_sym_screen_f32(const float* rep, int64_t n, int dim, int k,
                                      const float* d_sample, int64_t d_stride, int log_cap,
                                      int recv_cap, int pair_lo, int pair_hi, int span_end,
                                      int mid_pair, int extra_tiles, void* workspace,
                                      size_t workspace_bytes, exb_sym_hits* hits,
                                      exb_stream_t stream) {
  EXB_CHECK_ARG(rep && d_sample && workspace && hits, "null pointer");
  SymPlan pl;
  SymWs w;
  const SymSpan span = {pair_lo, pair_hi, span_end, mid_pair, extra_tiles, nullptr, false};
  const int rc = sym_screen_phase(rep, n, dim, k, d_sample, d_stride, log_cap, recv_cap, span,
                                  workspace, workspace_bytes, (cudaStream_t)stream, &pl, &w);
  if (rc) return rc;
  hits->target = w.tq_target; hits->val = w.tq_val; hits->src = w.tq_src;
  hits->count = w.tq_count; hits->capacity = w.tq_cap;
  return EXB_OK;
}

// Peer-memory variant: the kernel appends every hit to the buffer of the rank that owns its query
// row (peer_hits[r], mapped into this process), so no exchange step follows — after a barrier
// across the ranks every rank scatters ITS buffer (exb_knn_sym_scatter_hits) and reranks.
extern "C" int exb_knn_sym_screen_peer_f32(const float* rep, int64_t n, int dim, int k,
                                           const float* d_sample, int64_t d_stride, int log_cap,
                                           int recv_cap, int pair_lo, int pair_hi, int span_end,
                                           int mid_pair, int extra_tiles, int n_ranks,
                                           const exb_sym_hits* peer_hits, const int32_t* tile_lo,
                                           int operand_ready, void* workspace,
                                           size_t workspace_bytes, exb_stream_t stream) {
  EXB_CHECK_ARG(rep && d_sample && workspace && peer_hits && tile_lo, "null pointer");
  EXB_CHECK_ARG(n_ranks >= 1 && n_ranks <= EXB_SYM_MAX_RANKS, "1 <= n_ranks <= %d",
                EXB_SYM_MAX_RANKS);
  SymPeerTable tbl;
  memset(&tbl, 0, sizeof(tbl));
  tbl.ranks = n_ranks;
  for (int r = 0; r <= n_ranks; ++r) {
    EXB_CHECK_ARG(tile_lo[r] >= 0 && (r == 0 || tile_lo[r] >= tile_lo[r - 1]),
                  "tile_lo must be non-decreasing");
    tbl.tile_lo[r] = tile_lo[r];
  }
  EXB_CHECK_ARG(tile_lo[0] == 0 && (int64_t)tile_lo[n_ranks] * SCR_N >= n,
                "tile_lo must cover tiles [0, ceil(n / %d))", SCR_N);
  for (int r = 0; r < n_ranks; ++r) {
    const exb_sym_hits& h = peer_hits[r];
    EXB_CHECK_ARG(h.target && h.val && h.src && h.count && h.capacity > 0,
                  "peer_hits[%d] incomplete", r);
    tbl.target[r] = h.target; tbl.val[r] = h.val; tbl.src[r] = h.src;
    tbl.count[r] = h.count; tbl.cap[r] = h.capacity;
  }
  SymPlan pl;
  SymWs w;
  const SymSpan span = {pair_lo, pair_hi, span_end, mid_pair, extra_tiles, &tbl,
                        operand_ready != 0};
  return sym_screen_phase(rep, n, dim, k, d_sample, d_stride, log_cap, recv_cap, span, workspace,
                          workspace_bytes, (cudaStream_t)stream, &pl, &w);
}

// Where the symmetric screen keeps its operand inside the workspace, for callers that fill it
// themselves (every rank prepares ITS rows with exb_knn_prep_affine_f16 straight into these
// arrays, the shards are all-gathered over NVLink behind the sample pre-pass, the scalars are
// MAX-reduced) and then pass operand_ready = 1 to exb_knn_sym_screen_peer_f32.
extern "C" int exb_knn_sym_operand_slots(int64_t n, int dim, int k, int log_cap, int recv_cap,
                                         void* workspace, size_t workspace_bytes,
                                         exb_operand* slots, float** affine_slot) {
  EXB_CHECK_ARG(workspace && slots && affine_slot, "null pointer");
  SymPlan pl;
  SymWs w;
  EXB_CHECK_ARG(sym_plan(n, dim, k, &pl) && log_cap >= 32 && recv_cap >= 32, "bad sizes");
  const size_t need = carve_sym(nullptr, n, pl, k, log_cap, recv_cap, &w);
  EXB_CHECK_ARG(workspace_bytes >= need, "workspace too small: %zu < %zu", workspace_bytes, need);
  carve_sym(workspace, n, pl, k, log_cap, recv_cap, &w);
  slots->f16 = w.b16;
  slots->half_norm = w.hb;
  slots->round_err = w.eb;
  slots->scalars = w.bsc;
  *affine_slot = w.aff;
  return EXB_OK;
}

// Scatter of a hit buffer whose entry count lives on the device (this rank's peer-visible
// buffer after the barrier): count > capacity sets status flag bit 1 (invalid pass).
extern "C" int exb_knn_sym_scatter_hits(int64_t n, int dim, int k, int log_cap, int recv_cap,
                                        const exb_sym_hits* hits, void* workspace,
                                        exb_stream_t stream) {
  EXB_CHECK_ARG(workspace && hits && hits->target && hits->val && hits->src && hits->count &&
                hits->capacity > 0, "bad arguments");
  SymPlan pl;
  SymWs w;
  EXB_CHECK_ARG(sym_plan(n, dim, k, &pl), "bad sizes");
  carve_sym(workspace, n, pl, k, log_cap, recv_cap, &w);
  SymLists sl = sym_lists(w, log_cap, recv_cap);
  sl.tq_count = hits->count;
  sl.tq_cap = hits->capacity;
  sym_scatter_kernel<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(sl, hits->target, hits->val,
                                                                 hits->src, -1, w.status);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

extern "C" int exb_knn_sym_scatter(int64_t n, int dim, int k, int log_cap, int recv_cap,
                                   const int32_t* target, const float* val, const int32_t* src,
                                   int64_t n_hits, void* workspace, exb_stream_t stream) {
  EXB_CHECK_ARG(workspace && n_hits >= 0 && (n_hits == 0 || (target && val && src)),
                "bad arguments");
  SymPlan pl;
  SymWs w;
  EXB_CHECK_ARG(sym_plan(n, dim, k, &pl), "bad sizes");
  carve_sym(workspace, n, pl, k, log_cap, recv_cap, &w);
  if (n_hits == 0) return EXB_OK;
  const SymLists sl = sym_lists(w, log_cap, recv_cap);
  sym_scatter_kernel<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(sl, target, val, src, n_hits,
                                                                 w.status);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

extern "C" int exb_knn_sym_rerank_f32(const float* rep, int64_t n, int dim, int k,
                                      int32_t idx_base, int log_cap, int recv_cap, int64_t row_lo,
                                      int64_t row_hi, float* out_dist, int32_t* out_idx,
                                      void* workspace, exb_stream_t stream) {
  EXB_CHECK_ARG(rep && out_dist && out_idx && workspace && row_lo >= 0 && row_lo < row_hi &&
                row_hi <= n, "bad arguments");
  SymPlan pl;
  SymWs w;
  EXB_CHECK_ARG(sym_plan(n, dim, k, &pl), "bad sizes");
  carve_sym(workspace, n, pl, k, log_cap, recv_cap, &w);
  const SymLists sl = sym_lists(w, log_cap, recv_cap);
  return sym_rerank_phase(rep, n, dim, k, idx_base, row_lo, row_hi, out_dist, out_idx, pl, w, sl,
                          (cudaStream_t)stream);
}

extern "C" int exb_knn_screened_f32(const float* src, int64_t nq, const float* bank, int64_t nb,
                                    int dim, int k, int32_t idx_base, int src_is_bank,
                                    float* out_dist, int32_t* out_idx, void* workspace,
                                    size_t workspace_bytes, exb_stream_t stream) {
  return screened_impl(src, nq, bank, nb, dim, k, idx_base, src_is_bank, out_dist, out_idx,
                       workspace, workspace_bytes, stream, nullptr, nullptr);
}

extern "C" int exb_knn_screened_pre_f32(const float* src, int64_t nq, const exb_operand* src_op,
                                        const float* bank, int64_t nb, const exb_operand* bank_op,
                                        int dim, int k, int32_t idx_base, float* out_dist,
                                        int32_t* out_idx, void* workspace, size_t workspace_bytes,
                                        exb_stream_t stream) {
  EXB_CHECK_ARG(bank_op && bank_op->f16 && bank_op->half_norm && bank_op->round_err &&
                bank_op->scalars, "bank_op missing");
  const bool same = src == bank && nq == nb && (src_op == nullptr || src_op == bank_op ||
                                                src_op->f16 == bank_op->f16);
  if (!same)
    EXB_CHECK_ARG(src_op && src_op->f16 && src_op->half_norm && src_op->round_err &&
                  src_op->scalars, "src_op missing (src is not the bank)");
  EXB_CHECK_ARG((((uintptr_t)bank_op->f16 | (uintptr_t)(same ? bank_op : src_op)->f16) & 127) == 0,
                "prepared fp16 operands must be 128-byte aligned (TMA)");
  return screened_impl(src, nq, bank, nb, dim, k, idx_base, same ? 1 : 0, out_dist, out_idx,
                       workspace, workspace_bytes, stream, nullptr, nullptr, nullptr,
                       same ? bank_op : src_op, bank_op);
}

// ------------------------------------------------------------------------------------------
// Sharded relabel, "exchange the candidate lists" form: screen only, then (after the exchange)
// rerank the lists of all shards for this rank's query block.
// ------------------------------------------------------------------------------------------
extern "C" int exb_knn_screen_lists_count(int64_t nq, int64_t nb, int dim, int k) {
  ScreenPlan pl;
  if (!make_plan(nq, nb, dim, k, &pl) || pl.lat) return 0;
  return pl.splits * pl.trackers;
}

extern "C" int exb_knn_screen_lists_f32(const float* src, int64_t nq, const float* bank,
                                        int64_t nb, int dim, int k, int src_is_bank,
                                        float* out_val, int32_t* out_idx, int32_t* out_cnt,
                                        float* out_ha, float* out_ea, void* out_scalars,
                                        void* workspace, size_t workspace_bytes,
                                        exb_stream_t stream) {
  EXB_CHECK_ARG(out_val && out_idx && out_cnt && out_ha && out_ea && out_scalars, "null output");
  ListsOut lo = {out_val, out_idx, out_cnt, out_ha, out_ea, out_scalars};
  return screened_impl(src, nq, bank, nb, dim, k, 0, src_is_bank, nullptr, nullptr, workspace,
                       workspace_bytes, stream, nullptr, nullptr, &lo);
}

static int check_pre_ops(const float* src, int64_t nq, const exb_operand* src_op,
                         const float* bank, int64_t nb, const exb_operand* bank_op, bool* same) {
  EXB_CHECK_ARG(bank_op && bank_op->f16 && bank_op->half_norm && bank_op->round_err &&
                bank_op->scalars, "bank_op missing");
  *same = src == bank && nq == nb && (src_op == nullptr || src_op == bank_op ||
                                      src_op->f16 == bank_op->f16);
  if (!*same)
    EXB_CHECK_ARG(src_op && src_op->f16 && src_op->half_norm && src_op->round_err &&
                  src_op->scalars, "src_op missing (src is not the bank)");
  EXB_CHECK_ARG((((uintptr_t)bank_op->f16 | (uintptr_t)(*same ? bank_op : src_op)->f16) & 127) == 0,
                "prepared fp16 operands must be 128-byte aligned (TMA)");
  return EXB_OK;
}

extern "C" int exb_knn_screen_lists_pre_f32(const float* src, int64_t nq,
                                            const exb_operand* src_op, const float* bank,
                                            int64_t nb, const exb_operand* bank_op, int dim,
                                            int k, float* out_val, int32_t* out_idx,
                                            int32_t* out_cnt, float* out_ha, float* out_ea,
                                            void* out_scalars, void* workspace,
                                            size_t workspace_bytes, exb_stream_t stream) {
  EXB_CHECK_ARG(out_val && out_idx && out_cnt && out_ha && out_ea && out_scalars, "null output");
  bool same = false;
  const int rc = check_pre_ops(src, nq, src_op, bank, nb, bank_op, &same);
  if (rc) return rc;
  ListsOut lo = {out_val, out_idx, out_cnt, out_ha, out_ea, out_scalars};
  return screened_impl(src, nq, bank, nb, dim, k, 0, same ? 1 : 0, nullptr, nullptr, workspace,
                       workspace_bytes, stream, nullptr, nullptr, &lo, same ? bank_op : src_op,
                       bank_op);
}

struct RerankListsWs {
  RerankStatus* status;
  int* fb_rows;
  float* fb_d2;
  int* fb_idx;
  int fb_cap, fb_splits;
};

static size_t carve_rerank_lists(void* ws, int64_t nq, int64_t nb, int k, RerankListsWs* o) {
  Carver cv(ws ? ws : reinterpret_cast<void*>(256));
  o->fb_cap = (int)(nq < FALLBACK_CAP ? nq : FALLBACK_CAP);
  int fbs = (int)((nb + 127) / 128);
  if (fbs > FALLBACK_SPLITS) fbs = FALLBACK_SPLITS;
  if (fbs * k > 512) fbs = 512 / k;
  o->fb_splits = fbs < 1 ? 1 : fbs;
  o->status = cv.take<RerankStatus>(1);
  o->fb_rows = cv.take<int>(o->fb_cap);
  o->fb_d2 = cv.take<float>((size_t)o->fb_cap * o->fb_splits * k);
  o->fb_idx = cv.take<int>((size_t)o->fb_cap * o->fb_splits * k);
  return cv.off + 256;
}

extern "C" size_t exb_knn_rerank_lists_workspace_bytes(int64_t nq, int64_t nb, int k) {
  if (nq <= 0 || nb <= 0 || k < 1 || k > SCR_MAX_K) return 0;
  RerankListsWs w;
  return carve_rerank_lists(nullptr, nq, nb, k, &w);
}

extern "C" int exb_knn_rerank_lists_f32(const float* src, int64_t nq, const float* bank,
                                        int64_t nb, int dim, int k, int n_shards,
                                        int lists_per_shard, const float* cand_val,
                                        const int32_t* cand_idx, const int32_t* cand_cnt,
                                        const int32_t* host_shard_lo, const float* ha,
                                        const float* ea, const void* scalars, float* out_dist,
                                        int32_t* out_idx, void* workspace, size_t workspace_bytes,
                                        exb_stream_t stream) {
  EXB_CHECK_ARG(src && bank && cand_val && cand_idx && cand_cnt && host_shard_lo && ha && ea &&
                    scalars && out_dist && out_idx, "null pointer");
  EXB_CHECK_ARG(nq > 0 && nb > 0 && nq < (1LL << 31) && nb < (1LL << 31), "bad row counts");
  EXB_CHECK_ARG(dim >= 1 && dim <= EXB_SCREEN_MAX_DIM && k >= 1 && k <= SCR_MAX_K,
                "unsupported dim=%d or k=%d", dim, k);
  EXB_CHECK_ARG(n_shards >= 1 && lists_per_shard >= 1 && n_shards * lists_per_shard <= 16,
                "n_shards x lists_per_shard = %d x %d exceeds the 16 lists a row can gather",
                n_shards, lists_per_shard);
  if (k > nb) {
    set_error("selected index k out of range (k=%d > bank rows=%lld)", k, (long long)nb);
    return EXB_EKRANGE;
  }
  RerankListsWs w;
  const size_t need = carve_rerank_lists(nullptr, nq, nb, k, &w);
  EXB_CHECK_ARG(workspace && workspace_bytes >= need, "workspace too small: %zu < %zu",
                workspace_bytes, need);
  EXB_CHECK_ARG(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
  carve_rerank_lists(workspace, nq, nb, k, &w);
  cudaStream_t st = (cudaStream_t)stream;
  EXB_CUDA(cudaMemsetAsync(w.status, 0, 256, st));
  const bool vec = (dim % 4 == 0) && (((uintptr_t)src | (uintptr_t)bank) & 15) == 0;
  const int ni = vec ? (dim + 127) / 128 : 0;
  RerankParams rp;
  memset(&rp, 0, sizeof(rp));
  rp.Q = src; rp.B = bank; rp.nq = nq; rp.nb = nb; rp.dim = dim; rp.k = k; rp.idx_base = 0;
  rp.splits = n_shards * lists_per_shard;
  rp.lists_per_shard = lists_per_shard;
  rp.shard_rows = nq;
  for (int i = 0; i < 16; ++i) rp.shard_lo[i] = i < n_shards ? host_shard_lo[i] : 0;
  rp.cand_val = cand_val; rp.cand_idx = cand_idx; rp.cand_cnt = cand_cnt;
  rp.ha = ha; rp.ea = ea;
  rp.qsc = static_cast<const PrepScalars*>(scalars);
  rp.bsc = rp.qsc;
  rp.out_dist = out_dist; rp.out_idx = out_idx;
  rp.fallback_rows = w.fb_rows; rp.fallback_cap = w.fb_cap; rp.status = w.status;
  launch_rerank<false, true>(rp, (unsigned)((nq + RR_WARPS - 1) / RR_WARPS), ni, st);
  EXB_LAUNCH_CHECK();
  // undecided rows: exact kernel against the WHOLE bank (every rank holds it)
  return launch_exact(src, nq, bank, nb, dim, k, 0, out_dist, out_idx, w.fb_d2, w.fb_idx,
                      w.fb_splits, w.fb_rows, &w.status->n_fallback, w.fb_cap, st);
}

// ------------------------------------------------------------------------------------------
// PBE.__call__ in one call (utils/utils.py:289-319): kNN + RMS + epilogue
// ------------------------------------------------------------------------------------------
static size_t apt_knn_ws_bytes(int64_t nq, int64_t nb, int dim, int k) {
  const size_t a = exb_knn_screened_workspace_bytes(nq, nb, dim, k);
  const size_t b = exb_knn_exact_workspace_bytes(nq, nb, dim, k);
  return align_up(a > b ? a : b, 256);
}

extern "C" size_t exb_apt_reward_workspace_bytes(int64_t nq, int64_t nb, int dim, int k) {
  if (nq <= 0 || nb <= 0 || dim <= 0 || k <= 0 || k > EXB_MAX_K) return 0;
  return apt_knn_ws_bytes(nq, nb, dim, k) + align_up(exb_pbe_workspace_bytes(nq, k), 256);
}

static int apt_reward_impl(const float* src, int64_t nq, const float* bank, int64_t nb,
                           int dim, int k, int src_is_bank, int mode, int avg, int use_rms,
                           float clip, float* rms_m, float* rms_s, double* rms_n,
                           float* out_reward, float* out_dist, int32_t* out_idx,
                           void* workspace, size_t workspace_bytes, exb_stream_t stream,
                           const exb_operand* pre_q, const exb_operand* pre_b) {
  EXB_CHECK_ARG(src && bank && out_reward && out_dist && out_idx, "null pointer");
  EXB_CHECK_ARG(nq > 0 && nb > 0 && dim > 0, "empty input");
  EXB_CHECK_ARG(k >= 1 && k <= EXB_MAX_K, "k=%d outside [1,%d]", k, EXB_MAX_K);
  EXB_CHECK_ARG(mode >= 0 && mode <= 2, "mode must be 0 (auto), 1 (exact) or 2 (screened)");
  EXB_CHECK_ARG(!use_rms || (rms_m && rms_s && rms_n),
                "use_rms needs the RMS state pointers");
  if (k > nb) {
    set_error("selected index k out of range (k=%d > bank rows=%lld)", k, (long long)nb);
    return EXB_EKRANGE;
  }
  const size_t knn_bytes = apt_knn_ws_bytes(nq, nb, dim, k);
  const size_t pbe_bytes = exb_pbe_workspace_bytes(nq, k);
  EXB_CHECK_ARG(workspace && workspace_bytes >= knn_bytes + pbe_bytes,
                "workspace too small: %zu < %zu", workspace_bytes, knn_bytes + pbe_bytes);
  void* pbe_ws = static_cast<char*>(workspace) + knn_bytes;
  const bool can_screen = dim <= EXB_SCREEN_MAX_DIM && k <= SCR_MAX_K;
  if (mode == 2 && !can_screen) {
    set_error("screened kNN supports dim <= %d and k <= %d (got dim=%d k=%d)", EXB_SCREEN_MAX_DIM,
              SCR_MAX_K, dim, k);
    return EXB_EUNSUPPORTED;
  }
  const bool screen = mode == 2 || (mode == 0 && can_screen && nq >= 128 && nb >= 256);
  int rc;
  if (screen) {
    FusedPbe f;
    f.on = 1; f.raw_kth = 0; f.avg = avg; f.use_rms = use_rms; f.use_clip = clip >= 0.0f; f.clip = clip;
    f.rms = RmsParams{rms_m, rms_s, rms_n};
    f.out_reward = out_reward; f.partials = nullptr;
    bool fused = false;
    rc = screened_impl(src, nq, bank, nb, dim, k, 0, src_is_bank, out_dist, out_idx, workspace,
                       knn_bytes, stream, &f, &fused, nullptr, pre_q, pre_b);
    if (rc || fused) return rc;
  } else {
    rc = exb_knn_exact_f32(src, nq, bank, nb, dim, k, 0, out_dist, out_idx, workspace, knn_bytes,
                           stream);
    if (rc) return rc;
  }
  return exb_pbe_reward(out_dist, nq, k, avg, use_rms, clip, rms_m, rms_s, rms_n, out_reward, pbe_ws, pbe_bytes, stream);
}

extern "C" int exb_apt_reward_f32(const float* src, int64_t nq, const float* bank, int64_t nb,
                                  int dim, int k, int src_is_bank, int mode, int avg, int use_rms,
                                  float clip, float* rms_m, float* rms_s, double* rms_n,
                                  float* out_reward, float* out_dist, int32_t* out_idx,
                                  void* workspace, size_t workspace_bytes, exb_stream_t stream) {
  return apt_reward_impl(src, nq, bank, nb, dim, k, src_is_bank, mode, avg, use_rms, clip,
                         rms_m, rms_s, rms_n, out_reward, out_dist,
                         out_idx, workspace, workspace_bytes, stream, nullptr, nullptr);
}

extern "C" int exb_apt_reward_pre_f32(const float* src, int64_t nq, const exb_operand* src_op,
                                      const float* bank, int64_t nb, const exb_operand* bank_op,
                                      int dim, int k, int avg, int use_rms, float clip,
                                      float* rms_m, float* rms_s, double* rms_n, float* out_reward,
                                      float* out_dist, int32_t* out_idx, void* workspace,
                                      size_t workspace_bytes, exb_stream_t stream) {
  EXB_CHECK_ARG(bank_op && bank_op->f16 && bank_op->half_norm && bank_op->round_err &&
                bank_op->scalars, "bank_op missing");
  const bool same = src == bank && nq == nb && (src_op == nullptr || src_op == bank_op ||
                                                src_op->f16 == bank_op->f16);
  if (!same)
    EXB_CHECK_ARG(src_op && src_op->f16 && src_op->half_norm && src_op->round_err &&
                  src_op->scalars, "src_op missing (src is not the bank)");
  EXB_CHECK_ARG((((uintptr_t)bank_op->f16 | (uintptr_t)(same ? bank_op : src_op)->f16) & 127) == 0,
                "prepared fp16 operands must be 128-byte aligned (TMA)");
  return apt_reward_impl(src, nq, bank, nb, dim, k, same ? 1 : 0, 2, avg, use_rms, clip, rms_m, rms_s, rms_n, out_reward, out_dist, out_idx,
                         workspace, workspace_bytes, stream, same ? bank_op : src_op, bank_op);
}

// ProtoRL reward as one call: distance of every z row to its k-th nearest queue row
// (proto.py:115-118: no RMS, no clip, no log).  Same kernels as exb_apt_reward_f32 with the
// epilogue reduced to "take column k - 1" inside the rerank, so the per-update chain needs no
// extra launch.
extern "C" int exb_proto_reward_f32(const float* z, int64_t nq, const float* queue, int64_t nb,
                                    int dim, int k, int mode, float* out_reward, float* out_dist,
                                    int32_t* out_idx, void* workspace, size_t workspace_bytes,
                                    exb_stream_t stream) {
  EXB_CHECK_ARG(z && queue && out_reward && out_dist && out_idx, "null pointer");
  EXB_CHECK_ARG(nq > 0 && nb > 0 && dim > 0, "empty input");
  EXB_CHECK_ARG(k >= 1 && k <= EXB_MAX_K, "k=%d outside [1,%d]", k, EXB_MAX_K);
  EXB_CHECK_ARG(mode >= 0 && mode <= 2, "mode must be 0 (auto), 1 (exact) or 2 (screened)");
  if (k > nb) {
    set_error("selected index k out of range (k=%d > bank rows=%lld)", k, (long long)nb);
    return EXB_EKRANGE;
  }
  const size_t knn_bytes = apt_knn_ws_bytes(nq, nb, dim, k);
  EXB_CHECK_ARG(workspace && workspace_bytes >= knn_bytes, "workspace too small: %zu < %zu",
                workspace_bytes, knn_bytes);
  const bool can_screen = dim <= EXB_SCREEN_MAX_DIM && k <= SCR_MAX_K;
  if (mode == 2 && !can_screen) {
    set_error("screened kNN supports dim <= %d and k <= %d (got dim=%d k=%d)", EXB_SCREEN_MAX_DIM,
              SCR_MAX_K, dim, k);
    return EXB_EUNSUPPORTED;
  }
  int rc;
  if (mode == 2 || (mode == 0 && can_screen && nq >= 128 && nb >= 256)) {
    FusedPbe f;
    f.on = 1; f.raw_kth = 1; f.avg = 0; f.use_rms = 0; f.use_clip = 0; f.clip = -1.f;
    f.rms = RmsParams{nullptr, nullptr, nullptr};
    f.out_reward = out_reward; f.partials = nullptr;
    bool fused = false;
    rc = screened_impl(z, nq, queue, nb, dim, k, 0, z == queue && nq == nb, out_dist, out_idx,
                       workspace, knn_bytes, stream, &f, &fused, nullptr, nullptr, nullptr);
    if (rc || fused) return rc;
  } else {
    rc = exb_knn_exact_f32(z, nq, queue, nb, dim, k, 0, out_dist, out_idx, workspace, knn_bytes,
                           stream);
    if (rc) return rc;
  }
  return exb_take_kth(out_dist, nq, k, out_reward, stream);
}

extern "C" int exb_knn_set_per_update_scan(int block_cooperative) {
  return g_coop_scan.exchange(block_cooperative ? 1 : 0, std::memory_order_relaxed);
}

extern "C" int exb_knn_screened_status_sync(const void* workspace, int32_t host_status[4],
                                            exb_stream_t stream) {
  EXB_CHECK_ARG(workspace && host_status, "null pointer");
  RerankStatus s;
  EXB_CUDA(cudaMemcpyAsync(&s, workspace, sizeof(s), cudaMemcpyDeviceToHost,
                           (cudaStream_t)stream));
  EXB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  host_status[0] = s.n_fallback;
  host_status[1] = s.invalid;
  host_status[2] = (int32_t)(s.pairs > 0x7fffffffULL ? 0x7fffffff : s.pairs);
  host_status[3] = (int32_t)(s.pairs >> 31);
  return EXB_OK;
}

extern "C" int exb_debug_trace_sync(const void* workspace, int64_t host_ns[16],
                                    exb_stream_t stream) {
  EXB_CHECK_ARG(workspace && host_ns, "null pointer");
  // carve_screen: status, qsc, bsc, trace at 256-byte steps
  EXB_CUDA(cudaMemcpyAsync(host_ns, static_cast<const char*>(workspace) + 768, 16 * 8,
                           cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  EXB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return EXB_OK;
}

extern "C" int exb_profile_begin(int enable) {
  if (enable && !g_prof.created) {
    for (int i = 0; i < 64; ++i) {
      EXB_CUDA(cudaEventCreate(&g_prof.e0[i]));
      EXB_CUDA(cudaEventCreate(&g_prof.e1[i]));
    }
    g_prof.created = true;
  }
  g_prof.on = enable != 0;
  g_prof.n = 0;
  return EXB_OK;
}

extern "C" int exb_profile_collect_sync(float* host_ms, int cap, int* host_n) {
  EXB_CHECK_ARG(host_ms && host_n && cap >= 0, "bad arguments");
  int n = g_prof.n < cap ? g_prof.n : cap;
  for (int i = 0; i < n; ++i) {
    EXB_CUDA(cudaEventSynchronize(g_prof.e1[i]));
    EXB_CUDA(cudaEventElapsedTime(&host_ms[i], g_prof.e0[i], g_prof.e1[i]));
  }
  *host_n = n;
  return EXB_OK;
}
