// Running-moment (RMS) update + PBE reward epilogue.
//
// Replaces /root/reference/utils/utils.py:264-276 (RMS.__call__: batch mean, unbiased variance,
// Chan combination with the carried M, S, n) and :301-318 (divide by the UPDATED mean, clip,
// mean over k, log(1 + r)), which the reference runs as ~25 launch-bound ATen kernels.
//
// State convention (include/exorl_b200.h): M and S are the fp32 (1,) tensors the reference
// exposes, n is the reference's python float held as one fp64 word on the device; all three are
// updated in place (n += count, utils/utils.py:274), so a call has no by-value argument that
// changes from step to step and can be replayed from a CUDA graph.  Sums are accumulated in fp64
// and the updated M, S are rounded to fp32, like the reference's fp32 tensors.
#include "common.cuh"
#include "pbe_common.cuh"
#include <math.h>

namespace exb {

constexpr int PBE_SMALL_THREADS = 1024;
constexpr int64_t PBE_SMALL_LIMIT = 16384;  // values handled by the single-block kernel

// Partial sums per block, then the last block to finish (ticket) reduces them in a fixed order,
// optionally applies the RMS update, and re-arms the ticket.
__global__ void __launch_bounds__(PBE_THREADS)
pbe_moments_kernel(const float* __restrict__ x, int64_t count, int k, int avg,
                   double2* __restrict__ partials, unsigned* __restrict__ ticket,
                   double* __restrict__ out_sums, int apply, RmsParams p) {
  __shared__ double2 red[PBE_THREADS / 32];
  __shared__ bool is_last;
  double2 acc = make_double2(0.0, 0.0);
  for (int64_t i = (int64_t)blockIdx.x * PBE_THREADS + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * PBE_THREADS) {
    const double v = (double)x[value_index(i, k, avg)];
    acc.x += v;
    acc.y += v * v;
  }
  const double2 tot = block_sum2<PBE_THREADS>(acc, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = tot;
    __threadfence();
    const unsigned t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  double2 a = make_double2(0.0, 0.0);
  for (int i = threadIdx.x; i < (int)gridDim.x; i += PBE_THREADS) {
    const double2 t = partials[i];
    a.x += t.x;
    a.y += t.y;
  }
  const double2 all = block_sum2<PBE_THREADS>(a, red);
  if (threadIdx.x == 0) {
    out_sums[0] = (double)count;
    out_sums[1] = all.x;
    out_sums[2] = all.y;
    if (apply) rms_apply(out_sums, p);
    *ticket = 0u;
  }
}

__global__ void rms_apply_kernel(const double* __restrict__ sums, RmsParams p) {
  if (threadIdx.x == 0 && blockIdx.x == 0) rms_apply(sums, p);
}

__global__ void __launch_bounds__(PBE_THREADS)
pbe_finish_kernel(const float* __restrict__ dist, int64_t nq, int k, int avg, int use_rms,
                  int use_clip, float clip, const float* __restrict__ m_ptr,
                  float* __restrict__ out) {
  pdl_wait();   // dependent launch behind the fused rerank + moments kernel (no-op otherwise)
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  const float m = use_rms ? *m_ptr : 1.0f;
  out[i] = pbe_row(dist + i * k, k, avg, use_rms ? 1.f : 0.f, m, use_clip, clip);
}

// Whole epilogue in one block for the per-update sizes (B = 1024, k = 12).
__global__ void __launch_bounds__(PBE_SMALL_THREADS)
pbe_small_kernel(const float* __restrict__ dist, int64_t nq, int k, int avg, int use_rms,
                 int use_clip, float clip, RmsParams p, float* __restrict__ out) {
  __shared__ double2 red[PBE_SMALL_THREADS / 32];
  __shared__ double sums[3];
  if (use_rms) {
    const int64_t count = avg ? nq * k : nq;
    double2 acc = make_double2(0.0, 0.0);
    for (int64_t i = threadIdx.x; i < count; i += PBE_SMALL_THREADS) {
      const double v = (double)dist[value_index(i, k, avg)];
      acc.x += v;
      acc.y += v * v;
    }
    const double2 all = block_sum2<PBE_SMALL_THREADS>(acc, red);
    if (threadIdx.x == 0) {
      sums[0] = (double)count;
      sums[1] = all.x;
      sums[2] = all.y;
      rms_apply(sums, p);
      __threadfence_block();
    }
    __syncthreads();
  }
  const float m = use_rms ? *p.m : 1.0f;
  for (int64_t i = threadIdx.x; i < nq; i += PBE_SMALL_THREADS)
    out[i] = pbe_row(dist + i * k, k, avg, use_rms ? 1.f : 0.f, m, use_clip, clip);
}

__global__ void take_kth_kernel(const float* __restrict__ dist, int64_t nq, int k,
                                float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nq) out[i] = dist[i * k + (k - 1)];
}

// per-row finish after the fused rerank + moments kernel of the per-update chain (knn_screen.cu)
int launch_pbe_finish(const float* dist, int64_t nq, int k, int avg, int use_rms, int use_clip,
                      float clip, const float* m_ptr, float* out, cudaStream_t st) {
  static const int pdl_on = knob("EXB_PDL", 0);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)((nq + 127) / 128));
  cfg.blockDim = dim3(128);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_on ? 1 : 0;
  EXB_CUDA(cudaLaunchKernelEx(&cfg, pbe_finish_kernel, dist, nq, k, avg, use_rms, use_clip, clip,
                              m_ptr, out));
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

static int moments_blocks(int64_t count) {
  int64_t b = (count + PBE_THREADS * 8 - 1) / (PBE_THREADS * 8);
  const int64_t cap = 4LL * sm_count();
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace exb

using namespace exb;

extern "C" size_t exb_pbe_workspace_bytes(int64_t nq, int k) {
  (void)nq;
  (void)k;
  return 256 /*ticket*/ + 256 /*sums*/ + align_up(4 * 160 * sizeof(double2) * 2, 256);
}

namespace {
struct PbeWs {
  unsigned* ticket;
  double* sums;
  double2* partials;
};
int carve_pbe(void* ws, size_t bytes, PbeWs* o) {
  EXB_CHECK_ARG(ws && bytes >= exb_pbe_workspace_bytes(0, 0), "pbe workspace too small");
  EXB_CHECK_ARG(4 * sm_count() <= 4 * 160 * 2, "unexpected SM count");
  Carver cv(ws);
  o->ticket = cv.take<unsigned>(1);
  o->sums = cv.take<double>(4);
  o->partials = cv.take<double2>(4 * 160 * 2);
  return EXB_OK;
}
}  // namespace

// The ticket word must start at zero: the workspace is zero-initialised once by the caller
// (exb_pbe_workspace_init) and every launch re-arms it.
extern "C" int exb_pbe_workspace_init(void* workspace, size_t workspace_bytes,
                                      exb_stream_t stream) {
  EXB_CHECK_ARG(workspace && workspace_bytes >= exb_pbe_workspace_bytes(0, 0),
                "pbe workspace too small");
  EXB_CUDA(cudaMemsetAsync(workspace, 0, exb_pbe_workspace_bytes(0, 0), (cudaStream_t)stream));
  return EXB_OK;
}

extern "C" int exb_pbe_reward(const float* knn_dist, int64_t nq, int k, int avg, int use_rms,
                              float clip, float* rms_m, float* rms_s, double* rms_n,
                              float* out_reward, void* workspace, size_t workspace_bytes,
                              exb_stream_t stream) {
  EXB_CHECK_ARG(knn_dist && out_reward && nq > 0 && k >= 1, "bad arguments");
  EXB_CHECK_ARG(!use_rms || (rms_m && rms_s && rms_n),
                "use_rms needs the RMS state pointers");
  cudaStream_t st = (cudaStream_t)stream;
  RmsParams p{rms_m, rms_s, rms_n};
  const int use_clip = clip >= 0.0f;
  const int64_t count = avg ? nq * k : nq;
  if (nq * k <= PBE_SMALL_LIMIT) {
    pbe_small_kernel<<<1, PBE_SMALL_THREADS, 0, st>>>(knn_dist, nq, k, avg, use_rms, use_clip,
                                                      clip, p, out_reward);
    EXB_LAUNCH_CHECK();
    return EXB_OK;
  }
  if (use_rms) {
    PbeWs w;
    int rc = carve_pbe(workspace, workspace_bytes, &w);
    if (rc) return rc;
    pbe_moments_kernel<<<moments_blocks(count), PBE_THREADS, 0, st>>>(
        knn_dist, count, k, avg, w.partials, w.ticket, w.sums, 1, p);
    EXB_LAUNCH_CHECK();
  }
  pbe_finish_kernel<<<(unsigned)((nq + PBE_THREADS - 1) / PBE_THREADS), PBE_THREADS, 0, st>>>(
      knn_dist, nq, k, avg, use_rms, use_clip, clip, rms_m, out_reward);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

extern "C" int exb_pbe_moments(const float* knn_dist, int64_t nq, int k, int avg,
                               double* out_sums, void* workspace, size_t workspace_bytes,
                               exb_stream_t stream) {
  EXB_CHECK_ARG(knn_dist && out_sums && nq > 0 && k >= 1, "bad arguments");
  PbeWs w;
  int rc = carve_pbe(workspace, workspace_bytes, &w);
  if (rc) return rc;
  const int64_t count = avg ? nq * k : nq;
  RmsParams p{nullptr, nullptr, nullptr};
  pbe_moments_kernel<<<moments_blocks(count), PBE_THREADS, 0, (cudaStream_t)stream>>>(
      knn_dist, count, k, avg, w.partials, w.ticket, out_sums, 0, p);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

extern "C" int exb_pbe_finish(const float* knn_dist, int64_t nq, int k, int avg, int use_rms,
                              float clip, const double* sums, float* rms_m, float* rms_s, double* rms_n, float* out_reward, exb_stream_t stream) {
  EXB_CHECK_ARG(knn_dist && out_reward && nq > 0 && k >= 1, "bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (use_rms) {
    EXB_CHECK_ARG(sums && rms_m && rms_s && rms_n,
                  "use_rms needs sums and the RMS state pointers");
    RmsParams p{rms_m, rms_s, rms_n};
    rms_apply_kernel<<<1, 32, 0, st>>>(sums, p);
    EXB_LAUNCH_CHECK();
  }
  pbe_finish_kernel<<<(unsigned)((nq + PBE_THREADS - 1) / PBE_THREADS), PBE_THREADS, 0, st>>>(
      knn_dist, nq, k, avg, use_rms, clip >= 0.0f, clip, rms_m, out_reward);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

extern "C" int exb_rms_update(const float* x, int64_t n, float* rms_m, float* rms_s, double* rms_n, void* workspace, size_t workspace_bytes,
                              exb_stream_t stream) {
  EXB_CHECK_ARG(x && n > 0 && rms_m && rms_s && rms_n, "bad arguments");
  PbeWs w;
  int rc = carve_pbe(workspace, workspace_bytes, &w);
  if (rc) return rc;
  RmsParams p{rms_m, rms_s, rms_n};
  pbe_moments_kernel<<<moments_blocks(n), PBE_THREADS, 0, (cudaStream_t)stream>>>(
      x, n, 1, 1, w.partials, w.ticket, w.sums, 1, p);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

extern "C" int exb_take_kth(const float* knn_dist, int64_t nq, int k, float* out,
                            exb_stream_t stream) {
  EXB_CHECK_ARG(knn_dist && out && nq > 0 && k >= 1, "bad arguments");
  take_kth_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, (cudaStream_t)stream>>>(knn_dist, nq,
                                                                                  k, out);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

// ------------------------------------------------------------------------------------------
// RND intrinsic reward (the other consumer of the running moments):
//   prediction_error = mean_c (target - prediction)^2            rnd.py:52-54
//   _, var = RMS(prediction_error)                               rnd.py:96, utils/utils.py:264-276
//   reward = rnd_scale * prediction_error / (sqrt(var) + 1e-8)   rnd.py:97-98
// Launch 1: one warp per row computes the error, every block adds its rows' (sum, sumsq) in fp64,
// the last block (ticket) reduces the partials and applies the Chan update.  Launch 2: scaling by
// the UPDATED variance.
// ------------------------------------------------------------------------------------------
namespace exb {

__global__ void __launch_bounds__(PBE_THREADS)
rnd_error_moments_kernel(const float* __restrict__ pred, const float* __restrict__ target,
                         int64_t n, int dim, float* __restrict__ err_out,
                         double2* __restrict__ partials, unsigned* __restrict__ ticket,
                         double* __restrict__ out_sums, RmsParams p) {
  __shared__ double2 red[PBE_THREADS / 32];
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool vec = (dim & 3) == 0 &&
                   ((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(target)) & 15) == 0;
  const float inv_dim = 1.0f / (float)dim;
  double2 acc = make_double2(0.0, 0.0);
  for (int64_t row = (int64_t)blockIdx.x * (PBE_THREADS / 32) + warp; row < n;
       row += (int64_t)gridDim.x * (PBE_THREADS / 32)) {
    const float* a = pred + row * dim;
    const float* b = target + row * dim;
    float s = 0.f;
    if (vec) {
      const float4* a4 = reinterpret_cast<const float4*>(a);
      const float4* b4 = reinterpret_cast<const float4*>(b);
      for (int c = lane; c < (dim >> 2); c += 32) {
        const float4 x = __ldg(a4 + c), y = __ldg(b4 + c);
        const float d0 = y.x - x.x, d1 = y.y - x.y, d2 = y.z - x.z, d3 = y.w - x.w;
        s = fmaf(d0, d0, s); s = fmaf(d1, d1, s); s = fmaf(d2, d2, s); s = fmaf(d3, d3, s);
      }
    } else {
      for (int c = lane; c < dim; c += 32) {
        const float d = b[c] - a[c];
        s = fmaf(d, d, s);
      }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
    const float e = s * inv_dim;
    if (lane == 0) {
      err_out[row] = e;
      acc.x += (double)e;
      acc.y += (double)e * (double)e;
    }
  }
  const double2 tot = block_sum2<PBE_THREADS>(acc, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = tot;
    __threadfence();
    const unsigned t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  double2 a = make_double2(0.0, 0.0);
  for (int i = threadIdx.x; i < (int)gridDim.x; i += PBE_THREADS) {
    const double2 t = partials[i];
    a.x += t.x;
    a.y += t.y;
  }
  const double2 all = block_sum2<PBE_THREADS>(a, red);
  if (threadIdx.x == 0) {
    out_sums[0] = (double)n;
    out_sums[1] = all.x;
    out_sums[2] = all.y;
    rms_apply(out_sums, p);
    *ticket = 0u;
  }
}

__global__ void rnd_scale_kernel(const float* __restrict__ err, int64_t n, float scale,
                                 const float* __restrict__ s_ptr, float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float den = sqrtf(*s_ptr) + 1e-8f;
  out[i] = (scale * err[i]) / den;
}

}  // namespace exb

extern "C" int exb_rnd_reward_f32(const float* prediction, const float* target, int64_t n,
                                  int dim, float scale, float* rms_m, float* rms_s, double* rms_n, float* out_error, float* out_reward,
                                  void* workspace, size_t workspace_bytes, exb_stream_t stream) {
  EXB_CHECK_ARG(prediction && target && out_error && out_reward && n > 0 && dim > 0,
                "bad arguments");
  EXB_CHECK_ARG(rms_m && rms_s && rms_n, "RMS state pointers missing");
  PbeWs w;
  int rc = carve_pbe(workspace, workspace_bytes, &w);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  RmsParams p{rms_m, rms_s, rms_n};
  int64_t blocks = (n + PBE_THREADS / 32 - 1) / (PBE_THREADS / 32);
  const int64_t cap = 4LL * sm_count();
  if (blocks > cap) blocks = cap;
  rnd_error_moments_kernel<<<(unsigned)blocks, PBE_THREADS, 0, st>>>(
      prediction, target, n, dim, out_error, w.partials, w.ticket, w.sums, p);
  EXB_LAUNCH_CHECK();
  rnd_scale_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(out_error, n, scale, rms_s,
                                                               out_reward);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

// ------------------------------------------------------------------------------------------
// exb_pbe_reward for callers that may not allocate (no counterpart in the reference):
// the same two launches as exb_pbe_reward's large path
// (pbe_moments_kernel -> pbe_finish_kernel), with the moments kernel's scratch (ticket word,
// sums, per-block partials) in a file-scope device array instead of a caller workspace.  Every
// GPU that loads the library holds the array (PBE_NOALLOC_SCRATCH_BYTES = 16.5 KB) until exit.
// The host side never hands the array's NAME to a kernel (in host code that is a host-side
// stand-in, not a device address): the address comes from cudaGetSymbolAddress, once per device
// with that device current, and is cached here.
// ------------------------------------------------------------------------------------------
#include <mutex>

namespace {

constexpr int PBE_NOALLOC_MAX_BLOCKS = 1024;   // partial slots; a launch uses <= 4 x SM count
constexpr int PBE_NOALLOC_MAX_DEVICES = 64;
constexpr size_t PBE_NOALLOC_SCRATCH_BYTES =
    256 /*ticket*/ + 256 /*sums*/ + PBE_NOALLOC_MAX_BLOCKS * sizeof(double2);

static __device__ __align__(256) unsigned char g_pbe_noalloc_scratch[PBE_NOALLOC_SCRATCH_BYTES];

// Device address of the array on the CURRENT device (set-up on the first call per device).
int pbe_noalloc_scratch(unsigned char** out) {
  static std::mutex mu;
  static unsigned char* ptrs[PBE_NOALLOC_MAX_DEVICES] = {};
  int dev = -1;
  EXB_CUDA(cudaGetDevice(&dev));
  EXB_CHECK_ARG(dev >= 0 && dev < PBE_NOALLOC_MAX_DEVICES, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    EXB_CUDA(cudaGetSymbolAddress(&p, g_pbe_noalloc_scratch));
    EXB_CHECK_ARG(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    ptrs[dev] = static_cast<unsigned char*>(p);
  }
  *out = ptrs[dev];
  return EXB_OK;
}

}  // namespace

extern "C" int exb_pbe_reward_noalloc(const float* knn_dist, int64_t nq, int k, int avg,
                                      int use_rms, float clip, float* rms_m, float* rms_s,
                                      double* rms_n, float* out_reward, exb_stream_t stream) {
  EXB_CHECK_ARG(knn_dist && out_reward && nq > 0 && k >= 1, "bad arguments");
  EXB_CHECK_ARG(!use_rms || (rms_m && rms_s && rms_n),
                "use_rms needs the RMS state pointers");
  EXB_CHECK_ARG(nq <= EXB_PBE_NOALLOC_MAX_VALUES / k,
                "exb_pbe_reward_noalloc: nq * k = %lld x %d exceeds the static scratch's limit of "
                "%lld values (use exb_pbe_reward with a caller workspace)",
                (long long)nq, k, (long long)EXB_PBE_NOALLOC_MAX_VALUES);
  cudaStream_t st = (cudaStream_t)stream;
  const int use_clip = clip >= 0.0f;
  if (use_rms) {
    unsigned char* scratch = nullptr;
    int rc = pbe_noalloc_scratch(&scratch);
    if (rc) return rc;
    const int64_t count = avg ? nq * k : nq;
    const int blocks = moments_blocks(count);
    EXB_CHECK_ARG(blocks <= PBE_NOALLOC_MAX_BLOCKS, "unexpected SM count");
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
    double* sums = reinterpret_cast<double*>(scratch + 256);
    double2* partials = reinterpret_cast<double2*>(scratch + 512);
    // the array keeps the last call's contents: the ticket must start at zero (a call that
    // was aborted mid-kernel would otherwise poison every later one)
    EXB_CUDA(cudaMemsetAsync(ticket, 0, 256, st));
    RmsParams p{rms_m, rms_s, rms_n};
    pbe_moments_kernel<<<blocks, PBE_THREADS, 0, st>>>(knn_dist, count, k, avg, partials, ticket,
                                                       sums, 1, p);
    EXB_LAUNCH_CHECK();
  }
  pbe_finish_kernel<<<(unsigned)((nq + PBE_THREADS - 1) / PBE_THREADS), PBE_THREADS, 0, st>>>(
      knn_dist, nq, k, avg, use_rms, use_clip, clip, rms_m, out_reward);
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}
