// ProtoRL candidate sampling fused with the queue write.
//
// Replaces /root/reference/agents/unsupervised_learning/proto.py:105-112:
//   scores = protos(z).T                      # (P, B): prototype p against every batch row
//   prob   = softmax(scores, dim=1)           # over the batch, per prototype
//   candidates = Categorical(prob).sample()   # one batch row per prototype
//   queue[ptr:ptr+P] = z[candidates]; ptr = (ptr + P) % queue_size
// Categorical renormalises (probs / probs.sum(-1)) and torch.multinomial with one sample draws
// q ~ Exp(1) per entry and returns argmax(probs / q) — the exponential race — which is what the
// kernel evaluates, with q either given by the caller (parity mode: the reference's own noise
// tensor) or drawn from Philox4x32-10 keyed by (seed, call) like the replay index draw.
//
// The contraction is 2*P*B*D = 134 MFLOP at the reference's sizes (512 x 1024 x 128): launch- and
// L2-latency bound, so it runs on the CUDA cores; one block owns PS_PROTOS prototypes, keeps
// their weight rows in shared memory and streams z through registers (every z row is read once
// per block and used for all PS_PROTOS prototypes), then one warp per prototype does the two
// softmax reductions, the race and the queue row copy.
#include "common.cuh"
#include <float.h>
#include <math.h>

namespace exb {

constexpr int PS_PROTOS = 8;      // prototypes per block = warps per block
constexpr int PS_THREADS = 256;
constexpr int PS_MAX_BATCH = 6144;  // PS_PROTOS * B floats of scores in shared memory (192 KB)

__device__ __forceinline__ void ps_philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
  c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}

// Exp(1) variate of entry (p, b): Philox counter (b, p, call_lo, call_hi), key = seed, word 0;
// u = (x + 0.5) / 2^32 in (0, 1), q = -log(u) > 0.
__device__ __forceinline__ float ps_exponential(uint32_t p, uint32_t b, uint64_t seed,
                                                uint64_t call) {
  uint32_t c[4] = {b, p, (uint32_t)call, (uint32_t)(call >> 32)};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    ps_philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  const float u = ((float)(c[0] >> 8) + 0.5f) * (1.0f / 16777216.0f);   // 24 bits: exact in fp32
  return -logf(u);
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

template <bool VEC>
__global__ void __launch_bounds__(PS_THREADS)
proto_sample_kernel(const float* __restrict__ z, int64_t batch, const float* __restrict__ protos,
                    int n_protos, int dim, const float* __restrict__ noise, uint64_t seed,
                    uint64_t call, float* __restrict__ queue, int64_t queue_ptr,
                    int64_t* __restrict__ out_candidates, float* __restrict__ out_prob_max) {
  extern __shared__ float smem[];
  float* s_w = smem;                               // [PS_PROTOS][dim]
  float* s_sc = smem + PS_PROTOS * dim;            // [PS_PROTOS][batch]
  const int p0 = blockIdx.x * PS_PROTOS;
  const int np = min(PS_PROTOS, n_protos - p0);
  for (int i = threadIdx.x; i < PS_PROTOS * dim; i += PS_THREADS) {
    const int p = i / dim;
    s_w[i] = (p < np) ? __ldg(protos + (int64_t)(p0 + p) * dim + (i - p * dim)) : 0.f;
  }
  __syncthreads();

  // ---- scores[p][b] = <protos[p0 + p], z[b]>  (nn.Linear without bias, proto.py:105)
  for (int64_t b = threadIdx.x; b < batch; b += PS_THREADS) {
    float acc[PS_PROTOS];
#pragma unroll
    for (int p = 0; p < PS_PROTOS; ++p) acc[p] = 0.f;
    const float* zr = z + b * dim;
    if constexpr (VEC) {
      const float4* z4 = reinterpret_cast<const float4*>(zr);
      for (int c = 0; c < (dim >> 2); ++c) {
        const float4 v = __ldg(z4 + c);
#pragma unroll
        for (int p = 0; p < PS_PROTOS; ++p) {
          const float4 w = reinterpret_cast<const float4*>(s_w + p * dim)[c];
          acc[p] = fmaf(v.x, w.x, acc[p]);
          acc[p] = fmaf(v.y, w.y, acc[p]);
          acc[p] = fmaf(v.z, w.z, acc[p]);
          acc[p] = fmaf(v.w, w.w, acc[p]);
        }
      }
    } else {
      for (int c = 0; c < dim; ++c) {
        const float v = __ldg(zr + c);
#pragma unroll
        for (int p = 0; p < PS_PROTOS; ++p) acc[p] = fmaf(v, s_w[p * dim + c], acc[p]);
      }
    }
#pragma unroll
    for (int p = 0; p < PS_PROTOS; ++p) s_sc[(int64_t)p * batch + b] = acc[p];
  }
  __syncthreads();

  // ---- one warp per prototype: softmax over the batch (:106), Categorical's renormalisation,
  //      exponential race (:107), queue row write (:110-111)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp >= np) return;
  const int p = p0 + warp;
  float* sc = s_sc + (int64_t)warp * batch;
  float m = -FLT_MAX;
  for (int64_t b = lane; b < batch; b += 32) m = fmaxf(m, sc[b]);
  m = warp_max(m);
  float sum = 0.f;
  for (int64_t b = lane; b < batch; b += 32) {
    const float e = expf(sc[b] - m);
    sc[b] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  float sum2 = 0.f;
  for (int64_t b = lane; b < batch; b += 32) {
    const float pr = sc[b] / sum;          // softmax
    sc[b] = pr;
    sum2 += pr;
  }
  sum2 = warp_sum(sum2);
  float best = -1.f, best_p = 0.f;
  int64_t best_b = 0;
  for (int64_t b = lane; b < batch; b += 32) {
    const float pr = sc[b] / sum2;         // Categorical: probs / probs.sum(-1, keepdim=True)
    const float q = noise ? __ldg(noise + (int64_t)p * batch + b)
                          : ps_exponential((uint32_t)p, (uint32_t)b, seed, call);
    const float r = pr / q;
    if (r > best) { best = r; best_b = b; best_p = pr; }      // first maximum, like argmax
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, off);
    const int64_t oi = __shfl_xor_sync(0xffffffffu, best_b, off);
    const float op = __shfl_xor_sync(0xffffffffu, best_p, off);
    if (ob > best || (ob == best && oi < best_b)) { best = ob; best_b = oi; best_p = op; }
  }
  if (lane == 0) {
    out_candidates[p] = best_b;
    if (out_prob_max) out_prob_max[p] = best_p;
  }
  if (queue) {
    const float* src = z + best_b * dim;
    float* dst = queue + (queue_ptr + p) * dim;
    for (int c = lane; c < dim; c += 32) dst[c] = __ldg(src + c);
  }
}

}  // namespace exb

using namespace exb;

extern "C" int exb_proto_sample_f32(const float* z, int64_t batch, const float* protos,
                                    int n_protos, int dim, const float* noise, uint64_t seed,
                                    uint64_t call, float* queue, int64_t queue_rows,
                                    int64_t queue_ptr, int64_t* out_candidates,
                                    float* out_prob, exb_stream_t stream) {
  EXB_CHECK_ARG(z && protos && out_candidates, "null pointer");
  EXB_CHECK_ARG(batch > 0 && n_protos > 0 && dim > 0, "empty input");
  if (batch > PS_MAX_BATCH || dim > 1024) {
    set_error("proto candidate sampling supports batch <= %d and dim <= 1024 (got %lld, %d)",
              PS_MAX_BATCH, (long long)batch, dim);
    return EXB_EUNSUPPORTED;
  }
  if (queue)
    EXB_CHECK_ARG(queue_ptr >= 0 && queue_ptr + n_protos <= queue_rows,
                  "queue slice [%lld, %lld) exceeds the %lld queue rows", (long long)queue_ptr,
                  (long long)(queue_ptr + n_protos), (long long)queue_rows);
  const size_t smem = ((size_t)PS_PROTOS * dim + (size_t)PS_PROTOS * batch) * sizeof(float);
  const bool vec = (dim % 4 == 0) && (((uintptr_t)z | (uintptr_t)protos) & 15) == 0;
  const unsigned grid = (unsigned)((n_protos + PS_PROTOS - 1) / PS_PROTOS);
  cudaStream_t st = (cudaStream_t)stream;
  if (vec) {
    EXB_CUDA(cudaFuncSetAttribute(proto_sample_kernel<true>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    proto_sample_kernel<true><<<grid, PS_THREADS, smem, st>>>(
        z, batch, protos, n_protos, dim, noise, seed, call, queue, queue_ptr, out_candidates,
        out_prob);
  } else {
    EXB_CUDA(cudaFuncSetAttribute(proto_sample_kernel<false>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    proto_sample_kernel<false><<<grid, PS_THREADS, smem, st>>>(
        z, batch, protos, n_protos, dim, noise, seed, call, queue, queue_ptr, out_candidates,
        out_prob);
  }
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}
