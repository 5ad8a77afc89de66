// Rep producers: the step immediately in front of the kNN screen.  They
// compute the representation the reference feeds to PBE / the Proto kNN and emit, in the same
// pass, the fp32 rep (needed by the exact rerank) AND the prepared fp16 operand + row norms that
// exb_knn_prep_f16 would otherwise produce by re-reading the rep from HBM.
//
//   rep_trunk_kernel     ICM trunk  Linear(in_dim -> D) -> LayerNorm(D) -> Tanh
//                        (agents/unsupervised_learning/icm_apt.py:17-18, get_rep :50-52)
//   rep_ln_tanh_kernel   LayerNorm -> Tanh on given pre-activations (wide inputs: the Linear is a
//                        library GEMM, e.g. the 39200-wide pixel encoder output)
//   rep_normalize_kernel F.normalize(z, dim=1, p=2) (agents/unsupervised_learning/proto.py:104)
//
// All are HBM-write bound: per row 4*D (rep) + 2*Dpad + 8 (operand) bytes out, 4*in_dim in.
// The operand arithmetic (fp16 rounding, half norm, rounded-up error norm) is the SAME sequence of
// operations as knn_prep_kernel<true> in knn_screen.cu, so producer output == prep(rep) bitwise
// (tests/test_gpu_rep.py asserts it) and the screening certificate of DESIGN.md §4 carries over.
#include "common.cuh"
#include "prep_row.cuh"
#include "rep_gemm.cuh"
#include <algorithm>
#include <cuda_fp16.h>
#include <math.h>
#include <stdlib.h>

namespace exb {

constexpr int RP_WARPS = 8;
constexpr int RP_SMEM_LIMIT = 232448;   // 227 KB

// tanh of two values with packed fp32 arithmetic (FMUL2 / FFMA2 / FADD2 issue once per pair).
//   |y| <  0.55 : y + y t Q(t), t = y^2, Q = degree-4 least-squares fit on Chebyshev nodes of
//                 (tanh(sqrt t)/sqrt t - 1)/t over [0, 0.3025]; max error 0.83 ulp in fp32
//   |y| >= 0.55 : sign(y) (1 - 2 / (exp(2|y|) + 1)) with ex2.approx / rcp.approx: <= 2 ulp
//                 (exp overflow -> rcp(inf) = 0 -> 1; NaN propagates through both branches)
// ~11 issue slots per value against ~19 for tanhf, same accuracy class (tanhf: 2 ulp).
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float2 tanh2(float2 y) {
  const float2 t = __fmul2_rn(y, y);
  float2 q = make_float2(-0.006615746766328812f, -0.006615746766328812f);
  q = __ffma2_rn(q, t, make_float2(0.021312739700078964f, 0.021312739700078964f));
  q = __ffma2_rn(q, t, make_float2(-0.053910065442323685f, -0.053910065442323685f));
  q = __ffma2_rn(q, t, make_float2(0.13333117961883545f, 0.13333117961883545f));
  q = __ffma2_rn(q, t, make_float2(-0.3333333134651184f, -0.3333333134651184f));
  const float2 small = __ffma2_rn(y, __fmul2_rn(t, q), y);
  const float2 ay = make_float2(fabsf(y.x), fabsf(y.y));
  const float2 sc = __fmul2_rn(ay, make_float2(2.8853900817779268f, 2.8853900817779268f));
  const float2 e1 = __fadd2_rn(make_float2(ex2_approx(sc.x), ex2_approx(sc.y)),
                               make_float2(1.f, 1.f));
  const float2 big = __ffma2_rn(make_float2(rcp_approx(e1.x), rcp_approx(e1.y)),
                                make_float2(-2.f, -2.f), make_float2(1.f, 1.f));
  float2 r;
  r.x = ay.x < 0.55f ? small.x : copysignf(big.x, y.x);
  r.y = ay.y < 0.55f ? small.y : copysignf(big.y, y.y);
  return r;
}

// LayerNorm (biased variance, two-pass in registers) + affine + tanh on R warp-resident rows;
// the R reductions are interleaved so that their shuffle latencies overlap.
// torch.nn.LayerNorm semantics: y = (x - mean) / sqrt(var + eps) * weight + bias.
template <int R, int NCH, bool FULL>
__device__ __forceinline__ void ln_tanh_rows(float4 (&v)[R][NCH], int lane, int dim,
                                             const float4* __restrict__ g4,
                                             const float4* __restrict__ b4, float eps) {
  const int nv = dim >> 2;
  const float inv_dim = 1.0f / (float)dim;
  float s[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    float2 s2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < NCH; ++i)
      if (FULL || lane + 32 * i < nv)
        s2 = __fadd2_rn(s2, __fadd2_rn(make_float2(v[r][i].x, v[r][i].y),
                                       make_float2(v[r][i].z, v[r][i].w)));
    s[r] = s2.x + s2.y;
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1)
#pragma unroll
    for (int r = 0; r < R; ++r) s[r] += __shfl_xor_sync(0xffffffffu, s[r], off);
  float q[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const float mean = s[r] * inv_dim;
    const float2 nm = make_float2(-mean, -mean);
    float2 q2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < NCH; ++i)
      if (FULL || lane + 32 * i < nv) {
        const float2 a = __fadd2_rn(make_float2(v[r][i].x, v[r][i].y), nm);
        const float2 b = __fadd2_rn(make_float2(v[r][i].z, v[r][i].w), nm);
        v[r][i] = make_float4(a.x, a.y, b.x, b.y);
        q2 = __ffma2_rn(a, a, q2);
        q2 = __ffma2_rn(b, b, q2);
      }
    q[r] = q2.x + q2.y;
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1)
#pragma unroll
    for (int r = 0; r < R; ++r) q[r] += __shfl_xor_sync(0xffffffffu, q[r], off);
  float rstd[R];
#pragma unroll
  for (int r = 0; r < R; ++r) rstd[r] = rsqrtf(fmaf(q[r], inv_dim, eps));
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + 32 * i;
    if (FULL || c < nv) {
      const float4 g = __ldg(g4 + c), b = __ldg(b4 + c);
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float2 rs = make_float2(rstd[r], rstd[r]);
        const float2 lo = tanh2(__ffma2_rn(__fmul2_rn(make_float2(v[r][i].x, v[r][i].y), rs),
                                           make_float2(g.x, g.y), make_float2(b.x, b.y)));
        const float2 hi = tanh2(__ffma2_rn(__fmul2_rn(make_float2(v[r][i].z, v[r][i].w), rs),
                                           make_float2(g.z, g.w), make_float2(b.z, b.w)));
        v[r][i] = make_float4(lo.x, lo.y, hi.x, hi.y);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// ICM trunk, narrow input (in_dim * (D + 8 R) * 4 bytes of shared memory): persistent blocks,
// W^T resident in shared memory as [in_dim][D/4] float4, each warp computes R rows at a time so
// that one shared-memory read of W feeds R FMAs; the FMAs are packed (FFMA2: two fp32 FMAs per
// issue slot — the fp32 pipe takes one plain FFMA per two cycles and SMSP).  Linear = bias +
// sequential fma over the input columns (fp32, like the reference's addmm up to summation order).
// ------------------------------------------------------------------------------------------
template <int R, int NCH, bool FULL>
__global__ void __launch_bounds__(RP_WARPS * 32, R <= 4 ? 2 : 1)
rep_trunk_kernel(const float* __restrict__ obs, int64_t n, int in_dim,
                 const float* __restrict__ W, const float* __restrict__ bias,
                 const float* __restrict__ ln_w, const float* __restrict__ ln_b, float eps,
                 int dim, float* __restrict__ out_rep, OpOut op) {
  extern __shared__ float4 rp_smem[];
  const int nv = dim >> 2;
  float4* Wt = rp_smem;                                        // [in_dim][nv]
  float* obs_all = reinterpret_cast<float*>(Wt + (size_t)in_dim * nv);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* obs_s = obs_all + (size_t)warp * in_dim * R;          // [in_dim][R]
  {
    float* wt = reinterpret_cast<float*>(Wt);
    if ((in_dim & 3) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0) {
      // coalesced 16-byte reads of the row-major weight, all of a thread's loads in flight,
      // transposed scalar stores (the per-update call is latency-bound on this staging)
      const float4* W4 = reinterpret_cast<const float4*>(W);
      const int q = in_dim >> 2, total4 = q * dim;
      constexpr int SB = 12;   // loads in flight per thread: 24 x 512 weights = ONE round trip
      for (int i0 = threadIdx.x; i0 < total4; i0 += SB * blockDim.x) {
        float4 t[SB];
#pragma unroll
        for (int u = 0; u < SB; ++u) {
          const int i = i0 + u * blockDim.x;
          t[u] = (i < total4) ? __ldg(W4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < SB; ++u) {
          const int i = i0 + u * blockDim.x;
          if (i < total4) {
            const int d = i / q, kk = (i - d * q) * 4;
            wt[(kk + 0) * dim + d] = t[u].x;
            wt[(kk + 1) * dim + d] = t[u].y;
            wt[(kk + 2) * dim + d] = t[u].z;
            wt[(kk + 3) * dim + d] = t[u].w;
          }
        }
      }
    } else {
      const int total = in_dim * dim;
      for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int kk = i / dim, d = i - kk * dim;                // shared-memory-coalesced order
        wt[i] = __ldg(W + (size_t)d * in_dim + kk);
      }
    }
  }
  __syncthreads();
  const float4* b4 = reinterpret_cast<const float4*>(bias);
  const float4* g4 = reinterpret_cast<const float4*>(ln_w);
  const float4* be4 = reinterpret_cast<const float4*>(ln_b);
  const int64_t groups = (n + R - 1) / R;
  // group g -> block g % grid, warp (g / grid) % 8: a small batch spreads over all the SMs
  for (int64_t grp = (int64_t)blockIdx.x + (int64_t)gridDim.x * warp; grp < groups;
       grp += (int64_t)gridDim.x * RP_WARPS) {
    const int64_t row0 = grp * R;
    const int rows = (int)((n - row0) < R ? (n - row0) : R);
    const float* src = obs + row0 * in_dim;
    for (int i = lane; i < R * in_dim; i += 32) {
      const int r = i / in_dim, kk = i - r * in_dim;
      obs_s[kk * R + r] = (r < rows) ? __ldg(src + i) : 0.f;
    }
    __syncwarp();
    float4 acc[R][NCH];
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const int c = lane + 32 * i;
      const float4 bv = (FULL || c < nv) ? __ldg(b4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < R; ++r) acc[r][i] = bv;
    }
#pragma unroll 2
    for (int kk = 0; kk < in_dim; ++kk) {
      float4 w[NCH];
#pragma unroll
      for (int i = 0; i < NCH; ++i) {
        const int c = lane + 32 * i;
        w[i] = (FULL || c < nv) ? Wt[kk * nv + c] : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      float o[R];
      if constexpr (R % 4 == 0) {
#pragma unroll
        for (int r4 = 0; r4 < R / 4; ++r4) {
          const float4 t = reinterpret_cast<const float4*>(obs_s + kk * R)[r4];
          o[4 * r4] = t.x; o[4 * r4 + 1] = t.y; o[4 * r4 + 2] = t.z; o[4 * r4 + 3] = t.w;
        }
      } else {
#pragma unroll
        for (int r = 0; r < R; ++r) o[r] = obs_s[kk * R + r];
      }
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float2 oo = make_float2(o[r], o[r]);
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
          const float2 lo = __ffma2_rn(oo, make_float2(w[i].x, w[i].y),
                                       make_float2(acc[r][i].x, acc[r][i].y));
          const float2 hi = __ffma2_rn(oo, make_float2(w[i].z, w[i].w),
                                       make_float2(acc[r][i].z, acc[r][i].w));
          acc[r][i] = make_float4(lo.x, lo.y, hi.x, hi.y);
        }
      }
    }
    __syncwarp();   // obs_s is restaged by the next group
    ln_tanh_rows<R, NCH, FULL>(acc, lane, dim, g4, be4, eps);
    emit_rows<R, NCH, FULL>(acc, lane, dim, row0, rows, out_rep, op);
  }
}

// Per-update batches (latency-bound): one warp per row, NO shared-memory staging and no block
// barrier — every lane reads the rows of the row-major weight that belong to its 16 output
// columns straight from global memory (16-byte pieces, L1/L2-resident after the first warps) and
// accumulates even and odd input columns in the two halves of an FFMA2.  in_dim % 4 == 0.
template <int NCH, bool FULL>
__global__ void __launch_bounds__(128)
rep_trunk_small_kernel(const float* __restrict__ obs, int64_t n, int in_dim,
                       const float* __restrict__ W, const float* __restrict__ bias,
                       const float* __restrict__ ln_w, const float* __restrict__ ln_b, float eps,
                       int dim, float* __restrict__ out_rep, OpOut op) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (row >= n) return;
  const int nv = dim >> 2, q = in_dim >> 2;
  const float4* o4 = reinterpret_cast<const float4*>(obs + row * in_dim);
  float2 acc[NCH][4];
#pragma unroll
  for (int i = 0; i < NCH; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = make_float2(0.f, 0.f);
#pragma unroll 2
  for (int kc = 0; kc < q; ++kc) {
    const float4 o = __ldg(o4 + kc);
    float4 w[NCH][4];
#pragma unroll
    for (int i = 0; i < NCH; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int d = 4 * (lane + 32 * i) + j;
        w[i][j] = (FULL || d < dim)
                      ? __ldg(reinterpret_cast<const float4*>(W + (size_t)d * in_dim) + kc)
                      : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
    for (int i = 0; i < NCH; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[i][j] = __ffma2_rn(make_float2(o.x, o.y), make_float2(w[i][j].x, w[i][j].y), acc[i][j]);
        acc[i][j] = __ffma2_rn(make_float2(o.z, o.w), make_float2(w[i][j].z, w[i][j].w), acc[i][j]);
      }
  }
  float4 v[1][NCH];
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + 32 * i;
    const float4 bv = (FULL || c < nv) ? __ldg(reinterpret_cast<const float4*>(bias) + c)
                                       : make_float4(0.f, 0.f, 0.f, 0.f);
    v[0][i] = make_float4(bv.x + (acc[i][0].x + acc[i][0].y), bv.y + (acc[i][1].x + acc[i][1].y),
                          bv.z + (acc[i][2].x + acc[i][2].y), bv.w + (acc[i][3].x + acc[i][3].y));
  }
  ln_tanh_rows<1, NCH, FULL>(v, lane, dim, reinterpret_cast<const float4*>(ln_w),
                             reinterpret_cast<const float4*>(ln_b), eps);
  emit_rows<1, NCH, FULL>(v, lane, dim, row, 1, out_rep, op);
}

// LayerNorm -> Tanh on given pre-activations, one warp per row (in place allowed).  The
// pre-activation of a row is bias (optional) + the sum of `nparts` partial rows `part_stride`
// floats apart, added in part order (the split-K partial tiles of trunk_gemm_kernel).
template <int NCH>
__global__ void __launch_bounds__(RP_WARPS * 32)
rep_ln_tanh_kernel(const float* __restrict__ pre, int64_t n, int dim,
                   const float* __restrict__ ln_w, const float* __restrict__ ln_b, float eps,
                   float* out_rep, OpOut op, const float* __restrict__ bias = nullptr,
                   int nparts = 1, int64_t part_stride = 0) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * RP_WARPS + (threadIdx.x >> 5);
  if (row >= n) return;
  const int nv = dim >> 2;
  const float4* x4 = reinterpret_cast<const float4*>(pre + row * dim);
  float4 v[1][NCH];
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + 32 * i;
    v[0][i] = (c < nv) ? x4[c] : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (int s = 1; s < nparts; ++s) {
    const float4* y4 = reinterpret_cast<const float4*>(pre + (int64_t)s * part_stride + row * dim);
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const int c = lane + 32 * i;
      if (c < nv) {
        const float4 y = y4[c];
        v[0][i].x += y.x; v[0][i].y += y.y; v[0][i].z += y.z; v[0][i].w += y.w;
      }
    }
  }
  if (bias != nullptr) {
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const int c = lane + 32 * i;
      if (c < nv) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(bias) + c);
        v[0][i].x += b.x; v[0][i].y += b.y; v[0][i].z += b.z; v[0][i].w += b.w;
      }
    }
  }
  ln_tanh_rows<1, NCH, false>(v, lane, dim, reinterpret_cast<const float4*>(ln_w),
                              reinterpret_cast<const float4*>(ln_b), eps);
  emit_rows<1, NCH, false>(v, lane, dim, row, 1, out_rep, op);
}

// F.normalize(x, dim=1, p=2, eps): x / max(||x||_2, eps), one warp per row (in place allowed).
template <int NCH>
__global__ void __launch_bounds__(RP_WARPS * 32)
rep_normalize_kernel(const float* __restrict__ x, int64_t n, int dim, float eps, float* out_rep,
                     OpOut op) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * RP_WARPS + (threadIdx.x >> 5);
  if (row >= n) return;
  const int nv = dim >> 2;
  const float4* x4 = reinterpret_cast<const float4*>(x + row * dim);
  float4 v[1][NCH];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + 32 * i;
    v[0][i] = (c < nv) ? x4[c] : make_float4(0.f, 0.f, 0.f, 0.f);
    ss = fmaf(v[0][i].x, v[0][i].x, ss); ss = fmaf(v[0][i].y, v[0][i].y, ss);
    ss = fmaf(v[0][i].z, v[0][i].z, ss); ss = fmaf(v[0][i].w, v[0][i].w, ss);
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
  const float den = fmaxf(sqrtf(ss), eps);
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    v[0][i].x = v[0][i].x / den; v[0][i].y = v[0][i].y / den;
    v[0][i].z = v[0][i].z / den; v[0][i].w = v[0][i].w / den;
  }
  emit_rows<1, NCH, false>(v, lane, dim, row, 1, out_rep, op);
}

static int make_op(const exb_operand* o, int dim, OpOut* out) {
  out->f16 = nullptr; out->hn = nullptr; out->er = nullptr; out->sc = nullptr;
  out->dpad = (dim + 63) / 64 * 64;
  if (o == nullptr) return EXB_OK;
  EXB_CHECK_ARG(o->f16 && o->half_norm && o->round_err && o->scalars,
                "exb_operand with a null member");
  EXB_CHECK_ARG(((uintptr_t)o->f16 & 15) == 0, "operand rows must be 16-byte aligned");
  out->f16 = static_cast<__half*>(o->f16);
  out->hn = o->half_norm;
  out->er = o->round_err;
  out->sc = static_cast<PrepScalars*>(o->scalars);
  return EXB_OK;
}

static int check_rep_args(const void* in, int64_t n, int dim, float* out_rep,
                          const exb_operand* op) {
  EXB_CHECK_ARG(in && n > 0, "bad arguments");
  EXB_CHECK_ARG(out_rep || op, "nothing to produce: out_rep and out_op are both null");
  if (dim < 4 || dim > EXB_SCREEN_MAX_DIM || (dim & 3)) {
    set_error("rep producers support rep_dim %% 4 == 0, 4 <= rep_dim <= %d (got %d)",
              EXB_SCREEN_MAX_DIM, dim);
    return EXB_EUNSUPPORTED;
  }
  EXB_CHECK_ARG((((uintptr_t)in | (uintptr_t)out_rep) & 15) == 0, "rows must be 16-byte aligned");
  return EXB_OK;
}

}  // namespace exb

using namespace exb;

extern "C" size_t exb_rep_trunk_smem_bytes(int in_dim, int rep_dim) {
  return (size_t)in_dim * ((size_t)rep_dim + RP_WARPS * 8) * sizeof(float);
}

static int launch_trunk_small(const float* obs, int64_t n, int in_dim, const float* W,
                              const float* b, const float* g, const float* be, float eps, int dim,
                              float* out_rep, const OpOut& op, cudaStream_t st) {
  const int nch = (dim / 4 + 31) / 32;
  const unsigned grid = (unsigned)((n + 3) / 4);
#define EXB_SMALL_CASE(KEY, NCH, FULL)                                                         \
  case KEY:                                                                                    \
    rep_trunk_small_kernel<NCH, FULL><<<grid, 128, 0, st>>>(obs, n, in_dim, W, b, g, be, eps,  \
                                                            dim, out_rep, op);                 \
    break;
  switch (nch * 2 + (dim == nch * 128 ? 1 : 0)) {
    EXB_SMALL_CASE(2, 1, false)
    EXB_SMALL_CASE(3, 1, true)
    EXB_SMALL_CASE(4, 2, false)
    EXB_SMALL_CASE(5, 2, true)
    EXB_SMALL_CASE(6, 3, false)
    EXB_SMALL_CASE(7, 3, true)
    EXB_SMALL_CASE(8, 4, false)
    EXB_SMALL_CASE(9, 4, true)
    default: set_error("rep_dim %d too wide", dim); return EXB_EUNSUPPORTED;
  }
#undef EXB_SMALL_CASE
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

template <int R>
static int launch_trunk(const float* obs, int64_t n, int in_dim, const float* W, const float* b,
                        const float* g, const float* be, float eps, int dim, float* out_rep,
                        const OpOut& op, cudaStream_t st) {
  const size_t smem = (size_t)in_dim * ((size_t)dim + RP_WARPS * R) * sizeof(float);
  const int nch = (dim / 4 + 31) / 32;
  const int64_t groups = (n + R - 1) / R;
  int64_t blocks = groups;
  const int sms = sm_count();
  // persistent: as many blocks per SM as registers / shared memory allow (R = 4: two)
  const int per_sm = (R == 4 && 2 * (smem + 1024) <= (size_t)RP_SMEM_LIMIT) ? 2 : 1;
  if (blocks > (int64_t)sms * per_sm) blocks = (int64_t)sms * per_sm;
#define EXB_TRUNK_CASE(KEY, NCH, FULL)                                                          \
  case KEY: {                                                                                   \
    static bool attr = false;                                                                   \
    if (!attr) {                                                                                \
      EXB_CUDA(cudaFuncSetAttribute(rep_trunk_kernel<R, NCH, FULL>,                             \
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, RP_SMEM_LIMIT)); \
      attr = true;                                                                              \
    }                                                                                           \
    rep_trunk_kernel<R, NCH, FULL><<<(unsigned)blocks, RP_WARPS * 32, smem, st>>>(              \
        obs, n, in_dim, W, b, g, be, eps, dim, out_rep, op);                                    \
  } break;
  const int key = nch * 2 + (dim == nch * 128 ? 1 : 0);
  switch (key) {
    EXB_TRUNK_CASE(2, 1, false)
    EXB_TRUNK_CASE(3, 1, true)
    EXB_TRUNK_CASE(4, 2, false)
    EXB_TRUNK_CASE(5, 2, true)
    EXB_TRUNK_CASE(6, 3, false)
    EXB_TRUNK_CASE(7, 3, true)
    EXB_TRUNK_CASE(8, 4, false)
    EXB_TRUNK_CASE(9, 4, true)
    default: set_error("rep_dim %d too wide", dim); return EXB_EUNSUPPORTED;
  }
#undef EXB_TRUNK_CASE
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

extern "C" int exb_rep_trunk_f32(const float* obs, int64_t n, int in_dim, const float* weight,
                                 const float* bias, const float* ln_weight, const float* ln_bias,
                                 float ln_eps, int rep_dim, float* out_rep,
                                 const exb_operand* out_op, exb_stream_t stream) {
  int rc = check_rep_args(obs, n, rep_dim, out_rep, out_op);
  if (rc) return rc;
  EXB_CHECK_ARG(weight && bias && ln_weight && ln_bias && in_dim > 0, "bad arguments");
  EXB_CHECK_ARG((((uintptr_t)bias | (uintptr_t)ln_weight | (uintptr_t)ln_bias) & 15) == 0,
                "bias / LayerNorm vectors must be 16-byte aligned");
  if (exb_rep_trunk_smem_bytes(in_dim, rep_dim) > (size_t)RP_SMEM_LIMIT) {
    set_error("fused trunk keeps W in shared memory: in_dim %d x rep_dim %d does not fit "
              "(run the Linear as a GEMM and call exb_rep_ln_tanh_f32)", in_dim, rep_dim);
    return EXB_EUNSUPPORTED;
  }
  OpOut op;
  rc = make_op(out_op, rep_dim, &op);
  if (rc) return rc;
  // rows per warp step: 4 (128 registers, two blocks per SM) measured 1.24 ms per 10^6 walker
  // rows against 1.49 ms for 8 (255 registers, one block per SM); EXB_TRUNK_R=8 for A/B
  static const int r4 = [] {
    return knob("EXB_TRUNK_R", 4) != 8;
  }();
  // per-update batches: one row per warp so that a 1024-row batch runs as 1024 short warps over
  // all SMs instead of 256 long ones (16.4 us per 1024 x 24 -> 512 call with R = 4)
  // measured on B200, 1024 x 24 -> 512: the no-staging kernel takes 22.5 us against 14.3 us for
  // the shared-memory kernel with one row per warp (its per-lane weight reads cost 32 sectors per
  // load instruction), so it is an A/B knob only (EXB_TRUNK_SMALL=1)
  static const int small_on = knob("EXB_TRUNK_SMALL", 0);
  if (small_on && n <= 8192 && (in_dim & 3) == 0 &&
      (((uintptr_t)weight | (uintptr_t)obs) & 15) == 0)
    return launch_trunk_small(obs, n, in_dim, weight, bias, ln_weight, ln_bias, ln_eps, rep_dim,
                              out_rep, op, (cudaStream_t)stream);
  if (n <= 8192)
    return launch_trunk<1>(obs, n, in_dim, weight, bias, ln_weight, ln_bias, ln_eps, rep_dim,
                           out_rep, op, (cudaStream_t)stream);
  if (r4)
    return launch_trunk<4>(obs, n, in_dim, weight, bias, ln_weight, ln_bias, ln_eps, rep_dim,
                           out_rep, op, (cudaStream_t)stream);
  return launch_trunk<8>(obs, n, in_dim, weight, bias, ln_weight, ln_bias, ln_eps, rep_dim,
                         out_rep, op, (cudaStream_t)stream);
}

static int launch_ln_tanh(const float* pre, int64_t n, int rep_dim, const float* ln_weight,
                          const float* ln_bias, float ln_eps, float* out_rep, const OpOut& op,
                          const float* bias, int nparts, int64_t part_stride, cudaStream_t st) {
  const unsigned grid = (unsigned)((n + RP_WARPS - 1) / RP_WARPS);
#define EXB_LN_CASE(NCH)                                                                        \
  rep_ln_tanh_kernel<NCH><<<grid, RP_WARPS * 32, 0, st>>>(pre, n, rep_dim, ln_weight, ln_bias,  \
                                                          ln_eps, out_rep, op, bias, nparts,    \
                                                          part_stride)
  switch ((rep_dim / 4 + 31) / 32) {
    case 1: EXB_LN_CASE(1); break;
    case 2: EXB_LN_CASE(2); break;
    case 3: EXB_LN_CASE(3); break;
    default: EXB_LN_CASE(4); break;
  }
#undef EXB_LN_CASE
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}

extern "C" int exb_rep_ln_tanh_f32(const float* pre, int64_t n, int rep_dim,
                                   const float* ln_weight, const float* ln_bias, float ln_eps,
                                   float* out_rep, const exb_operand* out_op,
                                   exb_stream_t stream) {
  int rc = check_rep_args(pre, n, rep_dim, out_rep, out_op);
  if (rc) return rc;
  EXB_CHECK_ARG(ln_weight && ln_bias, "bad arguments");
  EXB_CHECK_ARG((((uintptr_t)ln_weight | (uintptr_t)ln_bias) & 15) == 0,
                "LayerNorm vectors must be 16-byte aligned");
  OpOut op;
  rc = make_op(out_op, rep_dim, &op);
  if (rc) return rc;
  return launch_ln_tanh(pre, n, rep_dim, ln_weight, ln_bias, ln_eps, out_rep, op, nullptr, 1, 0,
                        (cudaStream_t)stream);
}

// ---- wide trunk: tcgen05 3xTF32 GEMM (rep_gemm.cuh) + split sum / bias / LayerNorm / Tanh / emit
struct WidePlan {
  int m_tiles, n_tiles, chunks, splits, chunks_per_split;
  size_t wlo_floats, part_floats;
};

static bool wide_plan(int64_t n, int in_dim, int rep_dim, WidePlan* pl) {
  if (n < 1 || in_dim < 4 || (in_dim & 3) || rep_dim < 4 || rep_dim > EXB_SCREEN_MAX_DIM ||
      (rep_dim & 3) || n > (1LL << 31) - TG_M)
    return false;
  pl->m_tiles = (int)((n + TG_M - 1) / TG_M);
  pl->n_tiles = (rep_dim + TG_N - 1) / TG_N;
  pl->chunks = (in_dim + TG_KC - 1) / TG_KC;
  // split K until one wave of CTAs (one per SM: 193 KB of shared memory each) is full, with at
  // least two accumulation windows per split
  int64_t tiles = (int64_t)pl->m_tiles * pl->n_tiles;
  int splits = (int)(sm_count() / tiles);
  const int max_splits = pl->chunks / (2 * TG_WIN_CHUNKS);
  if (splits > max_splits) splits = max_splits;
  if (splits > 32) splits = 32;
  if (splits < 1) splits = 1;
  pl->chunks_per_split = (pl->chunks + splits - 1) / splits;
  pl->splits = (pl->chunks + pl->chunks_per_split - 1) / pl->chunks_per_split;
  pl->wlo_floats = (size_t)rep_dim * in_dim;
  pl->part_floats = (size_t)pl->splits * n * rep_dim;
  return true;
}

extern "C" size_t exb_rep_trunk_wide_workspace_bytes(int64_t n, int in_dim, int rep_dim) {
  WidePlan pl;
  if (!wide_plan(n, in_dim, rep_dim, &pl)) return 0;
  return align_up(pl.wlo_floats * 4, 256) + align_up(pl.part_floats * 4, 256) + 256;
}

extern "C" int exb_rep_trunk_wide_f32(const float* obs, int64_t n, int in_dim,
                                      const float* weight, const float* bias,
                                      const float* ln_weight, const float* ln_bias, float ln_eps,
                                      int rep_dim, float* out_rep, const exb_operand* out_op,
                                      void* workspace, size_t workspace_bytes,
                                      exb_stream_t stream) {
  int rc = check_rep_args(obs, n, rep_dim, out_rep, out_op);
  if (rc) return rc;
  EXB_CHECK_ARG(weight && bias && ln_weight && ln_bias && workspace, "bad arguments");
  EXB_CHECK_ARG((((uintptr_t)weight | (uintptr_t)bias | (uintptr_t)ln_weight |
                  (uintptr_t)ln_bias | (uintptr_t)workspace) & 15) == 0,
                "weight / bias / LayerNorm vectors / workspace must be 16-byte aligned");
  WidePlan pl;
  if (!wide_plan(n, in_dim, rep_dim, &pl)) {
    set_error("wide trunk needs in_dim %% 4 == 0 (TMA row pitch), got in_dim=%d rep_dim=%d", in_dim,
              rep_dim);
    return EXB_EUNSUPPORTED;
  }
  EXB_CHECK_ARG(workspace_bytes >= exb_rep_trunk_wide_workspace_bytes(n, in_dim, rep_dim),
                "workspace too small");
  OpOut op;
  rc = make_op(out_op, rep_dim, &op);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  Carver cv(workspace);
  float* wlo = cv.take<float>(pl.wlo_floats);
  float* part = cv.take<float>(pl.part_floats);
  const int64_t n4 = (int64_t)pl.wlo_floats / 4;
  tf32_lo_kernel<<<(unsigned)std::min<int64_t>((n4 + 255) / 256, 148 * 16), 256, 0, st>>>(
      reinterpret_cast<const float4*>(weight), n4, reinterpret_cast<float4*>(wlo));
  EXB_LAUNCH_CHECK();
  CUtensorMap t_obs, t_w, t_wlo;
  rc = tmap_encode_2d(&t_obs, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, obs, n, in_dim,
                      (int64_t)in_dim * 4, TG_KC, TG_M);
  if (rc) return rc;
  rc = tmap_encode_2d(&t_w, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, weight, rep_dim, in_dim,
                      (int64_t)in_dim * 4, TG_KC, TG_N);
  if (rc) return rc;
  rc = tmap_encode_2d(&t_wlo, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, wlo, rep_dim, in_dim,
                      (int64_t)in_dim * 4, TG_KC, TG_N);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    EXB_CUDA(cudaFuncSetAttribute(trunk_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  TG_SMEM_BYTES));
    attr_set = true;
  }
  TrunkGemmParams gp;
  gp.part = part; gp.n = n; gp.dim = rep_dim; gp.chunks = pl.chunks;
  gp.chunks_per_split = pl.chunks_per_split;
  const dim3 grid((unsigned)pl.m_tiles, (unsigned)pl.n_tiles, (unsigned)pl.splits);
  trunk_gemm_kernel<<<grid, TG_THREADS, TG_SMEM_BYTES, st>>>(t_obs, t_w, t_wlo, gp);
  EXB_LAUNCH_CHECK();
  return launch_ln_tanh(part, n, rep_dim, ln_weight, ln_bias, ln_eps, out_rep, op, bias, pl.splits,
                        (int64_t)n * rep_dim, st);
}

extern "C" int exb_rep_normalize_f32(const float* x, int64_t n, int dim, float eps,
                                     float* out_rep, const exb_operand* out_op,
                                     exb_stream_t stream) {
  int rc = check_rep_args(x, n, dim, out_rep, out_op);
  if (rc) return rc;
  OpOut op;
  rc = make_op(out_op, dim, &op);
  if (rc) return rc;
  const unsigned grid = (unsigned)((n + RP_WARPS - 1) / RP_WARPS);
  cudaStream_t st = (cudaStream_t)stream;
  switch ((dim / 4 + 31) / 32) {
    case 1: rep_normalize_kernel<1><<<grid, RP_WARPS * 32, 0, st>>>(x, n, dim, eps, out_rep, op); break;
    case 2: rep_normalize_kernel<2><<<grid, RP_WARPS * 32, 0, st>>>(x, n, dim, eps, out_rep, op); break;
    case 3: rep_normalize_kernel<3><<<grid, RP_WARPS * 32, 0, st>>>(x, n, dim, eps, out_rep, op); break;
    default: rep_normalize_kernel<4><<<grid, RP_WARPS * 32, 0, st>>>(x, n, dim, eps, out_rep, op); break;
  }
  EXB_LAUNCH_CHECK();
  return EXB_OK;
}
