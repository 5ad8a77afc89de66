// Shared helpers for the exorl_b200 CUDA sources (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <stdlib.h>
#include "../../include/exorl_b200.h"

namespace exb {

void set_error(const char* fmt, ...);

#define EXB_CHECK_ARG(cond, ...)            \
  do {                                      \
    if (!(cond)) {                          \
      exb::set_error(__VA_ARGS__);          \
      return EXB_EINVAL;                    \
    }                                       \
  } while (0)

#define EXB_CUDA(call)                                                              \
  do {                                                                              \
    cudaError_t e_ = (call);                                                        \
    if (e_ != cudaSuccess) {                                                        \
      exb::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call,                   \
                     cudaGetErrorString(e_));                                       \
      return EXB_ECUDA;                                                             \
    }                                                                               \
  } while (0)

#define EXB_LAUNCH_CHECK() EXB_CUDA(cudaGetLastError())

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Bump allocator over a caller-provided workspace.
struct Carver {
  char* base;
  size_t off;
  explicit Carver(void* p) : base(static_cast<char*>(p)), off(0) {}
  template <typename T>
  T* take(size_t count) {
    off = align_up(off, 256);
    T* r = reinterpret_cast<T*>(base + off);
    off += count * sizeof(T);
    return r;
  }
};

int sm_count();

// A/B switches of the bring-up work (EXB_SCREEN_*, EXB_PDL, EXB_TRUNK_*, EXB_GATHER_BULK,
// EXB_CHAIN_*).  The product library does not read the environment: knob() returns the value a
// measurement script set through exb_debug_knob_set() before the first use of the switch (values
// are latched at first use), else the default.  A library built with -DEXB_DEBUG_KNOBS
// (EXB_EXTRA_NVCC_FLAGS=-DEXB_DEBUG_KNOBS python -m exorl_b200._build --force) also reads the
// environment variable of the same name.
bool knob_lookup(const char* name, int* value);
static inline int knob(const char* name, int dflt) {
  int v = 0;
  if (knob_lookup(name, &v)) return v;
#ifdef EXB_DEBUG_KNOBS
  const char* e = getenv(name);
  if (e) return atoi(e);
#endif
  return dflt;
}

// Scalars of a prepared fp16 kNN operand (exb_knn_prep_f16 and the rep producers), max-accumulated
// over the rows of a call; device struct, zeroed by the caller.
struct PrepScalars {
  float err_max;       // max_j ||x_j - fp16(x_j)||
  float hn_max;        // max_j ||fp16(x_j)||^2 / 2
  int range_flag;      // a value does not fit fp16 / is not finite
  int pad;
};

// Programmatic dependent launch (the per-update chain is launch-latency bound): a kernel launched
// with cudaLaunchAttributeProgrammaticStreamSerialization may be scheduled while its predecessor
// still runs; it must execute pdl_wait() — in EVERY thread block, before any early return, before
// touching memory the predecessor reads or writes — which blocks until the predecessor grid has
// completed and its writes are visible.  pdl_launch_dependents() in the predecessor lets the
// dependent's blocks be scheduled early (they then sit in pdl_wait with their prologue done).
// Both are no-ops for ordinary launches.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// (value, index) ordering used everywhere: smaller value first, then smaller index.
__device__ __forceinline__ bool pair_less(float va, int ia, float vb, int ib) {
  return va < vb || (va == vb && ia < ib);
}

}  // namespace exb
