// Error plumbing + device info for the C ABI (include/exorl_b200.h).
#include "common.cuh"
#include <mutex>
#include <string.h>

namespace exb {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
namespace {
struct KnobEntry { char name[48]; int value; };
KnobEntry g_knobs[32];
int g_n_knobs = 0;
std::mutex g_knob_mu;
}  // namespace
bool knob_lookup(const char* name, int* value) {
  std::lock_guard<std::mutex> lock(g_knob_mu);
  for (int i = 0; i < g_n_knobs; ++i)
    if (strcmp(g_knobs[i].name, name) == 0) { *value = g_knobs[i].value; return true; }
  return false;
}
int sm_count() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    else
      return 148;
  }
  return cached;
}
}  // namespace exb

extern "C" int exb_abi_version(void) { return 2; }
extern "C" const char* exb_last_error(void) { return exb::g_err; }
extern "C" int exb_device_info(int* sms, int* cc) {
  int dev = 0, n = 0, maj = 0, min = 0;
  EXB_CUDA(cudaGetDevice(&dev));
  EXB_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
  EXB_CUDA(cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, dev));
  EXB_CUDA(cudaDeviceGetAttribute(&min, cudaDevAttrComputeCapabilityMinor, dev));
  if (sms) *sms = n;
  if (cc) *cc = maj * 10 + min;
  return EXB_OK;
}

extern "C" int exb_debug_knob_set(const char* name, int value) {
  EXB_CHECK_ARG(name && strlen(name) < 48, "bad knob name");
  std::lock_guard<std::mutex> lock(exb::g_knob_mu);
  for (int i = 0; i < exb::g_n_knobs; ++i)
    if (strcmp(exb::g_knobs[i].name, name) == 0) { exb::g_knobs[i].value = value; return EXB_OK; }
  EXB_CHECK_ARG(exb::g_n_knobs < 32, "knob table full");
  strcpy(exb::g_knobs[exb::g_n_knobs].name, name);
  exb::g_knobs[exb::g_n_knobs++].value = value;
  return EXB_OK;
}

// ------------------------------------------------------------------------------------------
// Peer-visible device buffers (multi-GPU symmetric screen, knn_screen.cu): plain cudaMalloc
// memory exported through CUDA IPC, so that the screen kernel of ANOTHER rank's process can
// append the hits it finds for this rank's rows straight into this rank's memory over NVLink.
// (Memory of torch's caching allocator cannot be used: its IPC export maps the whole allocator
// block and kernel stores through that mapping faulted on the B200 boxes.)
// ------------------------------------------------------------------------------------------
extern "C" int exb_peer_buffer_alloc(size_t bytes, void** ptr, unsigned char* handle64) {
  EXB_CHECK_ARG(bytes > 0 && ptr && handle64, "bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  EXB_CUDA(cudaMalloc(&p, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    exb::set_error("cudaIpcGetMemHandle -> %s", cudaGetErrorString(e));
    return EXB_ECUDA;
  }
  memcpy(handle64, &h, 64);
  *ptr = p;
  return EXB_OK;
}

extern "C" int exb_peer_buffer_open(const unsigned char* handle64, void** ptr) {
  EXB_CHECK_ARG(handle64 && ptr, "bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  EXB_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr = p;
  return EXB_OK;
}

extern "C" int exb_peer_buffer_close(void* ptr) {
  if (ptr) EXB_CUDA(cudaIpcCloseMemHandle(ptr));
  return EXB_OK;
}

extern "C" int exb_peer_buffer_free(void* ptr) {
  if (ptr) EXB_CUDA(cudaFree(ptr));
  return EXB_OK;
}

extern "C" int exb_peer_buffer_clear(void* ptr, size_t bytes, exb_stream_t stream) {
  EXB_CHECK_ARG(ptr && bytes > 0, "bad arguments");
  EXB_CUDA(cudaMemsetAsync(ptr, 0, bytes, (cudaStream_t)stream));
  return EXB_OK;
}

// Strided rows of a (pinned) host matrix into a dense device matrix: one 2-D DMA on `stream`
// (the sample of the symmetric screen = every stride-th rep row, uploaded ahead of the rows).
extern "C" int exb_h2d_rows_2d(void* dst, const void* src_host, int64_t rows, int64_t row_bytes,
                               int64_t src_pitch_bytes, exb_stream_t stream) {
  EXB_CHECK_ARG(dst && src_host && rows >= 0 && row_bytes > 0 && src_pitch_bytes >= row_bytes,
                "bad arguments");
  if (rows == 0) return EXB_OK;
  EXB_CUDA(cudaMemcpy2DAsync(dst, (size_t)row_bytes, src_host, (size_t)src_pitch_bytes,
                             (size_t)row_bytes, (size_t)rows, cudaMemcpyHostToDevice,
                             (cudaStream_t)stream));
  return EXB_OK;
}
