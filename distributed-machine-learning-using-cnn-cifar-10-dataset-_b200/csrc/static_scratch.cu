// b200_fc_bwd_f32_static: the fully-connected backward of b200_fc_bwd_f32 for callers that cannot
// hand the library a scratch buffer because they cannot allocate one.
//
// Who that is: a plug-in inside a training process whose host framework has already handed every
// free byte of HBM to its own caching pool (a cudaMalloc from the plug-in then fails, or forces the
// pool to flush), an operator that is only ever given tensors by its host application and has no
// allocator handle of its own, and latency-bound step loops that forbid cudaMalloc / cudaFree
// because those calls synchronise the whole device.  For them the scratch of the bias-gradient
// column sums lives in a fixed array inside the library image instead.
//
// Cost and limits: every GPU that loads libb200cnn.so carries B200_STATIC_SCRATCH_BYTES (1 MiB) of
// device memory until the process exits.  Requests whose scratch need exceeds it are refused with
// B200_E_INVALID.  All calls on a device share the one array and it keeps whatever the previous
// call left in it: the column-sum kernels overwrite every partial they later read, so nothing has
// to start at zero and nothing is cleared; but two calls that overlap on the GPU (different
// streams, or graph branches) race on it -- serialise them.
#include <mutex>

#include "launch.h"

// File-scope device array.  In HOST code this name is only a handle for the runtime's symbol
// table, not a device address: it must never be passed to a kernel.  The device address is
// fetched with cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_static_scratch[B200_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_mu;
void* g_scratch_addr[kMaxDevices] = {nullptr};

// once per device, with that device current: resolve the array's device address
int static_scratch_ptr(float** out) {
  int dev = -1;
  B200_CUDA_TRY(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < kMaxDevices, "fc_bwd_f32_static: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    B200_CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
    B200_REQUIRE(p != nullptr, "fc_bwd_f32_static: cudaGetSymbolAddress returned a null address");
    g_scratch_addr[dev] = p;
  }
  *out = reinterpret_cast<float*>(g_scratch_addr[dev]);
  return B200_OK;
}
}  // namespace

extern "C" {

// Same arithmetic, arguments and results as b200_fc_bwd_f32 minus the scratch pair.  Allocates
// nothing (no cudaMalloc / cudaFree / allocating library call).  The FIRST call on each device
// looks up the scratch array's address (cudaGetSymbolAddress), so make that first call -- with
// the device current -- before starting any stream capture.
int b200_fc_bwd_f32_static(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db,
                           int32_t batch, int32_t k, int32_t n, int32_t mask_x, void* stream) {
  B200_REQUIRE(x && w && dy && dw, "null pointer");
  B200_REQUIRE(batch > 0 && k > 0 && n > 0, "bad shape");
  const int64_t need = b200::colsum_scratch_bytes(batch, n);
  B200_REQUIRE(need <= B200_STATIC_SCRATCH_BYTES,
               "fc_bwd_f32_static: batch %d x n %d needs %lld bytes of scratch, the built-in array holds %d; "
               "use b200_fc_bwd_f32 with caller scratch",
               batch, n, (long long)need, B200_STATIC_SCRATCH_BYTES);
  float* scratch = nullptr;
  int rc = static_scratch_ptr(&scratch);
  if (rc) return rc;
  return b200::fc_bwd_f32(x, w, dy, dx, dw, db, scratch, batch, k, n, mask_x, (cudaStream_t)stream);
}

}  // extern "C"
