// Shared helpers for the b200cnn CUDA sources (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200cnn.h"

namespace b200 {

void set_error(const char* fmt, ...);

#define B200_CUDA_TRY(expr)                                                              \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess) {                                                            \
      ::b200::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); \
      return B200_E_CUDA;                                                                \
    }                                                                                    \
  } while (0)

#define B200_LAUNCH_CHECK() B200_CUDA_TRY(cudaGetLastError())

#define B200_REQUIRE(cond, ...)                 \
  do {                                          \
    if (!(cond)) {                              \
      ::b200::set_error(__VA_ARGS__);           \
      return B200_E_INVALID;                    \
    }                                           \
  } while (0)

#define B200_TRY(expr) do { int rc__ = (expr); if (rc__) return rc__; } while (0)

static inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }
static inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// ---- programmatic dependent launch ----------------------------------------------------------------------------
// Kernels of the bf16 step call pdl_launch_dependents() at their start: a consumer launched right behind them with
// programmatic stream serialization may then become resident and run its input-independent prologue (barrier /
// TMEM set-up, shared-memory clears, weight fetches) while the producer still executes.  Every thread of such a
// consumer calls pdl_wait() before it reads the producer's output or writes global memory; the wait returns when
// the whole producer grid has completed.  Both instructions do nothing in an ordinary launch.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif
#ifdef __CUDACC__
// one warp: sum over z < S of part[z * stride + i .. i+3], lane l taking z = l, l + 32, ... (all of its loads in flight at
// once) and a five-step shuffle tree -- THE summation order of conv1's gradient partials, shared by the update kernels
// (sgd_reduce_kernel, p2p_allreduce_sgd_kernel) and the stand-alone reduction, so that every route gives the same bits
__device__ __forceinline__ float4 warp_sum_partials(const float* __restrict__ part, int S, int stride, int i, int lane) {
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int z0 = lane; z0 < S; z0 += 128) {
    float4 v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int z = z0 + 32 * j;
      v[j] = z < S ? __ldg(reinterpret_cast<const float4*>(part + (int64_t)z * stride + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) { s.x += v[j].x; s.y += v[j].y; s.z += v[j].z; s.w += v[j].w; }
  }
#pragma unroll
  for (int m = 1; m <= 16; m <<= 1) {
    s.x += __shfl_xor_sync(0xffffffffu, s.x, m); s.y += __shfl_xor_sync(0xffffffffu, s.y, m);
    s.z += __shfl_xor_sync(0xffffffffu, s.z, m); s.w += __shfl_xor_sync(0xffffffffu, s.w, m);
  }
  return s;
}
#endif
// Small kernels that share the machine with the tensor-core kernels (side streams) ask for the maximum shared-memory
// carve-out even though they need little: an SM whose L1/shared split was set for a small-footprint kernel cannot
// take a 100-200 KB CTA until it has drained, which delayed FC1's dW behind the tail-gradient kernel by 3 us.
#ifdef __CUDACC__
template <typename K>
inline cudaError_t prefer_max_shared(K kernel) {
  return cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
}
#endif
// Function attributes (dynamic shared-memory limit, carve-out preference) are per DEVICE: a `static bool` guard would
// set them on the first GPU a process touches only.  PerDeviceOnce::first() is true once per device ordinal (races
// between host threads at worst set an attribute twice, which is harmless).
struct PerDeviceOnce {
  bool done[64] = {};
  bool first() {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= 64) return true;
    if (done[d]) return false;
    done[d] = true;
    return true;
  }
};
// The next launch through launch_kernel() on this host thread is made with programmatic stream serialization when
// this flag is set (the step driver in abi.cu sets it in front of the launches it wants overlapped); it resets itself.
extern thread_local bool g_pdl_next;
#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                 Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = g_pdl_next ? 1 : 0;
  g_pdl_next = false;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// ---- kernel-schedule trace (diagnostics build only: B200_TRACE=1 python build.py) -----------------------------
// Every instrumented kernel stamps %globaltimer when its first CTA starts and when its last CTA ends into
// buf[2*id], buf[2*id+1]; b200_debug_trace() hands the buffer to every translation unit.  Compiled out otherwise.
enum TraceId { TR_PACK = 0, TR_W2IMG, TR_CONV1_FWD, TR_CONV2_FWD, TR_FC1_FWD, TR_TAIL_ROWS, TR_TAIL_GRADS, TR_FC1_DW,
               TR_FC1_DX, TR_CONV2_DGRAD, TR_CONV2_WGRAD, TR_CONV2_REDUCE, TR_CONV1_WGRAD, TR_SGD, TR_P2P_FC, TR_P2P_CONV, TR_COUNT };
#ifdef B200_TRACE
static __device__ unsigned long long* b200_trace_buf;
struct TraceScope {
  int id;
  static __device__ __forceinline__ unsigned long long now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
  }
  // buf[2*TR_COUNT] = id of one kernel whose CTAs also log their own start / end / SM at buf[64 + 3*cta ...]
  __device__ __forceinline__ explicit TraceScope(int i) : id(i) {
    if (threadIdx.x == 0 && b200_trace_buf) {
      const unsigned long long t = now();
      atomicMin(b200_trace_buf + 2 * id, t);
      for (int w = 0; w < 2; ++w)       // up to two kernels log per CTA, in regions of 3 * 512 words
        if (b200_trace_buf[2 * TR_COUNT + w] == (unsigned long long)id) {
          const int cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
          unsigned sm;
          asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
          if (cta < 512) { b200_trace_buf[64 + 1536 * w + 3 * cta] = t; b200_trace_buf[64 + 1536 * w + 3 * cta + 2] = sm; }
        }
    }
  }
  __device__ __forceinline__ ~TraceScope() {
    if (threadIdx.x == 0 && b200_trace_buf) {
      const unsigned long long t = now();
      atomicMax(b200_trace_buf + 2 * id + 1, t);
      for (int w = 0; w < 2; ++w)
        if (b200_trace_buf[2 * TR_COUNT + w] == (unsigned long long)id) {
          const int cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
          if (cta < 512) b200_trace_buf[64 + 1536 * w + 3 * cta + 1] = t;
        }
    }
  }
};
#define B200_TRACE_SCOPE(id) ::b200::TraceScope trace_scope__(id)
#define B200_TRACE_SETTER(name) \
  void name(unsigned long long* p) { cudaMemcpyToSymbol(b200_trace_buf, &p, sizeof(p)); }
#else
#define B200_TRACE_SCOPE(id) do { } while (0)
#define B200_TRACE_SETTER(name) \
  void name(unsigned long long*) {}
#endif

constexpr int kNumSMs = 148;   // B200: 2 dies x 74 SMs; grids below are sized in multiples of it

}  // namespace b200
