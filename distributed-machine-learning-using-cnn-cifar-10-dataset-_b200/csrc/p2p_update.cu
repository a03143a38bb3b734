// The parameter-server push/apply of cifar10cnn.py (:162-163 executed on /job:ps, :195-196) as ONE kernel per
// GPU over NVLink peer memory: all-reduce of the flat gradient buffer fused with the SGD update.
//
// Every rank owns the shard [rank*S, (rank+1)*S) of the flat parameter vector.  Its kernel
//   A. waits until every peer has published "my gradients of this step are complete",
//   1. sums the shard's gradients over all ranks   -- direct NVLink loads from the peers' buffers, or one
//      multimem.ld_reduce per 16 bytes when the buffers have an NVSwitch multicast mapping (the switch adds),
//   2. applies  w -= lr * scale * sum  to its shard (fp32 master weights),
//   3. writes the updated shard into EVERY rank's parameter buffer (peer stores, or one multimem.st),
//   B. publishes "done" and waits for every peer's "done": after that all gradients may be overwritten and
//      all parameter buffers are complete and identical.
// Transfers per rank: 2*(N-1)/N of 4.27 MB (NVLS: 2/N), two flag round trips; no separate update launch, no
// NCCL call.  Flags live in each rank's signal pad (symmetric memory); the step number is a device-side
// epoch counter, so the launch is identical every step and can sit in a CUDA graph.
//
// Buffers (gradients, parameters, pads) must be peer-mapped symmetric allocations of the same size on all
// ranks; torch.distributed._symmetric_memory provides them (plumbing), see exchange.P2PExchange.
#include <stdlib.h>

#include <algorithm>

#include "launch.h"

namespace b200 {
namespace {
constexpr int kMaxRanks = 8;

struct P2PArgs {
  const float* grads[kMaxRanks];    // every rank's flat gradient buffer (index = rank)
  float* params[kMaxRanks];         // every rank's flat parameter buffer
  uint32_t* pads[kMaxRanks];        // every rank's flag pad (of this channel): [0,8) "ready" slots, [8,16) "done" slots
  const float* mc_grads;            // multicast mapping of the gradient buffers, or null
  float* mc_params;                 // multicast mapping of the parameter buffers, or null
  uint32_t* state;                  // local: [0] epoch, [1] CTA arrival counter, [2] CTAs done with the local reduction
  // optional local reduction in front of the exchange: out[i] = sum over z < red_S of red_part[z][i] for the first red_n
  // floats of this rank's gradient range (conv1's per-CTA partials), done by the first red_ctas CTAs
  const float* red_part; float* red_out; int red_S, red_n, red_ctas;
  int rank, world;
  int64_t total;                    // floats, divisible by 4 * world
  float step;                       // lr * scale
  long long timeout_cycles;         // bound of one flag wait (see flag_timeout_cycles)
  int trace_id;                     // diagnostics build: TR_P2P_FC / TR_P2P_CONV
};

__device__ __forceinline__ void st_flag(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_flag(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// bounded: a lost peer must fault this launch, not hang the GPU.  The bound has to outlast anything that may
// legitimately hold a peer back on the host between two steps (a checkpoint write, an evaluation pass, a lazy build,
// a slow batch): B200_P2P_TIMEOUT_S seconds, default 120 (include/b200cnn.h).  Past it every waiting rank traps,
// which leaves a sticky error on its context -- the job is lost either way, but the box is not.
__device__ __forceinline__ void wait_flag(const uint32_t* p, uint32_t epoch, long long timeout_cycles) {
  const long long t0 = clock64();
  while ((int32_t)(ld_flag(p) - epoch) < 0) {
    if (clock64() - t0 > timeout_cycles) {
      printf("b200cnn: peer flag wait timed out (block %d thread %d)\n", (int)blockIdx.x, (int)threadIdx.x);
      __trap();
    }
  }
}
__device__ __forceinline__ float4 mc_ld_reduce(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void mc_st(float* p, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

__global__ void __launch_bounds__(512) p2p_allreduce_sgd_kernel(P2PArgs a) {
  B200_TRACE_SCOPE(a.trace_id);
  const int tid = threadIdx.x;
  const uint32_t epoch = a.state[0] + 1;          // read by everyone before the last CTA bumps it at the very end
  // ---- 0: finish this rank's own gradients (fixed-order sum of conv1's partials): saves a kernel launch at the very
  //         end of the step.  CTA 0 publishes "ready" only once every reducing CTA has stored its share.
  if (a.red_ctas > 0) {
    if ((int)blockIdx.x < a.red_ctas) {
      const int t = blockIdx.x * blockDim.x + tid;
      const int i = min(t >> 5, a.red_n / 4 - 1) * 4;
      const float4 s = warp_sum_partials(a.red_part, a.red_S, a.red_n, i, t & 31);
      if ((t & 31) == 0 && (t >> 5) < a.red_n / 4) *reinterpret_cast<float4*>(a.red_out + i) = s;
      __syncthreads();
      if (tid == 0) { __threadfence(); atomicAdd(a.state + 2, 1u); }
    }
    if (blockIdx.x == 0) {
      if (tid == 0) {
        const long long t0 = clock64();
        while (atomicAdd(a.state + 2, 0u) < (uint32_t)a.red_ctas)
          if (clock64() - t0 > a.timeout_cycles) { printf("b200cnn: local reduction wait timed out\n"); __trap(); }
      }
      __syncthreads();
    }
  }
  // ---- A: gradients of all ranks complete ------------------------------------------------------------
  if (blockIdx.x == 0 && tid < a.world) {
    __threadfence_system();
    st_flag(a.pads[tid] + a.rank, epoch);         // "rank `a.rank` is ready", written into peer tid's pad
  }
  if (tid < a.world) wait_flag(a.pads[a.rank] + tid, epoch, a.timeout_cycles);
  __syncthreads();
  // ---- 1-3: reduce my shard, update, broadcast -----------------------------------------------------------
  const int64_t shard4 = a.total / 4 / a.world;
  const int64_t beg4 = (int64_t)a.rank * shard4;
  const bool nvls = a.mc_grads != nullptr && a.mc_params != nullptr;
  if (nvls) {
    // four switch reductions per thread in flight (a round trip through the switch is ~2-3 us; with few CTAs --
    // the channel that overlaps the backward pass is capped at 16 -- latency, not bandwidth, is the limit)
    constexpr int U = 4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = blockIdx.x * (int64_t)blockDim.x + tid; i0 < shard4; i0 += U * stride) {
      float4 g[U], w[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + u * stride;
        if (i < shard4) {
          g[u] = mc_ld_reduce(a.mc_grads + (beg4 + i) * 4);
          w[u] = *reinterpret_cast<const float4*>(a.params[a.rank] + (beg4 + i) * 4);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + u * stride;
        if (i < shard4) {
          w[u].x = fmaf(-a.step, g[u].x, w[u].x); w[u].y = fmaf(-a.step, g[u].y, w[u].y);
          w[u].z = fmaf(-a.step, g[u].z, w[u].z); w[u].w = fmaf(-a.step, g[u].w, w[u].w);
          mc_st(a.mc_params + (beg4 + i) * 4, w[u]);
        }
      }
    }
  } else
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + tid; i < shard4; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t o = (beg4 + i) * 4;
    float4 g;
    {
      float4 v[kMaxRanks];
#pragma unroll
      for (int p = 0; p < kMaxRanks; ++p)
        if (p < a.world) v[p] = *reinterpret_cast<const float4*>(a.grads[p] + o);     // all peer loads in flight
      g = v[0];
#pragma unroll
      for (int p = 1; p < kMaxRanks; ++p)
        if (p < a.world) { g.x += v[p].x; g.y += v[p].y; g.z += v[p].z; g.w += v[p].w; }   // rank order: same on all ranks
    }
    float4 w = *reinterpret_cast<const float4*>(a.params[a.rank] + o);
    w.x = fmaf(-a.step, g.x, w.x); w.y = fmaf(-a.step, g.y, w.y);
    w.z = fmaf(-a.step, g.z, w.z); w.w = fmaf(-a.step, g.w, w.w);
#pragma unroll
    for (int p = 0; p < kMaxRanks; ++p)
      if (p < a.world) *reinterpret_cast<float4*>(a.params[p] + o) = w;
  }
  // ---- B: everybody done ------------------------------------------------------------------------------------
  // one system-scope fence per CTA, by the thread that then counts the CTA in: the barrier makes the CTA's writes
  // visible to that thread, its fence (cumulative) publishes them to the peers before the count / the flags.
  // (512 fences per CTA, one per thread, were several microseconds of this kernel.)
  __syncthreads();
  __shared__ bool last;
  if (tid == 0) {
    __threadfence_system();
    last = atomicAdd(a.state + 1, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last) {
    if (tid < a.world) {
      __threadfence_system();
      st_flag(a.pads[tid] + 8 + a.rank, epoch);
      wait_flag(a.pads[a.rank] + 8 + tid, epoch, a.timeout_cycles);
    }
    __syncthreads();
    if (tid == 0) { a.state[1] = 0; a.state[2] = 0; a.state[0] = epoch; }
  }
}
// ---------------------------------------------------------------------------------------------------------------------
// One-shot variant for a SMALL range that is exposed at the end of the step (the conv gradients, 10 % of the buffer):
// every rank PUSHES its gradients into an inbox slot on every rank, tells them so, waits for the others' pushes and
// sums the slots of its own inbox in rank order (identical on all ranks) into its own parameters.  One flag wait instead
// of two barriers and no remote reads or parameter writes: a rank is done as soon as the last peer's push has landed.
//   inbox: [2 parities][world slots][n floats] of symmetric memory per rank.  A slot of parity p is rewritten two steps
//   later, which a peer can only do after this rank has published its flag of the step in between, i.e. after it has
//   finished reading this one -- so two parities make a trailing "done" barrier unnecessary.
// The first red_ctas CTAs compute the fixed-order sums of conv1's per-CTA partials for the head of the range themselves
// (and store them to the local gradient buffer), so nothing waits for a reduction launch.
struct PushArgs {
  float* grads;                     // local flat gradient buffer (range start)
  float* params;                    // local flat parameter buffer (range start)
  const float* inbox_local;         // this rank's inbox
  float* inbox_mc;                  // multicast mapping of all inboxes, or null
  float* inbox_peers[kMaxRanks];    // every rank's inbox (peer stores when there is no multicast mapping)
  uint32_t* pads[kMaxRanks];        // flag pads of this channel: slot r = "rank r has pushed epoch e"
  uint32_t* state;                  // local: [0] epoch, [1] CTAs finished, [3] CTAs that have pushed
  const float* red_part; int red_S, red_n, red_ctas;
  int rank, world;
  int64_t n;                        // floats in the range, a multiple of 4
  float step;
  long long timeout_cycles;
  int trace_id;
};
__global__ void __launch_bounds__(512) p2p_push_sum_sgd_kernel(PushArgs a) {
  B200_TRACE_SCOPE(a.trace_id);
  const int tid = threadIdx.x;
  const uint32_t epoch = a.state[0] + 1;
  const int64_t n4 = a.n >> 2;
  float* my_slot_mc = a.inbox_mc ? a.inbox_mc + ((int64_t)(epoch & 1) * a.world + a.rank) * a.n : nullptr;
  const int64_t slot_off = ((int64_t)(epoch & 1) * a.world + a.rank) * a.n;
  auto push = [&](int64_t i4, const float4& v) {
    if (my_slot_mc) mc_st(my_slot_mc + i4 * 4, v);
    else
      for (int p = 0; p < a.world; ++p) *reinterpret_cast<float4*>(a.inbox_peers[p] + slot_off + i4 * 4) = v;
  };
  // ---- 1: push my gradients (the head of the range straight out of the partial sums) ----------------------------
  const int64_t red4 = a.red_ctas > 0 ? a.red_n >> 2 : 0;
  if ((int)blockIdx.x < a.red_ctas) {
    const int t = blockIdx.x * blockDim.x + tid;
    const int g4 = min(t >> 5, (int)red4 - 1);
    const float4 s = warp_sum_partials(a.red_part, a.red_S, a.red_n, g4 * 4, t & 31);
    if ((t & 31) == 0 && (t >> 5) < red4) {
      *reinterpret_cast<float4*>(a.grads + g4 * 4) = s;
      push(g4, s);
    }
  } else {
    const int nb = gridDim.x - a.red_ctas;
    for (int64_t i = red4 + (int64_t)(blockIdx.x - a.red_ctas) * blockDim.x + tid; i < n4; i += (int64_t)nb * blockDim.x)
      push(i, *reinterpret_cast<const float4*>(a.grads + i * 4));
  }
  // ---- 2: "pushed": the last CTA to get here tells every rank (its own included) ---------------------------------
  __syncthreads();
  __shared__ bool last;
  if (tid == 0) {
    __threadfence_system();
    last = atomicAdd(a.state + 3, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last && tid < a.world) {
    __threadfence_system();
    st_flag(a.pads[tid] + a.rank, epoch);
  }
  // ---- 3: all pushes of this step have landed here ------------------------------------------------------------------
  if (tid < a.world) wait_flag(a.pads[a.rank] + tid, epoch, a.timeout_cycles);
  __syncthreads();
  // ---- 4: sum the slots in rank order, update -------------------------------------------------------------------------
  const float* in = a.inbox_local + (int64_t)(epoch & 1) * a.world * a.n;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + tid; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 v[kMaxRanks];
#pragma unroll
    for (int p = 0; p < kMaxRanks; ++p)
      if (p < a.world) v[p] = __ldcg(reinterpret_cast<const float4*>(in + (int64_t)p * a.n) + i);
    float4 g = v[0];
#pragma unroll
    for (int p = 1; p < kMaxRanks; ++p)
      if (p < a.world) { g.x += v[p].x; g.y += v[p].y; g.z += v[p].z; g.w += v[p].w; }
    float4 w = *reinterpret_cast<const float4*>(a.params + i * 4);
    w.x = fmaf(-a.step, g.x, w.x); w.y = fmaf(-a.step, g.y, w.y);
    w.z = fmaf(-a.step, g.z, w.z); w.w = fmaf(-a.step, g.w, w.w);
    *reinterpret_cast<float4*>(a.params + i * 4) = w;
  }
  // ---- 5: the last CTA to finish resets the counters and advances the epoch -------------------------------------------
  __syncthreads();
  if (tid == 0 && atomicAdd(a.state + 1, 1u) == gridDim.x - 1) { a.state[1] = 0; a.state[3] = 0; __threadfence(); a.state[0] = epoch; }
}

long long flag_timeout_cycles() {
  static long long cycles = 0;
  if (!cycles) {
    double sec = 120.0;
    if (const char* e = getenv("B200_P2P_TIMEOUT_S")) {
      const double v = atof(e);
      if (v > 0.0) sec = v;
    }
    cycles = (long long)(sec * 1.9e9);      // clock64 ticks at ~1.9 GHz
  }
  return cycles;
}
}  // namespace
B200_TRACE_SETTER(trace_set_p2p)

int p2p_push_sum_sgd(float* grads, float* params, const float* inbox_local, void* inbox_mc, void* const* inbox_peers,
                     void* const* peer_pads, uint32_t* state, int rank, int world, int64_t n, float lr, float scale,
                     cudaStream_t st, int channel, const float* red_part, int red_S, int red_n) {
  B200_REQUIRE(world >= 2 && world <= kMaxRanks && rank >= 0 && rank < world, "p2p_push_sum_sgd: bad rank %d / world %d", rank, world);
  B200_REQUIRE(grads && params && inbox_local && inbox_peers && peer_pads && state && n > 0 && n % 4 == 0,
               "p2p_push_sum_sgd: bad argument");
  PushArgs a{};
  a.grads = grads; a.params = params; a.inbox_local = inbox_local; a.inbox_mc = (float*)inbox_mc;
  for (int p = 0; p < world; ++p) {
    B200_REQUIRE(inbox_peers[p] && peer_pads[p], "p2p_push_sum_sgd: null peer pointer for rank %d", p);
    a.inbox_peers[p] = (float*)inbox_peers[p];
    a.pads[p] = (uint32_t*)peer_pads[p] + 16 * channel;
  }
  a.state = state + 4 * channel; a.rank = rank; a.world = world; a.n = n; a.step = lr * scale;
  a.timeout_cycles = flag_timeout_cycles();
  a.trace_id = channel == 0 ? TR_P2P_FC : TR_P2P_CONV;
  int grid = (int)std::min<int64_t>(kNumSMs, (n / 4 + 511) / 512);
  if (red_part) {
    B200_REQUIRE(red_S > 0 && red_n > 0 && red_n % 4 == 0 && red_n <= n && ((uintptr_t)red_part & 15) == 0,
                 "p2p_push_sum_sgd: bad partial range");
    a.red_part = red_part; a.red_S = red_S; a.red_n = red_n;
    a.red_ctas = (red_n / 4 * 32 + 511) / 512;
    grid = std::min(kNumSMs, a.red_ctas + std::max(1, (int)(((n - red_n) / 4 + 511) / 512)));
    B200_REQUIRE(grid > a.red_ctas, "p2p_push_sum_sgd: partial range too large");
  }
  p2p_push_sum_sgd_kernel<<<grid, 512, 0, st>>>(a);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int p2p_allreduce_sgd(const void* const* peer_grads, void* const* peer_params, void* const* peer_pads,
                      const void* mc_grads, void* mc_params, uint32_t* state, int rank, int world, int64_t total,
                      float lr, float scale, cudaStream_t st, int64_t offset, int channel, int max_ctas,
                      const float* red_part, int red_S, int red_n) {
  B200_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, "p2p_allreduce_sgd: bad rank %d / world %d",
               rank, world);
  B200_REQUIRE(total > 0 && total % (4 * world) == 0, "p2p_allreduce_sgd: %lld floats do not split into %d 16-byte shards",
               (long long)total, world);
  B200_REQUIRE(peer_grads && peer_params && peer_pads && state, "p2p_allreduce_sgd: null argument");
  P2PArgs a{};
  for (int p = 0; p < world; ++p) {
    B200_REQUIRE(peer_grads[p] && peer_params[p] && peer_pads[p], "p2p_allreduce_sgd: null peer pointer for rank %d", p);
    a.grads[p] = (const float*)peer_grads[p] + offset;
    a.params[p] = (float*)peer_params[p] + offset;
    a.pads[p] = (uint32_t*)peer_pads[p] + 16 * channel;
  }
  B200_REQUIRE(offset >= 0 && offset % 4 == 0 && channel >= 0 && channel < 2, "p2p_allreduce_sgd: bad sub-range / channel");
  a.mc_grads = mc_grads ? (const float*)mc_grads + offset : nullptr;
  a.mc_params = mc_params ? (float*)mc_params + offset : nullptr;
  a.state = state + 4 * channel; a.rank = rank; a.world = world; a.total = total; a.step = lr * scale;
  a.timeout_cycles = flag_timeout_cycles();
  a.trace_id = channel == 0 ? TR_P2P_FC : TR_P2P_CONV;
  const int64_t shard4 = total / 4 / world;
  // (with the multicast mapping a thread handles four 16-byte pieces per sweep)
  const int64_t per_cta = 512 * ((mc_grads && mc_params) ? 4 : 1);
  int grid = (int)std::max<int64_t>(1, std::min<int64_t>(max_ctas > 0 ? max_ctas : kNumSMs, (shard4 + per_cta - 1) / per_cta));
  if (red_part) {      // the first red_n floats of this range still sit in red_S partials: one warp per 16 bytes sums them
    B200_REQUIRE(red_S > 0 && red_n > 0 && red_n % 4 == 0 && red_n <= total && ((uintptr_t)red_part & 15) == 0,
                 "p2p_allreduce_sgd: bad partial range");
    a.red_part = red_part; a.red_out = const_cast<float*>(a.grads[rank]); a.red_S = red_S; a.red_n = red_n;
    a.red_ctas = (red_n / 4 * 32 + 511) / 512;
    B200_REQUIRE(a.red_ctas <= kNumSMs, "p2p_allreduce_sgd: partial range too large to reduce in front of the exchange");
    grid = std::max(grid, a.red_ctas);       // all of them resident at once (<= one CTA per SM): CTA 0 may wait for them
  }
  p2p_allreduce_sgd_kernel<<<grid, 512, 0, st>>>(a);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace b200
