// The small end of the network in two launches instead of ~25:
//   (+ one tiny kernel that rounds W2 to a bf16 image the rows kernel bulk-copies into shared memory)
//   rows kernel  (row-parallel):  full1 = relu(sum of FC1 split-K partials + b1)   [optional]
//                                 full2 = relu(full1 W2 + b2)   (W2 rounded to bf16) cifar10cnn.py:139
//                                 logits = relu(full2 W3 + b3)                      :145
//                                 softmax-xent rows, argmax, correct count          :150-157, :166-176
//                                 dlogits = (softmax - onehot)/B * [logits > 0]
//                                 dfull2 = (dlogits W3^T) * [full2 > 0]
//                                 dfull1 = (dfull2 W2^T) * [full1 > 0]
//   grads kernel (output-parallel, deterministic sums over the batch):
//                                 dW2 = full1^T dfull2, db2, dW3 = full2^T dlogits, db3,
//                                 db1 = colsum(dfull1), loss = mean, n_correct, n_bad
// SIMT with fp32 accumulation: 28 MFLOP per 128 images is nowhere near worth a tensor-core pipeline.
// W2 is held in shared memory as bf16 (it is read by two products; fp32 would not fit); activations,
// softmax, loss and every sum stay fp32.
#include <algorithm>

#include <cuda_bf16.h>

#include "launch.h"
#include "tc_common.cuh"

namespace b200 {
namespace {
constexpr int kMaxCls = 16;
constexpr int RT = 256;    // threads per CTA
constexpr int KT = 32;     // W2 rows / batch rows staged per shared-memory tile

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct TailArgs {
  int B, F1, F2, NC, R;            // batch, widths, classes, rows per CTA
  const float* fc1_part; int n_part;   // [n_part][B][F1] split-K partials of FC1 (or n_part = 0: full1 is an input)
  const float* b1;
  float* full1;                    // [B,F1]  in (n_part == 0) or out
  const float *w2, *b2, *w3, *b3;
  const __nv_bfloat16* w2b;        // W2 as the bf16 [F1][F2+2] image (w2_to_bf16_kernel)
  const int32_t* labels;
  float *full2, *logits, *dlogits, *dfull2, *dfull1;
  int32_t* pred;
  float* part;                     // [n_ctas][PV] per-CTA partial sums: db1 | db2 | db3(16) | dW3 | loss, ok, bad
  float* zero_ptr; int zero_n;     // buffer cleared here for later kernels' atomics (may be null)
  long long* timeline;             // diagnostics: clock64 stamps of CTA 0's phases, or null
};
__host__ __device__ inline int part_len(int F1, int F2, int NC) { return F1 + F2 + 16 + F2 * NC + 4; }

// contiguous `nfloat4` float4s global -> shared with cp.async (all threads)
__device__ __forceinline__ void stage_tile(float* dst, const float* src, int nfloat4, int tid) {
  for (int i = tid; i < nfloat4; i += RT) cp_async16(dst + 4 * i, src + 4 * i);
  cp_async_commit();
}

#define TL(i) do { if (a.timeline && blockIdx.x == 0 && threadIdx.x == 0) a.timeline[i] = clock64(); } while (0)

// W2 fp32 [F1][F2] -> bf16 image [F1][F2+2] (row stride odd in 32-bit words: conflict-free when every
// thread walks its own row), fetched by the rows kernel with bulk copies
__global__ void w2_to_bf16_kernel(const float* __restrict__ w2, __nv_bfloat16* __restrict__ out, int F1, int F2) {
  const int WS = F2 + 2;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < F1 * WS; i += gridDim.x * blockDim.x) {
    const int k = i / WS, n = i - k * WS;
    out[i] = __float2bfloat16(n < F2 ? __ldg(w2 + (size_t)k * F2 + n) : 0.f);
  }
}

template <int R>
__global__ void __launch_bounds__(RT) fc_tail_rows_kernel(TailArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int WS = a.F2 + 2;
  float* xsI = sm;                      // [F1][R]   full1 rows, row index innermost (one broadcast LDS.128 per k)
  float* d1s = xsI + R * a.F1;          // [R][F1]   dfull1 rows
  float* f2s = d1s + R * a.F1;          // [R][F2]
  float* d2I = f2s + R * a.F2;          // [F2][R]   dfull2, row index innermost
  float* lgs = d2I + R * a.F2;          // [R][16] logits
  float* dls = lgs + R * kMaxCls;       // [R][16] dlogits
  float* misc = dls + R * kMaxCls;      // [4] loss, ok, bad
  float* w3s = misc + 4;                // [F2][NC]
  uint64_t* bar = reinterpret_cast<uint64_t*>(w3s + ((a.F2 * a.NC + 3) & ~3));
  __nv_bfloat16* wb = reinterpret_cast<__nv_bfloat16*>(bar + 2);      // [F1][WS] W2 in bf16, resident for both passes
  const int tid = threadIdx.x;
  const int r0 = blockIdx.x * R;
  const int nr = min(R, a.B - r0);

  if (a.zero_ptr)     // every CTA clears its slice of the buffer a later kernel accumulates into
    for (int i = blockIdx.x * RT + tid; i < a.zero_n; i += gridDim.x * RT) a.zero_ptr[i] = 0.f;
  TL(0);
  // ---- W2 (bf16 image, 149 KB) -> shared memory with bulk copies; lands while full1 is assembled ------
  if (tid == 0) {
    tc::mbar_init(tc::smem_u32(bar), 1);
    tc::fence_mbar_init();
    const uint32_t total = (uint32_t)a.F1 * WS * 2;
    tc::mbar_arrive_expect_tx(tc::smem_u32(bar), total);
    for (uint32_t off = 0; off < total; off += 32768)
      tc::bulk_g2s(tc::smem_u32(wb) + off, reinterpret_cast<const uint8_t*>(a.w2b) + off, min(32768u, total - off),
                   tc::smem_u32(bar));
  }
  for (int i = tid; i < a.F2 * a.NC; i += RT) w3s[i] = __ldg(a.w3 + i);
  TL(1);
  // ---- full1 rows: add the split-K partials of FC1, bias, ReLU ------------------------------------
  const int f14 = a.F1 / 4;
  for (int i = tid; i < R * f14; i += RT) {
    const int r = i / f14, k4 = i - r * f14;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < nr) {
      float4* dst = reinterpret_cast<float4*>(a.full1 + (size_t)(r0 + r) * a.F1) + k4;
      if (a.n_part > 0) {
        v = __ldg(reinterpret_cast<const float4*>(a.b1) + k4);
        const float4* p = reinterpret_cast<const float4*>(a.fc1_part + (size_t)(r0 + r) * a.F1) + k4;
        const size_t stride4 = (size_t)a.B * a.F1 / 4;
#pragma unroll 4
        for (int s = 0; s < a.n_part; ++s) {          // loads are independent: several splits in flight
          const float4 t = __ldg(p + s * stride4);
          v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
        }
        v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
        *dst = v;
      } else {
        v = *dst;
      }
    }
    xsI[(4 * k4 + 0) * R + r] = v.x; xsI[(4 * k4 + 1) * R + r] = v.y;
    xsI[(4 * k4 + 2) * R + r] = v.z; xsI[(4 * k4 + 3) * R + r] = v.w;
  }
  __syncthreads();
  tc::mbar_wait(tc::smem_u32(bar), 0);
  TL(2);
  // ---- pass 1 over W2: full2 = relu(x W2 + b2), thread <-> column n ------------------------------------
  if (tid < a.F2) {
    float acc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = 0.f;
#pragma unroll 8
    for (int k = 0; k < a.F1; ++k) {
      const float wv = __bfloat162float(wb[k * WS + tid]);
      float xr[R];
#pragma unroll
      for (int q = 0; q < R / 4; ++q) *reinterpret_cast<float4*>(xr + 4 * q) = *reinterpret_cast<const float4*>(xsI + k * R + 4 * q);
#pragma unroll
      for (int r = 0; r < R; ++r) acc[r] = fmaf(xr[r], wv, acc[r]);
    }
    const float b = __ldg(a.b2 + tid);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float v = fmaxf(acc[r] + b, 0.f);
      f2s[r * a.F2 + tid] = v;
      if (r < nr) a.full2[(size_t)(r0 + r) * a.F2 + tid] = v;
    }
  }
  __syncthreads();
  TL(3);
  // ---- logits = relu(full2 W3 + b3): one thread per (row, class), W3 from shared memory ---------------
  if (tid < R * a.NC) {
    const int r = tid / a.NC, c = tid - r * a.NC;
    float s0 = 0.f, s1 = 0.f;
    for (int k = 0; k + 1 < a.F2; k += 2) {
      s0 = fmaf(f2s[r * a.F2 + k], w3s[k * a.NC + c], s0);
      s1 = fmaf(f2s[r * a.F2 + k + 1], w3s[(k + 1) * a.NC + c], s1);
    }
    if (a.F2 & 1) s0 = fmaf(f2s[r * a.F2 + a.F2 - 1], w3s[(a.F2 - 1) * a.NC + c], s0);
    lgs[r * kMaxCls + c] = fmaxf(s0 + s1 + __ldg(a.b3 + c), 0.f);
  }
  __syncthreads();
  TL(4);
  // ---- softmax cross-entropy per row (R <= 32 rows: one lane each in warp 0) -----------------------------
  if (tid < 32) {
    const int r = tid;
    float lossv = 0.f;
    int okv = 0, badv = 0;
    if (r < nr) {
      float m = -INFINITY;
      int am = 0;
      for (int c = 0; c < a.NC; ++c) {
        const float z = lgs[r * kMaxCls + c];
        if (z > m) { m = z; am = c; }                    // strict '>' = first maximal index (:173)
      }
      float s = 0.f;
      for (int c = 0; c < a.NC; ++c) s += expf(lgs[r * kMaxCls + c] - m);
      const int y = a.labels[r0 + r];
      const bool valid = (unsigned)y < (unsigned)a.NC;
      const float inv = 1.f / s, invB = 1.f / (float)a.B;
      for (int c = 0; c < a.NC; ++c) {
        const float lg = lgs[r * kMaxCls + c];
        const float d = (valid && lg > 0.f) ? (expf(lg - m) * inv - (c == y ? 1.f : 0.f)) * invB : 0.f;
        dls[r * kMaxCls + c] = d;
        a.logits[(size_t)(r0 + r) * a.NC + c] = lg;
        a.dlogits[(size_t)(r0 + r) * a.NC + c] = d;
      }
      if (valid) lossv = logf(s) - (lgs[r * kMaxCls + y] - m);
      else badv = 1;
      okv = (valid && am == y) ? 1 : 0;
      a.pred[r0 + r] = am;
    } else if (r < R) {
      for (int c = 0; c < a.NC; ++c) dls[r * kMaxCls + c] = 0.f;
    }
    lossv = warp_sum(lossv);
    for (int o = 16; o > 0; o >>= 1) {
      okv += __shfl_xor_sync(0xffffffffu, okv, o);
      badv += __shfl_xor_sync(0xffffffffu, badv, o);
    }
    if (tid == 0) { misc[0] = lossv; misc[1] = (float)okv; misc[2] = (float)badv; }
  }
  __syncthreads();
  TL(5);
  // ---- dfull2 = (dlogits W3^T) * [full2 > 0] -----------------------------------------------------------
  if (tid < a.F2) {
    float w[kMaxCls];
#pragma unroll
    for (int c = 0; c < kMaxCls; ++c) w[c] = c < a.NC ? w3s[tid * a.NC + c] : 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float s = 0.f;
#pragma unroll
      for (int c = 0; c < kMaxCls; ++c) s = fmaf(c < a.NC ? dls[r * kMaxCls + c] : 0.f, w[c], s);
      s = f2s[r * a.F2 + tid] > 0.f ? s : 0.f;
      d2I[tid * R + r] = s;
      if (r < nr) a.dfull2[(size_t)(r0 + r) * a.F2 + tid] = s;
    }
  }
  __syncthreads();
  TL(6);
  // ---- pass 2 over W2: dfull1 = (dfull2 W2^T) * [full1 > 0]; thread <-> its own W2 row k -----------------
  for (int k = tid; k < a.F1; k += RT) {
    float acc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = 0.f;
    const __nv_bfloat162* wrow = reinterpret_cast<const __nv_bfloat162*>(wb + k * WS);
#pragma unroll 4
    for (int n2 = 0; n2 < a.F2 / 2; ++n2) {
      const float2 wv = __bfloat1622float2(wrow[n2]);
      float d0[R], d1[R];
#pragma unroll
      for (int q = 0; q < R / 4; ++q) {
        *reinterpret_cast<float4*>(d0 + 4 * q) = *reinterpret_cast<const float4*>(d2I + (2 * n2) * R + 4 * q);
        *reinterpret_cast<float4*>(d1 + 4 * q) = *reinterpret_cast<const float4*>(d2I + (2 * n2 + 1) * R + 4 * q);
      }
#pragma unroll
      for (int r = 0; r < R; ++r) acc[r] = fmaf(d0[r], wv.x, fmaf(d1[r], wv.y, acc[r]));
    }
#pragma unroll
    for (int r = 0; r < R; ++r) d1s[r * a.F1 + k] = xsI[k * R + r] > 0.f ? acc[r] : 0.f;
  }
  __syncthreads();
  TL(7);
  // ---- write dfull1 rows and this CTA's partial sums over its rows ---------------------------------------
  for (int i = tid; i < nr * a.F1; i += RT) a.dfull1[(size_t)r0 * a.F1 + i] = d1s[i];
  float* part = a.part + (size_t)blockIdx.x * part_len(a.F1, a.F2, a.NC);
  for (int k = tid; k < a.F1; k += RT) {
    float s = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) s += d1s[r * a.F1 + k];     // rows >= nr are zero (dls = 0 -> d2 = 0 -> d1 = 0)
    part[k] = s;
  }
  if (tid < a.F2) {
    float s = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) s += d2I[tid * R + r];
    part[a.F1 + tid] = s;
    float* pw3 = part + a.F1 + a.F2 + 16 + tid * a.NC;
    for (int c = 0; c < a.NC; ++c) {
      float t3 = 0.f;
#pragma unroll
      for (int r = 0; r < R; ++r) t3 = fmaf(f2s[r * a.F2 + tid], dls[r * kMaxCls + c], t3);
      pw3[c] = t3;
    }
  }
  if (tid < 16) {
    float s = 0.f;
    if (tid < a.NC)
#pragma unroll
      for (int r = 0; r < R; ++r) s += dls[r * kMaxCls + tid];
    part[a.F1 + a.F2 + tid] = s;
  }
  if (tid < 4) part[a.F1 + a.F2 + 16 + a.F2 * a.NC + tid] = tid < 3 ? misc[tid] : 0.f;
  TL(8);
}

struct TailGradArgs {
  int B, F1, F2, NC, n_row_ctas;
  const float *full1, *dfull2, *part;
  float *dw2, *db2, *dw3, *db3, *db1;
  float* loss; int32_t* n_correct; int32_t* n_bad;
};

// CTAs [0, F1/4): dW2 rows k0..k0+3 = full1[:, k0..k0+3]^T dfull2, batch staged through shared memory.
// Remaining CTAs: fixed-order sums of the row-kernel partials (db1, db2, db3, dW3, loss, counts).
__global__ void __launch_bounds__(RT) fc_tail_grads_kernel(TailGradArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int tid = threadIdx.x;
  const int n_dw2 = a.F1 / 4;
  if ((int)blockIdx.x < n_dw2) {
    float* dt = sm;                          // [2][KT][F2]
    float* xt = dt + 2 * KT * a.F2;          // [2][KT][4]
    const int k0 = blockIdx.x * 4;
    const int NTb = (a.B + KT - 1) / KT;
    auto stage = [&](int t) {
      const int b0 = t * KT, nb = min(KT, a.B - b0);
      float* d = dt + (t & 1) * KT * a.F2;
      float* x = xt + (t & 1) * KT * 4;
      for (int i = tid; i < nb * a.F2 / 4; i += RT) cp_async16(d + 4 * i, a.dfull2 + (size_t)b0 * a.F2 + 4 * i);
      if (tid < nb) cp_async16(x + 4 * tid, a.full1 + (size_t)(b0 + tid) * a.F1 + k0);
      cp_async_commit();
    };
    stage(0);
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
    for (int t = 0; t < NTb; ++t) {
      if (t + 1 < NTb) stage(t + 1);
      else cp_async_commit();
      cp_async_wait<1>();
      __syncthreads();
      const int nb = min(KT, a.B - t * KT);
      const float* d = dt + (t & 1) * KT * a.F2;
      const float4* x = reinterpret_cast<const float4*>(xt + (t & 1) * KT * 4);
      if (tid < a.F2) {
#pragma unroll 8
        for (int bb = 0; bb < nb; ++bb) {
          const float4 xv = x[bb];
          const float dv = d[bb * a.F2 + tid];
          acc0 = fmaf(xv.x, dv, acc0); acc1 = fmaf(xv.y, dv, acc1); acc2 = fmaf(xv.z, dv, acc2); acc3 = fmaf(xv.w, dv, acc3);
        }
      }
      __syncthreads();
    }
    if (tid < a.F2) {
      a.dw2[(size_t)(k0 + 0) * a.F2 + tid] = acc0; a.dw2[(size_t)(k0 + 1) * a.F2 + tid] = acc1;
      a.dw2[(size_t)(k0 + 2) * a.F2 + tid] = acc2; a.dw2[(size_t)(k0 + 3) * a.F2 + tid] = acc3;
    }
    return;
  }
  const int PV = part_len(a.F1, a.F2, a.NC);
  const int v = (blockIdx.x - n_dw2) * RT + tid;
  if (v >= PV) return;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  int i = 0;
  for (; i + 4 <= a.n_row_ctas; i += 4) {      // four loads in flight; the summation order is fixed
    s0 += __ldg(a.part + (size_t)(i + 0) * PV + v); s1 += __ldg(a.part + (size_t)(i + 1) * PV + v);
    s2 += __ldg(a.part + (size_t)(i + 2) * PV + v); s3 += __ldg(a.part + (size_t)(i + 3) * PV + v);
  }
  for (; i < a.n_row_ctas; ++i) s0 += __ldg(a.part + (size_t)i * PV + v);
  const float s = (s0 + s1) + (s2 + s3);
  const int o1 = a.F1, o2 = o1 + a.F2, o3 = o2 + 16, o4 = o3 + a.F2 * a.NC;
  if (v < o1) a.db1[v] = s;
  else if (v < o2) a.db2[v - o1] = s;
  else if (v < o3) { if (v - o2 < a.NC) a.db3[v - o2] = s; }
  else if (v < o4) a.dw3[v - o3] = s;
  else if (v == o4) *a.loss = s / (float)a.B;
  else if (v == o4 + 1) *a.n_correct = (int)(s + 0.5f);
  else if (v == o4 + 2) *a.n_bad = (int)(s + 0.5f);
}

int rows_per_cta(int B) { return B <= 1024 ? 4 : 8; }
long long* g_tail_timeline = nullptr;
}  // namespace

void fc_tail_set_timeline(long long* dev_buf) { g_tail_timeline = dev_buf; }

static int64_t part_bytes(int B, int F1, int F2, int NC) {
  return round_up((int64_t)cdiv(B, rows_per_cta(B)) * part_len(F1, F2, NC) * 4, 256);
}
// per-CTA partial sums + the bf16 image of W2
int64_t fc_tail_scratch_bytes(int B, int F1, int F2, int NC) {
  return part_bytes(B, F1, F2, NC) + round_up((int64_t)F1 * (F2 + 2) * 2, 256);
}

// the bf16 image of W2 inside `scratch` (same layout fc_tail expects); lets a caller build it early, off
// the critical path, and then call fc_tail(..., w2_image_ready = true)
int fc_tail_prepare_w2(const float* w2, void* scratch, int B, int F1, int F2, int NC, cudaStream_t st) {
  __nv_bfloat16* w2b = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(scratch) + part_bytes(B, F1, F2, NC));
  w2_to_bf16_kernel<<<cdiv(F1 * (F2 + 2), 256), 256, 0, st>>>(w2, w2b, F1, F2);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int fc_tail(const float* fc1_part, int n_part, const float* b1, float* full1, const float* w2, const float* b2,
            const float* w3, const float* b3, const int32_t* labels, float* full2, float* logits, float* dlogits,
            float* dfull2, float* dfull1, int32_t* pred, float* loss, int32_t* n_correct, int32_t* n_bad, float* dw2,
            float* db2, float* dw3, float* db3, float* db1, void* scratch, float* zero_ptr, int zero_n, int B, int F1,
            int F2, int NC, cudaStream_t st, bool w2_image_ready, cudaStream_t grads_stream, cudaEvent_t rows_done) {
  B200_REQUIRE(B > 0 && NC > 0 && NC <= kMaxCls, "fc_tail: bad shape");
  B200_REQUIRE(F1 % KT == 0 && F2 % 4 == 0 && F2 <= RT && F2 > 0, "fc_tail: needs f1 %% 32 == 0, f2 %% 4 == 0, f2 <= 256");
  const int R = rows_per_cta(B);
  const int nrc = cdiv(B, R);
  float* part = reinterpret_cast<float*>(scratch);
  __nv_bfloat16* w2b = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(scratch) + part_bytes(B, F1, F2, NC));
  if (!w2_image_ready) {
    w2_to_bf16_kernel<<<cdiv(F1 * (F2 + 2), 256), 256, 0, st>>>(w2, w2b, F1, F2);
    B200_LAUNCH_CHECK();
  }
  TailArgs a{B, F1, F2, NC, R, fc1_part, n_part, b1, full1, w2, b2, w3, b3, w2b, labels, full2, logits, dlogits,
             dfull2, dfull1, pred, part, zero_ptr, zero_n, g_tail_timeline};
  const size_t smem = (size_t)(2 * R * F1 + 2 * R * F2 + 2 * R * kMaxCls + 4 + ((F2 * NC + 3) & ~3) + 4) * 4 +
                      (size_t)F1 * (F2 + 2) * 2;
  B200_REQUIRE(smem <= 200 * 1024, "fc_tail: layer widths too large (%zu bytes of shared memory)", smem);
  static bool attr = false;
  if (!attr) {
    B200_CUDA_TRY(cudaFuncSetAttribute(fc_tail_rows_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    B200_CUDA_TRY(cudaFuncSetAttribute(fc_tail_rows_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    B200_CUDA_TRY(cudaFuncSetAttribute(fc_tail_grads_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    attr = true;
  }
  if (R == 4) fc_tail_rows_kernel<4><<<nrc, RT, smem, st>>>(a);
  else fc_tail_rows_kernel<8><<<nrc, RT, smem, st>>>(a);
  B200_LAUNCH_CHECK();
  TailGradArgs ga{B, F1, F2, NC, nrc, full1, dfull2, part, dw2, db2, dw3, db3, db1, loss, n_correct, n_bad};
  const size_t gsmem = (size_t)(2 * KT * F2 + 2 * KT * 4) * 4;
  if (grads_stream && grads_stream != st) {       // the parameter gradients do not gate dfull1's consumer: let them overlap it
    B200_CUDA_TRY(cudaEventRecord(rows_done, st));
    B200_CUDA_TRY(cudaStreamWaitEvent(grads_stream, rows_done, 0));
    st = grads_stream;
  }
  fc_tail_grads_kernel<<<F1 / 4 + cdiv(part_len(F1, F2, NC), RT), RT, gsmem, st>>>(ga);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace b200
