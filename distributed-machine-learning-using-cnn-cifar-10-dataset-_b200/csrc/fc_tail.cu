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
// The rows kernel is row-parallel for latency: 8 (16 for large batches) batch rows per CTA, so 128 images
// spread over 16 SMs.  Its two products with W2 (FC2 forward and dfull1) run on warp-level bf16 MMAs
// (m16n8k16, fp32 accumulate) out of a bf16 image of W2 resident in shared memory -- rows this few are far
// below a tcgen05 tile (M >= 64).  Operands of those two products (full1, dfull2, W2) are rounded to bf16
// like every other tensor-core operand of the bf16 mode; FC3, softmax, loss and every sum stay fp32.
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
  const __nv_bfloat16* w2b;        // W2 as the bf16 [F1][F2+8] image (w2_to_bf16_kernel)
  const int32_t* labels;
  float *full2, *logits, *dlogits, *dfull2, *dfull1;
  int32_t* pred;
  float* part;                     // [n_ctas][PV] per-CTA partial sums: db1 | db2 | db3(16) | dW3 | loss, ok, bad
  float* zero_ptr; int zero_n;     // buffer cleared here for later kernels' atomics (may be null)
  long long* timeline;             // diagnostics: clock64 stamps of CTA 0's phases, or null
  uint8_t* dfull1_img;             // optional: dfull1 again as FC1's bf16 operand image [F1/8][B][8]
};
__host__ __device__ inline int part_len(int F1, int F2, int NC) { return F1 + F2 + 16 + F2 * NC + 4; }

// contiguous `nfloat4` float4s global -> shared with cp.async (all threads)
__device__ __forceinline__ void stage_tile(float* dst, const float* src, int nfloat4, int tid) {
  for (int i = tid; i < nfloat4; i += RT) cp_async16(dst + 4 * i, src + 4 * i);
  cp_async_commit();
}

#define TL(i) do { if (a.timeline && blockIdx.x == 0 && threadIdx.x == 0) a.timeline[i] = clock64(); } while (0)

// W2 fp32 [F1][F2] -> bf16 image [F1][F2+8]: rows stay 16-byte aligned (ldmatrix) and shift by four banks per
// row (conflict-free fragment loads); fetched by the rows kernel with bulk copies
__global__ void w2_to_bf16_kernel(const float* __restrict__ w2, __nv_bfloat16* __restrict__ out, int F1, int F2) {
  B200_TRACE_SCOPE(TR_W2IMG);
  pdl_launch_dependents();
  const int WS = F2 + 8;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < F1 * WS; i += gridDim.x * blockDim.x) {
    const int k = i / WS, n = i - k * WS;
    out[i] = __float2bfloat16(n < F2 ? __ldg(w2 + (size_t)k * F2 + n) : 0.f);
  }
}

// warp-level bf16 MMA, D[16x8] += A[16x16] B[16x8], fp32 accumulate.  The tail is row-parallel over many SMs
// for latency (8 or 16 batch rows per CTA), far below the 64-row minimum of a tcgen05 tile.
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// two 8x8 bf16 matrices, transposed on the way in: lanes 0-7 / 8-15 give the row addresses of matrix 0 / 1
__device__ __forceinline__ void ldmatrix_x2_trans(uint32_t& r0, uint32_t& r1, uint32_t saddr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(saddr));
}
// four 8x8 bf16 matrices; lane l gives the address of row l % 8 of matrix l / 8
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t saddr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(saddr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t saddr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(saddr));
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

struct TailSmem {          // byte offsets into dynamic shared memory
  int wb, xb, d2b, f2s, w3s, w3t, dls, stat, total;
};
__host__ __device__ inline TailSmem tail_smem(int RM, int F1, int F2, int NC) {
  TailSmem s;
  int o = 16;                                   // mbarrier first
  s.wb = o;  o += F1 * (F2 + 8) * 2;            // bf16 W2 image
  s.xb = o;  o += 16 * (F1 + 8) * 2;            // bf16 full1 rows (A operand of FC2); rows >= RM stay zero
  s.d2b = o; o += 16 * (F2 + 8) * 2;            // bf16 dfull2 rows (A operand of the dfull1 product)
  s.f2s = o; o += RM * F2 * 4;                  // fp32 full2 rows
  s.w3s = o; o += ((F2 * NC + 3) & ~3) * 4;     // W3
  s.w3t = o; o += kMaxCls * F2 * 4;             // W3 transposed [16][k], rows >= NC zero
  s.dls = o; o += RM * kMaxCls * 4;             // dlogits
  s.stat = o; o += RM * 4 * 4;                  // per row: loss, correct, bad label
  s.total = o;
  return s;
}

// RM: rows of the MMA tile that can hold batch rows (8: the upper half of the 16-row tile is zero; 16).  The CTA
// handles a.R <= RM batch rows.  T1: 8-column tiles of full2 per warp (F2 <= 64 * T1).
template <int RM, int T1>
__global__ void __launch_bounds__(RT) fc_tail_rows_kernel(TailArgs a) {
  B200_TRACE_SCOPE(TR_TAIL_ROWS);
  extern __shared__ __align__(16) uint8_t smraw[];
  const TailSmem L = tail_smem(RM, a.F1, a.F2, a.NC);
  const int WS = a.F2 + 8, XS = a.F1 + 8, DS = a.F2 + 8;      // pitches in bf16 elements
  uint64_t* bar = reinterpret_cast<uint64_t*>(smraw);
  __nv_bfloat16* wb = reinterpret_cast<__nv_bfloat16*>(smraw + L.wb);
  __nv_bfloat16* xb = reinterpret_cast<__nv_bfloat16*>(smraw + L.xb);
  __nv_bfloat16* d2b = reinterpret_cast<__nv_bfloat16*>(smraw + L.d2b);
  float* f2s = reinterpret_cast<float*>(smraw + L.f2s);
  float* w3s = reinterpret_cast<float*>(smraw + L.w3s);
  float* w3t = reinterpret_cast<float*>(smraw + L.w3t);
  float* dls = reinterpret_cast<float*>(smraw + L.dls);
  float* stat = reinterpret_cast<float*>(smraw + L.stat);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int r0 = blockIdx.x * a.R;
  const int nr = min(a.R, a.B - r0);
  float* part = a.part + (size_t)blockIdx.x * part_len(a.F1, a.F2, a.NC);

  if (a.zero_ptr)     // every CTA clears its slice of the buffer a later kernel accumulates into
    for (int i = blockIdx.x * RT + tid; i < a.zero_n; i += gridDim.x * RT) a.zero_ptr[i] = 0.f;
  TL(0);
  // ---- W2 (bf16 image, 150 KB) -> shared memory with bulk copies; lands while full1 is assembled ------
  if (tid == 0) {
    tc::mbar_init(tc::smem_u32(bar), 1);
    tc::fence_mbar_init();
    const uint32_t total = (uint32_t)a.F1 * WS * 2;
    tc::mbar_arrive_expect_tx(tc::smem_u32(bar), total);
    for (uint32_t off = 0; off < total; off += 32768)
      tc::bulk_g2s(tc::smem_u32(wb) + off, reinterpret_cast<const uint8_t*>(a.w2b) + off, min(32768u, total - off),
                   tc::smem_u32(bar));
  }
  // small operands used much later are requested now (nothing waits on them here)
  const float b3v = lane < a.NC ? __ldg(a.b3 + lane) : 0.f;
  int ylab[RM / 8];
#pragma unroll
  for (int q = 0; q < RM / 8; ++q) ylab[q] = warp + 8 * q < nr ? __ldg(a.labels + r0 + warp + 8 * q) : 0;
  float2 bias2[T1];
#pragma unroll
  for (int i = 0; i < T1; ++i)
    bias2[i] = __ldg(reinterpret_cast<const float2*>(a.b2 + min(warp + 8 * i, a.F2 / 8 - 1) * 8 + 2 * t));
  constexpr int W3B = 8;
  float w3v[W3B];
  const int n3 = a.F2 * a.NC;
#pragma unroll
  for (int u = 0; u < W3B; ++u) w3v[u] = tid + u * RT < n3 ? __ldg(a.w3 + tid + u * RT) : 0.f;
  TL(1);
  // ---- full1 rows: add the split-K partials of FC1, bias, ReLU; keep a bf16 copy as the MMA operand.  Three
  //      slots per thread at a time, the loads of all their splits in flight together ----------------------------
  const int f14 = a.F1 / 4;
  // launched with programmatic stream serialization, everything above ran while the producer of fc1_part (FC1)
  // was still executing; from here on its output is needed (no-op for an ordinary launch)
  pdl_wait();
  pdl_launch_dependents();
  if (RM == 8)      // upper half of the A tile
    for (int i = tid; i < XS; i += RT) reinterpret_cast<uint4*>(xb + 8 * XS)[i] = make_uint4(0, 0, 0, 0);
  if (a.R < RM)      // rows of the tile beyond this CTA's batch rows
    for (int i = tid; i < (RM - a.R) * XS / 8; i += RT) reinterpret_cast<uint4*>(xb + a.R * XS)[i] = make_uint4(0, 0, 0, 0);
  constexpr int SL = 2;      // slots (float4 of one row) per thread and batch; every split of both is in flight together
  for (int i0 = tid; i0 < a.R * f14; i0 += SL * RT) {
    // slots past the end / rows past the batch read a valid address and are not stored: no branch around the loads
    float4 v[SL];
    const float4* p[SL];
    int rr[SL], kk[SL];
    bool on[SL];
#pragma unroll
    for (int u = 0; u < SL; ++u) {
      const int i = min(i0 + u * RT, a.R * f14 - 1);
      rr[u] = i / f14; kk[u] = i - rr[u] * f14;
      on[u] = i0 + u * RT < a.R * f14;
      const int row = r0 + min(rr[u], nr - 1);
      if (a.n_part > 0) {
        p[u] = reinterpret_cast<const float4*>(a.fc1_part) + (size_t)row * f14 + kk[u];
        v[u] = __ldg(reinterpret_cast<const float4*>(a.b1) + kk[u]);
      } else {
        p[u] = reinterpret_cast<const float4*>(a.full1) + (size_t)row * f14 + kk[u];
        v[u] = *p[u];
      }
    }
    if (a.n_part > 0) {
      const size_t stride4 = (size_t)a.B * f14;
#pragma unroll 12
      for (int s = 0; s < a.n_part; ++s) {
#pragma unroll
        for (int u = 0; u < SL; ++u) {
          const float4 q = __ldg(p[u] + s * stride4);
          v[u].x += q.x; v[u].y += q.y; v[u].z += q.z; v[u].w += q.w;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < SL; ++u) {
      if (on[u]) {
        if (rr[u] < nr) {
          if (a.n_part > 0) {
            v[u].x = fmaxf(v[u].x, 0.f); v[u].y = fmaxf(v[u].y, 0.f); v[u].z = fmaxf(v[u].z, 0.f); v[u].w = fmaxf(v[u].w, 0.f);
            reinterpret_cast<float4*>(a.full1 + (size_t)(r0 + rr[u]) * a.F1)[kk[u]] = v[u];
          }
        } else {
          v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        // (a positive fp32 value too small for bf16 is stored as the smallest bf16 subnormal, not 0: the ReluGrad mask
        //  of the dfull1 product reads this copy, and `copy > 0` must mean `full1 > 0` exactly)
        uint2 pk = make_uint2(pack_bf16x2(v[u].x, v[u].y), pack_bf16x2(v[u].z, v[u].w));
        if (v[u].x > 0.f && (pk.x & 0xffffu) == 0u) pk.x |= 1u;
        if (v[u].y > 0.f && (pk.x >> 16) == 0u) pk.x |= 0x10000u;
        if (v[u].z > 0.f && (pk.y & 0xffffu) == 0u) pk.y |= 1u;
        if (v[u].w > 0.f && (pk.y >> 16) == 0u) pk.y |= 0x10000u;
        *reinterpret_cast<uint2*>(xb + rr[u] * XS + 4 * kk[u]) = pk;
      }
    }
  }
  // W3 in both orientations: [k][class] for the dfull2 product, [class][k] for the logits
#pragma unroll
  for (int u = 0; u < W3B; ++u) {
    const int i = tid + u * RT;
    if (i < n3) { const int k = i / a.NC; w3s[i] = w3v[u]; w3t[(i - k * a.NC) * a.F2 + k] = w3v[u]; }
  }
  for (int i = a.NC * a.F2 + tid; i < kMaxCls * a.F2; i += RT) w3t[i] = 0.f;
  for (int i = tid + W3B * RT; i < n3; i += RT) {        // wider layers than the reference's: the rest, unhurried
    const float v = __ldg(a.w3 + i);
    const int k = i / a.NC;
    w3s[i] = v; w3t[(i - k * a.NC) * a.F2 + k] = v;
  }
  __syncthreads();
  tc::mbar_wait(tc::smem_u32(bar), 0);
  TL(2);
  // ---- full2 = relu(full1 W2 + b2): warp w owns the 8-column tiles w, w+8, ...  (a tile index past the end is
  //      clamped: computed again, never stored).  Fragments of the next k-step are fetched before the MMAs of
  //      the current one ------------------------------------------------------------------------------------
  {
    const int nt = a.F2 / 8;
    float c[T1][4];
    uint32_t wrow[T1];
    // A fragments (16 x 16, both halves of the row tile: rows >= RM are zero in shared memory) and B fragments
    // come in with one ldmatrix.x4 per 32 k: lane l addresses row l % 8 of 8x8 matrix l / 8
    const uint32_t xrow = tc::smem_u32(xb) + (uint32_t)(((lane & 7) + ((lane >> 3) & 1) * 8) * XS + (lane >> 4) * 8) * 2;
#pragma unroll
    for (int i = 0; i < T1; ++i) {
      c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
      wrow[i] = tc::smem_u32(wb) + (uint32_t)(lane * WS + min(warp + 8 * i, nt - 1) * 8) * 2;
    }
    const uint32_t kstep = (uint32_t)(32 * WS) * 2;
    uint32_t al0[4], ah0[4], al1[4], ah1[4], b0[T1][4], b1[T1][4];
    auto load = [&](int kk, uint32_t (&al)[4], uint32_t (&ah)[4], uint32_t (&b)[T1][4]) {     // kk: index of the 32-k block
      ldmatrix_x4(al, xrow + kk * 64);
      ldmatrix_x4(ah, xrow + kk * 64 + 32);
#pragma unroll
      for (int i = 0; i < T1; ++i) ldmatrix_x4_trans(b[i], wrow[i] + kk * kstep);
    };
    auto mmas = [&](const uint32_t (&al)[4], const uint32_t (&ah)[4], const uint32_t (&b)[T1][4]) {
#pragma unroll
      for (int i = 0; i < T1; ++i) mma16816(c[i], al[0], al[1], al[2], al[3], b[i][0], b[i][1]);
#pragma unroll
      for (int i = 0; i < T1; ++i) mma16816(c[i], ah[0], ah[1], ah[2], ah[3], b[i][2], b[i][3]);
    };
    const int nkb = a.F1 / 32;
    load(0, al0, ah0, b0);
#pragma unroll 1
    for (int kk = 0; kk < nkb; kk += 2) {
      if (kk + 1 < nkb) load(kk + 1, al1, ah1, b1);
      mmas(al0, ah0, b0);
      if (kk + 2 < nkb) load(kk + 2, al0, ah0, b0);
      if (kk + 1 < nkb) mmas(al1, ah1, b1);
    }
#pragma unroll
    for (int i = 0; i < T1; ++i) {
      const int j = warp + 8 * i;
      if (j < nt) {
        const int n = j * 8 + 2 * t;
        const float2 v0 = make_float2(fmaxf(c[i][0] + bias2[i].x, 0.f), fmaxf(c[i][1] + bias2[i].y, 0.f));
        *reinterpret_cast<float2*>(f2s + g * a.F2 + n) = v0;
        if (g < nr) *reinterpret_cast<float2*>(a.full2 + (size_t)(r0 + g) * a.F2 + n) = v0;
        if (RM == 16) {
          const float2 v1 = make_float2(fmaxf(c[i][2] + bias2[i].x, 0.f), fmaxf(c[i][3] + bias2[i].y, 0.f));
          *reinterpret_cast<float2*>(f2s + (g + 8) * a.F2 + n) = v1;
          if (g + 8 < nr) *reinterpret_cast<float2*>(a.full2 + (size_t)(r0 + g + 8) * a.F2 + n) = v1;
        }
      }
    }
  }
  __syncthreads();
  TL(3);
  // ---- logits = relu(full2 W3 + b3) and the softmax cross-entropy of a row, all inside one warp: lanes split K,
  //      a butterfly leaves every class sum in every lane, lane c then owns class c -----------------------------
#pragma unroll
  for (int q = 0; q < RM / 8; ++q) {
    const int r = warp + 8 * q;
    float acc[kMaxCls];
#pragma unroll
    for (int c = 0; c < kMaxCls; ++c) acc[c] = 0.f;
#pragma unroll 2
    for (int k = lane; k < a.F2; k += 32) {
      const float f = f2s[r * a.F2 + k];
#pragma unroll
      for (int c = 0; c < kMaxCls; ++c) acc[c] = fmaf(f, w3t[c * a.F2 + k], acc[c]);
    }
    // transposing butterfly: each step adds the partner's half of the remaining classes, so 8+4+2+1 (+1) shuffles
    // leave the complete sum of class c in lanes 2c and 2c+1
#pragma unroll
    for (int h = kMaxCls / 2, o = 16; h >= 1; h >>= 1, o >>= 1) {
      const bool up = (lane & o) != 0;
#pragma unroll
      for (int i = 0; i < h; ++i) {
        const float send = up ? acc[i] : acc[i + h];
        const float keep = up ? acc[i + h] : acc[i];
        acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
      }
    }
    acc[0] += __shfl_xor_sync(0xffffffffu, acc[0], 1);
    float z = __shfl_sync(0xffffffffu, acc[0], (lane & 15) << 1);        // lane c <- class c
    const bool act = lane < a.NC;
    z = act ? fmaxf(z + b3v, 0.f) : 0.f;
    float m = act ? z : -INFINITY;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    const int am = __ffs(__ballot_sync(0xffffffffu, act && z == m)) - 1;     // first maximal index (:173)
    const float e = act ? expf(z - m) : 0.f;
    const float s = warp_sum(e);
    const int y = ylab[q];
    const bool valid = (unsigned)y < (unsigned)a.NC;
    const float zy = __shfl_sync(0xffffffffu, z, valid ? y : 0);
    float d = 0.f;
    if (r < nr) {
      if (act) {
        d = (valid && z > 0.f) ? (e / s - (lane == y ? 1.f : 0.f)) / (float)a.B : 0.f;
        a.logits[(size_t)(r0 + r) * a.NC + lane] = z;
        a.dlogits[(size_t)(r0 + r) * a.NC + lane] = d;
      }
      if (lane == 0) {
        stat[r * 4 + 0] = valid ? logf(s) - (zy - m) : 0.f;
        stat[r * 4 + 1] = (valid && am == y) ? 1.f : 0.f;
        stat[r * 4 + 2] = valid ? 0.f : 1.f;
        a.pred[r0 + r] = am;
      }
    } else if (lane == 0) {
      stat[r * 4 + 0] = stat[r * 4 + 1] = stat[r * 4 + 2] = 0.f;
    }
    if (lane < kMaxCls) dls[r * kMaxCls + lane] = d;       // zero for classes >= NC and rows >= nr
  }
  __syncthreads();
  TL(4);
  TL(5);
  // ---- dfull2 = (dlogits W3^T) * [full2 > 0]; thread <-> column n; also this CTA's db2 / dW3 / db3 sums ------
  if (tid < a.F2) {
    float w[kMaxCls], t3[kMaxCls];
#pragma unroll
    for (int c = 0; c < kMaxCls; ++c) { w[c] = c < a.NC ? w3s[tid * a.NC + c] : 0.f; t3[c] = 0.f; }
    float colsum = 0.f;
#pragma unroll
    for (int r = 0; r < RM; ++r) {
      if (r >= a.R) {            // (uniform) rows of the MMA tile that carry no batch row: dfull2 = 0
        d2b[r * DS + tid] = __float2bfloat16(0.f);
        continue;
      }
      const float f2v = f2s[r * a.F2 + tid];
      float dv[kMaxCls];
#pragma unroll
      for (int c4 = 0; c4 < kMaxCls / 4; ++c4)
        *reinterpret_cast<float4*>(dv + 4 * c4) = *reinterpret_cast<const float4*>(dls + r * kMaxCls + 4 * c4);
      float s = 0.f;
#pragma unroll
      for (int c = 0; c < kMaxCls; ++c) {
        s = fmaf(dv[c], w[c], s);
        t3[c] = fmaf(f2v, dv[c], t3[c]);
      }
      s = f2v > 0.f ? s : 0.f;
      d2b[r * DS + tid] = __float2bfloat16(s);
      if (r < nr) a.dfull2[(size_t)(r0 + r) * a.F2 + tid] = s;
      colsum += s;
    }
    if (RM == 8)
#pragma unroll
      for (int r = 8; r < 16; ++r) d2b[r * DS + tid] = __float2bfloat16(0.f);
    part[a.F1 + tid] = colsum;
    float* pw3 = part + a.F1 + a.F2 + 16 + tid * a.NC;
#pragma unroll
    for (int c = 0; c < kMaxCls; ++c)
      if (c < a.NC) pw3[c] = t3[c];
  }
  if (tid < 16) {
    float s = 0.f;
    if (tid < a.NC)
#pragma unroll
      for (int r = 0; r < RM; ++r) s += dls[r * kMaxCls + tid];
    part[a.F1 + a.F2 + tid] = s;
  }
  if (tid >= 32 && tid < 36) {
    float s = 0.f;
    if (tid < 35)
#pragma unroll
      for (int r = 0; r < RM; ++r) s += stat[r * 4 + (tid - 32)];
    part[a.F1 + a.F2 + 16 + a.F2 * a.NC + (tid - 32)] = s;
  }
  __syncthreads();
  TL(6);
  // ---- dfull1 = (dfull2 W2^T) * [full1 > 0]: warp w owns the 8-column tiles w, w+8, ... of each 48-tile round
  //      (tile indices past the end are clamped: computed again, never stored); the column sums over this CTA's
  //      rows (db1 partial) come out of the accumulator fragments --------------------------------------------------
  {
    constexpr int T2 = 6;
    const int nt = a.F1 / 8;
    const uint32_t drow = tc::smem_u32(d2b) + (uint32_t)(((lane & 7) + ((lane >> 3) & 1) * 8) * DS + (lane >> 4) * 8) * 2;
    const int nkb = a.F2 / 32;
    for (int jb = 0; jb < nt; jb += 8 * T2) {
      float c[T2][4];
      float2 m0[T2], m1[T2];
      uint32_t wr[T2];
#pragma unroll
      for (int i = 0; i < T2; ++i) {
        c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
        const int j = min(jb + warp + 8 * i, nt - 1);
        // B[k = n2][n = k1] = W2[k1][n2]: rows k1 of the image, 8 consecutive n2 per 16-byte row piece
        wr[i] = tc::smem_u32(wb) + (uint32_t)((j * 8 + (lane & 7)) * WS + (lane >> 3) * 8) * 2;
        m0[i] = m1[i] = make_float2(0.f, 0.f);
        // (the bf16 copy in shared memory has full1's signs: no trip to L2 for values this CTA wrote a moment ago)
        if (g < nr) m0[i] = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(xb + g * XS + j * 8 + 2 * t));
        if (RM == 16 && g + 8 < nr)
          m1[i] = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(xb + (g + 8) * XS + j * 8 + 2 * t));
      }
      uint32_t al0[4], ah0[4], al1[4], ah1[4], b0[T2][4], b1[T2][4];
      auto load = [&](int kk, uint32_t (&al)[4], uint32_t (&ah)[4], uint32_t (&b)[T2][4]) {
        ldmatrix_x4(al, drow + kk * 64);
        ldmatrix_x4(ah, drow + kk * 64 + 32);
#pragma unroll
        for (int i = 0; i < T2; ++i) ldmatrix_x4(b[i], wr[i] + kk * 64);
      };
      auto mmas = [&](const uint32_t (&al)[4], const uint32_t (&ah)[4], const uint32_t (&b)[T2][4]) {
#pragma unroll
        for (int i = 0; i < T2; ++i) mma16816(c[i], al[0], al[1], al[2], al[3], b[i][0], b[i][1]);
#pragma unroll
        for (int i = 0; i < T2; ++i) mma16816(c[i], ah[0], ah[1], ah[2], ah[3], b[i][2], b[i][3]);
      };
      load(0, al0, ah0, b0);
#pragma unroll 1
      for (int kk = 0; kk < nkb; kk += 2) {
        if (kk + 1 < nkb) load(kk + 1, al1, ah1, b1);
        mmas(al0, ah0, b0);
        if (kk + 2 < nkb) load(kk + 2, al0, ah0, b0);
        if (kk + 1 < nkb) mmas(al1, ah1, b1);
      }
#pragma unroll
      for (int i = 0; i < T2; ++i) {
        const int j = jb + warp + 8 * i;
        const int col = min(j, nt - 1) * 8 + 2 * t;
        const float2 v0 = make_float2(m0[i].x > 0.f ? c[i][0] : 0.f, m0[i].y > 0.f ? c[i][1] : 0.f);
        float s0 = v0.x, s1 = v0.y;
        if (j < nt && g < nr) {
          *reinterpret_cast<float2*>(a.dfull1 + (size_t)(r0 + g) * a.F1 + col) = v0;
          if (a.dfull1_img)      // column chunk j, batch row r0 + g, elements 2t, 2t+1 of the chunk
            *reinterpret_cast<uint32_t*>(a.dfull1_img + ((size_t)j * a.B + r0 + g) * 16 + 4 * t) = pack_bf16x2(v0.x, v0.y);
        }
        if (RM == 16) {
          const float2 v1 = make_float2(m1[i].x > 0.f ? c[i][2] : 0.f, m1[i].y > 0.f ? c[i][3] : 0.f);
          if (j < nt && g + 8 < nr) {
            *reinterpret_cast<float2*>(a.dfull1 + (size_t)(r0 + g + 8) * a.F1 + col) = v1;
            if (a.dfull1_img)
              *reinterpret_cast<uint32_t*>(a.dfull1_img + ((size_t)j * a.B + r0 + g + 8) * 16 + 4 * t) = pack_bf16x2(v1.x, v1.y);
          }
          s0 += v1.x; s1 += v1.y;
        }
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
          s0 += __shfl_xor_sync(0xffffffffu, s0, o);
          s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        }
        if (j < nt && g == 0) *reinterpret_cast<float2*>(part + col) = make_float2(s0, s1);
      }
    }
  }
  TL(7);
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
  B200_TRACE_SCOPE(TR_TAIL_GRADS);
  pdl_launch_dependents();
  pdl_wait();
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

int rows_per_cta(int B) { return B <= 256 ? 4 : B <= 1024 ? 8 : 16; }
long long* g_tail_timeline = nullptr;
}  // namespace

void fc_tail_set_timeline(long long* dev_buf) { g_tail_timeline = dev_buf; }
B200_TRACE_SETTER(trace_set_fc_tail)

static int64_t part_bytes(int B, int F1, int F2, int NC) {
  return round_up((int64_t)cdiv(B, rows_per_cta(B)) * part_len(F1, F2, NC) * 4, 256);
}
// per-CTA partial sums + the bf16 image of W2
int64_t fc_tail_scratch_bytes(int B, int F1, int F2, int NC) {
  return part_bytes(B, F1, F2, NC) + round_up((int64_t)F1 * (F2 + 8) * 2, 256);
}

// the bf16 image of W2 inside `scratch` (same layout fc_tail expects); lets a caller build it early, off
// the critical path, and then call fc_tail(..., w2_image_ready = true)
int fc_tail_prepare_w2(const float* w2, void* scratch, int B, int F1, int F2, int NC, cudaStream_t st) {
  __nv_bfloat16* w2b = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(scratch) + part_bytes(B, F1, F2, NC));
  B200_CUDA_TRY(launch_kernel(w2_to_bf16_kernel, dim3(cdiv(F1 * (F2 + 8), 256)), dim3(256), 0, st, w2, w2b, F1, F2));
  return B200_OK;
}

int fc_tail(const float* fc1_part, int n_part, const float* b1, float* full1, const float* w2, const float* b2,
            const float* w3, const float* b3, const int32_t* labels, float* full2, float* logits, float* dlogits,
            float* dfull2, float* dfull1, int32_t* pred, float* loss, int32_t* n_correct, int32_t* n_bad, float* dw2,
            float* db2, float* dw3, float* db3, float* db1, void* scratch, float* zero_ptr, int zero_n, int B, int F1,
            int F2, int NC, cudaStream_t st, bool w2_image_ready, cudaStream_t grads_stream, cudaEvent_t rows_done,
            bool overlap_producer, void* dfull1_img) {
  B200_REQUIRE(B > 0 && NC > 0 && NC <= kMaxCls, "fc_tail: bad shape");
  B200_REQUIRE(F1 % KT == 0 && F2 % 32 == 0 && F2 <= RT && F2 > 0, "fc_tail: needs f1 %% 32 == 0, f2 %% 32 == 0, f2 <= 256");
  const int R = rows_per_cta(B);
  const int nrc = cdiv(B, R);
  float* part = reinterpret_cast<float*>(scratch);
  __nv_bfloat16* w2b = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(scratch) + part_bytes(B, F1, F2, NC));
  if (!w2_image_ready) {
    w2_to_bf16_kernel<<<cdiv(F1 * (F2 + 8), 256), 256, 0, st>>>(w2, w2b, F1, F2);
    B200_LAUNCH_CHECK();
  }
  TailArgs a{B, F1, F2, NC, R, fc1_part, n_part, b1, full1, w2, b2, w3, b3, w2b, labels, full2, logits, dlogits,
             dfull2, dfull1, pred, part, zero_ptr, zero_n, g_tail_timeline, reinterpret_cast<uint8_t*>(dfull1_img)};
  const int RM = R <= 8 ? 8 : 16;              // rows of the MMA tile that carry batch rows
  const size_t smem = (size_t)tail_smem(RM, F1, F2, NC).total;
  B200_REQUIRE(smem <= 220 * 1024, "fc_tail: layer widths too large (%zu bytes of shared memory)", smem);
  static PerDeviceOnce grads_attr, rows_attr[2][5];
  if (grads_attr.first()) {
    B200_CUDA_TRY(cudaFuncSetAttribute(fc_tail_grads_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    B200_CUDA_TRY(prefer_max_shared(fc_tail_grads_kernel));
    B200_CUDA_TRY(prefer_max_shared(w2_to_bf16_kernel));
  }
  const int t1 = cdiv(F2, 64);
  // overlap_producer: the kernel in front of this one on `st` produced fc1_part and allows its dependents to start
  // early (griddepcontrol.launch_dependents); the rows kernel then fetches W2 / W3 / labels while it still runs
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(nrc); cfg.blockDim = dim3(RT); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = overlap_producer ? 1 : 0;
#define B200_TAIL_LAUNCH(RM_, T1_)                                                                                   \
  do {                                                                                                                 \
    if (rows_attr[RM_ / 16][T1_].first())                                                                              \
      B200_CUDA_TRY(cudaFuncSetAttribute(fc_tail_rows_kernel<RM_, T1_>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                         220 * 1024));                                                                 \
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, fc_tail_rows_kernel<RM_, T1_>, a));                                         \
  } while (0)
  if (RM == 8) {
    switch (t1) {
      case 1: B200_TAIL_LAUNCH(8, 1); break;
      case 2: B200_TAIL_LAUNCH(8, 2); break;
      case 3: B200_TAIL_LAUNCH(8, 3); break;
      default: B200_TAIL_LAUNCH(8, 4); break;
    }
  } else {
    switch (t1) {
      case 1: B200_TAIL_LAUNCH(16, 1); break;
      case 2: B200_TAIL_LAUNCH(16, 2); break;
      case 3: B200_TAIL_LAUNCH(16, 3); break;
      default: B200_TAIL_LAUNCH(16, 4); break;
    }
  }
#undef B200_TAIL_LAUNCH
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
