// fp32 "validation mode" kernels: plain SIMT CUDA-core arithmetic for every op on the
// cifar10cnn.py training step.  They exist so that logits / loss / gradients can be compared
// with the fp64 oracle at accumulation-order tolerance (tcgen05 has no fp32-input MMA); the
// bf16 tensor-core path reuses the memory-bound ones (pool, head, update).
//
// One generic 64x64x16 register-tiled GEMM kernel serves conv forward / dgrad / wgrad (as an
// implicit GEMM through gather loaders) and the FC forward / backward products.
#include <algorithm>

#include <cuda_bf16.h>

#include "launch.h"

namespace b200 {

// ------------------------------------------------------------------------------------------
// read_cifar_files crop + transpose + cast (cifar10cnn.py:57-68)
// ------------------------------------------------------------------------------------------
__global__ void crop_cast_kernel(const uint8_t* __restrict__ img, float* __restrict__ out, int B, int ih,
                                 int iw, int ch, int cw, int cin) {
  const int oh = (ih - ch) / 2, ow = (iw - cw) / 2;
  const int64_t total = (int64_t)B * ch * cw * cin;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % cin);
    int64_t t = i / cin;
    int x = (int)(t % cw);
    t /= cw;
    int y = (int)(t % ch);
    int b = (int)(t / ch);
    out[i] = (float)img[(((int64_t)b * cin + c) * ih + (y + oh)) * iw + (x + ow)];
  }
}

int crop_cast_f32(const uint8_t* img, float* out, int B, int ih, int iw, int ch, int cw, int cin,
                  cudaStream_t st) {
  B200_REQUIRE(B > 0 && ch <= ih && cw <= iw && ch > 0 && cw > 0 && cin > 0, "crop_cast: bad shape");
  int64_t total = (int64_t)B * ch * cw * cin;
  int grid = (int)std::min<int64_t>((total + 255) / 256, kNumSMs * 8);
  crop_cast_kernel<<<grid, 256, 0, st>>>(img, out, B, ih, iw, ch, cw, cin);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// ------------------------------------------------------------------------------------------
// Generic fp32 GEMM  C[M,N] = A[M,K] * B[K,N]  with functor loaders and epilogue.
// 256 threads, 64x64 tile, BK=16, each thread a 4x4 block.  blockIdx.z = split-K slice.
// ------------------------------------------------------------------------------------------
struct ALoadDenseK {     // A(m,k) = p[m*ld + k]   (k contiguous)
  const float* p; int64_t ld;
  static constexpr bool kContigK = true;
  __device__ __forceinline__ float operator()(int m, int k) const { return __ldg(p + (int64_t)m * ld + k); }
};
struct ALoadDenseM {     // A(m,k) = p[k*ld + m]   (m contiguous; a transposed operand)
  const float* p; int64_t ld;
  static constexpr bool kContigK = false;
  __device__ __forceinline__ float operator()(int m, int k) const { return __ldg(p + (int64_t)k * ld + m); }
};
struct Im2col {          // patch element (pixel, tap*C + c) of an NHWC tensor, SAME 5x5 zero pad
  const float* x; int H, W, C;
  __device__ __forceinline__ float at(int pix, int kk) const {
    int c = kk % C, tap = kk / C;
    int di = tap / 5, dj = tap - di * 5;
    int j = pix % W, t = pix / W;
    int i = t % H, n = t / H;
    int ii = i + di - 2, jj = j + dj - 2;
    if ((unsigned)ii >= (unsigned)H || (unsigned)jj >= (unsigned)W) return 0.f;
    return __ldg(x + (((int64_t)n * H + ii) * W + jj) * C + c);
  }
};
struct ALoadIm2col {     // A(m = pixel, k = tap*C+c): conv forward / dgrad
  Im2col g;
  static constexpr bool kContigK = true;
  __device__ __forceinline__ float operator()(int m, int k) const { return g.at(m, k); }
};
struct ALoadIm2colT {    // A(m = tap*C+c, k = pixel): conv wgrad
  Im2col g;
  static constexpr bool kContigK = false;
  __device__ __forceinline__ float operator()(int m, int k) const { return g.at(k, m); }
};
struct BLoadDenseN {     // B(k,n) = p[k*ld + n]   (n contiguous)
  const float* p; int64_t ld;
  static constexpr bool kContigN = true;
  __device__ __forceinline__ float operator()(int k, int n) const { return __ldg(p + (int64_t)k * ld + n); }
};
struct BLoadDenseK {     // B(k,n) = p[n*ld + k]   (k contiguous; a transposed operand)
  const float* p; int64_t ld;
  static constexpr bool kContigN = false;
  __device__ __forceinline__ float operator()(int k, int n) const { return __ldg(p + (int64_t)n * ld + k); }
};
struct Epilogue {
  float* out; int64_t ld;          // out[z*split_stride + m*ld + n]
  int64_t split_stride;
  const float* bias;               // per n, may be null
  int relu;
  const float* mask; int64_t ldmask;   // multiply by [mask(m,n) > 0], may be null
};

template <class AL, class BL>
__global__ void __launch_bounds__(256) gemm_f32_kernel(int M, int N, int K, int k_per_split, AL a, BL b,
                                                       Epilogue ep) {
  __shared__ __align__(16) float As[16][68];
  __shared__ __align__(16) float Bs[16][68];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  const int kbeg = blockIdx.z * k_per_split;
  const int kend = min(K, kbeg + k_per_split);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = kbeg; k0 < kend; k0 += 16) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      int mm, kk;
      if (AL::kContigK) { kk = tid & 15; mm = (tid >> 4) + 16 * r; }
      else              { mm = tid & 63; kk = (tid >> 6) + 4 * r; }
      float v = 0.f;
      if (m0 + mm < M && k0 + kk < kend) v = a(m0 + mm, k0 + kk);
      As[kk][mm] = v;
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      int nn, kk;
      if (BL::kContigN) { nn = tid & 63; kk = (tid >> 6) + 4 * r; }
      else              { kk = tid & 15; nn = (tid >> 4) + 16 * r; }
      float v = 0.f;
      if (n0 + nn < N && k0 + kk < kend) v = b(k0 + kk, n0 + nn);
      Bs[kk][nn] = v;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float4 av = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      float4 bv = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      float aa[4] = {av.x, av.y, av.z, av.w};
      float bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(aa[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* out = ep.out + (int64_t)blockIdx.z * ep.split_stride;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      float v = acc[i][j];
      if (ep.bias) v += __ldg(ep.bias + n);
      if (ep.relu) v = fmaxf(v, 0.f);
      if (ep.mask) v = (__ldg(ep.mask + (int64_t)m * ep.ldmask + n) > 0.f) ? v : 0.f;
      out[(int64_t)m * ep.ld + n] = v;
    }
  }
}

template <class AL, class BL>
static int launch_gemm(int M, int N, int K, int splits, AL a, BL b, Epilogue ep, cudaStream_t st) {
  int kps = (int)round_up(cdiv(K, splits), 16);
  dim3 grid(cdiv(N, 64), cdiv(M, 64), cdiv(K, kps));
  gemm_f32_kernel<AL, BL><<<grid, 256, 0, st>>>(M, N, K, kps, a, b, ep);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// sum of S partial buffers of n floats (deterministic order)
__global__ void reduce_partials_kernel(const float* __restrict__ part, float* __restrict__ out, int64_t n,
                                       int S) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int z = 0; z < S; ++z) s += part[(int64_t)z * n + i];
    out[i] = s;
  }
}

// ------------------------------------------------------------------------------------------
// conv2d SAME 5x5 stride 1 (+bias +relu), cifar10cnn.py:107-110 / :118-121
// ------------------------------------------------------------------------------------------
int conv5x5_fwd_f32(const float* x, const float* w, const float* bias, float* y, int B, int H, int W,
                    int Cin, int Cout, int relu, cudaStream_t st) {
  B200_REQUIRE(B > 0 && H > 0 && W > 0 && Cin > 0 && Cout > 0, "conv5x5_fwd: bad shape");
  B200_REQUIRE((int64_t)B * H * W < (1ll << 31), "conv5x5_fwd: too many pixels");
  ALoadIm2col a{{x, H, W, Cin}};
  BLoadDenseN b{w, Cout};
  Epilogue ep{y, Cout, 0, bias, relu, nullptr, 0};
  return launch_gemm(B * H * W, Cout, 25 * Cin, 1, a, b, ep, st);
}

// wr[e,f,k,c] = w[4-e,4-f,c,k]: the filter Conv2DBackpropInput correlates dy with.
__global__ void rot180_kernel(const float* __restrict__ w, float* __restrict__ wr, int Cin, int Cout) {
  int total = 25 * Cin * Cout;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    int c = i % Cin;
    int t = i / Cin;
    int k = t % Cout;
    int tap = t / Cout;
    wr[i] = w[((24 - tap) * Cin + c) * Cout + k];
  }
}
int conv5x5_rot180_f32(const float* w, float* wr, int Cin, int Cout, cudaStream_t st) {
  rot180_kernel<<<cdiv(25 * Cin * Cout, 256), 256, 0, st>>>(w, wr, Cin, Cout);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

static int wgrad_splits(int B, int H, int W, int Cin, int Cout) {
  int tiles = cdiv(25 * Cin, 64) * cdiv(Cout, 64);
  int64_t K = (int64_t)B * H * W;
  int s = cdiv(2 * kNumSMs, tiles);
  int smax = (int)std::max<int64_t>(1, K / 256);
  return max(1, min(s, min(smax, 256)));
}
int64_t conv5x5_wgrad_scratch_bytes(int B, int H, int W, int Cin, int Cout) {
  int s = wgrad_splits(B, H, W, Cin, Cout);
  return (int64_t)s * 25 * Cin * Cout * 4 + colsum_scratch_bytes(B * H * W, Cout);
}
int conv5x5_wgrad_f32(const float* x, const float* dy, float* dw, float* db, float* scratch, int B, int H,
                      int W, int Cin, int Cout, cudaStream_t st) {
  B200_REQUIRE((int64_t)B * H * W < (1ll << 31), "conv5x5_wgrad: too many pixels");
  int K = B * H * W, M = 25 * Cin;
  int s = wgrad_splits(B, H, W, Cin, Cout);
  int kps = (int)round_up(cdiv(K, s), 16);
  int nz = cdiv(K, kps);
  ALoadIm2colT a{{x, H, W, Cin}};
  BLoadDenseN b{dy, Cout};
  Epilogue ep{scratch, Cout, (int64_t)M * Cout, nullptr, 0, nullptr, 0};
  int rc = launch_gemm(M, Cout, K, s, a, b, ep, st);
  if (rc) return rc;
  int64_t n = (int64_t)M * Cout;
  reduce_partials_kernel<<<cdiv(n, 256), 256, 0, st>>>(scratch, dw, n, nz);
  B200_LAUNCH_CHECK();
  if (db) return colsum_f32(dy, db, scratch + (int64_t)s * M * Cout, K, Cout, st);
  return B200_OK;
}

// ------------------------------------------------------------------------------------------
// column sums (BiasAddGrad / Sum axis 0): two deterministic passes
// ------------------------------------------------------------------------------------------
static int colsum_chunks(int rows) { return max(1, min(2 * kNumSMs, rows / 64)); }
int64_t colsum_scratch_bytes(int rows, int cols) { return (int64_t)colsum_chunks(rows) * cols * 4; }

__global__ void colsum_partial_kernel(const float* __restrict__ a, float* __restrict__ part, int rows,
                                      int cols, int rows_per_chunk) {
  int r0 = blockIdx.x * rows_per_chunk, r1 = min(rows, r0 + rows_per_chunk);
  for (int c = threadIdx.x; c < cols; c += blockDim.x) {
    float s = 0.f;
    for (int r = r0; r < r1; ++r) s += a[(int64_t)r * cols + c];
    part[(int64_t)blockIdx.x * cols + c] = s;
  }
}
int colsum_f32(const float* a, float* out, float* scratch, int rows, int cols, cudaStream_t st) {
  int chunks = colsum_chunks(rows);
  int rpc = cdiv(rows, chunks);
  chunks = cdiv(rows, rpc);
  colsum_partial_kernel<<<chunks, min(256, (int)round_up(cols, 32)), 0, st>>>(a, scratch, rows, cols, rpc);
  B200_LAUNCH_CHECK();
  reduce_partials_kernel<<<cdiv(cols, 256), 256, 0, st>>>(scratch, out, cols, chunks);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// ------------------------------------------------------------------------------------------
// max_pool k3 s2 SAME (cifar10cnn.py:113 / :123): out = ceil(in/2), pad 0 before / rest after,
// padded cells never win, first maximal cell in (h,w) scan order is recorded.
// ------------------------------------------------------------------------------------------
__global__ void maxpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                   uint8_t* __restrict__ arg, int B, int H, int W, int C, int OH, int OW,
                                   int pbh, int pbw) {
  const int64_t total = (int64_t)B * OH * OW * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % C);
    int64_t t = i / C;
    int ow = (int)(t % OW);
    t /= OW;
    int oh = (int)(t % OH);
    int b = (int)(t / OH);
    float best = -INFINITY;
    int bi = 0;
    bool any = false;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      int h = 2 * oh + kh - pbh;
      if ((unsigned)h >= (unsigned)H) continue;
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        int w = 2 * ow + kw - pbw;
        if ((unsigned)w >= (unsigned)W) continue;
        float v = x[(((int64_t)b * H + h) * W + w) * C + c];
        if (!any || v > best) { best = v; bi = kh * 3 + kw; any = true; }
      }
    }
    y[i] = best;
    arg[i] = (uint8_t)bi;
  }
}
static void pool_geom(int H, int W, int& OH, int& OW, int& pbh, int& pbw) {
  OH = (H + 1) / 2; OW = (W + 1) / 2;
  int pth = max((OH - 1) * 2 + 3 - H, 0), ptw = max((OW - 1) * 2 + 3 - W, 0);
  pbh = pth / 2; pbw = ptw / 2;
}
int maxpool_fwd_f32(const float* x, float* y, uint8_t* arg, int B, int H, int W, int C, cudaStream_t st) {
  B200_REQUIRE(B > 0 && H > 1 && W > 1 && C > 0, "maxpool_fwd: bad shape");
  int OH, OW, pbh, pbw;
  pool_geom(H, W, OH, OW, pbh, pbw);
  int64_t total = (int64_t)B * OH * OW * C;
  int grid = (int)std::min<int64_t>((total + 255) / 256, kNumSMs * 16);
  maxpool_fwd_kernel<<<grid, 256, 0, st>>>(x, y, arg, B, H, W, C, OH, OW, pbh, pbw);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// MaxPoolGrad in gather form (no atomics): input cell (h,w) collects dy of every window whose
// recorded argmax is this cell; optionally multiplied by the ReluGrad mask [x > 0] of the
// tensor that was pooled.
__global__ void maxpool_bwd_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ arg,
                                   const float* __restrict__ x, float* __restrict__ dx, int B, int H, int W,
                                   int C, int OH, int OW, int pbh, int pbw, int relu_mask,
                                   float* __restrict__ colsum, __nv_bfloat16* __restrict__ dx_bf16) {
  const int64_t total = (int64_t)B * H * W * C;
  float csum = 0.f;   // with 256 % C == 0 a thread keeps one channel over the whole grid-stride loop
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % C);
    int64_t t = i / C;
    int w = (int)(t % W);
    t /= W;
    int h = (int)(t % H);
    int b = (int)(t / H);
    float s = 0.f;
    if (!relu_mask || x[i] > 0.f) {
      int hp = h + pbh, wp = w + pbw;            // padded coordinates
#pragma unroll
      for (int kh = 0; kh < 3; ++kh) {
        int r = hp - kh;
        if (r < 0 || (r & 1)) continue;
        int oh = r >> 1;
        if (oh >= OH) continue;
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
          int q = wp - kw;
          if (q < 0 || (q & 1)) continue;
          int ow = q >> 1;
          if (ow >= OW) continue;
          int64_t o = (((int64_t)b * OH + oh) * OW + ow) * C + c;
          const float d = __ldg(dy + o);                 // unconditional: the (<= 4) loads stay independent
          s += (__ldg(arg + o) == kh * 3 + kw) ? d : 0.f;
        }
      }
    }
    dx[i] = s;
    if (dx_bf16) dx_bf16[i] = __float2bfloat16_rn(s);      // operand copy for the tensor-core filter gradient
    csum += s;
  }
  if (colsum) {       // BiasAddGrad of the layer below, fused: per-CTA partial -> one atomic per channel
    __shared__ float red[256];
    red[threadIdx.x] = csum;
    __syncthreads();
    if ((int)threadIdx.x < C) {
      float t = 0.f;
      for (int j = threadIdx.x; j < 256; j += C) t += red[j];
      atomicAdd(colsum + threadIdx.x, t);
    }
  }
}
int maxpool_bwd_f32(const float* dy, const uint8_t* arg, const float* x, float* dx, int B, int H, int W,
                    int C, int relu_mask, cudaStream_t st, float* colsum_atomic, void* dx_bf16) {
  B200_REQUIRE(!colsum_atomic || (256 % C) == 0, "maxpool_bwd: fused bias gradient needs C to divide 256");
  B200_REQUIRE(B > 0 && H > 1 && W > 1 && C > 0, "maxpool_bwd: bad shape");
  B200_REQUIRE(!relu_mask || x, "maxpool_bwd: relu_mask needs x");
  int OH, OW, pbh, pbw;
  pool_geom(H, W, OH, OW, pbh, pbw);
  int64_t total = (int64_t)B * H * W * C;
  int grid = (int)std::min<int64_t>((total + 255) / 256, kNumSMs * 16);
  maxpool_bwd_kernel<<<grid, 256, 0, st>>>(dy, arg, x, dx, B, H, W, C, OH, OW, pbh, pbw, relu_mask, colsum_atomic,
                                           reinterpret_cast<__nv_bfloat16*>(dx_bf16));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// ------------------------------------------------------------------------------------------
// fully connected layers, cifar10cnn.py:133 / :139 / :145
// ------------------------------------------------------------------------------------------
int fc_fwd_f32(const float* x, const float* w, const float* b, float* y, int B, int K, int N, int relu,
               cudaStream_t st) {
  B200_REQUIRE(B > 0 && K > 0 && N > 0, "fc_fwd: bad shape");
  ALoadDenseK a{x, K};
  BLoadDenseN bl{w, N};
  Epilogue ep{y, N, 0, b, relu, nullptr, 0};
  return launch_gemm(B, N, K, 1, a, bl, ep, st);
}

int fc_bwd_f32(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db,
               float* scratch, int B, int K, int N, int mask_x, cudaStream_t st) {
  B200_REQUIRE(B > 0 && K > 0 && N > 0, "fc_bwd: bad shape");
  int rc;
  {  // dw[K,N] = x^T[K,B] * dy[B,N]
    ALoadDenseM a{x, K};
    BLoadDenseN bl{dy, N};
    Epilogue ep{dw, N, 0, nullptr, 0, nullptr, 0};
    if ((rc = launch_gemm(K, N, B, 1, a, bl, ep, st))) return rc;
  }
  if (db && (rc = colsum_f32(dy, db, scratch, B, N, st))) return rc;
  if (dx) {  // dx[B,K] = dy[B,N] * w^T[N,K]  (* [x > 0])
    ALoadDenseK a{dy, N};
    BLoadDenseK bl{w, N};
    Epilogue ep{dx, K, 0, nullptr, 0, mask_x ? x : nullptr, K};
    if ((rc = launch_gemm(B, K, N, 1, a, bl, ep, st))) return rc;
  }
  return B200_OK;
}

// ------------------------------------------------------------------------------------------
// cifar_loss (:150-157) + batch_accuracy (:166-176) + dlogits.  One CTA, one thread per row
// (n_cls <= 16 values in registers), block reduction with warp shuffles -> deterministic.
// ------------------------------------------------------------------------------------------
constexpr int kMaxCls = 16;
__global__ void __launch_bounds__(1024) head_kernel(const float* __restrict__ logits,
                                                    const int32_t* __restrict__ labels,
                                                    float* __restrict__ loss_mean,
                                                    float* __restrict__ dlogits, int32_t* __restrict__ pred,
                                                    int32_t* __restrict__ n_correct,
                                                    int32_t* __restrict__ n_bad, int B, int ncls) {
  float loss_acc = 0.f;
  int ok_acc = 0, bad_acc = 0;
  const float invB = 1.f / (float)B;
  for (int r = threadIdx.x; r < B; r += blockDim.x) {
    float z[kMaxCls];
    float m = -INFINITY;
    int am = 0;
#pragma unroll
    for (int c = 0; c < kMaxCls; ++c) {
      if (c < ncls) {
        z[c] = logits[(int64_t)r * ncls + c];
        if (z[c] > m) { m = z[c]; am = c; }        // strict '>' keeps the first maximal index
      }
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kMaxCls; ++c)
      if (c < ncls) { z[c] = expf(z[c] - m); s += z[c]; }
    int y = labels[r];
    bool valid = (unsigned)y < (unsigned)ncls;
    float inv = 1.f / s;
    float zy = 0.f;
#pragma unroll
    for (int c = 0; c < kMaxCls; ++c) {
      if (c < ncls) {
        float lg = logits[(int64_t)r * ncls + c];
        if (c == y) zy = lg;
        float d = (z[c] * inv - (c == y ? 1.f : 0.f)) * invB;
        if (dlogits) dlogits[(int64_t)r * ncls + c] = (valid && lg > 0.f) ? d : 0.f;   // ReluGrad of :145
      }
    }
    if (valid) loss_acc += logf(s) - (zy - m);
    else bad_acc += 1;
    if (pred) pred[r] = am;
    ok_acc += (valid && am == y) ? 1 : 0;
  }
  __shared__ float sl[32];
  __shared__ int so[32], sb[32];
  for (int o = 16; o > 0; o >>= 1) {
    loss_acc += __shfl_down_sync(0xffffffffu, loss_acc, o);
    ok_acc += __shfl_down_sync(0xffffffffu, ok_acc, o);
    bad_acc += __shfl_down_sync(0xffffffffu, bad_acc, o);
  }
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { sl[warp] = loss_acc; so[warp] = ok_acc; sb[warp] = bad_acc; }
  __syncthreads();
  if (warp == 0) {
    int nw = blockDim.x >> 5;
    loss_acc = lane < nw ? sl[lane] : 0.f;
    ok_acc = lane < nw ? so[lane] : 0;
    bad_acc = lane < nw ? sb[lane] : 0;
    for (int o = 16; o > 0; o >>= 1) {
      loss_acc += __shfl_down_sync(0xffffffffu, loss_acc, o);
      ok_acc += __shfl_down_sync(0xffffffffu, ok_acc, o);
      bad_acc += __shfl_down_sync(0xffffffffu, bad_acc, o);
    }
    if (lane == 0) {
      if (loss_mean) *loss_mean = loss_acc * invB;
      if (n_correct) *n_correct = ok_acc;
      if (n_bad) *n_bad = bad_acc;
    }
  }
}
int head_f32(const float* logits, const int32_t* labels, float* loss_mean, float* dlogits, int32_t* pred,
             int32_t* n_correct, int32_t* n_bad, int B, int ncls, cudaStream_t st) {
  B200_REQUIRE(B > 0 && ncls > 0 && ncls <= kMaxCls, "head: n_cls must be in 1..%d", kMaxCls);
  int threads = (int)std::min<int64_t>(1024, round_up(B, 32));
  head_kernel<<<1, threads, 0, st>>>(logits, labels, loss_mean, dlogits, pred, n_correct, n_bad, B, ncls);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// ------------------------------------------------------------------------------------------
// ApplyGradientDescent (cifar10cnn.py:162-163): p -= lr * scale * g
// ------------------------------------------------------------------------------------------
__global__ void sgd_kernel(float* __restrict__ p, const float* __restrict__ g, int64_t n, float step) {
  B200_TRACE_SCOPE(TR_SGD);
  pdl_wait();
  int64_t n4 = n >> 2;
  float4* p4 = reinterpret_cast<float4*>(p);
  const float4* g4 = reinterpret_cast<const float4*>(g);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4;
       i += (int64_t)gridDim.x * blockDim.x) {
    float4 a = p4[i], b = g4[i];
    a.x = fmaf(-step, b.x, a.x); a.y = fmaf(-step, b.y, a.y);
    a.z = fmaf(-step, b.z, a.z); a.w = fmaf(-step, b.w, a.w);
    p4[i] = a;
  }
  for (int64_t i = (n4 << 2) + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    p[i] = fmaf(-step, g[i], p[i]);
}
// The same update when the first n_part gradients still sit in S per-CTA partials (conv1's filter / bias gradient,
// tc_conv1.cu): the leading CTAs sum them in a fixed order -- one WARP per 16 bytes: lane l takes partials l, l+32, ...
// (all of its loads in flight at once: one round trip to L2), then a five-step shuffle tree -- store the finished
// gradient (the buffer is complete after the step) and apply it; the other CTAs update the rest.  Saves the launch
// of a reduction kernel at the very end of the step's critical path.
__global__ void reduce_partials_warp_kernel(const float* __restrict__ part, float* __restrict__ out, int n, int S, int stride) {
  pdl_wait();
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = min(t >> 5, n / 4 - 1) * 4;                   // surplus warps redo the last group
  const float4 s = warp_sum_partials(part, S, stride, i, t & 31);
  if ((t & 31) == 0 && (t >> 5) < n / 4) *reinterpret_cast<float4*>(out + i) = s;
}
__global__ void sgd_reduce_kernel(float* __restrict__ p, float* __restrict__ g, int64_t n, float step,
                                  const float* __restrict__ part, int S, int n_part) {
  B200_TRACE_SCOPE(TR_SGD);
  pdl_wait();
  const int lead = (n_part / 4 * 32 + 255) / 256;             // one warp per float4 of the partial range
  if ((int)blockIdx.x < lead) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = t & 31;
    const int i = min(t >> 5, n_part / 4 - 1) * 4;            // surplus warps redo the last group
    const float4 s = warp_sum_partials(part, S, n_part, i, lane);
    if (lane == 0 && (t >> 5) < n_part / 4) {
      *reinterpret_cast<float4*>(g + i) = s;
      float4 a = *reinterpret_cast<float4*>(p + i);
      a.x = fmaf(-step, s.x, a.x); a.y = fmaf(-step, s.y, a.y);
      a.z = fmaf(-step, s.z, a.z); a.w = fmaf(-step, s.w, a.w);
      *reinterpret_cast<float4*>(p + i) = a;
    }
    return;
  }
  const int64_t n4 = n >> 2, first4 = n_part >> 2;
  float4* p4 = reinterpret_cast<float4*>(p);
  const float4* g4 = reinterpret_cast<const float4*>(g);
  for (int64_t i = first4 + (blockIdx.x - lead) * (int64_t)blockDim.x + threadIdx.x; i < n4;
       i += (int64_t)(gridDim.x - lead) * blockDim.x) {
    float4 a = p4[i], b = g4[i];
    a.x = fmaf(-step, b.x, a.x); a.y = fmaf(-step, b.y, a.y);
    a.z = fmaf(-step, b.z, a.z); a.w = fmaf(-step, b.w, a.w);
    p4[i] = a;
  }
}
// n and n_part multiples of 4, the partial range = the first n_part floats of the flat buffers
int sgd_apply_reduce_f32(float* p, float* g, int64_t n, float lr, float scale, const float* part, int S, int n_part,
                         cudaStream_t st) {
  B200_REQUIRE(n > 0 && n % 4 == 0 && n_part > 0 && n_part % 4 == 0 && n_part <= n && S > 0 && part,
               "sgd_apply_reduce: bad sizes");
  B200_REQUIRE(((uintptr_t)p & 15) == 0 && ((uintptr_t)g & 15) == 0 && ((uintptr_t)part & 15) == 0,
               "sgd_apply_reduce: pointers must be 16-byte aligned");
  const int lead = (n_part / 4 * 32 + 255) / 256;
  int grid = lead + (int)std::min<int64_t>(((n - n_part) / 4 + 255) / 256 + 1, kNumSMs * 4);
  static PerDeviceOnce carve;
  if (carve.first()) B200_CUDA_TRY(prefer_max_shared(sgd_reduce_kernel));
  B200_CUDA_TRY(launch_kernel(sgd_reduce_kernel, dim3(grid), dim3(256), 0, st, p, g, n, lr * scale, part, S, n_part));
  B200_LAUNCH_CHECK();
  return B200_OK;
}
// y[b][n] = relu?(sum over splits of part[s][b][n] + bias[n]): closes a split-K product (FC1 of the layer-by-layer schedule)
__global__ void splitk_finish_kernel(const float* __restrict__ part, int nsplit, int64_t stride, const float* __restrict__ bias,
                                     float* __restrict__ y, int64_t total4, int N, int relu) {
  pdl_wait();
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 s = __ldg(reinterpret_cast<const float4*>(bias + (i * 4) % N));
    for (int z = 0; z < nsplit; ++z) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(part + z * stride) + i);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    if (relu) { s.x = fmaxf(s.x, 0.f); s.y = fmaxf(s.y, 0.f); s.z = fmaxf(s.z, 0.f); s.w = fmaxf(s.w, 0.f); }
    reinterpret_cast<float4*>(y)[i] = s;
  }
}
int splitk_finish_f32(const float* part, int nsplit, const float* bias, float* y, int B, int N, int relu, cudaStream_t st) {
  B200_REQUIRE(part && bias && y && nsplit > 0 && B > 0 && N > 0 && N % 4 == 0, "splitk_finish: bad argument");
  const int64_t total4 = (int64_t)B * N / 4;
  const int grid = (int)std::min<int64_t>((total4 + 255) / 256, kNumSMs * 4);
  B200_CUDA_TRY(launch_kernel(splitk_finish_kernel, dim3(grid), dim3(256), 0, st, part, nsplit, (int64_t)B * N, bias, y, total4, N, relu));
  return B200_OK;
}
int reduce_partials_warp(const float* part, float* out, int n, int S, int stride, cudaStream_t st) {
  B200_REQUIRE(part && out && n > 0 && n % 4 == 0 && S > 0 && stride % 4 == 0, "reduce_partials_warp: bad argument");
  B200_CUDA_TRY(launch_kernel(reduce_partials_warp_kernel, dim3((n / 4 * 32 + 255) / 256), dim3(256), 0, st, part, out, n, S, stride));
  return B200_OK;
}
int sgd_apply_f32(float* p, const float* g, int64_t n, float lr, float scale, cudaStream_t st) {
  B200_REQUIRE(n >= 0, "sgd_apply: negative size");
  B200_REQUIRE(((uintptr_t)p & 15) == 0 && ((uintptr_t)g & 15) == 0, "sgd_apply: pointers must be 16-byte aligned");
  if (n == 0) return B200_OK;
  int grid = (int)std::min<int64_t>((n / 4 + 255) / 256 + 1, kNumSMs * 4);
  static PerDeviceOnce carve;
  if (carve.first()) B200_CUDA_TRY(prefer_max_shared(sgd_kernel));
  B200_CUDA_TRY(launch_kernel(sgd_kernel, dim3(grid), dim3(256), 0, st, p, g, n, lr * scale));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_TRACE_SETTER(trace_set_f32)

}  // namespace b200
