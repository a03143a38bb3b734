// tcgen05 implicit-GEMM kernels for the 5x5 SAME convolutions of cifar10cnn.py (:107, :118) and
// their gradients (Conv2DBackpropInput / Conv2DBackpropFilter), bf16 operands, fp32 accumulation
// in TMEM.
//
// THE STRIP.  An image (or a few) is kept in shared memory as a zero-padded "strip": pixel (i,j)
// of image g sits in pixel slot  lead + (g*(H+2) + i)*P + j  with row pitch P = W+4, so every row
// is followed by 4 zero pixels (right pad of this row = left pad of the next) and every image by
// 2 zero rows.  Channels are split in chunks of 8 (16 B): byte address = chunk*CH_STRIDE + slot*16.
// That is exactly the no-swizzle canonical K-major operand layout with 8 consecutive pixel slots
// as a core matrix -- and because the layout is unswizzled, a filter tap (di,dj) is nothing but the
// start-address offset ((di-2)*P + (dj-2))*16 of the A descriptor.  The 25 taps x C/16 k-steps of
// the implicit GEMM therefore all read ONE resident copy of the input; no im2col is ever built.
// GEMM rows are pixel slots (pad slots compute garbage that the epilogue drops: 224/144 = 1.56x
// MMA work for 12x12 images), M = 128 rows per MMA at full tensor rate.
//
// The same strip read as an MN-major operand (8 channels contiguous, pixel slots along K) feeds
// the filter gradient: dW[tap] = X_strip(shifted by tap)^T * dY_strip, where the zero pads of the
// dY strip cancel every out-of-image product.
#include <cuda.h>
#include <stdlib.h>

#include <algorithm>

#include "launch.h"
#include "tc_common.cuh"

namespace b200 {
using namespace tc;

namespace {
constexpr int LOADERS = 128;
constexpr int WSTAGES = 3;      // filter stages in shared memory when a stage = the five taps of one filter row
constexpr int MAX_WSTAGES = 10; // ... and the most single-tap stages (two-CTAs-per-SM and wide variants)

struct ConvGeom {
  int B, H, W, KC, NC;      // batch, image size, contraction channels, output channels
  int P, rows_per_img;      // row pitch (W+4), H+2
  int ipc, tiles;           // images per CTA, 128-row M tiles per CTA
  int lead, npix;           // zero pixels before the first image row, total pixel slots
  int ch_stride;            // bytes between channel chunks of the strip (npix*16 + 16)
  int nstages;              // filter stages in the ring (of 5 taps or 1 tap each, a template parameter of the kernel)
  int tma_rows;             // > 0: the strip is written by the TMA (tensor-map box of P x tma_rows pixel slots per chunk)
  int tma_h0;               // first image row of a TMA box: -3 (one image: the box brings its own zero rows above) or 0
                            // (several images: boxes of H+2 rows tile the strip; the three zero rows above the first
                            // image are written once by the loader warps)
  int two_pass;             // KSPLIT == 2 with a TMA strip: the strip holds HALF the contraction channels at a time -- all
                            // 25 taps of channels [0, KC/2), a refill, all 25 taps of the rest -- which leaves twice the
                            // shared memory for filter stages in flight (the filter stream is latency-bound)
};

__host__ inline ConvGeom make_geom(int B, int H, int W, int KC, int NC, int ipc) {
  ConvGeom g;
  g.B = B; g.H = H; g.W = W; g.KC = KC; g.NC = NC;
  g.P = W + 4; g.rows_per_img = H + 2; g.ipc = ipc;
  int mtot = ipc * g.rows_per_img * g.P;
  g.tiles = (mtot + 127) / 128;
  g.lead = 2 * g.P + 2;
  g.npix = g.lead + g.tiles * 128 + 2 * g.P + 2;
  g.ch_stride = g.npix * 16 + 16;
  g.nstages = WSTAGES;
  g.tma_rows = 0;
  g.tma_h0 = 0;
  g.two_pass = 0;
  return g;
}
// The strip of ONE image written by the TMA: per 8-channel chunk one tensor-map box of P = W+4 columns x H+5 rows
// (image rows -3 .. H+1) of the dense bf16 NHWC tensor.  Everything outside the image is out of bounds for the tensor
// map and arrives as zeros, so the box IS the padded strip -- right pad of a row = left pad of the next, three zero
// rows above (two for the taps, one so that the slot "row -2, column -2" of tap (0,0) is still inside the box) and two
// below: no zeroing pass, no loads through registers.  Rows of the MMA tiles past the box read whatever follows in
// shared memory; they only feed output rows the epilogue drops (GEMM rows are independent).
// Several images per CTA (wide layers): one box of H+2 rows per image and chunk -- the image and the two zero rows
// below it, which are also the zero rows above the next image -- back to back from slot `lead` = 3 rows of zeros that the
// loader warps write once.
__host__ inline ConvGeom make_geom_tma(int B, int H, int W, int KC, int NC, int min_strip_bytes, int ipc = 1) {
  ConvGeom g = make_geom(B, H, W, KC, NC, ipc);
  g.lead = 3 * g.P;
  if (ipc == 1) {
    g.tma_rows = H + 5;
    g.tma_h0 = -3;
    g.npix = g.tma_rows * g.P;
  } else {
    g.tma_rows = H + 2;
    g.tma_h0 = 0;
    g.npix = g.lead + ipc * g.rows_per_img * g.P;
  }
  g.ch_stride = (g.npix * 16 + 127) & ~127;                         // TMA destinations are 128-byte aligned
  const int chunks = KC / 8;
  while (chunks * g.ch_stride < min_strip_bytes) g.ch_stride += 128;    // the fused pool parks its fp32 tile here
  return g;
}

// zero `bytes` (multiple of 16) of shared memory with the 128 loader threads
__device__ __forceinline__ void zero_smem(uint8_t* p, int bytes, int tid) {
  for (int o = tid * 16; o < bytes; o += LOADERS * 16) *reinterpret_cast<uint4*>(p + o) = make_uint4(0, 0, 0, 0);
}
__device__ __forceinline__ void loaders_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// copy the valid pixels of images [n0, n0+ipc) of a dense NHWC fp32 tensor into a strip (bf16)
__device__ __forceinline__ void fill_strip(uint8_t* strip, const float* __restrict__ x, const ConvGeom& g, int n0,
                                           int C, int ch_stride, int tid) {
  const int cpp = C >> 3;                       // chunks per pixel
  const int per_img = g.H * g.W * cpp;
  for (int img = 0; img < g.ipc; ++img) {
    const int n = n0 + img;
    if (n >= g.B) break;
    const float* src = x + (int64_t)n * g.H * g.W * C;
#pragma unroll 4
    for (int q = tid; q < per_img; q += LOADERS) {
      int c = q % cpp, pix = q / cpp;
      int j = pix % g.W, i = pix / g.W;
      uint4 v = load_chunk(src + (int64_t)pix * C + c * 8);
      int slot = g.lead + (img * g.rows_per_img + i) * g.P + j;
      *reinterpret_cast<uint4*>(strip + c * ch_stride + slot * 16) = v;
    }
  }
}

// the same from a tensor that is already bf16: one 16-byte load per chunk, no conversion
__device__ __forceinline__ void fill_strip_bf16(uint8_t* strip, const __nv_bfloat16* __restrict__ x, const ConvGeom& g,
                                                int n0, int C, int ch_stride, int tid) {
  const int cpp = C >> 3;
  const int per_img = g.H * g.W * cpp;
  for (int img = 0; img < g.ipc; ++img) {
    const int n = n0 + img;
    if (n >= g.B) break;
    const uint4* src = reinterpret_cast<const uint4*>(x + (int64_t)n * g.H * g.W * C);
#pragma unroll 9
    for (int q = tid; q < per_img; q += LOADERS) {
      int c = q % cpp, pix = q / cpp;
      int j = pix % g.W, i = pix / g.W;
      uint4 v = __ldg(src + q);
      int slot = g.lead + (img * g.rows_per_img + i) * g.P + j;
      *reinterpret_cast<uint4*>(strip + c * ch_stride + slot * 16) = v;
    }
  }
}

// ---- a dY strip built straight from the pooled gradient -------------------------------------------
// dconv = MaxPoolGrad(dpool, arg) * ReluGrad, the gradient w.r.t. the conv output that fed a 3x3/2 SAME
// max-pool (cifar10cnn.py:123), never materialised in HBM: the pooled gradient (masked by pool > 0, the
// ReluGrad of the selected cell) and the argmax of one image are staged in shared memory and every strip
// pixel gathers the <= 4 windows that cover it.  64 channels.
struct PoolGradSrc {
  const float* dpool;     // [B,OH,OW,64]
  const float* pool;      // [B,OH,OW,64]
  const uint8_t* arg;     // [B,OH,OW,64]
  int OH, OW, pbh, pbw;
};
__device__ __forceinline__ void stage_pool_grad(float* dpm, uint8_t* argb, const PoolGradSrc& p, int n, int tid) {
  const int ncell = p.OH * p.OW;
  const float4* dp4 = reinterpret_cast<const float4*>(p.dpool + (size_t)n * ncell * 64);
  const float4* pl4 = reinterpret_cast<const float4*>(p.pool + (size_t)n * ncell * 64);
  const uint32_t* ar4 = reinterpret_cast<const uint32_t*>(p.arg + (size_t)n * ncell * 64);
#pragma unroll 3
  for (int q = tid; q < ncell * 16; q += LOADERS) {
    float4 d = __ldg(dp4 + q), v = __ldg(pl4 + q);
    d.x = v.x > 0.f ? d.x : 0.f; d.y = v.y > 0.f ? d.y : 0.f;
    d.z = v.z > 0.f ? d.z : 0.f; d.w = v.w > 0.f ? d.w : 0.f;
    reinterpret_cast<float4*>(dpm)[q] = d;
    reinterpret_cast<uint32_t*>(argb)[q] = __ldg(ar4 + q);
  }
}
// gathers image slot `img` of the strip; bsum accumulates this thread's channel chunk (tid & 7) for the bias gradient
__device__ __forceinline__ void fill_strip_from_pool(uint8_t* strip, const float* dpm, const uint8_t* argb,
                                                     const ConvGeom& g, const PoolGradSrc& p, int img, int ch_stride,
                                                     int tid, float (&bsum)[8], float* __restrict__ dense_out = nullptr) {
  for (int q = tid; q < g.H * g.W * 8; q += LOADERS) {
    const int c8 = q & 7, pix = q >> 3;
    const int i = pix / g.W, j = pix - i * g.W;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int r = i + p.pbh - kh;
      if (r < 0 || (r & 1) || (r >> 1) >= p.OH) continue;
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int cc = j + p.pbw - kw;
        if (cc < 0 || (cc & 1) || (cc >> 1) >= p.OW) continue;
        const int cell = (r >> 1) * p.OW + (cc >> 1);
        const uint2 ab = *reinterpret_cast<const uint2*>(argb + cell * 64 + c8 * 8);
        const float4 d0 = *reinterpret_cast<const float4*>(dpm + cell * 64 + c8 * 8);
        const float4 d1 = *reinterpret_cast<const float4*>(dpm + cell * 64 + c8 * 8 + 4);
        const uint32_t code = kh * 3 + kw;
        const float dv[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const uint32_t ae = ((e < 4 ? ab.x : ab.y) >> (8 * (e & 3))) & 0xffu;
          if (ae == code) acc[e] += dv[e];
        }
      }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) bsum[e] += acc[e];
    const int slot = g.lead + (img * g.rows_per_img + i) * g.P + j;
    *reinterpret_cast<uint4*>(strip + c8 * ch_stride + slot * 16) = pack8(acc);
    if (dense_out) {          // dense NHWC fp32 copy for a consumer that runs later (8 lanes = one pixel's 256 B)
      float4* o = reinterpret_cast<float4*>(dense_out + (size_t)pix * 64 + c8 * 8);
      o[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
      o[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
    }
  }
}

// stage_pool_grad with NT threads and named barrier 3 (conv2's dgrad uses all eight non-TMA warps for its strip)
template <int NT>
__device__ __forceinline__ void nt_sync() { asm volatile("bar.sync 3, %0;" ::"n"(NT) : "memory"); }
template <int NT>
__device__ __forceinline__ void stage_pool_grad_nt(float* dpm, uint8_t* argb, const PoolGradSrc& p, int n, int tid) {
  const int ncell = p.OH * p.OW;
  const float4* dp4 = reinterpret_cast<const float4*>(p.dpool + (size_t)n * ncell * 64);
  const float4* pl4 = reinterpret_cast<const float4*>(p.pool + (size_t)n * ncell * 64);
  const uint32_t* ar4 = reinterpret_cast<const uint32_t*>(p.arg + (size_t)n * ncell * 64);
#pragma unroll 3
  for (int q = tid; q < ncell * 16; q += NT) {      // masked pooled gradient (ReluGrad of the selected cell) + argmax
    float4 d = __ldg(dp4 + q), v = __ldg(pl4 + q);
    d.x = v.x > 0.f ? d.x : 0.f; d.y = v.y > 0.f ? d.y : 0.f;
    d.z = v.z > 0.f ? d.z : 0.f; d.w = v.w > 0.f ? d.w : 0.f;
    reinterpret_cast<float4*>(dpm)[(q & ~15) + (q & 1) * 8 + ((q & 15) >> 1)] = d;      // layout of pool_grad_px
    reinterpret_cast<uint32_t*>(argb)[q] = __ldg(ar4 + q);
  }
}

struct ConvArgs {
  CUtensorMap tmap;    // g.tma_rows > 0: dense bf16 [B,H,W,KC] viewed as (8 channels, KC/8 chunks, W, H, B) for the strip boxes
  const float* x;      // [B,H,W,KC]
  const __nv_bfloat16* xb;   // the same, already bf16 (read instead of x when given)
  __nv_bfloat16* dconv_out_bf16;   // POOL dgrad: dense bf16 copy of the gathered dY (optional)
  uint8_t* pool_img;               // fused pool: bf16 operand image of the pooled output for FC1 (optional)
  const uint8_t* wp;   // filter packed by pack_conv_weights (bf16, ready-made shared-memory images per tap)
  int cin_f, cout_f;   // the FORWARD filter's channel counts (packing geometry)
  const float* bias;   // [NC] or null
  float* y;            // [B,H,W,NC]
  int relu;
  ConvGeom g;
  long long* timeline;  // optional device buffer [8]: clock64 stamps of CTA 0's phases (diagnostics), else null
  PoolGradSrc pg;       // POOL variant: x is unused, the strip is gathered from the pooled gradient
  float* dbias;         // POOL variant: += sum of dconv over pixels per channel (atomics; zero on entry), or null
  float* dbias_part;    // POOL variant: instead, this CTA's sum goes to dbias_part[blockIdx.x][64] (summed in a fixed
                        // order by the caller: run-to-run identical results), or null
  float* dconv_out;     // POOL variant: also store dconv as dense [B,H,W,64] fp32 (for the filter gradient), or null
  float* pool_out;      // forward, ipc == 1, NC == 64: fused 3x3/2 SAME max-pool of the output [B,OH,OW,64] (or null)
  uint8_t* arg_out;     //   ... and its argmax; y may then be null
};

// ---------------------------------------------------------------------------------------------------
// forward / input-gradient convolution.  grid = (ceil(B/ipc), NC/64), 288 threads: warps 0-3 stage
// the strip and run the epilogue, warps 4-7 issue MMAs (one accumulator tile each), warp 8 streams the
// packed filter taps with bulk (TMA) copies through a 6-stage ring.
// DGRAD = false: y = conv(x, w)            B operand (kc,nc) = w[tap][kc][nc]     (MN-major)
// DGRAD = true : y = conv(x, rot180(w)^T)  B operand (kc,nc) = w[24-tap][nc][kc]  (K-major)
// ---------------------------------------------------------------------------------------------------
// STAPS = filter taps per ring stage: 5 (one filter row; one wait / commit per row, the cheapest issue loop) or 1 (small
// stages, so that strip + ring stay under half an SM's shared memory).  MINB = 2 asks for two CTAs per SM: with more
// images than SMs one CTA's MMAs then overlap the other's strip fill and epilogue / pooling (every phase of a CTA is
// serial in itself), which is what large batches need; the single-wave batch 128 keeps the one-CTA shape.
// KSPLIT (with STAPS = 1): a stage holds 1/KSPLIT of a tap's contraction channels -- wide layers (256 channels: 33 KB
// per tap beside a 166 KB strip) stream the filter as (tap, channel-half) pieces.
template <bool DGRAD, int KSTEPS, bool POOL = false, int STAPS = 5, int MINB = 1, int KSPLIT = 1>
__global__ void __launch_bounds__(288, MINB) conv5x5_tc_kernel(const __grid_constant__ ConvArgs a) {
  static_assert(KSPLIT == 1 || STAPS == 1, "channel-split stages are single-tap stages");
  static_assert(KSTEPS % KSPLIT == 0, "KSPLIT must divide the k-steps");
  constexpr int KST = KSTEPS / KSPLIT;    // k-steps per ring stage and tap
  B200_TRACE_SCOPE(DGRAD ? TR_CONV2_DGRAD : TR_CONV2_FWD);
  constexpr int BN = 64;
  constexpr int NISS = 4;                 // MMA-issuing warps (warps 4..7); warp 8 streams the filter
  const ConvGeom& g = a.g;
  extern __shared__ __align__(128) uint8_t smem[];
  const int kchunks = KSTEPS * 2;         // KC = 16 * KSTEPS
  const bool two_pass = KSPLIT == 2 && g.two_pass;
  const int strip_bytes = (two_pass ? kchunks / 2 : kchunks) * g.ch_stride;
  // weight stage: MN-major [BN/8 groups][KC k][16B] (SBO = KC*16+16)  |  K-major [KC/8 chunks][BN rows][16B]
  const int w_mn_sbo = (g.KC / KSPLIT) * 16 + 16, w_k_lbo = BN * 16 + 16;
  const int wtap = ((DGRAD ? (kchunks / KSPLIT) * w_k_lbo : (BN / 8) * w_mn_sbo) + 127) & ~127;   // one tap's (piece of the) operand image
  const int wstage = STAPS * wtap;
  constexpr int ROUNDS = 25 / STAPS * KSPLIT;      // ring rounds per image
  uint8_t* strip = smem;
  uint8_t* wring = smem + ((strip_bytes + 127) & ~127);
  uint8_t* ctrl = wring + g.nstages * wstage;
  uint64_t* wfull = reinterpret_cast<uint64_t*>(ctrl);
  uint64_t* wempty = wfull + MAX_WSTAGES;
  uint64_t* strip_full = wempty + MAX_WSTAGES;
  uint64_t* accum = strip_full + 1;
  uint64_t* pass_done = accum + 1;        // two_pass: the MMAs on the first half of the channels have been issued and completed
  uint64_t* lead_ready = pass_done + 1;   // several images per TMA strip: the loader warps have zeroed the rows above the first image
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(lead_ready + 1);
  float* dpm = reinterpret_cast<float*>(ctrl + 256);                  // POOL: [OH*OW][64] masked pooled gradient
  uint8_t* argb = reinterpret_cast<uint8_t*>(dpm + (POOL ? a.pg.OH * a.pg.OW * 64 : 0));
  float* bred = reinterpret_cast<float*>(argb + (POOL ? ((a.pg.OH * a.pg.OW * 64 + 15) & ~15) : 0));   // [64]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n_img0 = blockIdx.x * g.ipc;
  const int nc0 = blockIdx.y * BN;
  const int tmem_cols = g.tiles * BN <= 64 ? 64 : g.tiles * BN <= 128 ? 128 : g.tiles * BN <= 256 ? 256 : 512;

  if (tid == 0) {
    for (int s = 0; s < g.nstages; ++s) { mbar_init(smem_u32(wfull + s), 1); mbar_init(smem_u32(wempty + s), NISS); }
    mbar_init(smem_u32(strip_full), POOL ? 256 : (g.tma_rows ? 1 : LOADERS));
    mbar_init(smem_u32(accum), NISS);
    mbar_init(smem_u32(pass_done), NISS);
    mbar_init(smem_u32(lead_ready), LOADERS);
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc(smem_u32(tmem_slot), tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (POOL && warp < 8) {
    // ---- dY strip from the pooled gradient: the four issuing warps help (they have nothing to issue before it) ----
    if (a.timeline && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) a.timeline[0] = clock64();
    for (int o = tid * 16; o < strip_bytes; o += 256 * 16) *reinterpret_cast<uint4*>(strip + o) = make_uint4(0, 0, 0, 0);
    pdl_wait();
    pdl_launch_dependents();
    if (a.timeline && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) a.timeline[1] = clock64();
    float bsum[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int img = 0; img < g.ipc && n_img0 + img < g.B; ++img) {
      const int n = n_img0 + img;
      if (img) nt_sync<256>();                       // the staging buffers are free again
      stage_pool_grad_nt<256>(dpm, argb, a.pg, n, tid);
      nt_sync<256>();                                // (also orders the clears above)
      float* dense = (a.dconv_out && blockIdx.y == 0) ? a.dconv_out + (size_t)n * g.H * g.W * 64 : nullptr;
      __nv_bfloat16* dense16 = (a.dconv_out_bf16 && blockIdx.y == 0) ? a.dconv_out_bf16 + (size_t)n * g.H * g.W * 64 : nullptr;
      pool_grad_rows<256>(dpm, argb, g.W, a.pg.OH, a.pg.OW, a.pg.pbh, a.pg.pbw, 0, g.H, tid,
                          [&](int i, int j, int c8, const float (&acc)[8]) {
#pragma unroll
                            for (int e = 0; e < 8; ++e) bsum[e] += acc[e];
                            const int slot = g.lead + (img * g.rows_per_img + i) * g.P + j;
                            const uint4 packed = pack8(acc);
                            *reinterpret_cast<uint4*>(strip + c8 * g.ch_stride + slot * 16) = packed;
                            if (dense16) *reinterpret_cast<uint4*>(dense16 + (size_t)(i * g.W + j) * 64 + c8 * 8) = packed;
                            if (dense) {     // dense NHWC fp32 copy for the filter gradient, which runs later
                              float4* o = reinterpret_cast<float4*>(dense + (size_t)(i * g.W + j) * 64 + c8 * 8);
                              o[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
                              o[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
                            }
                          });
    }
    fence_proxy_async();
    mbar_arrive(smem_u32(strip_full));
    if ((a.dbias || a.dbias_part) && blockIdx.y == 0) {        // BiasAddGrad of the conv layer: thread <-> channel chunk tid & 7
      // fixed order throughout: a thread's pixels in loop order, lanes by two shuffles, the eight warps one after the other
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        bsum[e] += __shfl_xor_sync(0xffffffffu, bsum[e], 8);
        bsum[e] += __shfl_xor_sync(0xffffffffu, bsum[e], 16);
      }
      if (lane < 8)
#pragma unroll
        for (int e = 0; e < 8; ++e) bred[warp * 64 + lane * 8 + e] = bsum[e];
      if (warp < 4) {
        nt_sync<256>();
        if (tid < 64) {
          float s = 0.f;
#pragma unroll
          for (int w8 = 0; w8 < 8; ++w8) s += bred[w8 * 64 + tid];
          if (a.dbias_part) a.dbias_part[(size_t)blockIdx.x * 64 + tid] = s;
          else atomicAdd(a.dbias + tid, s);
        }
      } else {
        asm volatile("bar.arrive 3, 256;" ::: "memory");      // the issuing warps do not wait for the sum
      }
    }
  }
  if (warp < 4) {
    if (!POOL && g.tma_rows) {       // the TMA warp fills the strip: nothing to stage here ...
      if (a.timeline && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) a.timeline[0] = a.timeline[1] = clock64();
      if (g.tma_h0 == 0) {           // ... but the zero rows above the first image (boxes start at the image's first row)
        const int nplanes = two_pass ? kchunks / 2 : kchunks;
        for (int q = tid; q < nplanes * g.lead; q += LOADERS) {
          const int c = q / g.lead, sl = q - c * g.lead;
          *reinterpret_cast<uint4*>(strip + c * g.ch_stride + sl * 16) = make_uint4(0, 0, 0, 0);
        }
        fence_proxy_async();
        mbar_arrive(smem_u32(lead_ready));
      }
      pdl_wait();
      pdl_launch_dependents();
    } else if (!POOL) {
      // ---- input strip --------------------------------------------------------------------------
      if (a.timeline && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) a.timeline[0] = clock64();
      zero_smem(strip, strip_bytes, tid);
      loaders_sync();
      pdl_wait();      // these warps are the only ones that touch global memory besides the filter fetch (TMA warp)
      pdl_launch_dependents();   // after our own wait: a dependent overlaps this kernel only, never our producer
      if (a.timeline && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) a.timeline[1] = clock64();
      if (a.xb) fill_strip_bf16(strip, a.xb, g, n_img0, g.KC, g.ch_stride, tid);
      else fill_strip(strip, a.x, g, n_img0, g.KC, g.ch_stride, tid);
      fence_proxy_async();
      mbar_arrive(smem_u32(strip_full));
    }
    if (a.timeline && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) a.timeline[2] = clock64();
    // ---- epilogue: TMEM -> registers -> (+bias, relu) -> dense NHWC fp32 ---------------------------
    mbar_wait(smem_u32(accum), 0);
    tc_fence_after();
    if (a.timeline && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) a.timeline[5] = clock64();
    const bool relu = a.relu != 0;
    for (int t = 0; t < g.tiles; ++t) {
      const int p = t * 128 + warp * 32 + lane;
      const int r = p / g.P, c = p - r * g.P;
      const int img = r / g.rows_per_img, i = r - img * g.rows_per_img;
      const int n = n_img0 + img;
      const bool valid = (img < g.ipc) && (n < g.B) && (i < g.H) && (c < g.W);
      float* out = a.y ? a.y + (((int64_t)n * g.H + i) * g.W + c) * g.NC + nc0 : nullptr;
      // fused pool: the finished MMAs no longer need the strip, so the fp32 conv tile of the (single) image
      // goes where the strip was, rows of 68 floats (conflict-free 16-byte stores)
      float* trow = a.pool_out ? reinterpret_cast<float*>(strip) + (size_t)(i * g.W + c) * 68 : nullptr;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + t * BN + c0, v);
        tmem_ld_wait();
        if (valid) {
          // (bias as vector loads and the ReLU as a select: the per-element `if (bias)` / `if (relu)`
          //  of the first version compiled to branches and cost more than the MMAs' tail)
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 b = a.bias ? __ldg(reinterpret_cast<const float4*>(a.bias + nc0 + c0 + j)) : make_float4(0.f, 0.f, 0.f, 0.f);
            const float x0 = __uint_as_float(v[j]) + b.x, x1 = __uint_as_float(v[j + 1]) + b.y;
            const float x2 = __uint_as_float(v[j + 2]) + b.z, x3 = __uint_as_float(v[j + 3]) + b.w;
            const float4 o = make_float4(relu ? fmaxf(x0, 0.f) : x0, relu ? fmaxf(x1, 0.f) : x1,      // selects, not branches
                                         relu ? fmaxf(x2, 0.f) : x2, relu ? fmaxf(x3, 0.f) : x3);
            if (out) *reinterpret_cast<float4*>(out + c0 + j) = o;
            if (trow) *reinterpret_cast<float4*>(trow + c0 + j) = o;
          }
        }
      }
    }
    if (a.timeline && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) a.timeline[6] = clock64();
    tc_fence_before();
  } else if (warp < 4 + NISS) {
    // ---- MMA issue.  One M=128,N=64,K=16 MMA takes 48 cycles but a single thread needs ~65-130 cycles
    // of scalar work to issue one (measured), so NISS warps issue in parallel, each feeding its own
    // accumulator tiles; the k-steps of a tap are a compile-time unrolled group.
    if (lane == 0) {
      const int w = warp - 4;
      const uint32_t idesc = make_idesc(128, BN, false, !DGRAD);
      const uint32_t a_hi = desc_hi(128), b_hi = DGRAD ? desc_hi(128) : desc_hi(w_mn_sbo);
      const uint32_t a_lo0 = desc_lo(smem_u32(strip) + g.lead * 16, g.ch_stride);
      const uint32_t b_lo0 = DGRAD ? desc_lo(smem_u32(wring), w_k_lbo) : desc_lo(smem_u32(wring), 128);
      const uint32_t a_kstep = 2 * (g.ch_stride >> 4), b_kstep = DGRAD ? 2 * (w_k_lbo >> 4) : 16;
      mbar_wait(smem_u32(strip_full), 0);
      if (!POOL && g.tma_rows && g.tma_h0 == 0) { mbar_wait(smem_u32(lead_ready), 0); tc_fence_after(); }
      if (a.timeline && blockIdx.x == 0 && blockIdx.y == 0 && w == 0) a.timeline[3] = clock64();
      // Every CTA streams the same 25 filter taps out of L2; each walks them in its own rotation so
      // that the CTAs do not all pull the same lines at once (the sum over taps is order-independent).
      // (single-tap stages rotate over all 25 taps: 25 phases instead of 5)
      const int rot = STAPS == 5 ? (int)((blockIdx.x + blockIdx.y) % 5u) * 5 : (int)((blockIdx.x * 7u + blockIdx.y) % 25u);
      int di = rot / 5;                                     // filter-row rotation (same in the producer warp)
      int dj1 = rot % 5;                                    // STAPS == 1: the tap inside the row
      uint32_t q = 0;                                       // stage pieces issued so far (0 = the accumulators are not initialised)
      int hh = 0;                                           // KSPLIT > 1: which piece of the tap's channels
      for (int r = 0; r < ROUNDS; ++r) {
        const int s = r % g.nstages;
        const uint32_t ph = (r / g.nstages) & 1;
        if (two_pass && r == 25) {                          // second half of the channels: the strip has been refilled
          mbar_wait(smem_u32(strip_full), 1);
          tc_fence_after();
        }
        mbar_wait(smem_u32(wfull + s), ph);               // one wait and one commit per stage: both cost
        tc_fence_after();                                 // ~100 cycles of the issuing thread (measured)
        const uint32_t b_row = b_lo0 + (uint32_t)s * (wstage >> 4);
#pragma unroll
        for (int tt = 0; tt < STAPS; ++tt, ++q) {
          const int dj = STAPS == 5 ? tt : dj1;
          const uint32_t a_tap = a_lo0 + (uint32_t)((di - 2) * g.P + (dj - 2)) +     // a tap is a slot offset
                                 (two_pass ? 0u : (uint32_t)(hh * KST) * a_kstep);
          // dgrad correlates with the filter rotated by 180 degrees: a five-tap stage holds forward row 4-di, read
          // backwards; a one-tap stage holds forward tap (4-di, 4-dj) itself
          const uint32_t b_tap = b_row + (uint32_t)(STAPS == 5 ? (DGRAD ? 4 - dj : dj) : 0) * (wtap >> 4);
          for (int t = w; t < g.tiles; t += NISS) {
            uint32_t a_lo = a_tap + t * 128, b_lo = b_tap;
            const uint32_t d = tmem + t * BN;
#pragma unroll
            for (int ks = 0; ks < KST; ++ks) {
              umma_bf16(d, desc_join(a_lo, a_hi), desc_join(b_lo, b_hi), idesc, q | (uint32_t)ks);
              a_lo += a_kstep; b_lo += b_kstep;
            }
          }
        }
        umma_commit(smem_u32(wempty + s));        // arrives when this warp's MMAs on the stage are done
        if (two_pass) {                           // pass-major: 25 taps of one half, then 25 taps of the other
          if (r == 24) { umma_commit(smem_u32(pass_done)); hh = 1; }
          if (++dj1 == 5) { dj1 = 0; if (++di == 5) di = 0; }
        } else if (STAPS == 5) { if (++di == 5) di = 0; }
        else if (++hh == KSPLIT) { hh = 0; if (++dj1 == 5) { dj1 = 0; if (++di == 5) di = 0; } }
      }
      umma_commit(smem_u32(accum));
      if (a.timeline && blockIdx.x == 0 && blockIdx.y == 0 && w == 0) a.timeline[4] = clock64();
    }
    __syncwarp();
  } else {
    // ---- filter taps: one thread, bulk copies global -> shared, completion counted in bytes ----------
    if (lane == 0) {
      if (!POOL && g.tma_rows) {
        // ---- the strip: one tensor-map box per 8-channel chunk, image rows -3 .. H+1, zeros outside the image ----
        pdl_wait();
        const uint32_t bar = smem_u32(strip_full);
        const int nc = two_pass ? kchunks / 2 : kchunks;
        const int nbox = g.tma_h0 == 0 ? g.ipc : 1;                  // one box per image and chunk
        const uint32_t box0 = g.tma_h0 == 0 ? (uint32_t)g.lead * 16 : 0u, boxstep = (uint32_t)g.rows_per_img * g.P * 16;
        mbar_arrive_expect_tx(bar, (uint32_t)(nc * nbox * g.P * g.tma_rows * 16));
        for (int im = 0; im < nbox; ++im)      // (an image index past the batch is out of bounds too: zeros)
          for (int c = 0; c < nc; ++c)
            tma_load_5d(smem_u32(strip + c * g.ch_stride) + box0 + im * boxstep, &a.tmap, 0, c, 0, g.tma_h0, n_img0 + im, bar);
      }
      const int img = (a.cout_f / 8) * (a.cin_f * 16 + 16);             // bytes of one packed tap
      const int rot = STAPS == 5 ? (int)((blockIdx.x + blockIdx.y) % 5u) * 5 : (int)((blockIdx.x * 7u + blockIdx.y) % 25u);
      int di = rot / 5, dj1 = rot % 5, hh = 0;                         // same rotation as the MMA warps
      const int gstride = a.cin_f * 16 + 16;                           // packed: bytes between cout groups of a tap
      for (int r = 0; r < ROUNDS; ++r) {
        const int s = r % g.nstages;
        const uint32_t ph = (r / g.nstages) & 1;
        if (two_pass && r == 25) {
          // every MMA on the first half of the channels has completed: the strip may take the second half
          mbar_wait(smem_u32(pass_done), 0);
          const uint32_t sbar = smem_u32(strip_full);
          const int nbox = g.tma_h0 == 0 ? g.ipc : 1;
          const uint32_t box0 = g.tma_h0 == 0 ? (uint32_t)g.lead * 16 : 0u, boxstep = (uint32_t)g.rows_per_img * g.P * 16;
          mbar_arrive_expect_tx(sbar, (uint32_t)((kchunks / 2) * nbox * g.P * g.tma_rows * 16));
          for (int im = 0; im < nbox; ++im)
            for (int c = 0; c < kchunks / 2; ++c)
              tma_load_5d(smem_u32(strip + c * g.ch_stride) + box0 + im * boxstep, &a.tmap, 0, kchunks / 2 + c, 0, g.tma_h0,
                          n_img0 + im, sbar);
          hh = 1;
        }
        mbar_wait(smem_u32(wempty + s), ph ^ 1);
        const uint32_t bar = smem_u32(wfull + s);
        // packed (forward) taps to fetch: a filter row in forward order, or the single tap the MMA warps want next
        const int ftap0 = STAPS == 5 ? (DGRAD ? 4 - di : di) * 5 : (DGRAD ? (4 - di) * 5 + (4 - dj1) : di * 5 + dj1);
        if (!DGRAD && KSPLIT == 1) {     // N-groups nc0/8 .. +8 of a tap are contiguous: one copy per tap
          const uint32_t bytes = (BN / 8) * w_mn_sbo;
          mbar_arrive_expect_tx(bar, STAPS * bytes);
          for (int j = 0; j < STAPS; ++j)
            bulk_g2s(smem_u32(wring + s * wstage + j * wtap),
                     a.wp + (size_t)(ftap0 + j) * img + (size_t)(nc0 / 8) * w_mn_sbo, bytes, bar);
        } else if (!DGRAD) {             // a piece of the contraction channels: one copy per cout group
          const uint32_t bytes = (uint32_t)(g.KC / KSPLIT) * 16;
          mbar_arrive_expect_tx(bar, (BN / 8) * bytes);
          const uint8_t* src = a.wp + (size_t)ftap0 * img + (size_t)(nc0 / 8) * gstride + (size_t)hh * bytes;
          for (int gi = 0; gi < BN / 8; ++gi)
            bulk_g2s(smem_u32(wring + s * wstage + gi * w_mn_sbo), src + (size_t)gi * gstride, bytes, bar);
        } else {          // read K-major: per k-chunk (forward cout group) the 64 rows nc0.. are contiguous
          constexpr int KCS = 2 * KSTEPS / KSPLIT;                      // chunks per stage and tap
          mbar_arrive_expect_tx(bar, (uint32_t)STAPS * KCS * BN * 16);
          for (int j = 0; j < STAPS; ++j) {
            const uint8_t* src = a.wp + (size_t)(ftap0 + j) * img + (size_t)nc0 * 16 + (size_t)(hh * KCS) * gstride;
            const uint32_t ws = smem_u32(wring + s * wstage + j * wtap);
            for (int c = 0; c < KCS; ++c) bulk_g2s(ws + c * w_k_lbo, src + (size_t)c * gstride, BN * 16, bar);
          }
        }
        if (two_pass) { if (++dj1 == 5) { dj1 = 0; if (++di == 5) di = 0; } }
        else if (STAPS == 5) { if (++di == 5) di = 0; }
        else if (++hh == KSPLIT) { hh = 0; if (++dj1 == 5) { dj1 = 0; if (++di == 5) di = 0; } }
      }
    }
    __syncwarp();
  }
  if (!DGRAD && !POOL && a.pool_out) {
    // ---- 3x3/2 SAME max-pool (pad 0 before / rest after), first max in (h,w) scan order (:123), by ALL nine warps
    //      (the issuing and the filter warps are done by now): 36 cells x 8 channel chunks = one item per thread at
    //      12x12, where the four epilogue warps alone needed three rounds ----
    __syncthreads();
    if (n_img0 < g.B) {
      const float* tile = reinterpret_cast<const float*>(strip);
      const int OH = (g.H + 1) / 2, OW = (g.W + 1) / 2;
      const int pbh = max((OH - 1) * 2 + 3 - g.H, 0) / 2, pbw = max((OW - 1) * 2 + 3 - g.W, 0) / 2;
      // one item = 8 channels of one pooled cell; window positions outside the image read a clamped address and enter
      // the scan as -FLT_MAX (no branch); strict '>' in (kh, kw) order = first maximum
      // (four channels per item: a pixel's sixteen lanes read 256 contiguous bytes, conflict-free; see conv1's pool)
      for (int q = tid; q < OH * OW * 16; q += (int)blockDim.x) {
        const int c4 = q & 15, cell = q >> 4;
        const int oh = cell / OW, ow = cell - oh * OW;
        float best[4];
        int bi[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) { best[e] = -3.402823466e38f; bi[e] = 0; }
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
          const int hh = 2 * oh + kh - pbh;
          const bool vh = (unsigned)hh < (unsigned)g.H;
#pragma unroll
          for (int kw = 0; kw < 3; ++kw) {
            const int ww = 2 * ow + kw - pbw;
            const bool ok = vh && (unsigned)ww < (unsigned)g.W;
            const float4 v0 = *reinterpret_cast<const float4*>(tile + (size_t)(ok ? hh * g.W + ww : 0) * 68 + c4 * 4);
            const float vv[4] = {v0.x, v0.y, v0.z, v0.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float x = ok ? vv[e] : -3.402823466e38f;
              const bool gt = x > best[e];
              best[e] = gt ? x : best[e];
              bi[e] = gt ? kh * 3 + kw : bi[e];
            }
          }
        }
        const size_t o = ((size_t)n_img0 * OH * OW + cell) * 64 + c4 * 4;
        *reinterpret_cast<float4*>(a.pool_out + o) = make_float4(best[0], best[1], best[2], best[3]);
        *reinterpret_cast<uint32_t*>(a.arg_out + o) = (uint32_t)(bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24));
        // the flattened pool output (:126, feature = cell * 64 + channel) as FC1's operand image [feature / 8][batch][8]
        if (a.pool_img) {
          const __nv_bfloat162 p0 = __floats2bfloat162_rn(best[0], best[1]), p1 = __floats2bfloat162_rn(best[2], best[3]);
          *reinterpret_cast<uint2*>(a.pool_img + ((size_t)(cell * 8 + (c4 >> 1)) * g.B + n_img0) * 16 + (c4 & 1) * 8) =
              make_uint2(*reinterpret_cast<const uint32_t*>(&p0), *reinterpret_cast<const uint32_t*>(&p1));
        }
      }
    }
  }
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc(tmem, tmem_cols);
  }
}

// fp32 HWIO filter [25][Cin][Cout] -> bf16 images: P[tap][g = cout/8][cin][cout%8], group stride
// Cin*16+16 bytes.  Read MN-major (groups along N) it is the forward B operand of tap `tap`; read
// K-major (groups along K) it is the input-gradient B operand of tap 24-tap.  One packed copy serves both.
__global__ void pack_conv_weights_kernel(const float* __restrict__ w, uint8_t* __restrict__ wp, int Cin, int Cout) {
  B200_TRACE_SCOPE(TR_PACK);
  pdl_launch_dependents();      // nothing downstream waits on this kernel programmatically: the next one may start at once
  const int gstride = Cin * 16 + 16, img = (Cout / 8) * gstride;
  const int total = 25 * Cin * (Cout / 8);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int g = i % (Cout / 8);
    const int t = i / (Cout / 8);
    const int cin = t % Cin, tap = t / Cin;
    *reinterpret_cast<uint4*>(wp + (size_t)tap * img + g * gstride + cin * 16) =
        load_chunk(w + ((size_t)tap * Cin + cin) * Cout + g * 8);
  }
}

int pick_ipc(int B, int H, int W, int KC) {
  // more images per CTA = fewer wasted rows and less filter re-streaming, but fewer CTAs; keep
  // the grid at or above one wave of the 148 SMs while TMEM (tiles*64 <= 512) and smem allow
  int best = 1;
  for (int ipc : {2, 4}) {
    ConvGeom g = make_geom(B, H, W, KC, 64, ipc);
    int64_t smem = (int64_t)(KC / 8) * g.ch_stride + WSTAGES * 5 * (int64_t)std::max((KC / 8) * 1040, 8 * (KC * 16 + 16)) + 1024;
    if (g.tiles * 64 <= 512 && smem <= 200 * 1024 && (B + ipc - 1) / ipc >= 2 * kNumSMs) best = ipc;
  }
  return best;
}

long long* g_conv_timeline = nullptr;
int g_conv_force_variant = -1;      // diagnostics / tests: 0 = one CTA per SM (five-tap stages), 1 = two per SM, -1 = automatic

// filter stages of the one-CTA-per-SM shape (five taps each); B200_CONV_STAGES overrides the default of 3 (experiments)
int conv_row_stages() {
  static int n = 0;
  if (!n) {
    const char* e = getenv("B200_CONV_STAGES");
    n = e ? std::max(2, std::min(atoi(e), 5)) : WSTAGES;
  }
  return n;
}

// dense bf16 [B,H,W,C] as the 5-D tensor (8 channels | C/8 chunks | W | H | B) with a box of one chunk x P x rows
int make_strip_tensor_map(CUtensorMap* tm, const void* x_bf16, int B, int H, int W, int C, int P, int rows) {
  const cuuint64_t gdim[5] = {8, (cuuint64_t)(C / 8), (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  const cuuint64_t gstr[4] = {16, (cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  const cuuint32_t box[5] = {8, 1, (cuuint32_t)P, (cuuint32_t)rows, 1};
  const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  // the driver entry point is looked up through the runtime (libcuda is not a link-time dependency: the library
  // must load on a machine without a driver, e.g. for the ABI checks)
  using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn encode = nullptr;
  if (!encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    B200_CUDA_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    B200_REQUIRE(fn && qres == cudaDriverEntryPointSuccess, "conv5x5_tc: the driver does not export cuTensorMapEncodeTiled");
    encode = reinterpret_cast<EncodeFn>(fn);
  }
  const CUresult rc = encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(x_bf16), gdim, gstr,
                                             box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) {
    set_error("conv5x5_tc: cuTensorMapEncodeTiled failed with code %d", (int)rc);
    return B200_E_CUDA;
  }
  return B200_OK;
}

template <bool DGRAD>
int launch_conv(const float* x, const uint8_t* wp, const float* bias, float* y, int B, int H, int W, int KC, int NC,
                int relu, cudaStream_t st, const PoolGradSrc* pg = nullptr, float* dbias = nullptr,
                float* dconv_out = nullptr, float* pool_out = nullptr, uint8_t* arg_out = nullptr,
                const void* x_bf16 = nullptr, void* dconv_out_bf16 = nullptr, void* pool_img = nullptr,
                float* dbias_part = nullptr, int* n_part_out = nullptr) {
  B200_REQUIRE(B > 0 && H > 0 && W > 0, "conv5x5_tc: bad shape");
  B200_REQUIRE(KC % 16 == 0 && NC % 64 == 0, "conv5x5_tc: channels must be multiples of 16 (in) / 64 (out), got %d/%d",
               KC, NC);
  B200_REQUIRE(y || pool_out, "conv5x5_tc: no output");
  B200_REQUIRE((pg || ((uintptr_t)x & 15) == 0) && ((uintptr_t)wp & 15) == 0 && ((uintptr_t)y & 15) == 0,
               "conv5x5_tc: pointers must be 16-byte aligned");
  const int kchunks = KC / 8;
  // wide layers (>= 128 contraction channels): single-tap stages, 256 channels in two pieces per tap
  const bool wide = KC >= 128;
  const int ksplit = KC == 256 ? 2 : 1;
  const int wtap = ((DGRAD ? (kchunks / ksplit) * (64 * 16 + 16) : 8 * ((KC / ksplit) * 16 + 16)) + 127) & ~127;
  const int extra = pg ? pg->OH * pg->OW * 64 * 5 + 16 + 8 * 64 * 4 : 0;
  // ---- shape of the CTA: two per SM with single-tap filter stages once the grid exceeds one wave, else one per SM ----
  constexpr int kHalfSm = 113 * 1024;                    // (228 KB per SM - 1 KB reserved per CTA) / 2
  const bool tma_ok = x_bf16 && !pg && W + 4 <= 256 && H + 5 <= 256 && ((uintptr_t)x_bf16 & 15) == 0;
  ConvGeom g2 = tma_ok ? make_geom_tma(B, H, W, KC, NC, pool_out ? H * W * 68 * 4 : 0) : make_geom(B, H, W, KC, NC, 1);
  const int strip2 = (kchunks * g2.ch_stride + 127) & ~127;
  const int nst2 = std::min(MAX_WSTAGES, (kHalfSm - strip2 - 256 - extra) / wtap);
  bool two = KC == 64 && B > kNumSMs && nst2 >= 4 && g2.tiles * 64 <= 256;
  if (g_conv_force_variant >= 0) two = g_conv_force_variant == 1 && KC == 64 && nst2 >= 4 && g2.tiles * 64 <= 256;
  ConvGeom g;
  if (two) {
    g = g2;
    g.nstages = nst2;
  } else if (wide) {
    g = g2;                                              // one image per CTA ...
    // ... or two, where the strip can be a TMA strip of half the channels at a time (256 channels, bf16 source): the
    // filter stream -- 819 KB per CTA at ~9 B/clk per SM, what bounds these kernels -- then feeds twice the MMAs.
    // (Two-pass strips alone, which only double the stages in flight, measured no faster: 189 vs 172 us for the wide
    //  conv2 forward.  B200_CONV_WIDE_IPC=1 switches back to one image per CTA.)
    static const int wide_ipc_env = getenv("B200_CONV_WIDE_IPC") ? atoi(getenv("B200_CONV_WIDE_IPC")) : 2;
    if (ksplit == 2 && tma_ok && wide_ipc_env >= 2 && B >= 2) {
      ConvGeom g3 = make_geom_tma(B, H, W, KC, NC, 0, 2);
      if (g3.tiles * 64 <= 512 && (((kchunks / 2) * g3.ch_stride + 127) & ~127) + 3 * wtap + 256 <= 227 * 1024) {
        g = g3;
        g.two_pass = 1;
      }
    }
    const int strip_w = g.two_pass ? (((kchunks / 2) * g.ch_stride + 127) & ~127) : strip2;
    g.nstages = std::min(MAX_WSTAGES, (227 * 1024 - strip_w - 256 - extra) / wtap);
    B200_REQUIRE(g.nstages >= 2, "conv5x5_tc: image %dx%d with %d channels does not leave room for the filter ring", H, W, KC);
  } else {
    const int ipc = pool_out ? 1 : pick_ipc(B, H, W, KC);       // the fused pool works on one image per CTA
    g = (tma_ok && ipc == 1) ? make_geom_tma(B, H, W, KC, NC, pool_out ? H * W * 68 * 4 : 0) : make_geom(B, H, W, KC, NC, ipc);
    g.nstages = conv_row_stages();
    while (g.nstages > 2 && ((kchunks * g.ch_stride + 127) & ~127) + g.nstages * 5 * wtap + 256 + extra > 227 * 1024) --g.nstages;
  }
  if (pool_out) {
    B200_REQUIRE(!DGRAD && NC == 64 && arg_out, "conv5x5_tc: fused pool needs the forward kernel with 64 output channels");
    // the epilogue parks the fp32 conv tile (68 floats per pixel) where the strip was: widen the chunk planes if the
    // strip alone is smaller (16x16 images from an fp32 source: 60 KB of strip, 70 KB of tile)
    while ((int64_t)H * W * 68 * 4 > (int64_t)(KC / 8) * g.ch_stride) g.ch_stride += 128;
  }
  B200_REQUIRE(g.tiles * 64 <= 512, "conv5x5_tc: image %dx%d too large for one CTA's TMEM", H, W);
  const int staps = (two || wide) ? 1 : 5;
  const int smem = ((((g.two_pass ? kchunks / 2 : kchunks)) * g.ch_stride + 127) & ~127) + g.nstages * staps * wtap + 256 + extra;
  B200_REQUIRE(smem <= 227 * 1024, "conv5x5_tc: %d bytes of shared memory needed", smem);
  void (*kern)(ConvArgs) = nullptr;
  if (pg) {
    B200_REQUIRE(DGRAD && KC == 64, "conv5x5_tc: the pooled-gradient source needs 64 channels");
    kern = two ? conv5x5_tc_kernel<true, 4, true, 1, 2> : conv5x5_tc_kernel<true, 4, true>;
  } else
  switch (KC / 16) {
    case 2: kern = conv5x5_tc_kernel<DGRAD, 2>; break;
    case 4: kern = two ? conv5x5_tc_kernel<DGRAD, 4, false, 1, 2> : conv5x5_tc_kernel<DGRAD, 4>; break;
    case 8: kern = conv5x5_tc_kernel<DGRAD, 8, false, 1, 1>; break;
    case 16: kern = conv5x5_tc_kernel<DGRAD, 16, false, 1, 1, 2>; break;
    default: B200_REQUIRE(false, "conv5x5_tc: contraction channels must be 32, 64, 128 or 256 (got %d)", KC);
  }
  B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  if (two) B200_CUDA_TRY(prefer_max_shared(kern));
  ConvArgs a{{}, x, reinterpret_cast<const __nv_bfloat16*>(x_bf16), reinterpret_cast<__nv_bfloat16*>(dconv_out_bf16),
             reinterpret_cast<uint8_t*>(pool_img), wp, DGRAD ? NC : KC, DGRAD ? KC : NC, bias, y, relu, g, g_conv_timeline, pg ? *pg : PoolGradSrc{}, dbias, dbias_part, dconv_out, pool_out, arg_out};
  if (n_part_out) *n_part_out = cdiv(B, g.ipc);
  if (g.tma_rows) B200_TRY(make_strip_tensor_map(&a.tmap, x_bf16, B, H, W, KC, g.P, g.tma_rows));
  dim3 grid(cdiv(B, g.ipc), NC / 64);
  B200_CUDA_TRY(launch_kernel(kern, grid, dim3(288), (size_t)smem, st, a));
  return B200_OK;
}

// ---------------------------------------------------------------------------------------------------
// filter gradient.  grid = (5 filter rows, S image splits); each CTA accumulates the 5 taps of
// one filter row (5 x [64 cin x 64 cout] fp32 in TMEM, M = 64 MMAs) over its share of the images,
// double-buffering the x / dy strips, and writes one partial; a second kernel sums the S partials
// in a fixed order (deterministic).
// ---------------------------------------------------------------------------------------------------
struct WgradArgs {
  const float* x;      // [B,H,W,64]
  const float* dy;     // [B,H,W,64]
  const __nv_bfloat16* xb;    // bf16 copies of x / dy (both or none): read instead, half the bytes of the 5x re-read
  const __nv_bfloat16* dyb;
  float* partial;      // [S][25][64][64]
  ConvGeom g;          // ipc = 1
  int cin_tot, cout_tot;   // channels of x / dy; a CTA works on the 64 x 64 block blockIdx.z = (cin block) * (cout_tot / 64) + (cout block)
  int imgs_per_cta;
  int nbuf;             // 2 = double-buffered strips, 1 when two pairs do not fit in shared memory
  PoolGradSrc pg;       // when pg.dpool != null, dy is unused and the dY strip is gathered from the pooled gradient
  long long* timeline;  // optional clock64 stamps of CTA (0,0) in slots [8,14) of the diagnostics buffer, else null
};

constexpr int WG_LOADERS = 256;     // the filter-gradient kernel is bound by its strip fills: eight loader warps
__device__ __forceinline__ void wg_loaders_sync() { asm volatile("bar.sync 2, 256;" ::: "memory"); }
// x and dY strips of one image (64 channels each) filled together: every thread has its whole share of both
// tensors (<= 5 + 5 chunks of 32 bytes) in flight before the first conversion, one memory round trip per image
__device__ __forceinline__ void fill_strip_pair(uint8_t* xs, uint8_t* x2, uint8_t* ds, const float* __restrict__ x,
                                                const float* __restrict__ dy, const ConvGeom& g, int n, int tid,
                                                int cx, int cd) {      // channels per pixel of x / dy (64-channel slice at x / dy)
  constexpr int U = 5;
  const int per_img = g.H * g.W * 8;
  const float* sx = x + (int64_t)n * g.H * g.W * cx;
  const float* sd = dy + (int64_t)n * g.H * g.W * cd;
  for (int q0 = tid; q0 < per_img; q0 += WG_LOADERS * U) {
    float4 vx[U][2], vd[U][2];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int q = q0 + u * WG_LOADERS;
      if (q < per_img) {
        const int64_t ox = (int64_t)(q >> 3) * cx + (q & 7) * 8, od = (int64_t)(q >> 3) * cd + (q & 7) * 8;
        vx[u][0] = __ldg(reinterpret_cast<const float4*>(sx + ox));
        vx[u][1] = __ldg(reinterpret_cast<const float4*>(sx + ox) + 1);
        vd[u][0] = __ldg(reinterpret_cast<const float4*>(sd + od));
        vd[u][1] = __ldg(reinterpret_cast<const float4*>(sd + od) + 1);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int q = q0 + u * WG_LOADERS;
      if (q < per_img) {
        const int c = q & 7, pix = q >> 3;
        const int i = pix / g.W, j = pix - i * g.W;
        const int off = c * g.ch_stride + (g.lead + i * g.P + j) * 16;
        const float fx[8] = {vx[u][0].x, vx[u][0].y, vx[u][0].z, vx[u][0].w, vx[u][1].x, vx[u][1].y, vx[u][1].z, vx[u][1].w};
        const float fd[8] = {vd[u][0].x, vd[u][0].y, vd[u][0].z, vd[u][0].w, vd[u][1].x, vd[u][1].y, vd[u][1].z, vd[u][1].w};
        const uint4 px = pack8(fx);
        *reinterpret_cast<uint4*>(xs + off) = px;
        *reinterpret_cast<uint4*>(x2 + off - 16) = px;          // the copy shifted down one slot
        *reinterpret_cast<uint4*>(ds + off) = pack8(fd);
      }
    }
  }
}

__device__ __forceinline__ void fill_strip_pair_bf16(uint8_t* xs, uint8_t* x2, uint8_t* ds, const __nv_bfloat16* __restrict__ x,
                                                     const __nv_bfloat16* __restrict__ dy, const ConvGeom& g, int n,
                                                     int tid, int cx, int cd) {
  constexpr int U = 5;
  const int per_img = g.H * g.W * 8;
  const uint4* sx = reinterpret_cast<const uint4*>(x + (int64_t)n * g.H * g.W * cx);
  const uint4* sd = reinterpret_cast<const uint4*>(dy + (int64_t)n * g.H * g.W * cd);
  const int cx8 = cx >> 3, cd8 = cd >> 3;
  for (int q0 = tid; q0 < per_img; q0 += WG_LOADERS * U) {
    uint4 vx[U], vd[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int q = min(q0 + u * WG_LOADERS, per_img - 1);      // clamped, not predicated: no branch between the loads
      vx[u] = __ldg(sx + (q >> 3) * cx8 + (q & 7)); vd[u] = __ldg(sd + (q >> 3) * cd8 + (q & 7));
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int q = q0 + u * WG_LOADERS;
      if (q < per_img) {
        const int c = q & 7, pix = q >> 3;
        const int i = pix / g.W, j = pix - i * g.W;
        const int off = c * g.ch_stride + (g.lead + i * g.P + j) * 16;
        *reinterpret_cast<uint4*>(xs + off) = vx[u];
        *reinterpret_cast<uint4*>(x2 + off - 16) = vx[u];       // the copy shifted down one slot
        *reinterpret_cast<uint4*>(ds + off) = vd[u];
      }
    }
  }
}

__global__ void __launch_bounds__(416, 1) conv5x5_wgrad_tc_kernel(WgradArgs a) {
  B200_TRACE_SCOPE(TR_CONV2_WGRAD);
  pdl_launch_dependents();
  constexpr int C = 64;
  const ConvGeom& g = a.g;
  extern __shared__ __align__(128) uint8_t smem[];
  // per buffer: [x: 8 chunks][x shifted down one slot: 8 chunks][dY: 8 chunks].  The two x copies are 16 consecutive
  // chunks at one stride, i.e. the 16 M-groups of ONE M = 128 MMA whose rows 0-63 are tap dj and rows 64-127 tap dj+1.
  const int strip_bytes = ((C / 8) * g.ch_stride + 127) & ~127;       // one 8-chunk strip (dY); x + x2 = 2 * 8 chunks
  const int xpair_bytes = (2 * (C / 8) * g.ch_stride + 127) & ~127;
  const int buf_bytes = xpair_bytes + strip_bytes;
  uint8_t* ctrl = smem + a.nbuf * buf_bytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(ctrl);
  uint64_t* empty = full + 2;
  uint64_t* accum = empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum + 1);
  float* dpm = reinterpret_cast<float*>(ctrl + 256);
  uint8_t* argb = reinterpret_cast<uint8_t*>(dpm + (a.pg.dpool ? a.pg.OH * a.pg.OW * 64 : 0));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int di = blockIdx.x;
  const int nbo = a.cout_tot / C;                               // 64-channel blocks of dy
  const int cb_in = blockIdx.z / nbo, cb_out = blockIdx.z - cb_in * nbo;
  const float* x_f = a.x ? a.x + cb_in * C : nullptr;           // this CTA's 64-channel slices
  const float* dy_f = a.dy ? a.dy + cb_out * C : nullptr;
  const __nv_bfloat16* x_b = a.xb ? a.xb + cb_in * C : nullptr;
  const __nv_bfloat16* dy_b = a.dyb ? a.dyb + cb_out * C : nullptr;
  const int img_beg = blockIdx.y * a.imgs_per_cta;
  const int img_end = min(g.B, img_beg + a.imgs_per_cta);
  const int nimg = max(0, img_end - img_beg);
  const int ksteps = (g.H * g.P + 15) / 16;      // dY is zero beyond the image's H rows of P slots

  if (tid == 0) {
    for (int s = 0; s < 2; ++s) { mbar_init(smem_u32(full + s), WG_LOADERS); mbar_init(smem_u32(empty + s), 3); }
    mbar_init(smem_u32(accum), 3);
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc(smem_u32(tmem_slot), 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const bool stamp = a.timeline && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0;
  if (warp < 8) {
    if (stamp) a.timeline[8] = clock64();
    // pads stay zero for every image: fills touch valid pixels only
    for (int o = tid * 16; o < a.nbuf * buf_bytes; o += WG_LOADERS * 16) *reinterpret_cast<uint4*>(smem + o) = make_uint4(0, 0, 0, 0);
    wg_loaders_sync();
    pdl_wait();
    if (stamp) a.timeline[9] = clock64();
    // bf16 sources, image fits one batch of 5 chunks per thread: the loads of image it+1 are in flight while image it
    // is stored (a round trip to L2 is ~3k cycles here whatever the size; the MMAs of an image take 1.7k)
    constexpr int PU = 5;
    const int per_img = g.H * g.W * 8;
    if (x_b && !a.pg.dpool && per_img <= PU * WG_LOADERS) {
      uint4 vxa[PU], vda[PU], vxb[PU], vdb[PU];
      const int cx8 = a.cin_tot >> 3, cd8 = a.cout_tot >> 3;
      auto load = [&](int it, uint4 (&vx)[PU], uint4 (&vd)[PU]) {
        const uint4* sx = reinterpret_cast<const uint4*>(x_b + (int64_t)(img_beg + it) * g.H * g.W * a.cin_tot);
        const uint4* sd = reinterpret_cast<const uint4*>(dy_b + (int64_t)(img_beg + it) * g.H * g.W * a.cout_tot);
#pragma unroll
        for (int u = 0; u < PU; ++u) {
          const int q = min(tid + u * WG_LOADERS, per_img - 1);      // clamped, not predicated
          vx[u] = __ldg(sx + (q >> 3) * cx8 + (q & 7)); vd[u] = __ldg(sd + (q >> 3) * cd8 + (q & 7));
        }
      };
      auto store = [&](int it, const uint4 (&vx)[PU], const uint4 (&vd)[PU]) {
        const int b = it % a.nbuf;
        mbar_wait(smem_u32(empty + b), ((it / a.nbuf) & 1) ^ 1);
        uint8_t* xs = smem + b * buf_bytes;
        uint8_t* x2 = xs + (C / 8) * g.ch_stride;
        uint8_t* ds = xs + xpair_bytes;
#pragma unroll
        for (int u = 0; u < PU; ++u) {
          const int q = tid + u * WG_LOADERS;
          if (q < per_img) {
            const int c = q & 7, pix = q >> 3;
            const int i = pix / g.W, j = pix - i * g.W;
            const int off = c * g.ch_stride + (g.lead + i * g.P + j) * 16;
            *reinterpret_cast<uint4*>(xs + off) = vx[u];
            *reinterpret_cast<uint4*>(x2 + off - 16) = vx[u];
            *reinterpret_cast<uint4*>(ds + off) = vd[u];
          }
        }
        fence_proxy_async();
        mbar_arrive(smem_u32(full + b));
        if (stamp && it == 0) a.timeline[10] = clock64();
      };
      if (nimg > 0) load(0, vxa, vda);
#pragma unroll 1
      for (int it = 0; it < nimg; it += 2) {
        if (it + 1 < nimg) load(it + 1, vxb, vdb);
        store(it, vxa, vda);
        if (it + 2 < nimg) load(it + 2, vxa, vda);
        if (it + 1 < nimg) store(it + 1, vxb, vdb);
      }
    } else
    for (int it = 0; it < nimg; ++it) {
      const int b = it % a.nbuf;
      const uint32_t ph = (it / a.nbuf) & 1;
      mbar_wait(smem_u32(empty + b), ph ^ 1);
      uint8_t* xs = smem + b * buf_bytes;
      uint8_t* x2 = xs + (C / 8) * g.ch_stride;      // x2[slot] = x[slot + 1]
      uint8_t* ds = xs + xpair_bytes;
      if (a.pg.dpool) {                       // the pooled-gradient gather keeps its 128-thread helpers (warps 0-3)
        if (warp < 4) {
          float unused[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
          fill_strip(xs, x_f, g, img_beg + it, C, g.ch_stride, tid);
          stage_pool_grad(dpm, argb, a.pg, img_beg + it, tid);
          loaders_sync();
          fill_strip_from_pool(ds, dpm, argb, g, a.pg, 0, g.ch_stride, tid, unused);
          loaders_sync();                     // nobody restages dpm while a slower loader still gathers from it
          for (int q = tid; q < (C / 8) * (g.npix - 1); q += LOADERS) {      // the shifted copy, from shared memory
            const int c = q / (g.npix - 1), sl = q - c * (g.npix - 1);
            *reinterpret_cast<uint4*>(x2 + c * g.ch_stride + sl * 16) = *reinterpret_cast<const uint4*>(xs + c * g.ch_stride + (sl + 1) * 16);
          }
        }
      } else {
        if (x_b) fill_strip_pair_bf16(xs, x2, ds, x_b, dy_b, g, img_beg + it, tid, a.cin_tot, a.cout_tot);
        else fill_strip_pair(xs, x2, ds, x_f, dy_f, g, img_beg + it, tid, a.cin_tot, a.cout_tot);
      }
      fence_proxy_async();
      mbar_arrive(smem_u32(full + b));
      if (stamp && it == 0) a.timeline[10] = clock64();
    }
    if (stamp) a.timeline[11] = clock64();
    // ---- epilogue ------------------------------------------------------------------------------------------
    mbar_wait(smem_u32(accum), 0);
    tc_fence_after();
    if (stamp) a.timeline[12] = clock64();
    // this block of the split's partial [25][cin_tot][cout_tot]
    float* part = a.partial + (int64_t)blockIdx.y * 25 * a.cin_tot * a.cout_tot + (int64_t)cb_in * C * a.cout_tot + cb_out * C;
    const int64_t tap_stride = (int64_t)a.cin_tot * a.cout_tot;
    const int q4 = warp & 3, c0 = (warp >> 2) * 32;       // TMEM lane quarter, column half of each accumulator
    // accumulators 0 / 1 (M = 128): lanes 0-63 = tap 2p, lanes 64-127 = tap 2p+1, lane % 64 = cin
#pragma unroll 1
    for (int p = 0; p < 2; ++p) {
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(q4 * 32) << 16) + p * C + c0, v);
      tmem_ld_wait();
      const int row = q4 * 32 + lane, dj = 2 * p + (row >> 6), cin = row & 63;
      float* out = part + (di * 5 + dj) * tap_stride + (int64_t)cin * a.cout_tot + c0;
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        *reinterpret_cast<float4*>(out + j) =
            (nimg > 0) ? make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                     __uint_as_float(v[j + 3]))
                       : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    {   // accumulator 2 (tap 4, M = 64): lanes 0-15 of each 32-lane quarter
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(q4 * 32) << 16) + 2 * C + c0, v);
      tmem_ld_wait();
      if (lane < 16) {
        const int cin = q4 * 16 + lane;
        float* out = part + (di * 5 + 4) * tap_stride + (int64_t)cin * a.cout_tot + c0;
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<float4*>(out + j) =
              (nimg > 0) ? make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                       __uint_as_float(v[j + 3]))
                         : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    tc_fence_before();
    if (stamp) a.timeline[13] = clock64();
  } else {
    // warps 8..10: one issuing thread per accumulator of this CTA's filter row: taps (0,1) and (2,3) as M = 128
    // MMAs over the x / shifted-x pair, tap 4 as an M = 64 MMA (48 + 48 + 46 cycles per k-step instead of 5 x 46)
    if (lane == 0 && warp < 11) {
      const int w = warp - 8;
      const uint32_t idesc = make_idesc(w < 2 ? 128 : 64, C, true, true);
      const uint32_t hi = desc_hi(g.ch_stride);                 // MN-major: SBO = channel-chunk stride, LBO = 128
      const uint32_t d = tmem + w * C;
      uint32_t acc = 0;
      for (int it = 0; it < nimg; ++it) {
        const int b = it % a.nbuf;
        const uint32_t ph = (it / a.nbuf) & 1;
        mbar_wait(smem_u32(full + b), ph);
        tc_fence_after();
        const uint32_t xs = smem_u32(smem + b * buf_bytes);
        uint32_t x_lo = desc_lo(xs + (g.lead + (di - 2) * g.P - 2 + 2 * w) * 16, 128);
        uint32_t d_lo = desc_lo(xs + xpair_bytes + g.lead * 16, 128);
        for (int ks = 0; ks < ksteps; ++ks) {                    // 16 pixel slots = 16 x 16 bytes per k-step
          umma_bf16(d, desc_join(x_lo, hi), desc_join(d_lo, hi), idesc, acc | (uint32_t)ks);
          x_lo += 16; d_lo += 16;
        }
        acc = 1;
        umma_commit(smem_u32(empty + b));
      }
      umma_commit(smem_u32(accum));   // with nimg == 0 nothing was issued: the commit arrives at once
    }
    __syncwarp();
  }
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

// sums the S partials of every 4-float group: four neighbouring lanes take every fourth partial each (all of a
// lane's loads in flight together) and combine with two shuffles -- a fixed order, so the result is deterministic.
// A launch can carry a second, independent segment (its blocks follow the first segment's): conv2's bias-gradient
// partials ride along with its filter-gradient partials.
struct ReduceSeg { const float* part; float* out; int n, S, stride; };
__global__ void __launch_bounds__(128) reduce_partials_tc(ReduceSeg s0, ReduceSeg s1) {
  B200_TRACE_SCOPE(TR_CONV2_REDUCE);
  pdl_wait();
  const int nb0 = (s0.n + 127) / 128;
  if ((int)blockIdx.x >= nb0) {
    // second segment: MANY partials of a short vector (conv2's bias gradient: one partial per dgrad CTA, up to the batch
    // size of them, 64 floats each) -- one BLOCK per 16 bytes: thread t sums partials t, t + 128, ... in order, the 32
    // lanes of a warp combine by a shuffle tree, the four warps' sums are added in warp order.  Fixed order again.
    __shared__ float4 wsum[4];
    const int i = (blockIdx.x - nb0) * 4;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int z0 = threadIdx.x; z0 < s1.S; z0 += 128 * 8) {
      float4 v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int z = z0 + 128 * j;
        v[j] = z < s1.S ? __ldg(reinterpret_cast<const float4*>(s1.part + (int64_t)z * s1.stride + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) { s.x += v[j].x; s.y += v[j].y; s.z += v[j].z; s.w += v[j].w; }
    }
#pragma unroll
    for (int m = 1; m <= 16; m <<= 1) {
      s.x += __shfl_xor_sync(0xffffffffu, s.x, m); s.y += __shfl_xor_sync(0xffffffffu, s.y, m);
      s.z += __shfl_xor_sync(0xffffffffu, s.z, m); s.w += __shfl_xor_sync(0xffffffffu, s.w, m);
    }
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      float4 r = wsum[0];
#pragma unroll
      for (int w = 1; w < 4; ++w) { r.x += wsum[w].x; r.y += wsum[w].y; r.z += wsum[w].z; r.w += wsum[w].w; }
      *reinterpret_cast<float4*>(s1.out + i) = r;
    }
    return;
  }
  const ReduceSeg& g = s0;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int sub = t & 3;
  const int i = min((t >> 2) * 4, g.n - 4);                 // surplus threads redo the last group (shuffles stay full-warp)
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int z0 = sub; z0 < g.S; z0 += 32) {
    float4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int z = z0 + 4 * j;
      v[j] = z < g.S ? __ldg(reinterpret_cast<const float4*>(g.part + (int64_t)z * g.stride + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) { s.x += v[j].x; s.y += v[j].y; s.z += v[j].z; s.w += v[j].w; }
  }
#pragma unroll
  for (int m = 1; m <= 2; m <<= 1) {
    s.x += __shfl_xor_sync(0xffffffffu, s.x, m); s.y += __shfl_xor_sync(0xffffffffu, s.y, m);
    s.z += __shfl_xor_sync(0xffffffffu, s.z, m); s.w += __shfl_xor_sync(0xffffffffu, s.w, m);
  }
  if (sub == 0 && (t >> 2) * 4 < g.n) *reinterpret_cast<float4*>(g.out + i) = s;
}

// image splits of the filter gradient: 5 filter rows x (64 x 64 channel blocks) x splits ~ one wave of CTAs.
// Up to one wave of images (batch <= 148) conv2's filter gradient SHARES the machine with conv1's (which follows dgrad
// on the main stream at the same moment, one CTA per SM each): conv1 takes ceil(B/2) CTAs (two images each), this kernel
// 5 x splits CTAs out of the rest -- 64 + 80 at batch 128.  Both then finish in one wave instead of one after the other:
// 105.6 vs 110.9 us per step back to back (118.7 -> 120.7 us with L2 flushed between steps; measured, scratch/gpu_call24.sh).
// (B200_WGRAD2_SPLITS / B200_WGRAD1_CTAS override both choices: experiments.)
int wgrad_splits_tc(int B, int nblk = 1) {
  static int env = -1;
  if (env < 0) { const char* e = getenv("B200_WGRAD2_SPLITS"); env = e ? atoi(e) : 0; }
  if (env > 0 && nblk == 1) return std::max(1, std::min(B, env));
  if (nblk == 1 && B <= kNumSMs) return std::max(1, std::min(B, (kNumSMs - (B + 1) / 2) / 5));
  if (nblk == 1) return std::max(1, std::min(B, (kNumSMs + 4) / 5));
  // several 64 x 64 channel blocks (wide layers): 5 * nblk CTAs per split do not fill a wave by themselves (80 of 148 for
  // 256 x 256 channels), so the splits are chosen by a small cost model -- waves x (images per CTA x 3.6k cycles measured
  // per image + 10k cycles of set-up and epilogue) + 2k cycles of partial-sum traffic per split
  int best = 1;
  long best_cost = -1;
  for (int s = 1; s <= std::min(B, 16); ++s) {
    const long waves = (5L * nblk * s + kNumSMs - 1) / kNumSMs;
    const long cost = waves * (((B + s - 1) / s) * 3600L + 10000L) + 2000L * s;
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = s; }
  }
  return best;
}
}  // namespace

int reduce_partials(const float* part, float* out, int n, int S, cudaStream_t st, int stride) {
  if (stride <= 0) stride = n;
  B200_REQUIRE(part && out && n > 0 && n % 4 == 0 && S > 0 && stride % 4 == 0, "reduce_partials: bad argument");
  B200_CUDA_TRY(launch_kernel(reduce_partials_tc, dim3(cdiv(n, 128)), dim3(128), 0, st, ReduceSeg{part, out, n, S, stride},
                              ReduceSeg{nullptr, nullptr, 0, 0, 0}));
  return B200_OK;
}
void conv5x5_set_timeline(long long* dev_buf) { g_conv_timeline = dev_buf; }
void conv5x5_force_variant(int v) { g_conv_force_variant = v; }
B200_TRACE_SETTER(trace_set_tc_conv)
int64_t conv5x5_packed_bytes(int Cin, int Cout) { return 25ll * (Cout / 8) * (Cin * 16 + 16); }
int pack_conv_weights(const float* w, void* wp, int Cin, int Cout, cudaStream_t st) {
  B200_REQUIRE(Cin % 8 == 0 && Cout % 8 == 0 && Cin > 0 && Cout > 0, "pack_conv_weights: channels must be multiples of 8");
  B200_REQUIRE(((uintptr_t)w & 15) == 0 && ((uintptr_t)wp & 15) == 0, "pack_conv_weights: pointers must be 16-byte aligned");
  const int total = 25 * Cin * (Cout / 8);
  static PerDeviceOnce carve;
  if (carve.first()) {
    B200_CUDA_TRY(prefer_max_shared(pack_conv_weights_kernel));
    B200_CUDA_TRY(prefer_max_shared(reduce_partials_tc));
  }
  B200_CUDA_TRY(launch_kernel(pack_conv_weights_kernel, dim3(std::min(cdiv(total, 256), 2 * kNumSMs)), dim3(256), 0, st, w,
                              (uint8_t*)wp, Cin, Cout));
  return B200_OK;
}
int conv5x5_fwd_tc(const float* x, const void* wp, const float* bias, float* y, int B, int H, int W, int Cin,
                   int Cout, int relu, cudaStream_t st, const void* x_bf16) {
  return launch_conv<false>(x, (const uint8_t*)wp, bias, y, B, H, W, Cin, Cout, relu, st, nullptr, nullptr, nullptr, nullptr,
                            nullptr, x_bf16);
}
// conv + bias + ReLU + 3x3/2 SAME max-pool in one kernel; y (the un-pooled activations) is optional
int conv5x5_pool_fwd_tc(const float* x, const void* wp, const float* bias, float* y, float* pool, uint8_t* arg, int B,
                        int H, int W, int Cin, int Cout, cudaStream_t st, const void* x_bf16, void* pool_img) {
  return launch_conv<false>(x, (const uint8_t*)wp, bias, y, B, H, W, Cin, Cout, 1, st, nullptr, nullptr, nullptr, pool, arg, x_bf16,
                            nullptr, pool_img);
}
// dx = Conv2DBackpropInput(dy, w): contraction over the forward Cout, output the forward Cin
int conv5x5_dgrad_tc(const float* dy, const void* wp, float* dx, int B, int H, int W, int Cin, int Cout,
                     cudaStream_t st, const void* dy_bf16) {
  return launch_conv<true>(dy, (const uint8_t*)wp, nullptr, dx, B, H, W, Cout, Cin, 0, st, nullptr, nullptr, nullptr, nullptr,
                           nullptr, dy_bf16);
}
static PoolGradSrc make_pool_src(const float* dpool, const float* pool, const uint8_t* arg, int H, int W) {
  PoolGradSrc p{dpool, pool, arg, (H + 1) / 2, (W + 1) / 2, 0, 0};
  p.pbh = std::max((p.OH - 1) * 2 + 3 - H, 0) / 2;
  p.pbw = std::max((p.OW - 1) * 2 + 3 - W, 0) / 2;
  return p;
}
// dx = Conv2DBackpropInput(dconv, w) with dconv = MaxPoolGrad(dpool, arg) * [pool > 0] gathered on the fly;
// dbias (optional, zero on entry) += sum of dconv over (B,H,W)
int conv5x5_dgrad_pool_tc(const float* dpool, const float* pool, const uint8_t* arg, const void* wp, float* dx,
                          float* dbias, float* dconv_out, int B, int H, int W, int Cin, int Cout, cudaStream_t st,
                          void* dconv_out_bf16, float* dbias_part, int* n_part) {
  B200_REQUIRE(Cin % 64 == 0 && Cout == 64, "conv5x5_dgrad_pool_tc: needs Cout == 64 (got %d -> %d)", Cin, Cout);
  PoolGradSrc p = make_pool_src(dpool, pool, arg, H, W);
  return launch_conv<true>(nullptr, (const uint8_t*)wp, nullptr, dx, B, H, W, Cout, Cin, 0, st, &p, dbias, dconv_out, nullptr,
                           nullptr, nullptr, dconv_out_bf16, nullptr, dbias_part, n_part);
}
int64_t conv5x5_wgrad_tc_scratch_bytes(int B, int Cin, int Cout) {
  return (int64_t)wgrad_splits_tc(B, (Cin / 64) * (Cout / 64)) * 25 * Cin * Cout * 4;
}
static int launch_wgrad(const float* x, const float* dy, const PoolGradSrc* pg, float* dw, float* scratch, int B, int H,
                        int W, int Cin, int Cout, cudaStream_t st, const void* x_bf16 = nullptr,
                        const void* dy_bf16 = nullptr, const float* db_part = nullptr, float* db = nullptr, int n_db_part = 0);
int conv5x5_wgrad_tc(const float* x, const float* dy, float* dw, float* scratch, int B, int H, int W, int Cin,
                     int Cout, cudaStream_t st, const void* x_bf16, const void* dy_bf16, const float* db_part, float* db,
                     int n_db_part) {
  B200_REQUIRE((x_bf16 == nullptr) == (dy_bf16 == nullptr), "conv5x5_wgrad_tc: give both bf16 copies or none");
  return launch_wgrad(x, dy, nullptr, dw, scratch, B, H, W, Cin, Cout, st, x_bf16, dy_bf16, db_part, db, n_db_part);
}
// filter gradient with dY = MaxPoolGrad(dpool, arg) * [pool > 0] gathered on the fly
int conv5x5_wgrad_pool_tc(const float* x, const float* dpool, const float* pool, const uint8_t* arg, float* dw,
                          float* scratch, int B, int H, int W, int Cin, int Cout, cudaStream_t st) {
  PoolGradSrc p = make_pool_src(dpool, pool, arg, H, W);
  return launch_wgrad(x, nullptr, &p, dw, scratch, B, H, W, Cin, Cout, st);
}
static int launch_wgrad(const float* x, const float* dy, const PoolGradSrc* pg, float* dw, float* scratch, int B, int H,
                        int W, int Cin, int Cout, cudaStream_t st, const void* x_bf16, const void* dy_bf16,
                        const float* db_part, float* db, int n_db_part) {
  B200_REQUIRE(Cin > 0 && Cout > 0 && Cin % 64 == 0 && Cout % 64 == 0,
               "conv5x5_wgrad_tc: channels must be multiples of 64 (got %d -> %d)", Cin, Cout);
  B200_REQUIRE(!pg || (Cin == 64 && Cout == 64), "conv5x5_wgrad_tc: the pooled-gradient source needs 64 -> 64 channels");
  B200_REQUIRE(B > 0 && H > 0 && W > 0, "conv5x5_wgrad_tc: bad shape");
  const int nblk = (Cin / 64) * (Cout / 64);
  ConvGeom g = make_geom(B, H, W, 64, 64, 1);
  // compact strips: the k-steps cover H*P slots from `lead`, the taps shift the x operand by at most +-(2P + 2)
  g.npix = g.lead + ((H * g.P + 15) / 16) * 16 + 2 * g.P + 2;
  g.ch_stride = g.npix * 16 + 16;
  const int strip_bytes = ((8 * g.ch_stride) + 127) & ~127;
  const int buf_bytes = ((16 * g.ch_stride + 127) & ~127) + strip_bytes;      // x, shifted x, dY
  const int extra = pg ? pg->OH * pg->OW * 64 * 5 + 16 : 0;
  const int nbuf = (2 * buf_bytes + 256 + extra <= 227 * 1024) ? 2 : 1;
  const int smem = nbuf * buf_bytes + 256 + extra;
  B200_REQUIRE(smem <= 227 * 1024, "conv5x5_wgrad_tc: image %dx%d needs %d bytes of shared memory", H, W, smem);
  static PerDeviceOnce attr;
  if (attr.first())
    B200_CUDA_TRY(cudaFuncSetAttribute(conv5x5_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       227 * 1024));
  int S = wgrad_splits_tc(B, nblk);
  int per = cdiv(B, S);
  S = cdiv(B, per);
  WgradArgs a{x, dy, reinterpret_cast<const __nv_bfloat16*>(x_bf16), reinterpret_cast<const __nv_bfloat16*>(dy_bf16), scratch, g, Cin, Cout, per, nbuf, pg ? *pg : PoolGradSrc{}, g_conv_timeline};
  const bool overlap = g_pdl_next;           // the step driver asked for overlapped launches: only the reduction is
  g_pdl_next = false;                        // programmatic (the kernel in front of this one is on another stream)
  B200_CUDA_TRY(launch_kernel(conv5x5_wgrad_tc_kernel, dim3(5, S, nblk), dim3(416), (size_t)smem, st, a));
  const int n = 25 * Cin * Cout;
  g_pdl_next = overlap;
  const bool with_db = db_part && db && n_db_part > 0;      // the bias-gradient partials of the dgrad kernel ride along
  B200_CUDA_TRY(launch_kernel(reduce_partials_tc, dim3(cdiv(n, 128) + (with_db ? 64 / 4 : 0)), dim3(128), 0, st,
                              ReduceSeg{scratch, dw, n, S, n},
                              with_db ? ReduceSeg{db_part, db, 64, n_db_part, 64} : ReduceSeg{nullptr, nullptr, 0, 0, 0}));
  return B200_OK;
}

}  // namespace b200
