// conv1 of cifar10cnn.py on tcgen05, fused with everything around it.
//
// Forward (`conv1_pool_fwd_tc`): raw uint8 CHW record -> centre crop (:57-68) -> 5x5 SAME conv 3->64
// + bias + ReLU (:107-110) -> 3x3/2 SAME max-pool with argmax (:113), one image per CTA, nothing
// but the pooled tensor and its argmax ever reaches HBM.
//
// The image is a one-chunk strip (see tc_conv.cu): pixel slot = 16 B = 8 bf16 channels, of which
// 3 are real (0..255 is exact in bf16) and 5 are zero.  A K=16 MMA consumes two chunks at distance
// LBO -- and LBO is free, so the second chunk is simply the NEXT FILTER TAP of the same strip:
// LBO = (slot offset of tap 2s+1 - slot offset of tap 2s) * 16.  13 MMAs per 128-slot tile cover the
// 25 taps (+1 dummy with zero weights).
//
// Filter gradient (`conv1_wgrad_tc`): MaxPoolGrad + ReluGrad are applied while building a dY strip
// in shared memory (gather form, deterministic), then  D_di[cout, dj*8+c] = sum_slots
// dY[slot,cout] * X[slot + tap(di,dj), c]  is ONE M=64 x N=40 MMA per filter row and k-step: the
// five taps of a row are the five N-groups of an MN-major B operand whose group stride SBO is one
// pixel slot (16 B).  Channel 3 of the x strip is the constant 1, so column dj*8+3 of the centre
// tap is the bias gradient sum(dY) for free.
#include <stdlib.h>

#include <algorithm>

#include "launch.h"
#include "tc_common.cuh"

namespace b200 {
using namespace tc;

namespace {

struct C1Geom {
  int B, ih, iw, H, W;     // batch, record image size, crop size
  int P, tiles, lead, npix;
  int OH, OW;
};
__host__ inline C1Geom make_c1geom(int B, int ih, int iw, int H, int W) {
  C1Geom g;
  g.B = B; g.ih = ih; g.iw = iw; g.H = H; g.W = W;
  g.P = W + 4;
  g.tiles = ((H + 2) * g.P + 127) / 128;
  g.lead = 2 * g.P + 2;
  g.npix = g.lead + g.tiles * 128 + 2 * g.P + 2;
  g.OH = (H + 1) / 2; g.OW = (W + 1) / 2;
  return g;
}

// one 16-byte slot: channels 0..2 = the pixel's bytes, channel 3 = `c3`, channels 4..7 = 0
__device__ __forceinline__ uint4 pixel_chunk(const uint8_t* __restrict__ rec, int ih, int iw, int y, int x, float c3) {
  const int plane = ih * iw;
  float f[8] = {(float)rec[y * iw + x], (float)rec[plane + y * iw + x], (float)rec[2 * plane + y * iw + x], c3,
                0.f, 0.f, 0.f, 0.f};
  return pack8(f);
}

struct Conv1FwdArgs {
  const uint8_t* img;   // [B,3,ih,iw]
  const float* w;       // [25][3][64] HWIO
  const float* bias;    // [64]
  float* pool;          // [B,OH,OW,64]
  uint8_t* arg;         // [B,OH,OW,64]
  __nv_bfloat16* pool_bf16;   // [B,OH,OW,64] or null: pool rounded to bf16 for the consumers that stage it as an MMA operand
  float* conv_dbg;      // [B,H,W,64] or null (tests: lets the pool fusion be checked on its own input)
  C1Geom g;
  int C1;               // output channels of the layer; a CTA computes the 64 starting at blockIdx.y * 64
  long long* timeline;  // diagnostics: clock64 stamps of CTA 0 in slots 14, 15, 30, 31 of the timeline buffer, or null
};

constexpr int kWLbo = 64 * 16 + 16;      // B operand: [26 k-chunks][64 cout][16 B], chunk stride padded
long long* g_conv1_timeline = nullptr;      // diagnostics (b200_debug_conv_timeline)
constexpr int kConvRow = 68;             // floats per conv-output row in smem (64 + 4: conflict-free STS.128)

// NW worker threads stage the strip / filter and move the conv tile out of TMEM; warps NW/32 .. NW/32+2 issue MMAs; all
// warps run the pool scan (280 instructions per item of four channels, most of them selects / integer compares)
template <int NW>
__global__ void __launch_bounds__(NW + 96, 1) conv1_pool_fwd_kernel(Conv1FwdArgs a) {
  constexpr int WORKERS = NW, MMA0 = NW / 32;
  auto workers_sync = [] { asm volatile("bar.sync 1, %0;" ::"n"(NW) : "memory"); };
  B200_TRACE_SCOPE(TR_CONV1_FWD);
  pdl_launch_dependents();
  const C1Geom& g = a.g;
  extern __shared__ __align__(128) uint8_t smem[];
  const int strip_bytes = (g.npix * 16 + 127) & ~127;
  uint8_t* strip = smem;
  uint8_t* wB = strip + strip_bytes;
  float* convbuf = reinterpret_cast<float*>(wB + 26 * kWLbo + 64);
  uint8_t* ctrl = reinterpret_cast<uint8_t*>(convbuf + (size_t)g.H * g.W * kConvRow);
  uint64_t* ready = reinterpret_cast<uint64_t*>(ctrl);
  uint64_t* accum = ready + 1;              // one per 128-pixel tile (<= 8): the epilogue of a tile overlaps later MMAs
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum + 8);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n = blockIdx.x;
  const int C1 = a.C1, nc0 = blockIdx.y * 64;
  const int tmem_cols = g.tiles * 64 <= 64 ? 64 : g.tiles * 64 <= 128 ? 128 : g.tiles * 64 <= 256 ? 256 : 512;

  if (tid == 0) {
    mbar_init(smem_u32(ready), WORKERS);
    for (int t = 0; t < 8; ++t) mbar_init(smem_u32(accum + t), 1);
    fence_mbar_init();
  }
  if (warp == MMA0) tmem_alloc(smem_u32(tmem_slot), tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const bool stamp = a.timeline && blockIdx.x == 0 && tid == 0;
  if (warp < MMA0) {
    if (stamp) a.timeline[14] = clock64();
    for (int o = tid * 16; o < strip_bytes; o += WORKERS * 16) *reinterpret_cast<uint4*>(strip + o) = make_uint4(0, 0, 0, 0);
    workers_sync();
    const uint8_t* rec = a.img + (size_t)n * 3 * g.ih * g.iw;
    const int oy = (g.ih - g.H) / 2, ox = (g.iw - g.W) / 2;
    // Every global load of this phase (the thread's <= PXB pixels x 3 planes and its <= WB filter chunks x 3
    // channels) is issued before the first conversion: one memory round trip instead of one per item.
    constexpr int PXB = NW >= 512 ? 2 : 3, WB = NW >= 512 ? 4 : 7;      // PXB * NW >= 24 * 24 pixels, WB * NW >= 26 * 64 chunks
    const int plane = g.ih * g.iw;
    for (int base = 0; base < g.H * g.W; base += PXB * WORKERS) {      // (every thread runs the first round: the filter)
      const int q0 = base + tid;
      uint8_t px[PXB][3];
      float wv[WB][3];
#pragma unroll
      for (int u = 0; u < PXB; ++u) {
        const int q = min(q0 + u * WORKERS, g.H * g.W - 1);
        const int i = q / g.W, j = q - i * g.W;
        const uint8_t* src = rec + (i + oy) * g.iw + (j + ox);
        px[u][0] = src[0]; px[u][1] = src[plane]; px[u][2] = src[2 * plane];
      }
      const bool first = base == 0;                      // the filter rides along with the first pixel batch
      if (first) {
#pragma unroll
        for (int u = 0; u < WB; ++u) {
          const int q = min(tid + u * WORKERS, 25 * 64 - 1), cout = q & 63, kc = q >> 6;
#pragma unroll
          for (int e = 0; e < 3; ++e) wv[u][e] = __ldg(a.w + (kc * 3 + e) * C1 + nc0 + cout);
        }
      }
#pragma unroll
      for (int u = 0; u < PXB; ++u) {
        const int q = q0 + u * WORKERS;
        if (q < g.H * g.W) {
          const int i = q / g.W, j = q - i * g.W;
          const float f[8] = {(float)px[u][0], (float)px[u][1], (float)px[u][2], 0.f, 0.f, 0.f, 0.f, 0.f};
          *reinterpret_cast<uint4*>(strip + (g.lead + i * g.P + j) * 16) = pack8(f);
        }
      }
      if (first) {
        // filter: k-chunk kc = tap kc (tap 25 = zeros); element e<3 of the chunk = w[tap][e][cout]
#pragma unroll
        for (int u = 0; u < WB; ++u) {
          const int q = tid + u * WORKERS, cout = q & 63, kc = q >> 6;
          if (q < 26 * 64) {
            const bool real = kc < 25;
            const float f[8] = {real ? wv[u][0] : 0.f, real ? wv[u][1] : 0.f, real ? wv[u][2] : 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            *reinterpret_cast<uint4*>(wB + kc * kWLbo + cout * 16) = pack8(f);
          }
        }
      }
    }
    fence_proxy_async();
    mbar_arrive(smem_u32(ready));
    if (stamp) a.timeline[15] = clock64();

    // ---- epilogue 1: TMEM -> +bias, ReLU -> fp32 conv tile in shared memory, tile by tile as the MMAs finish ----
    for (int t = (warp >> 2); t < g.tiles; t += NW / 128) {
      mbar_wait(smem_u32(accum + t), 0);
      tc_fence_after();
      if (stamp && t == 0) a.timeline[30] = clock64();
      const int p = t * 128 + (warp & 3) * 32 + lane;
      const int r = p / g.P, c = p - r * g.P;
      const bool valid = (r < g.H) && (c < g.W);
      float* row = convbuf + (size_t)(r * g.W + c) * kConvRow;
      float* dbg = a.conv_dbg ? a.conv_dbg + (((size_t)n * g.H + r) * g.W + c) * C1 + nc0 : nullptr;
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)((warp & 3) * 32) << 16) + t * 64 + c0, v);
        float bv[32];                                    // (half of the bias at a time: registers)
#pragma unroll
        for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(bv + j) = __ldg(reinterpret_cast<const float4*>(a.bias + nc0 + c0 + j));
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            float4 o;
            o.x = fmaxf(__uint_as_float(v[j + 0]) + bv[j + 0], 0.f);
            o.y = fmaxf(__uint_as_float(v[j + 1]) + bv[j + 1], 0.f);
            o.z = fmaxf(__uint_as_float(v[j + 2]) + bv[j + 2], 0.f);
            o.w = fmaxf(__uint_as_float(v[j + 3]) + bv[j + 3], 0.f);
            *reinterpret_cast<float4*>(row + c0 + j) = o;
            if (dbg) *reinterpret_cast<float4*>(dbg + c0 + j) = o;
          }
        }
      }
    }
    tc_fence_before();
  } else {
    // three warps with one issuing thread each: tiles t = w, w+3, ... (13 MMAs per tile, one commit per tile)
    if (lane == 0) {
      const int w = warp - MMA0;
      const uint32_t idesc = make_idesc(128, 64, false, false);
      mbar_wait(smem_u32(ready), 0);
      tc_fence_after();
      const uint32_t a_hi = desc_hi(128), b_hi = desc_hi(128);
      const uint32_t a_base = ((smem_u32(strip) >> 4) + g.lead) & 0x3FFFu;      // address field of slot `lead`
      for (int t = w; t < g.tiles; t += 3) {
        uint32_t b_lo = desc_lo(smem_u32(wB), kWLbo);
        const uint32_t d = tmem + t * 64;
#pragma unroll
        for (int ks = 0; ks < 13; ++ks) {
          const int t0 = 2 * ks, t1 = 2 * ks + 1;                               // compile-time after unrolling
          const int off0 = (t0 / 5 - 2) * g.P + (t0 % 5 - 2);
          const int off1 = (t1 < 25) ? (t1 / 5 - 2) * g.P + (t1 % 5 - 2) : off0 + 1;   // dummy tap: zero weights
          const uint32_t a_lo = (a_base + (uint32_t)(off0 + t * 128)) | ((uint32_t)(off1 - off0) << 16);   // LBO = tap distance
          umma_bf16(d, desc_join(a_lo, a_hi), desc_join(b_lo, b_hi), idesc, ks ? 1u : 0u);
          b_lo += 2 * (kWLbo >> 4);
        }
        umma_commit(smem_u32(accum + t));
      }
    }
    __syncwarp();
  }
  __syncthreads();       // the conv tile is complete; all warps pool it
  if (stamp) a.timeline[31] = clock64();
  {
    // ---- epilogue 2: 3x3/2 SAME max-pool (pad 0 before / 1 after), first max in (h,w) scan order ---
    const int pth = max((g.OH - 1) * 2 + 3 - g.H, 0), ptw = max((g.OW - 1) * 2 + 3 - g.W, 0);
    const int pbh = pth / 2, pbw = ptw / 2;
    // One item = 8 channels of one pooled cell (eight independent compare/select chains); window positions outside
    // the image read a clamped address and enter the scan as -FLT_MAX, so there is no branch in the scan.  The scan
    // runs in (kh, kw) order with a strict '>': the first maximum wins, as in the reference's MaxPoolWithArgmax use.
    // (one item = FOUR channels: the sixteen lanes of a pixel then read its 256 bytes as 16-byte pieces of distinct
    //  banks; with eight channels per item lanes c and c+4 of a quarter-warp hit the same banks, a 2-way conflict on
    //  every load of the scan)
#pragma unroll 1
    for (int q = tid; q < g.OH * g.OW * 16; q += (int)blockDim.x) {
      const int c4 = q & 15, cell = q >> 4;
      const int oh = cell / g.OW, ow = cell - oh * g.OW;
      float best[4];
      int bi[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) { best[e] = -3.402823466e38f; bi[e] = 0; }
#pragma unroll
      for (int kh = 0; kh < 3; ++kh) {
        const int h = 2 * oh + kh - pbh;
        const bool vh = (unsigned)h < (unsigned)g.H;
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
          const int w = 2 * ow + kw - pbw;
          const bool ok = vh && (unsigned)w < (unsigned)g.W;
          const float4 v0 = *reinterpret_cast<const float4*>(convbuf + (size_t)(ok ? h * g.W + w : 0) * kConvRow + c4 * 4);
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
      const size_t o = ((size_t)n * g.OH * g.OW + cell) * C1 + nc0 + c4 * 4;
      *reinterpret_cast<float4*>(a.pool + o) = make_float4(best[0], best[1], best[2], best[3]);
      if (a.pool_bf16) {
        const __nv_bfloat162 p0 = __floats2bfloat162_rn(best[0], best[1]), p1 = __floats2bfloat162_rn(best[2], best[3]);
        *reinterpret_cast<uint2*>(a.pool_bf16 + o) = make_uint2(*reinterpret_cast<const uint32_t*>(&p0), *reinterpret_cast<const uint32_t*>(&p1));
      }
      *reinterpret_cast<uint32_t*>(a.arg + o) = (uint32_t)(bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24));
    }
  }
  __syncthreads();
  if (warp == MMA0) {
    tc_fence_after();
    tmem_dealloc(tmem, tmem_cols);
  }
}
}  // namespace

bool conv1_tc_supported(int H, int W, int c_in, int c1) {
  if (c_in != 3 || c1 <= 0 || c1 % 64 != 0) return false;
  C1Geom g = make_c1geom(1, H, W, H, W);
  size_t smem = ((g.npix * 16 + 127) & ~127) + 26 * kWLbo + 64 + (size_t)H * W * kConvRow * 4 + 256;
  return g.tiles * 64 <= 512 && smem <= 227 * 1024;
}

int conv1_pool_fwd_tc(const uint8_t* img, const float* w, const float* bias, float* pool, uint8_t* arg,
                      float* conv_dbg, int B, int ih, int iw, int H, int W, cudaStream_t st, void* pool_bf16, int C1) {
  B200_REQUIRE(B > 0 && H <= ih && W <= iw && H >= 4 && W >= 4, "conv1_pool_fwd_tc: bad shape");
  B200_REQUIRE(conv1_tc_supported(H, W, 3, C1), "conv1_pool_fwd_tc: crop %dx%d x %d channels is not supported", H, W, C1);
  C1Geom g = make_c1geom(B, ih, iw, H, W);
  const int smem = ((g.npix * 16 + 127) & ~127) + 26 * kWLbo + 64 + H * W * kConvRow * 4 + 256;
  // (16 worker warps measured no faster than 8 here -- 0.1179 vs 0.1181 ms per step: the pool scan is bound by the
  //  integer / select pipe, not by latency -- so the smaller CTA stays the default)
  static const int nw = getenv("B200_CONV1_WORKERS") ? atoi(getenv("B200_CONV1_WORKERS")) : 256;
  static PerDeviceOnce attr;
  if (attr.first()) {
    B200_CUDA_TRY(cudaFuncSetAttribute(conv1_pool_fwd_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    B200_CUDA_TRY(cudaFuncSetAttribute(conv1_pool_fwd_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  }
  Conv1FwdArgs a{img, w, bias, pool, arg, reinterpret_cast<__nv_bfloat16*>(pool_bf16), conv_dbg, g, C1, g_conv1_timeline};
  if (nw == 256) conv1_pool_fwd_kernel<256><<<dim3(B, C1 / 64), 256 + 96, smem, st>>>(a);
  else conv1_pool_fwd_kernel<512><<<dim3(B, C1 / 64), 512 + 96, smem, st>>>(a);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace b200

// =====================================================================================================
// conv1 filter + bias gradient
// =====================================================================================================
namespace b200 {
namespace {

constexpr int C1W_BR = 8;      // image rows per band of the dY strip
constexpr int C1W_MAXB = 8;    // bands per image the barrier array allows (crops up to 96 rows)

struct Conv1WgradArgs {
  const uint8_t* img;     // [B,3,ih,iw]
  const float* dpool;     // [B,OH,OW,64]  gradient w.r.t. pool1
  const float* pool;      // [B,OH,OW,64]  pool1 itself (ReluGrad mask: the selected cell is > 0 iff pool1 > 0)
  const uint8_t* arg;     // [B,OH,OW,64]
  float* part;            // [slices][CTAs][4864]: this CTA's sum of dw [25][3][64] followed by db [64] of its channel slice
  C1Geom g;
  int C1;                 // channels of the layer; a CTA works on the 64 starting at blockIdx.y * 64
  long long* timeline;    // diagnostics: clock64 stamps of CTA 0 in slots [25,31) of the timeline buffer, or null
};

// NW worker threads build the strips (the gather is ALU- and latency-bound: with 8 warps it issued on a third of the
// cycles, and a band's 384 items per parity class left half of a second round idle); warps NW/32 .. NW/32+4 issue MMAs
template <int NW>
__global__ void __launch_bounds__(NW + 160, 1) conv1_wgrad_kernel(Conv1WgradArgs a) {
  constexpr int WORKERS = NW, MMA0 = NW / 32;
  auto workers_sync = [] { asm volatile("bar.sync 1, %0;" ::"n"(NW) : "memory"); };
  B200_TRACE_SCOPE(TR_CONV1_WGRAD);
  const C1Geom& g = a.g;
  extern __shared__ __align__(128) uint8_t smem[];
  const int xs_bytes = (g.npix * 16 + 127) & ~127;
  const int ch_stride = g.npix * 16 + 16;
  const int ncell = g.OH * g.OW;
  uint8_t* xs = smem;
  uint8_t* ds = xs + xs_bytes;
  float* dpm = reinterpret_cast<float*>(ds + ((8 * ch_stride + 127) & ~127));
  uint8_t* argb = reinterpret_cast<uint8_t*>(dpm + (size_t)ncell * 64);
  float* plr = reinterpret_cast<float*>(argb + (((size_t)ncell * 64 + 127) & ~127));      // pool1 of the image as fetched
  uint8_t* ctrl = reinterpret_cast<uint8_t*>(plr + (size_t)ncell * 64);
  uint64_t* full = reinterpret_cast<uint64_t*>(ctrl);       // one per row band
  uint64_t* empty = full + C1W_MAXB;
  uint64_t* accum = empty + 1;
  uint64_t* staged = accum + 1;                             // bulk copies of dpool / pool / arg have landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(staged + 1);
  const int nband = (g.H + C1W_BR - 1) / C1W_BR;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int C1 = a.C1, nc0 = blockIdx.y * 64;
  const int nimg = (g.B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int ksteps = (g.H * g.P + 15) / 16;

  if (tid == 0) {
    for (int b = 0; b < C1W_MAXB; ++b) mbar_init(smem_u32(full + b), WORKERS);
    mbar_init(smem_u32(empty), 5);
    mbar_init(smem_u32(accum), 5);
    mbar_init(smem_u32(staged), 1);
    fence_mbar_init();
  }
  if (warp == MMA0) tmem_alloc(smem_u32(tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const bool stamp = a.timeline && blockIdx.x == 0 && tid == 0;
  if (warp < MMA0) {
    if (stamp) a.timeline[25] = clock64();
    for (int o = tid * 16; o < 8 * ch_stride; o += WORKERS * 16) *reinterpret_cast<uint4*>(ds + o) = make_uint4(0, 0, 0, 0);
    const int pth = max((g.OH - 1) * 2 + 3 - g.H, 0), ptw = max((g.OW - 1) * 2 + 3 - g.W, 0);
    const int pbh = pth / 2, pbw = ptw / 2;
    const int oy = (g.ih - g.H) / 2, ox = (g.iw - g.W) / 2;
    for (int it = 0; it < nimg; ++it) {
      const int n = blockIdx.x + it * gridDim.x;
      mbar_wait(smem_u32(empty), (it & 1) ^ 1);          // MMAs of the previous image are done with the strips
      pdl_wait();
      pdl_launch_dependents();
      // pooled gradient, pool1 and argmax of this image: three contiguous blocks, fetched with bulk copies (TMA)
      // while the x strip is built -- as register loads this staging cost 5-8k cycles of round trips
      if (C1 == 64) {
        if (tid == 0) {
          const uint32_t nb4 = (uint32_t)ncell * 64 * 4, nb1 = (uint32_t)ncell * 64, bar = smem_u32(staged);
          const uint8_t* gd = reinterpret_cast<const uint8_t*>(a.dpool + (size_t)n * ncell * 64);
          const uint8_t* gp = reinterpret_cast<const uint8_t*>(a.pool + (size_t)n * ncell * 64);
          mbar_arrive_expect_tx(bar, 2 * nb4 + nb1);
          for (uint32_t off = 0; off < nb4; off += 16384) {
            bulk_g2s(smem_u32(dpm) + off, gd + off, min(16384u, nb4 - off), bar);
            bulk_g2s(smem_u32(plr) + off, gp + off, min(16384u, nb4 - off), bar);
          }
          bulk_g2s(smem_u32(argb), a.arg + (size_t)n * ncell * 64, nb1, bar);
        }
      } else if (warp == 0) {
        // a 64-channel slice of a wider layer: one 256 / 256 / 64-byte piece per pooled cell, issued by a whole warp
        const uint32_t bar = smem_u32(staged);
        if (lane == 0) mbar_arrive_expect_tx(bar, (uint32_t)ncell * (256 + 256 + 64));
        __syncwarp();
        const size_t base = (size_t)n * ncell * C1 + nc0;
        for (int cell = lane; cell < ncell; cell += 32) {
          bulk_g2s(smem_u32(dpm) + cell * 256, a.dpool + base + (size_t)cell * C1, 256, bar);
          bulk_g2s(smem_u32(plr) + cell * 256, a.pool + base + (size_t)cell * C1, 256, bar);
          bulk_g2s(smem_u32(argb) + cell * 64, a.arg + base + (size_t)cell * C1, 64, bar);
        }
      }
      // x strip: channels 0..2 = pixel bytes (0 in the pads), channel 3 = 1 in EVERY slot
      const uint8_t* rec = a.img + (size_t)n * 3 * g.ih * g.iw;
      for (int s = tid; s < g.npix; s += WORKERS) {
        const int p = s - g.lead;
        const int i = p >= 0 ? p / g.P : -1, j = p >= 0 ? p - i * g.P : 0;
        uint4 v;
        if (p >= 0 && i < g.H && j < g.W) v = pixel_chunk(rec, g.ih, g.iw, i + oy, j + ox, 1.f);
        else { float f[8] = {0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 0.f}; v = pack8(f); }
        *reinterpret_cast<uint4*>(xs + s * 16) = v;
      }
      mbar_wait(smem_u32(staged), it & 1);
      // ReluGrad mask (the selected cell is > 0 iff pool1 > 0) and the [half][chunk] order pool_grad_px reads, in
      // place: a batch is read into registers, then written (the permutation stays inside a cell's 16 float4)
      constexpr int MB = NW >= 512 ? 5 : 9;
      for (int base = 0; base < ncell * 16; base += MB * WORKERS) {      // (uniform trip count: the loop holds a barrier)
        const int q0 = base + tid;
        float4 dreg[MB];
#pragma unroll
        for (int k = 0; k < MB; ++k) {
          const int q = min(q0 + k * WORKERS, ncell * 16 - 1);
          float4 d = reinterpret_cast<const float4*>(dpm)[q];
          const float4 pv = reinterpret_cast<const float4*>(plr)[q];
          d.x = pv.x > 0.f ? d.x : 0.f; d.y = pv.y > 0.f ? d.y : 0.f;
          d.z = pv.z > 0.f ? d.z : 0.f; d.w = pv.w > 0.f ? d.w : 0.f;
          dreg[k] = d;
        }
        workers_sync();
#pragma unroll
        for (int k = 0; k < MB; ++k) {
          const int q = q0 + k * WORKERS;
          if (q < ncell * 16) reinterpret_cast<float4*>(dpm)[(q & ~15) + (q & 1) * 8 + ((q & 15) >> 1)] = dreg[k];
        }
      }
      workers_sync();
      if (stamp) a.timeline[26] = clock64();
      // dY strip = MaxPoolGrad of the masked gradient (parity-aware gather, tc_common.cuh) in bands of C1W_BR image
      // rows: each finished band is released to the MMA warps while the next one is built
      for (int b = 0; b < nband; ++b) {
        const int row0 = b * C1W_BR, rows = min(C1W_BR, g.H - row0);
        pool_grad_rows<WORKERS>(dpm, argb, g.W, g.OH, g.OW, pbh, pbw, row0, rows, tid,
                                [&](int i, int j, int c8, const float (&acc)[8]) {
                                  *reinterpret_cast<uint4*>(ds + c8 * ch_stride + (g.lead + i * g.P + j) * 16) = pack8(acc);
                                });
        fence_proxy_async();
        mbar_arrive(smem_u32(full + b));
      }
      if (stamp) a.timeline[27] = clock64();
    }
    // ---- epilogue: D_di[cout][dj*8 + c] (M = 64: lanes 0-15 of each subpartition) -> partial --------
    mbar_wait(smem_u32(accum), 0);
    tc_fence_after();
    if (stamp) a.timeline[28] = clock64();
    if (warp < 4 && nimg > 0) {
      // D_di[cout][dj*8 + c] -> this CTA's partial of dW1[(di*5+dj)][c][cout] and db1[cout]: 16 adjacent couts per
      // warp instruction.  The partials are summed in a fixed order afterwards (run-to-run identical gradients;
      // fp32 atomics here made the last bits depend on the order in which CTAs finished).
      const int cout = warp * 16 + lane;
      float* part = a.part + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 4864;
      for (int di = 0; di < 5; ++di) {
#pragma unroll 1
        for (int c0 = 0; c0 < 64; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + di * 64 + c0, v);
          tmem_ld_wait();
          if (lane < 16) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int col = c0 + j, dj = col >> 3, c = col & 7;
              if (col < 40 && c < 3) part[((di * 5 + dj) * 3 + c) * 64 + cout] = __uint_as_float(v[j]);
              if (di == 2 && col == 2 * 8 + 3) part[4800 + cout] = __uint_as_float(v[j]);   // constant-1 channel
            }
          }
        }
      }
    }
    tc_fence_before();
    if (stamp) a.timeline[29] = clock64();
  } else {
    // five warps: one issuing thread per filter row di (own accumulator each)
    if (lane == 0) {
      const int di = warp - MMA0;
      const uint32_t idesc = make_idesc(64, 40, true, true);
      const uint32_t d_hi = desc_hi(ch_stride), x_hi = desc_hi(16);     // x: N-group stride = one pixel slot
      const uint32_t d_lo0 = desc_lo(smem_u32(ds) + g.lead * 16, 128);
      const uint32_t x_lo0 = desc_lo(smem_u32(xs) + (g.lead + (di - 2) * g.P - 2) * 16, 128);
      const uint32_t d = tmem + di * 64;
      uint32_t acc = 0;
      for (int it = 0; it < nimg; ++it) {
        uint32_t d_lo = d_lo0, x_lo = x_lo0;
        int ks = 0;
        for (int b = 0; b < nband; ++b) {
          mbar_wait(smem_u32(full + b), it & 1);
          tc_fence_after();
          // k-steps (16 pixel slots each) that lie completely inside the rows finished so far
          const int kend = (b == nband - 1) ? ksteps : (min(g.H, (b + 1) * C1W_BR) * g.P) / 16;
          for (; ks < kend; ++ks) {
            umma_bf16(d, desc_join(d_lo, d_hi), desc_join(x_lo, x_hi), idesc, acc | (uint32_t)ks);
            d_lo += 16; x_lo += 16;
          }
        }
        acc = 1;
        umma_commit(smem_u32(empty));
      }
      umma_commit(smem_u32(accum));
    }
    __syncwarp();
  }
  __syncthreads();
  if (warp == MMA0) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

int conv1_wgrad_ctas(int B) {
  static int env = -1;      // B200_WGRAD1_CTAS: experiments (see wgrad_splits_tc in tc_conv.cu)
  if (env < 0) { const char* e = getenv("B200_WGRAD1_CTAS"); env = e ? atoi(e) : 0; }
  if (env > 0) return std::min(B, std::min(env, kNumSMs));
  // up to one wave of images: two images per CTA, so that conv2's filter gradient fits beside this kernel
  return B <= kNumSMs ? (B + 1) / 2 : kNumSMs;
}
size_t conv1_wgrad_smem(const C1Geom& g) {
  size_t xs = (g.npix * 16 + 127) & ~127, ch = g.npix * 16 + 16;
  return xs + ((8 * ch + 127) & ~127) + 2 * (size_t)g.OH * g.OW * 64 * 4 + (((size_t)g.OH * g.OW * 64 + 127) & ~127) +
         256;
}
}  // namespace

int conv1_wgrad_tc_partials(int B) { return conv1_wgrad_ctas(B); }
int64_t conv1_wgrad_tc_scratch_bytes(int B, int C1) { return (int64_t)conv1_wgrad_ctas(B) * (C1 / 64) * 4864 * 4; }

namespace {
// wider layers: the per-slice partials [slice][S][4864] summed (same warp order as reduce_partials_warp) into
// dw [25][3][C1] and db [C1]
__global__ void conv1_reduce_slices_kernel(const float* __restrict__ part, float* __restrict__ dw, float* __restrict__ db,
                                           int S, int C1) {
  pdl_wait();
  const int slice = blockIdx.y;
  const int t = blockIdx.x * blockDim.x + threadIdx.x, lane = t & 31;
  const int grp = min(t >> 5, 4864 / 4 - 1), i = grp * 4;
  const float* p = part + (size_t)slice * S * 4864;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int z0 = lane; z0 < S; z0 += 128) {
    float4 v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int z = z0 + 32 * j;
      v[j] = z < S ? __ldg(reinterpret_cast<const float4*>(p + (size_t)z * 4864 + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) { s.x += v[j].x; s.y += v[j].y; s.z += v[j].z; s.w += v[j].w; }
  }
#pragma unroll
  for (int m = 1; m <= 16; m <<= 1) {
    s.x += __shfl_xor_sync(0xffffffffu, s.x, m); s.y += __shfl_xor_sync(0xffffffffu, s.y, m);
    s.z += __shfl_xor_sync(0xffffffffu, s.z, m); s.w += __shfl_xor_sync(0xffffffffu, s.w, m);
  }
  if (lane == 0 && (t >> 5) < 4864 / 4) {
    if (i < 4800) *reinterpret_cast<float4*>(dw + (size_t)(i >> 6) * C1 + slice * 64 + (i & 63)) = s;
    else if (db) *reinterpret_cast<float4*>(db + slice * 64 + (i - 4800)) = s;
  }
}
}  // namespace

int conv1_wgrad_tc(const uint8_t* img, const float* dpool, const float* pool, const uint8_t* arg, float* dw, float* db,
                   float* scratch, int B, int ih, int iw, int H, int W, cudaStream_t st, bool defer_reduce, int C1) {
  B200_REQUIRE(B > 0 && H <= ih && W <= iw && H >= 4 && W >= 4, "conv1_wgrad_tc: bad shape");
  B200_REQUIRE(C1 > 0 && C1 % 64 == 0 && (C1 == 64 || !defer_reduce), "conv1_wgrad_tc: %d channels", C1);
  B200_REQUIRE(scratch && ((uintptr_t)scratch & 15) == 0, "conv1_wgrad_tc: scratch must be a 16-byte aligned buffer");
  C1Geom g = make_c1geom(B, ih, iw, H, W);
  const size_t smem = conv1_wgrad_smem(g);
  B200_REQUIRE(smem <= 227 * 1024, "conv1_wgrad_tc: crop %dx%d needs %zu bytes of shared memory", H, W, smem);
  static const int nw = getenv("B200_WGRAD1_WORKERS") ? atoi(getenv("B200_WGRAD1_WORKERS")) : 512;      // experiments: 256
  static PerDeviceOnce attr;
  if (attr.first()) {
    B200_CUDA_TRY(cudaFuncSetAttribute(conv1_wgrad_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    B200_CUDA_TRY(cudaFuncSetAttribute(conv1_wgrad_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  }
  const int S = conv1_wgrad_ctas(B);
  Conv1WgradArgs a{img, dpool, pool, arg, scratch, g, C1, g_conv1_timeline};
  const bool overlap = g_pdl_next;
  g_pdl_next = false;
  if (nw == 256) B200_CUDA_TRY(launch_kernel(conv1_wgrad_kernel<256>, dim3(S, C1 / 64), dim3(256 + 160), smem, st, a));
  else B200_CUDA_TRY(launch_kernel(conv1_wgrad_kernel<512>, dim3(S, C1 / 64), dim3(512 + 160), smem, st, a));
  if (defer_reduce) return B200_OK;
  g_pdl_next = overlap;                                   // (the reduction's prologue may overlap the kernel's tail)
  if (C1 != 64) {
    B200_CUDA_TRY(launch_kernel(conv1_reduce_slices_kernel, dim3((4864 / 4 * 32 + 255) / 256, C1 / 64), dim3(256), 0, st,
                                (const float*)scratch, dw, db, S, C1));
    return B200_OK;
  }
  if (db == dw + 4800) return reduce_partials_warp(scratch, dw, 4864, S, 4864, st);
  B200_TRY(reduce_partials_warp(scratch, dw, 4800, S, 4864, st));
  return db ? reduce_partials_warp(scratch + 4800, db, 64, S, 4864, st) : B200_OK;
}

void conv1_set_timeline(long long* dev_buf) { g_conv1_timeline = dev_buf; }
B200_TRACE_SETTER(trace_set_tc_conv1)

}  // namespace b200
