// Inline-PTX wrappers for the sm_100a tensor-core path: mbarrier, tcgen05 (alloc / mma / commit /
// ld / fences), bulk async copies, and the shared-memory matrix / instruction descriptors.
//
// All tensor-core kernels here use the NO-SWIZZLE canonical operand layouts, built from 16-byte
// "chunks" of 8 bf16 values:
//   K-major operand  (rows x K):  core matrix = 8 rows x 16 B, rows contiguous (128 B);
//       SBO = byte stride between 8-row groups, LBO = byte stride between 8-element K chunks.
//   MN-major operand (MN x K):    core matrix = 8 k-rows x 16 B (8 MN elements), contiguous;
//       SBO = byte stride between 8-element MN groups, LBO = byte stride between 8-row K groups.
// One tcgen05.mma.kind::f16 consumes K = 16, i.e. two chunks / groups along K (at +LBO).
// Without swizzling any 16-byte aligned start address is legal, which the implicit-GEMM
// convolutions exploit: a filter tap is a plain address offset into one resident image strip.
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

namespace b200 {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---- mbarrier ---------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must fault the launch, not hang the GPU (a hung box is a strike).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000ll) {   // ~2 s at 1.9 GHz
      printf("b200cnn: mbarrier wait timed out (block %d thread %d)\n", (int)blockIdx.x, (int)threadIdx.x);
      __trap();
    }
  }
}

// generic-proxy smem writes -> visible to the async proxy (tensor core / bulk copy engine)
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- bulk async copy global -> shared (TMA engine, 1-D) --------------------------------------
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

// ---- tensor-map (tiled) TMA load global -> shared, 5-D box; coordinates innermost first, may be negative / out of
// range: elements outside the tensor arrive as zeros.  `tmap` points at a CUtensorMap in kernel-parameter space
// (a member of a __grid_constant__ argument).  Only ever used on tensors in this GPU's own memory.
__device__ __forceinline__ void tma_load_5d(uint32_t dst_smem, const void* tmap, int c0, int c1, int c2, int c3, int c4,
                                            uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
      ::"r"(dst_smem), "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4), "r"(bar)
      : "memory");
}

// ---- tcgen05 ----------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// whole warp; writes the TMEM base address to *smem_dst.  ncols: power of two in 32..512
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// shared-memory matrix descriptor, no swizzle (layout_type 0), descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFFu);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// The same descriptor split in two 32-bit words so that an MMA issue loop only ADDS to the low word:
//   lo = start address (16-byte units, bits 0-13) | LBO (bits 16-29);   hi = SBO | version.
// The single issuing thread is the pacing resource of a tcgen05 kernel with small tiles (one
// M=128,N=64,K=16 MMA takes 48 cycles, measured), so per-MMA integer work must stay at a few adds.
__device__ __forceinline__ uint32_t desc_lo(uint32_t saddr, uint32_t lbo_bytes) {
  return ((saddr >> 4) & 0x3FFFu) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}
__device__ __forceinline__ uint32_t desc_hi(uint32_t sbo_bytes) { return ((sbo_bytes >> 4) & 0x3FFFu) | (1u << 14); }
__device__ __forceinline__ uint64_t desc_join(uint32_t lo, uint32_t hi) { return ((uint64_t)hi << 32) | lo; }

// instruction descriptor for kind::f16: bf16 x bf16 -> f32, M in {64,128}, N % 8 == 0 (16 for M=128)
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, bool a_mn_major, bool b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// one lane of a fully converged warp (the tcgen05 issue pattern: the WHOLE warp walks the loop with
// warp-uniform values, only the instruction itself is predicated on the elected lane -- a loop
// entered by a single lane makes the compiler wrap every UTCHMMA in an elect/retry loop)
__device__ __forceinline__ bool elect_one() {
  uint32_t p;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(p));
  return p != 0;
}
// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrives on `bar` once every previously issued MMA of this thread has completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// TMEM -> registers: this thread's lane (= its row of the 128-row accumulator), 32 consecutive columns
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// 8 values -> one 16-byte bf16 chunk
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 r;
  __nv_bfloat162 a = __floats2bfloat162_rn(f[0], f[1]), b = __floats2bfloat162_rn(f[2], f[3]);
  __nv_bfloat162 c = __floats2bfloat162_rn(f[4], f[5]), d = __floats2bfloat162_rn(f[6], f[7]);
  r.x = *reinterpret_cast<uint32_t*>(&a); r.y = *reinterpret_cast<uint32_t*>(&b);
  r.z = *reinterpret_cast<uint32_t*>(&c); r.w = *reinterpret_cast<uint32_t*>(&d);
  return r;
}
// load one chunk (8 consecutive elements) from global as bf16x8; source fp32 or bf16
__device__ __forceinline__ uint4 load_chunk(const float* p) {
  float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  return pack8(f);
}
__device__ __forceinline__ uint4 load_chunk(const __nv_bfloat16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
// the same chunk, but still in the form it left memory in: lets a kernel issue many loads before the first conversion
template <typename T> struct RawChunk;
template <> struct RawChunk<float> { float4 a, b; };
template <> struct RawChunk<__nv_bfloat16> { uint4 v; };
__device__ __forceinline__ RawChunk<float> load_raw(const float* p, bool ok) {
  RawChunk<float> r;
  r.a = r.b = make_float4(0.f, 0.f, 0.f, 0.f);
  if (ok) { r.a = __ldg(reinterpret_cast<const float4*>(p)); r.b = __ldg(reinterpret_cast<const float4*>(p) + 1); }
  return r;
}
__device__ __forceinline__ RawChunk<__nv_bfloat16> load_raw(const __nv_bfloat16* p, bool ok) {
  RawChunk<__nv_bfloat16> r;
  r.v = make_uint4(0, 0, 0, 0);
  if (ok) r.v = __ldg(reinterpret_cast<const uint4*>(p));
  return r;
}
__device__ __forceinline__ uint4 pack_raw(const RawChunk<float>& r) {
  const float f[8] = {r.a.x, r.a.y, r.a.z, r.a.w, r.b.x, r.b.y, r.b.z, r.b.w};
  return pack8(f);
}
__device__ __forceinline__ uint4 pack_raw(const RawChunk<__nv_bfloat16>& r) { return r.v; }


// ---- MaxPoolGrad by parity-aware gather ---------------------------------------------------------------------------
// 3x3 / stride-2 max-pool with (pbh, pbw) rows / columns of padding in front: conv pixel (i, j) lies in the windows
// (oh, kh) with 2*oh - pbh + kh = i -- ONE window row when i + pbh is odd (kh = 1), TWO when it is even (kh = 0, and
// kh = 2 of the window row above); columns alike.  A pixel's chunk of 8 channels is the sum, over those <= 4
// windows, of the masked pooled gradient of every channel whose argmax code (kh*3 + kw) points at the pixel.
// dpm: masked pooled gradient, fp32, per cell 16 float4 laid out [half][chunk] (channels 8*chunk + 4*half .. +3: the
// eight lanes of a pixel read 128 contiguous bytes), argb: [cells][64] argmax codes, both in shared memory.
template <int NA, int NB>       // window rows / columns covering the pixel (fixed by the parities of i+pbh, j+pbw)
__device__ __forceinline__ void pool_grad_px(const float* dpm, const uint8_t* argb, int OH, int OW, int oh0, int ow0,
                                             int c8, float (&acc)[8]) {
#pragma unroll
  for (int a = 0; a < NA; ++a) {
    const int oh = oh0 - a, kh = a ? 2 : (NA == 2 ? 0 : 1);
    const bool va = a ? oh >= 0 : oh < OH;
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      const int ow = ow0 - b, kw = b ? 2 : (NB == 2 ? 0 : 1);
      const bool v = va && (b ? ow >= 0 : ow < OW);
      const int cell = v ? oh * OW + ow : 0;
      const uint32_t code = v ? (uint32_t)(kh * 3 + kw) * 0x01010101u : 0xffffffffu;     // 0xff matches no code
      const uint2 ab = *reinterpret_cast<const uint2*>(argb + cell * 64 + c8 * 8);
      const float4 d0 = reinterpret_cast<const float4*>(dpm)[cell * 16 + c8];          // channels 8*c8 .. +3
      const float4 d1 = reinterpret_cast<const float4*>(dpm)[cell * 16 + 8 + c8];      // channels 8*c8+4 .. +7
      const uint32_t m0 = __vcmpeq4(ab.x, code), m1 = __vcmpeq4(ab.y, code);             // 0xff per matching byte
      acc[0] += __uint_as_float(__float_as_uint(d0.x) & __byte_perm(m0, 0, 0x0000));
      acc[1] += __uint_as_float(__float_as_uint(d0.y) & __byte_perm(m0, 0, 0x1111));
      acc[2] += __uint_as_float(__float_as_uint(d0.z) & __byte_perm(m0, 0, 0x2222));
      acc[3] += __uint_as_float(__float_as_uint(d0.w) & __byte_perm(m0, 0, 0x3333));
      acc[4] += __uint_as_float(__float_as_uint(d1.x) & __byte_perm(m1, 0, 0x0000));
      acc[5] += __uint_as_float(__float_as_uint(d1.y) & __byte_perm(m1, 0, 0x1111));
      acc[6] += __uint_as_float(__float_as_uint(d1.z) & __byte_perm(m1, 0, 0x2222));
      acc[7] += __uint_as_float(__float_as_uint(d1.w) & __byte_perm(m1, 0, 0x3333));
    }
  }
}
// all 8-channel chunks of image rows [row0, row0 + rows): emit(i, j, c8, acc) is called once per chunk.  NT threads
// (a multiple of 8, so a thread keeps the chunk index tid & 7); pixels are visited parity class by parity class so
// that the window counts above are compile-time constants.
template <int NT, typename Emit>
__device__ __forceinline__ void pool_grad_rows(const float* dpm, const uint8_t* argb, int W, int OH, int OW, int pbh,
                                               int pbw, int row0, int rows, int tid, Emit emit) {
#pragma unroll
  for (int cls = 0; cls < 4; ++cls) {
    const int pi = cls >> 1, pj = cls & 1;                       // parities of i + pbh, j + pbw
    const int i_first = row0 + ((((row0 + pbh) & 1) != pi) ? 1 : 0);
    const int ni = i_first < row0 + rows ? ((row0 + rows - 1 - i_first) >> 1) + 1 : 0;
    const int j_first = ((pbw & 1) != pj) ? 1 : 0;
    const int nj = j_first < W ? ((W - 1 - j_first) >> 1) + 1 : 0;
    const uint32_t magic = nj > 1 ? 0xffffffffu / (uint32_t)nj + 1u : 0u;      // pq / nj = umulhi(pq, magic), small pq
#pragma unroll 2
    for (int q = tid; q < ni * nj * 8; q += NT) {
      const int c8 = q & 7, pq = q >> 3;
      const int ii = nj > 1 ? (int)__umulhi((uint32_t)pq, magic) : pq, jj = pq - ii * nj;
      const int i = i_first + 2 * ii, j = j_first + 2 * jj;
      const int oh0 = (i + pbh) >> 1, ow0 = (j + pbw) >> 1;
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (pi == 0) {
        if (pj == 0) pool_grad_px<2, 2>(dpm, argb, OH, OW, oh0, ow0, c8, acc);
        else pool_grad_px<2, 1>(dpm, argb, OH, OW, oh0, ow0, c8, acc);
      } else {
        if (pj == 0) pool_grad_px<1, 2>(dpm, argb, OH, OW, oh0, ow0, c8, acc);
        else pool_grad_px<1, 1>(dpm, argb, OH, OW, oh0, ow0, c8, acc);
      }
      emit(i, j, c8, acc);
    }
  }
}

}  // namespace tc
}  // namespace b200
