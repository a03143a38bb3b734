// tcgen05 GEMM for the fully connected layers (cifar10cnn.py:133/:139 and their gradients):
//   C[M,N] (fp32) = A[M,K] * B[K,N], bf16 operands converted on the fly from fp32 or bf16
//   global tensors, fp32 accumulation in TMEM, fused epilogue (bias, ReLU, ReluGrad mask).
// Either operand may be K-major or MN-major in global memory, so forward (x W), dX (dY W^T) and
// dW (X^T dY) all read the row-major tensors of the TF layouts directly -- no transposes.
//
// CTA = 128 x BN output tile, 9 warps: warps 0-7 stage operands (LDG -> bf16 -> STS in the
// no-swizzle canonical layout, chunk stride padded by 16 B so the stores are conflict-free) and
// later warps 0-3 run the epilogue (warp w owns TMEM lanes 32w..32w+31); warp 8 allocates TMEM and issues
// the MMAs from one thread.  4-stage ring, full/empty mbarriers; tcgen05.commit frees a stage.
#include <algorithm>

#include "launch.h"
#include "tc_common.cuh"

namespace b200 {
using namespace tc;

namespace {
constexpr int BM = 128, BK = 64, STAGES = 4;
constexpr int LOADERS = 256;     // 8 staging warps (warps 0-3 also run the epilogue); warp 8 issues the MMAs

struct GemmArgs {
  int M, N, K;
  const void* A; int64_t lda;
  const void* B; int64_t ldb;
  float* C; int64_t ldc;
  const float* bias;
  int relu;
  const float* mask; int64_t ldmask;
  float* colsum;           // optional: colsum[n] = sum_m C[m,n] (after mask), needs gridDim.y == 1
  int k_per_split;         // multiple of BK; blockIdx.z covers K range [z*kps, (z+1)*kps)
  int64_t split_stride;    // floats between the C outputs of consecutive splits
  int trace_id;            // TraceId of this launch (diagnostics build)
  // Operand IMAGES (optional, both or none): the operand already in bf16 as [chunk][rows][8] -- for a K-major operand
  // chunk = k / 8 and rows = m (or n), for an MN-major one chunk = m / 8 (n / 8) and rows = k.  That is the shared-
  // memory layout of a stage, so one thread fetches a stage with 16-24 bulk copies (TMA) and no warp converts anything.
  const uint8_t* A_img; int A_rows;
  const uint8_t* B_img; int B_rows;
  long long* timeline;     // diagnostics: clock64 stamps of CTA (0,0,0) in slots [0,7) of the timeline buffer, or null
};

template <int BN>
struct Smem {
  // K-major: [8 chunks][rows][16 B], chunk stride rows*16+16.  MN-major: [groups][64 k][16 B], stride 1040.
  static constexpr int A_K_LBO = BM * 16 + 16, A_MN_SBO = BK * 16 + 16;
  static constexpr int B_K_LBO = BN * 16 + 16, B_MN_SBO = BK * 16 + 16;
  static constexpr int A_BYTES = (8 * A_K_LBO > (BM / 8) * A_MN_SBO) ? 8 * A_K_LBO : (BM / 8) * A_MN_SBO;
  static constexpr int B_BYTES = (8 * B_K_LBO > (BN / 8) * B_MN_SBO) ? 8 * B_K_LBO : (BN / 8) * B_MN_SBO;
  static constexpr int STAGE = ((A_BYTES + B_BYTES + 127) / 128) * 128;
  static constexpr int TOTAL = STAGES * STAGE + 256;
};

template <int BN, bool A_MN, bool B_MN, typename TA, typename TB>
// (register cap: warps 0, 4 and 8 share one SM sub-partition; at 136 registers they leave room there for two warps of
//  a small kernel -- the FC tail's gradient kernel runs beside FC1's dW and could not share an SM with it at 153)
__global__ void __maxnreg__(136) tc_gemm_kernel(GemmArgs g) {
  B200_TRACE_SCOPE(g.trace_id);
  using S = Smem<BN>;
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* ctrl = smem + STAGES * S::STAGE;
  uint64_t* full = reinterpret_cast<uint64_t*>(ctrl);
  uint64_t* empty = full + STAGES;
  uint64_t* accum = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int kbeg = blockIdx.z * g.k_per_split, kend = min(g.K, kbeg + g.k_per_split);
  const int nkb = (kend - kbeg + BK - 1) / BK;

  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(smem_u32(full + s), g.A_img ? 1 : LOADERS); mbar_init(smem_u32(empty + s), 1); }
    mbar_init(smem_u32(accum), 1);
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc(smem_u32(tmem_slot), BN < 32 ? 32 : BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const bool stamp = g.timeline && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
  if (stamp && tid == 0) g.timeline[0] = clock64();

  if (warp < 8) {
    // ------------------------------ operand staging ------------------------------------------
    pdl_wait();      // barrier / TMEM set-up above overlapped the producer; the operands are its output
    pdl_launch_dependents();   // after our own wait: a dependent overlaps this kernel only, never our producer
    if (stamp && tid == 0) g.timeline[1] = clock64();
    if (g.A_img) {
      // ---- operand images: warp 0 streams the stages with bulk copies, one copy per lane (a bulk copy has a fixed
      //      issue cost of ~100 cycles; 16-24 of them from one thread were most of a stage's latency) -----------------
      if (warp == 0) {
        const int mrows = min(BM, g.M - m0), nrows = min(BN, g.N - n0);
        const uint32_t a_bytes = A_MN ? (uint32_t)(BM / 8) * BK * 16 : 8u * mrows * 16;
        const uint32_t b_bytes = B_MN ? (uint32_t)(BN / 8) * BK * 16 : 8u * nrows * 16;
        constexpr int NCA = A_MN ? BM / 8 : 8, NCB = B_MN ? BN / 8 : 8;      // copies per stage: lanes [0, NCA), [NCA, NCA+NCB)
        static_assert(NCA + NCB <= 32, "one copy per lane");
        for (int kb = 0; kb < nkb; ++kb) {
          const int s = kb % STAGES;
          const uint32_t bar = smem_u32(full + s);
          if (lane == 0) {
            mbar_wait(smem_u32(empty + s), ((kb / STAGES) & 1) ^ 1);
            mbar_arrive_expect_tx(bar, a_bytes + b_bytes);
          }
          __syncwarp();
          const uint32_t sa = smem_u32(smem + s * S::STAGE), sb = sa + S::A_BYTES;
          const int k0 = kbeg + kb * BK;
          if (lane < NCA) {
            if (!A_MN)      // chunk = k / 8, rows = m
              bulk_g2s(sa + lane * S::A_K_LBO, g.A_img + ((int64_t)(k0 / 8 + lane) * g.A_rows + m0) * 16, mrows * 16, bar);
            else            // chunk = m / 8, rows = k
              bulk_g2s(sa + lane * S::A_MN_SBO, g.A_img + ((int64_t)(m0 / 8 + lane) * g.A_rows + k0) * 16, BK * 16, bar);
          } else if (lane < NCA + NCB) {
            const int l = lane - NCA;
            if (!B_MN)      // chunk = k / 8, rows = n
              bulk_g2s(sb + l * S::B_K_LBO, g.B_img + ((int64_t)(k0 / 8 + l) * g.B_rows + n0) * 16, nrows * 16, bar);
            else            // chunk = n / 8, rows = k
              bulk_g2s(sb + l * S::B_MN_SBO, g.B_img + ((int64_t)(n0 / 8 + l) * g.B_rows + k0) * 16, BK * 16, bar);
          }
        }
      }
    } else {
    const TA* A = reinterpret_cast<const TA*>(g.A);
    const TB* B = reinterpret_cast<const TB*>(g.B);
    // Each thread moves NA + NB chunks (8 elements) per k-block.  The global loads of a k-block are all issued
    // before anything is converted, and the NEXT k-block's loads are issued before the current one is converted
    // and stored: two k-blocks (96 KB per CTA) in flight instead of a round trip per operand per k-block -- with
    // one CTA per SM the load latency, not bandwidth, bounded this kernel.
    constexpr int NA = BM * 8 / LOADERS, NB = BN * 8 / LOADERS;
    static_assert(BM * 8 % LOADERS == 0 && BN * 8 % LOADERS == 0, "chunks must divide over the staging threads");
    auto issue = [&](int kb, RawChunk<TA> (&ra)[NA], RawChunk<TB> (&rb)[NB]) {
      const int k0 = kbeg + kb * BK;
#pragma unroll
      for (int i = 0; i < NA; ++i) {
        const int q = tid + i * LOADERS;
        if (!A_MN) {   // A(m,k) = A[m*lda + k]: 8 consecutive threads read one row's 8 chunks
          const int c = q & 7, r = q >> 3, m = m0 + r, k = k0 + c * 8;
          ra[i] = load_raw(A + (int64_t)m * g.lda + k, m < g.M && k < kend);
        } else {       // A(m,k) = A[k*lda + m]: consecutive threads read consecutive m-chunks of one k row
          const int gi = q % (BM / 8), kk = q / (BM / 8), m = m0 + gi * 8, k = k0 + kk;
          ra[i] = load_raw(A + (int64_t)k * g.lda + m, m < g.M && k < kend);
        }
      }
#pragma unroll
      for (int i = 0; i < NB; ++i) {
        const int q = tid + i * LOADERS;
        if (!B_MN) {   // B(k,n) = B[n*ldb + k]
          const int c = q & 7, r = q >> 3, n = n0 + r, k = k0 + c * 8;
          rb[i] = load_raw(B + (int64_t)n * g.ldb + k, n < g.N && k < kend);
        } else {       // B(k,n) = B[k*ldb + n]
          const int gi = q % (BN / 8), kk = q / (BN / 8), n = n0 + gi * 8, k = k0 + kk;
          rb[i] = load_raw(B + (int64_t)k * g.ldb + n, n < g.N && k < kend);
        }
      }
    };
    auto commit = [&](int kb, const RawChunk<TA> (&ra)[NA], const RawChunk<TB> (&rb)[NB]) {
      const int s = kb % STAGES;
      const uint32_t ph = (kb / STAGES) & 1;
      mbar_wait(smem_u32(empty + s), ph ^ 1);
      uint8_t* sa = smem + s * S::STAGE;
      uint8_t* sb = sa + S::A_BYTES;
#pragma unroll
      for (int i = 0; i < NA; ++i) {
        const int q = tid + i * LOADERS;
        if (!A_MN) *reinterpret_cast<uint4*>(sa + (q & 7) * S::A_K_LBO + (q >> 3) * 16) = pack_raw(ra[i]);
        else *reinterpret_cast<uint4*>(sa + (q % (BM / 8)) * S::A_MN_SBO + (q / (BM / 8)) * 16) = pack_raw(ra[i]);
      }
#pragma unroll
      for (int i = 0; i < NB; ++i) {
        const int q = tid + i * LOADERS;
        if (!B_MN) *reinterpret_cast<uint4*>(sb + (q & 7) * S::B_K_LBO + (q >> 3) * 16) = pack_raw(rb[i]);
        else *reinterpret_cast<uint4*>(sb + (q % (BN / 8)) * S::B_MN_SBO + (q / (BN / 8)) * 16) = pack_raw(rb[i]);
      }
      fence_proxy_async();
      mbar_arrive(smem_u32(full + s));
    };
    RawChunk<TA> ra0[NA], ra1[NA];
    RawChunk<TB> rb0[NB], rb1[NB];
    if (nkb > 0) issue(0, ra0, rb0);
#pragma unroll 1
    for (int kb = 0; kb < nkb; kb += 2) {
      if (kb + 1 < nkb) issue(kb + 1, ra1, rb1);
      commit(kb, ra0, rb0);
      if (kb + 2 < nkb) issue(kb + 2, ra0, rb0);
      if (kb + 1 < nkb) commit(kb + 1, ra1, rb1);
    }
    }
    // ------------------------------ epilogue (warps 0-3: one TMEM lane quarter each) -----------------
    if (warp < 4) {
    mbar_wait(smem_u32(accum), 0);
    tc_fence_after();
    if (stamp && tid == 0) g.timeline[4] = clock64();
    const int m = m0 + warp * 32 + lane;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
      tmem_ld_wait();
      if (stamp && tid == 0 && c0 == 0) g.timeline[6] = clock64();
      if (stamp && tid == 0 && c0 == 32) g.timeline[7] = clock64();
      // (conditions hoisted out of the per-element loops: as per-element branches this block cost 3.8k cycles)
      float o[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) o[j] = __uint_as_float(v[j]);
      const bool whole = n0 + c0 + 32 <= g.N;
      if (g.bias) {
        if (whole) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(g.bias + n0 + c0 + j));
            o[j] += b.x; o[j + 1] += b.y; o[j + 2] += b.z; o[j + 3] += b.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) o[j] += (n0 + c0 + j < g.N) ? __ldg(g.bias + min(n0 + c0 + j, g.N - 1)) : 0.f;
        }
      }
      if (g.relu) {
#pragma unroll
        for (int j = 0; j < 32; ++j) o[j] = fmaxf(o[j], 0.f);
      }
      if (g.mask && m < g.M) {
        const float* mk = g.mask + (int64_t)m * g.ldmask + n0 + c0;
        if (whole && (g.ldmask & 3) == 0) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 q = __ldg(reinterpret_cast<const float4*>(mk + j));
            o[j] = q.x > 0.f ? o[j] : 0.f; o[j + 1] = q.y > 0.f ? o[j + 1] : 0.f;
            o[j + 2] = q.z > 0.f ? o[j + 2] : 0.f; o[j + 3] = q.w > 0.f ? o[j + 3] : 0.f;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (n0 + c0 + j < g.N && !(__ldg(mk + j) > 0.f)) o[j] = 0.f;
        }
      }
      if (n0 + c0 + 32 <= g.N && ((g.ldc & 3) == 0)) {
        // A lane holds one ROW of the tile; storing it directly makes every store instruction 32 separate 16-byte
        // requests (measured: 8.6k cycles for this 32 KB tile).  The 32 x 32 block goes through shared memory (the
        // stage buffers are idle now; 36-float rows: conflict-free both ways) and leaves as 4 rows x 128 contiguous
        // bytes per instruction.
        float* tb = reinterpret_cast<float*>(smem) + warp * (32 * 36);
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<float4*>(tb + lane * 36 + j) = make_float4(o[j], o[j + 1], o[j + 2], o[j + 3]);
        __syncwarp();
        float* out = g.C + (int64_t)blockIdx.z * g.split_stride + (int64_t)(m0 + warp * 32) * g.ldc + n0 + c0;
#pragma unroll
        for (int r0 = 0; r0 < 32; r0 += 4) {
          const int r = r0 + (lane >> 3), cc = (lane & 7) * 4;
          const float4 x = *reinterpret_cast<const float4*>(tb + r * 36 + cc);
          if (m0 + warp * 32 + r < g.M) *reinterpret_cast<float4*>(out + (int64_t)r * g.ldc + cc) = x;
        }
      } else if (m < g.M) {
        float* out = g.C + (int64_t)blockIdx.z * g.split_stride + (int64_t)m * g.ldc + n0 + c0;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (n0 + c0 + j < g.N) out[j] = o[j];
      }
      if (g.colsum) {   // column sums over the rows of this tile (rows >= M contribute 0): warp shuffle tree
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float x = (m < g.M) ? o[j] : 0.f;
#pragma unroll
          for (int off = 16; off > 0; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
          if (lane == 0 && n0 + c0 + j < g.N) atomicAdd(g.colsum + n0 + c0 + j, x);
        }
      }
    }
    }
    tc_fence_before();
    if (stamp && tid == 0) g.timeline[5] = clock64();
  } else {
    // ------------------------------ MMA issue (one thread) -----------------------------------------
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(BM, BN, A_MN, B_MN);
      const uint32_t a_hi = A_MN ? desc_hi(S::A_MN_SBO) : desc_hi(128), b_hi = B_MN ? desc_hi(S::B_MN_SBO) : desc_hi(128);
      const uint32_t a_lo0 = A_MN ? desc_lo(smem_u32(smem), 128) : desc_lo(smem_u32(smem), S::A_K_LBO);
      const uint32_t b_lo0 = B_MN ? desc_lo(smem_u32(smem) + S::A_BYTES, 128) : desc_lo(smem_u32(smem) + S::A_BYTES, S::B_K_LBO);
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % STAGES;
        const uint32_t ph = (kb / STAGES) & 1;
        mbar_wait(smem_u32(full + s), ph);
        tc_fence_after();
        if (stamp && kb == 0) g.timeline[2] = clock64();
        uint32_t a_lo = a_lo0 + (uint32_t)s * (S::STAGE >> 4), b_lo = b_lo0 + (uint32_t)s * (S::STAGE >> 4);
#pragma unroll
        for (int ks = 0; ks < BK / 16; ++ks) {
          umma_bf16(tmem, desc_join(a_lo, a_hi), desc_join(b_lo, b_hi), idesc, (kb | ks) ? 1u : 0u);
          a_lo += A_MN ? 16 : 2 * (S::A_K_LBO >> 4);
          b_lo += B_MN ? 16 : 2 * (S::B_K_LBO >> 4);
        }
        umma_commit(smem_u32(empty + s));
      }
      umma_commit(smem_u32(accum));
      if (stamp) g.timeline[3] = clock64();
    }
    __syncwarp();
  }
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc(tmem, BN < 32 ? 32 : BN);
  }
}

long long* g_gemm_timeline = nullptr;

template <int BN, bool A_MN, bool B_MN, typename TA, typename TB>
int launch(const GemmArgs& g_in, cudaStream_t st) {
  GemmArgs g = g_in;
  g.timeline = g_gemm_timeline;
  using S = Smem<BN>;
  auto kern = tc_gemm_kernel<BN, A_MN, B_MN, TA, TB>;
  static PerDeviceOnce attr_done;      // per instantiation and per device
  if (attr_done.first()) B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
  dim3 grid(cdiv(g.N, BN), cdiv(g.M, BM), cdiv(g.K, g.k_per_split));
  B200_CUDA_TRY(launch_kernel(kern, grid, dim3(288), (size_t)S::TOTAL, st, g));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <bool A_MN, bool B_MN>
int dispatch(const GemmArgs& g, cudaStream_t st) {
  // BN = 64 keeps more CTAs in flight for these small problems
  return launch<64, A_MN, B_MN, float, float>(g, st);
}

int check_align(const GemmArgs& g, bool a_mn, bool b_mn) {
  B200_REQUIRE(g.M > 0 && g.N > 0 && g.K > 0, "tc_gemm: bad shape");
  B200_REQUIRE(((uintptr_t)g.A & 15) == 0 && ((uintptr_t)g.B & 15) == 0 && (g.lda % 4) == 0 && (g.ldb % 4) == 0,
               "tc_gemm: operands must be 16-byte aligned with leading dimensions that are multiples of 4");
  B200_REQUIRE((a_mn ? g.M : g.K) % 8 == 0 && (b_mn ? g.N : g.K) % 8 == 0,
               "tc_gemm: contiguous operand dimensions must be multiples of 8 (M=%d N=%d K=%d)", g.M, g.N, g.K);
  return B200_OK;
}
}  // namespace

// y = relu?(x @ w + b): A = x [B,K] K-major, B = w [K,N] MN-major
int fc_fwd_tc(const float* x, const float* w, const float* b, float* y, int Bt, int K, int N, int relu,
              cudaStream_t st) {
  GemmArgs g{Bt, N, K, x, K, w, N, y, N, b, relu, nullptr, 0, nullptr, (int)round_up(K, BK), 0, TR_FC1_FWD};
  int rc = check_align(g, false, true);
  if (rc) return rc;
  return dispatch<false, true>(g, st);
}

// split-K forward product without epilogue: part[s][B][N] = x[:, ks] @ w[ks, :]; the consumer
// (fc_tail) adds the partials, the bias and the ReLU.  Returns the number of splits used.
int fc_fwd_tc_splits(int Bt, int K, int N) {
  int tiles = cdiv(N, 64) * cdiv(Bt, BM);
  int s = std::max(1, std::min(K / (3 * BK), (kNumSMs + tiles - 1) / tiles));
  int kps = (int)round_up(cdiv(K, s), BK);
  return cdiv(K, kps);
}
int fc_fwd_tc_splitk(const float* x, const float* w, float* part, int Bt, int K, int N, cudaStream_t st) {
  int s = fc_fwd_tc_splits(Bt, K, N);
  int kps = (int)round_up(cdiv(K, s), BK);
  GemmArgs g{Bt, N, K, x, K, w, N, part, N, nullptr, 0, nullptr, 0, nullptr, kps, (int64_t)Bt * N, TR_FC1_FWD};
  int rc = check_align(g, false, true);
  if (rc) return rc;
  return dispatch<false, true>(g, st);
}

// dw[K,N] = x^T[K,Bt] dy[Bt,N]: A(m=k_feat, k=batch) = x[batch*K + m] MN-major; B(k=batch, n) = dy MN-major
int fc_dw_tc(const float* x, const float* dy, float* dw, int Bt, int K, int N, cudaStream_t st) {
  GemmArgs g{K, N, Bt, x, K, dy, N, dw, N, nullptr, 0, nullptr, 0, nullptr, (int)round_up(Bt, BK), 0, TR_FC1_DW};
  int rc = check_align(g, true, true);
  if (rc) return rc;
  return dispatch<true, true>(g, st);
}
// dx[Bt,K] = (dy[Bt,N] w^T) * [x > 0]?: A = dy K-major; B(k=n_out, n=k_feat) = w[n*N + k] K-major
int fc_dx_tc(const float* x, const float* w, const float* dy, float* dx, int Bt, int K, int N, int mask_x,
             cudaStream_t st) {
  GemmArgs g{Bt, K, N, dy, N, w, N, dx, K, nullptr, 0, mask_x ? x : nullptr, K, nullptr, (int)round_up(N, BK), 0, TR_FC1_DX};
  int rc = check_align(g, false, false);
  if (rc) return rc;
  return dispatch<false, false>(g, st);
}
// both gradients of one FC layer (bias gradient: colsum_f32 / fc_tail)
int fc_bwd_tc(const float* x, const float* w, const float* dy, float* dx, float* dw, int Bt, int K, int N,
              int mask_x, cudaStream_t st) {
  int rc = fc_dw_tc(x, dy, dw, Bt, K, N, st);
  if (rc || !dx) return rc;
  return fc_dx_tc(x, w, dy, dx, Bt, K, N, mask_x, st);
}

// ---- operand images -------------------------------------------------------------------------------------------
// w [K][N] fp32 -> bf16 image [N/8][K][8]: serves the forward product (B MN-major: chunk = n/8, rows = k) and dX
// (B K-major with k = n_out: chunk = n_out/8, rows = k_feat) alike.  Rebuilt once per step on a side stream.
namespace {
__global__ void pack_fc_weights_kernel(const float* __restrict__ w, uint4* __restrict__ img, int K, int N) {
  pdl_launch_dependents();
  const int ng = N / 8;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < K * ng; idx += gridDim.x * blockDim.x) {
    const int g8 = idx % ng, k = idx / ng;                       // neighbouring threads read neighbouring 32 bytes
    img[(size_t)g8 * K + k] = load_chunk(w + (size_t)k * N + g8 * 8);
  }
}
int img_ok(const void* a, const void* b) { return a && b && ((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0; }
}  // namespace
int64_t fc_weight_image_bytes(int K, int N) { return (int64_t)K * N * 2; }
int pack_fc_weights(const float* w, void* img, int K, int N, cudaStream_t st) {
  B200_REQUIRE(K > 0 && N > 0 && N % 8 == 0 && ((uintptr_t)w & 15) == 0 && ((uintptr_t)img & 15) == 0,
               "pack_fc_weights: bad shape / alignment");
  B200_CUDA_TRY(launch_kernel(pack_fc_weights_kernel, dim3(std::min(cdiv((int64_t)K * (N / 8), 256), 4 * kNumSMs)), dim3(256), 0,
                              st, w, (uint4*)img, K, N));
  return B200_OK;
}
bool fc_images_supported(int Bt, int K, int N) { return Bt % 64 == 0 && K % 128 == 0 && N % 64 == 0; }
// the three FC1 products on operand images: x_img [K/8][Bt][8], w_img [N/8][K][8], dy_img [N/8][Bt][8] (all bf16)
int fc_fwd_tc_splitk_img(const void* x_img, const void* w_img, float* part, int Bt, int K, int N, cudaStream_t st) {
  B200_REQUIRE(img_ok(x_img, w_img) && fc_images_supported(Bt, K, N), "fc_fwd_tc_splitk_img: bad operands");
  int s = fc_fwd_tc_splits(Bt, K, N);
  int kps = (int)round_up(cdiv(K, s), BK);
  GemmArgs g{Bt, N, K, nullptr, K, nullptr, N, part, N, nullptr, 0, nullptr, 0, nullptr, kps, (int64_t)Bt * N, TR_FC1_FWD,
             (const uint8_t*)x_img, Bt, (const uint8_t*)w_img, K};
  return dispatch<false, true>(g, st);
}
int fc_dw_tc_img(const void* x_img, const void* dy_img, float* dw, int Bt, int K, int N, cudaStream_t st) {
  B200_REQUIRE(img_ok(x_img, dy_img) && fc_images_supported(Bt, K, N), "fc_dw_tc_img: bad operands");
  GemmArgs g{K, N, Bt, nullptr, K, nullptr, N, dw, N, nullptr, 0, nullptr, 0, nullptr, (int)round_up(Bt, BK), 0, TR_FC1_DW,
             (const uint8_t*)x_img, Bt, (const uint8_t*)dy_img, Bt};
  return dispatch<true, true>(g, st);
}
int fc_dx_tc_img(const void* dy_img, const void* w_img, float* dx, int Bt, int K, int N, cudaStream_t st) {
  B200_REQUIRE(img_ok(dy_img, w_img) && fc_images_supported(Bt, K, N), "fc_dx_tc_img: bad operands");
  GemmArgs g{Bt, K, N, nullptr, N, nullptr, N, dx, K, nullptr, 0, nullptr, K, nullptr, (int)round_up(N, BK), 0, TR_FC1_DX,
             (const uint8_t*)dy_img, Bt, (const uint8_t*)w_img, K};
  return dispatch<false, false>(g, st);
}

void gemm_set_timeline(long long* dev_buf) { g_gemm_timeline = dev_buf; }
B200_TRACE_SETTER(trace_set_tc_gemm)

}  // namespace b200
