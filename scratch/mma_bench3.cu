// microbenchmark 3: several issuing warps, each feeding its own accumulator tile
#include <cstdio>
#include <cuda_runtime.h>
#include "../distributed-machine-learning-using-cnn-cifar-10-dataset-_b200/csrc/tc_common.cuh"
using namespace b200::tc;

__global__ void __launch_bounds__(320, 1) k(int issuers, int per_issuer, long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bars[8];
  __shared__ uint32_t slot;
  int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int o = tid * 16; o < 200 * 1024; o += 320 * 16) *reinterpret_cast<uint4*>(smem + o) = make_uint4(0, 0, 0, 0);
  if (tid == 0) { for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&bars[i]), 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&slot), 512);
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  uint32_t tmem = slot;
  const int ch_stride = 5200;
  if (warp >= 2 && warp < 2 + issuers && lane == 0) {
    const int w = warp - 2;
    const uint32_t idesc = make_idesc(128, 64, false, false);
    const uint32_t a_hi = desc_hi(128), b_hi = desc_hi(128);
    const uint32_t a_lo0 = desc_lo(smem_u32(smem) + 34 * 16 + w * 2048, ch_stride);
    const uint32_t b_lo0 = desc_lo(smem_u32(smem) + 48 * 1024, 1040);
    const uint32_t a_kstep = 2 * (ch_stride >> 4), b_kstep = 2 * (1040 >> 4);
    long long t0 = clock64();
    uint32_t acc = 0;
    for (int i = 0; i < per_issuer / 4; ++i) {
      uint32_t a_lo = a_lo0 + (i % 25), b_lo = b_lo0 + (i % 6) * 520;
#pragma unroll 4
      for (int ks = 0; ks < 4; ++ks) {
        umma_bf16(tmem + w * 64, desc_join(a_lo, a_hi), desc_join(b_lo, b_hi), idesc, acc | (uint32_t)ks);
        a_lo += a_kstep; b_lo += b_kstep;
      }
      acc = 1;
    }
    umma_commit(smem_u32(&bars[w]));
    mbar_wait(smem_u32(&bars[w]), 0);
    long long t2 = clock64();
    out[w] = t2 - t0;
  }
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main() {
  long long* out; cudaMallocManaged(&out, 64);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int issuers : {1, 2, 3, 4, 6}) {
    const int per = 400;
    k<<<1, 320, 200 * 1024>>>(issuers, per, out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    long long mx = 0; for (int i = 0; i < issuers; ++i) mx = out[i] > mx ? out[i] : mx;
    printf("issuers=%d: %lld cycles for %d MMAs -> %.1f cycles per MMA overall\n", issuers, mx, issuers * per, mx / (double)(issuers * per));
  }
  return 0;
}
