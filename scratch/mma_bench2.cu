// microbenchmark 2: the per-tap issue pattern of the conv kernel, feature by feature
#include <cstdio>
#include <cuda_runtime.h>
#include "../distributed-machine-learning-using-cnn-cifar-10-dataset-_b200/csrc/tc_common.cuh"
using namespace b200::tc;

struct Cfg { int commit, wait, fence, bmn, elect, tiles, mul_k, mul_tap, mul_t, lead; };

__global__ void __launch_bounds__(192, 1) k(Cfg c, long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bars[16];
  __shared__ uint32_t slot;
  int tid = threadIdx.x, lane = tid & 31; const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  for (int o = tid * 16; o < 200 * 1024; o += 192 * 16) *reinterpret_cast<uint4*>(smem + o) = make_uint4(0, 0, 0, 0);
  if (tid == 0) { for (int i = 0; i < 16; ++i) mbar_init(smem_u32(&bars[i]), i < 8 ? 1 : 1000000); fence_mbar_init(); }
  if (warp == 4) tmem_alloc(smem_u32(&slot), 512);
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  if (tid == 0) for (int i = 0; i < 7; ++i) mbar_arrive(smem_u32(&bars[i]));   // bars 0..6: phase 0 complete
  __syncthreads();
  uint32_t tmem = slot;
  const int P = 16, lead = c.lead, ch_stride = 5200, wstage = 8320;
  if (warp == 4 && (c.elect || lane == 0)) {
    const uint32_t tmem_u = c.elect ? __shfl_sync(0xffffffffu, tmem, 0) : tmem;
    const uint32_t idesc = make_idesc(128, 64, false, c.bmn);
    const uint32_t a_hi = desc_hi(128), b_hi = c.bmn ? desc_hi(1040) : desc_hi(128);
    const uint32_t a_lo0 = desc_lo(smem_u32(smem) + lead * 16, ch_stride);
    const uint32_t b_lo0 = c.bmn ? desc_lo(smem_u32(smem) + 48 * 1024, 128) : desc_lo(smem_u32(smem) + 48 * 1024, 1040);
    const uint32_t a_kstep = c.mul_k * 2 * (ch_stride >> 4), b_kstep = c.bmn ? 16 : 2 * (1040 >> 4);
    long long t0 = clock64();
    uint32_t acc = 0;
    int di = 0, dj = 0;
    for (int rep = 0; rep < 4; ++rep)
    for (int i = 0; i < 25; ++i) {
      const int s = i % 6;
      if (c.wait) mbar_wait(smem_u32(&bars[s]), 0);
      if (c.fence) tc_fence_after();
      const uint32_t a_tap = a_lo0 + (uint32_t)(c.mul_tap * ((di - 2) * P + (dj - 2)));
      const uint32_t b_tap = b_lo0 + (uint32_t)s * (wstage >> 4);
      for (int t = 0; t < c.tiles; ++t) {
        uint32_t a_lo = a_tap + c.mul_t * t * 128, b_lo = b_tap;
        const uint32_t d = tmem_u + t * 64;
#pragma unroll 4
        for (int ks = 0; ks < 4; ++ks) {
          if (c.elect) { if (elect_one()) umma_bf16(d, desc_join(a_lo, a_hi), desc_join(b_lo, b_hi), idesc, acc | (uint32_t)ks); }
          else umma_bf16(d, desc_join(a_lo, a_hi), desc_join(b_lo, b_hi), idesc, acc | (uint32_t)ks);
          a_lo += a_kstep; b_lo += b_kstep;
        }
      }
      acc = 1;
      if (c.commit) { if (c.elect) { if (elect_one()) umma_commit(smem_u32(&bars[8 + s])); } else umma_commit(smem_u32(&bars[8 + s])); }
      if (++dj == 5) { dj = 0; if (++di == 5) di = 0; }
    }
    if (!c.elect || elect_one()) umma_commit(smem_u32(&bars[7]));
    long long t1 = clock64();
    mbar_wait(smem_u32(&bars[7]), 0);
    long long t2 = clock64();
    if (lane == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  __syncthreads();
  if (warp == 4) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main() {
  long long* out; cudaMallocManaged(&out, 16);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  Cfg cfgs[] = {
    // commit wait fence bmn elect tiles mul_k mul_tap mul_t lead
    {0, 0, 0, 0, 0, 2, 1, 1, 1, 34},
    {0, 0, 0, 0, 1, 2, 1, 1, 1, 34},
    {1, 1, 1, 0, 1, 2, 1, 1, 1, 34},
    {1, 1, 1, 1, 1, 7, 1, 1, 1, 34},
  };


  for (auto& c : cfgs) {
    k<<<1, 192, 200 * 1024>>>(c, out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    int n = 4 * 25 * c.tiles * 4;
    printf("mul_k=%d mul_tap=%d mul_t=%d lead=%d | commit=%d wait=%d fence=%d bmn=%d elect=%d tiles=%d : issue %.1f cyc/mma, complete %.1f cyc/mma (%.0f per tap)\n", c.mul_k, c.mul_tap, c.mul_t, c.lead, c.commit, c.wait, c.fence,
           c.bmn, c.elect, c.tiles, out[0] / (double)n, out[1] / (double)n, out[1] / 100.0);
  }
  return 0;
}
