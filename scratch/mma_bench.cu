// microbenchmark: tcgen05.mma issue/throughput for no-swizzle operand layouts with various strides
#include <cstdio>
#include <cuda_runtime.h>
#include "../distributed-machine-learning-using-cnn-cifar-10-dataset-_b200/csrc/tc_common.cuh"
using namespace b200::tc;

struct Cfg { int M, N, a_mn, b_mn, a_lbo, a_sbo, b_lbo, b_sbo, a_step, dep, commit_every, spinners; };

__global__ void __launch_bounds__(160, 1) k(Cfg c, int iters, long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint64_t bars2[8];
  __shared__ uint32_t slot;
  int tid = threadIdx.x, warp = tid >> 5;
  for (int o = tid * 16; o < 200 * 1024; o += 160 * 16) *reinterpret_cast<uint4*>(smem + o) = make_uint4(0, 0, 0, 0);
  if (tid == 0) { mbar_init(smem_u32(&bar), 1); for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&bars2[i]), 1000000); fence_mbar_init(); }
  if (warp == 4) tmem_alloc(smem_u32(&slot), 512);
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  uint32_t tmem = slot;
  if (tid == 128) {
    uint32_t idesc = make_idesc(c.M, c.N, c.a_mn, c.b_mn);
    uint32_t sa = smem_u32(smem), sb = sa + 120 * 1024;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      uint64_t da = make_desc(sa + (i & 3) * c.a_step, c.a_lbo, c.a_sbo);
      uint64_t db = make_desc(sb, c.b_lbo, c.b_sbo);
      umma_bf16(tmem + (c.dep ? 0 : (i & 3) * c.N), da, db, idesc, 1);
      if (c.commit_every && (i % c.commit_every) == c.commit_every - 1) umma_commit(smem_u32(&bars2[(i / c.commit_every) & 7]));
    }
    umma_commit(smem_u32(&bar));
    long long t1 = clock64();
    mbar_wait(smem_u32(&bar), 0);
    long long t2 = clock64();
    out[0] = t1 - t0; out[1] = t2 - t0;
  }
  if (c.spinners && tid < 128) mbar_wait(smem_u32(&bar), 0);
  __syncthreads();
  if (warp == 4) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main() {
  long long* out; cudaMallocManaged(&out, 16);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  Cfg cfgs[] = {
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 0, 0, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 0, 1, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 0, 2, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 0, 4, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 0, 8, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 0, 16, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 0, 32, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 0, 64, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 1, 0, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 1, 1, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 1, 2, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 1, 4, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 1, 8, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 1, 16, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 1, 32, 0},
    {128, 64, 0, 0, 5200, 128, 1040, 128, 16, 1, 64, 0},
    {128, 128, 0, 0, 2048, 128, 4096, 128, 0, 1, 0, 0},
    {128, 128, 0, 0, 2048, 128, 4096, 128, 0, 1, 4, 0},
    {128, 128, 0, 0, 2048, 128, 4096, 128, 0, 1, 8, 0},
    {128, 256, 0, 0, 2048, 128, 4096, 128, 0, 1, 0, 0},
    {128, 256, 0, 0, 2048, 128, 4096, 128, 0, 1, 4, 0},
    {128, 256, 0, 0, 2048, 128, 4096, 128, 0, 1, 8, 0},
  };


  const int iters = 2000;
  for (auto& c : cfgs) {
    k<<<1, 160, 200 * 1024>>>(c, iters, out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    printf("N=%d dep=%d commit_every=%d spinners=%d : issue %.1f cyc/mma, complete %.1f cyc/mma\n", c.N, c.dep, c.commit_every, c.spinners, out[0] / (double)iters, out[1] / (double)iters);
  }
  return 0;
}
