// throughput / latency of mma.sync.m16n8k16 bf16 on sm_100a: W warps per CTA, C independent accumulator chains
#include <cstdio>
#include <cuda_runtime.h>
template <int C>
__global__ void k(float* out, long long* cyc, int iters) {
  float c[C][4];
  for (int i = 0; i < C; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
  unsigned a0 = threadIdx.x, a1 = 1, a2 = 2, a3 = 3, b0 = 4, b1 = 5;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < C; ++i)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  __syncthreads();
  long long t1 = clock64();
  float s = 0; for (int i = 0; i < C; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 8);
  const int iters = 1000;
  for (int warps : {1, 4, 8, 16}) {
    long long h;
#define RUN(C) k<C><<<1, warps * 32>>>(out, cyc, iters); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); \
    printf("warps %2d chains %d: %.1f cycles per HMMA per warp, %.2f cycles per HMMA per SM\n", warps, C, (double)h / (iters * C), (double)h / (iters * C * warps));
    RUN(1) RUN(3) RUN(6)
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
