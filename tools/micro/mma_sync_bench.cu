// microbenchmark: legacy mma.sync m16n8k16 bf16 issue rate on sm_100a (how fast is the HMMA path really?)
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
template <int ACC>
__global__ void k(float* out, int iters, uint32_t seed) {
  float d[ACC][4];
  for (int j = 0; j < ACC; ++j) d[j][0] = d[j][1] = d[j][2] = d[j][3] = 0.f;
  uint32_t a[4] = {seed, seed + 1, seed + 2, seed + 3}, b0 = seed * 3, b1 = seed * 5;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < ACC; ++j)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                   : "+f"(d[j][0]), "+f"(d[j][1]), "+f"(d[j][2]), "+f"(d[j][3])
                   : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  float s = 0;
  for (int j = 0; j < ACC; ++j) s += d[j][0] + d[j][1] + d[j][2] + d[j][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  float* out;
  cudaMalloc(&out, 148 * 8 * 1024 * 4);
  for (int warps = 4; warps <= 16; warps *= 2) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 4096;
    k<8><<<148, warps * 32>>>(out, iters, 1);
    cudaEventRecord(e0);
    k<8><<<148, warps * 32>>>(out, iters, 1);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double mmas = 148.0 * warps * iters * 8;
    printf("warps/SM %d: %.3f ms, %.1f TFLOP/s, %.1f cycles@1.9GHz per mma per SMSP\n", warps, ms, mmas * 4096 / ms / 1e9,
           ms * 1e-3 * 1.9e9 / (mmas / (148.0 * 4)));
  }
  return 0;
}
