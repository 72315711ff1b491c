// microbenchmark (development only): throughput of fp32 reductions to global memory - scalar red, red.v2, red.v4 -
// in the access pattern of a dQ drain: a thread owns a row and adds 128 consecutive floats, rows 16 KB apart.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_bench red_bench.cu && ./red_bench
#include <cstdio>
#include <cuda_runtime.h>

template <int V>
__global__ void __launch_bounds__(128) red_kernel(float* acc, int64_t ld, int tiles_per_cta, int n_tiles) {
  // CTA handles tiles (64 rows x 128 cols) round-robin; thread t < 64 owns row t of the tile... use 128 threads: 2 per row (64 cols each)
  const int row = threadIdx.x >> 1, c0 = (threadIdx.x & 1) * 64;
  for (int i = 0; i < tiles_per_cta; ++i) {
    const int tile = (blockIdx.x + i * gridDim.x) % n_tiles;
    float* p = acc + ((int64_t)(tile / 32) * 64 + row) * ld + (tile % 32) * 128 + c0;
    const float v = (float)i;
    if (V == 1) {
#pragma unroll
      for (int c = 0; c < 64; ++c) asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p + c), "f"(v) : "memory");
    } else if (V == 2) {
#pragma unroll
      for (int c = 0; c < 64; c += 2) asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(p + c), "f"(v), "f"(v) : "memory");
    } else {
#pragma unroll
      for (int c = 0; c < 64; c += 4)
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p + c), "f"(v), "f"(v), "f"(v), "f"(v) : "memory");
    }
  }
}

template <int V>
void run(float* acc, int64_t ld, int n_tiles) {
  const int ctas = 148, per = 128;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  red_kernel<V><<<ctas, 128>>>(acc, ld, 8, n_tiles);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  red_kernel<V><<<ctas, 128>>>(acc, ld, per, n_tiles);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double floats = (double)ctas * per * 64 * 128;
  printf("V=%d: %.3f ms, %.1f Gfloat/s, %.2f floats/clk/SM (1.965 GHz), per 64x128 tile %.0f clk, err=%s\n", V, ms, floats / ms / 1e6,
         floats / (ms * 1e-3) / 148 / 1.965e9, (ms * 1e-3) * 1.965e9 / per, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  const int64_t ld = 4096;
  const int n_tiles = 128 * 32;  // 8192 rows x 4096 cols fp32 = 134 MB (the C3 dq accumulator)
  float* acc;
  cudaMalloc(&acc, (size_t)8192 * ld * 4);
  cudaMemset(acc, 0, (size_t)8192 * ld * 4);
  run<1>(acc, ld, n_tiles);
  run<2>(acc, ld, n_tiles);
  run<4>(acc, ld, n_tiles);
  run<4>(acc, ld, n_tiles);
  return 0;
}
