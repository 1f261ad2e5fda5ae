// Probe: what an [event, kernel, event] bracket costs for an (almost) empty kernel, as a function of parameter bytes,
// dynamic shared memory and TMEM allocation -- the fixed per-launch overhead the attention launches carry in bench.py.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/launch_overhead tools/probe/launch_overhead.cu && /tmp/launch_overhead
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

template <int kBytes>
struct Blob { unsigned char b[kBytes]; };

template <int kBytes>
__global__ void k_params(const __grid_constant__ Blob<kBytes> blob, int* sink) {
  extern __shared__ unsigned char sm[];
  if (threadIdx.x == 0 && blob.b[threadIdx.x % kBytes] == 123 && sink) sink[0] = sm[0];
}

__global__ void k_fill(float* p, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = 1.f;
}

template <typename F>
static float bracket(F launch, cudaStream_t st, float* big, size_t nbig, bool pre) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  std::vector<float> t;
  for (int i = 0; i < 30; ++i) {
    if (pre) k_fill<<<148 * 4, 256, 0, st>>>(big, nbig);   // a "previous kernel" with a different smem configuration
    cudaEventRecord(a, st);
    launch();
    cudaEventRecord(b, st);
    cudaStreamSynchronize(st);
    float ms; cudaEventElapsedTime(&ms, a, b);
    t.push_back(ms * 1e3f);
  }
  std::sort(t.begin(), t.end());
  return t[t.size() / 2];
}

int main() {
  cudaStream_t st; cudaStreamCreate(&st);
  float* big; size_t nbig = 8 << 20; cudaMalloc(&big, nbig * 4);
  int* sink; cudaMalloc(&sink, 4);
  const int smem_big = 211 * 1024;
  cudaFuncSetAttribute(k_params<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_big);
  cudaFuncSetAttribute(k_params<3584>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_big);
  cudaFuncSetAttribute(k_params<6400>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_big);
  Blob<64> b64{}; Blob<3584> b3k{}; Blob<6400> b6k{};
  for (int pre = 0; pre < 2; ++pre) {
    printf("previous kernel in the stream: %s\n", pre ? "a fill kernel (8 Mi floats)" : "none (idle GPU)");
    for (int smem : {0, smem_big}) {
      for (int grid : {148}) {
        float t0 = bracket([&] { k_params<64><<<grid, 384, smem, st>>>(b64, sink); }, st, big, nbig, pre);
        float t1 = bracket([&] { k_params<3584><<<grid, 384, smem, st>>>(b3k, sink); }, st, big, nbig, pre);
        float t2 = bracket([&] { k_params<6400><<<grid, 384, smem, st>>>(b6k, sink); }, st, big, nbig, pre);
        printf("  grid %d, 384 thr, dyn smem %6d B: bracket median  64 B params %.2f us | 3584 B %.2f us | 6400 B %.2f us\n", grid, smem, t0, t1, t2);
      }
    }
    float tm = bracket([&] { cudaMemsetAsync(sink, 0, 4, st); }, st, big, nbig, pre);
    float te = bracket([&] {}, st, big, nbig, pre);
    printf("  memset(4 B) alone %.2f us | empty bracket %.2f us\n", tm, te);
    float tmk = bracket([&] { cudaMemsetAsync(big, 0, 20000, st); k_params<6400><<<148, 384, smem_big, st>>>(b6k, sink); }, st, big, nbig, pre);
    printf("  memset(20 KB) + kernel(6400 B params, 211 KB smem) %.2f us\n", tmk);
  }
  return 0;
}
