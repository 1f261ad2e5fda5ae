// Event-bracketed duration of (nearly) empty kernels on the current GPU: what a CUDA-event roofline pays per launch
// beyond the kernel's own execution.  Variants: parameter bytes (64 B vs 3.6 KB __grid_constant__), dynamic shared
// memory (0 vs 206 KB), block size (128 vs 640), grid (1 vs 144), and what precedes the launch (memset kernel or same).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/launch_overhead tools/micro/launch_overhead.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <algorithm>
#include <vector>

struct Big { unsigned char b[3584]; };
__global__ void k_small(int* p) { if (p && threadIdx.x == 9999) *p = 1; }
__global__ void k_big(const __grid_constant__ Big b, int* p) { if (p && threadIdx.x == 9999) *p = b.b[5]; }
extern __shared__ unsigned char dyn[];
__global__ void k_smem(int* p) { if (p && threadIdx.x == 9999) *p = dyn[3]; }
__global__ void k_smem_big(const __grid_constant__ Big b, int* p) { if (p && threadIdx.x == 9999) *p = dyn[3] + b.b[5]; }

template <typename F>
static float bracket(F launch, bool flush, void* fbuf, size_t fbytes) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  std::vector<float> t;
  for (int i = 0; i < 60; ++i) {
    if (flush) cudaMemsetAsync(fbuf, 0, fbytes);
    cudaEventRecord(a);
    launch();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    if (i >= 10) t.push_back(ms * 1e3f);
  }
  std::sort(t.begin(), t.end());
  return t[t.size() / 2];
}

int main() {
  Big big = {};
  void* fbuf; size_t fbytes = 256u << 20;
  cudaMalloc(&fbuf, fbytes);
  const int smem = 206 * 1024;
  cudaFuncSetAttribute(k_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(k_smem_big, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int flush = 0; flush < 2; ++flush) {
    printf("preceded by %s\n", flush ? "a 256 MB memset (cold L2, as the microbench)" : "nothing (idle stream)");
    printf("  events only                        %6.2f us\n", bracket([] {}, flush, fbuf, fbytes));
    printf("  1x128,   64 B params, 0 smem       %6.2f us\n", bracket([] { k_small<<<1, 128>>>(nullptr); }, flush, fbuf, fbytes));
    printf("  144x640, 64 B params, 0 smem       %6.2f us\n", bracket([] { k_small<<<144, 640>>>(nullptr); }, flush, fbuf, fbytes));
    printf("  144x640, 3.6 KB params, 0 smem     %6.2f us\n", bracket([&] { k_big<<<144, 640>>>(big, nullptr); }, flush, fbuf, fbytes));
    printf("  144x640, 64 B params, 206 KB smem  %6.2f us\n", bracket([&] { k_smem<<<144, 640, smem>>>(nullptr); }, flush, fbuf, fbytes));
    printf("  144x640, 3.6 KB params, 206 KB     %6.2f us\n", bracket([&] { k_smem_big<<<144, 640, smem>>>(big, nullptr); }, flush, fbuf, fbytes));
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
