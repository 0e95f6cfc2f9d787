// Aggregate throughput of fire-and-forget global reductions (RED) on B200: the primitive the
// per-run flush of the voxel pass (voxel_runs.cu) is built from.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o red_tput red_tput.cu ; run: ./red_tput
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(void* buf, size_t nwords, int iters, int lanes) {
  uint32_t x = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  const int lane = threadIdx.x & 31;
  if (lane >= lanes) return;
  for (int it = 0; it < iters; ++it) {
    x = x * 1664525u + 1013904223u;
    size_t i = ((size_t)(x >> 4) * 2654435761ull >> 7) % nwords;
    if (MODE == 0) atomicAdd((uint32_t*)buf + i, 1u);
    if (MODE == 1) atomicAdd((unsigned long long*)buf + i, 1ull);
    if (MODE == 2) atomicAdd((double*)buf + i, 1.0);
    if (MODE == 3) atomicMax((uint32_t*)buf + i, x);
    if (MODE == 4) {  // record-like: 8 words of one 32-byte sector chosen at random
      size_t r = (i & ~(size_t)7);
      atomicAdd((uint32_t*)buf + r + (it & 7), 1u);
    }
    if (MODE == 5) {  // neighbouring lanes share an address (groups of 8)
      size_t j = ((size_t)((x >> 4) & ~7u) * 2654435761ull >> 7) % nwords;
      uint32_t y = __shfl_sync(__activemask(), (uint32_t)j, lane & ~7);
      atomicAdd((double*)buf + y, 1.0);
    }
  }
}

template <int MODE>
void run(const char* name, size_t bytes, int lanes, int wordBytes) {
  void* buf; cudaMalloc(&buf, bytes); cudaMemset(buf, 0, bytes);
  size_t nwords = bytes / wordBytes;
  int iters = 512, blocks = 148 * 16, threads = 256;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k<MODE><<<blocks, threads>>>(buf, nwords, 16, lanes);
  cudaEventRecord(a);
  k<MODE><<<blocks, threads>>>(buf, nwords, iters, lanes);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double ops = (double)blocks * threads / 32 * lanes * iters;
  printf("%-34s table %7.0f MB lanes %2d : %8.1f G ops/s  (%.3f ms)\n", name, bytes / 1048576.0, lanes, ops / ms / 1e6, ms);
  cudaFree(buf);
}

int main() {
  for (size_t mb : {16, 2048}) {
    for (int lanes : {32, 4}) {
      run<0>("RED.ADD.u32 random", mb << 20, lanes, 4);
      run<1>("RED.ADD.u64 random", mb << 20, lanes, 8);
      run<2>("RED.ADD.f64 random", mb << 20, lanes, 8);
      run<3>("RED.MAX.u32 random", mb << 20, lanes, 4);
      run<4>("RED.ADD.u32 record sector", mb << 20, lanes, 4);
      run<5>("RED.ADD.f64 groups of 8 lanes", mb << 20, lanes, 8);
    }
    printf("\n");
  }
  return 0;
}
