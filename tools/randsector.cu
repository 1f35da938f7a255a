// randsector.cu — measured ceiling for the access pattern of chain_step_kernel: independent 4-byte loads that
// each touch a different random 32-byte sector of a buffer far larger than L2 (the chains' orientation words
// and container entries).  Prints one JSON line: sectors/s and the equivalent GB/s at 32 B per sector, for a
// range of loads in flight per thread, next to a streaming copy of the same buffer.
// Measurement tool only (bench.py reads its output as the "random-sector" roofline denominator).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/randsector tools/randsector.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t mix(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}

// every thread: `iters` rounds of MLP independent loads, one random sector each
template <int MLP>
__global__ void rand_kernel(const uint32_t* __restrict__ buf, uint64_t nsectors, int iters, uint32_t* out) {
  const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t acc = 0;
  uint64_t s = mix(tid + 1);
  for (int it = 0; it < iters; it++) {
    uint32_t v[MLP];
#pragma unroll
    for (int k = 0; k < MLP; k++) {
      s = s * 6364136223846793005ULL + 1442695040888963407ULL;
      const uint64_t sec = __umul64hi(mix(s), nsectors);
      v[k] = __ldcg(&buf[sec * 8 + (s & 7)]);
    }
#pragma unroll
    for (int k = 0; k < MLP; k++) acc += v[k];
  }
  if (acc == 0x12345678u) out[0] = acc;
}

// dependent chain: one load in flight per thread (latency-bound pointer chase through random sectors)
__global__ void chase_kernel(const uint32_t* __restrict__ buf, uint64_t nsectors, int iters, uint32_t* out) {
  const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t s = mix(tid + 1);
  uint32_t acc = 0;
  for (int it = 0; it < iters; it++) {
    const uint64_t sec = __umul64hi(mix(s + acc), nsectors);
    acc = __ldcg(&buf[sec * 8]);
    s += 0x9E3779B97F4A7C15ULL;
  }
  if (acc == 0x12345678u) out[0] = acc;
}

__global__ void copy_kernel(const uint4* __restrict__ a, uint4* __restrict__ b, uint64_t n) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) b[i] = a[i];
}

template <int MLP>
static double run_rand(const uint32_t* buf, uint64_t nsectors, int blocks, int threads, int iters, uint32_t* out) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  rand_kernel<MLP><<<blocks, threads>>>(buf, nsectors, iters / 4, out);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  rand_kernel<MLP><<<blocks, threads>>>(buf, nsectors, iters, out);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  return (double)blocks * threads * iters * MLP / (ms * 1e-3);
}

int main(int argc, char** argv) {
  const double gb = argc > 1 ? atof(argv[1]) : 16.0;
  if (argc > 2) CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(argv[2])));   // L2 fetch granularity hint (32 / 64 / 128)
  size_t fetch = 0;
  CK(cudaDeviceGetLimit(&fetch, cudaLimitMaxL2FetchGranularity));
  const uint64_t bytes = (uint64_t)(gb * (1ULL << 30)) & ~255ULL;
  const uint64_t nsectors = bytes / 32;
  uint32_t *buf, *out;
  uint4* dst;
  CK(cudaMalloc(&buf, bytes));
  CK(cudaMalloc(&dst, 1ULL << 30));
  CK(cudaMalloc(&out, 64));
  CK(cudaMemset(buf, 1, bytes));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  const int sm = prop.multiProcessorCount;
  printf("{\"gpu\": \"%s\", \"buffer_gb\": %.1f, \"l2_fetch_granularity\": %zu, \"rows\": [", prop.name, gb, fetch);
  bool first = true;
  // warps per SM x loads in flight per thread
  const int wps[] = {8, 16, 28, 32, 64};
  for (int wi = 0; wi < 5; wi++) {
    const int threads = 256, blocks = sm * wps[wi] * 32 / threads;
    double r[4];
    r[0] = run_rand<1>(buf, nsectors, blocks, threads, 2048, out);
    r[1] = run_rand<2>(buf, nsectors, blocks, threads, 1024, out);
    r[2] = run_rand<4>(buf, nsectors, blocks, threads, 512, out);
    r[3] = run_rand<8>(buf, nsectors, blocks, threads, 256, out);
    const int mlp[] = {1, 2, 4, 8};
    for (int k = 0; k < 4; k++) {
      printf("%s{\"warps_per_sm\": %d, \"loads_in_flight_per_thread\": %d, \"gsectors_per_s\": %.3f, \"gb_per_s_32B\": %.1f}", first ? "" : ", ",
             wps[wi], mlp[k], r[k] / 1e9, r[k] * 32 / 1e9);
      first = false;
    }
  }
  printf("], ");
  // dependent chase: latency of one random DRAM access under 28 warps/SM of the same
  {
    const int threads = 128, blocks = sm * 7;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    chase_kernel<<<blocks, threads>>>(buf, nsectors, 256, out);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    chase_kernel<<<blocks, threads>>>(buf, nsectors, 4096, out);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("\"chase_ns_per_load_28warps\": %.1f, ", ms * 1e6 / 4096);
    chase_kernel<<<1, 32>>>(buf, nsectors, 256, out);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    chase_kernel<<<1, 32>>>(buf, nsectors, 4096, out);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("\"chase_ns_per_load_1warp\": %.1f, ", ms * 1e6 / 4096);
  }
  {
    const uint64_t n16 = (1ULL << 30) / 16;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    copy_kernel<<<sm * 16, 256>>>((const uint4*)buf, dst, n16);
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
      CK(cudaEventRecord(e0));
      copy_kernel<<<sm * 16, 256>>>((const uint4*)buf + (uint64_t)(rep + 1) * n16, dst, n16);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (ms < best) best = ms;
    }
    printf("\"copy_gb_per_s\": %.1f}\n", 2.0 * (1ULL << 30) / (best * 1e-3) / 1e9);
  }
  return 0;
}
