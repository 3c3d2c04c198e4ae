// Microbenchmark: issue rate of cp.(reduce.)async.bulk shared->global per SM, as a function of the number of issuing
// warps and the copy size.  Developer tool (informs the Corr epilogue design); not part of the library.
// NB the GB/s it prints is cycles counted inside the kernel x the nominal clock; for the chip-level rate by the wall
// clock see bulk_pattern.cu (2.6 TB/s of reduction payload).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <bool RED>
__global__ void __launch_bounds__(512) k(int32_t* out, int nwarps, int bytes, int ops_per_warp, size_t cta_stride,
                                         long long* cycles) {
  extern __shared__ __align__(128) uint8_t sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) reinterpret_cast<int*>(sm)[i] = 1;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  long long t0 = clock64();
  if (warp < nwarps && lane == 0) {
    int32_t* base = out + blockIdx.x * (cta_stride / 4) + (size_t)warp * ops_per_warp * (bytes / 4);
    const uint32_t src = smem_u32(sm) + (warp & 7) * 2048;
    for (int i = 0; i < ops_per_warp; ++i) {
      int32_t* dst = base + (size_t)i * (bytes / 4);
      if (RED)
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.s32 [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
      else
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
      if ((i & 15) == 15) {
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
      }
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

int main() {
  const int nsm = 148;
  const size_t cta_stride = 8u << 20;   // 8 MB of output per CTA
  int32_t* out;
  long long* cyc;
  cudaMalloc(&out, cta_stride * nsm);
  cudaMemset(out, 0, cta_stride * nsm);
  cudaMalloc(&cyc, nsm * sizeof(long long));
  cudaFuncSetAttribute(k<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  long long h[nsm];
  for (int red = 0; red < 2; ++red)
    for (int bytes : {256, 512, 1024, 2048})
      for (int nw : {1, 2, 4, 8, 16}) {
        const int total_ops = (4 << 20) / bytes;            // 4 MB per CTA in total
        const int opw = total_ops / nw;
        for (int rep = 0; rep < 2; ++rep) {
          if (red) k<true><<<nsm, 512, 65536>>>(out, nw, bytes, opw, cta_stride, cyc);
          else k<false><<<nsm, 512, 65536>>>(out, nw, bytes, opw, cta_stride, cyc);
          cudaDeviceSynchronize();
        }
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        double mean = 0;
        for (int i = 0; i < nsm; ++i) mean += h[i];
        mean /= nsm;
        printf("%s bytes=%4d warps=%2d : %.1f cycles/op/SM  (%.1f B/cycle/SM, %.0f GB/s chip @1.965GHz)\n",
               red ? "reduce" : "store ", bytes, nw, mean / (opw * nw), bytes * (double)(opw * nw) / mean,
               bytes * (double)(opw * nw) / mean * 148 * 1.965);
      }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
