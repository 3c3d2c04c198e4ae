// Microbenchmark: what in the correlator epilogue's access pattern costs cp.reduce.async.bulk its sequential rate?
// 148 CTAs x 8 issuing warps, 512-byte reductions from shared memory, > L2 footprint.  Variants:
//   seq        every warp streams its own contiguous region, deep queue (wait_group.read 2 every 16 ops)  [bulk_rate.cu]
//   seq_drain  the same addresses, but a warp issues 16 ops and waits until ALL have read their source before the next
//              16 -- the epilogue's constraint (one staging buffer per warp)
//   tile_drain the epilogue's addresses: a CTA owns 512 B of each 2 KB row (4 neighbouring CTAs complete a row), a
//              warp's 16 ops are 8 consecutive rows in each of two regions far apart, 8 warps = 4 row groups x 2
//              region pairs; drain as above
//   tile       the same addresses with the deep queue
// Developer tool; not part of the library.
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(256) k(int32_t* out, int mode, int rounds, size_t region_words, long long* cycles) {
  extern __shared__ __align__(128) uint8_t sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<int*>(sm)[i] = 1;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  const bool tile = mode >= 2, drain = (mode & 1);
  long long t0 = clock64();
  if (lane == 0) {
    const uint32_t src = smem_u32(sm) + warp * 8192;
    for (int g = 0; g < rounds; ++g) {
      for (int kk = 0; kk < 16; ++kk) {
        int32_t* dst;
        if (!tile) {
          // contiguous per warp: (cta, warp) owns rounds * 16 * 512 bytes
          dst = out + ((size_t)(blockIdx.x * 8 + warp) * rounds * 16 + (size_t)g * 16 + kk) * 128;
        } else {
          const int region = 2 * (warp >> 2) + (kk >> 3) + 4 * (g & 1);          // 8 regions (2 planes x 4 quadrants)
          const int row = (warp & 3) * 8 + (kk & 7);
          const size_t rowblock = (size_t)(blockIdx.x >> 2) * (rounds / 2) + (g >> 1);   // 32 rows of 2 KB
          dst = out + (size_t)region * region_words + (rowblock * 32 + row) * 512 + (blockIdx.x & 3) * 128;
        }
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.s32 [%0], [%1], %2;" ::"l"(dst),
                     "r"(src + (kk & 15) * 512), "r"(512) : "memory");
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      if (drain) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      else asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

int main() {
  const int nsm = 148, rounds = 64;                     // 64 rounds x 8 warps x 16 ops x 512 B = 4 MB per CTA
  const size_t total = (size_t)nsm * rounds * 8 * 16 * 512;   // 592 MB
  const size_t region_words = total / 8 / 4 + 4096;
  int32_t* out;
  long long* cyc;
  cudaMalloc(&out, region_words * 4 * 8 + (1 << 20));
  cudaMemset(out, 0, region_words * 4 * 8 + (1 << 20));
  cudaMalloc(&cyc, nsm * sizeof(long long));
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  const char* names[4] = {"seq", "seq_drain", "tile", "tile_drain"};
  long long h[nsm];
  for (int rep = 0; rep < 2; ++rep)
    for (int mode = 0; mode < 4; ++mode) {
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      k<<<nsm, 256, 65536>>>(out, mode, rounds, region_words, cyc);
      cudaDeviceSynchronize();
      cudaEventRecord(e0);
      k<<<nsm, 256, 65536>>>(out, mode, rounds, region_words, cyc);
      cudaEventRecord(e1);
      cudaDeviceSynchronize();
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
      double mean = 0;
      for (int i = 0; i < nsm; ++i) mean += h[i];
      mean /= nsm;
      printf("%-10s : %.3f ms, %.0f GB/s of payload (chip), %.1f B/cycle/SM\n", names[mode], ms, total / ms / 1e6,
             (double)rounds * 8 * 16 * 512 / mean);
    }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
