// Bandwidth-bound visibility kernels: CorrAcc accumulate, CorrSubsel gather, xGPU-order -> logical reorder.
#include "xeng_internal.h"

namespace xeng {
namespace {

// ---------------------------------------------------------------------------------------------
// CorrAcc: acc = in (first span of an integration) or acc += in
// (/root/reference/pipeline/lwa352_pipeline/blocks/corr_acc_block.py:299-306, BFMap "a = b" / "a += b").
// 12 B of HBM traffic per int32 word (8 B on the first span).  128-bit accesses, 4 in flight per thread.
template <bool FIRST>
__global__ void __launch_bounds__(256) acc_kernel(int4* __restrict__ acc, const int4* __restrict__ in, size_t nvec) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  for (; i + 3 * stride < nvec; i += 4 * stride) {
    int4 b[4], a[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) b[k] = __ldcs(in + i + k * stride);
    if (!FIRST) {
#pragma unroll
      for (int k = 0; k < 4; ++k) a[k] = acc[i + k * stride];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        b[k].x += a[k].x; b[k].y += a[k].y; b[k].z += a[k].z; b[k].w += a[k].w;
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) acc[i + k * stride] = b[k];
  }
  for (; i < nvec; i += stride) {
    int4 b = __ldcs(in + i);
    if (!FIRST) {
      int4 a = acc[i];
      b.x += a.x; b.y += a.y; b.z += a.z; b.w += a.w;
    }
    acc[i] = b;
  }
}
template <bool FIRST>
__global__ void acc_tail_kernel(int32_t* __restrict__ acc, const int32_t* __restrict__ in, size_t begin, size_t n) {
  size_t i = begin + blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (i < n) acc[i] = FIRST ? in[i] : acc[i] + in[i];
}

// ---------------------------------------------------------------------------------------------
// CorrSubsel: out[c', v] = sum_{k < nchan_sum} X[c'*nchan_sum + k][vismap[v]], imag negated where conj[v]
// (bfXgpuSubSelect, call site corr_subsel_block.py:298; semantics pinned by
// pipeline/verification/test_corr_part_rx.py:38-49,76-84).  Output re/im interleaved (corr_subsel_block.py:302).
__global__ void __launch_bounds__(256) subsel_kernel(const int32_t* __restrict__ in, int2* __restrict__ out,
                                                     const int32_t* __restrict__ vismap,
                                                     const int32_t* __restrict__ conj, int nvis, int nchan_out,
                                                     int nchan_sum, long W, long plane_words) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  const int co = blockIdx.y;
  if (v >= nvis || co >= nchan_out) return;
  const long idx = vismap[v];
  const int32_t* p = in + static_cast<long>(co) * nchan_sum * W + idx;
  int re = 0, im = 0;
  for (int k = 0; k < nchan_sum; ++k) {
    re += __ldg(p + k * W);
    im += __ldg(p + plane_words + k * W);
  }
  if (conj[v]) im = -im;
  out[static_cast<long>(co) * nvis + v] = make_int2(re, im);
}

// ---------------------------------------------------------------------------------------------
// bfXgpuReorder (call site corr_output_full_block.py:669): xGPU order -> [nstand][nstand][npol][npol][nchan][2]
// holding v(s0,p0) conj(v(s1,p1)) (corr_output_full_block.py:386-392,582-591).
// 32 logical entries x 32 channels per block, transposed through shared memory so the 8-byte (re,im)
// writes are contiguous along channel.
__global__ void __launch_bounds__(1024) reorder_kernel(const int32_t* __restrict__ in, int2* __restrict__ out,
                                                       const int32_t* __restrict__ a2bl,
                                                       const int32_t* __restrict__ conj, long nent, int nchan, long W,
                                                       long plane_words) {
  __shared__ int2 tile[32][33];
  const long e0 = static_cast<long>(blockIdx.x) * 32;
  const int c0 = blockIdx.y * 32;
  {
    const long e = e0 + threadIdx.x;
    const int c = c0 + threadIdx.y;
    if (e < nent && c < nchan) {
      const long idx = a2bl[e];
      const int re = __ldg(in + static_cast<long>(c) * W + idx);
      int im = __ldg(in + plane_words + static_cast<long>(c) * W + idx);
      if (conj[e]) im = -im;
      tile[threadIdx.x][threadIdx.y] = make_int2(re, im);
    }
  }
  __syncthreads();
  {
    const long e = e0 + threadIdx.y;
    const int c = c0 + threadIdx.x;
    if (e < nent && c < nchan) out[e * nchan + c] = tile[threadIdx.y][threadIdx.x];
  }
}

}  // namespace

cudaError_t launch_acc(int32_t* acc, const int32_t* in, size_t n, int first, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const bool aligned = ((reinterpret_cast<uintptr_t>(acc) | reinterpret_cast<uintptr_t>(in)) & 15) == 0;
  const size_t nvec = aligned ? n / 4 : 0;
  if (nvec) {
    size_t blocks = (nvec + 256 * 4 - 1) / (256 * 4);
    const size_t cap = 148 * 8;  // 8 resident CTAs of 256 threads per SM
    if (blocks > cap) blocks = cap;
    if (first) acc_kernel<true><<<static_cast<int>(blocks), 256, 0, stream>>>(reinterpret_cast<int4*>(acc), reinterpret_cast<const int4*>(in), nvec);
    else acc_kernel<false><<<static_cast<int>(blocks), 256, 0, stream>>>(reinterpret_cast<int4*>(acc), reinterpret_cast<const int4*>(in), nvec);
  }
  const size_t done = nvec * 4;
  if (done < n) {
    const int blocks = static_cast<int>((n - done + 255) / 256);
    if (first) acc_tail_kernel<true><<<blocks, 256, 0, stream>>>(acc, in, done, n);
    else acc_tail_kernel<false><<<blocks, 256, 0, stream>>>(acc, in, done, n);
  }
  return cudaGetLastError();
}

cudaError_t launch_subsel(const int32_t* in, int32_t* out, const int32_t* vismap, const int32_t* conj, int nvis,
                          int nchan, int nchan_sum, long W, cudaStream_t stream) {
  const int nchan_out = nchan / nchan_sum;
  if (nvis <= 0 || nchan_out <= 0) return cudaSuccess;
  dim3 grid((nvis + 255) / 256, nchan_out);
  subsel_kernel<<<grid, 256, 0, stream>>>(in, reinterpret_cast<int2*>(out), vismap, conj, nvis, nchan_out, nchan_sum, W,
                                          static_cast<long>(nchan) * W);
  return cudaGetLastError();
}

cudaError_t launch_reorder(const int32_t* in, int32_t* out, const int32_t* a2bl, const int32_t* conj, int nstand,
                           int npol, int nchan, long W, cudaStream_t stream) {
  const long nent = static_cast<long>(nstand) * nstand * npol * npol;
  if (nent == 0 || nchan == 0) return cudaSuccess;
  dim3 grid(static_cast<unsigned>((nent + 31) / 32), (nchan + 31) / 32);
  reorder_kernel<<<grid, dim3(32, 32), 0, stream>>>(in, reinterpret_cast<int2*>(out), a2bl, conj, nent, nchan, W,
                                                    static_cast<long>(nchan) * W);
  return cudaGetLastError();
}

}  // namespace xeng
