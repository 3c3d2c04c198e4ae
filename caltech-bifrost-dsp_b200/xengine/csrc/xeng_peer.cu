// Per-step products to a peer GPU (or to local memory) by SM stores, not by the copy engines.
//
// The reference ships CorrSubsel's fast dump and BeamformSum's powers to a central receiver as UDP packets every
// integration (corr_subsel_block.py:298-305, beamform_sum_beams_block.py:245-247).  Here the receiver is rank 0's HBM,
// mapped into every rank through torch symmetric memory, and a product is a few MB every step.  Sent through
// cudaMemcpyAsync those small copies share the copy-engine queue with the 766 MB long dump that leaves every
// acc_len gulps and sit behind it (profiles/r02v: the step that needs the push slot back waits ~1 ms at N=2).
// These two kernels keep the per-step product off that queue: 128 threads and 32 registers per CTA, no shared
// memory, so a CTA fits on an SM beside the persistent correlator's (832 threads x 72 registers, 216 KB).
#include "xeng_internal.h"

namespace xeng {
namespace {

constexpr int PUSH_THREADS = 128;

// CLEAR: the product is MOVED -- its source is zero afterwards, ready to be accumulated into again (the fused fast
// dump adds into its buffer: the per-step cudaMemsetAsync in front of the correlator launch goes away)
template <bool CLEAR>
__global__ void __launch_bounds__(PUSH_THREADS) peer_push_kernel(int4* __restrict__ dst, int4* __restrict__ src,
                                                                 size_t nvec) {
  const size_t stride = static_cast<size_t>(gridDim.x) * PUSH_THREADS;
  const int4 zero = make_int4(0, 0, 0, 0);
  size_t i = blockIdx.x * static_cast<size_t>(PUSH_THREADS) + threadIdx.x;
  for (; i + 3 * stride < nvec; i += 4 * stride) {
    int4 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = __ldcs(src + i + k * stride);
#pragma unroll
    for (int k = 0; k < 4; ++k) dst[i + k * stride] = v[k];
    if (CLEAR) {
#pragma unroll
      for (int k = 0; k < 4; ++k) src[i + k * stride] = zero;
    }
  }
  for (; i < nvec; i += stride) {
    dst[i] = __ldcs(src + i);
    if (CLEAR) src[i] = zero;
  }
}

// The sequence word follows its payload in stream order: a consumer that reads `seq` here finds the payload landed.
__global__ void peer_flag_kernel(long long* dst_seq, long long seq) {
  __threadfence_system();
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(dst_seq), "l"(seq) : "memory");
}

}  // namespace

cudaError_t launch_peer_push(void* dst, void* src, size_t nbytes, long long* dst_seq, long long seq, int nsm,
                             int clear_src, cudaStream_t stream) {
  const size_t nvec = nbytes / 16;
  if (nvec) {
    size_t blocks = (nvec + PUSH_THREADS * 4 - 1) / (PUSH_THREADS * 4);
    if (blocks > static_cast<size_t>(nsm)) blocks = nsm;        // one CTA per SM: what fits beside the correlator
    if (clear_src)
      peer_push_kernel<true><<<static_cast<int>(blocks), PUSH_THREADS, 0, stream>>>(static_cast<int4*>(dst),
                                                                                    static_cast<int4*>(src), nvec);
    else
      peer_push_kernel<false><<<static_cast<int>(blocks), PUSH_THREADS, 0, stream>>>(static_cast<int4*>(dst),
                                                                                     static_cast<int4*>(src), nvec);
  }
  if (dst_seq) peer_flag_kernel<<<1, 1, 0, stream>>>(dst_seq, seq);
  return cudaGetLastError();
}

}  // namespace xeng
