// Offline imaging path (second half): cross-correlation of the UPCHANNELISED voltages.
//
// scripts/lwa352-upchan-imag.py:93-106 feeds xeng_upchan_fft's FREQ_MERGED output ([time][freq][stand][pol], cf32)
// through a frequency selection into stock bifrost `blocks.correlate(nframe_per_integration=...)` -- a complex fp32
// Hermitian rank-nframe update per frequency, accumulated over the frames of an integration.  bifrost is an absent
// submodule, so the output convention is taken from the reference's own correlator checks (check_golden_inputs.py:64-92,
// corr_output_full_block.py:582-591): X[f][i][j] = sum_t v[t][f][i] * conj(v[t][f][j]), lower triangle (j <= i)
// -- PARITY UNPINNED by reference vectors, checked against the float64 oracle.
//
// The voltages are FFT outputs (not integers), so this runs on the fp32 FMA pipe: 64 x 64 output tiles, 256 threads,
// 4 x 4 complex accumulators per thread, 16-frame chunks double-buffered in shared memory with the next chunk's global
// loads in flight.  Only tiles on or below the diagonal are computed and written (diagonal tiles are written whole).
#include "xeng_internal.h"

namespace xeng {
namespace {

constexpr int UC_T = 64;      // tile edge (inputs)
constexpr int UC_K = 16;      // frames per shared-memory chunk
constexpr int UC_THREADS = 256;

__global__ void __launch_bounds__(UC_THREADS) upchan_correlate_kernel(const float2* __restrict__ in, float2* __restrict__ out,
                                                                       int nframe, int nfreq, int ninput, int first) {
  __shared__ float2 As[2][UC_K][UC_T];      // double buffered: one barrier per chunk
  __shared__ float2 Bs[2][UC_K][UC_T];
  int I = 0, p = blockIdx.x;            // tile pair index p = I (I + 1) / 2 + J, J <= I
  while (p > I) {
    p -= I + 1;
    ++I;
  }
  const int J = p;
  const int f = blockIdx.y;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  // loader: thread -> (frame of the chunk, 4 consecutive inputs); a frame's 64 inputs are 512 contiguous bytes
  const int lk = threadIdx.x >> 4, li = (threadIdx.x & 15) * 4;
  const long fstride = static_cast<long>(nfreq) * ninput;
  const float2* pa = in + static_cast<long>(f) * ninput + I * UC_T + li;
  const float2* pb = in + static_cast<long>(f) * ninput + J * UC_T + li;
  const float2 zero = make_float2(0.f, 0.f);
  float2 ga[4], gb[4];
  auto fetch = [&](int k0) {
    const int t = k0 + lk;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const bool oka = t < nframe && I * UC_T + li + e < ninput, okb = t < nframe && J * UC_T + li + e < ninput;
      ga[e] = oka ? __ldg(pa + t * fstride + e) : zero;
      gb[e] = okb ? __ldg(pb + t * fstride + e) : zero;
    }
  };
  float2 acc[4][4];
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[r][c] = zero;

  fetch(0);
  int buf = 0;
  for (int k0 = 0; k0 < nframe; k0 += UC_K, buf ^= 1) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      As[buf][lk][li + e] = ga[e];
      Bs[buf][lk][li + e] = gb[e];
    }
    __syncthreads();   // chunk k0 is in As/Bs[buf]; every thread has finished reading [buf] two chunks ago (the
                       // barrier of the previous iteration lies between those reads and these writes)
    if (k0 + UC_K < nframe) fetch(k0 + UC_K);   // in flight while this chunk is multiplied
#pragma unroll
    for (int k = 0; k < UC_K; ++k) {
      float2 a[4], b[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) a[r] = As[buf][k][ty + 16 * r];
#pragma unroll
      for (int c = 0; c < 4; ++c) b[c] = Bs[buf][k][tx + 16 * c];
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) {   // a * conj(b)
          acc[r][c].x = fmaf(a[r].x, b[c].x, fmaf(a[r].y, b[c].y, acc[r][c].x));
          acc[r][c].y = fmaf(a[r].y, b[c].x, fmaf(-a[r].x, b[c].y, acc[r][c].y));
        }
    }
  }
  float2* o = out + static_cast<long>(f) * ninput * ninput;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int i = I * UC_T + ty + 16 * r;
    if (i >= ninput) continue;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int j = J * UC_T + tx + 16 * c;
      if (j >= ninput) continue;
      float2* q = o + static_cast<long>(i) * ninput + j;
      float2 v = acc[r][c];
      if (!first) {
        const float2 old = *q;
        v.x += old.x;
        v.y += old.y;
      }
      *q = v;
    }
  }
}

}  // namespace

cudaError_t launch_upchan_correlate(const float2* in, float2* out, int nframe, int nfreq, int ninput, int first,
                                    cudaStream_t stream) {
  if (nframe <= 0 || nfreq <= 0 || ninput <= 0) return cudaSuccess;
  const int T = (ninput + UC_T - 1) / UC_T;
  dim3 grid(T * (T + 1) / 2, nfreq);
  upchan_correlate_kernel<<<grid, UC_THREADS, 0, stream>>>(in, out, nframe, nfreq, ninput, first);
  return cudaGetLastError();
}

}  // namespace xeng
