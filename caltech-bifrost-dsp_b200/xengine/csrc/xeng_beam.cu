// Beamform + beam-power integration for sm_100a.
//
// bfBeamformRun  (call site /root/reference/pipeline/lwa352_pipeline/blocks/beamform_block.py:449):
//     out[c][b][t] = sum_i w[c][b][i] * v[t][c][i]         complex fp32, no conjugation
//     (pinned by pipeline/verification/tests/beamformer/beamformer_test.py:76-84)
// bfBeamformIntegrate / ...SingleBeam (beamform_sum_beams_block.py:245, beamform_sum_single_beam_block.py:114):
//     out[p][T][c][0..3] = sum_{t in T} [ |X|^2, |Y|^2, Re X Y*, Im X Y* ],  X = beam 2p, Y = beam 2p+1
//     (pinned by pipeline/verification/tests/beamformer/beamformer_sum_test.py:64-77)
//
// The contraction is done in fp32 on the FMA pipe (bit-for-bit fp32 weights, exact integer voltages):
// a CTA owns (channel, 128 time samples, 32 beams); 4+4-bit voltages are unpacked once per 64-input
// chunk into shared memory as complex fp32 and reused by all 32 beams; each thread keeps a
// 4-beam x 4-time complex register tile (64 FFMA per 4 LDS.128).
#include "xeng_internal.h"

namespace xeng {
namespace {

constexpr int BB = 32;    // beams per CTA
constexpr int TT = 128;   // time samples per CTA
constexpr int IK = 64;    // inputs per shared-memory chunk
constexpr int BF_THREADS = 256;

struct BeamSmem {
  float2 v[IK][TT];   // 64 KB
  float2 w[BB][IK];   // 16 KB
};

__device__ __forceinline__ void load_v32(const BeamArgs& a, int c, int t, int i0, bool vec_ok, uint4 (&raw)[2]) {
  raw[0] = raw[1] = make_uint4(0, 0, 0, 0);
  if (t >= a.ntime) return;
  const uint8_t* p = a.in + (static_cast<long>(t) * a.nchan + c) * a.ninput;
  if (vec_ok && i0 + 32 <= a.ninput) {
    raw[0] = __ldg(reinterpret_cast<const uint4*>(p + i0));
    raw[1] = __ldg(reinterpret_cast<const uint4*>(p + i0 + 16));
  } else {
    uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int k = 0; k < 32; ++k)
      if (i0 + k < a.ninput) w[k >> 2] |= static_cast<uint32_t>(p[i0 + k]) << (8 * (k & 3));
    raw[0] = make_uint4(w[0], w[1], w[2], w[3]);
    raw[1] = make_uint4(w[4], w[5], w[6], w[7]);
  }
}

__device__ __forceinline__ float2 decode44(uint32_t byte) {
  const int b = static_cast<int>(byte);
  const int re = (b << 24) >> 28;   // high nibble, sign-extended
  const int im = (b << 28) >> 28;   // low nibble, sign-extended
  return make_float2(static_cast<float>(re), static_cast<float>(im));
}

// FUSED: the |B|^2 integration of BeamformSumBeams runs in the epilogue and only the powers leave the SM.  The CTA
// then owns tile_t = the largest multiple of ntime_sum that fits in TT samples (120 of 128 for ntime_sum = 24: the
// same number of CTAs along time for a 960-sample gulp) so that every integration block lies inside one CTA; the
// per-sample products and the time-ordered sums are the expressions of beam_integrate_kernel below.
template <bool FUSED>
__global__ void __launch_bounds__(BF_THREADS, 2) beamform_kernel(BeamArgs a) {
  extern __shared__ __align__(16) uint8_t bf_smem[];
  BeamSmem& S = *reinterpret_cast<BeamSmem*>(bf_smem);
  const int tid = threadIdx.x;
  const int t0 = blockIdx.x * (FUSED ? a.tile_t : TT), c = blockIdx.y, b0 = blockIdx.z * BB;
  const int tb = tid >> 5, tt = tid & 31;

  // staging roles
  const int vt = tid & (TT - 1), vhalf = tid >> 7;     // v: time sample, which 32-input half of the chunk
  const bool vec_ok = (a.ninput % 16 == 0) && ((reinterpret_cast<uintptr_t>(a.in) & 15) == 0);
  const bool wvec_ok = (a.ninput % 2 == 0) && ((reinterpret_cast<uintptr_t>(a.w) & 15) == 0);

  float2 acc[4][4];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int k = 0; k < 4; ++k) acc[j][k] = make_float2(0.f, 0.f);

  const int nchunk = (a.ninput + IK - 1) / IK;
  uint4 vraw[2];
  float4 wraw[4];
  auto load_w = [&](int i0) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int idx = j * BF_THREADS + tid;
      const int b = idx >> 5, i2 = idx & 31;       // 32 float4 (= 64 inputs) per beam row
      const int i = i0 + 2 * i2;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (b0 + b < a.nbeam && i < a.ninput) {
        const float2* wp = a.w + (static_cast<long>(c) * a.nbeam + (b0 + b)) * a.ninput + i;
        if (wvec_ok && i + 1 < a.ninput) {
          v = __ldg(reinterpret_cast<const float4*>(wp));
        } else {
          const float2 x = wp[0];
          v.x = x.x; v.y = x.y;
          if (i + 1 < a.ninput) { const float2 y = wp[1]; v.z = y.x; v.w = y.y; }
        }
      }
      wraw[j] = v;
    }
  };
  load_v32(a, c, t0 + vt, vhalf * 32, vec_ok, vraw);
  load_w(0);

  for (int ch = 0; ch < nchunk; ++ch) {
    __syncthreads();  // previous chunk fully consumed
    {
      const uint32_t words[8] = {vraw[0].x, vraw[0].y, vraw[0].z, vraw[0].w, vraw[1].x, vraw[1].y, vraw[1].z, vraw[1].w};
#pragma unroll
      for (int k = 0; k < 32; ++k) S.v[vhalf * 32 + k][vt] = decode44((words[k >> 2] >> (8 * (k & 3))) & 0xFF);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int idx = j * BF_THREADS + tid;
        *reinterpret_cast<float4*>(&S.w[idx >> 5][2 * (idx & 31)]) = wraw[j];
      }
    }
    __syncthreads();
    if (ch + 1 < nchunk) {  // prefetch next chunk into registers while computing this one
      load_v32(a, c, t0 + vt, (ch + 1) * IK + vhalf * 32, vec_ok, vraw);
      load_w((ch + 1) * IK);
    }
#pragma unroll 4
    for (int i = 0; i < IK; i += 2) {
      float4 wv[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) wv[j] = *reinterpret_cast<const float4*>(&S.w[4 * tb + j][i]);
      // a thread's four samples are {2tt, 2tt+1, 64+2tt, 64+2tt+1}: each LDS.128 of a warp is 512 contiguous bytes
      // (4 wavefronts; samples 4tt..4tt+3 would put the lanes 32 B apart and cost 8)
      const float4 va0 = *reinterpret_cast<const float4*>(&S.v[i][2 * tt]);
      const float4 va1 = *reinterpret_cast<const float4*>(&S.v[i][TT / 2 + 2 * tt]);
      const float4 vb0 = *reinterpret_cast<const float4*>(&S.v[i + 1][2 * tt]);
      const float4 vb1 = *reinterpret_cast<const float4*>(&S.v[i + 1][TT / 2 + 2 * tt]);
      const float2 va[4] = {make_float2(va0.x, va0.y), make_float2(va0.z, va0.w), make_float2(va1.x, va1.y),
                            make_float2(va1.z, va1.w)};
      const float2 vb[4] = {make_float2(vb0.x, vb0.y), make_float2(vb0.z, vb0.w), make_float2(vb1.x, vb1.y),
                            make_float2(vb1.z, vb1.w)};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float re = acc[j][k].x, im = acc[j][k].y;
          re = fmaf(wv[j].x, va[k].x, re);
          re = fmaf(-wv[j].y, va[k].y, re);
          im = fmaf(wv[j].x, va[k].y, im);
          im = fmaf(wv[j].y, va[k].x, im);
          re = fmaf(wv[j].z, vb[k].x, re);
          re = fmaf(-wv[j].w, vb[k].y, re);
          im = fmaf(wv[j].z, vb[k].y, im);
          im = fmaf(wv[j].w, vb[k].x, im);
          acc[j][k] = make_float2(re, im);
        }
      }
    }
  }

  if (FUSED) {
    __syncthreads();                                       // every warp is done with S.v: reuse it for the products
    float4* prod = reinterpret_cast<float4*>(&S.v[0][0]);  // [BB / 2 dual-pol beams][TT samples], 32 KB
#pragma unroll
    for (int jp = 0; jp < 2; ++jp) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 x = acc[2 * jp][k], y = acc[2 * jp + 1][k];
        prod[(2 * tb + jp) * TT + (k >> 1) * (TT / 2) + 2 * tt + (k & 1)] = beam_products(x, y);
      }
    }
    __syncthreads();
    const int nblk = a.ntime / a.ntime_sum, per_cta = a.tile_t / a.ntime_sum, T0 = t0 / a.ntime_sum;
    float4* pout = reinterpret_cast<float4*>(a.pow_out);
    for (int idx = tid; idx < (BB / 2) * per_cta; idx += BF_THREADS) {
      const int pair = idx / per_cta, Tl = idx - pair * per_cta;
      if (T0 + Tl >= nblk || b0 + 2 * pair + 1 >= a.nbeam) continue;
      float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
      const float4* q = prod + pair * TT + Tl * a.ntime_sum;
      for (int k = 0; k < a.ntime_sum; ++k) {
        sum.x += q[k].x; sum.y += q[k].y; sum.z += q[k].z; sum.w += q[k].w;
      }
      pout[(static_cast<long>(b0 / 2 + pair) * nblk + T0 + Tl) * a.nchan + c] = sum;
    }
    return;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int b = b0 + 4 * tb + j;
    if (b >= a.nbeam) continue;
#pragma unroll
    for (int h = 0; h < 2; ++h) {                          // the two sample pairs of the thread
      const int t = t0 + h * (TT / 2) + 2 * tt;
      float2* o = a.out + (static_cast<long>(c) * a.nbeam + b) * a.ntime + t;
      if (t + 1 < a.ntime && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
        *reinterpret_cast<float4*>(o) = make_float4(acc[j][2 * h].x, acc[j][2 * h].y, acc[j][2 * h + 1].x, acc[j][2 * h + 1].y);
      } else {
#pragma unroll
        for (int k = 0; k < 2; ++k)
          if (t + k < a.ntime) o[k] = acc[j][2 * h + k];
      }
    }
  }
}

// One block per (channel, dual-pol beam): per-sample products go to shared memory with coalesced
// loads, then one thread per output time block sums ntime_sum of them in time order.
constexpr int PW_T = 1024;
__global__ void __launch_bounds__(256) beam_integrate_kernel(const float2* __restrict__ in, float4* __restrict__ out,
                                                             int nchan, int nbeam, int ntime, int ntime_sum,
                                                             int beam0, int nbeam_out) {
  // products of output block T start at element T * (ntime_sum + 1): the one-element skew keeps the lanes of the
  // summing pass (one per output block, stride ntime_sum float4) off each other's banks for even ntime_sum
  __shared__ float4 prod[2 * PW_T];
  const int c = blockIdx.x, p = blockIdx.y;
  const float2* X = in + (static_cast<long>(c) * nbeam + 2 * (beam0 + p)) * ntime;
  const float2* Y = X + ntime;
  const int nblk = ntime / ntime_sum;
  const int span = (PW_T / ntime_sum) * ntime_sum;   // whole output blocks per pass
  if (span == 0) {
    // ntime_sum > PW_T: one thread per output block walks global memory directly
    for (int T = threadIdx.x; T < nblk; T += blockDim.x) {
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k = 0; k < ntime_sum; ++k) {
        const float4 pr = beam_products(X[T * ntime_sum + k], Y[T * ntime_sum + k]);
        s.x += pr.x;
        s.y += pr.y;
        s.z += pr.z;
        s.w += pr.w;
      }
      out[(static_cast<long>(p) * nblk + T) * nchan + c] = s;
    }
    return;
  }
  for (int tbase = 0; tbase < nblk * ntime_sum; tbase += span) {
    const int n = min(span, nblk * ntime_sum - tbase);
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
      prod[k + k / ntime_sum] = beam_products(X[tbase + k], Y[tbase + k]);
    }
    __syncthreads();
    // one thread per (output block, Stokes-like component): 4x the threads of a float4-per-thread pass, each
    // component still summed in time order
    const float* prodf = reinterpret_cast<const float*>(prod);
    float* outf = reinterpret_cast<float*>(out);
    for (int e = threadIdx.x; e < 4 * (n / ntime_sum); e += blockDim.x) {
      const int T = e >> 2, comp = e & 3;
      float s = 0.f;
      for (int k = 0; k < ntime_sum; ++k) s += prodf[(T * (ntime_sum + 1) + k) * 4 + comp];
      outf[((static_cast<long>(p) * nblk + tbase / ntime_sum + T) * nchan + c) * 4 + comp] = s;
    }
    __syncthreads();
  }
}

// BeamformVlbiOutput's tap: the first nsend voltage beams, [chan][beam][time] -> [time][chan][beam]
// (idata[:, 0:nsend, :].transpose([2,0,1]), beamform_vlbi_output_block.py:260-264).  32 channels x 32 samples per
// block through shared memory: reads run along time, writes along (chan, beam).
__global__ void __launch_bounds__(1024) vlbi_tap_kernel(const float2* __restrict__ in, float2* __restrict__ out,
                                                        int nchan, int nbeam, int ntime, int nsend) {
  __shared__ float2 tile[32][33];
  const int c0 = blockIdx.y * 32, t0 = blockIdx.x * 32;
  for (int b = 0; b < nsend; ++b) {
    const int c = c0 + threadIdx.y, t = t0 + threadIdx.x;
    if (c < nchan && t < ntime) tile[threadIdx.y][threadIdx.x] = in[(static_cast<long>(c) * nbeam + b) * ntime + t];
    __syncthreads();
    const int co = c0 + threadIdx.x, to = t0 + threadIdx.y;
    if (co < nchan && to < ntime) out[(static_cast<long>(to) * nchan + co) * nsend + b] = tile[threadIdx.x][threadIdx.y];
    __syncthreads();
  }
}

}  // namespace

cudaError_t launch_vlbi_tap(const float2* in, float2* out, int nchan, int nbeam, int ntime, int nsend,
                            cudaStream_t stream) {
  if (nchan <= 0 || ntime <= 0 || nsend <= 0) return cudaSuccess;
  dim3 grid((ntime + 31) / 32, (nchan + 31) / 32);
  vlbi_tap_kernel<<<grid, dim3(32, 32), 0, stream>>>(in, out, nchan, nbeam, ntime, nsend);
  return cudaGetLastError();
}

cudaError_t launch_beamform(const BeamArgs& a, cudaStream_t stream) {
  cudaError_t ce = configure_once_per_device([]() -> cudaError_t {
    cudaError_t e = cudaFuncSetAttribute(beamform_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(sizeof(BeamSmem)));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(beamform_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(sizeof(BeamSmem)));
    return e;
  }, kCfgBeam);
  if (ce != cudaSuccess) return ce;
  if (a.ntime <= 0 || a.nchan <= 0 || a.nbeam <= 0) return cudaSuccess;
  if (a.pow_out) {                                          // fused power integration
    if (a.ntime_sum <= 0 || a.ntime_sum > TT || (a.nbeam & 1)) return cudaErrorInvalidValue;
    BeamArgs f = a;
    f.tile_t = (TT / a.ntime_sum) * a.ntime_sum;
    const int nt = (a.ntime / a.ntime_sum) * a.ntime_sum;   // whole integration blocks only
    if (nt == 0) return cudaSuccess;
    dim3 grid((nt + f.tile_t - 1) / f.tile_t, a.nchan, (a.nbeam + BB - 1) / BB);
    beamform_kernel<true><<<grid, BF_THREADS, sizeof(BeamSmem), stream>>>(f);
    return cudaGetLastError();
  }
  dim3 grid((a.ntime + TT - 1) / TT, a.nchan, (a.nbeam + BB - 1) / BB);
  beamform_kernel<false><<<grid, BF_THREADS, sizeof(BeamSmem), stream>>>(a);
  return cudaGetLastError();
}
int beamform_fused_max_ntime_sum() { return TT; }

cudaError_t launch_beam_integrate(const float2* in, float* out, int nchan, int nbeam, int ntime, int ntime_sum,
                                  int beam0, int nbeam_out, cudaStream_t stream) {
  if (nchan <= 0 || nbeam_out <= 0 || ntime / ntime_sum <= 0) return cudaSuccess;
  dim3 grid(nchan, nbeam_out);
  beam_integrate_kernel<<<grid, 256, 0, stream>>>(in, reinterpret_cast<float4*>(out), nchan, nbeam, ntime, ntime_sum,
                                                  beam0, nbeam_out);
  return cudaGetLastError();
}

}  // namespace xeng
