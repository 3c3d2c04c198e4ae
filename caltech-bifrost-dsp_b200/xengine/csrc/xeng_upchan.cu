// Offline upchannelise path: a short FFT along blocks of consecutive time samples of the raw 4+4-bit
// voltages, then either the upchannelised voltages themselves or the offline beamformer's weighted sum over inputs.
//
// In the reference this is a chain of stock bifrost blocks working through HBM-sized intermediates
// (/root/reference/pipeline/scripts/lwa352-upchan-bf.py:94-113 and lwa352-upchan-imag.py:93-98):
//   TrigBufSourceBlock  ci4 [time][fine_time][freq][stand][pol]        (trigger_source_block.py:37-41)
//   blocks.transpose -> [time][freq][stand][pol][fine_time]
//   blocks.fft(axes='fine_time')  -> cf32 [time][freq][stand][pol][fine_freq]   (forward, unnormalised, no fftshift)
//   imag:  blocks.transpose -> [time][freq][fine_freq][stand][pol], merge (freq, fine_freq)
//   bf:    BfOfflineWeightsBlock: x *= cal * coeffs, coeffs [beam][freq][stand][pol][fine_freq]
//          (beamform_offline_block.py:218-240), blocks.reduce over stand, then over pol
//          -> cf32 [time][beam][freq][fine_freq]
// Here each is ONE kernel: a thread transforms one (frame, chan, input) series in registers (radix-2 DIF, compile-time
// unrolled, twiddles as immediates), so the transposes never exist and, for the beamformer, neither do the
// upchannelised voltages (8 B per sample) nor the per-beam weighted copy of them (8 B per sample per beam).
// Algorithmic bytes: fft  1 B read + 8 B written per sample;  beamform  1 B read per sample + weights once
// (8 B * nbeam * nchan * ninput * nfine) + 8 B * nbeam / ninput per sample written.  Both HBM-bound.
#include "xeng_internal.h"

namespace xeng {
namespace {

// cos / sin of 2*pi*t/32, t = 0..15 (a switch over a compile-time t folds to an immediate)
__device__ __forceinline__ float tw_cos(int t) {
  switch (t) {
    case 0: return 1.0f;
    case 1: return 0.98078528040323043f;
    case 2: return 0.92387953251128674f;
    case 3: return 0.83146961230254524f;
    case 4: return 0.70710678118654757f;
    case 5: return 0.55557023301960229f;
    case 6: return 0.38268343236508984f;
    case 7: return 0.19509032201612833f;
    case 8: return 0.0f;
    case 9: return -0.19509032201612833f;
    case 10: return -0.38268343236508984f;
    case 11: return -0.55557023301960229f;
    case 12: return -0.70710678118654757f;
    case 13: return -0.83146961230254524f;
    case 14: return -0.92387953251128674f;
    default: return -0.98078528040323043f;
  }
}
__device__ __forceinline__ float tw_sin(int t) { return t <= 8 ? tw_cos(8 - t) : tw_cos(t - 8); }

__host__ __device__ constexpr int bitrev(int v, int n) {
  int r = 0;
  for (int b = 1; b < n; b <<= 1) { r = (r << 1) | (v & 1); v >>= 1; }
  return r;
}

// Forward DFT X[k] = sum_n x[n] exp(-2 pi i n k / N) in place, decimation in frequency: element j of the result
// holds X[bitrev(j)].
template <int N>
__device__ __forceinline__ void fft_dif(float (&re)[N], float (&im)[N]) {
#pragma unroll
  for (int half = N / 2; half >= 1; half >>= 1) {
#pragma unroll
    for (int base = 0; base < N; base += 2 * half) {
#pragma unroll
      for (int j = 0; j < half; ++j) {
        const int a = base + j, b = a + half;
        const float dr = re[a] - re[b], di = im[a] - im[b];
        re[a] += re[b];
        im[a] += im[b];
        const int t = j * (16 / half);   // twiddle exp(-2 pi i t / 32)
        if (t == 0) {
          re[b] = dr; im[b] = di;
        } else if (t == 8) {             // times -i
          re[b] = di; im[b] = -dr;
        } else {
          const float c = tw_cos(t), s = tw_sin(t);
          re[b] = dr * c + di * s;
          im[b] = di * c - dr * s;
        }
      }
    }
  }
}

// 4+4-bit complex byte -> (re, im), the unpack every block of the path uses (corr_block.py:270-275)
__device__ __forceinline__ void unpack4(uint32_t b, float& r, float& i) {
  r = static_cast<float>(static_cast<int>(static_cast<int8_t>(b)) >> 4);
  i = static_cast<float>(static_cast<int>(static_cast<int8_t>(b << 4)) >> 4);
}

// Raw samples of series (frame, chan, input), one per register (left unpacked so that nothing waits on the loads
// until the transform reads them).  in: u8 [nframe][N][nchan][ninput].
template <int N>
__device__ __forceinline__ void load_raw(const uint8_t* __restrict__ in, long frame, int chan, int input, int nchan,
                                         int ninput, uint32_t (&raw)[N]) {
  const uint8_t* p = in + ((frame * N) * nchan + chan) * static_cast<long>(ninput) + input;
  const long step = static_cast<long>(nchan) * ninput;
#pragma unroll
  for (int n = 0; n < N; ++n) raw[n] = __ldg(p + n * step);
}

template <int N>
__device__ __forceinline__ void fft_raw(const uint32_t (&raw)[N], float (&re)[N], float (&im)[N]) {
#pragma unroll
  for (int n = 0; n < N; ++n) unpack4(raw[n], re[n], im[n]);
  fft_dif<N>(re, im);
}

// Loads the N samples of series (frame, chan, input) and transforms them.
template <int N>
__device__ __forceinline__ void load_fft(const uint8_t* __restrict__ in, long frame, int chan, int input, int nchan,
                                         int ninput, float (&re)[N], float (&im)[N]) {
  uint32_t raw[N];
  load_raw<N>(in, frame, chan, input, nchan, ninput, raw);
  fft_raw<N>(raw, re, im);
}

constexpr int UPCHAN_WARPS = 4;
constexpr int TILE_F2 = 32 * 32 + 32;   // a warp's 32 inputs x N bins as float2, one pad element per 32

// flat element f = input_in_warp * N + k of a warp's tile; one pad per 32 keeps both access patterns conflict-free
__device__ __forceinline__ int tile_idx(int f) { return f + (f >> 5); }

// grid (nframe * nchan, ceil(ninput / 128)), 128 threads; lane = input.
// layout 0: out cf32 [nframe][nchan][ninput][N]   (bf script order)   -- through a per-warp shared-memory transpose
// layout 1: out cf32 [nframe][nchan][N][ninput]   (imag script order, = merged freq axis [nchan * N])
template <int N>
__global__ void __launch_bounds__(UPCHAN_WARPS * 32) upchan_fft_kernel(const uint8_t* __restrict__ in,
                                                                       float2* __restrict__ out, int nchan, int ninput,
                                                                       int layout) {
  __shared__ float2 tile[UPCHAN_WARPS][TILE_F2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long row = blockIdx.x;   // frame * nchan + chan
  const long frame = row / nchan;
  const int chan = static_cast<int>(row - frame * nchan);
  const int i0 = (blockIdx.y * UPCHAN_WARPS + warp) * 32;
  if (i0 >= ninput) return;
  const int nv = min(32, ninput - i0);
  float re[N], im[N];
  if (lane < nv) load_fft<N>(in, frame, chan, i0 + lane, nchan, ninput, re, im);
  if (layout == 1) {
    if (lane < nv) {
      float2* o = out + (row * N) * ninput + i0 + lane;
#pragma unroll
      for (int j = 0; j < N; ++j) __stcs(o + static_cast<long>(bitrev(j, N)) * ninput, make_float2(re[j], im[j]));
    }
    return;
  }
  if (lane < nv) {
#pragma unroll
    for (int j = 0; j < N; ++j) tile[warp][tile_idx(lane * N + bitrev(j, N))] = make_float2(re[j], im[j]);
  }
  __syncwarp();
  float2* o = out + (row * ninput + i0) * N;   // the warp's nv * N outputs are contiguous
#pragma unroll
  for (int j = 0; j < N; ++j) {
    const int f = j * 32 + lane;
    if (f < nv * N) __stcs(o + f, tile[warp][tile_idx(f)]);
  }
}

constexpr int BF_BC = 4;   // beams per pass over the inputs
constexpr int BF_FR = 5;   // frames per CTA = warps per CTA (a 480-sample gulp is 15 frames of 32)
#ifndef XENG_UPCHAN_MINB
#define XENG_UPCHAN_MINB 4
#endif
constexpr int BF_MINB = XENG_UPCHAN_MINB;   // resident CTAs per SM the register budget is set for

// grid (nchan, ceil(nframe / BF_FR)), BF_FR warps.  The CTA walks blocks of 32 inputs.  Phase 1: warp q transforms
// frame frame0 + q of the block (lane = input) into its shared-memory tile.  Phase 2: the warps split the block's
// 32 * N (input, bin) elements; a lane loads the BF_BC weights of its element ONCE (contiguous 256-byte rows of
// w[beam][chan][input][0..N)) and applies them to all BF_FR frames, so a weight is read nframe / BF_FR times in all
// instead of nframe times.  Element f = input * N + k sits in lane f % 32, hence a lane always works on bin
// k = lane % N and keeps BF_BC x BF_FR running sums; warps and the 32 / N input groups are folded at the end.
// w: cf32 [nbeam][nchan][ninput][N];  out: cf32 [nframe][nbeam][nchan][N].
template <int N>
__global__ void __launch_bounds__(BF_FR * 32, BF_MINB) upchan_beamform_kernel(const uint8_t* __restrict__ in,
                                                                     const float2* __restrict__ w,
                                                                     float2* __restrict__ out, int nframe, int nchan,
                                                                     int ninput, int nbeam) {
  __shared__ float2 tile[BF_FR * TILE_F2];
  float2* red = tile;   // [warp][BF_BC][BF_FR][32], aliased onto the tiles once a pass over the inputs is done
  static_assert(BF_FR * BF_BC * BF_FR * 32 <= BF_FR * TILE_F2, "reduction scratch must fit in the tiles");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int chan = blockIdx.x;
  const long frame0 = static_cast<long>(blockIdx.y) * BF_FR;
  const bool valid = frame0 + warp < nframe;
  const int nblk = (ninput + 31) >> 5;
  // gridDim.z = 2 splits the inputs between two CTAs that add their sums into the (zeroed) output: two addends
  // commute, so the result does not depend on which CTA arrives first
  const int blk_lo = static_cast<int>(static_cast<long>(nblk) * blockIdx.z / gridDim.z);
  const int blk_hi = static_cast<int>(static_cast<long>(nblk) * (blockIdx.z + 1) / gridDim.z);
  const long wbeam = static_cast<long>(nchan) * ninput * N;
  for (int b0 = 0; b0 < nbeam; b0 += BF_BC) {
    if (!valid)   // a frame past the end contributes zeros (its outputs are not written)
      for (int i = lane; i < TILE_F2; i += 32) tile[warp * TILE_F2 + i] = make_float2(0.f, 0.f);
    float2 acc[BF_BC][BF_FR];
#pragma unroll
    for (int b = 0; b < BF_BC; ++b)
#pragma unroll
      for (int q = 0; q < BF_FR; ++q) acc[b][q] = make_float2(0.f, 0.f);
    // weights of this lane's element j of block blk for the BF_BC beams of this pass (zeros outside the data)
    auto load_w = [&](int blk, int j, float2 (&c)[BF_BC]) {
      const int f = j * 32 + lane;
      const bool ok = f < min(32, ninput - blk * 32) * N;
      const float2* p = w + (static_cast<long>(b0) * nchan + chan) * ninput * N + static_cast<long>(blk) * 32 * N + f;
#pragma unroll
      for (int b = 0; b < BF_BC; ++b) c[b] = (ok && b0 + b < nbeam) ? __ldg(p + b * wbeam) : make_float2(0.f, 0.f);
    };
    // the weight loads run one element ahead of the arithmetic, across block boundaries, so their latency hides
    // behind the previous element's work (or the next block's transform)
    float2 cn[BF_BC];
    if (warp < N && blk_lo < blk_hi) load_w(blk_lo, warp, cn);
    for (int blk = blk_lo; blk < blk_hi; ++blk) {
      const int i0 = blk * 32;
      const int nv = min(32, ninput - i0);
      __syncthreads();   // phase 2 of the previous block has finished with the tiles
      if (valid && lane < nv) {
        float re[N], im[N];
        load_fft<N>(in, frame0 + warp, chan, i0 + lane, nchan, ninput, re, im);
#pragma unroll
        for (int j = 0; j < N; ++j) tile[warp * TILE_F2 + tile_idx(lane * N + bitrev(j, N))] = make_float2(re[j], im[j]);
      }
      __syncthreads();
      for (int j = warp; j < N; j += BF_FR) {
        float2 c[BF_BC];
#pragma unroll
        for (int b = 0; b < BF_BC; ++b) c[b] = cn[b];
        if (j + BF_FR < N) load_w(blk, j + BF_FR, cn);
        else if (blk + 1 < blk_hi) load_w(blk + 1, warp, cn);
        const int f = j * 32 + lane;   // = input_in_block * N + k
        if (f < nv * N) {
          const int ti = tile_idx(f);
#pragma unroll
          for (int q = 0; q < BF_FR; ++q) {
            const float2 x = tile[q * TILE_F2 + ti];
#pragma unroll
            for (int b = 0; b < BF_BC; ++b) {
              acc[b][q].x = fmaf(c[b].x, x.x, fmaf(-c[b].y, x.y, acc[b][q].x));
              acc[b][q].y = fmaf(c[b].x, x.y, fmaf(c[b].y, x.x, acc[b][q].y));
            }
          }
        }
      }
    }
    __syncthreads();   // tiles are dead: reuse them as reduction scratch
#pragma unroll
    for (int b = 0; b < BF_BC; ++b) {
#pragma unroll
      for (int q = 0; q < BF_FR; ++q) {
        float2 a = acc[b][q];
#pragma unroll
        for (int off = 16; off >= N; off >>= 1) {   // lanes sharing bin k = lane % N hold different input groups
          a.x += __shfl_xor_sync(0xffffffffu, a.x, off);
          a.y += __shfl_xor_sync(0xffffffffu, a.y, off);
        }
        red[((warp * BF_BC + b) * BF_FR + q) * 32 + lane] = a;
      }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < BF_BC * BF_FR * N; e += BF_FR * 32) {
      const int k = e % N, q = (e / N) % BF_FR, b = e / (N * BF_FR);
      if (b0 + b < nbeam && frame0 + q < nframe) {
        float2 s = make_float2(0.f, 0.f);
#pragma unroll
        for (int wq = 0; wq < BF_FR; ++wq) {
          const float2 r = red[((wq * BF_BC + b) * BF_FR + q) * 32 + k];
          s.x += r.x;
          s.y += r.y;
        }
        float2* o = out + (((frame0 + q) * nbeam + b0 + b) * nchan + chan) * N + k;
        if (gridDim.z == 1) {
          *o = s;
        } else {
          atomicAdd(&o->x, s.x);
          atomicAdd(&o->y, s.y);
        }
      }
    }
    __syncthreads();
  }
}

template <int N>
cudaError_t fft_n(const uint8_t* in, float2* out, long nframe, int nchan, int ninput, int layout, cudaStream_t st) {
  dim3 grid(static_cast<unsigned>(nframe * nchan), (ninput + UPCHAN_WARPS * 32 - 1) / (UPCHAN_WARPS * 32));
  upchan_fft_kernel<N><<<grid, UPCHAN_WARPS * 32, 0, st>>>(in, out, nchan, ninput, layout);
  return cudaGetLastError();
}
template <int N>
cudaError_t bf_n(const uint8_t* in, const float2* w, float2* out, long nframe, int nchan, int ninput, int nbeam,
                 cudaStream_t st) {
  const int split = ninput >= 128 ? 2 : 1;
  if (split > 1) {
    cudaError_t e = cudaMemsetAsync(out, 0, sizeof(float2) * nframe * nbeam * nchan * N, st);
    if (e != cudaSuccess) return e;
  }
  dim3 grid(nchan, static_cast<unsigned>((nframe + BF_FR - 1) / BF_FR), split);
  upchan_beamform_kernel<N><<<grid, BF_FR * 32, 0, st>>>(in, w, out, static_cast<int>(nframe), nchan, ninput, nbeam);
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_upchan_fft(const uint8_t* in, float2* out, long nframe, int nfine, int nchan, int ninput, int layout,
                              cudaStream_t st) {
  if (nframe <= 0) return cudaSuccess;
  switch (nfine) {
    case 2: return fft_n<2>(in, out, nframe, nchan, ninput, layout, st);
    case 4: return fft_n<4>(in, out, nframe, nchan, ninput, layout, st);
    case 8: return fft_n<8>(in, out, nframe, nchan, ninput, layout, st);
    case 16: return fft_n<16>(in, out, nframe, nchan, ninput, layout, st);
    case 32: return fft_n<32>(in, out, nframe, nchan, ninput, layout, st);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t launch_upchan_beamform(const uint8_t* in, const float2* w, float2* out, long nframe, int nfine, int nchan,
                                   int ninput, int nbeam, cudaStream_t st) {
  if (nframe <= 0) return cudaSuccess;
  switch (nfine) {
    case 2: return bf_n<2>(in, w, out, nframe, nchan, ninput, nbeam, st);
    case 4: return bf_n<4>(in, w, out, nframe, nchan, ninput, nbeam, st);
    case 8: return bf_n<8>(in, w, out, nframe, nchan, ninput, nbeam, st);
    case 16: return bf_n<16>(in, w, out, nframe, nchan, ninput, nbeam, st);
    case 32: return bf_n<32>(in, w, out, nframe, nchan, ninput, nbeam, st);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace xeng
