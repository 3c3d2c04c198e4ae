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
// LAYOUT 0: [chan'][vis][2], the ring layout above.  The other two are the payloads CorrOutputPart builds from it on
// the host: 1 = "COR" packets [vis/4][chan'][4][2] (send_packets_bf's dview.transpose([1,0,2,3]),
// corr_output_part_block.py:378-391); 2 = plain packets [vis][chan'][2] in big-endian words (dout, dtype '>i',
// corr_output_part_block.py:429,460-463).
template <int LAYOUT>
__global__ void __launch_bounds__(256) subsel_kernel(const int32_t* __restrict__ in, int2* __restrict__ out,
                                                     const int32_t* __restrict__ vismap,
                                                     const int32_t* __restrict__ conj, int nvis, int nchan_out,
                                                     int nchan_sum, long W, long plane_words) {
  const int v0 = blockIdx.x * blockDim.x + threadIdx.x;
  const int co = blockIdx.y;
  if (co >= nchan_out) return;
  const bool live = v0 < nvis;
  const int v = live ? v0 : nvis - 1;                  // idle lanes shadow the last entry: the warp stays whole
  const long idx = vismap[v];
  // The four pol products of a stand pair are one 16-byte cell, and a real selection (CorrSubsel's 48 x 49 / 2 stand
  // pairs x 4 products, corr_subsel_block.py:185) takes whole cells on neighbouring entries: the lanes of a warp
  // that want words of the same cell elect one of them to read it -- one 128-bit load per (cell, channel, plane)
  // instead of up to four 32-bit loads that each cost a full DRAM sector.
  const long cell = idx >> 2;
  const unsigned peers = __match_any_sync(0xffffffffu, cell);
  const int leader = __ffs(peers) - 1;
  const bool aligned = ((reinterpret_cast<uintptr_t>(in) | (static_cast<unsigned long>(W) << 2) |
                         (static_cast<unsigned long>(plane_words) << 2)) & 15) == 0;
  int re = 0, im = 0;
  if (aligned) {
    int4 r4 = make_int4(0, 0, 0, 0), i4 = make_int4(0, 0, 0, 0);
    if ((threadIdx.x & 31) == leader) {
      const int32_t* p = in + static_cast<long>(co) * nchan_sum * W + cell * 4;
      for (int k = 0; k < nchan_sum; ++k) {
        const int4 a = __ldg(reinterpret_cast<const int4*>(p + k * W));
        const int4 b = __ldg(reinterpret_cast<const int4*>(p + plane_words + k * W));
        r4.x += a.x; r4.y += a.y; r4.z += a.z; r4.w += a.w;
        i4.x += b.x; i4.y += b.y; i4.z += b.z; i4.w += b.w;
      }
    }
    r4.x = __shfl_sync(0xffffffffu, r4.x, leader); r4.y = __shfl_sync(0xffffffffu, r4.y, leader);
    r4.z = __shfl_sync(0xffffffffu, r4.z, leader); r4.w = __shfl_sync(0xffffffffu, r4.w, leader);
    i4.x = __shfl_sync(0xffffffffu, i4.x, leader); i4.y = __shfl_sync(0xffffffffu, i4.y, leader);
    i4.z = __shfl_sync(0xffffffffu, i4.z, leader); i4.w = __shfl_sync(0xffffffffu, i4.w, leader);
    const int w4 = static_cast<int>(idx & 3);
    re = w4 == 0 ? r4.x : w4 == 1 ? r4.y : w4 == 2 ? r4.z : r4.w;
    im = w4 == 0 ? i4.x : w4 == 1 ? i4.y : w4 == 2 ? i4.z : i4.w;
  } else {
    const int32_t* p = in + static_cast<long>(co) * nchan_sum * W + idx;
    for (int k = 0; k < nchan_sum; ++k) {
      re += __ldg(p + k * W);
      im += __ldg(p + plane_words + k * W);
    }
  }
  if (!live) return;
  if (conj[v]) im = -im;
  if (LAYOUT == 0) out[static_cast<long>(co) * nvis + v] = make_int2(re, im);
  if (LAYOUT == 1) out[(static_cast<long>(v >> 2) * nchan_out + co) * 4 + (v & 3)] = make_int2(re, im);
  if (LAYOUT == 2) out[static_cast<long>(v) * nchan_out + co] = make_int2(__byte_perm(re, 0, 0x0123), __byte_perm(im, 0, 0x0123));
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


// ---------------------------------------------------------------------------------------------
// Tiled reorder for npol == 2.  One block = one second stand s1, 32 first stands s0 of one parity
// (s0 = par + 2 (32 jc + k) <= s1) and 32 channels.  With the usual antenna map those 32 baselines are 32
// neighbouring 16-byte cells of one xGPU register-tile row, so the gathers of a warp cover whole sectors; the
// tile is turned through shared memory so that the writes run along channel.  Still table-driven (a2bl / conj),
// so any antenna map gives the same answer.
//   RO_FULL: out[s0][s1][p0][p1][chan][2] and the mirrored entry [s1][s0][...] (bfXgpuReorder,
//            corr_output_full_block.py:669, shape :392)
//   RO_COR : out[bl][chan][p0][p1][2], bl = upper-triangle baseline number, s0-major -- the "COR" packet
//            payload send_packets_bf() builds with transpose([0,3,1,2,4]) (corr_output_full_block.py:505-511)
//   RO_PY  : out[bl][p0][p1][chan][2] -- the payload of send_packets_py() (corr_output_full_block.py:457-460)
enum { RO_FULL = 0, RO_COR = 1, RO_PY = 2 };
constexpr int RO_ROW = 4 * 32 + 1;   // int2 per k row: 4 pol products x 32 channels, +1 keeps STS.64 conflict-free

constexpr int RO_TY = 8;              // warps per block; each walks 32 / RO_TY rows of the 32 x 32 tile
constexpr int RO_R = 32 / RO_TY;

template <int MODE>
__global__ void __launch_bounds__(32 * RO_TY) reorder_tile_kernel(const int32_t* __restrict__ in, int2* __restrict__ out,
                                                                  const int32_t* __restrict__ a2bl,
                                                                  const int32_t* __restrict__ conj, int nstand, int nchan,
                                                                  long W, long plane_words, int nchunk) {
  __shared__ int2 tile[32 * RO_ROW];
  __shared__ int tab[32 * 4];                              // word index | conj << 31 of the block's 32 x 4 entries
  __shared__ int tab2[32 * 4];                             // RO_FULL: tile slot | flip << 31 of the mirrored entries
  // blockIdx.x walks the (parity, chunk) groups; group g = 2 jc + par holds the second stands s1 >= its first s0
  int s0_first = 0, s1 = blockIdx.x;
  for (int g = 0; g < 2 * nchunk; ++g) {
    s0_first = (g & 1) + 64 * (g >> 1);
    const int n = nstand - s0_first;
    if (s1 < n) break;
    s1 -= n;
  }
  s1 += s0_first;
  const int c0 = blockIdx.y * 32;
#pragma unroll 1
  for (int mirror = 0; mirror < (MODE == RO_FULL ? 2 : 1); ++mirror) {
    if (MODE == RO_FULL && mirror) {
      // The mirrored entry [s1][s0][p][q] normally names a word the first pass already gathered (the same product,
      // conjugated).  Look it up among the 4 words of this stand pair instead of gathering again; a map for which
      // that fails anywhere in the block takes the general path below.
      int ok = 1;
      if (threadIdx.y < 4) {
        const int s0 = s0_first + 2 * threadIdx.x;
        if (s0 <= s1) {
          const long e = (static_cast<long>(s1) * nstand + s0) * 4 + threadIdx.y;
          const int wm = __ldg(a2bl + e), cm = __ldg(conj + e) ? 1 : 0;
          int found = -1;
#pragma unroll
          for (int pp = 0; pp < 4; ++pp) {
            const int t1 = tab[threadIdx.x * 4 + pp];
            if ((t1 & 0x7fffffff) == wm) found = pp | ((cm != ((t1 >> 31) & 1)) ? 0x80000000 : 0);
          }
          ok = found != -1;
          tab2[threadIdx.x * 4 + threadIdx.y] = found;
        }
      }
      if (__syncthreads_and(ok)) {
        const int c = c0 + threadIdx.x;
#pragma unroll
        for (int r = 0; r < RO_R; ++r) {
          const int ky = threadIdx.y + RO_TY * r;
          const int s0 = s0_first + 2 * ky;
          if (s0 < s1 && c < nchan) {
            const int2* t = tile + ky * RO_ROW + threadIdx.x;
            const long e = (static_cast<long>(s1) * nstand + s0) * 4;
#pragma unroll
            for (int pp = 0; pp < 4; ++pp) {
              const int m = tab2[ky * 4 + pp];
              int2 v = t[(m & 3) * 32];
              if (m < 0) v.y = -v.y;
              __stcs(out + (e + pp) * nchan + c, v);
            }
          }
        }
        break;
      }
    }
    if (mirror) __syncthreads();
    if (threadIdx.y < 4) {  // table entries once per block, not once per channel
      const int s0 = s0_first + 2 * threadIdx.x;
      if (s0 <= s1) {
        const long e = ((mirror ? static_cast<long>(s1) * nstand + s0 : static_cast<long>(s0) * nstand + s1)) * 4 +
                       threadIdx.y;
        tab[threadIdx.x * 4 + threadIdx.y] = __ldg(a2bl + e) | (__ldg(conj + e) ? 0x80000000 : 0);
      }
    }
    __syncthreads();
    {  // gather: lanes along the 32 first stands, warps along channel; all 8 * RO_R loads of a thread are in flight
       // before the first one is used
      const int s0 = s0_first + 2 * threadIdx.x;
      if (s0 <= s1) {
        const int4 t4 = *reinterpret_cast<const int4*>(tab + threadIdx.x * 4);
        const int tt[4] = {t4.x, t4.y, t4.z, t4.w};
        int re[RO_R][4], im[RO_R][4];
#pragma unroll
        for (int r = 0; r < RO_R; ++r) {
          const int c = c0 + threadIdx.y + RO_TY * r;
          if (c < nchan) {
            const int32_t* pre = in + static_cast<long>(c) * W;
            const int32_t* pim = pre + plane_words;
#pragma unroll
            for (int pp = 0; pp < 4; ++pp) {
              re[r][pp] = __ldg(pre + (tt[pp] & 0x7fffffff));
              im[r][pp] = __ldg(pim + (tt[pp] & 0x7fffffff));
            }
          }
        }
#pragma unroll
        for (int r = 0; r < RO_R; ++r) {
          const int cy = threadIdx.y + RO_TY * r;
          if (c0 + cy < nchan) {
#pragma unroll
            for (int pp = 0; pp < 4; ++pp)
              tile[threadIdx.x * RO_ROW + pp * 32 + cy] = make_int2(re[r][pp], tt[pp] < 0 ? -im[r][pp] : im[r][pp]);
          }
        }
      }
    }
    __syncthreads();
    // scatter: lanes along channel, warps along the 32 first stands
    const int c = c0 + threadIdx.x;
#pragma unroll
    for (int r = 0; r < RO_R; ++r) {
      const int ky = threadIdx.y + RO_TY * r;
      const int s0 = s0_first + 2 * ky;
      if (s0 <= s1 && c < nchan && !(mirror && s0 == s1)) {
        const int2* t = tile + ky * RO_ROW + threadIdx.x;
        if (MODE == RO_FULL) {
          const long e = ((mirror ? static_cast<long>(s1) * nstand + s0 : static_cast<long>(s0) * nstand + s1)) * 4;
#pragma unroll
          for (int pp = 0; pp < 4; ++pp) __stcs(out + (e + pp) * nchan + c, t[pp * 32]);
        } else {
          const long bl = static_cast<long>(s0) * nstand - (static_cast<long>(s0) * (s0 - 1)) / 2 + (s1 - s0);
          if (MODE == RO_PY) {
#pragma unroll
            for (int pp = 0; pp < 4; ++pp) __stcs(out + (bl * 4 + pp) * nchan + c, t[pp * 32]);
          } else {
            int4* o = reinterpret_cast<int4*>(out + (bl * nchan + c) * 4);
            const int2 v0 = t[0], v1 = t[32], v2 = t[64], v3 = t[96];
            __stcs(o, make_int4(v0.x, v0.y, v1.x, v1.y));
            __stcs(o + 1, make_int4(v2.x, v2.y, v3.x, v3.y));
          }
        }
      }
    }
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
                          int nchan, int nchan_sum, long W, int layout, cudaStream_t stream) {
  const int nchan_out = nchan / nchan_sum;
  if (nvis <= 0 || nchan_out <= 0) return cudaSuccess;
  dim3 grid((nvis + 255) / 256, nchan_out);
  int2* o = reinterpret_cast<int2*>(out);
  const long plane = static_cast<long>(nchan) * W;
  switch (layout) {
    case 0: subsel_kernel<0><<<grid, 256, 0, stream>>>(in, o, vismap, conj, nvis, nchan_out, nchan_sum, W, plane); break;
    case 1: subsel_kernel<1><<<grid, 256, 0, stream>>>(in, o, vismap, conj, nvis, nchan_out, nchan_sum, W, plane); break;
    case 2: subsel_kernel<2><<<grid, 256, 0, stream>>>(in, o, vismap, conj, nvis, nchan_out, nchan_sum, W, plane); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_reorder(const int32_t* in, int32_t* out, const int32_t* a2bl, const int32_t* conj, int nstand,
                           int npol, int nchan, long W, int layout, cudaStream_t stream) {
  const long nent = static_cast<long>(nstand) * nstand * npol * npol;
  if (nent == 0 || nchan == 0) return cudaSuccess;
  if (npol != 2) {
    if (layout != 0) return cudaErrorInvalidValue;          // the triangular payload layouts are dual-pol only
    dim3 grid(static_cast<unsigned>((nent + 31) / 32), (nchan + 31) / 32);
    reorder_kernel<<<grid, dim3(32, 32), 0, stream>>>(in, reinterpret_cast<int2*>(out), a2bl, conj, nent, nchan, W,
                                                      static_cast<long>(nchan) * W);
    return cudaGetLastError();
  }
  const int nchunk = ((nstand + 1) / 2 + 31) / 32;          // 32-stand chunks per parity class
  unsigned nblk = 0;                                        // (s1, parity, chunk) triples on or below the diagonal
  for (int g = 0; g < 2 * nchunk; ++g) {
    const int n = nstand - ((g & 1) + 64 * (g >> 1));
    if (n > 0) nblk += n;
  }
  dim3 grid(nblk, (nchan + 31) / 32);
  int2* o = reinterpret_cast<int2*>(out);
  const long plane = static_cast<long>(nchan) * W;
  switch (layout) {
    case RO_FULL: reorder_tile_kernel<RO_FULL><<<grid, dim3(32, RO_TY), 0, stream>>>(in, o, a2bl, conj, nstand, nchan, W, plane, nchunk); break;
    case RO_COR: reorder_tile_kernel<RO_COR><<<grid, dim3(32, RO_TY), 0, stream>>>(in, o, a2bl, conj, nstand, nchan, W, plane, nchunk); break;
    case RO_PY: reorder_tile_kernel<RO_PY><<<grid, dim3(32, RO_TY), 0, stream>>>(in, o, a2bl, conj, nstand, nchan, W, plane, nchunk); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

}  // namespace xeng
