// Corr: the LWA-352 X-engine cross-multiply-accumulate for sm_100a.
//
// Replaces bfXgpuKernel (call site /root/reference/pipeline/lwa352_pipeline/blocks/corr_block.py:445).
// For every channel c and input pair (in0, in1) with station(in1) >= station(in0) it forms
//     X[c][in0,in1] = sum_t conj(v[t,c,in0]) * v[t,c,in1]
// and writes it at xGPU's register-tile index (corr_block.py:37-58), real plane first, imaginary plane
// `matlen` words later (corr_block.py:292-293, xgpu_test.py:94).
//
// Design (one persistent CTA per SM, warp-specialised):
//   * output tile = 128 "row" inputs (in1) x 128 "column" inputs (in0) of one channel; only tiles on or
//     below the block diagonal are computed.
//   * producer warps read the 4+4-bit samples straight from the [t][chan][input] gulp with 128-bit
//     loads, sign-extend the nibbles with SWAR arithmetic and store int8 operand panels into shared
//     memory in tcgen05's MN-major no-swizzle canonical layout -- the gulp's input-contiguous layout IS
//     MN-major for a contraction over time, so no transpose is ever done.
//   * K stacks (re, im):  A row  = [ r1 | i1 ],  B rows = [ r0 | i0 ] (real part) and [ -i0 | r0 ]
//     (imaginary part), so ONE tcgen05.mma.kind::i8 of shape M=128, N=256 per 16 time samples yields
//     Re in TMEM columns 0..127 and Im in columns 128..255, int32 accumulators.
//   * the accumulator is double-buffered in TMEM (2 x 256 columns); epilogue warps drain one tile while
//     the MMA warp fills the other.  The epilogue lays the tile out in shared memory as the contiguous
//     runs of xGPU's triangular order and hands each run to the bulk-copy engine: a plain bulk store,
//     or cp.reduce.async.bulk .add.s32 when accumulating into an existing integration (this is how the
//     long-term CorrAcc integration is fused: the read-modify-write happens in L2, not in the SMs).
#include "xeng_ptx.cuh"
#include "xeng_internal.h"

namespace xeng {

namespace {

constexpr int TILE = 128;        // inputs per tile edge (= 64 stations = 32 station pairs)
constexpr int PAIRS = TILE / 4;  // station pairs per tile edge
constexpr int ST_T = 32;         // time samples per pipeline stage (2 MMA K-steps of 16 samples)
constexpr int NSTAGE = 3;
constexpr int KGRP_BYTES = 128;                            // one core matrix: 8 k-rows x 16 B
constexpr int SBO = (2 * ST_T / 8) * KGRP_BYTES + 16;      // 16-input group stride, +16 B: conflict-free STS
constexpr int PANEL = (TILE / 16) * SBO;                   // 128 inputs x 64 k-rows
constexpr int STAGE_BYTES = 3 * PANEL;                     // A | B_re | B_im  (B_re,B_im contiguous: N = 256)
constexpr int RUN_BYTES = PAIRS * 16;                      // one output run: 32 cells x 4 words
constexpr int RUN_STRIDE = RUN_BYTES + 16;                 // padded: conflict-free STS.64
constexpr int PLANE_RUNS = 4 * PAIRS;                      // 4 quadrants x 32 row pairs
constexpr int STAGING_BYTES = PLANE_RUNS * RUN_STRIDE;     // one plane (re or im) of a tile
constexpr int N_EPI_WARPS = 4, N_PROD_WARPS = 8;
constexpr int N_THREADS = (N_EPI_WARPS + N_PROD_WARPS + 1) * 32;
constexpr int N_PROD = N_PROD_WARPS * 32;
constexpr int TMEM_COLS = 512;

struct SmemLayout {
  uint64_t full[NSTAGE];
  uint64_t empty[NSTAGE];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
  uint32_t pad[3];
};
constexpr int SMEM_HDR = 128;
static_assert(sizeof(SmemLayout) <= SMEM_HDR, "header");
constexpr int SMEM_BYTES = SMEM_HDR + NSTAGE * STAGE_BYTES + 2 * STAGING_BYTES + 128;

constexpr uint32_t IDESC = umma_idesc(/*S32*/ 2, /*INT8*/ 1, /*INT8*/ 1, /*A MN-major*/ 1, /*B MN-major*/ 1, 128, 256);

struct Tile {
  int c, I, J;
};
__device__ __forceinline__ Tile decode_tile(int tile, int tiles_per_chan) {
  Tile t;
  t.c = tile / tiles_per_chan;
  int r = tile - t.c * tiles_per_chan;
  int I = static_cast<int>((sqrtf(8.f * r + 1.f) - 1.f) * 0.5f);
  while ((I + 1) * (I + 2) / 2 <= r) ++I;
  while (I * (I + 1) / 2 > r) --I;
  t.I = I;
  t.J = r - I * (I + 1) / 2;
  return t;
}

// 16 consecutive inputs of one (t, chan) row; zero outside [0, ninput) and outside [0, T).
__device__ __forceinline__ uint4 load_raw16(const uint8_t* __restrict__ in, long row_off, int n, int ninput, bool tvalid,
                                            bool vec_ok) {
  uint4 z = make_uint4(0, 0, 0, 0);
  if (!tvalid || n >= ninput) return z;
  if (vec_ok && n + 16 <= ninput) return ldg_nc_v4(in + row_off + n);
  uint32_t w[4] = {0, 0, 0, 0};
  for (int k = 0; k < 16; ++k)
    if (n + k < ninput) w[k >> 2] |= static_cast<uint32_t>(in[row_off + n + k]) << (8 * (k & 3));
  return make_uint4(w[0], w[1], w[2], w[3]);
}

__global__ void __launch_bounds__(N_THREADS, 1) corr_tc_kernel(CorrArgs a) {
  extern __shared__ uint8_t smem_raw[];
  // 128-B align the carve-up (descriptor start addresses need 16 B; keep it generous)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  SmemLayout* S = reinterpret_cast<SmemLayout*>(smem);
  uint8_t* stages = smem + SMEM_HDR;
  uint8_t* staging = stages + NSTAGE * STAGE_BYTES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_kstage = (a.ntime + ST_T - 1) / ST_T;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NSTAGE; ++s) {
      mbar_init(&S->full[s], N_PROD);
      mbar_init(&S->empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&S->tmem_full[b], 1);
      mbar_init(&S->tmem_empty[b], N_EPI_WARPS * 32);
    }
    mbar_fence_init();
  }
  if (warp == N_EPI_WARPS + N_PROD_WARPS) {
    tmem_alloc(&S->tmem_base, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = S->tmem_base;

  if (warp < N_EPI_WARPS) {
    // ===================================================================== epilogue
    const int m = threadIdx.x;  // TMEM lane == row of the tile == in1 - 128*I
    const int A1loc = m >> 2, b1 = (m >> 1) & 1, p1 = m & 1;
    const long plane_words = static_cast<long>(a.nchan) * a.words_per_chan;
    int it = 0;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
      const Tile tl = decode_tile(tile, a.tiles_per_chan);
      const int buf = it & 1;
      mbar_wait(&S->tmem_full[buf], (it >> 1) & 1, a.err, 100 + buf);
      tc_fence_after();
      const bool diag = (tl.I == tl.J);
      const int A1 = tl.I * PAIRS + A1loc;       // global station-pair index of this thread's row
      const int A0base = tl.J * PAIRS;
      // run handled by this thread when it issues bulk copies: quadrant q = m / 32, row pair m % 32
      const int rq = m >> 5, rA1loc = m & 31, rA1 = tl.I * PAIRS + rA1loc;
      int rcells = diag ? (rA1loc + 1) : PAIRS;
      rcells = min(rcells, a.npairs - A0base);
      if (rA1 >= a.npairs) rcells = 0;
      const long rword = (static_cast<long>(rq) * a.quad_cells + (static_cast<long>(rA1) * (rA1 + 1)) / 2 + A0base) * 4;
#pragma unroll 1
      for (int plane = 0; plane < 2; ++plane) {
        uint8_t* stg = staging + plane * STAGING_BYTES;
        // the bulk copies issued two chunks ago read this staging buffer: drain them (per thread), then sync
        bulk_wait_read<1>();
        named_bar_sync(1, N_EPI_WARPS * 32);
#pragma unroll 1
        for (int cc = 0; cc < 4; ++cc) {
          uint32_t r[32];
          const uint32_t taddr = tmem_base + (static_cast<uint32_t>(warp * 32) << 16) +
                                 static_cast<uint32_t>(buf * 256 + plane * 128 + cc * 32);
          tmem_ld_32x32b_x32(taddr, r);
          tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            // columns j = cc*32 + 2k, 2k+1  ->  in0 local = 4*a0 + 2*b0 + p0
            const int a0 = cc * 8 + (k >> 1), b0 = k & 1;
            uint32_t v0 = r[2 * k], v1 = r[2 * k + 1];
            if (diag && b0 == 1 && b1 == 0 && a0 == A1loc) { v0 = 0; v1 = 0; }  // unused xGPU padding cell
            const int run = (2 * b0 + b1) * PAIRS + A1loc;
            sts_v2(smem_u32(stg + run * RUN_STRIDE + a0 * 16 + p1 * 8), v0, v1);
          }
        }
        if (plane == 1) {  // accumulator fully read: hand the TMEM buffer back to the MMA warp
          tc_fence_before();
          mbar_arrive(&S->tmem_empty[buf]);
        }
        fence_proxy_async_smem();
        named_bar_sync(1, N_EPI_WARPS * 32);
        if (rcells > 0) {
          int32_t* dst = a.out + plane * plane_words + static_cast<long>(tl.c) * a.words_per_chan + rword;
          const uint32_t src = smem_u32(stg + m * RUN_STRIDE);
          if (a.accumulate) bulk_reduce_add_s32(dst, src, rcells * 16);
          else bulk_store(dst, src, rcells * 16);
        }
        bulk_commit();
      }
    }
    bulk_wait<0>();
  } else if (warp < N_EPI_WARPS + N_PROD_WARPS) {
    // ===================================================================== producers (unpack)
    const int pw = warp - N_EPI_WARPS;
    const int g = lane & 7;                 // 16-input group inside the 128-input panel
    const int tloc = pw * 4 + (lane >> 3);  // time sample inside the stage, 0..31
    const uint32_t sts_off = g * SBO + ((tloc >> 3) * 2) * KGRP_BYTES + (tloc & 7) * 16;
    const long row_bytes = static_cast<long>(a.ninput);
    const bool vec_ok = (a.ninput % 16 == 0) && ((reinterpret_cast<uintptr_t>(a.in) & 15) == 0);
    int slot = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
      const Tile tl = decode_tile(tile, a.tiles_per_chan);
      const int nA = tl.I * TILE + g * 16, nB = tl.J * TILE + g * 16;
      uint4 rawA, rawB;
      {
        const int t = tloc;
        const long off = (static_cast<long>(t) * a.nchan + tl.c) * row_bytes;
        rawA = load_raw16(a.in, off, nA, a.ninput, t < a.ntime, vec_ok);
        rawB = load_raw16(a.in, off, nB, a.ninput, t < a.ntime, vec_ok);
      }
      for (int ks = 0; ks < n_kstage; ++ks) {
        // prefetch the next stage of this tile while the current one is unpacked
        uint4 nxtA = make_uint4(0, 0, 0, 0), nxtB = nxtA;
        if (ks + 1 < n_kstage) {
          const int t = (ks + 1) * ST_T + tloc;
          const long off = (static_cast<long>(t) * a.nchan + tl.c) * row_bytes;
          nxtA = load_raw16(a.in, off, nA, a.ninput, t < a.ntime, vec_ok);
          nxtB = load_raw16(a.in, off, nB, a.ninput, t < a.ntime, vec_ok);
        }
        mbar_wait(&S->empty[slot], phase ^ 1, a.err, 200 + slot);
        const uint32_t base = smem_u32(stages + slot * STAGE_BYTES) + sts_off;
        uint4 re, im;
        // A panel: rows (in1):  k-group 2h = re, 2h+1 = im
        re = make_uint4(unpack_re4(rawA.x), unpack_re4(rawA.y), unpack_re4(rawA.z), unpack_re4(rawA.w));
        im = make_uint4(unpack_im4(rawA.x), unpack_im4(rawA.y), unpack_im4(rawA.z), unpack_im4(rawA.w));
        sts_v4(base, re);
        sts_v4(base + KGRP_BYTES, im);
        // B panels: columns (in0):  real part [ re | im ], imaginary part [ -im | re ]
        re = make_uint4(unpack_re4(rawB.x), unpack_re4(rawB.y), unpack_re4(rawB.z), unpack_re4(rawB.w));
        im = make_uint4(unpack_im4(rawB.x), unpack_im4(rawB.y), unpack_im4(rawB.z), unpack_im4(rawB.w));
        sts_v4(base + PANEL, re);
        sts_v4(base + PANEL + KGRP_BYTES, im);
        sts_v4(base + 2 * PANEL, make_uint4(neg4(im.x), neg4(im.y), neg4(im.z), neg4(im.w)));
        sts_v4(base + 2 * PANEL + KGRP_BYTES, re);
        fence_proxy_async_smem();
        mbar_arrive(&S->full[slot]);
        rawA = nxtA;
        rawB = nxtB;
        if (++slot == NSTAGE) { slot = 0; phase ^= 1; }
      }
    }
  } else {
    // ===================================================================== MMA issuer (one lane)
    if (lane == 0) {
      int slot = 0, it = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        mbar_wait(&S->tmem_empty[buf], ((it >> 1) & 1) ^ 1, a.err, 300 + buf);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + buf * 256;
        for (int ks = 0; ks < n_kstage; ++ks) {
          mbar_wait(&S->full[slot], phase, a.err, 400 + slot);
          tc_fence_after();
          const uint32_t sbase = smem_u32(stages + slot * STAGE_BYTES);
          const int trem = a.ntime - ks * ST_T;
          const int nk = trem >= ST_T ? ST_T / 16 : (trem + 15) / 16;
          for (int k = 0; k < nk; ++k) {
            const uint64_t ad = umma_smem_desc(sbase + k * 4 * KGRP_BYTES, KGRP_BYTES, SBO);
            const uint64_t bd = umma_smem_desc(sbase + PANEL + k * 4 * KGRP_BYTES, KGRP_BYTES, SBO);
            umma_i8_ss(tmem_d, ad, bd, IDESC, (ks | k) != 0);
          }
          umma_commit(&S->empty[slot]);
          if (ks == n_kstage - 1) umma_commit(&S->tmem_full[buf]);
          if (++slot == NSTAGE) { slot = 0; phase ^= 1; }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == N_EPI_WARPS + N_PROD_WARPS) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------------------------------------
// Plain CUDA-core correlator (one thread per visibility word pair).  Bring-up / cross-check kernel:
// exercised by the GPU tests against the tensor-core kernel at sizes the CPU oracle cannot reach.
__global__ void corr_simt_kernel(CorrArgs a) {
  const long ncell = static_cast<long>(a.npairs) * (a.npairs + 1) / 2;
  const long total = ncell * 16 * a.nchan;
  const long plane_words = static_cast<long>(a.nchan) * a.words_per_chan;
  for (long idx = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long>(gridDim.x) * blockDim.x) {
    const int sub = idx & 15;
    long rest = idx >> 4;
    const long cell = rest % ncell;
    const int c = static_cast<int>(rest / ncell);
    int A1 = static_cast<int>((sqrt(8.0 * cell + 1.0) - 1.0) * 0.5);
    while (static_cast<long>(A1 + 1) * (A1 + 2) / 2 <= cell) ++A1;
    while (static_cast<long>(A1) * (A1 + 1) / 2 > cell) --A1;
    const int A0 = static_cast<int>(cell - static_cast<long>(A1) * (A1 + 1) / 2);
    const int q = sub >> 2, p1 = (sub >> 1) & 1, p0 = sub & 1;
    const int b0 = q >> 1, b1 = q & 1;
    const int in0 = 4 * A0 + 2 * b0 + p0, in1 = 4 * A1 + 2 * b1 + p1;
    int re = 0, im = 0;
    if (!(A0 == A1 && b0 == 1 && b1 == 0)) {
      for (int t = 0; t < a.ntime; ++t) {
        const uint8_t* row = a.in + (static_cast<long>(t) * a.nchan + c) * a.ninput;
        const int x0 = row[in0], x1 = row[in1];
        const int r0 = static_cast<int8_t>(x0) >> 4, i0 = static_cast<int8_t>(x0 << 4) >> 4;
        const int r1 = static_cast<int8_t>(x1) >> 4, i1 = static_cast<int8_t>(x1 << 4) >> 4;
        re += r0 * r1 + i0 * i1;
        im += r0 * i1 - i0 * r1;
      }
    }
    const long w = static_cast<long>(c) * a.words_per_chan + (static_cast<long>(q) * a.quad_cells + cell) * 4 + 2 * p1 + p0;
    if (a.accumulate) {
      a.out[w] += re;
      a.out[plane_words + w] += im;
    } else {
      a.out[w] = re;
      a.out[plane_words + w] = im;
    }
  }
}

}  // namespace

int corr_tc_smem_bytes() { return SMEM_BYTES; }

cudaError_t launch_corr_tc(const CorrArgs& a, int nsm, cudaStream_t stream) {
  cudaError_t ce = configure_once_per_device([]() -> cudaError_t {
    return cudaFuncSetAttribute(corr_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
  }, kCfgCorrTc);
  if (ce != cudaSuccess) return ce;
  int grid = a.ntiles < nsm ? a.ntiles : nsm;
  if (grid <= 0) return cudaSuccess;
  corr_tc_kernel<<<grid, N_THREADS, SMEM_BYTES, stream>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_corr_simt(const CorrArgs& a, cudaStream_t stream) {
  const long ncell = static_cast<long>(a.npairs) * (a.npairs + 1) / 2;
  const long total = ncell * 16 * a.nchan;
  if (total <= 0) return cudaSuccess;
  long blocks = (total + 255) / 256;
  if (blocks > 148L * 32) blocks = 148L * 32;
  corr_simt_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace xeng
