// Corr, TMA-fed variant (the production kernel when ninput % 16 == 0 and the gulp is 16-byte aligned).
//
// Same contraction, tiling and output order as xeng_corr.cu (see the header there), with a different
// data path into the tensor cores:
//   TMA warp      : cp.async.bulk.tensor 3-D boxes {128 inputs, 1 channel, 32 samples} of raw 4+4-bit
//                   bytes into a deep shared-memory ring (RAW_STAGES x 8 KB) -- global-load latency is
//                   hidden by the ring depth instead of by per-thread register prefetch.
//   unpack warps  : LDS.128 raw bytes -> SWAR nibble sign-extension -> int8 operand panels in tcgen05's
//                   MN-major no-swizzle layout (A = [re|im] rows, B = [re|im] and [-im|re] columns).
//   MMA warp      : one lane issues tcgen05.mma.kind::i8 M=128 N=256 K=32, int32 accumulators in TMEM,
//                   double-buffered (2 x 256 columns).
//   epilogue warps: tcgen05.ld -> shared-memory staging laid out as runs of xGPU's triangular order ->
//                   bulk store, or bulk reduce-add (.s32) for the fused CorrAcc accumulation.
// B columns are ordered n = b0*64 + a0*2 + p0 (b0 = station parity inside the station pair) so that one
// 64-column slice of the accumulator is exactly two xGPU quadrants: staging is 33 KB per slice.
// Diagonal tiles load their raw panel once.
#include <cuda.h>

#include "xeng_ptx.cuh"
#include "xeng_internal.h"

namespace xeng {
namespace {

constexpr int TILE = 128;
constexpr int PAIRS = TILE / 4;
constexpr int ST_T = 32;                                   // samples per stage (2 MMA K-steps)
constexpr int NSTAGE = 3;                                  // operand ring depth
constexpr int RAW_STAGES = 10;                             // raw ring depth
constexpr int RAW_PANEL = ST_T * TILE;                     // 4 KB: [32 samples][128 inputs]
constexpr int RAW_STAGE_BYTES = 2 * RAW_PANEL;             // A | B
constexpr int KGRP_BYTES = 128;
constexpr int SBO = (2 * ST_T / 8) * KGRP_BYTES + 16;      // 1040: +16 B keeps the STS.128 conflict-free
constexpr int PANEL = (TILE / 16) * SBO;
constexpr int STAGE_BYTES = 3 * PANEL;                     // A | B_re | B_im
constexpr int RUN_BYTES = PAIRS * 16;                      // 32 cells x 4 words
constexpr int RUN_STRIDE = RUN_BYTES + 16;
constexpr int SLICE_RUNS = 2 * PAIRS;                      // 2 quadrants x 32 row pairs
constexpr int STAGING_BYTES = SLICE_RUNS * RUN_STRIDE;     // 33 792
constexpr int N_EPI_WARPS = 4, N_PROD_WARPS = 8;
constexpr int WARP_TMA = N_EPI_WARPS + N_PROD_WARPS, WARP_MMA = WARP_TMA + 1;
constexpr int N_THREADS = (WARP_MMA + 1) * 32;
constexpr int N_PROD = N_PROD_WARPS * 32;
constexpr int TMEM_COLS = 512;

struct SmemHdr {
  uint64_t raw_full[RAW_STAGES];
  uint64_t raw_empty[RAW_STAGES];
  uint64_t full[NSTAGE];
  uint64_t empty[NSTAGE];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};
constexpr int SMEM_HDR = 256;
static_assert(sizeof(SmemHdr) <= SMEM_HDR, "header");
constexpr int SMEM_BYTES = SMEM_HDR + RAW_STAGES * RAW_STAGE_BYTES + NSTAGE * STAGE_BYTES + 2 * STAGING_BYTES + 128;
static_assert(SMEM_BYTES <= 232448, "shared memory budget");

constexpr uint32_t IDESC = umma_idesc(2, 1, 1, 1, 1, 128, 256);

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

__global__ void __launch_bounds__(N_THREADS, 1) corr_tma_kernel(const __grid_constant__ CUtensorMap tmap, CorrArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  SmemHdr* S = reinterpret_cast<SmemHdr*>(smem);
  uint8_t* raw = smem + SMEM_HDR;
  uint8_t* stages = raw + RAW_STAGES * RAW_STAGE_BYTES;
  uint8_t* staging = stages + NSTAGE * STAGE_BYTES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_kstage = (a.ntime + ST_T - 1) / ST_T;

  if (threadIdx.x == 0) {
    for (int s = 0; s < RAW_STAGES; ++s) {
      mbar_init(&S->raw_full[s], 1);
      mbar_init(&S->raw_empty[s], N_PROD);
    }
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
  if (warp == WARP_MMA) {
    tmem_alloc(&S->tmem_base, TMEM_COLS);
    tmem_relinquish();
  }
  if (warp == WARP_TMA && lane == 0) tma_prefetch_desc(&tmap);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = S->tmem_base;

  if (warp < N_EPI_WARPS) {
    // ===================================================================== epilogue
    const int m = threadIdx.x;  // TMEM lane == tile row == in1 - 128*I = 4*A1loc + 2*b1 + p1
    const int A1loc = m >> 2, b1 = (m >> 1) & 1, p1 = m & 1;
    const long plane_words = static_cast<long>(a.nchan) * a.words_per_chan;
    const uint32_t my_run = smem_u32(staging) + (b1 * PAIRS + A1loc) * RUN_STRIDE + p1 * 8;
    // bulk-copy issue role (threads 0..63): run (quadrant bit b1 = m >> 5, row pair m & 31) of the slice
    const int rb1 = (m >> 5) & 1, rA1loc = m & 31;
    int it = 0;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
      const Tile tl = decode_tile(tile, a.tiles_per_chan);
      const int buf = it & 1;
      mbar_wait(&S->tmem_full[buf], (it >> 1) & 1, a.err, 100 + buf);
      tc_fence_after();
      const bool diag = (tl.I == tl.J);
      const int A0base = tl.J * PAIRS;
      const int rA1 = tl.I * PAIRS + rA1loc;
      int rcells = diag ? (rA1loc + 1) : PAIRS;
      rcells = min(rcells, a.npairs - A0base);
      if (rA1 >= a.npairs || m >= SLICE_RUNS) rcells = 0;
      const long rcell0 = (static_cast<long>(rA1) * (rA1 + 1)) / 2 + A0base;
#pragma unroll 1
      for (int sl = 0; sl < 4; ++sl) {        // slice = (plane, b0): 64 accumulator columns
        const int plane = sl >> 1, b0 = sl & 1;
        const uint32_t stg_off = (sl & 1) * STAGING_BYTES;
        // the copies issued two slices ago read this staging buffer: drain them, then sync the warps
        if (m < SLICE_RUNS) bulk_wait_read<1>();
        named_bar_sync(1, N_EPI_WARPS * 32);
        uint32_t r0[32], r1[32];
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(warp * 32) << 16) +
                               static_cast<uint32_t>(buf * 256 + plane * 128 + b0 * 64);
        tmem_ld_32x32b_x32(taddr, r0);
        tmem_ld_32x32b_x32(taddr + 32, r1);
        tmem_ld_wait();
        if (sl == 3) {  // accumulator fully read: hand the TMEM buffer back to the MMA warp
          tc_fence_before();
          mbar_arrive(&S->tmem_empty[buf]);
        }
        const bool pad = diag && b0 == 1 && b1 == 0;   // unused xGPU padding cell (a0 == a1 swapped order)
#pragma unroll
        for (int a0 = 0; a0 < 16; ++a0) {
          uint32_t v0 = r0[2 * a0], v1 = r0[2 * a0 + 1];
          if (pad && a0 == A1loc) { v0 = 0; v1 = 0; }
          sts_v2(my_run + stg_off + a0 * 16, v0, v1);
        }
#pragma unroll
        for (int a0 = 0; a0 < 16; ++a0) {
          uint32_t v0 = r1[2 * a0], v1 = r1[2 * a0 + 1];
          if (pad && a0 + 16 == A1loc) { v0 = 0; v1 = 0; }
          sts_v2(my_run + stg_off + (a0 + 16) * 16, v0, v1);
        }
        fence_proxy_async_smem();
        named_bar_sync(1, N_EPI_WARPS * 32);
        if (m < SLICE_RUNS) {
          if (rcells > 0) {
            const long word = ((static_cast<long>(2 * b0 + rb1) * a.quad_cells + rcell0) << 2);
            int32_t* dst = a.out + plane * plane_words + static_cast<long>(tl.c) * a.words_per_chan + word;
            const uint32_t src = smem_u32(staging) + stg_off + m * RUN_STRIDE;
            if (a.accumulate) bulk_reduce_add_s32(dst, src, rcells * 16);
            else bulk_store(dst, src, rcells * 16);
          }
          bulk_commit();
        }
      }
    }
    bulk_wait<0>();
  } else if (warp < WARP_TMA) {
    // ===================================================================== unpack: raw ring -> operand ring
    const int pw = warp - N_EPI_WARPS;
    const int g = lane & 7;                 // 16-row group of the A panel == 16-column group of the B panel
    const int tloc = pw * 4 + (lane >> 3);  // sample inside the stage, 0..31
    const uint32_t sts_off = g * SBO + ((tloc >> 3) * 2) * KGRP_BYTES + (tloc & 7) * 16;
    const uint32_t ldsA_off = tloc * TILE + g * 16;
    const uint32_t ldsB_off = tloc * TILE + (g & 3) * 32;   // 8 station pairs = 32 raw bytes
    const uint32_t bsel = (g & 4) ? 0x7632u : 0x5410u;      // b0 = 1: high half-words, b0 = 0: low half-words
    int rs = 0, slot = 0;
    uint32_t rphase = 0, phase = 0;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
      const Tile tl = decode_tile(tile, a.tiles_per_chan);
      const uint32_t boff = (tl.I == tl.J) ? 0u : static_cast<uint32_t>(RAW_PANEL);   // diagonal: B raw == A raw
      for (int ks = 0; ks < n_kstage; ++ks) {
        mbar_wait(&S->raw_full[rs], rphase, a.err, 500 + rs);
        const uint32_t rbase = smem_u32(raw + rs * RAW_STAGE_BYTES);
        const uint4 rawA = lds_v4(rbase + ldsA_off);
        const uint4 b_lo = lds_v4(rbase + boff + ldsB_off);
        const uint4 b_hi = lds_v4(rbase + boff + ldsB_off + 16);
        // B columns of this group: n = a0*2 + p0 for 8 station pairs, one station parity (b0)
        const uint4 rawB = make_uint4(__byte_perm(b_lo.x, b_lo.y, bsel), __byte_perm(b_lo.z, b_lo.w, bsel),
                                      __byte_perm(b_hi.x, b_hi.y, bsel), __byte_perm(b_hi.z, b_hi.w, bsel));
        const uint4 reA = make_uint4(unpack_re4(rawA.x), unpack_re4(rawA.y), unpack_re4(rawA.z), unpack_re4(rawA.w));
        const uint4 imA = make_uint4(unpack_im4(rawA.x), unpack_im4(rawA.y), unpack_im4(rawA.z), unpack_im4(rawA.w));
        const uint4 reB = make_uint4(unpack_re4(rawB.x), unpack_re4(rawB.y), unpack_re4(rawB.z), unpack_re4(rawB.w));
        const uint4 imB = make_uint4(unpack_im4(rawB.x), unpack_im4(rawB.y), unpack_im4(rawB.z), unpack_im4(rawB.w));
        const uint4 nimB = make_uint4(neg4(imB.x), neg4(imB.y), neg4(imB.z), neg4(imB.w));
        mbar_arrive(&S->raw_empty[rs]);     // raw bytes are in registers: the TMA warp may refill the slot
        if (++rs == RAW_STAGES) { rs = 0; rphase ^= 1; }

        mbar_wait(&S->empty[slot], phase ^ 1, a.err, 200 + slot);
        const uint32_t base = smem_u32(stages + slot * STAGE_BYTES) + sts_off;
        sts_v4(base, reA);                          // A rows:     [ re | im ]
        sts_v4(base + KGRP_BYTES, imA);
        sts_v4(base + PANEL, reB);                  // B real:     [ re | im ]
        sts_v4(base + PANEL + KGRP_BYTES, imB);
        sts_v4(base + 2 * PANEL, nimB);             // B imag:     [ -im | re ]
        sts_v4(base + 2 * PANEL + KGRP_BYTES, reB);
        fence_proxy_async_smem();
        mbar_arrive(&S->full[slot]);
        if (++slot == NSTAGE) { slot = 0; phase ^= 1; }
      }
    }
  } else if (warp == WARP_TMA) {
    // ===================================================================== TMA producer (one lane)
    if (lane == 0) {
      int rs = 0;
      uint32_t rphase = 0;
      for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
        const Tile tl = decode_tile(tile, a.tiles_per_chan);
        const bool diag = (tl.I == tl.J);
        for (int ks = 0; ks < n_kstage; ++ks) {
          mbar_wait(&S->raw_empty[rs], rphase ^ 1, a.err, 600 + rs);
          const uint32_t dst = smem_u32(raw + rs * RAW_STAGE_BYTES);
          mbar_arrive_expect_tx(&S->raw_full[rs], diag ? RAW_PANEL : RAW_STAGE_BYTES);
          tma_load_3d(dst, &tmap, &S->raw_full[rs], tl.I * TILE, tl.c, ks * ST_T);
          if (!diag) tma_load_3d(dst + RAW_PANEL, &tmap, &S->raw_full[rs], tl.J * TILE, tl.c, ks * ST_T);
          if (++rs == RAW_STAGES) { rs = 0; rphase ^= 1; }
        }
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
  if (warp == WARP_MMA) tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

bool corr_tma_supported(const CorrArgs& a) {
  return (a.ninput % 16 == 0) && ((reinterpret_cast<uintptr_t>(a.in) & 15) == 0) &&
         ((reinterpret_cast<uintptr_t>(a.out) & 15) == 0);
}

cudaError_t launch_corr_tma(const CorrArgs& a, int nsm, cudaStream_t stream) {
  static EncodeTiledFn encode = nullptr;
  static bool configured = false;
  if (!configured) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess) return e;
    if (qres != cudaDriverEntryPointSuccess || !fn) return cudaErrorNotSupported;
    encode = reinterpret_cast<EncodeTiledFn>(fn);
    e = cudaFuncSetAttribute(corr_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  int grid = a.ntiles < nsm ? a.ntiles : nsm;
  if (grid <= 0) return cudaSuccess;
  alignas(64) CUtensorMap tmap;
  const cuuint64_t gdim[3] = {static_cast<cuuint64_t>(a.ninput), static_cast<cuuint64_t>(a.nchan),
                              static_cast<cuuint64_t>(a.ntime)};
  const cuuint64_t gstride[2] = {static_cast<cuuint64_t>(a.ninput),
                                 static_cast<cuuint64_t>(a.ninput) * static_cast<cuuint64_t>(a.nchan)};
  const cuuint32_t box[3] = {TILE, 1, ST_T};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t*>(a.in), gdim, gstride, box, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
  corr_tma_kernel<<<grid, N_THREADS, SMEM_BYTES, stream>>>(tmap, a);
  return cudaGetLastError();
}

}  // namespace xeng
