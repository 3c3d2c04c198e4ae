// Corr, plane-split TMA-fed variant (the production kernel when ninput % 16 == 0 and buffers are 16-byte aligned).
//
// Same contraction, tiling and output order as xeng_corr.cu / xeng_corr2.cu; what changes is how the int8 operands
// are laid out and produced, because ncu + ablation showed the kernel to be bound by shared-memory traffic and by
// the latency chain of the unpack stage (profiles/r01b_*):
//
//   * operand planes instead of interleaved [re|im] panels.  Per 32-sample stage the unpack warps write five
//     planes of 128 x 32 int8: A_re, A_im (tile rows) and B_im, B_re, B_nim = -B_im (tile columns), 20 KB instead
//     of 24 KB.  The planes are contiguous in the MN direction of the UMMA descriptor, so one N = 256 operand can
//     start at B_re and run on into B_nim, another can start at B_im and run on into B_re:
//         D[128 x 256] += A_re x [B_re | B_nim]^T      ->  [ re1 re0 | -re1 im0 ]
//         D[128 x 256] += A_im x [B_im | B_re ]^T      ->  [ im1 im0 |  im1 re0 ]
//     columns 0..127 accumulate Re, 128..255 accumulate Im of conj(v0) v1.  B_re is stored once.
//   * two unpack groups of 4 warps work on alternate stages, each thread handles two samples per stage, so the
//     raw-wait -> LDS -> ALU -> operand-slot-wait -> STS -> fence -> arrive chain of one stage overlaps the next.
//   * all shared-memory accesses of the unpack and epilogue warps are bank-conflict free (row-wise LDS.128 of the
//     raw box, 528-byte group stride for the operand STS.128, +64 B skew between the two quadrant halves of the
//     epilogue staging for the STS.64).
//
// Roles (448 threads): warps 0-3 epilogue, 4-7 unpack group 0, 8-11 unpack group 1, 12 TMA producer, 13 MMA issuer.
#include <cuda.h>

#include "xeng_ptx.cuh"
#include "xeng_internal.h"

namespace xeng {
namespace {

constexpr int TILE = 128;
constexpr int PAIRS = TILE / 4;                            // stand pairs per tile side
constexpr int ST_T = 32;                                   // samples per stage = K of one MMA
constexpr int NSTAGE = 4;                                  // operand ring depth
constexpr int RAW_STAGES = 8;                              // raw ring depth
constexpr int RAW_PANEL = ST_T * TILE;                     // 4 KB: [32 samples][128 inputs]
constexpr int RAW_STAGE_BYTES = 2 * RAW_PANEL;             // A | B
constexpr int KGRP_BYTES = 128;                            // core matrix: 8 samples x 16 inputs
constexpr int SBO = (ST_T / 8) * KGRP_BYTES + 16;          // 528: +16 B keeps the STS.128 conflict-free
constexpr int PLANE = (TILE / 16) * SBO;                   // 4224
constexpr int STAGE_BYTES = 5 * PLANE;                     // A_re | A_im | B_im | B_re | B_nim
constexpr int OFF_A_RE = 0, OFF_A_IM = PLANE, OFF_B_IM = 2 * PLANE, OFF_B_RE = 3 * PLANE, OFF_B_NIM = 4 * PLANE;
constexpr int RUN_BYTES = PAIRS * 16;                      // 32 cells x 4 words
constexpr int RUN_STRIDE = RUN_BYTES + 16;
constexpr int SLICE_RUNS = 2 * PAIRS;                      // 2 quadrants x 32 row pairs
constexpr int HALF_SKEW = 64;                              // quadrant half b1 = 1 starts 64 B later (bank skew)
constexpr int STAGING_BYTES = SLICE_RUNS * RUN_STRIDE + HALF_SKEW + 64;   // 33 920, multiple of 128
constexpr int N_EPI_WARPS = 4, N_GROUP_WARPS = 4, N_GROUPS = 2;
constexpr int WARP_TMA = N_EPI_WARPS + N_GROUPS * N_GROUP_WARPS, WARP_MMA = WARP_TMA + 1;
constexpr int N_THREADS = (WARP_MMA + 1) * 32;
constexpr int N_GROUP = N_GROUP_WARPS * 32;                // threads per unpack group
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
static_assert(STAGE_BYTES % 128 == 0 && STAGING_BYTES % 128 == 0, "alignment");
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

__device__ __forceinline__ uint4 re16(uint4 x) {
  return make_uint4(unpack_re4(x.x), unpack_re4(x.y), unpack_re4(x.z), unpack_re4(x.w));
}
__device__ __forceinline__ uint4 im16(uint4 x) {
  return make_uint4(unpack_im4(x.x), unpack_im4(x.y), unpack_im4(x.z), unpack_im4(x.w));
}
__device__ __forceinline__ uint4 neg16(uint4 x) { return make_uint4(neg4(x.x), neg4(x.y), neg4(x.z), neg4(x.w)); }

// Column group of the B operand held by this lane: 16 raw bytes = 4 stand pairs x (b0, p0); the lane with
// half == h keeps the half-words of stand parity b0 == h of its own 4 pairs and receives those of the partner
// lane (lane ^ 4), giving 8 stand pairs x 2 pols = 16 columns n = a0*2 + p0 of one parity.
__device__ __forceinline__ uint4 gather_parity(uint4 raw, int half) {
  const uint32_t lo01 = __byte_perm(raw.x, raw.y, 0x5410), lo23 = __byte_perm(raw.z, raw.w, 0x5410);
  const uint32_t hi01 = __byte_perm(raw.x, raw.y, 0x7632), hi23 = __byte_perm(raw.z, raw.w, 0x7632);
  const uint32_t send0 = half ? lo01 : hi01, send1 = half ? lo23 : hi23;
  const uint32_t recv0 = __shfl_xor_sync(0xffffffffu, send0, 4), recv1 = __shfl_xor_sync(0xffffffffu, send1, 4);
  return half ? make_uint4(recv0, recv1, hi01, hi23) : make_uint4(lo01, lo23, recv0, recv1);
}

__global__ void __launch_bounds__(N_THREADS, 1) corr_planes_kernel(const __grid_constant__ CUtensorMap tmap, CorrArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  SmemHdr* S = reinterpret_cast<SmemHdr*>(smem);
  uint8_t* raw = smem + SMEM_HDR;
  uint8_t* stages = raw + RAW_STAGES * RAW_STAGE_BYTES;
  uint8_t* staging = stages + NSTAGE * STAGE_BYTES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_kstage = (a.ntime + ST_T - 1) / ST_T;     // the TMA zero-fills samples past ntime

  if (threadIdx.x == 0) {
    for (int s = 0; s < RAW_STAGES; ++s) {
      mbar_init(&S->raw_full[s], 1);
      mbar_init(&S->raw_empty[s], N_GROUP);
    }
    for (int s = 0; s < NSTAGE; ++s) {
      mbar_init(&S->full[s], N_GROUP);
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
    const uint32_t my_run = smem_u32(staging) + (b1 * PAIRS + A1loc) * RUN_STRIDE + b1 * HALF_SKEW + p1 * 8;
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
      if (a.ablate & 1) rcells = 0;
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
        if (!(a.ablate & 2)) {
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
        }
        fence_proxy_async_smem();
        named_bar_sync(1, N_EPI_WARPS * 32);
        if (m < SLICE_RUNS) {
          if (rcells > 0) {
            const long word = ((static_cast<long>(2 * b0 + rb1) * a.quad_cells + rcell0) << 2);
            int32_t* dst = a.out + plane * plane_words + static_cast<long>(tl.c) * a.words_per_chan + word;
            const uint32_t src = smem_u32(staging) + stg_off + m * RUN_STRIDE + rb1 * HALF_SKEW;
            if (a.accumulate) bulk_reduce_add_s32(dst, src, rcells * 16);
            else bulk_store(dst, src, rcells * 16);
          }
          bulk_commit();
        }
      }
    }
    bulk_wait<0>();
  } else if (warp < WARP_TMA) {
    // ===================================================================== unpack: raw ring -> operand planes
    const int grp = (warp - N_EPI_WARPS) / N_GROUP_WARPS;          // stages with (index & 1) == grp
    const int wg = (warp - N_EPI_WARPS) % N_GROUP_WARPS;
    const int q = lane & 3, half = (lane >> 2) & 1;                  // 16-byte chunk 2q+half of the 128-input row
    const int t0 = wg * 4 + (lane >> 3), t1 = t0 + 16;               // the two samples of the stage this thread owns
    const uint32_t lds0 = t0 * TILE + (2 * q + half) * 16, lds1 = t1 * TILE + (2 * q + half) * 16;
    const uint32_t k0 = (t0 >> 3) * KGRP_BYTES + (t0 & 7) * 16, k1 = (t1 >> 3) * KGRP_BYTES + (t1 & 7) * 16;
    const uint32_t stsA = (2 * q + half) * SBO;                      // tile rows 16*(2q+half) .. +15
    const uint32_t stsB = (4 * half + q) * SBO;                      // tile columns n = half*64 + q*16 .. +15
    int gs = 0;                                                      // running stage index over all tiles
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
      const Tile tl = decode_tile(tile, a.tiles_per_chan);
      const uint32_t boff = (tl.I == tl.J) ? 0u : static_cast<uint32_t>(RAW_PANEL);   // diagonal: B raw == A raw
      for (int ks = 0; ks < n_kstage; ++ks, ++gs) {
        if ((gs & 1) != grp) continue;
        const int rs = gs % RAW_STAGES, slot = gs % NSTAGE;
        mbar_wait(&S->raw_full[rs], (gs / RAW_STAGES) & 1, a.err, 500 + rs);
        const uint32_t rbase = smem_u32(raw + rs * RAW_STAGE_BYTES);
        const uint4 rawA0 = lds_v4(rbase + lds0), rawA1 = lds_v4(rbase + lds1);
        const uint4 rawB0 = lds_v4(rbase + boff + lds0), rawB1 = lds_v4(rbase + boff + lds1);
        const uint4 colB0 = gather_parity(rawB0, half), colB1 = gather_parity(rawB1, half);
        mbar_arrive(&S->raw_empty[rs]);     // raw bytes are in registers: the TMA warp may refill the slot
        uint4 reA0, imA0, reA1, imA1, reB0, imB0, reB1, imB1, nimB0, nimB1;
        if (a.ablate & 16) {
          reA0 = imA0 = rawA0; reA1 = imA1 = rawA1; reB0 = imB0 = nimB0 = colB0; reB1 = imB1 = nimB1 = colB1;
        } else {
          reA0 = re16(rawA0); imA0 = im16(rawA0); reA1 = re16(rawA1); imA1 = im16(rawA1);
          reB0 = re16(colB0); imB0 = im16(colB0); reB1 = re16(colB1); imB1 = im16(colB1);
          nimB0 = neg16(imB0); nimB1 = neg16(imB1);
        }

        mbar_wait(&S->empty[slot], ((gs / NSTAGE) & 1) ^ 1, a.err, 200 + slot);
        if (!(a.ablate & 8)) {
          const uint32_t base = smem_u32(stages + slot * STAGE_BYTES);
          sts_v4(base + OFF_A_RE + stsA + k0, reA0);
          sts_v4(base + OFF_A_RE + stsA + k1, reA1);
          sts_v4(base + OFF_A_IM + stsA + k0, imA0);
          sts_v4(base + OFF_A_IM + stsA + k1, imA1);
          sts_v4(base + OFF_B_IM + stsB + k0, imB0);
          sts_v4(base + OFF_B_IM + stsB + k1, imB1);
          sts_v4(base + OFF_B_RE + stsB + k0, reB0);
          sts_v4(base + OFF_B_RE + stsB + k1, reB1);
          sts_v4(base + OFF_B_NIM + stsB + k0, nimB0);
          sts_v4(base + OFF_B_NIM + stsB + k1, nimB1);
        }
        if (!(a.ablate & 64)) fence_proxy_async_smem();
        mbar_arrive(&S->full[slot]);
      }
    }
  } else if (warp == WARP_TMA) {
    // ===================================================================== TMA producer (one lane)
    if (lane == 0) {
      int gs = 0;
      for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
        const Tile tl = decode_tile(tile, a.tiles_per_chan);
        const bool diag = (tl.I == tl.J);
        for (int ks = 0; ks < n_kstage; ++ks, ++gs) {
          const int rs = gs % RAW_STAGES;
          mbar_wait(&S->raw_empty[rs], ((gs / RAW_STAGES) & 1) ^ 1, a.err, 600 + rs);
          const uint32_t dst = smem_u32(raw + rs * RAW_STAGE_BYTES);
          if (a.ablate & 32) { mbar_arrive(&S->raw_full[rs]); continue; }
          mbar_arrive_expect_tx(&S->raw_full[rs], diag ? RAW_PANEL : RAW_STAGE_BYTES);
          tma_load_3d(dst, &tmap, &S->raw_full[rs], tl.I * TILE, tl.c, ks * ST_T);
          if (!diag) tma_load_3d(dst + RAW_PANEL, &tmap, &S->raw_full[rs], tl.J * TILE, tl.c, ks * ST_T);
        }
      }
    }
  } else {
    // ===================================================================== MMA issuer (one lane)
    if (lane == 0) {
      int gs = 0, it = 0;
      for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        mbar_wait(&S->tmem_empty[buf], ((it >> 1) & 1) ^ 1, a.err, 300 + buf);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + buf * 256;
        for (int ks = 0; ks < n_kstage; ++ks, ++gs) {
          const int slot = gs % NSTAGE;
          mbar_wait(&S->full[slot], (gs / NSTAGE) & 1, a.err, 400 + slot);
          tc_fence_after();
          const uint32_t sbase = smem_u32(stages + slot * STAGE_BYTES);
          if (!(a.ablate & 4)) {
            // [ re1 re0 | -re1 im0 ] += A_re x [B_re | B_nim]^T ;  [ im1 im0 | im1 re0 ] += A_im x [B_im | B_re]^T
            umma_i8_ss(tmem_d, umma_smem_desc(sbase + OFF_A_RE, KGRP_BYTES, SBO),
                       umma_smem_desc(sbase + OFF_B_RE, KGRP_BYTES, SBO), IDESC, ks != 0);
            umma_i8_ss(tmem_d, umma_smem_desc(sbase + OFF_A_IM, KGRP_BYTES, SBO),
                       umma_smem_desc(sbase + OFF_B_IM, KGRP_BYTES, SBO), IDESC, 1);
          }
          umma_commit(&S->empty[slot]);
          if (ks == n_kstage - 1) umma_commit(&S->tmem_full[buf]);
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

cudaError_t launch_corr_planes(const CorrArgs& a, int nsm, cudaStream_t stream) {
  static EncodeTiledFn encode = nullptr;
  static bool configured = false;
  if (!configured) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess) return e;
    if (qres != cudaDriverEntryPointSuccess || !fn) return cudaErrorNotSupported;
    encode = reinterpret_cast<EncodeTiledFn>(fn);
    e = cudaFuncSetAttribute(corr_planes_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
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
  corr_planes_kernel<<<grid, N_THREADS, SMEM_BYTES, stream>>>(tmap, a);
  return cudaGetLastError();
}

}  // namespace xeng
