// Corr (+ fused CorrAcc): the production kernel, TMA-fed tcgen05 (ninput % 16 == 0, 16-byte aligned buffers).
//
// Contraction (corr_block.py:445 / xgpu_test.py:102-136): per channel, word regtile_index(in0, in1) = sum_t conj(v[in0]) v[in1]
// for every station pair a1 >= a0, written in xGPU's register-tile order.  Here: a K-stacked int8 GEMM on the 5th-gen
// tensor cores with int32 accumulators in TMEM.
//
//   tile      (channel, 128 "row" inputs I, 128 "column" inputs J <= I); persistent grid, one CTA per SM, static
//             round-robin over tiles; TMEM holds two 256-column accumulators [Re(128) | Im(128)] (double buffered).
//   TMA warp  cp.async.bulk.tensor.3d boxes {128 inputs, 1 channel, 64 samples} of RAW 4+4-bit bytes, SWIZZLE_128B,
//             into a 4-deep ring (diagonal tiles fetch one panel).
//   unpack    2 groups x 4 warps on alternate stages.  A thread owns 32 consecutive inputs of one sample per
//             32-sample K-block (two LDS.128 of one swizzled row: conflict-free), sign-extends nibbles with SWAR
//             arithmetic and writes five int8 operand planes in tcgen05's MN-major no-swizzle core-matrix layout:
//             A_re, A_im (rows) and B_im, B_re, B_nim = -B_im (columns, reordered n = b0*64 + a0*2 + p0 with 8 PRMT).
//             Planes are contiguous along MN, so an N = 256 operand runs from one plane into the next:
//                 D[128 x 256] += A_re x [B_re | B_nim]^T      ->  [ re1 re0 | -re1 im0 ]
//                 D[128 x 256] += A_im x [B_im | B_re ]^T      ->  [ im1 im0 |  im1 re0 ]
//             Lanes run sample-fastest: the STS.128 of a quarter-warp are 128 contiguous bytes (conflict-free).
//             The raw slot is handed back with an arrive that is ordered after the LDS *data* (mbar_arrive_after).
//   MMA warp  one lane issues tcgen05.mma.kind::i8 M=128 N=256 K=32 (4 per stage), tcgen05.commit -> mbarriers.
//   epilogue  8 independent warps.  Warp w reads TMEM lanes 32*(w&3).. and the columns of stand parity b0 = w>>2
//             (tcgen05.ld.32x32b.x32), stages its own 16 runs (run = 32 cells x 16 B, contiguous in xGPU order) in a
//             private 8.4 KB buffer and issues their bulk copies itself: cp.async.bulk (store) or
//             cp.reduce.async.bulk.add.s32 (fused CorrAcc: the read-modify-write happens in L2).  No cross-warp
//             barrier; operands are warp-uniform so one elected lane issues without ptxas' per-lane waterfall.
//
// Why this shape (measurements: profiles/r01c_*, scripts/trace.py, scripts/ubench/bulk_rate.cu): a warp can issue one
// bulk copy per ~80 cycles, so the 220 copies of a tile need >= 4 warps issuing concurrently; one MMA-issuing thread
// needs >= 4 MMAs per barrier round trip to keep the tensor pipe fed; the bulk-reduce path itself sustains
// ~2.6 TB/s of payload by the wall clock (= 5.2 TB/s of HBM read + write; profiles/r03o_bulk_pattern.txt).
//
// Roles (832 threads): warps 0-7 epilogue, 8-15 unpack group 0, 16-23 unpack group 1, 24 TMA producer, 25 MMA issuer.
#include <cuda.h>

#include "xeng_ptx.cuh"
#include "xeng_internal.h"

#ifndef XENG_CT_ABLATE
#define XENG_CT_ABLATE 0
#endif
// developer timing builds only (results invalid): where the epilogue's bulk copies go.  1: every run starts on a
// 128-byte line; 2: a warp's 8 runs of one quadrant are contiguous (4 KB) instead of one per triangle row
#ifndef XENG_CT_PATTERN
#define XENG_CT_PATTERN 0
#endif
// fused CorrAcc: tiles ahead whose accumulator lines an epilogue warp pulls into L2 (prefetch.global.L2) before it
// needs them, so that the bulk reduction finds them there instead of holding its staging buffer through an HBM read
#ifndef XENG_CT_PREFETCH
#define XENG_CT_PREFETCH 0
#endif

namespace xeng {
namespace {

constexpr int TILE = 128;
constexpr int PAIRS = TILE / 4;                            // stand pairs per tile side
constexpr int ST_T = 64;                                   // samples per stage
constexpr int KB = ST_T / 32;                              // MMA K-blocks (32 samples) per stage
constexpr int NSTAGE = 2;                                  // operand ring depth
constexpr int RAW_STAGES = 4;                              // raw ring depth
constexpr int RAW_PANEL = ST_T * TILE;                     // 8 KB: [64 samples][128 inputs], 128B-swizzled
constexpr int RAW_STAGE_BYTES = 2 * RAW_PANEL;             // A | B
constexpr int KGRP_BYTES = 128;                            // core matrix: 8 samples x 16 inputs
constexpr int SBO = (ST_T / 8) * KGRP_BYTES;               // 1024
constexpr int PLANE = (TILE / 16) * SBO;                   // 8192
constexpr int STAGE_BYTES = 5 * PLANE;                     // A_re | A_im | B_im | B_re | B_nim
constexpr int OFF_A_RE = 0, OFF_A_IM = PLANE, OFF_B_IM = 2 * PLANE, OFF_B_RE = 3 * PLANE, OFF_B_NIM = 4 * PLANE;
constexpr int RUN_BYTES = PAIRS * 16;                      // 32 cells x 4 words
constexpr int RUN_STRIDE = RUN_BYTES + 16;
constexpr int HALF_SKEW = 64;                              // quadrant half b1 = 1 starts 64 B later (bank skew)
constexpr int WARP_STAGING = 16 * RUN_STRIDE + HALF_SKEW + 64;   // 8576: the 16 runs one epilogue warp stages
constexpr int N_EPI_WARPS = 8, N_GROUP_WARPS = 4 * KB, N_GROUPS = 2;   // a group = 4 warps per 32-sample K-block
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
constexpr int SMEM_HDR = 4096;                             // barriers + 2 KB scratch; raw panels stay 1024-byte aligned
constexpr int SCRATCH_OFF = 2048;                          // one dummy word per unpack thread (mbar_arrive_after)
static_assert(sizeof(SmemHdr) <= SMEM_HDR, "header");
static_assert(STAGE_BYTES % 128 == 0 && WARP_STAGING % 128 == 0, "alignment");
constexpr int SMEM_BYTES = SMEM_HDR + RAW_STAGES * RAW_STAGE_BYTES + NSTAGE * STAGE_BYTES + N_EPI_WARPS * WARP_STAGING + 1024;
static_assert(SMEM_BYTES <= 232448, "shared memory budget");

constexpr uint32_t IDESC = umma_idesc(2, 1, 1, 1, 1, 128, 256);
constexpr uint32_t IDESC_N128 = umma_idesc(2, 1, 1, 1, 1, 128, 128);   // ablation 128 only

// Tile order inside a channel: r = I*(I+1)/2 + J (J <= I).  (I, J) come from a constant table and (channel, r) are
// walked incrementally, so that every value derived from the tile index is computed on the uniform datapath (no
// division / sqrt): with the warp index taken from %warpid this keeps the bulk-copy operands in uniform registers.
constexpr int MAX_T = 64;                                  // row blocks: ninput <= 8192
constexpr int MAX_TPC = MAX_T * (MAX_T + 1) / 2;
__constant__ uint8_t c_tile_I[MAX_TPC];
__constant__ uint8_t c_tile_J[MAX_TPC];

struct Tile {
  int c, I, J;
};
struct TileWalk {
  int tile, c, r;
  __device__ __forceinline__ explicit TileWalk(const CorrArgs& a) {
    tile = blockIdx.x;
    c = 0;
    r = blockIdx.x;
    while (r >= a.tiles_per_chan) {
      r -= a.tiles_per_chan;
      ++c;
    }
  }
  __device__ __forceinline__ bool valid(const CorrArgs& a) const { return tile < a.ntiles; }
  __device__ __forceinline__ void next(const CorrArgs& a) {
    tile += gridDim.x;
    c += a.step_c;
    r += a.step_r;
    if (r >= a.tiles_per_chan) {
      r -= a.tiles_per_chan;
      ++c;
    }
  }
  __device__ __forceinline__ Tile get() const {
    Tile t;
    t.c = c;
    t.I = c_tile_I[r];
    t.J = c_tile_J[r];
    return t;
  }
};

__device__ __forceinline__ uint4 re16(uint4 x) {
  return make_uint4(unpack_re4(x.x), unpack_re4(x.y), unpack_re4(x.z), unpack_re4(x.w));
}
__device__ __forceinline__ uint4 im16(uint4 x) {
  return make_uint4(unpack_im4(x.x), unpack_im4(x.y), unpack_im4(x.z), unpack_im4(x.w));
}
__device__ __forceinline__ uint4 neg16(uint4 x) { return make_uint4(neg4(x.x), neg4(x.y), neg4(x.z), neg4(x.w)); }

// 32 raw bytes of one sample = 8 stand pairs x (b0, p0).  Column unit of stand parity b0: the half-words b0 of the
// eight words, i.e. 8 pairs x 2 pols = 16 columns n = a0*2 + p0.
__device__ __forceinline__ uint4 parity_lo(uint4 a, uint4 b) {
  return make_uint4(__byte_perm(a.x, a.y, 0x5410), __byte_perm(a.z, a.w, 0x5410), __byte_perm(b.x, b.y, 0x5410),
                    __byte_perm(b.z, b.w, 0x5410));
}
__device__ __forceinline__ uint4 parity_hi(uint4 a, uint4 b) {
  return make_uint4(__byte_perm(a.x, a.y, 0x7632), __byte_perm(a.z, a.w, 0x7632), __byte_perm(b.x, b.y, 0x7632),
                    __byte_perm(b.z, b.w, 0x7632));
}

// developer phase timer: time since the previous mark is charged to bucket i of the calling thread; the buckets of
// one thread per role of CTA 0 are written to a.trace at the end (ablation instantiation only, registers only)
#define XTRACE(role, i)                                                                    \
  do {                                                                                     \
    if (ABL && a.trace) {                                                                  \
      const uint32_t now__ = static_cast<uint32_t>(clock64());                             \
      tacc[(i) & 7] += now__ - tlast;                                                      \
      tlast = now__;                                                                       \
    }                                                                                      \
  } while (0)
#define XTRACE_DUMP(role)                                                                  \
  do {                                                                                     \
    if (ABL && a.trace && blockIdx.x == 0)                                                 \
      for (int i__ = 0; i__ < 8; ++i__) a.trace[(role) * 8 + i__] = tacc[i__];             \
  } while (0)

constexpr uint32_t SUB_NONE = 0xFFFFFFFFu;
// one selected word: staging -> (negate the imaginary part where conj) -> out[v][plane] += value
#ifndef XENG_CT_SUBX
#define XENG_CT_SUBX 0   // developer timing builds (results invalid): 1 = entries fetched, nothing read or reduced; 2 = no reduction
#endif
__device__ __forceinline__ void sub_emit(int32_t* so, uint32_t stg_w, uint32_t en, int plane) {
  if (XENG_CT_SUBX == 1) {
    if (en == 0x7FFFFFF1u) asm volatile("red.global.add.s32 [%0], %1;" ::"l"(so), "r"(1) : "memory");   // keeps `en` live
    return;
  }
  int val = static_cast<int>(lds_u32(stg_w + (en & 0x3FFFu)));
  if (plane == 1 && ((en >> 14) & 1u)) val = -val;
  if (XENG_CT_SUBX == 2) {
    if (val == 0x7FFFFFF1) asm volatile("red.global.add.s32 [%0], %1;" ::"l"(so), "r"(val) : "memory");
    return;
  }
  asm volatile("red.global.add.s32 [%0], %1;" ::"l"(so + static_cast<long>(en >> 15) * 2), "r"(val) : "memory");
}

template <bool ABL, bool SUB>
__global__ void __launch_bounds__(N_THREADS, 1) corr_swz_kernel(const __grid_constant__ CUtensorMap tmap, CorrArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  SmemHdr* S = reinterpret_cast<SmemHdr*>(smem);
  uint8_t* raw = smem + SMEM_HDR;
  uint8_t* stages = raw + RAW_STAGES * RAW_STAGE_BYTES;
  uint8_t* staging = stages + NSTAGE * STAGE_BYTES;

  // Warp index through redux.sync (CREDUX writes a uniform register): ptxas then knows it is warp-uniform -- which it
  // does not for threadIdx.x >> 5 or a shuffle -- so the role dispatch stays convergent and the bulk-copy operands
  // of the epilogue are computed on the uniform datapath (no per-lane R2UR waterfall).
  const int warp = static_cast<int>(__reduce_min_sync(0xffffffffu, threadIdx.x >> 5));
  const int lane = threadIdx.x & 31;
  const int n_kstage = (a.ntime + ST_T - 1) / ST_T;     // the TMA zero-fills samples past ntime
  const int ablate = ABL ? a.ablate : XENG_CT_ABLATE;   // XENG_CT_ABLATE: compile-time ablation of developer timing builds
  uint32_t tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  uint32_t tlast = static_cast<uint32_t>(clock64());
  (void)tacc; (void)tlast;

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
    // ===================================================================== epilogue (8 independent warps)
    // Warp w reads TMEM lanes 32*(w & 3) .. +31 (tile rows m = 4*A1loc + 2*b1 + p1) and the accumulator columns of
    // stand parity b0 = w >> 2, plane 0 then plane 1.  A lane stages only its own row, so the 16 runs a warp
    // writes ((b1, 8 row pairs) x 32 cells x 16 B) are private to it: no cross-warp barrier, the warp issues the
    // bulk copies of exactly the runs it staged and waits only for its own copies before reusing its buffer.
    const int q = warp & 3, b0 = warp >> 2;
    const int m = q * 32 + lane;
    const int A1l = lane >> 2, b1 = (lane >> 1) & 1, p1 = lane & 1;   // A1loc = 8*q + A1l
    const long plane_words = static_cast<long>(a.nchan) * a.words_per_chan;
    const uint32_t stg_w = smem_u32(staging) + warp * WARP_STAGING;
    const uint32_t my_run = stg_w + (b1 * 8 + A1l) * RUN_STRIDE + b1 * HALF_SKEW + p1 * 8;
    int it = 0;
    for (TileWalk tw(a); tw.valid(a); tw.next(a), ++it) {
      const Tile tl = tw.get();
      const int buf = it & 1;
      int sub_lb = 0, sub_le = 0;
      uint32_t sub_e0 = SUB_NONE, sub_e1 = SUB_NONE;
      if (SUB) {     // this (tile, warp)'s selection list; its global-load latency hides behind the accumulator wait
        sub_lb = a.sub_begin[tw.r * N_EPI_WARPS + warp];
        sub_le = a.sub_begin[tw.r * N_EPI_WARPS + warp + 1];
        if (sub_lb + lane < sub_le) sub_e0 = a.sub_entries[sub_lb + lane];
        if (sub_lb + 32 + lane < sub_le) sub_e1 = a.sub_entries[sub_lb + 32 + lane];
      }
      if (XENG_CT_PREFETCH > 0 && a.accumulate) {
        // lane -> one of this warp's 32 runs (plane, quadrant half rb1, row pair k) of the tile XENG_CT_PREFETCH ahead
        TileWalk pw = tw;
#pragma unroll
        for (int i = 0; i < XENG_CT_PREFETCH; ++i) pw.next(a);
        if (pw.valid(a)) {
          const Tile pt = pw.get();
          const int pl = lane >> 4, rb1 = (lane >> 3) & 1, k = lane & 7;
          const int pA0 = pt.J * PAIRS, prow = pt.I * PAIRS + 8 * q + k;
          const int pn = min(pt.I == pt.J ? (8 * q + k + 1) : PAIRS, a.npairs - pA0);
          if (prow < a.npairs && pn > 0) {
            const long cell = (static_cast<long>(prow) * (prow + 1)) / 2 + pA0;
            const char* p0 = reinterpret_cast<const char*>(a.out + static_cast<long>(pt.c) * a.words_per_chan +
                                                           pl * plane_words +
                                                           ((static_cast<long>(2 * b0 + rb1) * a.quad_cells + cell) << 2));
            const char* pe = p0 + pn * 16;
            const char* ln = reinterpret_cast<const char*>(reinterpret_cast<uintptr_t>(p0) & ~uintptr_t(127));
#pragma unroll
            for (int j = 0; j < 5; ++j, ln += 128)
              if (ln < pe) asm volatile("prefetch.global.L2 [%0];" ::"l"(ln));
          }
        }
      }
      XTRACE(4, 0);
      mbar_wait(&S->tmem_full[buf], (it >> 1) & 1, a.err, 100 + buf);
      XTRACE(4, 1);
      tc_fence_after();
      const bool diag = (tl.I == tl.J);
      const int A0base = tl.J * PAIRS;
      const int rA1_0 = tl.I * PAIRS + 8 * q;                            // first row pair of this warp
      const int ncell_max = min(PAIRS, a.npairs - A0base);
      const long cell_0 = (static_cast<long>(rA1_0) * (rA1_0 + 1)) / 2 + A0base;
      int nrow = min(8, a.npairs - rA1_0);                               // row pairs of this warp that exist
      if (ablate & 1) nrow = 0;
      int32_t* const out_c = a.out + static_cast<long>(tl.c) * a.words_per_chan;
#pragma unroll 1
      for (int plane = 0; plane < 2; ++plane) {
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                               static_cast<uint32_t>(buf * 256 + plane * 128 + b0 * 64);
        uint32_t r0[32];
        tmem_ld_32x32b_x32(taddr, r0);
        bulk_wait_read<0>();     // this warp's previous copies have left its staging buffer (groups: elected lane)
        __syncwarp();
        tmem_ld_wait();
        XTRACE(4, 4);
        if (!(ablate & 2)) {
#pragma unroll
          for (int a0 = 0; a0 < 16; ++a0) sts_v2(my_run + a0 * 16, r0[2 * a0], r0[2 * a0 + 1]);
        }
        tmem_ld_32x32b_x32(taddr + 32, r0);
        tmem_ld_wait();
        if (plane == 1) {  // accumulator fully read by this warp: hand the TMEM buffer back to the MMA warp
          tc_fence_before();
          mbar_arrive(&S->tmem_empty[buf]);
        }
        if (!(ablate & 2)) {
#pragma unroll
          for (int a0 = 0; a0 < 16; ++a0) sts_v2(my_run + (a0 + 16) * 16, r0[2 * a0], r0[2 * a0 + 1]);
          // unused xGPU padding cell of a diagonal tile (a0 == a1 in swapped order) is written as 0
          if (diag && b0 == 1 && b1 == 0) sts_v2(my_run + (8 * q + A1l) * 16, 0u, 0u);
        }
        XTRACE(4, 5);
        fence_proxy_async_smem();
        __syncwarp();
        XTRACE(4, 6);
        // warp-uniform: every lane walks the 16 runs (b1 = 0: 8 row pairs, b1 = 1: 8 row pairs), the elected lane issues
#pragma unroll
        for (int rb1 = 0; rb1 < 2; ++rb1) {
          int32_t* dst = out_c + plane * plane_words + ((static_cast<long>(2 * b0 + rb1) * a.quad_cells + cell_0) << 2);
          uint32_t src = stg_w + rb1 * (8 * RUN_STRIDE + HALF_SKEW);
          int rA1 = rA1_0;
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int n = min(diag ? (8 * q + k + 1) : PAIRS, ncell_max);
            if (k < nrow) {
              int32_t* d = dst;
              if (XENG_CT_PATTERN == 1) d = reinterpret_cast<int32_t*>(reinterpret_cast<uintptr_t>(dst) & ~uintptr_t(127));
              if (a.accumulate) bulk_reduce_add_s32_elect(d, src, n * 16);
              else bulk_store_elect(d, src, n * 16);
            }
            ++rA1;
            if (XENG_CT_PATTERN == 2) dst += PAIRS << 2;
            else
            dst += rA1 << 2;           // next row pair of the triangle starts rA1+1 cells later
            src += RUN_STRIDE;
          }
        }
        bulk_commit();
        if (SUB) {
          // fused CorrSubsel (corr_subsel_block.py:298): the selected words of this warp's 16 runs, straight from the
          // staging buffer; the nchan_sum channels of one output row come from different tiles -> red.global.add.
          // The first 64 entries of the list were fetched before the accumulator wait (sub_e0 / sub_e1).
          int32_t* const so = a.sub_out + (static_cast<long>(tl.c / a.sub_nsum) * a.sub_nvis) * 2 + plane;
          if (sub_e0 != SUB_NONE) sub_emit(so, stg_w, sub_e0, plane);
          if (sub_e1 != SUB_NONE) sub_emit(so, stg_w, sub_e1, plane);
          for (int e = sub_lb + 64 + lane; e < sub_le; e += 32) sub_emit(so, stg_w, a.sub_entries[e], plane);
        }
        XTRACE(4, 7);
      }
    }
    bulk_wait<0>();
    if (warp == 0 && lane == 0) XTRACE_DUMP(4);
  } else if (warp < WARP_TMA) {
    // ===================================================================== unpack: raw ring -> operand planes
    const int grp = (warp - N_EPI_WARPS) / N_GROUP_WARPS;          // stages with (index & 1) == grp
    const int kb = ((warp - N_EPI_WARPS) % N_GROUP_WARPS) >> 2;     // the 32-sample K-block of the stage this warp unpacks
    const int wg = (warp - N_EPI_WARPS) & 3;
    const int tlow = lane & 7, c = lane >> 3;                        // sample 32*kb + wg*8 + tlow, inputs 32c .. 32c+31
    // swizzled raw row: logical 16-byte chunk j of row t sits at chunk j ^ (t & 7)
    const uint32_t lds_a = (kb * 32 + wg * 8 + tlow) * TILE + (((2 * c) ^ tlow) << 4), lds_b = lds_a ^ 16u;
    const uint32_t kofs = (kb * 4 + wg) * KGRP_BYTES + tlow * 16;    // K position inside a 16-input group
    const uint32_t stsA = (2 * c) * SBO + kofs;                      // tile rows 32c .. 32c+15, next group +SBO
    const uint32_t stsB = c * SBO + kofs;                            // columns n = 16c .. (b0 = 0), +4*SBO for b0 = 1
    const uint32_t raw0 = smem_u32(raw), stages0 = smem_u32(stages);
    const uint32_t scratch = smem_u32(smem) + SCRATCH_OFF + (threadIdx.x - N_EPI_WARPS * 32) * 4;
    int gs0 = 0;                                                     // running stage index at the tile start
    for (TileWalk tw(a); tw.valid(a); tw.next(a), gs0 += n_kstage) {
      const Tile tl = tw.get();
      const uint32_t boff = (tl.I == tl.J) ? 0u : static_cast<uint32_t>(RAW_PANEL);   // diagonal: B raw == A raw
      for (int ks = (gs0 ^ grp) & 1; ks < n_kstage; ks += 2) {
        const int gs = gs0 + ks;
        const int rs = gs % RAW_STAGES, slot = gs % NSTAGE;
        XTRACE(grp, 0);
        const bool ldg_path = ABL && (ablate & 64);                    // ablation: global -> registers, no raw ring
        // ablation 128 (timing only, results invalid): what a CTA pair with the B columns split would leave one CTA
        // to do -- B loads / stores for half the columns, MMAs that read half of B
        const bool halfB = (ablate & 128) != 0;
        if (!ldg_path) mbar_wait(&S->raw_full[rs], (gs / RAW_STAGES) & 1, a.err, 500 + rs);
        XTRACE(grp, 1);
        const uint32_t rbase = raw0 + rs * RAW_STAGE_BYTES;
        const bool live = kb * 32 < a.ntime - ks * ST_T;               // this K-block holds samples (else: not issued)
        uint4 rawA0, rawA1, rawB0, rawB1;
        rawA0 = rawA1 = rawB0 = rawB1 = make_uint4(0u, 0u, 0u, 0u);
        if (ldg_path) {
          if (live) {   // timing only: no bounds handling (rows past ninput read the next channel's bytes)
            const long t = ks * ST_T + kb * 32 + wg * 8 + tlow;
            const uint8_t* g = a.in + (min(t, static_cast<long>(a.ntime - 1)) * a.nchan + tl.c) * a.ninput + 32 * c;
            const int ia = min(tl.I * TILE, a.ninput - TILE), ib = min(tl.J * TILE, a.ninput - TILE);
            rawA0 = __ldg(reinterpret_cast<const uint4*>(g + ia));
            rawA1 = __ldg(reinterpret_cast<const uint4*>(g + ia + 16));
            rawB0 = __ldg(reinterpret_cast<const uint4*>(g + ib));
            rawB1 = __ldg(reinterpret_cast<const uint4*>(g + ib + 16));
          }
        } else {
        if (live) {
          rawA0 = lds_v4(rbase + lds_a);
          rawA1 = lds_v4(rbase + lds_b);
          if (!halfB || c < 2) {
            rawB0 = lds_v4(rbase + boff + lds_a);
            rawB1 = lds_v4(rbase + boff + lds_b);
          }
        }
        // The TMA warp may refill the raw slot once the loads have RETURNED, not merely issued (SYNCS.ARRIVE does
        // not wait for outstanding LDS): make the arrive depend on one word of every load.
        mbar_arrive_after(&S->raw_empty[rs], scratch, rawA0.x ^ rawA1.x ^ rawB0.x ^ rawB1.x);
        }
        const uint32_t base = stages0 + slot * STAGE_BYTES;
        uint4 reA0, imA0, reA1, imA1, reB0, imB0, reB1, imB1, nimB0, nimB1;
        if (live) {
          const uint4 colB0 = parity_lo(rawB0, rawB1), colB1 = parity_hi(rawB0, rawB1);
          if (ABL && (ablate & 16)) {
            reA0 = imA0 = rawA0; reA1 = imA1 = rawA1;
            reB0 = imB0 = nimB0 = colB0; reB1 = imB1 = nimB1 = colB1;
          } else {
            reA0 = re16(rawA0); imA0 = im16(rawA0); reA1 = re16(rawA1); imA1 = im16(rawA1);
            reB0 = re16(colB0); imB0 = im16(colB0); reB1 = re16(colB1); imB1 = im16(colB1);
            nimB0 = neg16(imB0); nimB1 = neg16(imB1);
          }
        }
        // every thread waits, live or not: a parity wait that is skipped leaves the thread a phase behind, and its next
        // wait on this barrier would then be satisfied one MMA too early
        XTRACE(grp, 2);
        mbar_wait(&S->empty[slot], ((gs / NSTAGE) & 1) ^ 1, a.err, 200 + slot);
        XTRACE(grp, 3);
        if (live) {
          if (!(ablate & 8)) {
            sts_v4(base + OFF_A_RE + stsA, reA0);
            sts_v4(base + OFF_A_RE + stsA + SBO, reA1);
            sts_v4(base + OFF_A_IM + stsA, imA0);
            sts_v4(base + OFF_A_IM + stsA + SBO, imA1);
            if (!halfB || c < 2) {
            sts_v4(base + OFF_B_IM + stsB, imB0);
            sts_v4(base + OFF_B_IM + stsB + 4 * SBO, imB1);
            sts_v4(base + OFF_B_RE + stsB, reB0);
            sts_v4(base + OFF_B_RE + stsB + 4 * SBO, reB1);
            sts_v4(base + OFF_B_NIM + stsB, nimB0);
            sts_v4(base + OFF_B_NIM + stsB + 4 * SBO, nimB1);
            }
          }
        }
        XTRACE(grp, 4);
        fence_proxy_async_smem();
        XTRACE(grp, 5);
        mbar_arrive(&S->full[slot]);
      }
    }
    if (kb == 0 && wg == 0 && lane == 0) XTRACE_DUMP(grp);
  } else if (warp == WARP_TMA) {
    // ===================================================================== TMA producer (one lane)
    if (lane == 0 && !(ABL && (ablate & 64))) {
      int gs = 0;
      for (TileWalk tw(a); tw.valid(a); tw.next(a)) {
        const Tile tl = tw.get();
        const bool diag = (tl.I == tl.J);
        for (int ks = 0; ks < n_kstage; ++ks, ++gs) {
          const int rs = gs % RAW_STAGES;
          XTRACE(3, 0);
          mbar_wait(&S->raw_empty[rs], ((gs / RAW_STAGES) & 1) ^ 1, a.err, 600 + rs);
          XTRACE(3, 1);
          const uint32_t dst = smem_u32(raw + rs * RAW_STAGE_BYTES);
          if (ABL && (ablate & 32)) { mbar_arrive(&S->raw_full[rs]); continue; }
          mbar_arrive_expect_tx(&S->raw_full[rs], diag ? RAW_PANEL : RAW_STAGE_BYTES);
          tma_load_3d(dst, &tmap, &S->raw_full[rs], tl.I * TILE, tl.c, ks * ST_T);
          if (!diag) tma_load_3d(dst + RAW_PANEL, &tmap, &S->raw_full[rs], tl.J * TILE, tl.c, ks * ST_T);
        }
      }
      XTRACE_DUMP(3);
    }
  } else {
    // ===================================================================== MMA issuer (one lane)
    if (lane == 0) {
      int gs = 0, it = 0;
      for (TileWalk tw(a); tw.valid(a); tw.next(a), ++it) {
        const int buf = it & 1;
        XTRACE(2, 0);
        mbar_wait(&S->tmem_empty[buf], ((it >> 1) & 1) ^ 1, a.err, 300 + buf);
        XTRACE(2, 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + buf * 256;
        for (int ks = 0; ks < n_kstage; ++ks, ++gs) {
          const int slot = gs % NSTAGE;
          XTRACE(2, 2);
          mbar_wait(&S->full[slot], (gs / NSTAGE) & 1, a.err, 400 + slot);
          XTRACE(2, 3);
          tc_fence_after();
          const uint32_t sbase = smem_u32(stages + slot * STAGE_BYTES);
          const int nkb = min(KB, (a.ntime - ks * ST_T + 31) >> 5);
          if (!(ablate & 4)) {
            // [ re1 re0 | -re1 im0 ] += A_re x [B_re | B_nim]^T ;  [ im1 im0 | im1 re0 ] += A_im x [B_im | B_re]^T
#pragma unroll
            for (int j = 0; j < KB; ++j) {
              if (j < nkb) {
                const uint32_t sj = sbase + j * (4 * KGRP_BYTES);
                const uint32_t idesc = (ablate & 128) ? IDESC_N128 : IDESC;
                umma_i8_ss(tmem_d, umma_smem_desc(sj + OFF_A_RE, KGRP_BYTES, SBO),
                           umma_smem_desc(sj + OFF_B_RE, KGRP_BYTES, SBO), idesc, (ks | j) != 0);
                umma_i8_ss(tmem_d, umma_smem_desc(sj + OFF_A_IM, KGRP_BYTES, SBO),
                           umma_smem_desc(sj + OFF_B_IM, KGRP_BYTES, SBO), idesc, 1);
              }
            }
          }
          umma_commit(&S->empty[slot]);
          XTRACE(2, 4);
          if (ks == n_kstage - 1) umma_commit(&S->tmem_full[buf]);
        }
      }
      XTRACE_DUMP(2);
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

int corr_swz_staging_offset(int A1loc, int b1, int p1, int a0, int p0) {
  // must match the epilogue: lane (A1loc & 7, b1, p1) writes {column 2*a0, 2*a0+1} at my_run + a0*16
  return (b1 * 8 + (A1loc & 7)) * RUN_STRIDE + b1 * HALF_SKEW + p1 * 8 + a0 * 16 + p0 * 4;
}

bool corr_tma_supported(const CorrArgs& a) {
  return (a.ninput % 16 == 0) && (a.ninput <= MAX_T * TILE) && ((reinterpret_cast<uintptr_t>(a.in) & 15) == 0) &&
         ((reinterpret_cast<uintptr_t>(a.out) & 15) == 0);
}

cudaError_t launch_corr_swz(const CorrArgs& a_in, int nsm, cudaStream_t stream) {
  CorrArgs a = a_in;
  static EncodeTiledFn encode = nullptr;     // driver entry point: process-wide, not per device
  static std::once_flag encode_once;
  static cudaError_t encode_err = cudaSuccess;
  std::call_once(encode_once, []() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    encode_err = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (encode_err == cudaSuccess && (qres != cudaDriverEntryPointSuccess || !fn)) encode_err = cudaErrorNotSupported;
    if (encode_err == cudaSuccess) encode = reinterpret_cast<EncodeTiledFn>(fn);
  });
  if (encode_err != cudaSuccess) return encode_err;
  // function attributes and the __constant__ tile table live in the device's context: once per device
  cudaError_t ce = configure_once_per_device([]() -> cudaError_t {
    cudaError_t e = cudaFuncSetAttribute(corr_swz_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(corr_swz_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(corr_swz_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    static uint8_t tab_i[MAX_TPC], tab_j[MAX_TPC];   // r = I*(I+1)/2 + J  ->  (I, J); the same for every shape
    for (int I = 0, r = 0; I < MAX_T; ++I)
      for (int J = 0; J <= I; ++J, ++r) {
        tab_i[r] = static_cast<uint8_t>(I);
        tab_j[r] = static_cast<uint8_t>(J);
      }
    e = cudaMemcpyToSymbol(c_tile_I, tab_i, sizeof(tab_i));
    if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_tile_J, tab_j, sizeof(tab_j));
    return e;
  }, kCfgCorrSwz);
  if (ce != cudaSuccess) return ce;
  int grid = a.ntiles < nsm ? a.ntiles : nsm;
  if (grid <= 0) return cudaSuccess;
  a.step_c = grid / a.tiles_per_chan;
  a.step_r = grid % a.tiles_per_chan;
  alignas(64) CUtensorMap tmap;
  const cuuint64_t gdim[3] = {static_cast<cuuint64_t>(a.ninput), static_cast<cuuint64_t>(a.nchan),
                              static_cast<cuuint64_t>(a.ntime)};
  const cuuint64_t gstride[2] = {static_cast<cuuint64_t>(a.ninput),
                                 static_cast<cuuint64_t>(a.ninput) * static_cast<cuuint64_t>(a.nchan)};
  const cuuint32_t box[3] = {TILE, 1, ST_T};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t*>(a.in), gdim, gstride, box, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
  if (a.sub_out) corr_swz_kernel<false, true><<<grid, N_THREADS, SMEM_BYTES, stream>>>(tmap, a);
  else if (a.ablate) corr_swz_kernel<true, false><<<grid, N_THREADS, SMEM_BYTES, stream>>>(tmap, a);
  else corr_swz_kernel<false, false><<<grid, N_THREADS, SMEM_BYTES, stream>>>(tmap, a);
  return cudaGetLastError();
}

}  // namespace xeng
