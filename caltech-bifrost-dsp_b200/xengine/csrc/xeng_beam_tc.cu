// Beamform (+ fused BeamformSum power integration) on the 5th-gen tensor cores.
//
// bfBeamformRun (call site /root/reference/pipeline/lwa352_pipeline/blocks/beamform_block.py:449):
//     out[c][b][t] = sum_i w[c][b][i] * v[t][c][i]         complex, fp32 weights, 4+4-bit integer voltages
//     (pinned by pipeline/verification/tests/beamformer/beamformer_test.py:76-84)
//
// The contraction is integer on one side: v is exact in int8.  The fp32 weights of one (channel, beam) row are
// rewritten once per weight update (beam_weight_digits_kernel) as a 24-bit block-floating-point integer
// W = rint(w * 2^(23-e)), 2^e just above the largest |component| of the row, split into three signed radix-256
// digits W = d0 + 256 d1 + 65536 d2, d in [-128, 127].  Each digit plane is an int8 GEMM operand, so
//
//     out = 2^(e-23) * (D0 + 256 D1 + 65536 D2),    Dg = sum_i v_i * dg_i    exact int32 (|Dg| <= 2 ninput * 1024)
//
// runs as tcgen05.mma.kind::i8 with int32 accumulators in TMEM: exact products, exact order-independent sums,
// and the only roundings are the weight quantisation (<= 2^-24 of the row's largest component: what fp32 storage of
// that component already costs) and two fp32 roundings when the three digit sums are recombined.  int8 runs at
// twice the bf16 rate with half the operand bytes of a 3-way bf16 split.
//
//   GEMM     per (channel c, 128 time samples): D[128 t x N] = A[128 x K] * B[N x K]^T,
//            K = 2 ninput stacked [re | im], N = 3 digits x 2 (re, im) x nb beams (192 for 32 beams), B rows
//                out_re(b): [ w_re | -w_im ]      out_im(b): [ w_im | w_re ]
//   A        lives in TENSOR MEMORY (tcgen05.mma with the A operand in TMEM, K-major).  The producer warp brings RAW
//            4+4-bit boxes {64 inputs, 1 channel, 128 samples} into a 6-deep shared-memory ring with TMA
//            (SWIZZLE_64B; a thread reading its own sample's 64 bytes straight from global memory costs a warp 32
//            sector requests per 16-byte load -- the first version of this kernel was L1-request bound on exactly
//            that); an unpack thread takes its sample's row with four conflict-free LDS.128, sign-extends the
//            nibbles with SWAR arithmetic and tcgen05.st's them as 16 + 16 columns (re | im).  The UNPACKED voltages
//            never touch shared memory.
//   B        digit planes are precomputed in global memory in exactly the K-major no-swizzle core-matrix layout the
//            MMA reads ([n/8][k/16][n%8][16 B] per 32-deep K-block), so one stage (64 inputs = 4 MMAs) is ONE 1-D
//            bulk copy (cp.async.bulk, 24 KB) into a 6-deep shared-memory ring of its own (an L2 -> SM copy needs
//            ~1300 cycles to land; sharing the A slots' depth of 4 left the MMA warp waiting for digits).
//   D        two accumulators of N columns (double buffered): the epilogue of tile i overlaps the MMAs of tile i+1.
//   epilogue 4 warps (one per TMEM lane quarter; lane = time sample): tcgen05.ld the three digit sums of 8 beams,
//            recombine in fp32, scale, and either store cf32 voltages [c][b][t] (a warp writes 256 contiguous bytes
//            per beam) or, FUSED, form [|X|^2, |Y|^2, Re XY*, Im XY*] per dual-pol beam in shared memory and sum
//            ntime_sum samples in time order (beamform_sum_beams_block.py:245; beamformer_sum_test.py:64-77) --
//            the same expressions and order as beam_integrate_kernel, so both paths agree bit for bit.
//
// Roles (608 threads): warps 0-3 epilogue, 4-15 unpack (3 groups x 4 lane quarters, a group per stage; A slots in TMEM
// columns 384..511, 4 deep), 16 digit-plane producer, 17 raw-voltage producer, 18 MMA issuer.  Persistent grid, one CTA per SM,
// tiles ordered (beam block, channel, time tile) so the CTAs working at one moment share a channel's digits in L2.
#include <cuda.h>

#include "xeng_ptx.cuh"
#include "xeng_internal.h"

namespace xeng {
namespace {

constexpr int BT_M = 128;                    // time samples per tile (TMEM lanes)
constexpr int BT_KIN = 64;                   // inputs per stage: 4 MMAs of K = 32 (re 0-31, re 32-63, im 0-31, im 32-63)
constexpr int BT_SLOTS = 4;                  // A-stage ring depth (TMEM columns 384..511)
constexpr int BT_BSLOTS = 6;                 // digit-plane (B) ring depth in shared memory: an L2 -> SM bulk copy takes
                                             // ~1300 cycles to land against 384 cycles of MMAs per stage
constexpr int BT_ACOLS = 32;                 // TMEM columns of one A stage: 16 re + 16 im
constexpr int BT_DCOLS = 192;                // column stride between the two accumulators
constexpr int BT_A_COL0 = 2 * BT_DCOLS;      // 384: A stages live in columns 384..511
constexpr int BT_NB_MAX = 32;                // beams per beam block
constexpr int BT_EPI_WARPS = 4, BT_GROUPS = 3, BT_UNPACK_WARPS = 4 * BT_GROUPS;
constexpr int BT_WARP_PROD = BT_EPI_WARPS + BT_UNPACK_WARPS, BT_WARP_RAW = BT_WARP_PROD + 1, BT_WARP_MMA = BT_WARP_RAW + 1;
constexpr int BT_THREADS = (BT_WARP_MMA + 1) * 32;
constexpr int BT_STAGE_MAX = 6 * BT_NB_MAX * 2 * BT_KIN;          // 24576 B: N rows x 128 K-bytes
constexpr int BT_HDR = 4096;                                      // barriers + one scratch word per unpack thread
constexpr int BT_SCRATCH_OFF = 1024;
constexpr int BT_RAW_SLOTS = 6;                                   // raw ring depth: 48 KB of voltages in flight per SM
constexpr int BT_RAW_BYTES = BT_M * BT_KIN;                       // 8 KB: [128 samples][64 inputs], 64B-swizzled
constexpr int BT_PROD_BYTES = 4 * 4 * BT_M * 4;                   // 8 KB: [4 pairs][component][sample] floats
constexpr int BT_SMEM = BT_HDR + BT_RAW_SLOTS * BT_RAW_BYTES + BT_BSLOTS * BT_STAGE_MAX + BT_PROD_BYTES + 1024;
static_assert(BT_SMEM <= 232448, "shared memory budget");
static_assert(BT_A_COL0 + BT_SLOTS * BT_ACOLS <= 512, "TMEM budget");

struct BtHdr {
  uint64_t full[BT_SLOTS];       // A stage written to TMEM (4 unpack warps)
  uint64_t empty[BT_SLOTS];      // the MMAs reading the A slot have completed (tcgen05.commit)
  uint64_t b_full[BT_BSLOTS];    // digit-plane stage landed (expect_tx)
  uint64_t b_empty[BT_BSLOTS];   // the MMAs reading the B slot have completed (tcgen05.commit)
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint64_t raw_full[BT_RAW_SLOTS];    // TMA box landed
  uint64_t raw_empty[BT_RAW_SLOTS];   // the 4 unpack warps of the stage have read it
  uint32_t tmem_base;
};

struct BeamTcArgs {
  const uint8_t* in;       // [ntime][nchan][ninput] 4+4-bit
  const int8_t* wdig;      // [nchan][nbb][nst][stage_bytes] digit planes
  const float* wscale;     // [nchan][nbb][32] 2^(e-23)
  float2* out;             // [nchan][nbeam][ntime]
  float* pow_out;          // FUSED: [nbeam/2][ntime/ntime_sum][nchan][4]
  int ntime, nchan, ninput, nbeam;
  int nb, nbb, nst;        // beams per block (8/16/24/32), beam blocks, 64-input stages
  int ntt, tile_t;         // time tiles, samples owned by one tile (128, FUSED: the largest multiple of ntime_sum <= 128)
  int nt_valid;            // samples processed (ntime; FUSED: whole integration blocks only)
  int ntime_sum, ntiles;
};

__device__ __forceinline__ float digit_sum_to_float(uint32_t d) {
  // exact for |d| < 2^22: the integer is added into the mantissa of 1.5 * 2^23
  return __int_as_float(0x4B400000 + static_cast<int>(d)) - 12582912.0f;
}

template <bool FUSED>
__global__ void __launch_bounds__(BT_THREADS, 1) beamform_tc_kernel(const __grid_constant__ CUtensorMap tmap, BeamTcArgs a) {
  extern __shared__ uint8_t bt_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(bt_smem_raw) + 1023) & ~uintptr_t(1023));
  BtHdr* S = reinterpret_cast<BtHdr*>(smem);
  uint8_t* rawring = smem + BT_HDR;
  uint8_t* bst = rawring + BT_RAW_SLOTS * BT_RAW_BYTES;
  float* prod = reinterpret_cast<float*>(bst + BT_BSLOTS * BT_STAGE_MAX);

  const int warp = static_cast<int>(__reduce_min_sync(0xffffffffu, threadIdx.x >> 5));
  const int lane = threadIdx.x & 31;
  const int N = 6 * a.nb;
  const uint32_t stage_bytes = static_cast<uint32_t>(N) * (2 * BT_KIN);

  if (threadIdx.x == 0) {
    for (int s = 0; s < BT_SLOTS; ++s) {
      mbar_init(&S->full[s], 4);
      mbar_init(&S->empty[s], 1);
    }
    for (int s = 0; s < BT_BSLOTS; ++s) {
      mbar_init(&S->b_full[s], 1);
      mbar_init(&S->b_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&S->tmem_full[b], 1);
      mbar_init(&S->tmem_empty[b], BT_EPI_WARPS * 32);
    }
    for (int s = 0; s < BT_RAW_SLOTS; ++s) {
      mbar_init(&S->raw_full[s], 1);
      mbar_init(&S->raw_empty[s], 4 * 32);
    }
    mbar_fence_init();
  }
  if (warp == BT_WARP_MMA) {
    tmem_alloc(&S->tmem_base, 512);
    tmem_relinquish();
  }
  if (warp == BT_WARP_RAW && lane == 0) tma_prefetch_desc(&tmap);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = S->tmem_base;

  if (warp < BT_EPI_WARPS) {
    // ===================================================================== epilogue: lane = time sample
    const int q = warp, m = q * 32 + lane;
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    int it = 0;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
      const int tt = tile % a.ntt, cb = tile / a.ntt, c = cb % a.nchan, bb = cb / a.nchan;
      const int t = tt * a.tile_t + m, buf = it & 1;
      const bool row_ok = m < a.tile_t && t < a.nt_valid;
      const float* sc = a.wscale + (static_cast<long>(c) * a.nbb + bb) * BT_NB_MAX;
      mbar_wait(&S->tmem_full[buf], (it >> 1) & 1, nullptr, 100 + buf);
      tc_fence_after();
      const int nchunk = a.nb >> 3;                       // 8 beams = 16 accumulator columns per pass
      for (int ch = 0; ch < nchunk; ++ch) {
        uint32_t d0[16], d1[16], d2[16];
        const uint32_t taddr = lane_addr + static_cast<uint32_t>(buf * BT_DCOLS + ch * 16);
        tmem_ld_32x32b_x16(taddr, d0);
        tmem_ld_32x32b_x16(taddr + 2 * a.nb, d1);
        tmem_ld_32x32b_x16(taddr + 4 * a.nb, d2);
        tmem_ld_wait();
        if (ch == nchunk - 1) {      // accumulator fully read by this thread: hand it back to the MMA warp
          tc_fence_before();
          mbar_arrive(&S->tmem_empty[buf]);
        }
        float v[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          const float f0 = digit_sum_to_float(d0[k]), f1 = digit_sum_to_float(d1[k]), f2 = digit_sum_to_float(d2[k]);
          v[k] = fmaf(fmaf(f2, 256.0f, f1), 256.0f, f0) * __ldg(sc + ch * 8 + (k >> 1));
        }
        if (!FUSED) {
          if (row_ok) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const int b = bb * a.nb + ch * 8 + k;
              if (b < a.nbeam) a.out[(static_cast<long>(c) * a.nbeam + b) * a.ntime + t] = make_float2(v[2 * k], v[2 * k + 1]);
            }
          }
        } else {
          // per 8-beam chunk: the 4 dual-pol beams' per-sample products go to shared memory, then ntime_sum of them
          // are summed in time order (same expressions and order as beam_integrate_kernel)
#pragma unroll
          for (int p = 0; p < 4; ++p) {
            const float4 pr = beam_products(make_float2(v[4 * p], v[4 * p + 1]), make_float2(v[4 * p + 2], v[4 * p + 3]));
            float* dst = prod + (p * 4) * BT_M + m;
            dst[0] = pr.x;
            dst[BT_M] = pr.y;
            dst[2 * BT_M] = pr.z;
            dst[3 * BT_M] = pr.w;
          }
          named_bar_sync(1, BT_EPI_WARPS * 32);            // every sample's products of this chunk are in shared memory
          const int nblk = a.ntime / a.ntime_sum, per_tile = a.tile_t / a.ntime_sum, T0 = tt * per_tile;
          for (int e = threadIdx.x; e < 4 * per_tile * 4; e += BT_EPI_WARPS * 32) {
            const int comp = e & 3, rest = e >> 2, Tl = rest % per_tile, pair = rest / per_tile;
            const int pg = (bb * a.nb) / 2 + ch * 4 + pair;
            if (T0 + Tl >= nblk || 2 * pg + 1 >= a.nbeam) continue;
            const float* src = prod + (pair * 4 + comp) * BT_M + Tl * a.ntime_sum;
            float s = 0.f;
            for (int k = 0; k < a.ntime_sum; ++k) s += src[k];
            a.pow_out[((static_cast<long>(pg) * nblk + T0 + Tl) * a.nchan + c) * 4 + comp] = s;
          }
          named_bar_sync(1, BT_EPI_WARPS * 32);            // sums done before the next chunk's products land
        }
      }
    }
  } else if (warp < BT_WARP_PROD) {
    // ===================================================================== unpack: raw ring -> registers -> TMEM
    const int u = warp - BT_EPI_WARPS, q = u & 3, g = u >> 2;
    const int m = q * 32 + lane;
    const uint32_t a_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + BT_A_COL0;
    // row m of a 64B-swizzled box: logical 16-byte chunk j sits at chunk j ^ ((m >> 1) & 3)
    const uint32_t row_addr = smem_u32(rawring) + m * BT_KIN, sw = (m >> 1) & 3;
    const uint32_t scratch = smem_u32(smem) + BT_SCRATCH_OFF + (threadIdx.x - BT_EPI_WARPS * 32) * 4;
    const int nstage_total = ((a.ntiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) /
                              static_cast<int>(gridDim.x)) * a.nst;      // stages this CTA runs
    for (int gs = g; gs < nstage_total; gs += BT_GROUPS) {
      const int rs = gs % BT_RAW_SLOTS;
      mbar_wait(&S->raw_full[rs], (gs / BT_RAW_SLOTS) & 1, nullptr, 600 + rs);
      const uint32_t base = row_addr + rs * BT_RAW_BYTES;
      uint4 cur[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) cur[j] = lds_v4(base + ((j ^ sw) << 4));
      uint32_t r[32];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        r[4 * j + 0] = unpack_re4(cur[j].x); r[16 + 4 * j + 0] = unpack_im4(cur[j].x);
        r[4 * j + 1] = unpack_re4(cur[j].y); r[16 + 4 * j + 1] = unpack_im4(cur[j].y);
        r[4 * j + 2] = unpack_re4(cur[j].z); r[16 + 4 * j + 2] = unpack_im4(cur[j].z);
        r[4 * j + 3] = unpack_re4(cur[j].w); r[16 + 4 * j + 3] = unpack_im4(cur[j].w);
      }
      // The raw slot may be refilled once the loads have RETURNED, not merely issued (an mbarrier arrive does not
      // wait for outstanding LDS: xeng_corr_tma.cu found that race the hard way): the arrive is ordered behind a
      // store of a word that depends on every load.
      mbar_arrive_after(&S->raw_empty[rs], scratch, cur[0].x ^ cur[1].y ^ cur[2].z ^ cur[3].w);
      const int slot = gs % BT_SLOTS;
      mbar_wait(&S->empty[slot], ((gs / BT_SLOTS) & 1) ^ 1, nullptr, 200 + slot);
      tc_fence_after();
      tmem_st_32x32b_x32(a_addr + slot * BT_ACOLS, r);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&S->full[slot]);
    }
  } else if (warp == BT_WARP_PROD) {
    // ===================================================================== digit-plane producer (one lane)
    if (lane == 0) {
      int gs = 0;
      for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
        const int cb = tile / a.ntt;                       // = bb * nchan + c: digit planes are stored [c][bb]
        const int c = cb % a.nchan, bb = cb / a.nchan;
        const int8_t* src = a.wdig + (static_cast<long>(c) * a.nbb + bb) * a.nst * stage_bytes;
        for (int s = 0; s < a.nst; ++s, ++gs) {
          const int slot = gs % BT_BSLOTS;
          mbar_wait(&S->b_empty[slot], ((gs / BT_BSLOTS) & 1) ^ 1, nullptr, 300 + slot);
          mbar_arrive_expect_tx(&S->b_full[slot], stage_bytes);
          bulk_load(smem_u32(bst + slot * BT_STAGE_MAX), src + static_cast<long>(s) * stage_bytes, stage_bytes,
                    &S->b_full[slot]);
        }
      }
    }
  } else if (warp == BT_WARP_RAW) {
    // ===================================================================== raw-voltage producer (one lane)
    if (lane == 0) {
      int gs = 0;
      for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
        const int tt = tile % a.ntt, c = (tile / a.ntt) % a.nchan;
        for (int s = 0; s < a.nst; ++s, ++gs) {
          const int rs = gs % BT_RAW_SLOTS;
          mbar_wait(&S->raw_empty[rs], ((gs / BT_RAW_SLOTS) & 1) ^ 1, nullptr, 700 + rs);
          mbar_arrive_expect_tx(&S->raw_full[rs], BT_RAW_BYTES);
          tma_load_3d(smem_u32(rawring + rs * BT_RAW_BYTES), &tmap, &S->raw_full[rs], s * BT_KIN, c, tt * a.tile_t);
        }
      }
    }
  } else {
    // ===================================================================== MMA issuer (one lane)
    if (lane == 0) {
      const uint32_t idesc = umma_idesc(/*S32*/ 2, /*INT8*/ 1, /*INT8*/ 1, /*A K-major*/ 0, /*B K-major*/ 0, 128, N);
      int gs = 0, it = 0;
      for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        mbar_wait(&S->tmem_empty[buf], ((it >> 1) & 1) ^ 1, nullptr, 400 + buf);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + buf * BT_DCOLS;
        for (int s = 0; s < a.nst; ++s, ++gs) {
          const int slot = gs % BT_SLOTS, bslot = gs % BT_BSLOTS;
          mbar_wait(&S->b_full[bslot], (gs / BT_BSLOTS) & 1, nullptr, 550 + bslot);
          mbar_wait(&S->full[slot], (gs / BT_SLOTS) & 1, nullptr, 500 + slot);
          tc_fence_after();
          const uint32_t sb = smem_u32(bst + bslot * BT_STAGE_MAX);
          const uint32_t ta = tmem_base + BT_A_COL0 + slot * BT_ACOLS;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            umma_i8_ts(tmem_d, ta + j * 8, umma_smem_desc(sb + j * (N * 32), 128, 256), idesc, (s | j) != 0);
          umma_commit(&S->empty[slot]);
          umma_commit(&S->b_empty[bslot]);
        }
        umma_commit(&S->tmem_full[buf]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == BT_WARP_MMA) tmem_dealloc(tmem_base, 512);
}

// ----------------------------------------------------------------------------------------------------------------
// fp32 weights -> digit planes.  One block per (channel, 4 beams): thread (chunk, bq) owns inputs 16 chunk .. +15 of beam
// 4 quad + bq.  The four beams of a block are the eight rows (n & 7) of one core-matrix row group, and a thread
// writes its beam's two adjacent rows (out_re, out_im) as 32 contiguous bytes, so four neighbouring lanes fill a whole
// 128-byte core matrix: every store sector is complete.
__device__ __forceinline__ void split_digits(int w, int (&d)[3]) {
  d[0] = static_cast<int8_t>(w & 0xFF);   // signed low byte
  w = (w - d[0]) / 256;                   // exact
  d[1] = static_cast<int8_t>(w & 0xFF);
  d[2] = (w - d[1]) / 256;                // in [-128, 127] for |W| <= 8 355 711
}
// x * 2^k for any int k, exact unless the result leaves the normal range (two clamped power-of-two factors)
__device__ __forceinline__ float scale_pow2(float x, int k) {
  const int k1 = max(-126, min(127, k)), k2 = max(-126, min(127, k - k1));
  return x * __int_as_float((k1 + 127) << 23) * __int_as_float((k2 + 127) << 23);
}

constexpr int WD_THREADS = 128;
constexpr int WD_CHUNK_F2 = 18;            // a 16-input chunk (16 float2 = 128 B) padded to 144 B: conflict-free LDS.128
__global__ void __launch_bounds__(WD_THREADS) beam_weight_digits_kernel(const float2* __restrict__ w, int8_t* __restrict__ dig,
                                                                        float* __restrict__ scale, int nchan, int nbeam,
                                                                        int ninput, int nb, int nbb, int nst) {
  extern __shared__ __align__(16) float2 wtile[];          // [4 beams][nchunk][WD_CHUNK_F2]
  __shared__ int smax[4];                  // per beam: max |component| as float bits (non-negative floats order as ints)
  __shared__ int sbad;
  const int c = blockIdx.x, quad = blockIdx.y;
  const int tid = threadIdx.x;
  const int nchunk = nst * 4, row_f2 = nchunk * WD_CHUNK_F2;
  if (tid < 4) smax[tid] = 0;
  if (tid == 0) sbad = 0;
  for (int i = tid; i < 4 * row_f2; i += WD_THREADS) wtile[i] = make_float2(0.f, 0.f);   // K padding, padded beams
  __syncthreads();
  // ---- pass 1: the four weight rows into shared memory with coalesced 16-byte loads (a warp reads 512 contiguous
  // bytes), row maxima on the way
  const int f4_per_row = ninput / 2;
  for (int idx = tid; idx < 4 * f4_per_row; idx += WD_THREADS) {
    const int bq = idx / f4_per_row, pair = idx - bq * f4_per_row, i = 2 * pair;
    const int bg = quad * 4 + bq;
    if (bg >= nbeam) continue;
    const float4 v = __ldg(reinterpret_cast<const float4*>(w + (static_cast<long>(c) * nbeam + bg) * ninput + i));
    *reinterpret_cast<float4*>(&wtile[bq * row_f2 + (i >> 4) * WD_CHUNK_F2 + (i & 15)]) = v;
    const float a0 = fabsf(v.x), a1 = fabsf(v.y), a2 = fabsf(v.z), a3 = fabsf(v.w);
    if (!(a0 <= 3.0e38f) || !(a1 <= 3.0e38f) || !(a2 <= 3.0e38f) || !(a3 <= 3.0e38f)) atomicOr(&sbad, 1 << bq);   // inf / nan
    atomicMax(&smax[bq], __float_as_int(fmaxf(fmaxf(a0, a1), fmaxf(a2, a3))));
  }
  __syncthreads();
  // ---- pass 2: thread (chunk, bq) quantises and splits its 16 inputs, writes 32 contiguous bytes per digit block
  const int bq = tid & 3;
  const int bg = quad * 4 + bq, bb = bg / nb, bl = bg - bb * nb;      // padded beam index, its block, index in block
  const float mx = __int_as_float(smax[bq]);
  const bool rowbad = (sbad >> bq) & 1;    // inf / nan weights poison their whole row, as they would in fp32
  int e = 0;
  if (mx > 0.f) {
    const float fr = frexpf(mx, &e);       // mx = fr * 2^e, fr in [0.5, 1)
    if (fr > 0.99607f) ++e;                // keep |W| <= 127 * 65793 = 8 355 711 (three signed digits)
  }
  if (tid < 4)
    scale[(static_cast<long>(c) * nbb + bb) * BT_NB_MAX + bl] = rowbad ? __int_as_float(0x7fc00000) : scale_pow2(1.0f, e - 23);
  const int N = 6 * nb;
  for (int chunk = tid >> 2; chunk < nchunk; chunk += WD_THREADS / 4) {
    uint32_t pre[3][4], pim[3][4], pnim[3][4];               // 16 digit bytes each: w_re, w_im, -w_im
#pragma unroll
    for (int gdig = 0; gdig < 3; ++gdig)
#pragma unroll
      for (int j = 0; j < 4; ++j) pre[gdig][j] = pim[gdig][j] = pnim[gdig][j] = 0u;
    if (!rowbad) {
      const float4* src = reinterpret_cast<const float4*>(&wtile[bq * row_f2 + chunk * WD_CHUNK_F2]);
#pragma unroll
      for (int k2 = 0; k2 < 8; ++k2) {
        const float4 x2 = src[k2];
        const float xs[4] = {x2.x, x2.y, x2.z, x2.w};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int k = 2 * k2 + h;
          const int wr = __float2int_rn(scale_pow2(xs[2 * h], 23 - e)), wi = __float2int_rn(scale_pow2(xs[2 * h + 1], 23 - e));
          int dr[3], di[3], dn[3];
          split_digits(wr, dr);
          split_digits(wi, di);
          split_digits(-wi, dn);
#pragma unroll
          for (int gdig = 0; gdig < 3; ++gdig) {
            pre[gdig][k >> 2] |= static_cast<uint32_t>(dr[gdig] & 0xFF) << (8 * (k & 3));
            pim[gdig][k >> 2] |= static_cast<uint32_t>(di[gdig] & 0xFF) << (8 * (k & 3));
            pnim[gdig][k >> 2] |= static_cast<uint32_t>(dn[gdig] & 0xFF) << (8 * (k & 3));
          }
        }
      }
    }
    const int s = chunk >> 2, kq = chunk & 3;                // stage, 16-input chunk inside the stage
    const int jre = kq >> 1, jim = 2 + (kq >> 1), k16 = kq & 1;
    int8_t* stage = dig + ((static_cast<long>(c) * nbb + bb) * nst + s) * (static_cast<long>(N) * 2 * BT_KIN);
#pragma unroll
    for (int gdig = 0; gdig < 3; ++gdig) {
      const int n_re = gdig * 2 * nb + 2 * bl;               // rows n_re (out_re) and n_re + 1 (out_im): 32 adjacent bytes
      uint4* pj = reinterpret_cast<uint4*>(stage + jre * (N * 32) + (n_re >> 3) * 256 + k16 * 128 + (n_re & 7) * 16);
      pj[0] = make_uint4(pre[gdig][0], pre[gdig][1], pre[gdig][2], pre[gdig][3]);        // out_re += v_re * w_re
      pj[1] = make_uint4(pim[gdig][0], pim[gdig][1], pim[gdig][2], pim[gdig][3]);        // out_im += v_re * w_im
      uint4* pi = reinterpret_cast<uint4*>(stage + jim * (N * 32) + (n_re >> 3) * 256 + k16 * 128 + (n_re & 7) * 16);
      pi[0] = make_uint4(pnim[gdig][0], pnim[gdig][1], pnim[gdig][2], pnim[gdig][3]);    // out_re += v_im * (-w_im)
      pi[1] = make_uint4(pre[gdig][0], pre[gdig][1], pre[gdig][2], pre[gdig][3]);        // out_im += v_im * w_re
    }
  }
}

}  // namespace

bool beam_tc_geometry(int ninput, int nchan, int nbeam, BeamTcGeom* g) {
  if (ninput <= 0 || nchan <= 0 || nbeam <= 0) return false;
  if (ninput % 16 != 0 || ninput > 2032) return false;     // 16-byte loads; |digit sums| < 2^22 (digit_sum_to_float)
  g->nb = nbeam >= BT_NB_MAX ? BT_NB_MAX : ((nbeam + 7) / 8) * 8;
  g->nbb = (nbeam + g->nb - 1) / g->nb;
  g->nst = (ninput + BT_KIN - 1) / BT_KIN;
  g->dig_bytes = static_cast<size_t>(nchan) * g->nbb * g->nst * (6 * g->nb * 2 * BT_KIN);
  g->scale_floats = static_cast<size_t>(nchan) * g->nbb * BT_NB_MAX;
  return true;
}

bool beam_tc_supported(const BeamArgs& a) {
  BeamTcGeom g;
  if (!beam_tc_geometry(a.ninput, a.nchan, a.nbeam, &g)) return false;
  if (reinterpret_cast<uintptr_t>(a.in) & 15) return false;
  if (a.pow_out && (a.ntime_sum <= 0 || a.ntime_sum > BT_M || (a.nbeam & 1))) return false;
  return true;
}

cudaError_t launch_beam_weight_digits(const float2* w, int8_t* dig, float* scale, int nchan, int nbeam, int ninput,
                                      cudaStream_t stream) {
  BeamTcGeom g;
  if (!beam_tc_geometry(ninput, nchan, nbeam, &g)) return cudaErrorInvalidValue;
  if (reinterpret_cast<uintptr_t>(w) & 15) return cudaErrorInvalidValue;     // float4 loads of the weight rows
  dim3 grid(nchan, g.nbb * g.nb / 4);
  const size_t smem = static_cast<size_t>(4) * g.nst * 4 * WD_CHUNK_F2 * sizeof(float2);     // <= 73.7 KB (ninput 2032)
  cudaError_t ce = configure_once_per_device([]() -> cudaError_t {
    return cudaFuncSetAttribute(beam_weight_digits_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
  }, kCfgBeamDigits);
  if (ce != cudaSuccess) return ce;
  beam_weight_digits_kernel<<<grid, WD_THREADS, smem, stream>>>(w, dig, scale, nchan, nbeam, ninput, g.nb, g.nbb, g.nst);
  return cudaGetLastError();
}

cudaError_t launch_beamform_tc(const BeamArgs& b, const int8_t* dig, const float* scale, int nsm, cudaStream_t stream) {
  BeamTcGeom g;
  if (!beam_tc_geometry(b.ninput, b.nchan, b.nbeam, &g)) return cudaErrorInvalidValue;
  cudaError_t e = configure_once_per_device([]() -> cudaError_t {
    cudaError_t e2 = cudaFuncSetAttribute(beamform_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, BT_SMEM);
    if (e2 == cudaSuccess)
      e2 = cudaFuncSetAttribute(beamform_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BT_SMEM);
    return e2;
  }, kCfgBeamTc);
  if (e != cudaSuccess) return e;
  if (b.ntime <= 0) return cudaSuccess;
  BeamTcArgs a;
  a.in = b.in; a.wdig = dig; a.wscale = scale; a.out = b.out; a.pow_out = b.pow_out;
  a.ntime = b.ntime; a.nchan = b.nchan; a.ninput = b.ninput; a.nbeam = b.nbeam;
  a.nb = g.nb; a.nbb = g.nbb; a.nst = g.nst;
  a.ntime_sum = b.ntime_sum;
  if (b.pow_out) {
    a.tile_t = (BT_M / b.ntime_sum) * b.ntime_sum;
    a.nt_valid = (b.ntime / b.ntime_sum) * b.ntime_sum;      // whole integration blocks only
  } else {
    a.tile_t = BT_M;
    a.nt_valid = b.ntime;
  }
  if (a.nt_valid <= 0) return cudaSuccess;
  a.ntt = (a.nt_valid + a.tile_t - 1) / a.tile_t;
  const long ntiles = static_cast<long>(a.ntt) * a.nchan * a.nbb;
  if (ntiles > 0x7fffffffL) return cudaErrorInvalidValue;
  a.ntiles = static_cast<int>(ntiles);
  const int grid = a.ntiles < nsm ? a.ntiles : nsm;
  TensorMapEncodeFn encode = nullptr;
  e = tensor_map_encoder(&encode);
  if (e != cudaSuccess) return e;
  alignas(64) CUtensorMap tmap;
  const cuuint64_t gdim[3] = {static_cast<cuuint64_t>(a.ninput), static_cast<cuuint64_t>(a.nchan),
                              static_cast<cuuint64_t>(a.nt_valid)};       // rows past nt_valid read as zero
  const cuuint64_t gstride[2] = {static_cast<cuuint64_t>(a.ninput),
                                 static_cast<cuuint64_t>(a.ninput) * static_cast<cuuint64_t>(a.nchan)};
  const cuuint32_t box[3] = {BT_KIN, 1, BT_M};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t*>(a.in), gdim, gstride, box, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
  if (b.pow_out) beamform_tc_kernel<true><<<grid, BT_THREADS, BT_SMEM, stream>>>(tmap, a);
  else beamform_tc_kernel<false><<<grid, BT_THREADS, BT_SMEM, stream>>>(tmap, a);
  return cudaGetLastError();
}

}  // namespace xeng
