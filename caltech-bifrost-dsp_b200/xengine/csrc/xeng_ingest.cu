// Ingest side: SNAP2 F-engine packets -> the [time][chan][input] voltage buffer Corr / Beamform read.
//
// In the reference bifrost's UDP capture does this on the host, one memcpy per packet row
// (/root/reference/pipeline/lwa352_pipeline/blocks/capture_block.py:219-305 drives it; the packet is defined in
// pipeline/test_transmitters/test_tx.c:27-42 and built by test_tx_vectors.py:37-47,103-109):
//   header, big-endian: u64 seq | u32 sync_time | u16 npol | u16 npol_tot | u16 nchan | u16 nchan_tot |
//                       u32 chan_block_id | u32 chan0 | u32 pol0            (32 bytes)
//   payload: u8 [nchan][npol], one 4+4-bit complex sample per byte, one spectrum (seq) per packet.
// Here the receive thread only lands raw packets in a pinned buffer; one kernel scatters a batch of them:
//   out[seq - seq0][chan0 - first_chan + c][pol0 + p] = payload[c][p].
// Pure byte movement: 1 B read + 1 B written per sample, HBM-bound.
#include "xeng_internal.h"

namespace xeng {
namespace {

__device__ __forceinline__ uint32_t be32(const uint8_t* p) {
  return (static_cast<uint32_t>(p[0]) << 24) | (static_cast<uint32_t>(p[1]) << 16) | (static_cast<uint32_t>(p[2]) << 8) | p[3];
}
__device__ __forceinline__ uint32_t be16(const uint8_t* p) { return (static_cast<uint32_t>(p[0]) << 8) | p[1]; }

constexpr int SNAP2_HDR = 32;

// One CTA per packet.  Rows of npol bytes are contiguous on both sides; when everything is 16-byte aligned a thread
// moves 16 bytes at a time, else bytes.
__global__ void __launch_bounds__(128) snap2_assemble_kernel(const uint8_t* __restrict__ pkts, int npkt, long pkt_stride,
                                                             uint8_t* __restrict__ out, unsigned long long seq0,
                                                             int ntime, int nchan, int ninput, int first_chan,
                                                             int* __restrict__ nvalid) {
  const int pk = blockIdx.x;
  if (pk >= npkt) return;
  const uint8_t* h = pkts + static_cast<long>(pk) * pkt_stride;
  const unsigned long long seq = (static_cast<unsigned long long>(be32(h)) << 32) | be32(h + 4);
  const int npol = be16(h + 12), pn = be16(h + 16);
  const long chan0 = be32(h + 24), pol0 = be32(h + 28);
  // a packet outside the window being assembled (late, early, other pipeline's channels, malformed) is dropped
  if (seq < seq0 || seq - seq0 >= static_cast<unsigned long long>(ntime)) return;
  if (npol <= 0 || pn <= 0 || chan0 < first_chan || chan0 + pn > static_cast<long>(first_chan) + nchan ||
      pol0 + npol > ninput || SNAP2_HDR + static_cast<long>(pn) * npol > pkt_stride)
    return;
  if (threadIdx.x == 0 && nvalid) atomicAdd(nvalid, 1);
  const uint8_t* src = h + SNAP2_HDR;
  uint8_t* dst = out + ((seq - seq0) * nchan + (chan0 - first_chan)) * static_cast<long>(ninput) + pol0;
  const bool vec = ((npol | ninput) & 15) == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
  if (vec) {
    const int per_row = npol >> 4, n = pn * per_row;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int c = i / per_row, q = i - c * per_row;
      reinterpret_cast<uint4*>(dst + static_cast<long>(c) * ninput)[q] =
          __ldcs(reinterpret_cast<const uint4*>(src + static_cast<long>(c) * npol) + q);
    }
  } else {
    const int n = pn * npol;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int c = i / npol, q = i - c * npol;
      dst[static_cast<long>(c) * ninput + q] = src[i];
    }
  }
}

}  // namespace

cudaError_t launch_snap2_assemble(const uint8_t* pkts, int npkt, long pkt_stride, uint8_t* out, unsigned long long seq0,
                                  int ntime, int nchan, int ninput, int first_chan, int* nvalid, cudaStream_t stream) {
  if (npkt <= 0) return cudaSuccess;
  snap2_assemble_kernel<<<npkt, 128, 0, stream>>>(pkts, npkt, pkt_stride, out, seq0, ntime, nchan, ninput, first_chan,
                                                  nvalid);
  return cudaGetLastError();
}

}  // namespace xeng
