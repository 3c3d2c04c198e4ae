/*
 * libxeng -- C ABI of the B200-native LWA-352 X-engine hot path.
 *
 * This is the drop-in boundary: every entry point replaces one of the ctypes `_bf.*` calls the
 * reference blocks make (paths below are relative to /root/reference/pipeline/lwa352_pipeline/blocks).
 * Plain pointers and sizes only; the caller owns every buffer, the library owns opaque handles.
 * All functions return 0 (XENG_OK) on success and a negative code on failure; the message of the most
 * recent failure on the calling thread is returned by xeng_last_error().  No exceptions cross this
 * boundary.  Work is enqueued on the handle's CUDA stream (default: the legacy default stream, which
 * is also torch's default stream); call the *_sync function before committing an output span, as the
 * reference does with stream_synchronize()/BFSync().  There is no CPU fallback anywhere in this library.
 */
#ifndef XENG_H_
#define XENG_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XENG_OK 0
#define XENG_ERR_ARG (-1)     /* invalid argument / unsupported shape */
#define XENG_ERR_CUDA (-2)    /* a CUDA runtime call or kernel failed */
#define XENG_ERR_NOMEM (-3)   /* device allocation failed */
#define XENG_ERR_WATCHDOG (-4)/* in-kernel pipeline watchdog fired */

const char* xeng_last_error(void);
int xeng_version(void);
/* Number of CUDA devices visible; <0 on error.  The library fails loudly when there is none. */
int xeng_device_count(void);

/* ------------------------------------------------------------------ Corr (xGPU replacement) */
typedef struct xeng_corr_s* xeng_corr_t;

/* Replaces bfXgpuInitialize(in, out, gpu) -- corr_block.py:253.  gpu = -1 keeps the current device.
 * npol must be 2 and nstand a multiple of 4 (xGPU register-tile order, corr_block.py:37-58). */
int xeng_corr_create(xeng_corr_t* h, int nstand, int npol, int nchan, int ntime_gulp, int gpu);
int xeng_corr_destroy(xeng_corr_t h);
int xeng_corr_set_stream(xeng_corr_t h, void* cuda_stream);
/* 0 = tcgen05 tensor-core kernel (default: TMA-fed when ninput % 16 == 0 and buffers are 16-byte aligned, else
 * register-fed), 1 = CUDA-core cross-check kernel (tests, Corr(test=True)), 2 = register-fed tcgen05. */
int xeng_corr_set_impl(xeng_corr_t h, int impl);
/* int32 words per plane: nchan*(nstand/2+1)*(nstand/4)*npol*npol*4 -- corr_block.py:231.
 * The output buffer holds 2*matlen int32: real plane, then imaginary plane. */
int64_t xeng_corr_matlen(xeng_corr_t h);

/* Replaces bfXgpuKernel(in, out, doDump) -- corr_block.py:445.
 * d_in : device u8 [ntime_gulp][nchan][nstand][npol], byte = (re<<4)|(im&0xF)  (corr_block.py:270-275)
 * d_out: device int32 [2][nchan][matlen/nchan] in xGPU register-tile order; written only when do_dump != 0.
 * Integrates internally across calls; on do_dump the integration is written to d_out and cleared.
 * Word regtile_index(in0,in1) holds sum_t conj(v[in0]) v[in1] (pipeline/verification/xgpu_test.py:102-136). */
int xeng_corr_run(xeng_corr_t h, const void* d_in, void* d_out, int do_dump);

/* Replaces bfXgpuCorrelate(in, out, doDump) (host-memory variant, pipeline/verification/xgpu_test.py:86-89):
 * same as xeng_corr_run but h_in / h_out are HOST buffers; copies are done inside, and the call
 * returns after the result (if do_dump) is in h_out. */
int xeng_corr_run_host(xeng_corr_t h, const void* h_in, void* h_out, int do_dump);

/* Fused Corr + CorrAcc step (corr_block.py:445 followed by corr_acc_block.py:299-306):
 * d_acc = X(d_in) when first != 0, else d_acc += X(d_in); the read-modify-write is done by the
 * bulk-reduce engine in L2, no intermediate visibility buffer is written. */
int xeng_corr_accumulate(xeng_corr_t h, const void* d_in, void* d_acc, int first);
/* Fused Corr + CorrAcc + CorrSubsel step (BASELINE configs[4]: corr_block.py:445, corr_acc_block.py:299-306 and
 * corr_subsel_block.py:298 in ONE launch).  xeng_corr_set_subsel takes the HOST arrays CorrSubsel.update_subsel
 * builds (corr_subsel_block.py:237-250): vismap[v] = word index inside one channel of the xGPU buffer, conj[v] != 0
 * where the imaginary part is negated; nvis = 0 clears the selection.  xeng_corr_accumulate_subsel then does
 * d_acc (=|+=) X(d_in) as xeng_corr_accumulate and puts this gulp's selection, summed over nchan_sum adjacent
 * channels, into d_sub_out: device int32 [nchan/nchan_sum][nvis][2] (re, im) -- the bfXgpuSubSelect output layout;
 * sub_first != 0: d_sub_out = sel(X), else d_sub_out += sel(X) (a fast dump that spans several gulps).
 * The selected words are taken from the epilogue's staging buffers (no per-gulp visibility matrix is written);
 * shapes the TMA kernel cannot take run Corr -> scratch, xeng_acc, xeng_subsel instead. */
int xeng_corr_set_subsel(xeng_corr_t h, const int32_t* vismap, const int32_t* conj, int nvis, int nchan_sum);
int xeng_corr_accumulate_subsel(xeng_corr_t h, const void* d_in, void* d_acc, int first, void* d_sub_out,
                                int sub_first);
int xeng_corr_sync(xeng_corr_t h);

/* Replaces bfXgpuGetOrder(antpol_to_input, antpol_to_bl, bl_is_conj) -- corr_block.py:331-333.  HOST arrays.
 * a2i [nstand][npol]; a2bl, conj [nstand][nstand][npol][npol].  a2bl = word index inside one channel of the
 * xGPU buffer, conj = 1 if that word must be conjugated to read as v(s0,p0) conj(v(s1,p1)). */
int xeng_get_order(const int32_t* a2i, int32_t* a2bl, int32_t* conj, int nstand, int npol);

/* ------------------------------------------------------------------ CorrAcc / CorrSubsel / reorder */
/* Replaces BFMap("a = b") / BFMap("a += b") -- corr_acc_block.py:299-306.  Device pointers, n int32 words. */
int xeng_acc(void* d_acc, const void* d_in, size_t n, int first, void* cuda_stream);

/* Replaces bfXgpuSubSelect(in, out, vismap, conj, nchan_sum, 0) -- corr_subsel_block.py:298.
 * d_in int32 [2][nchan][words_per_chan]; d_out int32 [nchan/nchan_sum][nvis][2] (re,im interleaved);
 * d_vismap, d_conj device int32 [nvis]. */
int xeng_subsel(const void* d_in, void* d_out, const void* d_vismap, const void* d_conj, int nvis, int nchan,
                int nchan_sum, int64_t words_per_chan, void* cuda_stream);

/* Replaces bfXgpuReorder(in, out, a2bl, conj) -- corr_output_full_block.py:669 (a CPU loop in the reference;
 * a device kernel here).  d_out int32 [nstand][nstand][npol][npol][nchan][2] = v(s0,p0) conj(v(s1,p1)). */
int xeng_reorder(const void* d_in, void* d_out, const void* d_a2bl, const void* d_conj, int nstand, int npol,
                 int nchan, int64_t words_per_chan, void* cuda_stream);

/* The same reorder, written straight into CorrOutputFull's packet payloads: one payload per baseline s0 <= s1,
 * baselines s0-major (bl = s0*nstand - s0*(s0-1)/2 + s1 - s0), dual-pol only, d_out 16-byte aligned.
 *   XENG_PAYLOAD_COR: int32 [nbl][nchan][npol][npol][2] -- what send_packets_bf() hands to the "cor" packet writer
 *                     after its transpose([0,3,1,2,4]) (corr_output_full_block.py:505-511)
 *   XENG_PAYLOAD_PY : int32 [nbl][npol][npol][nchan][2] -- reordered_data[s0, s1].tobytes() of send_packets_py()
 *                     (corr_output_full_block.py:457-460)
 * Half the bytes of xeng_reorder's full square matrix; nothing is left to transpose on the host. */
#define XENG_PAYLOAD_COR 1
#define XENG_PAYLOAD_PY 2
/* xeng_subsel written straight into CorrOutputPart's packet payloads (same XENG_PAYLOAD_* constants):
 *   XENG_PAYLOAD_COR: int32 [nvis/4][nchan/nchan_sum][4][2] -- send_packets_bf()'s dview.transpose([1,0,2,3])
 *                     (corr_output_part_block.py:378-391); needs 4 | nvis
 *   XENG_PAYLOAD_PY : big-endian int32 [nvis][nchan/nchan_sum][2] -- `dout`, dtype '>i'
 *                     (corr_output_part_block.py:429,460-463) */
int xeng_subsel_packets(const void* d_in, void* d_out, const void* d_vismap, const void* d_conj, int nvis, int nchan,
                        int nchan_sum, int64_t words_per_chan, int layout, void* cuda_stream);
int xeng_reorder_packets(const void* d_in, void* d_out, const void* d_a2bl, const void* d_conj, int nstand, int npol,
                         int nchan, int64_t words_per_chan, int layout, void* cuda_stream);

/* ------------------------------------------------------------------ ingest */
/* SNAP2 F-engine packets -> the u8 [ntime][nchan][ninput] voltage buffer, on the device.  Replaces the per-row
 * host memcpy of bifrost's UDP capture behind Capture (capture_block.py:219-305); packet = 32-byte big-endian header
 * (u64 seq, u32 sync_time, u16 npol, u16 npol_tot, u16 nchan, u16 nchan_tot, u32 chan_block_id, u32 chan0, u32 pol0)
 * + u8 [nchan][npol] (test_transmitters/test_tx.c:27-42, test_tx_vectors.py:37-47,103-109).
 * d_pkts: npkt raw packets pkt_stride bytes apart (device memory); out[seq - seq0][chan0 - first_chan + c][pol0 + p]
 * = payload[c][p].  Packets outside [seq0, seq0 + ntime) x [first_chan, first_chan + nchan) x [0, ninput) are
 * dropped; d_nvalid (device int32, optional) counts the accepted ones.  zero_first clears d_out (and d_nvalid)
 * first, so samples of lost packets read as zero. */
int xeng_snap2_assemble(const void* d_pkts, int npkt, int64_t pkt_stride, void* d_out, uint64_t seq0, int ntime,
                        int nchan, int ninput, int first_chan, int zero_first, void* d_nvalid, void* cuda_stream);

/* ------------------------------------------------------------------ offline upchannelise */
/* Replaces the stock bifrost chain transpose -> fft(axes='fine_time') [-> transpose] of the offline scripts
 * (scripts/lwa352-upchan-bf.py:94-96, scripts/lwa352-upchan-imag.py:93-98) with one kernel.
 * d_in u8 [nframe][nfine][nchan][ninput] (TrigBufSourceBlock's ci4 [time][fine_time][freq][stand][pol],
 * trigger_source_block.py:37-41); forward unnormalised DFT along fine_time, nfine in {2,4,8,16,32}, bins in natural
 * order (no fftshift).  d_out cf32: XENG_UPCHAN_FINE_LAST [nframe][nchan][ninput][nfine] (bf script),
 * XENG_UPCHAN_FREQ_MERGED [nframe][nchan][nfine][ninput] = [nframe][nchan*nfine][stand][pol] (imag script). */
#define XENG_UPCHAN_FINE_LAST 0
#define XENG_UPCHAN_FREQ_MERGED 1
int xeng_upchan_fft(const void* d_in, void* d_out, int nframe, int nfine, int nchan, int ninput, int layout,
                    void* cuda_stream);
/* The whole of lwa352-upchan-bf.py:94-113 in one kernel: upchannelise, multiply by the per-(beam, chan, input, bin)
 * weights BfOfflineWeightsBlock applies (beamform_offline_block.py:218-240: calibration x geometric coefficients,
 * constant over a gulp) and reduce over stand and pol.  d_weights cf32 [nbeam][nchan][ninput][nfine];
 * d_out cf32 [nframe][nbeam][nchan][nfine]. */
int xeng_upchan_beamform(const void* d_in, const void* d_weights, void* d_out, int nframe, int nfine, int nchan,
                         int ninput, int nbeam, void* cuda_stream);
/* The correlate half of the offline imaging script (scripts/lwa352-upchan-imag.py:93-106): stock bifrost
 * blocks.correlate(nframe_per_integration=...) on the upchannelised, frequency-merged voltages.  d_in cf32
 * [nframe][nfreq][ninput] (xeng_upchan_fft's XENG_UPCHAN_FREQ_MERGED output, or a frequency slice of it);
 * d_out cf32 [nfreq][ninput][ninput]: element (i, j), j <= i, (=|+=) sum_t v[t][f][i] * conj(v[t][f][j]) -- first != 0
 * starts an integration, first == 0 adds this gulp's frames to it.  64 x 64 tiles on or below the diagonal are written
 * (diagonal tiles whole); the rest of the upper triangle is not touched.  bifrost is an absent submodule: the
 * conjugation convention is the reference correlator's own (corr_output_full_block.py:582-591), parity unpinned. */
int xeng_upchan_correlate(const void* d_in, void* d_out, int nframe, int nfreq, int ninput, int first,
                          void* cuda_stream);

/* Plain asynchronous copy between any two addresses of the unified address space (device, pinned host, a peer
 * GPU's memory mapped into this process) on `cuda_stream`: cudaMemcpyAsync(cudaMemcpyDefault), i.e. the COPY ENGINES.
 * The gathers of finished products to rank 0 (xengine/sharding.py; the reference sends them as UDP packets,
 * corr_subsel_block.py:298-305, corr_acc_block.py:312-317) use it because a framework-level tensor copy onto a
 * peer-mapped buffer is an SM kernel, which would have to share the SMs with the persistent correlator. */
int xeng_memcpy_async(void* dst, const void* src, size_t nbytes, void* cuda_stream);

/* One per-step product to its receiver by SM stores: `nbytes` (a multiple of 16; both pointers 16-byte aligned) from
 * `src` to `dst`, then -- when dst_seq is not NULL -- the 64-bit word `seq` to `dst_seq`, released at system scope
 * after the payload.  dst / dst_seq may be a peer GPU's memory mapped into this process (torch symmetric memory):
 * this is the path's replacement for the reference's per-integration UDP sends of CorrSubsel's fast dump
 * (corr_subsel_block.py:298-305) and BeamformSum's powers (beamform_sum_beams_block.py:245-247).  Small CTAs that fit
 * beside the persistent correlator; nothing on the copy-engine queue the long dump occupies. */
int xeng_peer_push(void* dst, const void* src, size_t nbytes, void* dst_seq, long long seq, void* cuda_stream);
/* xeng_peer_push that MOVES: `src` is zero afterwards.  The fused fast dump adds into its buffer
 * (xeng_corr_accumulate_subsel with sub_first = 0), so a buffer that is moved to its receiver needs no
 * cudaMemsetAsync in front of the next correlator launch. */
int xeng_peer_move(void* dst, void* src, size_t nbytes, void* dst_seq, long long seq, void* cuda_stream);

/* ------------------------------------------------------------------ Beamform / BeamformSum */
typedef struct xeng_bf_s* xeng_bf_t;

/* Replaces bfBeamformInitialize(gpu, ninput, nchan, ntime, nbeam, ntime_blocks) -- beamform_block.py:251-253. */
int xeng_bf_create(xeng_bf_t* h, int gpu, int ninput, int nchan, int ntime, int nbeam, int ntime_blocks);
int xeng_bf_destroy(xeng_bf_t h);
int xeng_bf_set_stream(xeng_bf_t h, void* cuda_stream);
/* Replaces bfBeamformRun(in, out, weights) -- beamform_block.py:449.
 * d_in u8 [ntime][nchan][ninput]; d_weights cf32 [nchan][nbeam][ninput]; d_out cf32 [nchan][nbeam][ntime]. */
int xeng_bf_run(xeng_bf_t h, const void* d_in, void* d_out, const void* d_weights);
/* Where bfBeamformRun receives the weights with every call (beamform_block.py:449), the tensor-core kernel reads
 * them as int8 digit planes (24-bit block floating point per (channel, beam) row; csrc/xeng_beam_tc.cu).
 * xeng_bf_run(..., d_weights != NULL) converts them on the stream before every launch (drop-in semantics: the
 * caller may have rewritten the buffer, as Beamform does at beamform_block.py:431).  A caller that knows when its
 * weights change converts once with xeng_bf_set_weights and then passes d_weights = NULL to xeng_bf_run /
 * xeng_bf_run_integrate.  The d_weights buffer must stay valid until the next set_weights (the fp32 kernel reads it). */
int xeng_bf_set_weights(xeng_bf_t h, const void* d_weights);
/* 0 (default): tcgen05 int8 kernel whenever ninput % 16 == 0, ninput <= 2032 and d_in is 16-byte aligned, else the
 * fp32 FMA kernel; 1: always the fp32 FMA kernel (independent cross-check, as xeng_corr_set_impl). */
int xeng_bf_set_impl(xeng_bf_t h, int impl);
/* Replaces bfBeamformIntegrate(in, out, ntime_sum) -- beamform_sum_beams_block.py:245, beamform_sum_block.py:116.
 * d_in cf32 [nchan][nbeam][ntime]; d_out f32 [nbeam/2][ntime/ntime_sum][nchan][4]. */
int xeng_bf_integrate(xeng_bf_t h, const void* d_in, void* d_out, int ntime_sum);
/* Replaces bfBeamformIntegrateSingleBeam(in, out, ntime_sum, beam_id) -- beamform_sum_single_beam_block.py:114.
 * d_out f32 [ntime/ntime_sum][nchan][4] for dual-pol beam beam_id. */
int xeng_bf_integrate_single(xeng_bf_t h, const void* d_in, void* d_out, int ntime_sum, int beam_id);
/* Beamform's ntime_sum != None mode (beamform_block.py:216-218,249-251: bfBeamformInitialize with ntime_blocks > 0,
 * then bfBeamformRun emits integrated powers): d_out f32 [nbeam/2][ntime/ntime_sum][nchan][4].  For ntime_sum <= 128
 * this is ONE kernel: the power integration runs in the beamformer's epilogue and the voltage beams never reach
 * HBM.  With the tap on (or a longer ntime_sum) it is the beamformer followed by the integration kernel, the voltage
 * beams staying in a handle-owned scratch buffer readable through xeng_bf_voltages (a VLBI-style tap). */
int xeng_bf_run_integrate(xeng_bf_t h, const void* d_in, void* d_out, const void* d_weights, int ntime_sum);

/* No counterpart in the reference: added for callers that cannot allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Beamform followed by
 * BeamformSum on the CURRENT device, no handle: d_in 4+4-bit [ntime][nchan][ninput], d_weights cf32
 * [nchan][nbeam][ninput], d_out f32 [nbeam/2][ntime/ntime_sum][nchan][4] -- the result of xeng_bf_run_integrate with the
 * fp32 kernel.  The intermediate beam voltages live in a `static __device__` array inside the library
 * (XENG_STATIC_SCRATCH_BYTES on every GPU that loads libxeng, until the process exits); the call makes no cudaMalloc /
 * cudaFree and returns XENG_ERR_ARG when nchan*nbeam*ntime > XENG_STATIC_SCRATCH_FLOAT2.  Every call on a device
 * shares that array and finds the previous call's voltages in it (nothing needs to start at zero: the beamformer
 * writes every voltage the integration reads); calls that overlap on the GPU -- two streams with no event between
 * them -- race on it. */
#define XENG_STATIC_SCRATCH_FLOAT2 (2u << 20)                       /* complex beam voltages */
#define XENG_STATIC_SCRATCH_BYTES (8u * XENG_STATIC_SCRATCH_FLOAT2) /* 16 MiB */
int xeng_beam_power_static(const void* d_in, const void* d_weights, void* d_out, int ninput, int nchan, int ntime,
                           int nbeam, int ntime_sum, void* cuda_stream);
int xeng_bf_set_tap(xeng_bf_t h, int on);
/* Device pointer of the handle-owned cf32 [nchan][nbeam][ntime] voltage scratch (NULL until an
 * xeng_bf_run_integrate with the tap on has run). */
const void* xeng_bf_voltages(xeng_bf_t h);
/* BeamformVlbiOutput's voltage tap (beamform_vlbi_output_block.py:260-264: idata[:, 0:nsend, :] copied to the host
 * and transposed to time x chan x beam there): d_in cf32 [nchan][nbeam][ntime] -> d_out cf32 [ntime][nchan][nbeam_send],
 * on the device. */
int xeng_bf_vlbi_tap(xeng_bf_t h, const void* d_in, void* d_out, int nbeam_send);
int xeng_bf_sync(xeng_bf_t h);

#ifdef __cplusplus
}
#endif
#endif /* XENG_H_ */
