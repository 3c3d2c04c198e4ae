// Internal (non-ABI) declarations shared by the kernel translation units and xeng_api.cu.
#pragma once
#include <cstdint>
#include <mutex>
#include <cuda_runtime.h>

namespace xeng {

// Per-device one-time configuration (cudaFuncSetAttribute, __constant__ tables): function attributes and constant
// memory belong to the device's context, so a process that holds handles on several GPUs must configure each one.
// `fn` runs once per (key, current device), under a lock (two block threads may make their first launch together).
enum CfgKey { kCfgCorrSwz = 0, kCfgCorrTc, kCfgBeam, kCfgBeamTc, kCfgBeamDigits, kCfgCount };
constexpr int kCfgMaxDevices = 64;
template <class F>
inline cudaError_t configure_once_per_device(F&& fn, int key) {
  static std::mutex mu;
  static bool done[kCfgCount][kCfgMaxDevices] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kCfgMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(mu);
  if (done[key][dev]) return cudaSuccess;
  e = fn();
  if (e == cudaSuccess) done[key][dev] = true;
  return e;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point lookup (no -lcuda): process-wide, fetched once
#ifdef CUDA_VERSION
typedef CUresult (*TensorMapEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                      const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                      CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline cudaError_t tensor_map_encoder(TensorMapEncodeFn* out) {
  static TensorMapEncodeFn encode = nullptr;
  static std::once_flag once;
  static cudaError_t err = cudaSuccess;
  std::call_once(once, []() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    err = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (err == cudaSuccess && (qres != cudaDriverEntryPointSuccess || !fn)) err = cudaErrorNotSupported;
    if (err == cudaSuccess) encode = reinterpret_cast<TensorMapEncodeFn>(fn);
  });
  *out = encode;
  return err;
}
#endif

struct CorrArgs {
  const uint8_t* in;    // [ntime][nchan][ninput] 4+4-bit complex
  int32_t* out;         // [2][nchan][words_per_chan]
  int* err;             // device watchdog word (0 = ok)
  int ntime, nchan, ninput;
  int npairs;           // nstand / 2 (station pairs)
  int tiles_per_chan, ntiles;
  int step_c, step_r;   // grid / tiles_per_chan, grid % tiles_per_chan (filled by the TMA kernel's launcher)
  int accumulate;       // 0: out = X, 1: out += X
  int ablate;           // profiling ablation mask (XENG_CORR_ABLATE, results invalid when != 0)
  long long* trace;     // developer timeline buffer (XENG_CORR_TRACE), null in production
  long words_per_chan;  // (nstand/2+1)*(nstand/4)*16
  long quad_cells;      // (nstand/2+1)*(nstand/4): cells per xGPU quadrant
  // fused CorrSubsel (TMA kernel only; sub_out == nullptr: off).  For tile-in-channel r and epilogue warp w the entries
  // sub_entries[sub_begin[r*8+w] .. sub_begin[r*8+w+1]) name the selected words that warp stages:
  // bits 0..13 byte offset in the warp's staging buffer, bit 14 conjugate, bits 15.. index v of the selection.
  const int* sub_begin;
  const uint32_t* sub_entries;
  int32_t* sub_out;     // [nchan / sub_nsum][sub_nvis][2], zeroed by the launcher, accumulated with red.global.add
  int sub_nvis, sub_nsum;
};

int corr_tc_smem_bytes();
cudaError_t launch_corr_tc(const CorrArgs& a, int nsm, cudaStream_t stream);
cudaError_t launch_corr_simt(const CorrArgs& a, cudaStream_t stream);
// TMA-fed tensor-core kernel (xeng_corr_tma.cu): needs ninput % 16 == 0 and 16-byte aligned buffers; the default
bool corr_tma_supported(const CorrArgs& a);
cudaError_t launch_corr_swz(const CorrArgs& a, int nsm, cudaStream_t stream);
// staging-buffer geometry of the TMA kernel's epilogue, for the host-side sub-select table (xeng_api.cu)
int corr_swz_staging_offset(int A1loc, int b1, int p1, int a0, int p0);

// bandwidth kernels (xeng_vis.cu)
cudaError_t launch_acc(int32_t* acc, const int32_t* in, size_t n, int first, cudaStream_t stream);
cudaError_t launch_peer_push(void* dst, void* src, size_t nbytes, long long* dst_seq, long long seq, int nsm,
                             int clear_src, cudaStream_t stream);
cudaError_t launch_subsel(const int32_t* in, int32_t* out, const int32_t* vismap, const int32_t* conj, int nvis,
                          int nchan, int nchan_sum, long W, int layout, cudaStream_t stream);
cudaError_t launch_reorder(const int32_t* in, int32_t* out, const int32_t* a2bl, const int32_t* conj, int nstand,
                           int npol, int nchan, long W, int layout, cudaStream_t stream);

// beamformer (xeng_beam.cu)
struct BeamArgs {
  const uint8_t* in;   // [ntime][nchan][ninput]
  const float2* w;     // [nchan][nbeam][ninput]
  float2* out;         // [nchan][nbeam][ntime]
  int ntime, nchan, ninput, nbeam;
  float* pow_out = nullptr;   // non-null: fused power integration, [nbeam/2][ntime/ntime_sum][nchan][4]; `out` unused
  int ntime_sum = 0, tile_t = 0;
};
// tensor-core beamformer (xeng_beam_tc.cu): int8 digit planes of the weights x int8 voltages, exact int32 sums
struct BeamTcGeom {
  int nb, nbb, nst;              // beams per block, beam blocks, 64-input K stages
  size_t dig_bytes, scale_floats;
};
bool beam_tc_geometry(int ninput, int nchan, int nbeam, BeamTcGeom* g);
bool beam_tc_supported(const BeamArgs& a);
cudaError_t launch_beam_weight_digits(const float2* w, int8_t* dig, float* scale, int nchan, int nbeam, int ninput,
                                      cudaStream_t stream);
cudaError_t launch_beamform_tc(const BeamArgs& a, const int8_t* dig, const float* scale, int nsm, cudaStream_t stream);
// no-allocation Beamform + BeamformSum (xeng_static.cu): beam voltages in a static __device__ array
size_t beam_power_static_capacity();
cudaError_t launch_beam_power_static(const uint8_t* in, const float2* w, float* pow_out, int ninput, int nchan,
                                     int ntime, int nbeam, int ntime_sum, cudaStream_t stream);
#ifdef __CUDACC__
// [|X|^2, |Y|^2, Re X Y*, Im X Y*] of one sample (beamformer_sum_test.py:64-77).  Explicit fma/mul so that every
// kernel forming these products (stand-alone integrate, both fused beamformer epilogues) rounds identically.
__device__ __forceinline__ float4 beam_products(float2 x, float2 y) {
  return make_float4(__fmaf_rn(x.y, x.y, __fmul_rn(x.x, x.x)), __fmaf_rn(y.y, y.y, __fmul_rn(y.x, y.x)),
                     __fmaf_rn(x.y, y.y, __fmul_rn(x.x, y.x)), __fmaf_rn(x.y, y.x, -__fmul_rn(x.x, y.y)));
}
#endif
cudaError_t launch_snap2_assemble(const uint8_t* pkts, int npkt, long pkt_stride, uint8_t* out, unsigned long long seq0,
                                  int ntime, int nchan, int ninput, int first_chan, int* nvalid, cudaStream_t stream);
cudaError_t launch_beamform(const BeamArgs& a, cudaStream_t stream);
int beamform_fused_max_ntime_sum();
cudaError_t launch_vlbi_tap(const float2* in, float2* out, int nchan, int nbeam, int ntime, int nsend,
                            cudaStream_t stream);
cudaError_t launch_beam_integrate(const float2* in, float* out, int nchan, int nbeam, int ntime, int ntime_sum,
                                  int beam0, int nbeam_out, cudaStream_t stream);

// offline upchannelise path (xeng_upchan.cu)
cudaError_t launch_upchan_fft(const uint8_t* in, float2* out, long nframe, int nfine, int nchan, int ninput, int layout,
                              cudaStream_t stream);
cudaError_t launch_upchan_beamform(const uint8_t* in, const float2* w, float2* out, long nframe, int nfine, int nchan,
                                   int ninput, int nbeam, cudaStream_t stream);
cudaError_t launch_upchan_correlate(const float2* in, float2* out, int nframe, int nfreq, int ninput, int first,
                                    cudaStream_t stream);

}  // namespace xeng
