// Internal (non-ABI) declarations shared by the kernel translation units and xeng_api.cu.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace xeng {

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
};

int corr_tc_smem_bytes();
cudaError_t launch_corr_tc(const CorrArgs& a, int nsm, cudaStream_t stream);
cudaError_t launch_corr_simt(const CorrArgs& a, cudaStream_t stream);
// TMA-fed tensor-core kernel (xeng_corr_tma.cu): needs ninput % 16 == 0 and 16-byte aligned buffers; the default
bool corr_tma_supported(const CorrArgs& a);
cudaError_t launch_corr_swz(const CorrArgs& a, int nsm, cudaStream_t stream);

// bandwidth kernels (xeng_vis.cu)
cudaError_t launch_acc(int32_t* acc, const int32_t* in, size_t n, int first, cudaStream_t stream);
cudaError_t launch_subsel(const int32_t* in, int32_t* out, const int32_t* vismap, const int32_t* conj, int nvis,
                          int nchan, int nchan_sum, long words_per_chan, cudaStream_t stream);
cudaError_t launch_reorder(const int32_t* in, int32_t* out, const int32_t* a2bl, const int32_t* conj, int nstand,
                           int npol, int nchan, long words_per_chan, cudaStream_t stream);

// beamformer (xeng_beam.cu)
struct BeamArgs {
  const uint8_t* in;   // [ntime][nchan][ninput]
  const float2* w;     // [nchan][nbeam][ninput]
  float2* out;         // [nchan][nbeam][ntime]
  int ntime, nchan, ninput, nbeam;
};
cudaError_t launch_beamform(const BeamArgs& a, cudaStream_t stream);
cudaError_t launch_beam_integrate(const float2* in, float* out, int nchan, int nbeam, int ntime, int ntime_sum,
                                  int beam0, int nbeam_out, cudaStream_t stream);

}  // namespace xeng
