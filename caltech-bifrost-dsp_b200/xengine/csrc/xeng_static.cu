// Beamform + BeamformSum for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc): the two-kernel form of xeng_bf_run_integrate (fp32 beamformer -> beam voltages -> power
// integration) with the intermediate beam voltages in a file-scope __device__ array instead of the handle's
// cudaMalloc'ed buffer.  No counterpart in the reference; see include/xeng.h (xeng_beam_power_static).
//
// The array is part of the library image: every GPU that loads libxeng.so holds XENG_STATIC_SCRATCH_BYTES of it until
// the process exits.  Every call on a device shares it and finds the previous call's voltages in it; nothing has to
// start at zero (the beamformer writes every voltage the integration reads), and two calls that overlap on the GPU
// (different streams, no event between them) race on it.
#include <mutex>

#include "../../../include/xeng.h"
#include "xeng_internal.h"

namespace xeng {
namespace {

constexpr size_t kStaticVolt = XENG_STATIC_SCRATCH_FLOAT2;                 // complex beam voltages the array holds
static __device__ __align__(256) float2 g_static_volt[kStaticVolt];        // 16 MiB

// In host code `g_static_volt` is the host-side shadow of the symbol, not a device address: the address a kernel may
// use comes from cudaGetSymbolAddress, asked once per device with that device current.
cudaError_t static_volt_address(float2** out) {
  static std::mutex mu;
  static float2* addr[kCfgMaxDevices] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kCfgMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_volt);
    if (e != cudaSuccess) return e;
    addr[dev] = static_cast<float2*>(p);
  }
  *out = addr[dev];
  return cudaSuccess;
}

}  // namespace

size_t beam_power_static_capacity() { return kStaticVolt; }

cudaError_t launch_beam_power_static(const uint8_t* in, const float2* w, float* pow_out, int ninput, int nchan,
                                     int ntime, int nbeam, int ntime_sum, cudaStream_t stream) {
  float2* volt = nullptr;
  cudaError_t e = static_volt_address(&volt);
  if (e != cudaSuccess) return e;
  BeamArgs a;
  a.in = in;
  a.w = w;
  a.out = volt;
  a.ntime = ntime; a.nchan = nchan; a.ninput = ninput; a.nbeam = nbeam;
  e = launch_beamform(a, stream);
  if (e != cudaSuccess) return e;
  return launch_beam_integrate(volt, pow_out, nchan, nbeam, ntime, ntime_sum, 0, nbeam / 2, stream);
}

}  // namespace xeng
