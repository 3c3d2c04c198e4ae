// libxeng C ABI (see include/xeng.h).  Host-side glue only: argument checks, handle state, launches.
// No CPU arithmetic path exists here: every entry point either launches a CUDA kernel or fails.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <new>

#include "../../../include/xeng.h"
#include "xeng_internal.h"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define XENG_CUDA(expr)                                                                       \
  do {                                                                                        \
    cudaError_t e__ = (expr);                                                                 \
    if (e__ != cudaSuccess)                                                                   \
      return fail(e__ == cudaErrorMemoryAllocation ? XENG_ERR_NOMEM : XENG_ERR_CUDA, "%s: %s", \
                  #expr, cudaGetErrorString(e__));                                            \
  } while (0)

int select_device(int gpu) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0)
    return fail(XENG_ERR_CUDA, "no CUDA device available (%s); libxeng has no CPU path",
                e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  if (gpu >= n) return fail(XENG_ERR_ARG, "gpu %d out of range (%d devices)", gpu, n);
  if (gpu >= 0) XENG_CUDA(cudaSetDevice(gpu));
  return XENG_OK;
}

}  // namespace

struct xeng_corr_s {
  int nstand, npol, nchan, ntime, ninput, device, nsm, impl;
  long words_per_chan, quad_cells;
  int64_t matlen;
  cudaStream_t stream;
  int32_t* d_acc;      // internal integration (2*matlen words), allocated on first multi-call integration
  bool acc_live;       // d_acc holds a partial integration
  int* h_err;          // mapped pinned watchdog word
  int* d_err;
  uint8_t* d_stage_in; // staging for the host-memory variant
  int32_t* d_stage_out;
  // fused CorrSubsel selection (xeng_corr_set_subsel)
  int sub_nvis, sub_nsum;
  int* d_sub_begin;          // [tiles_per_chan * 8 + 1]
  uint32_t* d_sub_entries;   // [nvis] packed (staging offset | conj << 14 | v << 15), grouped by (tile, warp)
  int32_t* d_sub_vismap;     // [nvis] device copies for the unfused fallback
  int32_t* d_sub_conj;
  int32_t* d_sub_scratch;    // per-gulp visibilities, unfused fallback only
  // developer-only profiling switches, read ONCE at create time and announced on stderr (results are invalid
  // while ablate != 0); never consulted per launch
  int ablate;
  char trace_path[256];
  long long* d_trace;
};

struct xeng_bf_s {
  int device, ninput, nchan, ntime, nbeam, ntime_blocks, nsm;
  cudaStream_t stream;
  float2* d_volt;   // scratch voltage beams for xeng_bf_run_integrate when the voltage tap is on
  int tap;          // xeng_bf_set_tap: keep the voltage beams (two kernels) instead of the fused power epilogue
  int impl;         // 0: tensor cores when the shape allows (default), 1: fp32 FMA kernel (cross-check)
  // tensor-core path: the weights as int8 digit planes + per-(channel, beam) scales (xeng_beam_tc.cu)
  int8_t* d_wdig;
  float* d_wscale;
  bool w_loaded;            // d_wdig / d_wscale hold the weights last given to set_weights / run
  const float2* w_f32;      // the caller's fp32 weights last given to set_weights (fp32 path with d_w == NULL)
};

extern "C" {

const char* xeng_last_error(void) { return g_err; }
int xeng_version(void) { return 100; }
int xeng_device_count(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) return fail(XENG_ERR_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
  return n;
}

// ------------------------------------------------------------------------------------------ Corr
int xeng_corr_create(xeng_corr_t* out, int nstand, int npol, int nchan, int ntime_gulp, int gpu) {
  if (!out) return fail(XENG_ERR_ARG, "null handle pointer");
  *out = nullptr;
  if (npol != 2) return fail(XENG_ERR_ARG, "npol must be 2 (xGPU register-tile order), got %d", npol);
  if (nstand <= 0 || (nstand & 3))   // corr_block.py:231 and :50 agree on the buffer size only when 4 | nstand
    return fail(XENG_ERR_ARG, "nstand must be a positive multiple of 4, got %d", nstand);
  if (nchan <= 0 || ntime_gulp <= 0) return fail(XENG_ERR_ARG, "nchan and ntime_gulp must be positive");
  int rc = select_device(gpu);
  if (rc) return rc;
  xeng_corr_s* h = new (std::nothrow) xeng_corr_s();
  if (!h) return fail(XENG_ERR_NOMEM, "host allocation failed");
  memset(h, 0, sizeof(*h));
  h->nstand = nstand; h->npol = npol; h->nchan = nchan; h->ntime = ntime_gulp;
  h->ninput = nstand * npol;
  h->quad_cells = static_cast<long>(nstand / 2 + 1) * nstand / 4;
  h->words_per_chan = h->quad_cells * 16;
  h->matlen = static_cast<int64_t>(nchan) * h->words_per_chan;
  cudaError_t e = cudaGetDevice(&h->device);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&h->nsm, cudaDevAttrMultiProcessorCount, h->device);
  if (e == cudaSuccess) e = cudaHostAlloc(reinterpret_cast<void**>(&h->h_err), sizeof(int), cudaHostAllocMapped);
  if (e == cudaSuccess) {
    *h->h_err = 0;
    e = cudaHostGetDevicePointer(reinterpret_cast<void**>(&h->d_err), h->h_err, 0);
  }
  if (e != cudaSuccess) {
    if (h->h_err) cudaFreeHost(h->h_err);
    delete h;
    return fail(XENG_ERR_CUDA, "xeng_corr_create: %s", cudaGetErrorString(e));
  }
  if (const char* ab = getenv("XENG_CORR_ABLATE")) h->ablate = atoi(ab);
  if (const char* tp = getenv("XENG_CORR_TRACE")) snprintf(h->trace_path, sizeof(h->trace_path), "%s", tp);
  if (h->ablate || h->trace_path[0])
    fprintf(stderr, "libxeng: DEVELOPER PROFILING MODE on this Corr handle (XENG_CORR_ABLATE=%d, XENG_CORR_TRACE=%s): "
                    "visibilities are INVALID while the ablation mask is non-zero\n", h->ablate,
            h->trace_path[0] ? h->trace_path : "");
  *out = h;
  return XENG_OK;
}

int xeng_corr_destroy(xeng_corr_t h) {
  if (!h) return XENG_OK;
  cudaSetDevice(h->device);
  if (h->d_acc) cudaFree(h->d_acc);
  if (h->d_stage_in) cudaFree(h->d_stage_in);
  if (h->d_stage_out) cudaFree(h->d_stage_out);
  if (h->d_sub_begin) cudaFree(h->d_sub_begin);
  if (h->d_sub_entries) cudaFree(h->d_sub_entries);
  if (h->d_sub_vismap) cudaFree(h->d_sub_vismap);
  if (h->d_sub_conj) cudaFree(h->d_sub_conj);
  if (h->d_sub_scratch) cudaFree(h->d_sub_scratch);
  if (h->d_trace) cudaFree(h->d_trace);
  if (h->h_err) cudaFreeHost(h->h_err);
  delete h;
  return XENG_OK;
}

int xeng_corr_set_stream(xeng_corr_t h, void* s) {
  if (!h) return fail(XENG_ERR_ARG, "null handle");
  h->stream = static_cast<cudaStream_t>(s);
  return XENG_OK;
}
int xeng_corr_set_impl(xeng_corr_t h, int impl) {
  if (!h) return fail(XENG_ERR_ARG, "null handle");
  if (impl < 0 || impl > 2)
    return fail(XENG_ERR_ARG, "impl must be 0 (tcgen05, auto: TMA-fed when the shape allows), 1 (CUDA-core cross-check) or 2 (register-fed tcgen05)");
  h->impl = impl;
  return XENG_OK;
}
int64_t xeng_corr_matlen(xeng_corr_t h) { return h ? h->matlen : 0; }

static int corr_launch(xeng_corr_t h, const void* d_in, int32_t* d_out, int accumulate, int32_t* d_sub_out = nullptr) {
  xeng::CorrArgs a;
  a.in = static_cast<const uint8_t*>(d_in);
  a.out = d_out;
  a.err = h->d_err;
  a.ntime = h->ntime; a.nchan = h->nchan; a.ninput = h->ninput;
  a.npairs = h->nstand / 2;
  const int T = (h->ninput + 127) / 128;
  a.tiles_per_chan = T * (T + 1) / 2;
  a.ntiles = a.tiles_per_chan * h->nchan;
  a.accumulate = accumulate;
  a.ablate = h->ablate;
  a.trace = nullptr;
  a.sub_begin = h->d_sub_begin; a.sub_entries = h->d_sub_entries; a.sub_out = d_sub_out;
  a.sub_nvis = h->sub_nvis; a.sub_nsum = h->sub_nsum;
  const char* trace_path = h->trace_path[0] ? h->trace_path : nullptr;   // developer-only: per-role timeline of CTA 0
  const size_t trace_bytes = 5 * 4096 * 2 * sizeof(long long);
  if (trace_path && a.ablate) {
    if (!h->d_trace) XENG_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->d_trace), trace_bytes));
    XENG_CUDA(cudaMemsetAsync(h->d_trace, 0, trace_bytes, h->stream));
    a.trace = h->d_trace;
  }
  a.words_per_chan = h->words_per_chan;
  a.quad_cells = h->quad_cells;
  if ((reinterpret_cast<uintptr_t>(d_out) & 15) != 0) return fail(XENG_ERR_ARG, "visibility buffer must be 16-byte aligned");
  if (h->impl == 1) XENG_CUDA(xeng::launch_corr_simt(a, h->stream));
  else if (h->impl == 0 && xeng::corr_tma_supported(a)) XENG_CUDA(xeng::launch_corr_swz(a, h->nsm, h->stream));
  else XENG_CUDA(xeng::launch_corr_tc(a, h->nsm, h->stream));   // register-fed tcgen05 kernel: any ninput / alignment
  if (a.trace) {
    XENG_CUDA(cudaStreamSynchronize(h->stream));
    long long* host = static_cast<long long*>(malloc(trace_bytes));
    if (host) {
      XENG_CUDA(cudaMemcpy(host, h->d_trace, trace_bytes, cudaMemcpyDeviceToHost));
      FILE* f = fopen(trace_path, "wb");
      if (f) { fwrite(host, 1, trace_bytes, f); fclose(f); }
      free(host);
    }
  }
  return XENG_OK;
}

static int corr_check(xeng_corr_t h, const void* in, const void* out, bool need_out) {
  if (!h) return fail(XENG_ERR_ARG, "null handle");
  if (!in) return fail(XENG_ERR_ARG, "null input pointer");
  if (need_out && !out) return fail(XENG_ERR_ARG, "null output pointer");
  XENG_CUDA(cudaSetDevice(h->device));
  return XENG_OK;
}

int xeng_corr_run(xeng_corr_t h, const void* d_in, void* d_out, int do_dump) {
  int rc = corr_check(h, d_in, d_out, do_dump != 0);
  if (rc) return rc;
  const size_t bytes = static_cast<size_t>(2 * h->matlen) * sizeof(int32_t);
  if (do_dump && !h->acc_live)  // single-gulp integration: straight into the caller's buffer
    return corr_launch(h, d_in, static_cast<int32_t*>(d_out), 0);
  if (!h->d_acc) XENG_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->d_acc), bytes));
  rc = corr_launch(h, d_in, h->d_acc, h->acc_live ? 1 : 0);
  if (rc) return rc;
  h->acc_live = true;
  if (do_dump) {
    XENG_CUDA(cudaMemcpyAsync(d_out, h->d_acc, bytes, cudaMemcpyDeviceToDevice, h->stream));
    h->acc_live = false;  // next call overwrites: this is xGPU's "clear on dump"
  }
  return XENG_OK;
}

int xeng_corr_accumulate(xeng_corr_t h, const void* d_in, void* d_acc, int first) {
  int rc = corr_check(h, d_in, d_acc, true);
  if (rc) return rc;
  return corr_launch(h, d_in, static_cast<int32_t*>(d_acc), first ? 0 : 1);
}

static bool corr_uses_tma(xeng_corr_t h, const void* d_in, const void* d_out) {
  xeng::CorrArgs a;
  a.in = static_cast<const uint8_t*>(d_in);
  a.out = static_cast<int32_t*>(const_cast<void*>(d_out));
  a.ninput = h->ninput;
  return h->impl == 0 && xeng::corr_tma_supported(a);
}

int xeng_corr_set_subsel(xeng_corr_t h, const int32_t* vismap, const int32_t* conj, int nvis, int nchan_sum) {
  if (!h) return fail(XENG_ERR_ARG, "null handle");
  XENG_CUDA(cudaSetDevice(h->device));
  XENG_CUDA(cudaStreamSynchronize(h->stream));       // a launch may still be reading the old tables
  if (h->d_sub_begin) { cudaFree(h->d_sub_begin); h->d_sub_begin = nullptr; }
  if (h->d_sub_entries) { cudaFree(h->d_sub_entries); h->d_sub_entries = nullptr; }
  if (h->d_sub_vismap) { cudaFree(h->d_sub_vismap); h->d_sub_vismap = nullptr; }
  if (h->d_sub_conj) { cudaFree(h->d_sub_conj); h->d_sub_conj = nullptr; }
  h->sub_nvis = 0; h->sub_nsum = 0;
  if (nvis == 0) return XENG_OK;
  if (!vismap || !conj) return fail(XENG_ERR_ARG, "null pointer");
  if (nvis < 0 || nvis >= (1 << 17)) return fail(XENG_ERR_ARG, "nvis must be in [0, 131072), got %d", nvis);
  if (nchan_sum <= 0 || h->nchan % nchan_sum) return fail(XENG_ERR_ARG, "nchan_sum must divide nchan");
  const int T = (h->ninput + 127) / 128, tpc = T * (T + 1) / 2, nlist = tpc * 8;
  int* begin = static_cast<int*>(calloc(nlist + 1, sizeof(int)));
  int* key = static_cast<int*>(malloc(sizeof(int) * nvis));
  uint32_t* packed = static_cast<uint32_t*>(malloc(sizeof(uint32_t) * nvis));
  uint32_t* sorted = static_cast<uint32_t*>(malloc(sizeof(uint32_t) * nvis));
  if (!begin || !key || !packed || !sorted) { free(begin); free(key); free(packed); free(sorted); return fail(XENG_ERR_NOMEM, "host allocation failed"); }
  int rc = XENG_OK;
  for (int v = 0; v < nvis && rc == XENG_OK; ++v) {
    const long w = vismap[v];
    if (w < 0 || w >= h->words_per_chan) { rc = fail(XENG_ERR_ARG, "vismap[%d] = %ld outside [0, %ld)", v, w, h->words_per_chan); break; }
    const long cell = w >> 2, quad = cell / h->quad_cells, t = cell % h->quad_cells;
    const int p1 = static_cast<int>((w >> 1) & 1), p0 = static_cast<int>(w & 1);
    const int b0 = static_cast<int>(quad >> 1), b1 = static_cast<int>(quad & 1);
    long A1 = static_cast<long>((sqrt(8.0 * static_cast<double>(t) + 1.0) - 1.0) * 0.5);
    while ((A1 + 1) * (A1 + 2) / 2 <= t) ++A1;
    while (A1 * (A1 + 1) / 2 > t) --A1;
    const long A0 = t - A1 * (A1 + 1) / 2;
    const int I = static_cast<int>(A1 / 32), J = static_cast<int>(A0 / 32), A1loc = static_cast<int>(A1 % 32);
    const int list = (I * (I + 1) / 2 + J) * 8 + b0 * 4 + A1loc / 8;
    key[v] = list;
    packed[v] = static_cast<uint32_t>(xeng::corr_swz_staging_offset(A1loc, b1, p1, static_cast<int>(A0 % 32), p0)) |
                (conj[v] ? (1u << 14) : 0u) | (static_cast<uint32_t>(v) << 15);
    ++begin[list + 1];
  }
  if (rc == XENG_OK) {
    for (int i = 0; i < nlist; ++i) begin[i + 1] += begin[i];
    int* fill = static_cast<int*>(calloc(nlist, sizeof(int)));
    if (!fill) rc = fail(XENG_ERR_NOMEM, "host allocation failed");
    else {
      for (int v = 0; v < nvis; ++v) sorted[begin[key[v]] + fill[key[v]]++] = packed[v];
      free(fill);
    }
  }
  cudaError_t e = cudaSuccess;
  if (rc == XENG_OK) {
    e = cudaMalloc(reinterpret_cast<void**>(&h->d_sub_begin), sizeof(int) * (nlist + 1));
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&h->d_sub_entries), sizeof(uint32_t) * nvis);
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&h->d_sub_vismap), sizeof(int32_t) * nvis);
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&h->d_sub_conj), sizeof(int32_t) * nvis);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_sub_begin, begin, sizeof(int) * (nlist + 1), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_sub_entries, sorted, sizeof(uint32_t) * nvis, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_sub_vismap, vismap, sizeof(int32_t) * nvis, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_sub_conj, conj, sizeof(int32_t) * nvis, cudaMemcpyHostToDevice);
  }
  free(begin); free(key); free(packed); free(sorted);
  if (rc != XENG_OK) return rc;
  if (e != cudaSuccess) return fail(XENG_ERR_CUDA, "xeng_corr_set_subsel: %s", cudaGetErrorString(e));
  h->sub_nvis = nvis; h->sub_nsum = nchan_sum;
  return XENG_OK;
}

int xeng_corr_accumulate_subsel(xeng_corr_t h, const void* d_in, void* d_acc, int first, void* d_sub_out,
                                int sub_first) {
  int rc = corr_check(h, d_in, d_acc, true);
  if (rc) return rc;
  if (!d_sub_out) return fail(XENG_ERR_ARG, "null sub-select output pointer");
  if (h->sub_nvis <= 0) return fail(XENG_ERR_ARG, "no selection set: call xeng_corr_set_subsel first");
  int32_t* so = static_cast<int32_t*>(d_sub_out);
  if (corr_uses_tma(h, d_in, d_acc)) {   // one launch: correlate, accumulate in L2, sub-select from the staging buffers
    if (sub_first)
      XENG_CUDA(cudaMemsetAsync(so, 0, sizeof(int32_t) * 2 * h->sub_nvis * (h->nchan / h->sub_nsum), h->stream));
    return corr_launch(h, d_in, static_cast<int32_t*>(d_acc), first ? 0 : 1, so);
  }
  // shapes the TMA kernel cannot take: per-gulp visibilities into a scratch buffer, then the stand-alone kernels
  const size_t words = static_cast<size_t>(2 * h->matlen);
  if (!h->d_sub_scratch) XENG_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->d_sub_scratch), words * sizeof(int32_t)));
  rc = corr_launch(h, d_in, h->d_sub_scratch, 0);
  if (rc) return rc;
  XENG_CUDA(xeng::launch_acc(static_cast<int32_t*>(d_acc), h->d_sub_scratch, words, first ? 1 : 0, h->stream));
  if (sub_first) {
    XENG_CUDA(xeng::launch_subsel(h->d_sub_scratch, so, h->d_sub_vismap, h->d_sub_conj, h->sub_nvis, h->nchan,
                                  h->sub_nsum, h->words_per_chan, 0, h->stream));
  } else {      // sub_out += sel(X): select into the (now free) head of the scratch buffer, then add
    const size_t nsel = static_cast<size_t>(2) * h->sub_nvis * (h->nchan / h->sub_nsum);
    int32_t* tmp = nullptr;
    XENG_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&tmp), nsel * sizeof(int32_t), h->stream));
    XENG_CUDA(xeng::launch_subsel(h->d_sub_scratch, tmp, h->d_sub_vismap, h->d_sub_conj, h->sub_nvis, h->nchan,
                                  h->sub_nsum, h->words_per_chan, 0, h->stream));
    XENG_CUDA(xeng::launch_acc(so, tmp, nsel, 0, h->stream));
    XENG_CUDA(cudaFreeAsync(tmp, h->stream));
  }
  return XENG_OK;
}

int xeng_corr_sync(xeng_corr_t h) {
  if (!h) return fail(XENG_ERR_ARG, "null handle");
  cudaError_t e = cudaStreamSynchronize(h->stream);
  if (h->h_err && *h->h_err != 0)
    return fail(XENG_ERR_WATCHDOG, "corr kernel watchdog fired at barrier tag %d (%s)", *h->h_err,
                cudaGetErrorString(e));
  if (e != cudaSuccess) return fail(XENG_ERR_CUDA, "cudaStreamSynchronize: %s", cudaGetErrorString(e));
  return XENG_OK;
}

int xeng_corr_run_host(xeng_corr_t h, const void* h_in, void* h_out, int do_dump) {
  int rc = corr_check(h, h_in, h_out, do_dump != 0);
  if (rc) return rc;
  const size_t in_bytes = static_cast<size_t>(h->ntime) * h->nchan * h->ninput;
  const size_t out_bytes = static_cast<size_t>(2 * h->matlen) * sizeof(int32_t);
  if (!h->d_stage_in) XENG_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->d_stage_in), in_bytes));
  if (do_dump && !h->d_stage_out) XENG_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->d_stage_out), out_bytes));
  XENG_CUDA(cudaMemcpyAsync(h->d_stage_in, h_in, in_bytes, cudaMemcpyHostToDevice, h->stream));
  rc = xeng_corr_run(h, h->d_stage_in, h->d_stage_out, do_dump);
  if (rc) return rc;
  if (do_dump) XENG_CUDA(cudaMemcpyAsync(h_out, h->d_stage_out, out_bytes, cudaMemcpyDeviceToHost, h->stream));
  return xeng_corr_sync(h);
}

// bfXgpuGetOrder: host-only index bookkeeping (the reference computes it on the CPU as well).
int xeng_get_order(const int32_t* a2i, int32_t* a2bl, int32_t* conj, int nstand, int npol) {
  if (!a2i || !a2bl || !conj) return fail(XENG_ERR_ARG, "null pointer");
  if (npol != 2 || nstand <= 0 || (nstand & 3)) return fail(XENG_ERR_ARG, "need npol == 2 and 4 | nstand");
  const long Q = static_cast<long>(nstand / 2 + 1) * nstand / 4;
  for (int i = 0; i < nstand * npol; ++i)
    if (a2i[i] < 0 || a2i[i] >= nstand * npol) return fail(XENG_ERR_ARG, "ant_to_input[%d] = %d out of range", i, a2i[i]);
  for (int s0 = 0; s0 < nstand; ++s0)
    for (int s1 = 0; s1 < nstand; ++s1)
      for (int p0 = 0; p0 < npol; ++p0)
        for (int p1 = 0; p1 < npol; ++p1) {
          const int i0 = a2i[s0 * npol + p0], i1 = a2i[s1 * npol + p1];
          const bool stored = (i1 >> 1) >= (i0 >> 1);   // word (i0,i1) exists and holds conj(v_i0) v_i1
          const int lo = stored ? i0 : i1, hi = stored ? i1 : i0;
          const int a0 = lo >> 1, a1 = hi >> 1;
          const long tri = static_cast<long>(a1 / 2) * (a1 / 2 + 1) / 2 + a0 / 2;
          const long cell = (2 * (a0 & 1) + (a1 & 1)) * Q + tri;
          const long o = ((static_cast<long>(s0) * nstand + s1) * npol + p0) * npol + p1;
          a2bl[o] = static_cast<int32_t>(cell * 4 + 2 * (hi & 1) + (lo & 1));
          conj[o] = stored ? 1 : 0;
        }
  return XENG_OK;
}

// ------------------------------------------------------------------ CorrAcc / CorrSubsel / reorder
int xeng_acc(void* d_acc, const void* d_in, size_t n, int first, void* stream) {
  if (!d_acc || !d_in) return fail(XENG_ERR_ARG, "null pointer");
  XENG_CUDA(xeng::launch_acc(static_cast<int32_t*>(d_acc), static_cast<const int32_t*>(d_in), n, first,
                             static_cast<cudaStream_t>(stream)));
  return XENG_OK;
}

int xeng_subsel(const void* d_in, void* d_out, const void* d_vismap, const void* d_conj, int nvis, int nchan,
                int nchan_sum, int64_t words_per_chan, void* stream) {
  if (!d_in || !d_out || !d_vismap || !d_conj) return fail(XENG_ERR_ARG, "null pointer");
  if (nvis < 0 || nchan <= 0 || nchan_sum <= 0 || words_per_chan <= 0) return fail(XENG_ERR_ARG, "bad sizes");
  XENG_CUDA(xeng::launch_subsel(static_cast<const int32_t*>(d_in), static_cast<int32_t*>(d_out),
                                static_cast<const int32_t*>(d_vismap), static_cast<const int32_t*>(d_conj), nvis,
                                nchan, nchan_sum, words_per_chan, 0, static_cast<cudaStream_t>(stream)));
  return XENG_OK;
}

int xeng_subsel_packets(const void* d_in, void* d_out, const void* d_vismap, const void* d_conj, int nvis, int nchan,
                        int nchan_sum, int64_t words_per_chan, int layout, void* stream) {
  if (!d_in || !d_out || !d_vismap || !d_conj) return fail(XENG_ERR_ARG, "null pointer");
  if (nvis < 0 || nchan <= 0 || nchan_sum <= 0 || words_per_chan <= 0) return fail(XENG_ERR_ARG, "bad sizes");
  if (layout != XENG_PAYLOAD_COR && layout != XENG_PAYLOAD_PY) return fail(XENG_ERR_ARG, "unknown payload layout");
  if (layout == XENG_PAYLOAD_COR && (nvis & 3)) return fail(XENG_ERR_ARG, "COR packets hold 4 visibilities: 4 | nvis");
  XENG_CUDA(xeng::launch_subsel(static_cast<const int32_t*>(d_in), static_cast<int32_t*>(d_out),
                                static_cast<const int32_t*>(d_vismap), static_cast<const int32_t*>(d_conj), nvis,
                                nchan, nchan_sum, words_per_chan, layout, static_cast<cudaStream_t>(stream)));
  return XENG_OK;
}

int xeng_reorder(const void* d_in, void* d_out, const void* d_a2bl, const void* d_conj, int nstand, int npol,
                 int nchan, int64_t words_per_chan, void* stream) {
  if (!d_in || !d_out || !d_a2bl || !d_conj) return fail(XENG_ERR_ARG, "null pointer");
  if (nstand <= 0 || npol <= 0 || nchan <= 0) return fail(XENG_ERR_ARG, "bad sizes");
  XENG_CUDA(xeng::launch_reorder(static_cast<const int32_t*>(d_in), static_cast<int32_t*>(d_out),
                                 static_cast<const int32_t*>(d_a2bl), static_cast<const int32_t*>(d_conj), nstand,
                                 npol, nchan, words_per_chan, 0, static_cast<cudaStream_t>(stream)));
  return XENG_OK;
}

int xeng_reorder_packets(const void* d_in, void* d_out, const void* d_a2bl, const void* d_conj, int nstand, int npol,
                         int nchan, int64_t words_per_chan, int layout, void* stream) {
  if (!d_in || !d_out || !d_a2bl || !d_conj) return fail(XENG_ERR_ARG, "null pointer");
  if (nstand <= 0 || nchan <= 0) return fail(XENG_ERR_ARG, "bad sizes");
  if (npol != 2) return fail(XENG_ERR_ARG, "packet payload layouts are dual-pol only");
  if (layout != XENG_PAYLOAD_COR && layout != XENG_PAYLOAD_PY) return fail(XENG_ERR_ARG, "unknown payload layout");
  if (reinterpret_cast<uintptr_t>(d_out) & 15) return fail(XENG_ERR_ARG, "d_out must be 16-byte aligned");
  XENG_CUDA(xeng::launch_reorder(static_cast<const int32_t*>(d_in), static_cast<int32_t*>(d_out),
                                 static_cast<const int32_t*>(d_a2bl), static_cast<const int32_t*>(d_conj), nstand,
                                 npol, nchan, words_per_chan, layout, static_cast<cudaStream_t>(stream)));
  return XENG_OK;
}

// ------------------------------------------------------------------------------------ ingest
int xeng_snap2_assemble(const void* d_pkts, int npkt, int64_t pkt_stride, void* d_out, uint64_t seq0, int ntime,
                        int nchan, int ninput, int first_chan, int zero_first, void* d_nvalid, void* stream) {
  if (!d_out || (npkt > 0 && !d_pkts)) return fail(XENG_ERR_ARG, "null pointer");
  if (npkt < 0 || pkt_stride < 32 || ntime <= 0 || nchan <= 0 || ninput <= 0 || first_chan < 0)
    return fail(XENG_ERR_ARG, "bad sizes");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (zero_first) {      // samples of packets that never arrive read as zero, as bifrost's capture leaves them
    XENG_CUDA(cudaMemsetAsync(d_out, 0, static_cast<size_t>(ntime) * nchan * ninput, st));
    if (d_nvalid) XENG_CUDA(cudaMemsetAsync(d_nvalid, 0, sizeof(int), st));
  }
  XENG_CUDA(xeng::launch_snap2_assemble(static_cast<const uint8_t*>(d_pkts), npkt, pkt_stride,
                                        static_cast<uint8_t*>(d_out), seq0, ntime, nchan, ninput, first_chan,
                                        static_cast<int*>(d_nvalid), st));
  return XENG_OK;
}

// ------------------------------------------------------------------------------------ offline upchannelise
static int upchan_check(const void* a, const void* b, int nframe, int nfine, int nchan, int ninput) {
  if (!a || !b) return fail(XENG_ERR_ARG, "null pointer");
  if (nframe < 0 || nchan <= 0 || ninput <= 0) return fail(XENG_ERR_ARG, "bad sizes");
  if (nfine != 2 && nfine != 4 && nfine != 8 && nfine != 16 && nfine != 32)
    return fail(XENG_ERR_ARG, "nfine must be 2, 4, 8, 16 or 32");
  return XENG_OK;
}

int xeng_upchan_fft(const void* d_in, void* d_out, int nframe, int nfine, int nchan, int ninput, int layout,
                    void* stream) {
  int rc = upchan_check(d_in, d_out, nframe, nfine, nchan, ninput);
  if (rc) return rc;
  if (layout != XENG_UPCHAN_FINE_LAST && layout != XENG_UPCHAN_FREQ_MERGED) return fail(XENG_ERR_ARG, "bad layout");
  if (static_cast<long>(nframe) * nchan > 2147483647L) return fail(XENG_ERR_ARG, "nframe * nchan too large");
  XENG_CUDA(xeng::launch_upchan_fft(static_cast<const uint8_t*>(d_in), static_cast<float2*>(d_out), nframe, nfine,
                                    nchan, ninput, layout, static_cast<cudaStream_t>(stream)));
  return XENG_OK;
}

int xeng_upchan_beamform(const void* d_in, const void* d_weights, void* d_out, int nframe, int nfine, int nchan,
                         int ninput, int nbeam, void* stream) {
  int rc = upchan_check(d_in, d_out, nframe, nfine, nchan, ninput);
  if (rc) return rc;
  if (!d_weights) return fail(XENG_ERR_ARG, "null pointer");
  if (nbeam <= 0 || nframe > 65535) return fail(XENG_ERR_ARG, "bad sizes");
  XENG_CUDA(xeng::launch_upchan_beamform(static_cast<const uint8_t*>(d_in), static_cast<const float2*>(d_weights),
                                         static_cast<float2*>(d_out), nframe, nfine, nchan, ninput, nbeam,
                                         static_cast<cudaStream_t>(stream)));
  return XENG_OK;
}

int xeng_upchan_correlate(const void* d_in, void* d_out, int nframe, int nfreq, int ninput, int first, void* stream) {
  if (!d_in || !d_out) return fail(XENG_ERR_ARG, "null pointer");
  if (nframe < 0 || nfreq <= 0 || ninput <= 0) return fail(XENG_ERR_ARG, "bad sizes");
  if (nfreq > 65535) return fail(XENG_ERR_ARG, "at most 65535 frequencies per call, got %d", nfreq);
  if ((reinterpret_cast<uintptr_t>(d_in) & 7) || (reinterpret_cast<uintptr_t>(d_out) & 7))
    return fail(XENG_ERR_ARG, "cf32 buffers must be 8-byte aligned");
  XENG_CUDA(xeng::launch_upchan_correlate(static_cast<const float2*>(d_in), static_cast<float2*>(d_out), nframe, nfreq,
                                          ninput, first ? 1 : 0, static_cast<cudaStream_t>(stream)));
  return XENG_OK;
}

int xeng_memcpy_async(void* dst, const void* src, size_t nbytes, void* stream) {
  if (!dst || !src) return fail(XENG_ERR_ARG, "null pointer");
  if (nbytes == 0) return XENG_OK;
  XENG_CUDA(cudaMemcpyAsync(dst, src, nbytes, cudaMemcpyDefault, static_cast<cudaStream_t>(stream)));
  return XENG_OK;
}

static int peer_push_impl(void* dst, void* src, size_t nbytes, void* dst_seq, long long seq, int clear_src, void* stream);
int xeng_peer_push(void* dst, const void* src, size_t nbytes, void* dst_seq, long long seq, void* stream) {
  return peer_push_impl(dst, const_cast<void*>(src), nbytes, dst_seq, seq, 0, stream);
}
int xeng_peer_move(void* dst, void* src, size_t nbytes, void* dst_seq, long long seq, void* stream) {
  return peer_push_impl(dst, src, nbytes, dst_seq, seq, 1, stream);
}
static int peer_push_impl(void* dst, void* src, size_t nbytes, void* dst_seq, long long seq, int clear_src, void* stream) {
  if (!dst || !src) return fail(XENG_ERR_ARG, "null pointer");
  if ((nbytes & 15) || (reinterpret_cast<uintptr_t>(dst) & 15) || (reinterpret_cast<uintptr_t>(src) & 15))
    return fail(XENG_ERR_ARG, "xeng_peer_push wants 16-byte aligned pointers and a multiple of 16 bytes");
  if (dst_seq && (reinterpret_cast<uintptr_t>(dst_seq) & 7)) return fail(XENG_ERR_ARG, "dst_seq must be 8-byte aligned");
  static int nsm_of[64] = {0};
  int dev = 0;
  XENG_CUDA(cudaGetDevice(&dev));
  int nsm = (dev >= 0 && dev < 64) ? nsm_of[dev] : 0;
  if (!nsm) {
    XENG_CUDA(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
    if (dev >= 0 && dev < 64) nsm_of[dev] = nsm;
  }
  XENG_CUDA(xeng::launch_peer_push(dst, src, nbytes, static_cast<long long*>(dst_seq), seq, nsm, clear_src,
                                   static_cast<cudaStream_t>(stream)));
  return XENG_OK;
}

// ------------------------------------------------------------------------------------ Beamform
int xeng_bf_create(xeng_bf_t* out, int gpu, int ninput, int nchan, int ntime, int nbeam, int ntime_blocks) {
  if (!out) return fail(XENG_ERR_ARG, "null handle pointer");
  *out = nullptr;
  if (ninput <= 0 || nchan <= 0 || ntime <= 0 || nbeam <= 0) return fail(XENG_ERR_ARG, "sizes must be positive");
  int rc = select_device(gpu);
  if (rc) return rc;
  xeng_bf_s* h = new (std::nothrow) xeng_bf_s();
  if (!h) return fail(XENG_ERR_NOMEM, "host allocation failed");
  h->ninput = ninput; h->nchan = nchan; h->ntime = ntime; h->nbeam = nbeam; h->ntime_blocks = ntime_blocks;
  h->stream = nullptr;
  h->d_volt = nullptr;
  h->tap = 0;
  h->impl = 0;
  h->d_wdig = nullptr; h->d_wscale = nullptr; h->w_loaded = false; h->w_f32 = nullptr;
  cudaError_t e = cudaGetDevice(&h->device);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&h->nsm, cudaDevAttrMultiProcessorCount, h->device);
  if (e != cudaSuccess) {
    delete h;
    return fail(XENG_ERR_CUDA, "xeng_bf_create: %s", cudaGetErrorString(e));
  }
  *out = h;
  return XENG_OK;
}
int xeng_bf_destroy(xeng_bf_t h) {
  if (h) {
    cudaSetDevice(h->device);
    if (h->d_volt) cudaFree(h->d_volt);
    if (h->d_wdig) cudaFree(h->d_wdig);
    if (h->d_wscale) cudaFree(h->d_wscale);
  }
  delete h;
  return XENG_OK;
}
int xeng_bf_set_impl(xeng_bf_t h, int impl) {
  if (!h) return fail(XENG_ERR_ARG, "null handle");
  if (impl < 0 || impl > 1) return fail(XENG_ERR_ARG, "impl must be 0 (tcgen05 int8 digit planes when the shape allows) or 1 (fp32 FMA cross-check)");
  h->impl = impl;
  return XENG_OK;
}

// fp32 weights -> int8 digit planes (+ scales) in the handle; the digits are what the tensor-core kernel reads
static int bf_convert_weights(xeng_bf_t h, const void* d_w) {
  xeng::BeamTcGeom g;
  if (!xeng::beam_tc_geometry(h->ninput, h->nchan, h->nbeam, &g)) return XENG_OK;   // fp32 path only for this shape
  if (!h->d_wdig) {
    XENG_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->d_wdig), g.dig_bytes));
    XENG_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->d_wscale), g.scale_floats * sizeof(float)));
  }
  XENG_CUDA(xeng::launch_beam_weight_digits(static_cast<const float2*>(d_w), h->d_wdig, h->d_wscale, h->nchan,
                                            h->nbeam, h->ninput, h->stream));
  return XENG_OK;
}

int xeng_bf_set_weights(xeng_bf_t h, const void* d_w) {
  if (!h || !d_w) return fail(XENG_ERR_ARG, "null pointer");
  XENG_CUDA(cudaSetDevice(h->device));
  int rc = bf_convert_weights(h, d_w);
  if (rc) return rc;
  h->w_loaded = true;
  h->w_f32 = static_cast<const float2*>(d_w);
  return XENG_OK;
}

// One beamformer launch: tensor cores when the shape allows, else (or impl == 1) the fp32 FMA kernel.
// d_w != NULL: the reference's calling convention (weights with every call) -- converted on the stream first.
// d_w == NULL: the weights last loaded with xeng_bf_set_weights.
static int bf_launch(xeng_bf_t h, xeng::BeamArgs& a, const void* d_w) {
  if (h->impl == 0 && xeng::beam_tc_supported(a)) {
    if (d_w) {
      int rc = bf_convert_weights(h, d_w);
      if (rc) return rc;
      h->w_loaded = true;
      h->w_f32 = static_cast<const float2*>(d_w);
    } else if (!h->w_loaded) {
      return fail(XENG_ERR_ARG, "no weights: pass d_w or call xeng_bf_set_weights first");
    }
    XENG_CUDA(xeng::launch_beamform_tc(a, h->d_wdig, h->d_wscale, h->nsm, h->stream));
    return XENG_OK;
  }
  a.w = d_w ? static_cast<const float2*>(d_w) : h->w_f32;
  if (!a.w) return fail(XENG_ERR_ARG, "no weights: pass d_w or call xeng_bf_set_weights first");
  XENG_CUDA(xeng::launch_beamform(a, h->stream));
  return XENG_OK;
}
int xeng_bf_set_stream(xeng_bf_t h, void* s) {
  if (!h) return fail(XENG_ERR_ARG, "null handle");
  h->stream = static_cast<cudaStream_t>(s);
  return XENG_OK;
}

int xeng_bf_run(xeng_bf_t h, const void* d_in, void* d_out, const void* d_w) {
  if (!h || !d_in || !d_out) return fail(XENG_ERR_ARG, "null pointer");
  XENG_CUDA(cudaSetDevice(h->device));
  xeng::BeamArgs a;
  a.in = static_cast<const uint8_t*>(d_in);
  a.w = nullptr;
  a.out = static_cast<float2*>(d_out);
  a.ntime = h->ntime; a.nchan = h->nchan; a.ninput = h->ninput; a.nbeam = h->nbeam;
  return bf_launch(h, a, d_w);
}

int xeng_bf_run_integrate(xeng_bf_t h, const void* d_in, void* d_out, const void* d_w, int ntime_sum) {
  if (!h || !d_in || !d_out) return fail(XENG_ERR_ARG, "null pointer");
  if (ntime_sum <= 0 || (h->nbeam & 1) || h->ntime % ntime_sum) return fail(XENG_ERR_ARG, "need ntime_sum | ntime and an even number of beams");
  XENG_CUDA(cudaSetDevice(h->device));
  if (!h->tap && ntime_sum <= xeng::beamform_fused_max_ntime_sum()) {   // one kernel, only the powers leave the SM
    xeng::BeamArgs a;
    a.in = static_cast<const uint8_t*>(d_in);
    a.w = nullptr;
    a.out = nullptr;
    a.ntime = h->ntime; a.nchan = h->nchan; a.ninput = h->ninput; a.nbeam = h->nbeam;
    a.pow_out = static_cast<float*>(d_out);
    a.ntime_sum = ntime_sum;
    return bf_launch(h, a, d_w);
  }
  if (!h->d_volt) XENG_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->d_volt), sizeof(float2) * h->nchan * h->nbeam * static_cast<size_t>(h->ntime)));
  int rc = xeng_bf_run(h, d_in, h->d_volt, d_w);
  if (rc) return rc;
  return xeng_bf_integrate(h, h->d_volt, d_out, ntime_sum);
}
const void* xeng_bf_voltages(xeng_bf_t h) { return h ? h->d_volt : nullptr; }
int xeng_bf_set_tap(xeng_bf_t h, int on) {
  if (!h) return fail(XENG_ERR_ARG, "null handle");
  h->tap = on ? 1 : 0;
  return XENG_OK;
}

int xeng_bf_integrate(xeng_bf_t h, const void* d_in, void* d_out, int ntime_sum) {
  if (!h || !d_in || !d_out) return fail(XENG_ERR_ARG, "null pointer");
  if (ntime_sum <= 0 || (h->nbeam & 1)) return fail(XENG_ERR_ARG, "need ntime_sum > 0 and an even number of beams");
  XENG_CUDA(cudaSetDevice(h->device));
  XENG_CUDA(xeng::launch_beam_integrate(static_cast<const float2*>(d_in), static_cast<float*>(d_out), h->nchan,
                                        h->nbeam, h->ntime, ntime_sum, 0, h->nbeam / 2, h->stream));
  return XENG_OK;
}

int xeng_bf_vlbi_tap(xeng_bf_t h, const void* d_in, void* d_out, int nbeam_send) {
  if (!h || !d_in || !d_out) return fail(XENG_ERR_ARG, "null pointer");
  if (nbeam_send <= 0 || nbeam_send > h->nbeam) return fail(XENG_ERR_ARG, "need 0 < nbeam_send <= nbeam");
  XENG_CUDA(cudaSetDevice(h->device));
  XENG_CUDA(xeng::launch_vlbi_tap(static_cast<const float2*>(d_in), static_cast<float2*>(d_out), h->nchan, h->nbeam,
                                  h->ntime, nbeam_send, h->stream));
  return XENG_OK;
}

int xeng_bf_integrate_single(xeng_bf_t h, const void* d_in, void* d_out, int ntime_sum, int beam_id) {
  if (!h || !d_in || !d_out) return fail(XENG_ERR_ARG, "null pointer");
  if (ntime_sum <= 0 || beam_id < 0 || 2 * beam_id + 1 >= h->nbeam) return fail(XENG_ERR_ARG, "bad ntime_sum / beam_id");
  XENG_CUDA(cudaSetDevice(h->device));
  XENG_CUDA(xeng::launch_beam_integrate(static_cast<const float2*>(d_in), static_cast<float*>(d_out), h->nchan,
                                        h->nbeam, h->ntime, ntime_sum, beam_id, 1, h->stream));
  return XENG_OK;
}

int xeng_beam_power_static(const void* d_in, const void* d_w, void* d_out, int ninput, int nchan, int ntime, int nbeam,
                           int ntime_sum, void* stream) {
  if (!d_in || !d_w || !d_out) return fail(XENG_ERR_ARG, "null pointer");
  if (ninput <= 0 || nchan <= 0 || ntime <= 0 || nbeam <= 0) return fail(XENG_ERR_ARG, "sizes must be positive");
  if (ntime_sum <= 0 || (nbeam & 1) || ntime % ntime_sum)
    return fail(XENG_ERR_ARG, "need ntime_sum | ntime and an even number of beams");
  const unsigned long long nvolt = 1ull * nchan * nbeam * static_cast<unsigned long long>(ntime);
  if (nvolt > xeng::beam_power_static_capacity())
    return fail(XENG_ERR_ARG, "xeng_beam_power_static: nchan*nbeam*ntime = %llu beam voltages exceed the static scratch "
                "array (XENG_STATIC_SCRATCH_FLOAT2 = %llu); use xeng_bf_run_integrate", nvolt,
                static_cast<unsigned long long>(xeng::beam_power_static_capacity()));
  XENG_CUDA(xeng::launch_beam_power_static(static_cast<const uint8_t*>(d_in), static_cast<const float2*>(d_w),
                                           static_cast<float*>(d_out), ninput, nchan, ntime, nbeam, ntime_sum,
                                           static_cast<cudaStream_t>(stream)));
  return XENG_OK;
}

int xeng_bf_sync(xeng_bf_t h) {
  if (!h) return fail(XENG_ERR_ARG, "null handle");
  cudaError_t e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) return fail(XENG_ERR_CUDA, "cudaStreamSynchronize: %s", cudaGetErrorString(e));
  return XENG_OK;
}

}  // extern "C"
