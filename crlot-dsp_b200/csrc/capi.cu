// capi.cu -- implementation of include/crlot_cuda.h: the C-ABI over the sm_100a kernels.
// No CPU fallback lives here: every compute entry point launches CUDA kernels or fails.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <new>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/crlot_cuda.h"
#include "kernels.h"
#include "tables.h"

using namespace crlot;

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
int fail_cuda(cudaError_t e, const char* where) {
  g_err = std::string(where) + ": " + cudaGetErrorString(e);
  cudaGetLastError();
  return CRLOT_ERR_CUDA;
}
#define CU(call)                                          \
  do {                                                    \
    cudaError_t e__ = (call);                             \
    if (e__ != cudaSuccess) return fail_cuda(e__, #call); \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

bool supported_nfft(int n) { return n == 256 || n == 512 || n == 1024 || n == 2048 || n == 4096; }

template <class T> struct DevArr {
  T* p = nullptr;
  size_t n = 0;
  ~DevArr() { if (p) cudaFree(p); }
  cudaError_t alloc(size_t count) {
    if (p) { cudaFree(p); p = nullptr; }
    n = count;
    return count ? cudaMalloc(&p, count * sizeof(T)) : cudaSuccess;
  }
  cudaError_t upload(const T* h, size_t count) {
    cudaError_t e = alloc(count);
    if (e != cudaSuccess) return e;
    return count ? cudaMemcpy(p, h, count * sizeof(T), cudaMemcpyHostToDevice) : cudaSuccess;
  }
};

// page-locked host staging (true asynchronous copies, no driver bounce buffer)
template <class T> struct PinnedArr {
  T* p = nullptr;
  size_t n = 0;
  ~PinnedArr() { if (p) cudaFreeHost(p); }
  cudaError_t resize(size_t count) {
    if (p) { cudaFreeHost(p); p = nullptr; }
    n = count;
    return count ? cudaHostAlloc(reinterpret_cast<void**>(&p), count * sizeof(T), cudaHostAllocDefault) : cudaSuccess;
  }
  T* data() { return p; }
  T& operator[](size_t i) { return p[i]; }
};

// "This host array is the unmodified output of my previous call and still lives on the device": the per-frame
// drop-in loop (bench/e2e_benchmark.cc:152-166) hands each call's output array straight to the next call --
// forward -> inverse -> push_frame_AoS.  A facade object records (caller's pointer, bytes, its device buffer, its
// page-locked copy of what it delivered); the next call finds the entry by pointer, checks with one memcmp against
// the page-locked copy that the caller has not touched the data, and then reads the device buffer instead of
// uploading the same bytes again.  Entries die with their owner.
struct Resident { const void* host; size_t bytes; const void* dev; const void* pinned; int device; const void* owner; };
thread_local Resident g_resident[2] = {};
void resident_publish(int slot, const void* host, size_t bytes, const void* dev, const void* pinned, int device, const void* owner) {
  g_resident[slot] = Resident{host, bytes, dev, pinned, device, owner};
}
const void* resident_lookup(const void* host, size_t bytes, int device) {
  for (const Resident& r : g_resident)
    if (r.host == host && r.dev && r.device == device && r.bytes >= bytes && std::memcmp(r.pinned, host, bytes) == 0) return r.dev;
  return nullptr;
}
void resident_forget(const void* owner) {
  for (Resident& r : g_resident) if (r.owner == owner) r = Resident{};
}

}  // namespace

struct crlot_buf_s {
  void* ptr = nullptr;
  size_t bytes = 0;
  int device = 0;
  bool owned = false;
};

// ---------------------------------------------------------------------------------------------
// host table builders
// ---------------------------------------------------------------------------------------------
static int window_make_impl(int type, size_t n, int periodic, int norm, float* out) {
  if (n == 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Window size must be greater than 0");
  if (!out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "window output pointer is null");
  if (type == CRLOT_WIN_BLACKMAN_HARRIS) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Blackman-Harris window not yet implemented");
  if (type < CRLOT_WIN_HANN || type > CRLOT_WIN_BLACKMAN_HARRIS) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Unknown window type");
  if (norm < CRLOT_NORM_NONE || norm > CRLOT_NORM_OLA_SUM_WSQ) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Unknown normalization type");
  const double den = periodic ? static_cast<double>(n) : static_cast<double>(n - 1);
  const double step = n > 1 ? 2.0 * M_PI / den : 0.0;
  for (size_t i = 0; i < n; ++i) {
    const double ang = step * static_cast<double>(i);
    double v = 1.0;
    if (n > 1) {
      switch (type) {
        case CRLOT_WIN_HANN: v = 0.5 * (1.0 - std::cos(ang)); break;
        case CRLOT_WIN_HAMMING: v = 0.54 - 0.46 * std::cos(ang); break;
        case CRLOT_WIN_BLACKMAN: { const double c1 = std::cos(ang), c2 = std::cos(2.0 * ang); v = 0.42 - 0.5 * c1 + 0.08 * c2; break; }
        default: v = 1.0; break;
      }
    }
    out[i] = static_cast<float>(v);
  }
  if (norm != CRLOT_NORM_NONE) {
    double acc = 0.0;
    if (norm == CRLOT_NORM_SUM_TO_ONE) for (size_t i = 0; i < n; ++i) acc += static_cast<double>(out[i]);
    else for (size_t i = 0; i < n; ++i) acc += static_cast<double>(out[i]) * static_cast<double>(out[i]);
    if (acc > 0.0) {
      const float k = static_cast<float>(norm == CRLOT_NORM_SUM_TO_ONE ? 1.0 / acc : 1.0 / std::sqrt(acc));
      for (size_t i = 0; i < n; ++i) out[i] *= k;
    }
  }
  return CRLOT_OK;
}

static void norm_linear_impl(float* norm, const float* w, long long ring, long long N, long long H) {
  std::fill(norm, norm + ring, 0.0f);
  if (ring <= 0 || N <= 0 || H <= 0) return;
  const long long k_first = -((N + H - 1) / H), k_last = (ring + N - 1 + H - 1) / H;
  for (long long k = k_first; k <= k_last; ++k) {
    long long s = (k * H) % ring;
    if (s < 0) s += ring;
    const long long head = std::min(N, ring - s);
    for (long long i = 0; i < head; ++i) norm[s + i] += w[i];      // fp32 adds, k then i ascending
    for (long long i = head; i < N; ++i) norm[i - head] += w[i];   // second span starts at ring[0]
  }
}

static void envelope_impl(float* norm, const float* w, size_t ring, size_t N, size_t H, int inside, float eps) {
  if (!w || !inside) {
    std::fill(norm, norm + ring, 1.0f);
  } else if (H == N) {
    for (size_t i = 0; i < ring; ++i) norm[i] = std::max(w[i % N], eps);
  } else {
    norm_linear_impl(norm, w, static_cast<long long>(ring), static_cast<long long>(N), static_cast<long long>(H));
  }
}

// ---------------------------------------------------------------------------------------------
// plan
// ---------------------------------------------------------------------------------------------
struct HostSlot {
  cudaStream_t stream = nullptr;
  DevArr<float> in, frames, out, aos;
  DevArr<cpx> spec;
};

struct crlot_plan_s {
  crlot_stft_desc d{};
  int bins = 0;
  size_t ring_len = 0;
  bool syn_active = false;
  DevArr<float> win_a, win_s, win_ones, env;
  DevArr<cpx> tw1, sa, pre;
  // Work counters of the dynamically scheduled kernels (K1, fused K2+K3).  A launch zeroes its counter word on its
  // own stream and the kernel pulls chunks from it, so launches that may overlap must not share a word: every
  // stream that ever launches on this plan gets its own (launches on one stream are ordered, so K1 and the fused
  // kernel of that stream share it).  kCounterWords streams are tracked; beyond that words are reused round-robin.
  static constexpr int kCounterWords = 1024;
  DevArr<unsigned int> counter;
  std::mutex counter_mu;
  std::unordered_map<cudaStream_t, int> counter_of;
  int counter_next = 0;
  unsigned int* counter_for(cudaStream_t st) {
    std::lock_guard<std::mutex> lk(counter_mu);
    auto it = counter_of.find(st);
    if (it == counter_of.end()) {
      if (counter_of.size() >= static_cast<size_t>(kCounterWords)) counter_of.clear();
      it = counter_of.emplace(st, counter_next).first;
      counter_next = (counter_next + 1) % kCounterWords;
    }
    return counter.p + it->second;
  }
  DevArr<float2> den_rcp;      // {max(env, eps), refined reciprocal}: division table of the fused kernel 2+3
  bool generic = false;        // any-size kernels (k_generic.cu): nfft outside the register-FFT set, or hop % 4 != 0
  int k1_scrub = 0;            // kernel-1 clean-up mode (kernels.h)
  bool k23_win_ok = true;      // synthesis window can be pre-scaled by 2/N exactly (fused kernel 2+3)
  const float* sgain = nullptr;   // spectral-domain hook (caller-owned device table), see crlot_stft_plan_set_spectral_gain
  long long sgain_row_stride = 0;
  unsigned int tiny2 = 0;
  std::vector<float> env_host;
  static constexpr int kHostSlots = 3;   // H2D / kernels / D2H of a slab are serial on its stream: three slabs in flight keep both copy engines busy
  HostSlot slot[kHostSlots];
  long long slot_frames = 0;   // capacity of the host-path workspace, in frames
  int slot_group = 0;          // ... and in channels per slab (interleaved host path)
  bool slot_aos = false;
};

static cudaError_t upload_cpx(DevArr<cpx>& dst, const std::vector<float>& v) {
  return dst.upload(reinterpret_cast<const cpx*>(v.data()), v.size() / 2);
}

extern "C" {

int crlot_cuda_abi_version(void) { return CRLOT_ABI_VERSION; }
// internal: lets the other translation units of the library (framer.cu) set the last-error text
void crlot_cuda_set_error_(const char* msg) { g_err = msg ? msg : ""; }
const char* crlot_cuda_last_error(void) { return g_err.c_str(); }

int crlot_cuda_device_count(int* count) {
  if (!count) return fail(CRLOT_ERR_INVALID_ARGUMENT, "count is null");
  CU(cudaGetDeviceCount(count));
  return CRLOT_OK;
}
int crlot_cuda_stream_sync(int device, void* stream) {
  DeviceGuard g(device);
  if (!g.ok) return fail(CRLOT_ERR_CUDA, "cannot select device");
  CU(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}

// ---- buffers --------------------------------------------------------------------------------
int crlot_buf_create(size_t bytes, int device, crlot_buf_t* out) {
  if (!out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "out is null");
  DeviceGuard g(device);
  if (!g.ok) return fail(CRLOT_ERR_CUDA, "cannot select device");
  auto* b = new (std::nothrow) crlot_buf_s;
  if (!b) return fail(CRLOT_ERR_ALLOC, "host allocation failed");
  if (bytes) {
    cudaError_t e = cudaMalloc(&b->ptr, bytes);
    if (e != cudaSuccess) { delete b; fail_cuda(e, "cudaMalloc"); return e == cudaErrorMemoryAllocation ? CRLOT_ERR_ALLOC : CRLOT_ERR_CUDA; }
  }
  b->bytes = bytes; b->device = device; b->owned = true;
  *out = b;
  return CRLOT_OK;
}
int crlot_buf_wrap(void* device_ptr, size_t bytes, int device, crlot_buf_t* out) {
  if (!out || (!device_ptr && bytes)) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  auto* b = new (std::nothrow) crlot_buf_s;
  if (!b) return fail(CRLOT_ERR_ALLOC, "host allocation failed");
  b->ptr = device_ptr; b->bytes = bytes; b->device = device; b->owned = false;
  *out = b;
  return CRLOT_OK;
}
int crlot_buf_destroy(crlot_buf_t buf) {
  if (!buf) return CRLOT_OK;
  if (buf->owned && buf->ptr) { DeviceGuard g(buf->device); cudaFree(buf->ptr); }
  delete buf;
  return CRLOT_OK;
}
int crlot_buf_upload(crlot_buf_t buf, size_t off, const void* host, size_t bytes, void* stream) {
  if (!buf || (!host && bytes)) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (off + bytes > buf->bytes) return fail(CRLOT_ERR_INVALID_ARGUMENT, "upload exceeds buffer");
  DeviceGuard g(buf->device);
  CU(cudaMemcpyAsync(static_cast<char*>(buf->ptr) + off, host, bytes, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}
int crlot_buf_download(crlot_buf_t buf, size_t off, void* host, size_t bytes, void* stream) {
  if (!buf || (!host && bytes)) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (off + bytes > buf->bytes) return fail(CRLOT_ERR_INVALID_ARGUMENT, "download exceeds buffer");
  DeviceGuard g(buf->device);
  CU(cudaMemcpyAsync(host, static_cast<const char*>(buf->ptr) + off, bytes, cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}
int crlot_buf_zero(crlot_buf_t buf, size_t off, size_t bytes, void* stream) {
  if (!buf) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null buffer");
  if (off + bytes > buf->bytes) return fail(CRLOT_ERR_INVALID_ARGUMENT, "range exceeds buffer");
  DeviceGuard g(buf->device);
  CU(cudaMemsetAsync(static_cast<char*>(buf->ptr) + off, 0, bytes, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}
void* crlot_buf_device_ptr(crlot_buf_t buf) { return buf ? buf->ptr : nullptr; }
size_t crlot_buf_bytes(crlot_buf_t buf) { return buf ? buf->bytes : 0; }
int crlot_host_alloc(size_t bytes, void** out) {
  if (!out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "out is null");
  cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault);
  if (e != cudaSuccess) { fail_cuda(e, "cudaHostAlloc"); return CRLOT_ERR_ALLOC; }
  return CRLOT_OK;
}
int crlot_host_free(void* p) {
  if (p) CU(cudaFreeHost(p));
  return CRLOT_OK;
}

// ---- tables ---------------------------------------------------------------------------------
int crlot_window_make(int type, size_t n, int periodic, int norm, float* out) {
  return window_make_impl(type, n, periodic, norm, out);
}
size_t crlot_ola_ring_len(size_t frame, size_t hop) {
  if (hop == 0) return 0;
  return ((frame + hop - 1) / hop + 20) * hop;
}
int crlot_build_norm_linear(float* norm, const float* window, size_t ring_len, size_t frame, size_t hop) {
  if (!norm || (!window && frame)) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  norm_linear_impl(norm, window, static_cast<long long>(ring_len), static_cast<long long>(frame), static_cast<long long>(hop));
  return CRLOT_OK;
}
int crlot_ola_envelope(float* norm, const float* w, size_t ring_len, size_t frame, size_t hop, int inside, float eps) {
  if (!norm) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  envelope_impl(norm, w, ring_len, frame, hop, inside, eps);
  return CRLOT_OK;
}
long long crlot_frame_count(long long len, int frame, int hop, int mode) {
  if (len <= 0 || frame <= 0 || hop <= 0) return 0;
  if (mode == CRLOT_DROP) return len < frame ? 0 : (len - frame) / hop + 1;
  return (len + hop - 1) / hop;
}

// ---- plan -----------------------------------------------------------------------------------
int crlot_stft_plan_create(const crlot_stft_desc* desc, crlot_plan_t* out) {
  if (!desc || !out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (desc->nfft <= 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Frame size must be greater than 0");
  if (desc->hop <= 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Hop size must be greater than 0");
  if (!(desc->eps > 0.0f)) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Invalid OLA configuration");
  if (desc->boundary_mode != CRLOT_ZERO_PAD && desc->boundary_mode != CRLOT_DROP) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad boundary mode");
  if (desc->nfft % 2 != 0) return fail(CRLOT_ERR_RUNTIME, "FFT size must be even for real FFT");
  // register-FFT kernels: nfft in {256..4096} and hop % 4 == 0; everything else the reference's plan accepts
  // runs on the any-size kernels (k_generic.cu), up to 16384 points
  const bool generic = !supported_nfft(desc->nfft) || desc->hop % 4 != 0;
  if (generic && !generic_len_supported(desc->nfft / 2))
    return fail(CRLOT_ERR_UNSUPPORTED, "nfft must be an even number <= 16384 on the GPU path");
  if (desc->hop > desc->nfft) return fail(CRLOT_ERR_UNSUPPORTED, "hop must be <= nfft on the GPU path");
  DeviceGuard g(desc->device);
  if (!g.ok) return fail(CRLOT_ERR_CUDA, "cannot select device");
  std::unique_ptr<crlot_plan_s> p(new (std::nothrow) crlot_plan_s);
  if (!p) return fail(CRLOT_ERR_ALLOC, "host allocation failed");
  p->d = *desc;
  p->generic = generic;
  p->bins = desc->nfft / 2 + 1;
  const int N = desc->nfft;
  // analysis window (ones = none)
  std::vector<float> wa(N, 1.0f);
  if (desc->analysis_window) std::copy(desc->analysis_window, desc->analysis_window + N, wa.begin());
  CU(p->win_a.upload(wa.data(), N));
  // Kernel-1 clean-up mode.  Screening (mode 2) is exact when every windowed product that the
  // reference rule would alter comes from a raw sample that is non-finite or 0 < |x| < thr with
  // thr = 1e-30 / min non-zero |w| (plus a relative margin for the rounding of x*w); it needs
  // |w| <= 1 (no overflow of finite samples) and finite window values.
  p->k1_scrub = desc->scrub ? 1 : 0;
  if (desc->scrub) {
    float wmin = 0.0f, wmax = 0.0f;
    bool finite = true;
    for (float v : wa) {
      const float a = std::fabs(v);
      if (!std::isfinite(v)) finite = false;
      if (a > 0.0f && (wmin == 0.0f || a < wmin)) wmin = a;
      wmax = std::max(wmax, a);
    }
    if (finite && wmax <= 1.0f && wmin > 0.0f) {
      const double thr = 1e-30 / static_cast<double>(wmin) * (1.0 + 1e-5);
      if (thr < 1e30) {
        const float t = static_cast<float>(thr);
        unsigned int bits;
        std::memcpy(&bits, &t, 4);
        p->tiny2 = 2u * bits - 2u;
        p->k1_scrub = 2;
      }
    }
  }
  // synthesis window policy, ref: OLAAccumulator.cc:82-83
  const float* ws = desc->apply_window_inside ? desc->ola_window : desc->push_window;
  p->syn_active = ws != nullptr;
  if (ws) CU(p->win_s.upload(ws, N));
  if (ws)
    for (int i = 0; i < N; ++i)
      if (!std::isfinite(ws[i]) || (ws[i] != 0.0f && std::fabs(ws[i]) < 1e-25f)) p->k23_win_ok = false;
  {
    const std::vector<float> ones(N, 1.0f);          // the fused kernel always multiplies: y * 1.0f == y
    CU(p->win_ones.upload(ones.data(), N));
  }
  // envelope
  if (desc->envelope) {
    if (desc->envelope_len == 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "envelope_len is 0");
    p->ring_len = desc->envelope_len;
    p->env_host.assign(desc->envelope, desc->envelope + desc->envelope_len);
  } else {
    p->ring_len = crlot_ola_ring_len(N, desc->hop);
    p->env_host.resize(p->ring_len);
    envelope_impl(p->env_host.data(), desc->ola_window, p->ring_len, N, desc->hop, desc->apply_window_inside, desc->eps);
  }
  CU(p->env.upload(p->env_host.data(), p->env_host.size()));
  {
    // the fused kernel's table-driven division needs every denominator well inside the normal range
    bool ok = true;
    for (float v : p->env_host) {
      const float b = std::max(v, desc->eps);
      if (!(b >= 9.094947017729282e-13f && b <= 1.099511627776e12f)) ok = false;   // [2^-40, 2^40]; NaN fails
    }
    if (ok) {
      CU(p->den_rcp.alloc(p->env_host.size()));
      CU(launch_build_den_rcp(p->env.p, desc->eps, static_cast<long long>(p->env_host.size()), p->den_rcp.p, nullptr));
      CU(cudaStreamSynchronize(nullptr));
    }
  }
  const FftTables T = generic ? make_generic_tables(N) : make_fft_tables(N);
  CU(upload_cpx(p->tw1, T.tw1));
  CU(upload_cpx(p->sa, T.sa));
  if (!T.pre.empty()) CU(upload_cpx(p->pre, T.pre));
  CU(p->counter.alloc(crlot_plan_s::kCounterWords));
  p->d.analysis_window = p->d.ola_window = p->d.push_window = p->d.envelope = nullptr;  // not retained
  *out = p.release();
  return CRLOT_OK;
}

int crlot_stft_plan_destroy(crlot_plan_t plan) {
  if (!plan) return CRLOT_OK;
  DeviceGuard g(plan->d.device);
  for (auto& s : plan->slot) if (s.stream) cudaStreamDestroy(s.stream);
  delete plan;
  return CRLOT_OK;
}
int crlot_stft_plan_info(crlot_plan_t plan, crlot_plan_info* info) {
  if (!plan || !info) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  info->nfft = plan->d.nfft; info->hop = plan->d.hop; info->bins = plan->bins;
  info->ring_len = plan->ring_len; info->synthesis_window_active = plan->syn_active; info->device = plan->d.device;
  return CRLOT_OK;
}
int crlot_stft_plan_envelope(crlot_plan_t plan, float* host_out) {
  if (!plan || !host_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  std::copy(plan->env_host.begin(), plan->env_host.end(), host_out);
  return CRLOT_OK;
}

// ---- spectral-domain hook -----------------------------------------------------------
int crlot_stft_plan_set_spectral_gain(crlot_plan_t plan, const float* d_gain, long long row_stride) {
  if (!plan) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (d_gain && row_stride != 0 && row_stride < plan->bins)
    return fail(CRLOT_ERR_INVALID_ARGUMENT, "row_stride must be 0 (one gain row for every frame) or >= bins");
  if (d_gain && reinterpret_cast<uintptr_t>(d_gain) % 4 != 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "gain table must be 4-byte aligned");
  plan->sgain = d_gain;
  plan->sgain_row_stride = d_gain ? row_stride : 0;
  return CRLOT_OK;
}

int crlot_spectral_gain_apply(crlot_plan_t plan, void* d_spec, long long rows, const float* d_gain, long long row_stride,
                              void* stream) {
  if (!plan || !d_spec || !d_gain) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (rows < 0 || (row_stride != 0 && row_stride < plan->bins)) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  DeviceGuard g(plan->d.device);
  CU(launch_spectral_gain(static_cast<cpx*>(d_spec), rows, plan->bins, d_gain, row_stride, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}

// ---- execute --------------------------------------------------------------------------------
int crlot_stft_execute(crlot_plan_t plan, const float* d_signal, long long chan_stride, long long len, int channels,
                       long long n_frames, void* d_spec, void* stream) {
  if (!plan || !d_signal || !d_spec) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (channels <= 0 || n_frames < 0 || len < 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  if (reinterpret_cast<uintptr_t>(d_spec) % 8 != 0) return fail(CRLOT_ERR_UNSUPPORTED, "buffers must be 8-byte aligned");
  if (!plan->generic && (chan_stride % 4 != 0 || reinterpret_cast<uintptr_t>(d_signal) % 16 != 0))
    return fail(CRLOT_ERR_UNSUPPORTED, "signal must be 16-byte aligned with chan_stride a multiple of 4 floats");
  DeviceGuard g(plan->d.device);
  K1Params k{};
  k.sig = d_signal; k.chan_stride = chan_stride; k.len = len; k.channels = channels; k.frames = n_frames;
  k.nfft = plan->d.nfft; k.hop = plan->d.hop;
  k.spec = static_cast<cpx*>(d_spec); k.spec_chan_rows = n_frames;
  k.window = plan->win_a.p; k.tw1 = plan->tw1.p; k.sa = plan->sa.p; k.pre = plan->pre.p;
  k.counter = plan->counter_for(static_cast<cudaStream_t>(stream)); k.chunk_steps = 64; k.scrub = plan->k1_scrub; k.tiny2 = plan->tiny2;
  k.generic = plan->generic;
  CU(launch_stft_fwd(k, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}

// ---- FrameQueue centre / pad front end ---------------------------------------------
long long crlot_framequeue_count(long long len, int frame, int hop, int center) {
  // FrameQueue::calculateNumFrames over the padded length (FrameQueue.cc:75,100-115)
  if (frame <= 0 || hop <= 0 || len < 0) return -1;
  const long long padded = center ? len + 2ll * (frame / 2) : len;
  if (padded < frame) return 0;
  const long long tail = frame > hop ? frame - hop : 0;
  if (padded < tail) return 0;
  return (padded - tail) / hop;
}

int crlot_framequeue_extract(const float* d_signal, long long len, int frame, int hop, int center, int pad_mode,
                             long long n_frames, float* d_frames, int device, void* stream) {
  // same checks and messages as the reference constructor (FrameQueue.cc:13-21)
  if (frame <= 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Frame size must be greater than 0");
  if (hop <= 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Hop size must be greater than 0");
  if (!d_signal && len > 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Input pointer cannot be null when length > 0");
  if (len < 0 || len >= (1ll << 31)) return fail(CRLOT_ERR_INVALID_ARGUMENT, "length out of range (the reference indexes with int)");
  if (pad_mode < CRLOT_PAD_CONSTANT || pad_mode > CRLOT_PAD_EDGE) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad pad mode");
  const long long F = crlot_framequeue_count(len, frame, hop, center);
  if (n_frames < 0 || n_frames > F) return fail(CRLOT_ERR_INVALID_ARGUMENT, "n_frames exceeds crlot_framequeue_count");
  if (n_frames > 0 && !d_frames) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  DeviceGuard g(device);
  CU(launch_framequeue_extract(d_signal, len, frame, hop, center ? frame / 2 : 0, pad_mode, n_frames, d_frames,
                               static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}

int crlot_stft_execute_framequeue(crlot_plan_t plan, const float* d_signal, long long chan_stride, long long len,
                                  int channels, long long n_frames, int center, int pad_mode, void* d_spec, void* stream) {
  if (!plan || !d_signal || !d_spec) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (channels <= 0 || n_frames < 0 || len < 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  if (pad_mode < CRLOT_PAD_CONSTANT || pad_mode > CRLOT_PAD_EDGE) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad pad mode");
  if (len >= (1ll << 31)) return fail(CRLOT_ERR_INVALID_ARGUMENT, "length out of range (the reference indexes with int)");
  if (n_frames > crlot_framequeue_count(len, plan->d.nfft, plan->d.hop, center))
    return fail(CRLOT_ERR_INVALID_ARGUMENT, "n_frames exceeds crlot_framequeue_count");
  if (reinterpret_cast<uintptr_t>(d_spec) % 8 != 0) return fail(CRLOT_ERR_UNSUPPORTED, "buffers must be 8-byte aligned");
  if (!plan->generic && (chan_stride % 4 != 0 || reinterpret_cast<uintptr_t>(d_signal) % 16 != 0))
    return fail(CRLOT_ERR_UNSUPPORTED, "signal must be 16-byte aligned with chan_stride a multiple of 4 floats");
  DeviceGuard g(plan->d.device);
  K1Params k{};
  k.sig = d_signal; k.chan_stride = chan_stride; k.len = len; k.channels = channels; k.frames = n_frames;
  k.pad = center ? plan->d.nfft / 2 : 0; k.pad_mode = pad_mode;
  k.nfft = plan->d.nfft; k.hop = plan->d.hop;
  k.spec = static_cast<cpx*>(d_spec); k.spec_chan_rows = n_frames;
  k.window = plan->win_a.p; k.tw1 = plan->tw1.p; k.sa = plan->sa.p; k.pre = plan->pre.p;
  k.counter = plan->counter_for(static_cast<cudaStream_t>(stream)); k.chunk_steps = 64; k.scrub = plan->k1_scrub; k.tiny2 = plan->tiny2;
  k.generic = plan->generic;
  CU(launch_stft_fwd(k, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}

int crlot_istft_execute(crlot_plan_t plan, const void* d_spec, long long rows, float* d_frames, int fuse_win, void* stream) {
  if (!plan || !d_spec || !d_frames) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (rows < 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  if (reinterpret_cast<uintptr_t>(d_spec) % 8 != 0 || reinterpret_cast<uintptr_t>(d_frames) % 8 != 0)
    return fail(CRLOT_ERR_UNSUPPORTED, "buffers must be 8-byte aligned");
  DeviceGuard g(plan->d.device);
  K2Params k{};
  k.spec = static_cast<const cpx*>(d_spec); k.rows = rows; k.nfft = plan->d.nfft; k.frames = d_frames;
  k.window = (fuse_win && plan->syn_active) ? plan->win_s.p : nullptr;
  k.tw1 = plan->tw1.p; k.sa = plan->sa.p; k.pre = plan->pre.p; k.scrub = plan->d.scrub;
  k.sgain = plan->sgain; k.sgain_row_stride = plan->sgain_row_stride; k.generic = plan->generic;
  CU(launch_istft_inv(k, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}

int crlot_ola_accumulate(crlot_plan_t plan, const float* d_frames, long long frames_chan_rows, long long frames_per_chan,
                         int channels, int prewindowed, long long t_offset, long long frame_t0, float* d_out,
                         long long out_chan_stride, long long n_out, void* stream) {
  if (!plan || !d_frames || !d_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (channels <= 0 || frames_per_chan < 0 || n_out < 0 || t_offset < 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  DeviceGuard g(plan->d.device);
  K3Params k{};
  k.frames = d_frames; k.frames_chan_rows = frames_chan_rows; k.frames_per_chan = frames_per_chan; k.channels = channels;
  k.nfft = plan->d.nfft; k.hop = plan->d.hop;
  k.window = (!prewindowed && plan->syn_active) ? plan->win_s.p : nullptr;
  k.gain = plan->d.gain; k.eps = plan->d.eps; k.norm = plan->env.p; k.ring_len = static_cast<long long>(plan->ring_len);
  k.t_offset = t_offset; k.frame_t0 = frame_t0; k.out = d_out; k.out_chan_stride = out_chan_stride; k.n_out = n_out;
  CU(launch_ola_gather(k, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}

static void fill_k23(K23Params& k, crlot_plan_t plan, cudaStream_t st) {
  k.nfft = plan->d.nfft; k.hop = plan->d.hop;
  k.window = plan->syn_active ? plan->win_s.p : plan->win_ones.p;
  k.tw1 = plan->tw1.p; k.sa = plan->sa.p; k.pre = plan->pre.p; k.scrub = plan->d.scrub;
  k.gain = plan->d.gain; k.eps = plan->d.eps; k.norm = plan->env.p; k.ring_len = static_cast<long long>(plan->ring_len);
  k.den_rcp = plan->den_rcp.p;
  k.counter = plan->counter_for(st); k.chunk_blocks = 0;
  k.sgain = plan->sgain; k.sgain_row_stride = plan->sgain_row_stride;
}

int crlot_selftest_division(crlot_plan_t plan, unsigned long long n, unsigned long long seed, unsigned long long* mismatches) {
  if (!plan || !mismatches) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  DeviceGuard g(plan->d.device);
  DevArr<unsigned long long> d;
  CU(d.alloc(1));
  CU(cudaMemset(d.p, 0, sizeof(unsigned long long)));
  CU(launch_division_selftest(plan->den_rcp.p, plan->den_rcp.p ? static_cast<long long>(plan->ring_len) : 0, n, seed, d.p, nullptr));
  CU(cudaMemcpy(mismatches, d.p, sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return CRLOT_OK;
}

int crlot_istft_ola_supported(crlot_plan_t plan, long long t_offset, long long frame_t0, const float* d_out,
                              long long out_chan_stride) {
  if (!plan || !plan->k23_win_ok || plan->generic) return 0;
  return istft_ola_supported(plan->d.nfft, plan->d.hop, static_cast<long long>(plan->ring_len), t_offset, frame_t0,
                             out_chan_stride, d_out, plan->env.p) ? 1 : 0;
}

int crlot_istft_ola_execute(crlot_plan_t plan, const void* d_spec, long long spec_chan_rows, long long frames_per_chan,
                            int channels, long long t_offset, long long frame_t0, float* d_out,
                            long long out_chan_stride, long long n_out, void* stream) {
  if (!plan || !d_spec || !d_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (channels <= 0 || frames_per_chan < 0 || n_out < 0 || t_offset < 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  if (reinterpret_cast<uintptr_t>(d_spec) % 8 != 0) return fail(CRLOT_ERR_UNSUPPORTED, "buffers must be 8-byte aligned");
  if (!crlot_istft_ola_supported(plan, t_offset, frame_t0, d_out, out_chan_stride))
    return fail(CRLOT_ERR_UNSUPPORTED,
                "fused inverse+overlap-add needs hop a multiple of 64 (16 for nfft 256, 32 for 512/1024), "
                "even offsets/strides; use crlot_istft_execute + crlot_ola_accumulate");
  DeviceGuard g(plan->d.device);
  K23Params k{};
  fill_k23(k, plan, static_cast<cudaStream_t>(stream));
  k.spec = static_cast<const cpx*>(d_spec); k.spec_chan_rows = spec_chan_rows; k.frames_per_chan = frames_per_chan;
  k.channels = channels; k.t_offset = t_offset; k.frame_t0 = frame_t0; k.out = d_out; k.out_chan_stride = out_chan_stride;
  k.n_out = n_out;
  CU(launch_istft_ola(k, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}

// ---- fully fused round trip: K1 -> [spectral hook] -> K2+K3 in ONE kernel, spectra never materialised -----------
int crlot_roundtrip_fused_supported(crlot_plan_t plan, const float* d_signal, long long chan_stride, long long t_offset,
                                    long long frame_t0, const float* d_out, long long out_chan_stride) {
  if (!plan || !plan->k23_win_ok || plan->generic || !plan->den_rcp.p) return 0;
  return roundtrip_fused_supported(plan->d.nfft, plan->d.hop, static_cast<long long>(plan->ring_len), t_offset, frame_t0,
                                   out_chan_stride, d_out, plan->env.p, d_signal, chan_stride) ? 1 : 0;
}

int crlot_roundtrip_fused_execute(crlot_plan_t plan, const float* d_signal, long long chan_stride, long long len, int channels,
                                  long long n_frames, long long t_offset, long long frame_t0, float* d_out,
                                  long long out_chan_stride, long long n_out, void* stream) {
  if (!plan || !d_signal || !d_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (channels <= 0 || n_frames < 0 || n_out < 0 || t_offset < 0 || len < 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  if (!crlot_roundtrip_fused_supported(plan, d_signal, chan_stride, t_offset, frame_t0, d_out, out_chan_stride))
    return fail(CRLOT_ERR_UNSUPPORTED,
                "the one-kernel round trip exists for nfft 2048 / hop 512 with even offsets and strides; "
                "use crlot_roundtrip_execute with a spectra buffer");
  DeviceGuard g(plan->d.device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  K23Params k{};
  fill_k23(k, plan, st);
  k.spec = nullptr; k.spec_chan_rows = n_frames; k.frames_per_chan = n_frames; k.channels = channels;
  k.t_offset = t_offset; k.frame_t0 = frame_t0; k.out = d_out; k.out_chan_stride = out_chan_stride; k.n_out = n_out;
  k.sig = d_signal; k.sig_chan_stride = chan_stride; k.sig_len = len; k.awindow = plan->win_a.p;
  k.k1_scrub = plan->k1_scrub; k.tiny2 = plan->tiny2;
  CU(launch_istft_ola(k, st));
  return CRLOT_OK;
}

int crlot_roundtrip_execute(crlot_plan_t plan, const float* d_signal, long long chan_stride, long long len, int channels,
                            void* d_spec, float* d_frames, float* d_out, long long out_chan_stride, void* stream) {
  if (!plan) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  const long long F = crlot_frame_count(len, plan->d.nfft, plan->d.hop, plan->d.boundary_mode);
  if (F == 0) return CRLOT_OK;
  if (!d_spec && !d_frames)   // neither spectra nor iFFT frames wanted: the whole round trip in one kernel
    return crlot_roundtrip_fused_execute(plan, d_signal, chan_stride, len, channels, F, 0, 0, d_out, out_chan_stride,
                                         (F - 1) * plan->d.hop + plan->d.nfft, stream);
  int rc = crlot_stft_execute(plan, d_signal, chan_stride, len, channels, F, d_spec, stream);
  if (rc) return rc;
  if (!d_frames)   // no materialisation of the iFFT frames requested: kernels 2+3 fused
    return crlot_istft_ola_execute(plan, d_spec, F, F, channels, 0, 0, d_out, out_chan_stride,
                                   (F - 1) * plan->d.hop + plan->d.nfft, stream);
  rc = crlot_istft_execute(plan, d_spec, F * channels, d_frames, 1, stream);
  if (rc) return rc;
  const long long n_out = (F - 1) * plan->d.hop + plan->d.nfft;
  return crlot_ola_accumulate(plan, d_frames, F, F, channels, 1, 0, 0, d_out, out_chan_stride, n_out, stream);
}

// The three workspace slots of the host-buffer entry points (one stream each), grown on demand.
static int ensure_host_slots(crlot_plan_t plan, long long slab_frames, int G, bool aos, bool fused, long long ist, long long ost) {
  const int N = plan->d.nfft;
  if (plan->slot_frames < slab_frames || plan->slot_group < G || (aos && !plan->slot_aos)) {
    for (auto& s : plan->slot) {
      if (!s.stream) CU(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
      CU(s.in.alloc(static_cast<size_t>(ist) * G));
      CU(s.spec.alloc(static_cast<size_t>(slab_frames) * plan->bins * G));
      if (!fused) CU(s.frames.alloc(static_cast<size_t>(slab_frames) * N * G));
      CU(s.out.alloc(static_cast<size_t>(ost) * G));
      if (aos) CU(s.aos.alloc(static_cast<size_t>(ist) * G));
    }
    plan->slot_frames = slab_frames; plan->slot_group = G; plan->slot_aos = aos;
  }
  return CRLOT_OK;
}

// Host-buffer round trip: time slabs through three workspace slots, one stream each, so the
// H2D copy of slab i+1 and the D2H copy of slab i-1 overlap the kernels of slab i.
// aos: h_in / h_out are interleaved [n][channels] (what Framer::push and WavReader hand over,
// ref: dsp/frame/framer.cc:133-141, io/wav.cc:90-123); every slab carries all channels, is
// de-interleaved on the device after the copy and re-interleaved before the copy back.
static int roundtrip_host_impl(crlot_plan_t plan, const float* h_in, long long chan_stride, long long len, int channels,
                               float* h_out, long long out_chan_stride, long long* n_out_per_chan, bool aos) {
  if (!plan || !h_in || !h_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (channels <= 0 || len < 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  const int N = plan->d.nfft, H = plan->d.hop;
  const long long F = crlot_frame_count(len, N, H, plan->d.boundary_mode);
  const long long n_out = F > 0 ? (F - 1) * H + N : 0;
  if (n_out_per_chan) *n_out_per_chan = n_out;
  if (F == 0) return CRLOT_OK;
  const int G = aos ? channels : 1;            // channels carried by one slab
  if (G > 1 && plan->sgain && plan->sgain_row_stride != 0)
    return fail(CRLOT_ERR_UNSUPPORTED, "a time-frequency mask with the interleaved multi-channel host path: use the planar entry point");
  DeviceGuard g(plan->d.device);
  // slab = `slab_hops` hops of output; its frames are the ones overlapping that range
  const long long overlap = (N + H - 1) / H - 1;
  const long long slab_hops = std::min<long long>(std::max<long long>((1ll << 24) / H / G, 4 * (overlap + 1)), F + overlap);
  const long long slab_frames = slab_hops + overlap;
  const long long ist = ((slab_frames - 1) * H + N + 3) / 4 * 4;      // planar strides inside the workspace
  const long long ost = (slab_hops * H + N + 3) / 4 * 4;
  // slab offsets are multiples of the hop, so one eligibility test covers every slab
  const bool fused = !plan->generic && plan->k23_win_ok && istft_ola_supported(N, H, static_cast<long long>(plan->ring_len), 0, 0, ost, nullptr, plan->env.p);
  { const int rc = ensure_host_slots(plan, slab_frames, G, aos, fused, ist, ost); if (rc) return rc; }
  // every slot launches on its own stream, hence with its own work-counter word (crlot_plan_s::counter_for)
  int it = 0;
  for (int c = 0; c < (aos ? 1 : channels); ++c) {
    const float* xin = h_in + static_cast<long long>(c) * chan_stride;
    float* yout = h_out + static_cast<long long>(c) * out_chan_stride;
    for (long long t0 = 0; t0 < n_out; t0 += slab_hops * H, ++it) {
      HostSlot& s = plan->slot[it % crlot_plan_s::kHostSlots];
      const long long t1 = std::min(n_out, t0 + slab_hops * H);
      long long k_lo = t0 - N + 1 <= 0 ? 0 : (t0 - N + 1 + H - 1) / H;
      long long k_hi = std::min(F - 1, (t1 - 1) / H);
      const long long nf = k_hi - k_lo + 1;
      const long long s0 = k_lo * H;                                   // first input sample of the slab
      const long long s1 = std::min(len, k_hi * H + N);                // one past the last that exists
      if (aos) {
        CU(cudaMemcpyAsync(s.aos.p, h_in + s0 * G, static_cast<size_t>(s1 - s0) * G * sizeof(float), cudaMemcpyHostToDevice, s.stream));
        CU(launch_deinterleave(s.aos.p, s1 - s0, G, s.in.p, ist, s.stream));
      } else {
        CU(cudaMemcpyAsync(s.in.p, xin + s0, static_cast<size_t>(s1 - s0) * sizeof(float), cudaMemcpyHostToDevice, s.stream));
      }
      K1Params k1{};
      k1.sig = s.in.p; k1.chan_stride = G > 1 ? ist : 0; k1.len = s1 - s0; k1.channels = G; k1.frames = nf; k1.nfft = N; k1.hop = H;
      k1.spec = s.spec.p; k1.spec_chan_rows = nf; k1.window = plan->win_a.p; k1.tw1 = plan->tw1.p; k1.sa = plan->sa.p;
      k1.pre = plan->pre.p; k1.counter = plan->counter_for(s.stream); k1.chunk_steps = 64; k1.scrub = plan->k1_scrub; k1.tiny2 = plan->tiny2;
      k1.generic = plan->generic;
      CU(launch_stft_fwd(k1, s.stream));
      const long long row0 = static_cast<long long>(c) * F + k_lo;     // first spectrum row of the slab (planar mode)
      if (fused) {
        K23Params k{};
        fill_k23(k, plan, s.stream);
        k.spec = s.spec.p; k.spec_chan_rows = nf; k.frames_per_chan = nf; k.channels = G; k.t_offset = t0;
        k.frame_t0 = s0 - t0; k.out = s.out.p; k.out_chan_stride = G > 1 ? ost : 0; k.n_out = t1 - t0;
        if (k.sgain) k.sgain += (G > 1 ? 0 : row0) * k.sgain_row_stride;
        CU(launch_istft_ola(k, s.stream));
      } else {
        K2Params k2{};
        k2.spec = s.spec.p; k2.rows = nf * G; k2.nfft = N; k2.frames = s.frames.p;
        k2.window = plan->syn_active ? plan->win_s.p : nullptr; k2.tw1 = plan->tw1.p; k2.sa = plan->sa.p; k2.pre = plan->pre.p;
        k2.scrub = plan->d.scrub;
        k2.sgain = plan->sgain ? plan->sgain + (G > 1 ? 0 : row0) * plan->sgain_row_stride : nullptr;
        k2.sgain_row_stride = plan->sgain_row_stride; k2.generic = plan->generic;
        CU(launch_istft_inv(k2, s.stream));
        K3Params k3{};
        k3.frames = s.frames.p; k3.frames_chan_rows = nf; k3.frames_per_chan = nf; k3.channels = G; k3.nfft = N; k3.hop = H;
        k3.window = nullptr; k3.gain = plan->d.gain; k3.eps = plan->d.eps; k3.norm = plan->env.p;
        k3.ring_len = static_cast<long long>(plan->ring_len); k3.t_offset = t0; k3.frame_t0 = s0 - t0; k3.out = s.out.p;
        k3.out_chan_stride = G > 1 ? ost : 0; k3.n_out = t1 - t0;
        CU(launch_ola_gather(k3, s.stream));
      }
      if (aos) {
        CU(launch_interleave(s.out.p, ost, t1 - t0, G, s.aos.p, s.stream));
        CU(cudaMemcpyAsync(h_out + t0 * G, s.aos.p, static_cast<size_t>(t1 - t0) * G * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
      } else {
        CU(cudaMemcpyAsync(yout + t0, s.out.p, static_cast<size_t>(t1 - t0) * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
      }
    }
  }
  for (auto& s : plan->slot) if (s.stream) CU(cudaStreamSynchronize(s.stream));
  return CRLOT_OK;
}

int crlot_roundtrip_host(crlot_plan_t plan, const float* h_in, long long chan_stride, long long len, int channels,
                         float* h_out, long long out_chan_stride, long long* n_out_per_chan) {
  return roundtrip_host_impl(plan, h_in, chan_stride, len, channels, h_out, out_chan_stride, n_out_per_chan, false);
}

int crlot_roundtrip_host_interleaved(crlot_plan_t plan, const float* h_in, long long len, int channels, float* h_out,
                                     long long* n_out_per_chan) {
  return roundtrip_host_impl(plan, h_in, 0, len, channels, h_out, 0, n_out_per_chan, true);
}

// FrameQueue front end from host buffers: kernel 1 over FrameQueue's frames, slab by slab through the same three
// workspace slots (H2D of slab i+1 and D2H of the spectra of slab i-1 overlap kernel 1 of slab i).  A slab carries the
// signal range its frames touch; the frame positions are shifted by giving kernel 1 the pad that is left at the slab's
// first frame (centre pad minus the samples skipped -- zero for interior slabs), so interior slabs run the plain bulk-copy
// path and only the slabs that contain a signal end compute pad positions (a reflection never reaches further than
// nfft/2 < one slab from the end it bounces off).
int crlot_stft_host_framequeue(crlot_plan_t plan, const float* h_signal, long long chan_stride, long long len, int channels,
                               int center, int pad_mode, void* h_spec, long long* n_frames_out) {
  if (!plan || !h_signal || !h_spec) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (channels <= 0 || len < 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  if (pad_mode < CRLOT_PAD_CONSTANT || pad_mode > CRLOT_PAD_EDGE) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad pad mode");
  if (len >= (1ll << 31)) return fail(CRLOT_ERR_INVALID_ARGUMENT, "length out of range (the reference indexes with int)");
  const int N = plan->d.nfft, H = plan->d.hop;
  const long long F = crlot_framequeue_count(len, N, H, center);
  if (n_frames_out) *n_frames_out = F < 0 ? 0 : F;
  if (F <= 0) return CRLOT_OK;
  if (!plan->generic && H % 4 != 0) return fail(CRLOT_ERR_UNSUPPORTED, "hop must be a multiple of 4");
  DeviceGuard g(plan->d.device);
  const long long pad = center ? N / 2 : 0;
  // same slab geometry as the round trip, so the two entry points share their slots
  const long long overlap = (N + H - 1) / H - 1;
  const long long slab_hops = std::min<long long>(std::max<long long>((1ll << 24) / H, 4 * (overlap + 1)), F + overlap);
  const long long slab_frames = slab_hops + overlap;
  const long long ist = ((slab_frames - 1) * H + N + 3) / 4 * 4;
  const long long ost = (slab_hops * H + N + 3) / 4 * 4;
  const bool fused = !plan->generic && plan->k23_win_ok && istft_ola_supported(N, H, static_cast<long long>(plan->ring_len), 0, 0, ost, nullptr, plan->env.p);
  { const int rc = ensure_host_slots(plan, slab_frames, 1, false, fused, ist, ost); if (rc) return rc; }
  // (a signal short enough for a reflection to bounce off both ends has fewer frames than one slab holds: it goes as one
  // slab with both true ends in it)
  cpx* spec_out = static_cast<cpx*>(h_spec);
  int it = 0;
  for (int c = 0; c < channels; ++c) {
    const float* xin = h_signal + static_cast<long long>(c) * chan_stride;
    for (long long k_lo = 0; k_lo < F; k_lo += slab_frames, ++it) {
      HostSlot& s = plan->slot[it % crlot_plan_s::kHostSlots];
      const long long nf = std::min(slab_frames, F - k_lo);
      // signal range the slab's frames touch, clipped to what exists; a slab that reaches an end keeps that end
      long long s0 = k_lo * H - pad, s1 = (k_lo + nf - 1) * H + N - pad;
      if (s0 < 0) s0 = 0;
      if (s0 > len) s0 = len / 4 * 4;                                  // frames entirely inside the right pad
      if (s1 > len) s1 = len;
      if (pad_mode != CRLOT_PAD_CONSTANT && s1 == len) {
        // right-pad positions mirror / repeat samples up to nfft/2 before the end: keep them in the slab
        s0 = std::min(s0, std::max<long long>(0, (len - N) / 4 * 4));
      }
      if (s1 > s0)
        CU(cudaMemcpyAsync(s.in.p, xin + s0, static_cast<size_t>(s1 - s0) * sizeof(float), cudaMemcpyHostToDevice, s.stream));
      K1Params k1{};
      k1.sig = s.in.p; k1.chan_stride = 0; k1.len = s1 - s0; k1.channels = 1; k1.frames = nf; k1.nfft = N; k1.hop = H;
      k1.pad = static_cast<int>(pad + s0 - k_lo * H); k1.pad_mode = pad_mode;
      k1.spec = s.spec.p; k1.spec_chan_rows = nf; k1.window = plan->win_a.p; k1.tw1 = plan->tw1.p; k1.sa = plan->sa.p;
      k1.pre = plan->pre.p; k1.counter = plan->counter_for(s.stream); k1.chunk_steps = 64; k1.scrub = plan->k1_scrub; k1.tiny2 = plan->tiny2;
      k1.generic = plan->generic;
      CU(launch_stft_fwd(k1, s.stream));
      CU(cudaMemcpyAsync(spec_out + (static_cast<long long>(c) * F + k_lo) * plan->bins, s.spec.p,
                         static_cast<size_t>(nf) * plan->bins * sizeof(cpx), cudaMemcpyDeviceToHost, s.stream));
    }
  }
  for (auto& s : plan->slot) if (s.stream) CU(cudaStreamSynchronize(s.stream));
  return CRLOT_OK;
}

int crlot_deinterleave(const float* d_interleaved, long long n, int channels, float* d_planar, long long chan_stride,
                       int device, void* stream) {
  if (n < 0 || channels <= 0 || chan_stride < n) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  if (n > 0 && (!d_interleaved || !d_planar)) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  DeviceGuard g(device);
  CU(launch_deinterleave(d_interleaved, n, channels, d_planar, chan_stride, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}

int crlot_interleave(const float* d_planar, long long chan_stride, long long n, int channels, float* d_interleaved,
                     int device, void* stream) {
  if (n < 0 || channels <= 0 || chan_stride < n) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  if (n > 0 && (!d_interleaved || !d_planar)) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  DeviceGuard g(device);
  CU(launch_interleave(d_planar, chan_stride, n, channels, d_interleaved, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}

int crlot_synth_uniform(float* d_out, long long chan_stride, int channels, long long n, unsigned long long seed,
                        unsigned long long channel0, unsigned long long i0, int device, void* stream) {
  if (!d_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  DeviceGuard g(device);
  CU(launch_synth_uniform(d_out, chan_stride, channels, n, seed, channel0, i0, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// drop-in FFT plan (IFftPlan over host pointers)
// ---------------------------------------------------------------------------------------------
struct crlot_fft_s {
  int nfft = 0, batch = 0, si = 0, so = 0, device = 0;
  bool complex_domain = false, generic = false;
  DevArr<cpx> ctw, d_cin, d_cout;   // complex-domain plan: table of the 2n-point geometry, batch buffers
  crlot_plan_t plan = nullptr;
  DevArr<float> d_time;
  DevArr<cpx> d_freq;
  PinnedArr<float> h_time, h_freq;   // page-locked staging: copies are asynchronous, one synchronize per call
};

extern "C" {

int crlot_fft_plan_create(int nfft, int in_place, int batch, int stride_in, int stride_out, int device, crlot_fft_t* out) {
  if (!out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  // same order of checks and messages as the reference adapter (kissfft_adapter.cc:13-39)
  if (batch < 1 || batch > 16) return fail(CRLOT_ERR_RUNTIME, "Batch size must be between 1 and 16");
  if (stride_in < 1 || stride_out < 1) return fail(CRLOT_ERR_RUNTIME, "Stride must be at least 1");
  if (in_place) return fail(CRLOT_ERR_RUNTIME, "In-place FFT is not yet supported");
  if (nfft <= 0) return fail(CRLOT_ERR_RUNTIME, "FFT size must be greater than 0");
  if (nfft % 2 != 0) return fail(CRLOT_ERR_RUNTIME, "FFT size must be even for real FFT");
  if (!supported_nfft(nfft) && !generic_len_supported(nfft / 2))
    return fail(CRLOT_ERR_UNSUPPORTED, "real nfft must be <= 16384 on the GPU path");
  std::unique_ptr<crlot_fft_s> f(new (std::nothrow) crlot_fft_s);
  if (!f) return fail(CRLOT_ERR_ALLOC, "host allocation failed");
  f->nfft = nfft; f->batch = batch; f->si = stride_in; f->so = stride_out; f->device = device;
  crlot_stft_desc d{};
  d.nfft = nfft; d.hop = nfft; d.boundary_mode = CRLOT_DROP; d.apply_window_inside = 0; d.gain = 1.0f; d.eps = 1e-8f;
  d.scrub = 1; d.device = device;
  int rc = crlot_stft_plan_create(&d, &f->plan);
  if (rc) return rc;
  DeviceGuard g(device);
  cudaError_t e = f->d_time.alloc(static_cast<size_t>(batch) * nfft);
  if (e == cudaSuccess) e = f->d_freq.alloc(static_cast<size_t>(batch) * (nfft / 2 + 1));
  if (e == cudaSuccess) e = f->h_time.resize(static_cast<size_t>(batch) * nfft);
  if (e == cudaSuccess) e = f->h_freq.resize(static_cast<size_t>(batch) * (nfft / 2 + 1) * 2);
  if (e != cudaSuccess) { crlot_stft_plan_destroy(f->plan); return fail_cuda(e, "cudaHostAlloc"); }
  *out = f.release();
  return CRLOT_OK;
}
int crlot_fft_plan_destroy(crlot_fft_t f) {
  if (!f) return CRLOT_OK;
  resident_forget(f);
  if (f->plan) crlot_stft_plan_destroy(f->plan);
  { DeviceGuard g(f->device); delete f; }
  return CRLOT_OK;
}
// ---- complex-domain plan (ref: kissfft_adapter.cc:13-28,56-66,171-246) ------------------
int crlot_fft_plan_create_complex(int nfft, int in_place, int batch, int stride_in, int stride_out, int device,
                                  crlot_fft_t* out) {
  if (!out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (batch < 1 || batch > 16) return fail(CRLOT_ERR_RUNTIME, "Batch size must be between 1 and 16");
  if (stride_in < 1 || stride_out < 1) return fail(CRLOT_ERR_RUNTIME, "Stride must be at least 1");
  if (in_place) return fail(CRLOT_ERR_RUNTIME, "In-place FFT is not yet supported");
  if (nfft <= 0) return fail(CRLOT_ERR_RUNTIME, "FFT size must be greater than 0");
  const bool generic = nfft != 128 && nfft != 256 && nfft != 512 && nfft != 1024 && nfft != 2048;
  if (generic && !generic_len_supported(nfft))
    return fail(CRLOT_ERR_UNSUPPORTED, "complex nfft must be <= 8192 on the GPU path");
  std::unique_ptr<crlot_fft_s> f(new (std::nothrow) crlot_fft_s);
  if (!f) return fail(CRLOT_ERR_ALLOC, "host allocation failed");
  f->nfft = nfft; f->batch = batch; f->si = stride_in; f->so = stride_out; f->device = device; f->complex_domain = true; f->generic = generic;
  DeviceGuard g(device);
  if (!g.ok) return fail(CRLOT_ERR_CUDA, "cannot select device");
  const FftTables T = generic ? make_generic_tables(2 * nfft) : make_fft_tables(2 * nfft);
  cudaError_t e = upload_cpx(f->ctw, T.tw1);
  if (e == cudaSuccess) e = f->d_cin.alloc(static_cast<size_t>(batch) * nfft);
  if (e == cudaSuccess) e = f->d_cout.alloc(static_cast<size_t>(batch) * nfft);
  if (e == cudaSuccess) e = f->h_freq.resize(static_cast<size_t>(batch) * nfft * 2);
  if (e != cudaSuccess) return fail_cuda(e, "cudaMalloc");
  *out = f.release();
  return CRLOT_OK;
}

// contiguous rows on the device (B200 extension: any number of rows per call)
int crlot_fft_complex_device(crlot_fft_t f, const void* d_in, void* d_out, long long rows, int inverse, void* stream) {
  if (!f || !d_in || !d_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (!f->complex_domain) return fail(CRLOT_ERR_RUNTIME, "Complex FFT not supported for Real domain plan");
  if (rows < 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  if (reinterpret_cast<uintptr_t>(d_in) % 8 != 0 || reinterpret_cast<uintptr_t>(d_out) % 8 != 0)
    return fail(CRLOT_ERR_UNSUPPORTED, "buffers must be 8-byte aligned");
  DeviceGuard g(f->device);
  CfftParams k{};
  k.in = static_cast<const cpx*>(d_in); k.out = static_cast<cpx*>(d_out); k.rows = rows; k.n = f->nfft;
  k.inverse = inverse ? 1 : 0; k.tw1 = f->ctw.p; k.generic = f->generic ? 1 : 0;
  CU(launch_cfft(k, static_cast<cudaStream_t>(stream)));
  return CRLOT_OK;
}

// batch b reads in + b*si*n with element stride si, writes out + b*so*n with element stride so
// (ref: kissfft_adapter.cc:183-201,219-243)
static int fft_complex_host(crlot_fft_t f, const float* h_in_ri, float* h_out_ri, int batch, int inverse) {
  if (!f || !h_in_ri || !h_out_ri) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (!f->complex_domain) return fail(CRLOT_ERR_RUNTIME, "Complex FFT not supported for Real domain plan");
  if (batch < 1 || batch > f->batch) return fail(CRLOT_ERR_RUNTIME, "Invalid batch size");
  const int n = f->nfft;
  for (int b = 0; b < batch; ++b)
    for (int i = 0; i < n; ++i) {
      const long long s = 2 * (static_cast<long long>(b) * f->si * n + static_cast<long long>(i) * f->si);
      f->h_freq[2 * (static_cast<size_t>(b) * n + i)] = h_in_ri[s];
      f->h_freq[2 * (static_cast<size_t>(b) * n + i) + 1] = h_in_ri[s + 1];
    }
  DeviceGuard g(f->device);
  CU(cudaMemcpy(f->d_cin.p, f->h_freq.data(), sizeof(float) * 2 * batch * n, cudaMemcpyHostToDevice));
  int rc = crlot_fft_complex_device(f, f->d_cin.p, f->d_cout.p, batch, inverse, nullptr);
  if (rc) return rc;
  CU(cudaMemcpy(f->h_freq.data(), f->d_cout.p, sizeof(float) * 2 * batch * n, cudaMemcpyDeviceToHost));
  for (int b = 0; b < batch; ++b)
    for (int i = 0; i < n; ++i) {
      const long long o = 2 * (static_cast<long long>(b) * f->so * n + static_cast<long long>(i) * f->so);
      h_out_ri[o] = f->h_freq[2 * (static_cast<size_t>(b) * n + i)];
      h_out_ri[o + 1] = f->h_freq[2 * (static_cast<size_t>(b) * n + i) + 1];
    }
  return CRLOT_OK;
}
int crlot_fft_forward_complex(crlot_fft_t f, const float* h_in_ri, float* h_out_ri, int batch) {
  return fft_complex_host(f, h_in_ri, h_out_ri, batch, 0);
}
int crlot_fft_inverse_complex(crlot_fft_t f, const float* h_in_ri, float* h_out_ri, int batch) {
  return fft_complex_host(f, h_in_ri, h_out_ri, batch, 1);
}

// forward: batch b reads in + b*si*n with element stride si; writes out + b*so*(n/2+1) with
// element stride so (ref: kissfft_adapter.cc:97-98,103,118).  One launch for the whole batch; copies through
// page-locked staging on the default stream, ONE synchronize at the end.
int crlot_fft_forward(crlot_fft_t f, const float* h_in, float* h_out_ri, int batch) {
  if (!f || !h_in || !h_out_ri) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (f->complex_domain) return fail(CRLOT_ERR_RUNTIME, "Real FFT not supported for Complex domain plan");
  if (batch < 1 || batch > f->batch) return fail(CRLOT_ERR_RUNTIME, "Invalid batch size");
  const int n = f->nfft, nb = n / 2 + 1;
  DeviceGuard g(f->device);
  const size_t in_bytes = sizeof(float) * batch * n, out_bytes = sizeof(float) * 2 * batch * nb;
  const void* dev_in = f->si == 1 ? resident_lookup(h_in, in_bytes, f->device) : nullptr;   // an earlier call's output?
  if (dev_in) {
    if (dev_in != f->d_time.p) {       // another object's buffer: our page-locked mirror no longer matches d_time
      CU(cudaMemcpyAsync(f->d_time.p, dev_in, in_bytes, cudaMemcpyDeviceToDevice, nullptr));
      resident_forget(f);
    }
  } else {
    if (f->si == 1) std::memcpy(f->h_time.data(), h_in, in_bytes);
    else
      for (int b = 0; b < batch; ++b)
        for (int i = 0; i < n; ++i) f->h_time[static_cast<size_t>(b) * n + i] = h_in[static_cast<long long>(b) * f->si * n + static_cast<long long>(i) * f->si];
    CU(cudaMemcpyAsync(f->d_time.p, f->h_time.data(), in_bytes, cudaMemcpyHostToDevice, nullptr));
  }
  int rc = crlot_stft_execute(f->plan, f->d_time.p, 0, static_cast<long long>(batch) * n, 1, batch, f->d_freq.p, nullptr);
  if (rc) return rc;
  CU(cudaMemcpyAsync(f->h_freq.data(), f->d_freq.p, out_bytes, cudaMemcpyDeviceToHost, nullptr));
  CU(cudaStreamSynchronize(nullptr));
  if (f->so == 1) {
    std::memcpy(h_out_ri, f->h_freq.data(), out_bytes);
    resident_publish(0, h_out_ri, out_bytes, f->d_freq.p, f->h_freq.data(), f->device, f);
  } else {
    for (int b = 0; b < batch; ++b)
      for (int i = 0; i < nb; ++i) {
        const long long o = 2 * (static_cast<long long>(b) * f->so * nb + static_cast<long long>(i) * f->so);
        h_out_ri[o] = f->h_freq[2 * (static_cast<size_t>(b) * nb + i)];
        h_out_ri[o + 1] = f->h_freq[2 * (static_cast<size_t>(b) * nb + i) + 1];
      }
  }
  return CRLOT_OK;
}
// inverse: batch b reads in + b*si*(n/2+1) with element stride si; writes out + b*so*n with
// element stride so (ref: kissfft_adapter.cc:139-147,154-165)
int crlot_fft_inverse(crlot_fft_t f, const float* h_in_ri, float* h_out, int batch) {
  if (!f || !h_in_ri || !h_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (f->complex_domain) return fail(CRLOT_ERR_RUNTIME, "Real FFT not supported for Complex domain plan");
  if (batch < 1 || batch > f->batch) return fail(CRLOT_ERR_RUNTIME, "Invalid batch size");
  const int n = f->nfft, nb = n / 2 + 1;
  DeviceGuard g(f->device);
  const size_t in_bytes = sizeof(float) * 2 * batch * nb, out_bytes = sizeof(float) * batch * n;
  const void* dev_in = f->si == 1 ? resident_lookup(h_in_ri, in_bytes, f->device) : nullptr;   // untouched forward() output?
  if (dev_in) {
    if (dev_in != f->d_freq.p) {
      CU(cudaMemcpyAsync(f->d_freq.p, dev_in, in_bytes, cudaMemcpyDeviceToDevice, nullptr));
      resident_forget(f);
    }
  } else {
    if (f->si == 1) std::memcpy(f->h_freq.data(), h_in_ri, in_bytes);
    else
      for (int b = 0; b < batch; ++b)
        for (int i = 0; i < nb; ++i) {
          const long long q = 2 * (static_cast<long long>(b) * f->si * nb + static_cast<long long>(i) * f->si);
          f->h_freq[2 * (static_cast<size_t>(b) * nb + i)] = h_in_ri[q];
          f->h_freq[2 * (static_cast<size_t>(b) * nb + i) + 1] = h_in_ri[q + 1];
        }
    CU(cudaMemcpyAsync(f->d_freq.p, f->h_freq.data(), in_bytes, cudaMemcpyHostToDevice, nullptr));
  }
  int rc = crlot_istft_execute(f->plan, f->d_freq.p, batch, f->d_time.p, 0, nullptr);
  if (rc) return rc;
  CU(cudaMemcpyAsync(f->h_time.data(), f->d_time.p, out_bytes, cudaMemcpyDeviceToHost, nullptr));
  CU(cudaStreamSynchronize(nullptr));
  if (f->so == 1) {
    std::memcpy(h_out, f->h_time.data(), out_bytes);
    resident_publish(1, h_out, out_bytes, f->d_time.p, f->h_time.data(), f->device, f);
  } else {
    for (int b = 0; b < batch; ++b)
      for (int i = 0; i < n; ++i) h_out[static_cast<long long>(b) * f->so * n + static_cast<long long>(i) * f->so] = f->h_time[static_cast<size_t>(b) * n + i];
  }
  return CRLOT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// drop-in OLAAccumulator state: device rings, reference bookkeeping on the host
// ---------------------------------------------------------------------------------------------
struct crlot_ola_s {
  int sample_rate = 0, device = 0;
  size_t N = 0, H = 0, C = 0, ring_len = 0;
  float eps = 1e-8f;
  bool inside = false, shadow = false, has_window = false;
  size_t read_pos = 0, produced = 0;
  float peak = 0.0f;
  DevArr<float> ring, norm, win, call_win, stage;
  PinnedArr<float> h_stage;           // page-locked staging of frames in / samples out
  bool stage_busy = false;            // an asynchronous upload from h_stage may still be in flight
  std::vector<float> h_norm, h_win;
  void rebuild_norm() {   // ref: OLAAccumulator::initialize_normalization
    h_norm.resize(ring_len);
    envelope_impl(h_norm.data(), has_window ? h_win.data() : nullptr, ring_len, N, H, inside, eps);
  }
};

extern "C" {

int crlot_ola_create(int sample_rate, size_t frame_size, size_t hop_size, size_t channels, float eps,
                     int apply_window_inside, int shadow_ring, int device, crlot_ola_t* out) {
  if (!out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (!(sample_rate > 0 && frame_size > 0 && hop_size > 0 && channels > 0 && eps > 0.0f))
    return fail(CRLOT_ERR_INVALID_ARGUMENT, "Invalid OLA configuration");
  std::unique_ptr<crlot_ola_s> o(new (std::nothrow) crlot_ola_s);
  if (!o) return fail(CRLOT_ERR_ALLOC, "host allocation failed");
  o->sample_rate = sample_rate; o->device = device; o->N = frame_size; o->H = hop_size; o->C = channels;
  o->eps = eps; o->inside = apply_window_inside != 0; o->shadow = shadow_ring != 0;
  o->ring_len = crlot_ola_ring_len(frame_size, hop_size);
  DeviceGuard g(device);
  if (!g.ok) return fail(CRLOT_ERR_CUDA, "cannot select device");
  CU(o->ring.alloc(o->ring_len * channels));
  CU(cudaMemset(o->ring.p, 0, o->ring_len * channels * sizeof(float)));
  o->rebuild_norm();
  CU(o->norm.upload(o->h_norm.data(), o->ring_len));
  CU(o->win.alloc(frame_size));
  CU(o->call_win.alloc(frame_size));
  CU(o->stage.alloc(std::max(o->ring_len, frame_size) * channels));
  CU(o->h_stage.resize(std::max(o->ring_len, frame_size) * channels));
  *out = o.release();
  return CRLOT_OK;
}
int crlot_ola_destroy(crlot_ola_t o) {
  if (!o) return CRLOT_OK;
  { DeviceGuard g(o->device); delete o; }
  return CRLOT_OK;
}
int crlot_ola_set_window(crlot_ola_t o, const float* w, int wlen) {
  if (!o) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (!w) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Window pointer cannot be null");
  if (wlen != static_cast<int>(o->N)) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Window size must match frame size");
  DeviceGuard g(o->device);
  o->h_win.assign(w, w + wlen);
  o->has_window = true;
  CU(cudaMemcpy(o->win.p, w, sizeof(float) * wlen, cudaMemcpyHostToDevice));
  o->rebuild_norm();
  CU(cudaMemcpy(o->norm.p, o->h_norm.data(), sizeof(float) * o->ring_len, cudaMemcpyHostToDevice));
  return CRLOT_OK;
}

// Frames are accumulated ASYNCHRONOUSLY (nothing comes back to the host): the upload from the page-locked stage and
// the ring kernel are enqueued on the default stream and the call returns; the next call that rewrites the stage,
// and produce(), order themselves behind them.  d_src != nullptr: the samples are already on the device (the
// caller handed over an untouched IFftPlan::inverse output, see Resident above) -- no upload at all.
static int ola_wait_stage(crlot_ola_t o) {
  if (o->stage_busy) { CU(cudaStreamSynchronize(nullptr)); o->stage_busy = false; }
  return CRLOT_OK;
}
static int ola_add_staged(crlot_ola_t o, const float* window, size_t start, size_t off, size_t size, float gain,
                          const float* d_src) {
  // h_stage (or d_src) holds [C][size] planar samples already offset by `off`
  const bool use = o->inside ? o->has_window : (window != nullptr);
  const float* dwin = nullptr;
  if (use) {
    if (o->inside) dwin = o->win.p + off;
    else {
      CU(cudaMemcpyAsync(o->call_win.p, window + off, sizeof(float) * size, cudaMemcpyHostToDevice, nullptr));
      CU(cudaStreamSynchronize(nullptr));        // `window` is the caller's (pageable) memory
      dwin = o->call_win.p;
    }
  }
  const size_t n = std::min(size, o->ring_len);    // RingBuffer::split clamps to capacity
  if (!d_src) {
    CU(cudaMemcpyAsync(o->stage.p, o->h_stage.data(), sizeof(float) * size * o->C, cudaMemcpyHostToDevice, nullptr));
    o->stage_busy = true;
    d_src = o->stage.p;
  }
  CU(launch_ring_axpy(o->ring.p, static_cast<long long>(o->ring_len), static_cast<int>(o->C), d_src,
                      static_cast<long long>(size), dwin, gain, static_cast<long long>(start % o->ring_len),
                      static_cast<int>(n), nullptr));
  o->produced = std::max(o->produced, start + size);
  return CRLOT_OK;
}

int crlot_ola_add_frame_soa(crlot_ola_t o, const float* const* ch_frames, const float* window, size_t start_sample,
                            size_t start_off, size_t size, float gain) {
  if (!o) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (!ch_frames) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Channel frames pointer cannot be null");
  if (size == 0 || start_off >= o->N) return CRLOT_OK;
  if (start_off + size > o->N) size = o->N - start_off;
  for (size_t c = 0; c < o->C; ++c)
    if (!ch_frames[c]) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Channel frame pointer cannot be null");
  DeviceGuard g(o->device);
  if (o->C == 1) {
    const void* dev = resident_lookup(ch_frames[0], sizeof(float) * o->N, o->device);
    if (dev) return ola_add_staged(o, window, start_sample, start_off, size, gain, static_cast<const float*>(dev) + start_off);
  }
  int rc = ola_wait_stage(o);
  if (rc) return rc;
  for (size_t c = 0; c < o->C; ++c) std::memcpy(o->h_stage.data() + c * size, ch_frames[c] + start_off, sizeof(float) * size);
  return ola_add_staged(o, window, start_sample, start_off, size, gain, nullptr);
}
int crlot_ola_push_frame_aos(crlot_ola_t o, const float* interleaved, const float* window, size_t start_sample,
                             size_t start_off, size_t size, float gain) {
  if (!o) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (!interleaved) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Interleaved input pointer cannot be null");
  if (size == 0 || start_off >= o->N) return CRLOT_OK;
  if (start_off + size > o->N) size = o->N - start_off;
  DeviceGuard g(o->device);
  // the reference re-enters add_frame_SoA with start_off = 0 on the de-interleaved scratch, so the
  // window (stored or per-call) is indexed from 0 there (OLAAccumulator.cc:159)
  if (o->C == 1) {                                          // one channel: interleaved == planar
    const void* dev = resident_lookup(interleaved, sizeof(float) * o->N, o->device);
    if (dev) return ola_add_staged(o, window, start_sample, 0, size, gain, static_cast<const float*>(dev) + start_off);
  }
  int rc = ola_wait_stage(o);
  if (rc) return rc;
  const float* src = interleaved + start_off * o->C;      // ref: aos_to_soa.cc:13-17
  for (size_t c = 0; c < o->C; ++c)
    for (size_t i = 0; i < size; ++i) o->h_stage[c * size + i] = src[i * o->C + c];
  return ola_add_staged(o, window, start_sample, 0, size, gain, nullptr);
}
int crlot_ola_produce(crlot_ola_t o, float* const* ch_out, size_t n, size_t* produced_out) {
  if (produced_out) *produced_out = 0;
  if (!o) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (!ch_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Output channel buffer cannot be null");
  if (n == 0) return CRLOT_OK;
  for (size_t c = 0; c < o->C; ++c)
    if (!ch_out[c]) return fail(CRLOT_ERR_INVALID_ARGUMENT, "Output channel buffer cannot be null");
  const size_t avail = o->produced > o->read_pos ? o->produced - o->read_pos : 0;
  if (avail == 0) return CRLOT_OK;
  if (avail < n) n = avail;
  const size_t m = std::min(n, o->ring_len);
  DeviceGuard g(o->device);
  CU(launch_ring_produce(o->ring.p, static_cast<long long>(o->ring_len), static_cast<int>(o->C), o->norm.p, o->eps,
                         static_cast<long long>(o->read_pos % o->ring_len), static_cast<int>(m), o->stage.p,
                         static_cast<long long>(m), nullptr));
  CU(cudaMemcpyAsync(o->h_stage.data(), o->stage.p, sizeof(float) * m * o->C, cudaMemcpyDeviceToHost, nullptr));
  CU(cudaStreamSynchronize(nullptr));
  o->stage_busy = false;
  for (size_t c = 0; c < o->C; ++c) std::memcpy(ch_out[c], o->h_stage.data() + c * m, sizeof(float) * m);
  o->read_pos = (o->read_pos + n) % o->ring_len;
  for (size_t i = 0; i < m; ++i) o->peak = std::max(o->peak, std::fabs(ch_out[0][i]));
  if (produced_out) *produced_out = n;
  return CRLOT_OK;
}
int crlot_ola_flush(crlot_ola_t o) {
  if (!o) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  o->produced = std::max(o->produced, o->read_pos + o->N);
  return CRLOT_OK;
}
int crlot_ola_reset(crlot_ola_t o) {
  if (!o) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  DeviceGuard g(o->device);
  CU(cudaMemset(o->ring.p, 0, o->ring_len * o->C * sizeof(float)));
  o->read_pos = 0; o->produced = 0; o->peak = 0.0f;
  o->has_window = false; o->h_win.clear();               // reset() also forgets the window (OLAAccumulator.cc:244-246)
  o->rebuild_norm();
  CU(cudaMemcpy(o->norm.p, o->h_norm.data(), sizeof(float) * o->ring_len, cudaMemcpyHostToDevice));
  return CRLOT_OK;
}
size_t crlot_ola_get(crlot_ola_t o, int what) {
  if (!o) return 0;
  switch (what) {
    case 0: return o->produced;
    case 1: return o->read_pos;
    case 2: return o->ring_len;
    case 3: return o->has_window ? 1 : 0;
    default: return 0;
  }
}
float crlot_ola_meter_peak(crlot_ola_t o) { return o ? o->peak : 0.0f; }

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// streaming state (config 5): device rings + counters, host mirror of the bookkeeping
// ---------------------------------------------------------------------------------------------
struct crlot_stream_s {
  crlot_plan_t plan = nullptr;
  int S = 0, max_chunk = 0, rcap = 0, max_nf = 0;
  long long w = 0, k = 0;            // host mirror of the device counters
  int last_nf = 0;
  DevArr<StreamCounters> ctr;
  DevArr<unsigned int> done;         // block ticket of the one-launch tick
  DevArr<float> ring, acc, frames;
  DevArr<cpx> spec;
  DevArr<float> h2d, d2h;            // staging of crlot_stream_tick_host (pageable host buffers)
  cudaStream_t host_stream = nullptr;
  static constexpr int kTickets = 64;  // completion events of crlot_stream_submit_host, a ring
  cudaEvent_t done_ev[kTickets] = {};
  unsigned long long submitted = 0;
  // three-stream pipeline of crlot_stream_submit_host: H2D of tick t+1 (copy_in) || kernel of tick t (host_stream) ||
  // D2H of tick t-1 (copy_out) through kPipe device staging slots
  static constexpr int kPipe = 4;
  cudaStream_t copy_in = nullptr, copy_out = nullptr;
  DevArr<float> pin[kPipe], pout[kPipe];
  cudaEvent_t ev_in[kPipe] = {}, ev_k[kPipe] = {}, ev_out[kPipe] = {};
  unsigned long long piped = 0;
  bool poisoned = false;             // a launch failed mid-tick: host mirrors and device counters may disagree
  // instantiated graphs of crlot_stream_run, kept across calls and re-pointed with cudaGraphExecUpdate
  struct CachedGraph { int key; cudaGraphExec_t exec; };
  std::vector<CachedGraph> graphs;
};

namespace {

// CRLOT_STREAM_NO_FUSED_TICK=1 keeps every tick on the five-launch path (A/B timing and the
// bit-identity test of the one-kernel tick); read once
bool stream_fused_tick_enabled() {
  static const bool on = getenv("CRLOT_STREAM_NO_FUSED_TICK") == nullptr;
  return on;
}

// CRLOT_STREAM_NO_PDL=1: one-kernel ticks without programmatic dependent launch (A/B timing)
bool stream_pdl_enabled() {
  static const bool on = getenv("CRLOT_STREAM_NO_PDL") == nullptr;
  return on;
}

int stream_nf(const crlot_stream_s* s, long long w_new) {
  const int N = s->plan->d.nfft, H = s->plan->d.hop;
  const long long total = w_new >= N ? (w_new - N) / H + 1 : 0;   // Framer DROP count, framer.cc:106-115
  return static_cast<int>(total - s->k);
}

// the launches of one tick; every argument is tick-independent given (w0, k_base, nf)
cudaError_t stream_tick_launches(crlot_stream_s* s, const float* d_in, long long in_stride, int chunk_len, long long w0,
                                 float* d_out, long long out_stride, long long k_base, int nf, cudaStream_t st) {
  crlot_plan_s* p = s->plan;
  const int N = p->d.nfft, H = p->d.hop;
  // steady state of a low-latency stream (one new frame per stream): the whole tick in one launch
  if (nf == 1 && stream_fused_tick_enabled() &&
      stream_tick1_supported(N, H, chunk_len, static_cast<long long>(p->ring_len), out_stride, d_out)) {
    StreamTick1Params t{};
    t.ctr = s->ctr.p; t.done = s->done.p; t.chunk = d_in; t.chunk_stride = in_stride; t.w0 = w0; t.chunk_len = chunk_len;
    t.ring = s->ring.p; t.rcap = s->rcap; t.n_streams = s->S; t.nfft = N; t.hop = H; t.spec = s->spec.p;
    t.window = p->win_a.p; t.syn_window = p->syn_active ? p->win_s.p : nullptr; t.tw1 = p->tw1.p; t.scrub = p->d.scrub;
    t.gain = p->d.gain; t.eps = p->d.eps; t.norm = p->env.p; t.ring_len = static_cast<long long>(p->ring_len);
    t.acc = s->acc.p; t.out = d_out; t.out_stride = out_stride; t.k_base = k_base;
    t.pdl = stream_pdl_enabled() ? 1 : 0;
    return launch_stream_tick1(t, st);
  }
  cudaError_t e = launch_stream_ingest(s->ctr.p, d_in, in_stride, chunk_len, s->ring.p, s->rcap, s->S, w0, st);
  if (e != cudaSuccess) return e;
  if (nf > 0) {
    if (!p->generic && stream_fwd_supported(N)) {
      StreamFwdParams f{};
      f.ctr = s->ctr.p; f.ring = s->ring.p; f.rcap = s->rcap; f.n_streams = s->S; f.nf = nf; f.nfft = N; f.hop = H;
      f.spec = s->spec.p; f.window = p->win_a.p; f.tw1 = p->tw1.p; f.sa = p->sa.p; f.scrub = p->d.scrub;
      e = launch_stream_fwd(f, st);
    } else {
      // nfft = 4096 and the any-size plans: gather the frames out of the rings (into the buffer kernel 2 overwrites
      // afterwards) and run kernel 1 over them as non-overlapping frames
      e = launch_stream_gather(s->ctr.p, s->ring.p, s->rcap, s->S, nf, N, H, s->frames.p, st);
      if (e != cudaSuccess) return e;
      K1Params k{};
      k.sig = s->frames.p; k.chan_stride = 0; k.len = static_cast<long long>(s->S) * nf * N; k.channels = 1;
      k.frames = static_cast<long long>(s->S) * nf; k.nfft = N; k.hop = N; k.spec = s->spec.p; k.spec_chan_rows = k.frames;
      k.window = p->win_a.p; k.tw1 = p->tw1.p; k.sa = p->sa.p; k.pre = p->pre.p; k.counter = p->counter_for(st);
      k.chunk_steps = 64; k.scrub = p->k1_scrub; k.tiny2 = p->tiny2; k.generic = p->generic;
      e = launch_stft_fwd(k, st);
    }
    if (e != cudaSuccess) return e;
    K2Params k2{};
    k2.spec = s->spec.p; k2.rows = static_cast<long long>(s->S) * nf; k2.nfft = N; k2.frames = s->frames.p;
    k2.window = p->syn_active ? p->win_s.p : nullptr; k2.tw1 = p->tw1.p; k2.sa = p->sa.p; k2.pre = p->pre.p;
    k2.scrub = p->d.scrub; k2.generic = p->generic;
    e = launch_istft_inv(k2, st);
    if (e != cudaSuccess) return e;
    StreamOlaParams o{};
    o.ctr = s->ctr.p; o.frames = s->frames.p; o.n_streams = s->S; o.nf = nf; o.nfft = N; o.hop = H;
    o.span = N + (nf - 1) * H; o.n_prod = nf * H; o.window = nullptr; o.gain = p->d.gain; o.eps = p->d.eps;
    o.norm = p->env.p; o.ring_len = static_cast<long long>(p->ring_len); o.acc = s->acc.p; o.out = d_out;
    o.out_stride = out_stride; o.k_base = k_base;
    e = launch_stream_ola(o, st);
    if (e != cudaSuccess) return e;
  }
  return launch_stream_advance(s->ctr.p, chunk_len, nf, st);
}

int stream_check_io(const crlot_stream_s* s, const void* d_in, long long in_stride, int chunk_len, const void* d_out,
                    long long out_stride) {
  if (!s || !d_in || !d_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (s->poisoned) return fail(CRLOT_ERR_RUNTIME, "streaming state is inconsistent after a failed launch: call crlot_stream_reset");
  if (chunk_len < 0 || chunk_len > s->max_chunk) return fail(CRLOT_ERR_INVALID_ARGUMENT, "chunk_len exceeds max_chunk");
  if (out_stride % 4 != 0 || reinterpret_cast<uintptr_t>(d_out) % 16 != 0)
    return fail(CRLOT_ERR_UNSUPPORTED, "output must be 16-byte aligned with out_stride a multiple of 4 floats");
  (void)in_stride;
  return CRLOT_OK;
}

}  // namespace

extern "C" {

int crlot_stream_create(crlot_plan_t plan, int n_streams, int max_chunk, crlot_stream_t* out) {
  if (!plan || !out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (n_streams <= 0 || max_chunk <= 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "n_streams and max_chunk must be positive");
  if (plan->d.boundary_mode != CRLOT_DROP)
    return fail(CRLOT_ERR_UNSUPPORTED, "streaming needs BoundaryMode::DROP (chunked ZERO_PAD emits premature frames)");
  // every plan shape runs (the register-FFT sizes 256..2048 on the ring-reading kernels, nfft = 4096 and the any-size
  // plans through a gather + kernel 1); the ring kernels move 16-byte groups
  if (plan->d.nfft % 4 != 0 || plan->d.hop % 4 != 0)
    return fail(CRLOT_ERR_UNSUPPORTED, "streaming needs nfft % 4 == 0 and hop % 4 == 0");
  const int N = plan->d.nfft, H = plan->d.hop;
  std::unique_ptr<crlot_stream_s> s(new (std::nothrow) crlot_stream_s);
  if (!s) return fail(CRLOT_ERR_ALLOC, "host allocation failed");
  s->plan = plan; s->S = n_streams; s->max_chunk = max_chunk;
  int rc = 4;
  while (rc < N + max_chunk) rc <<= 1;
  s->rcap = rc;
  s->max_nf = max_chunk / H + 1;
  // the accumulate span of one tick must fit the accumulator ring without lapping itself
  if (static_cast<size_t>(N) + static_cast<size_t>(s->max_nf - 1) * H > plan->ring_len)
    return fail(CRLOT_ERR_UNSUPPORTED, "max_chunk too large for the OLA ring (more than 20 hops per tick)");
  DeviceGuard g(plan->d.device);
  CU(s->ctr.alloc(1));
  CU(s->done.alloc(1));
  CU(cudaMemset(s->done.p, 0, sizeof(unsigned int)));
  CU(s->ring.alloc(static_cast<size_t>(n_streams) * rc));
  CU(s->acc.alloc(static_cast<size_t>(n_streams) * plan->ring_len));
  CU(s->spec.alloc(static_cast<size_t>(n_streams) * s->max_nf * plan->bins));
  CU(s->frames.alloc(static_cast<size_t>(n_streams) * s->max_nf * N));
  CU(cudaMemset(s->ctr.p, 0, sizeof(StreamCounters)));
  CU(cudaMemset(s->ring.p, 0, sizeof(float) * n_streams * rc));
  CU(cudaMemset(s->acc.p, 0, sizeof(float) * n_streams * plan->ring_len));
  *out = s.release();
  return CRLOT_OK;
}

int crlot_stream_destroy(crlot_stream_t s) {
  if (!s) return CRLOT_OK;
  {
    DeviceGuard g(s->plan->d.device);
    cudaDeviceSynchronize();                       // graph launches / submitted ticks may still be in flight
    for (auto& c : s->graphs) cudaGraphExecDestroy(c.exec);
    for (auto& e : s->done_ev) if (e) cudaEventDestroy(e);
    for (int i = 0; i < crlot_stream_s::kPipe; ++i)
      for (cudaEvent_t e : {s->ev_in[i], s->ev_k[i], s->ev_out[i]}) if (e) cudaEventDestroy(e);
    if (s->copy_in) cudaStreamDestroy(s->copy_in);
    if (s->copy_out) cudaStreamDestroy(s->copy_out);
    if (s->host_stream) cudaStreamDestroy(s->host_stream);
    delete s;
  }
  return CRLOT_OK;
}

int crlot_stream_reset(crlot_stream_t s, void* stream) {
  if (!s) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  DeviceGuard g(s->plan->d.device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CU(cudaMemsetAsync(s->ctr.p, 0, sizeof(StreamCounters), st));
  CU(cudaMemsetAsync(s->ring.p, 0, sizeof(float) * s->S * s->rcap, st));
  CU(cudaMemsetAsync(s->acc.p, 0, sizeof(float) * s->S * s->plan->ring_len, st));
  s->w = s->k = 0; s->last_nf = 0; s->poisoned = false;
  return CRLOT_OK;
}

int crlot_stream_frames_ready(crlot_stream_t s, int chunk_len) {
  if (!s || chunk_len < 0) return 0;
  return stream_nf(s, s->w + chunk_len);
}

int crlot_stream_tick(crlot_stream_t s, const float* d_chunk, long long chunk_stride, int chunk_len, float* d_out,
                      long long out_stride, int* n_out, void* stream) {
  if (n_out) *n_out = 0;
  int rc = stream_check_io(s, d_chunk, chunk_stride, chunk_len, d_out, out_stride);
  if (rc) return rc;
  DeviceGuard g(s->plan->d.device);
  const int nf = stream_nf(s, s->w + chunk_len);
  const cudaError_t e = stream_tick_launches(s, d_chunk, chunk_stride, chunk_len, s->w, d_out, out_stride, s->k, nf,
                                             static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) { s->poisoned = true; return fail_cuda(e, "crlot_stream_tick"); }
  s->w += chunk_len; s->k += nf; s->last_nf = nf;
  if (n_out) *n_out = nf * s->plan->d.hop;
  return CRLOT_OK;
}

int crlot_stream_run(crlot_stream_t s, const float* d_in, long long in_stride, int chunk_len, long long n_ticks,
                     float* d_out, long long out_stride, long long* n_out, void* stream) {
  if (n_out) *n_out = 0;
  int rc = stream_check_io(s, d_in, in_stride, chunk_len, d_out, out_stride);
  if (rc) return rc;
  if (n_ticks < 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  DeviceGuard g(s->plan->d.device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long w0 = s->w, k_base = s->k;
  // One instantiated graph per distinct frames-per-tick value (the launches of a tick depend on the tick only
  // through the device counters).  The executables live on the stream object: a later call re-captures the same
  // launches with its own pointers / offsets and re-points the executable with cudaGraphExecUpdate (the topology
  // is identical), which costs microseconds where instantiation costs hundreds; launches already enqueued are not
  // affected by an update.  The call is asynchronous with respect to `stream`.
  const bool use_graph = n_ticks >= 16;
  std::vector<int> fresh;                      // keys already re-pointed at this call's arguments
  int status = CRLOT_OK;
  for (long long t = 0; t < n_ticks && status == CRLOT_OK; ++t) {
    const int nf = stream_nf(s, s->w + chunk_len);
    cudaError_t e = cudaSuccess;
    // steady state (chunk == hop, one frame per stream per tick from here on): the tick is a single
    // kernel whose arguments do not depend on the tick, so kBurst consecutive ticks are replayed as ONE
    // graph of kBurst kernel nodes -- the host pays one graph launch per burst instead of per tick
    constexpr int kBurst = 32;
    int burst = 1;
    if (use_graph && nf == 1 && chunk_len == s->plan->d.hop && n_ticks - t >= kBurst && stream_fused_tick_enabled() &&
        !s->plan->generic &&
        stream_tick1_supported(s->plan->d.nfft, s->plan->d.hop, chunk_len, static_cast<long long>(s->plan->ring_len), out_stride, d_out))
      burst = kBurst;
    const int key = burst > 1 ? -burst : nf;
    if (!use_graph) {
      e = stream_tick_launches(s, d_in, in_stride, chunk_len, w0, d_out, out_stride, k_base, nf, st);
    } else {
      cudaGraphExec_t exec = nullptr;
      size_t slot = s->graphs.size();
      for (size_t i = 0; i < s->graphs.size(); ++i) if (s->graphs[i].key == key) { exec = s->graphs[i].exec; slot = i; }
      bool is_fresh = false;
      for (int k : fresh) if (k == key) is_fresh = true;
      if (!exec || !is_fresh) {
        // captured on a private stream (the caller's may be the legacy default stream, which
        // cannot capture); capturing only records the launches, the replay goes to `st`
        cudaGraph_t graph = nullptr;
        if (!s->host_stream) e = cudaStreamCreateWithFlags(&s->host_stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaStreamBeginCapture(s->host_stream, cudaStreamCaptureModeRelaxed);
        if (e == cudaSuccess) {
          cudaError_t e1 = cudaSuccess;
          for (int b = 0; b < burst && e1 == cudaSuccess; ++b)
            e1 = stream_tick_launches(s, d_in, in_stride, chunk_len, w0, d_out, out_stride, k_base, nf, s->host_stream);
          e = cudaStreamEndCapture(s->host_stream, &graph);
          if (e1 != cudaSuccess) e = e1;
        }
        if (e == cudaSuccess && exec) {
          cudaGraphExecUpdateResultInfo info{};
          if (cudaGraphExecUpdate(exec, graph, &info) != cudaSuccess) {      // topology changed: rebuild
            cudaGetLastError();
            cudaStreamSynchronize(st);
            cudaGraphExecDestroy(exec);
            exec = nullptr;
            s->graphs.erase(s->graphs.begin() + static_cast<long>(slot));
          }
        }
        if (e == cudaSuccess && !exec) {
          e = cudaGraphInstantiate(&exec, graph, 0);
          if (e == cudaSuccess) s->graphs.push_back({key, exec});
        }
        if (graph) cudaGraphDestroy(graph);
        if (e == cudaSuccess) fresh.push_back(key);
      }
      if (e == cudaSuccess) e = cudaGraphLaunch(exec, st);
    }
    if (e != cudaSuccess) { s->poisoned = true; status = fail_cuda(e, "crlot_stream_run"); break; }
    s->w += static_cast<long long>(burst) * chunk_len; s->k += static_cast<long long>(burst) * nf; s->last_nf = nf;
    t += burst - 1;
  }
  if (n_out) *n_out = (s->k - k_base) * s->plan->d.hop;
  return status;
}

// Device-visible alias of a page-locked host pointer (cudaHostAlloc / cudaHostRegister memory), else nullptr.
static void* mapped_alias(const void* h) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, h) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
}

// One tick with host buffers, enqueued on the state's own stream.  Page-locked buffers with suitable alignment are
// handed to the tick kernel AS THEY ARE (zero-copy: the kernel reads the chunk and writes the hop across PCIe
// itself, so a tick is one launch and no copy-engine round trips); anything else is staged through device buffers
// with two copies.
static int stream_tick_host_enqueue(crlot_stream_t s, const float* h_chunk, long long chunk_stride, int chunk_len,
                                    float* h_out, long long out_stride, int* n_out) {
  if (n_out) *n_out = 0;
  if (!s || !h_chunk || !h_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (s->poisoned) return fail(CRLOT_ERR_RUNTIME, "streaming state is inconsistent after a failed launch: call crlot_stream_reset");
  if (chunk_len < 0 || chunk_len > s->max_chunk) return fail(CRLOT_ERR_INVALID_ARGUMENT, "chunk_len exceeds max_chunk");
  const int H = s->plan->d.hop;
  if (!s->host_stream) CU(cudaStreamCreateWithFlags(&s->host_stream, cudaStreamNonBlocking));
  cudaStream_t st = s->host_stream;
  const int nf = stream_nf(s, s->w + chunk_len);
  static const bool zero_copy_on = getenv("CRLOT_STREAM_NO_ZERO_COPY") == nullptr;
  const float* zc_in = zero_copy_on ? static_cast<const float*>(mapped_alias(h_chunk)) : nullptr;
  float* zc_out = zero_copy_on ? static_cast<float*>(mapped_alias(h_out)) : nullptr;
  if (zc_in && zc_out && out_stride % 4 == 0 && reinterpret_cast<uintptr_t>(zc_out) % 16 == 0) {
    const cudaError_t e = stream_tick_launches(s, zc_in, chunk_stride, chunk_len, s->w, zc_out, out_stride, s->k, nf, st);
    if (e != cudaSuccess) { s->poisoned = true; return fail_cuda(e, "crlot_stream_tick_host"); }
  } else {
    const long long in_pitch = (s->max_chunk + 3) / 4 * 4, out_pitch = static_cast<long long>(s->max_nf) * H;
    if (!s->h2d.p) {
      CU(s->h2d.alloc(static_cast<size_t>(s->S) * in_pitch));
      CU(s->d2h.alloc(static_cast<size_t>(s->S) * out_pitch));
    }
    CU(cudaMemcpy2DAsync(s->h2d.p, in_pitch * sizeof(float), h_chunk, chunk_stride * sizeof(float),
                         static_cast<size_t>(chunk_len) * sizeof(float), s->S, cudaMemcpyHostToDevice, st));
    const cudaError_t e = stream_tick_launches(s, s->h2d.p, in_pitch, chunk_len, s->w, s->d2h.p, out_pitch, s->k, nf, st);
    if (e != cudaSuccess) { s->poisoned = true; return fail_cuda(e, "crlot_stream_tick_host"); }
    if (nf > 0)
      CU(cudaMemcpy2DAsync(h_out, out_stride * sizeof(float), s->d2h.p, out_pitch * sizeof(float),
                           static_cast<size_t>(nf) * H * sizeof(float), s->S, cudaMemcpyDeviceToHost, st));
  }
  s->w += chunk_len; s->k += nf; s->last_nf = nf;
  if (n_out) *n_out = nf * H;
  return CRLOT_OK;
}

int crlot_stream_tick_host(crlot_stream_t s, const float* h_chunk, long long chunk_stride, int chunk_len, float* h_out,
                           long long out_stride, int* n_out) {
  if (!s) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  DeviceGuard g(s->plan->d.device);
  const int rc = stream_tick_host_enqueue(s, h_chunk, chunk_stride, chunk_len, h_out, out_stride, n_out);
  if (rc) return rc;
  CU(cudaStreamSynchronize(s->host_stream));
  return CRLOT_OK;
}

// submit on three streams: the chunk's H2D copy, the tick and the hop's D2H copy of consecutive tickets overlap
static int stream_submit_pipelined(crlot_stream_t s, const float* h_chunk, long long chunk_stride, int chunk_len,
                                   float* h_out, long long out_stride, int* n_out, cudaEvent_t done) {
  const int H = s->plan->d.hop;
  const long long in_pitch = (s->max_chunk + 3) / 4 * 4, out_pitch = static_cast<long long>(s->max_nf) * H;
  if (!s->copy_in) {
    CU(cudaStreamCreateWithFlags(&s->copy_in, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&s->copy_out, cudaStreamNonBlocking));
    for (int i = 0; i < crlot_stream_s::kPipe; ++i) {
      CU(s->pin[i].alloc(static_cast<size_t>(s->S) * in_pitch));
      CU(s->pout[i].alloc(static_cast<size_t>(s->S) * out_pitch));
      CU(cudaEventCreateWithFlags(&s->ev_in[i], cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&s->ev_k[i], cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&s->ev_out[i], cudaEventDisableTiming));
    }
  }
  const int slot = static_cast<int>(s->piped % crlot_stream_s::kPipe);
  const bool reused = s->piped >= crlot_stream_s::kPipe;
  const int nf = stream_nf(s, s->w + chunk_len);
  // H2D: the slot's previous tick must have finished reading its chunk
  if (reused) CU(cudaStreamWaitEvent(s->copy_in, s->ev_k[slot], 0));
  if (chunk_stride == chunk_len && in_pitch == chunk_len)
    CU(cudaMemcpyAsync(s->pin[slot].p, h_chunk, static_cast<size_t>(s->S) * chunk_len * sizeof(float), cudaMemcpyHostToDevice, s->copy_in));
  else
    CU(cudaMemcpy2DAsync(s->pin[slot].p, in_pitch * sizeof(float), h_chunk, chunk_stride * sizeof(float),
                         static_cast<size_t>(chunk_len) * sizeof(float), s->S, cudaMemcpyHostToDevice, s->copy_in));
  CU(cudaEventRecord(s->ev_in[slot], s->copy_in));
  // tick: after its chunk has landed and the slot's previous output has left
  CU(cudaStreamWaitEvent(s->host_stream, s->ev_in[slot], 0));
  if (reused) CU(cudaStreamWaitEvent(s->host_stream, s->ev_out[slot], 0));
  const cudaError_t e = stream_tick_launches(s, s->pin[slot].p, in_pitch, chunk_len, s->w, s->pout[slot].p, out_pitch, s->k, nf, s->host_stream);
  if (e != cudaSuccess) { s->poisoned = true; return fail_cuda(e, "crlot_stream_submit_host"); }
  CU(cudaEventRecord(s->ev_k[slot], s->host_stream));
  // D2H
  CU(cudaStreamWaitEvent(s->copy_out, s->ev_k[slot], 0));
  if (nf > 0) {
    const size_t row = static_cast<size_t>(nf) * H;
    if (out_stride == static_cast<long long>(row) && out_pitch == static_cast<long long>(row))
      CU(cudaMemcpyAsync(h_out, s->pout[slot].p, static_cast<size_t>(s->S) * row * sizeof(float), cudaMemcpyDeviceToHost, s->copy_out));
    else
      CU(cudaMemcpy2DAsync(h_out, out_stride * sizeof(float), s->pout[slot].p, out_pitch * sizeof(float), row * sizeof(float), s->S,
                           cudaMemcpyDeviceToHost, s->copy_out));
  }
  CU(cudaEventRecord(s->ev_out[slot], s->copy_out));
  CU(cudaEventRecord(done, s->copy_out));
  ++s->piped;
  s->w += chunk_len; s->k += nf; s->last_nf = nf;
  if (n_out) *n_out = nf * H;
  return CRLOT_OK;
}

int crlot_stream_submit_host(crlot_stream_t s, const float* h_chunk, long long chunk_stride, int chunk_len, float* h_out,
                             long long out_stride, int* n_out, unsigned long long* ticket) {
  if (n_out) *n_out = 0;
  if (!s || !ticket || !h_chunk || !h_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (s->poisoned) return fail(CRLOT_ERR_RUNTIME, "streaming state is inconsistent after a failed launch: call crlot_stream_reset");
  if (chunk_len < 0 || chunk_len > s->max_chunk) return fail(CRLOT_ERR_INVALID_ARGUMENT, "chunk_len exceeds max_chunk");
  DeviceGuard g(s->plan->d.device);
  if (!s->host_stream) CU(cudaStreamCreateWithFlags(&s->host_stream, cudaStreamNonBlocking));
  cudaEvent_t& ev = s->done_ev[s->submitted % crlot_stream_s::kTickets];
  if (!ev) CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  // CRLOT_STREAM_SUBMIT=zerocopy: the tick kernel reads / writes the page-locked buffers itself (one stream);
  // default: the three-stream copy pipeline (copies of neighbouring ticks overlap the tick kernel)
  static const bool zc = getenv("CRLOT_STREAM_SUBMIT") != nullptr && std::string(getenv("CRLOT_STREAM_SUBMIT")) == "zerocopy";
  if (zc) {
    const int rc = stream_tick_host_enqueue(s, h_chunk, chunk_stride, chunk_len, h_out, out_stride, n_out);
    if (rc) return rc;
    CU(cudaEventRecord(ev, s->host_stream));
  } else {
    const int rc = stream_submit_pipelined(s, h_chunk, chunk_stride, chunk_len, h_out, out_stride, n_out, ev);
    if (rc) return rc;
  }
  *ticket = s->submitted++;
  return CRLOT_OK;
}

int crlot_stream_wait_host(crlot_stream_t s, unsigned long long ticket) {
  if (!s) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (ticket >= s->submitted) return fail(CRLOT_ERR_INVALID_ARGUMENT, "no such ticket");
  DeviceGuard g(s->plan->d.device);
  // a ticket older than the event ring has certainly completed once its slot's current occupant has
  const unsigned long long newest = s->submitted - 1;
  const unsigned long long t = newest - ticket >= crlot_stream_s::kTickets ? newest : ticket;
  CU(cudaEventSynchronize(s->done_ev[t % crlot_stream_s::kTickets]));
  return CRLOT_OK;
}

int crlot_stream_drain(crlot_stream_t s, float* d_out, long long out_stride, int* n_out, void* stream) {
  if (n_out) *n_out = 0;
  if (!s || !d_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  const int N = s->plan->d.nfft, H = s->plan->d.hop;
  if (s->k == 0 || N == H) return CRLOT_OK;
  if (out_stride % 4 != 0 || reinterpret_cast<uintptr_t>(d_out) % 16 != 0)
    return fail(CRLOT_ERR_UNSUPPORTED, "output must be 16-byte aligned with out_stride a multiple of 4 floats");
  DeviceGuard g(s->plan->d.device);
  StreamOlaParams o{};
  o.ctr = s->ctr.p; o.frames = s->frames.p; o.n_streams = s->S; o.nf = 0; o.nfft = N; o.hop = H;
  o.span = N - H; o.n_prod = N - H; o.window = nullptr; o.gain = s->plan->d.gain; o.eps = s->plan->d.eps;
  o.norm = s->plan->env.p; o.ring_len = static_cast<long long>(s->plan->ring_len); o.acc = s->acc.p; o.out = d_out;
  o.out_stride = out_stride; o.k_base = s->k;
  CU(launch_stream_ola(o, static_cast<cudaStream_t>(stream)));
  if (n_out) *n_out = N - H;
  return CRLOT_OK;
}

int crlot_stream_last_spectra(crlot_stream_t s, void** d_spec, int* frames_per_stream) {
  if (!s || !d_spec || !frames_per_stream) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  *d_spec = s->spec.p;
  *frames_per_stream = s->last_nf;
  return CRLOT_OK;
}

long long crlot_stream_get(crlot_stream_t s, int what) {
  if (!s) return 0;
  switch (what) {
    case 0: return s->w;
    case 1: return s->k;
    case 2: return s->rcap;
    default: return 0;
  }
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// round trip on STATIC scratch: no allocation inside the call (no counterpart in the reference; added
// for callers that cannot call cudaMalloc -- code under CUDA graph capture, tight serving
// loops, harnesses that forbid it).  The work counter of the dynamically scheduled kernels, the spectra and (when the
// fused kernel 2+3 does not cover the plan) the iFFT frames live in one file-scope device array that every GPU which
// loads the library holds until exit.  All calls share it: calls that overlap on the GPU race.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[CRLOT_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxStaticDevices = 64;
constexpr size_t kStaticHeader = 256;       // [0, 256): work-counter words; the sections after it are 256-byte aligned
std::mutex g_static_mu;
unsigned char* g_static_ptr[kMaxStaticDevices] = {};

// The device address of g_static_scratch on the CURRENT device, resolved once per device.  In host code the array's
// name is a host-side stand-in, never a device address: only cudaGetSymbolAddress yields what a kernel may be given.
int static_scratch_for(int device, unsigned char** out) {
  if (device < 0 || device >= kMaxStaticDevices) return fail(CRLOT_ERR_UNSUPPORTED, "device ordinal beyond the static-scratch table");
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetSymbolAddress(static scratch)");
    if (!p) return fail(CRLOT_ERR_CUDA, "cudaGetSymbolAddress returned a null device address");
    g_static_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_static_ptr[device];
  return CRLOT_OK;
}
size_t round256(size_t b) { return (b + 255) & ~static_cast<size_t>(255); }
}  // namespace

extern "C" int crlot_roundtrip_execute_static(crlot_plan_t plan, const float* d_signal, long long chan_stride, long long len,
                                              int channels, float* d_out, long long out_chan_stride, void* stream) {
  if (!plan || !d_signal || !d_out) return fail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (channels <= 0 || len < 0) return fail(CRLOT_ERR_INVALID_ARGUMENT, "bad sizes");
  if (!plan->generic && (chan_stride % 4 != 0 || reinterpret_cast<uintptr_t>(d_signal) % 16 != 0))
    return fail(CRLOT_ERR_UNSUPPORTED, "signal must be 16-byte aligned with chan_stride a multiple of 4 floats");
  const int N = plan->d.nfft, H = plan->d.hop;
  const long long F = crlot_frame_count(len, N, H, plan->d.boundary_mode);
  if (F == 0) return CRLOT_OK;
  const long long n_out = (F - 1) * H + N;
  const bool fused = crlot_istft_ola_supported(plan, 0, 0, d_out, out_chan_stride) != 0;
  const size_t rows = static_cast<size_t>(F) * static_cast<size_t>(channels);
  const size_t spec_bytes = round256(rows * static_cast<size_t>(plan->bins) * sizeof(cpx));
  const size_t frame_bytes = fused ? 0 : round256(rows * static_cast<size_t>(N) * sizeof(float));
  const size_t need = kStaticHeader + spec_bytes + frame_bytes;
  if (need > static_cast<size_t>(CRLOT_STATIC_SCRATCH_BYTES))
    return fail(CRLOT_ERR_INVALID_ARGUMENT,
                "crlot_roundtrip_execute_static: " + std::to_string(rows) + " frames need " + std::to_string(need) +
                    " bytes of scratch, the static scratch holds CRLOT_STATIC_SCRATCH_BYTES = " +
                    std::to_string(static_cast<size_t>(CRLOT_STATIC_SCRATCH_BYTES)) +
                    "; split the signal or use crlot_roundtrip_execute with caller-provided buffers");
  DeviceGuard g(plan->d.device);
  if (!g.ok) return fail(CRLOT_ERR_CUDA, "cannot select device");
  unsigned char* base = nullptr;
  if (int rc = static_scratch_for(plan->d.device, &base)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned int* counter = reinterpret_cast<unsigned int*>(base);   // zeroed on `st` by each launcher before its kernel
  cpx* spec = reinterpret_cast<cpx*>(base + kStaticHeader);
  float* frames = reinterpret_cast<float*>(base + kStaticHeader + spec_bytes);

  K1Params k1{};
  k1.sig = d_signal; k1.chan_stride = chan_stride; k1.len = len; k1.channels = channels; k1.frames = F;
  k1.nfft = N; k1.hop = H; k1.spec = spec; k1.spec_chan_rows = F;
  k1.window = plan->win_a.p; k1.tw1 = plan->tw1.p; k1.sa = plan->sa.p; k1.pre = plan->pre.p;
  k1.counter = counter; k1.chunk_steps = 64; k1.scrub = plan->k1_scrub; k1.tiny2 = plan->tiny2; k1.generic = plan->generic;
  CU(launch_stft_fwd(k1, st));
  if (fused) {
    K23Params k{};
    k.nfft = N; k.hop = H; k.window = plan->syn_active ? plan->win_s.p : plan->win_ones.p;
    k.tw1 = plan->tw1.p; k.sa = plan->sa.p; k.pre = plan->pre.p; k.scrub = plan->d.scrub;
    k.gain = plan->d.gain; k.eps = plan->d.eps; k.norm = plan->env.p; k.ring_len = static_cast<long long>(plan->ring_len);
    k.den_rcp = plan->den_rcp.p; k.counter = counter; k.chunk_blocks = 0;
    k.sgain = plan->sgain; k.sgain_row_stride = plan->sgain_row_stride;
    k.spec = spec; k.spec_chan_rows = F; k.frames_per_chan = F; k.channels = channels;
    k.t_offset = 0; k.frame_t0 = 0; k.out = d_out; k.out_chan_stride = out_chan_stride; k.n_out = n_out;
    CU(launch_istft_ola(k, st));
    return CRLOT_OK;
  }
  K2Params k2{};
  k2.spec = spec; k2.rows = static_cast<long long>(rows); k2.nfft = N; k2.frames = frames;
  k2.window = plan->syn_active ? plan->win_s.p : nullptr;
  k2.tw1 = plan->tw1.p; k2.sa = plan->sa.p; k2.pre = plan->pre.p; k2.scrub = plan->d.scrub;
  k2.sgain = plan->sgain; k2.sgain_row_stride = plan->sgain_row_stride; k2.generic = plan->generic;
  CU(launch_istft_inv(k2, st));
  K3Params k3{};
  k3.frames = frames; k3.frames_chan_rows = F; k3.frames_per_chan = F; k3.channels = channels; k3.nfft = N; k3.hop = H;
  k3.window = nullptr; k3.gain = plan->d.gain; k3.eps = plan->d.eps; k3.norm = plan->env.p;
  k3.ring_len = static_cast<long long>(plan->ring_len); k3.t_offset = 0; k3.frame_t0 = 0;
  k3.out = d_out; k3.out_chan_stride = out_chan_stride; k3.n_out = n_out;
  CU(launch_ola_gather(k3, st));
  return CRLOT_OK;
}
