// framer.cu -- device-resident dsp::Framer (C-ABI crlot_framer_*).
//
// The reference Framer (ref: dsp/frame/framer.cc:15-182) is a growable interleaved host vector with
// a read and a write cursor.  Here the samples live in HBM, de-interleaved once on entry into a
// planar [channel][capacity] buffer, so that kernel 1 (k1_stft.cu) can address frame k of channel
// c as buf[c*cap + read + k*hop + (0..N)) with no further copy: "frame extraction" is addressing.
// pop() to a host pointer is kept for drop-in compatibility (one small D2H copy per frame);
// crlot_framer_stft() is the fast path that consumes every available frame through kernel 1.
//
// Bookkeeping (available frames, cursor advance, the clamp at the write cursor, ZERO_PAD's
// one-frame-at-a-time tail, DROP) follows the reference line by line in per-channel sample units;
// the reference's vector capacity rule is mirrored in integers only so that buffer_size() reports
// what the reference would.
#include <cuda_runtime.h>

#include <algorithm>
#include <new>
#include <string>
#include <vector>

#include "../../include/crlot_cuda.h"

namespace crlot_framer_detail {

// [frames][channels] interleaved -> planar [channels][cap] at offset w
__global__ void deinterleave_kernel(const float* __restrict__ src, float* __restrict__ dst, long long cap,
                                    long long w, long long frames, int channels) {
  const long long total = frames * channels;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long t = i / channels;
    const int c = (int)(i - t * channels);
    dst[c * cap + w + t] = src[i];          // coalesced read; writes are C interleaved streams
  }
}

// planar [channels][cap] from offset r, `avail` valid samples -> [n][channels] interleaved, zero beyond avail
__global__ void interleave_kernel(const float* __restrict__ src, float* __restrict__ dst, long long cap,
                                  long long r, long long avail, long long n, int channels) {
  const long long total = n * channels;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long t = i / channels;
    const int c = (int)(i - t * channels);
    dst[i] = t < avail ? src[c * cap + r + t] : 0.0f;
  }
}

}  // namespace crlot_framer_detail

using namespace crlot_framer_detail;

struct crlot_framer_s {
  size_t N = 0, H = 0, C = 1;
  int mode = CRLOT_ZERO_PAD;
  int device = 0;
  // device storage: planar [C][cap]; r, w in per-channel samples
  float* buf = nullptr;
  long long cap = 0, r = 0, w = 0;
  float* stage = nullptr;       // device staging for interleaved transfers
  size_t stage_floats = 0;
  // the reference's own cursors/capacity in interleaved samples (buffer_size() parity only)
  size_t ref_size = 0, ref_r = 0, ref_w = 0;
  // Stream ordering: push()/pop() run on the legacy default stream, push_device()/stft() on the caller's, which
  // may be non-blocking.  Every operation that touches the store first makes its stream wait for the event the
  // previous operation recorded (order_after_last) and records its own on completion (mark_done), so that e.g. the
  // compaction of ensure_room never overwrites samples a kernel 1 launched earlier on another stream still reads.
  cudaEvent_t ev = nullptr;
  bool ev_pending = false;
  // pop() read-ahead (one channel): the per-frame drop-in loop pops frame after frame, each only hop samples
  // further on.  One D2H copy brings kAheadHops hops of the store into page-locked host memory and the following
  // pops are served from it (the samples of [lo, hi) of the device store, which only ever grows behind w).
  static constexpr long long kAheadHops = 64;
  float* ahead = nullptr;
  long long ahead_cap = 0, ahead_lo = 0, ahead_hi = 0;   // store indices covered; empty when lo == hi
};

// the library-wide last-error slot lives in capi.cu
extern "C" void crlot_cuda_set_error_(const char* msg);

namespace {

int ffail(int code, const char* msg) {
  crlot_cuda_set_error_(msg);
  return code;
}
int ffail_cuda(cudaError_t e, const char* where) {
  std::string m = std::string(where) + ": " + cudaGetErrorString(e);
  cudaGetLastError();
  crlot_cuda_set_error_(m.c_str());
  return e == cudaErrorMemoryAllocation ? CRLOT_ERR_ALLOC : CRLOT_ERR_CUDA;
}
#define FCU(call)                                           \
  do {                                                      \
    cudaError_t e__ = (call);                               \
    if (e__ != cudaSuccess) return ffail_cuda(e__, #call);  \
  } while (0)

struct Dev {
  int prev = -1;
  explicit Dev(int d) { cudaGetDevice(&prev); if (prev != d) cudaSetDevice(d); }
  ~Dev() { if (prev >= 0) cudaSetDevice(prev); }
};

long long round_up4(long long v) { return (v + 3) & ~3LL; }

int order_after_last(crlot_framer_s* f, cudaStream_t st) {
  if (f->ev_pending) FCU(cudaStreamWaitEvent(st, f->ev, 0));
  return CRLOT_OK;
}
int mark_done(crlot_framer_s* f, cudaStream_t st) {
  FCU(cudaEventRecord(f->ev, st));
  f->ev_pending = true;
  return CRLOT_OK;
}

// ref: Framer::calculate_available_frames, framer.cc:88-118 (per-channel sample units)
size_t available_now(const crlot_framer_s* f) {
  if (f->w <= f->r) return 0;
  const size_t avail = (size_t)(f->w - f->r);
  if (avail < f->N) return (f->mode == CRLOT_ZERO_PAD && avail > 0) ? 1 : 0;
  return (avail - f->N) / f->H + 1;      // the DROP correction of :110-115 can never trigger here
}

// frames that repeated pop() would deliver from the current state
long long drainable_now(const crlot_framer_s* f) {
  return crlot_frame_count(f->w - f->r, (int)f->N, (int)f->H, f->mode);
}

int ensure_stage(crlot_framer_s* f, size_t floats) {
  if (f->stage_floats >= floats) return CRLOT_OK;
  if (f->stage) cudaFree(f->stage);
  f->stage = nullptr;
  f->stage_floats = 0;
  FCU(cudaMalloc(&f->stage, floats * sizeof(float)));
  f->stage_floats = floats;
  return CRLOT_OK;
}

// make room for `add` more samples per channel behind w: compact to the front (cursor stays a
// multiple of 4 so that kernel 1's 16-byte accesses remain aligned), then grow by doubling
int ensure_room(crlot_framer_s* f, long long add, cudaStream_t st) {
  if (f->w + add <= f->cap) return CRLOT_OK;
  const long long r4 = f->r & ~3LL;
  const long long keep = f->w - r4;
  long long need = round_up4(keep + add);
  if (need <= f->cap && r4 >= keep) {                      // non-overlapping move inside the buffer
    if (keep > 0)
      FCU(cudaMemcpy2DAsync(f->buf, f->cap * sizeof(float), f->buf + r4, f->cap * sizeof(float),
                            keep * sizeof(float), f->C, cudaMemcpyDeviceToDevice, st));
    f->r -= r4; f->w -= r4;
    f->ahead_lo = f->ahead_hi = 0;
    return CRLOT_OK;
  }
  long long ncap = std::max<long long>(need, 2 * f->cap);
  ncap = round_up4(ncap);
  float* nb = nullptr;
  FCU(cudaMalloc(&nb, (size_t)ncap * f->C * sizeof(float)));
  if (keep > 0) {
    cudaError_t e = cudaMemcpy2DAsync(nb, ncap * sizeof(float), f->buf + r4, f->cap * sizeof(float),
                                      keep * sizeof(float), f->C, cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) { cudaFree(nb); return ffail_cuda(e, "cudaMemcpy2DAsync"); }
  }
  FCU(cudaStreamSynchronize(st));
  if (f->buf) cudaFree(f->buf);
  f->buf = nb; f->cap = ncap;
  f->r -= r4; f->w -= r4;
  f->ahead_lo = f->ahead_hi = 0;
  return CRLOT_OK;
}

// move [r, w) to the front of a fresh buffer so that the read cursor is 16-byte aligned again
int realign(crlot_framer_s* f, cudaStream_t st) {
  float* nb = nullptr;
  FCU(cudaMalloc(&nb, (size_t)f->cap * f->C * sizeof(float)));
  const long long keep = f->w - f->r;
  if (keep > 0) {
    cudaError_t e = cudaMemcpy2DAsync(nb, f->cap * sizeof(float), f->buf + f->r, f->cap * sizeof(float),
                                      keep * sizeof(float), f->C, cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) { cudaFree(nb); return ffail_cuda(e, "cudaMemcpy2DAsync"); }
  }
  FCU(cudaStreamSynchronize(st));
  cudaFree(f->buf);
  f->buf = nb;
  f->w = keep; f->r = 0;
  f->ahead_lo = f->ahead_hi = 0;
  return CRLOT_OK;
}

void ref_reset(crlot_framer_s* f) {                        // framer.cc:76-86
  f->ref_r = f->ref_w = 0;
  f->ref_size = f->N * f->C * 2;
}
void ref_push(crlot_framer_s* f, size_t samples) {         // framer.cc:50-57,120-126
  const size_t req = f->ref_w + samples;
  if (f->ref_size < req) f->ref_size = std::max(req, f->ref_size * 2);
  f->ref_w += samples;
}
void ref_pop(crlot_framer_s* f) {                          // framer.cc:162-179
  f->ref_r += f->H * f->C;
  if (f->ref_r > f->ref_w) f->ref_r = f->ref_w;
  if (f->ref_r > f->ref_size / 2) {
    f->ref_w = f->ref_r < f->ref_w ? f->ref_w - f->ref_r : 0;
    f->ref_r = 0;
  }
}

}  // namespace

extern "C" {

int crlot_framer_create(size_t frame_size, size_t hop_size, size_t channels, int boundary_mode, int device,
                        crlot_framer_t* out) {
  if (!out) return ffail(CRLOT_ERR_INVALID_ARGUMENT, "out is null");
  if (frame_size == 0) return ffail(CRLOT_ERR_INVALID_ARGUMENT, "Frame size must be greater than 0");
  if (hop_size == 0) return ffail(CRLOT_ERR_INVALID_ARGUMENT, "Hop size must be greater than 0");
  if (channels == 0) return ffail(CRLOT_ERR_INVALID_ARGUMENT, "Channels must be greater than 0");
  if (boundary_mode != CRLOT_ZERO_PAD && boundary_mode != CRLOT_DROP)
    return ffail(CRLOT_ERR_INVALID_ARGUMENT, "bad boundary mode");
  int ndev = 0;
  FCU(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return ffail(CRLOT_ERR_CUDA, "no such CUDA device");
  auto* f = new (std::nothrow) crlot_framer_s;
  if (!f) return ffail(CRLOT_ERR_ALLOC, "host allocation failed");
  f->N = frame_size; f->H = hop_size; f->C = channels; f->mode = boundary_mode; f->device = device;
  Dev g(device);
  f->cap = round_up4((long long)frame_size * 2);
  cudaError_t e = cudaMalloc(&f->buf, (size_t)f->cap * channels * sizeof(float));
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&f->ev, cudaEventDisableTiming);
  if (e != cudaSuccess) { if (f->buf) cudaFree(f->buf); delete f; return ffail_cuda(e, "cudaMalloc"); }
  ref_reset(f);
  *out = f;
  return CRLOT_OK;
}

int crlot_framer_destroy(crlot_framer_t f) {
  if (!f) return CRLOT_OK;
  Dev g(f->device);
  if (f->ev_pending) cudaEventSynchronize(f->ev);
  if (f->ev) cudaEventDestroy(f->ev);
  if (f->buf) cudaFree(f->buf);
  if (f->stage) cudaFree(f->stage);
  if (f->ahead) cudaFreeHost(f->ahead);
  delete f;
  return CRLOT_OK;
}

int crlot_framer_reset(crlot_framer_t f) {
  if (!f) return ffail(CRLOT_ERR_INVALID_ARGUMENT, "null framer");
  f->r = f->w = 0;
  f->ahead_lo = f->ahead_hi = 0;
  ref_reset(f);
  return CRLOT_OK;
}

int crlot_framer_push(crlot_framer_t f, const float* h_interleaved, size_t frames) {
  if (!f) return ffail(CRLOT_ERR_INVALID_ARGUMENT, "null framer");
  if (!h_interleaved && frames > 0) return ffail(CRLOT_ERR_INVALID_ARGUMENT, "null input");   // framer.cc:41-43 -> false
  if (frames == 0) return CRLOT_OK;
  Dev g(f->device);
  cudaStream_t st = nullptr;
  int rc = order_after_last(f, st);
  if (rc) return rc;
  rc = ensure_room(f, (long long)frames, st);
  if (rc) return rc;
  if (f->C == 1) {
    FCU(cudaMemcpyAsync(f->buf + f->w, h_interleaved, frames * sizeof(float), cudaMemcpyHostToDevice, st));
  } else {
    rc = ensure_stage(f, frames * f->C);
    if (rc) return rc;
    FCU(cudaMemcpyAsync(f->stage, h_interleaved, frames * f->C * sizeof(float), cudaMemcpyHostToDevice, st));
    const long long total = (long long)frames * (long long)f->C;
    const int blocks = (int)std::min<long long>((total + 255) / 256, 148 * 8);
    deinterleave_kernel<<<blocks, 256, 0, st>>>(f->stage, f->buf, f->cap, f->w, (long long)frames, (int)f->C);
    FCU(cudaGetLastError());
  }
  FCU(cudaStreamSynchronize(st));     // the caller may reuse its host buffer on return, like std::copy
  f->ev_pending = false;              // everything issued so far has completed
  f->w += (long long)frames;
  ref_push(f, frames * f->C);
  return CRLOT_OK;
}

int crlot_framer_push_device(crlot_framer_t f, const float* d_planar, long long chan_stride, size_t frames,
                             void* stream) {
  if (!f) return ffail(CRLOT_ERR_INVALID_ARGUMENT, "null framer");
  if (!d_planar && frames > 0) return ffail(CRLOT_ERR_INVALID_ARGUMENT, "null input");
  if (frames == 0) return CRLOT_OK;
  Dev g(f->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = order_after_last(f, st);
  if (rc) return rc;
  rc = ensure_room(f, (long long)frames, st);
  if (rc) return rc;
  FCU(cudaMemcpy2DAsync(f->buf + f->w, f->cap * sizeof(float), d_planar, chan_stride * sizeof(float),
                        frames * sizeof(float), f->C, cudaMemcpyDeviceToDevice, st));
  rc = mark_done(f, st);
  if (rc) return rc;
  f->w += (long long)frames;
  ref_push(f, frames * f->C);
  return CRLOT_OK;
}

size_t crlot_framer_available(crlot_framer_t f) { return f ? available_now(f) : 0; }

int crlot_framer_pop(crlot_framer_t f, float* h_out_frame, int* popped) {
  if (popped) *popped = 0;
  if (!f) return ffail(CRLOT_ERR_INVALID_ARGUMENT, "null framer");
  if (!h_out_frame) return CRLOT_OK;                        // framer.cc:62-64 -> false
  if (available_now(f) == 0) return CRLOT_OK;               // framer.cc:129-131 -> false
  Dev g(f->device);
  cudaStream_t st = nullptr;
  {
    const int rc0 = order_after_last(f, st);
    if (rc0) return rc0;
  }
  const long long avail = std::min<long long>(f->w - f->r, (long long)f->N);
  if (f->C == 1) {
    if (f->r < f->ahead_lo || f->r + avail > f->ahead_hi) {       // not covered by the read-ahead: refill it
      const long long want = std::min<long long>(f->w - f->r, (long long)f->N + crlot_framer_s::kAheadHops * (long long)f->H);
      if (!f->ahead) {
        const long long cap = (long long)f->N + crlot_framer_s::kAheadHops * (long long)f->H;
        FCU(cudaHostAlloc((void**)&f->ahead, (size_t)cap * sizeof(float), cudaHostAllocDefault));
        f->ahead_cap = cap;
      }
      FCU(cudaMemcpyAsync(f->ahead, f->buf + f->r, (size_t)want * sizeof(float), cudaMemcpyDeviceToHost, st));
      FCU(cudaStreamSynchronize(st));
      f->ahead_lo = f->r; f->ahead_hi = f->r + want;
    }
    std::copy(f->ahead + (f->r - f->ahead_lo), f->ahead + (f->r - f->ahead_lo) + avail, h_out_frame);
    std::fill(h_out_frame + avail, h_out_frame + f->N, 0.0f);    // ZERO_PAD tail, framer.cc:157-159
  } else {
    int rc = ensure_stage(f, f->N * f->C);
    if (rc) return rc;
    const long long total = (long long)f->N * (long long)f->C;
    interleave_kernel<<<(int)((total + 255) / 256), 256, 0, st>>>(f->buf, f->stage, f->cap, f->r, avail,
                                                                   (long long)f->N, (int)f->C);
    FCU(cudaGetLastError());
    FCU(cudaMemcpyAsync(h_out_frame, f->stage, (size_t)total * sizeof(float), cudaMemcpyDeviceToHost, st));
    FCU(cudaStreamSynchronize(st));
  }
  f->r = std::min(f->r + (long long)f->H, f->w);            // framer.cc:162-165
  ref_pop(f);
  if (popped) *popped = 1;
  return CRLOT_OK;
}

int crlot_framer_stft(crlot_framer_t f, crlot_plan_t plan, long long max_frames, void* d_spec, long long* n_frames,
                      void* stream) {
  if (n_frames) *n_frames = 0;
  if (!f || !plan || !n_frames) return ffail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  crlot_plan_info info;
  int rc = crlot_stft_plan_info(plan, &info);
  if (rc) return rc;
  if ((size_t)info.nfft != f->N || (size_t)info.hop != f->H || info.device != f->device)
    return ffail(CRLOT_ERR_INVALID_ARGUMENT, "plan and framer disagree on frame size, hop or device");
  long long k = drainable_now(f);
  if (max_frames >= 0) k = std::min(k, max_frames);
  if (k == 0) return CRLOT_OK;
  if (!d_spec) return ffail(CRLOT_ERR_INVALID_ARGUMENT, "null spectra buffer");
  Dev g(f->device);
  rc = order_after_last(f, static_cast<cudaStream_t>(stream));
  if (rc) return rc;
  if (f->r & 3) {                                           // after a clamped ZERO_PAD tail: re-base once
    rc = realign(f, static_cast<cudaStream_t>(stream));
    if (rc) return rc;
  }
  // frame j of channel c = buf[c*cap + r + j*H + (0..N)), zero at and beyond (w - r): kernel 1's contract
  rc = crlot_stft_execute(plan, f->buf + f->r, f->cap, f->w - f->r, (int)f->C, k, d_spec, stream);
  if (rc) return rc;
  rc = mark_done(f, static_cast<cudaStream_t>(stream));
  if (rc) return rc;
  for (long long j = 0; j < k; ++j) {                       // cursor arithmetic of k pops, O(k) integers
    f->r = std::min(f->r + (long long)f->H, f->w);
    ref_pop(f);
  }
  *n_frames = k;
  return CRLOT_OK;
}

size_t crlot_framer_get(crlot_framer_t f, int what) {
  if (!f) return 0;
  switch (what) {
    case 0: return f->N;
    case 1: return f->H;
    case 2: return f->C;
    case 3: return (size_t)f->mode;
    case 4: return f->ref_size;                  // Framer::buffer_size()
    case 5: return (size_t)f->r;                 // per-channel samples
    case 6: return (size_t)f->w;
    case 7: return (size_t)f->cap;               // device capacity per channel
    case 8: return (size_t)drainable_now(f);     // frames that repeated pop() would deliver
    default: return 0;
  }
}

const float* crlot_framer_device_ptr(crlot_framer_t f, long long* chan_stride, long long* valid) {
  if (!f) return nullptr;
  if (chan_stride) *chan_stride = f->cap;
  if (valid) *valid = f->w - f->r;
  return f->buf + f->r;
}

}  // extern "C"
