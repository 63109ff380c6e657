// ring.cu -- (1) dsp::ring::RingBuffer<T> behind the C-ABI (crlot_ring_*): unified-memory cells with the
// reference's index arithmetic (ref: dsp/ring/ring_buffer.cc:15-205), and (2) the kernel-dispatch seam
// axpy / axpy_windowed / normalize_and_clear on device pointers (ref: dsp/ola/kernels.h:28-53,
// dsp/ola/kernels.cc:18-36) -- the same roundings as the reference's scalar kernels.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <limits>
#include <new>
#include <string>

#include "../../include/crlot_cuda.h"

extern "C" void crlot_cuda_set_error_(const char* msg);

namespace {

int rfail(int code, const char* msg) {
  crlot_cuda_set_error_(msg);
  return code;
}
int rfail_cuda(cudaError_t e, const char* where) {
  const std::string m = std::string(where) + ": " + cudaGetErrorString(e);
  cudaGetLastError();
  crlot_cuda_set_error_(m.c_str());
  return e == cudaErrorMemoryAllocation ? CRLOT_ERR_ALLOC : CRLOT_ERR_CUDA;
}
#define RCU(call)                                           \
  do {                                                      \
    cudaError_t e__ = (call);                               \
    if (e__ != cudaSuccess) return rfail_cuda(e__, #call);  \
  } while (0)

struct Dev {
  int prev = -1;
  explicit Dev(int d) { cudaGetDevice(&prev); if (prev != d) cudaSetDevice(d); }
  ~Dev() { if (prev >= 0) cudaSetDevice(prev); }
};

// dst[i] = fma(src[i] (* win[i], rounded), g, dst[i])     ref: kernels.cc:18-28
template <bool WIN>
__global__ void axpy_kernel(float* __restrict__ dst, const float* __restrict__ src, const float* __restrict__ win, float g,
                            size_t n) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    float f = src[i];
    if (WIN) f = __fmul_rn(f, win[i]);
    dst[i] = __fmaf_rn(f, g, dst[i]);
  }
}
// out[i] = acc[i] / max(norm[i], eps); acc[i] = 0          ref: kernels.cc:30-36
__global__ void normalize_and_clear_kernel(float* __restrict__ out, float* __restrict__ acc, const float* __restrict__ norm,
                                           float eps, size_t n) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    out[i] = __fdiv_rn(acc[i], fmaxf(norm[i], eps));
    acc[i] = 0.0f;
  }
}

int grid_for(size_t n) { return static_cast<int>(std::min<size_t>((n + 255) / 256, 148 * 8)); }

}  // namespace

struct crlot_ring_s {
  unsigned char* buf = nullptr;
  size_t capacity = 0, elem = 0, write_pos = 0;
  bool shadow = false;
  int device = 0;
};

extern "C" {

// ---- kernel-dispatch seam ----------------------------------------------------------------------------
int crlot_axpy(float* d_dst, const float* d_src, float g, size_t n, int device, void* stream) {
  if (n == 0) return CRLOT_OK;
  if (!d_dst || !d_src) return rfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  Dev d(device);
  axpy_kernel<false><<<grid_for(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(d_dst, d_src, nullptr, g, n);
  RCU(cudaGetLastError());
  return CRLOT_OK;
}
int crlot_axpy_windowed(float* d_dst, const float* d_src, const float* d_win, float g, size_t n, int device, void* stream) {
  if (n == 0) return CRLOT_OK;
  if (!d_dst || !d_src || !d_win) return rfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  Dev d(device);
  axpy_kernel<true><<<grid_for(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(d_dst, d_src, d_win, g, n);
  RCU(cudaGetLastError());
  return CRLOT_OK;
}
int crlot_normalize_and_clear(float* d_out, float* d_acc, const float* d_norm, float eps, size_t n, int device, void* stream) {
  if (n == 0) return CRLOT_OK;
  if (!d_out || !d_acc || !d_norm) return rfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  Dev d(device);
  normalize_and_clear_kernel<<<grid_for(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(d_out, d_acc, d_norm, eps, n);
  RCU(cudaGetLastError());
  return CRLOT_OK;
}

// ---- RingBuffer ---------------------------------------------------------------------------------------
int crlot_ring_create(size_t capacity, size_t elem_bytes, int shadow, int device, crlot_ring_t* out) {
  if (!out) return rfail(CRLOT_ERR_INVALID_ARGUMENT, "out is null");
  if (capacity == 0) return rfail(CRLOT_ERR_INVALID_ARGUMENT, "RingBuffer capacity must be > 0");   // ring_buffer.cc:21-23
  if (elem_bytes == 0) return rfail(CRLOT_ERR_INVALID_ARGUMENT, "element size must be > 0");
  auto* r = new (std::nothrow) crlot_ring_s;
  if (!r) return rfail(CRLOT_ERR_ALLOC, "host allocation failed");
  r->capacity = capacity; r->elem = elem_bytes; r->shadow = shadow != 0; r->device = device;
  Dev d(device);
  const size_t bytes = (r->shadow ? 2 : 1) * capacity * elem_bytes;
  cudaError_t e = cudaMallocManaged(reinterpret_cast<void**>(&r->buf), bytes, cudaMemAttachGlobal);
  if (e == cudaSuccess) e = cudaMemset(r->buf, 0, bytes);            // the reference zero-fills the physical buffer (:31-32)
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    if (r->buf) cudaFree(r->buf);
    delete r;
    return rfail_cuda(e, "cudaMallocManaged");
  }
  *out = r;
  return CRLOT_OK;
}
int crlot_ring_destroy(crlot_ring_t r) {
  if (!r) return CRLOT_OK;
  { Dev d(r->device); if (r->buf) cudaFree(r->buf); }
  delete r;
  return CRLOT_OK;
}
void* crlot_ring_data(crlot_ring_t r) { return r ? r->buf : nullptr; }
size_t crlot_ring_capacity(crlot_ring_t r) { return r ? r->capacity : 0; }
size_t crlot_ring_physical_capacity(crlot_ring_t r) { return r ? (r->shadow ? 2 * r->capacity : r->capacity) : 0; }
size_t crlot_ring_write_pos(crlot_ring_t r) { return r ? r->write_pos : 0; }
int crlot_ring_has_shadow(crlot_ring_t r) { return r && r->shadow ? 1 : 0; }

int crlot_ring_split(crlot_ring_t r, size_t start, size_t len, crlot_ring_span* out) {
  if (!r || !out) return rfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  out->off1 = out->len1 = out->off2 = out->len2 = 0;
  if (len == 0) return CRLOT_OK;                                                          // :46-48
  if (len > std::numeric_limits<size_t>::max() - start) len = std::numeric_limits<size_t>::max() - start;   // :51-53
  if (len > r->capacity) len = r->capacity;                                               // :56-58
  start %= r->capacity;                                                                   // :61
  if (start + len <= r->capacity) {                                                       // :73-75
    out->off1 = start; out->len1 = len;
    return CRLOT_OK;
  }
  out->off1 = start; out->len1 = r->capacity - start;                                     // :78-84
  out->off2 = 0; out->len2 = len - out->len1;
  return CRLOT_OK;
}

int crlot_ring_shadow_sync(crlot_ring_t r, size_t n) {
  if (!r) return rfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (!r->shadow || n == 0) return CRLOT_OK;                                              // :133-135
  Dev d(r->device);
  const size_t to_copy = std::min(n, r->capacity);
  RCU(cudaMemcpy(r->buf + r->capacity * r->elem, r->buf, to_copy * r->elem, cudaMemcpyDefault));
  return CRLOT_OK;
}

size_t crlot_ring_contiguous_offset(crlot_ring_t r, size_t read_pos) {
  if (!r) return 0;
  return r->shadow ? read_pos : read_pos % r->capacity;                                   // :144-152
}

static int ring_write_impl(crlot_ring_t r, const void* src, size_t n, size_t* written, cudaStream_t st, bool sync) {
  if (written) *written = 0;
  if (!r) return rfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  if (!src || n == 0) return CRLOT_OK;                                                    // :167-169
  if (n > r->capacity) n = r->capacity;               // the reference would overrun its buffer here
  Dev d(r->device);
  const unsigned char* s = static_cast<const unsigned char*>(src);
  const size_t tail = r->capacity - r->write_pos;
  if (n <= tail) {                                                                        // :174-181
    RCU(cudaMemcpyAsync(r->buf + r->write_pos * r->elem, s, n * r->elem, cudaMemcpyDefault, st));
    r->write_pos += n;
    if (r->write_pos == r->capacity) r->write_pos = 0;
  } else {                                                                                // :182-201
    RCU(cudaMemcpyAsync(r->buf + r->write_pos * r->elem, s, tail * r->elem, cudaMemcpyDefault, st));
    const size_t head = n - tail;
    RCU(cudaMemcpyAsync(r->buf, s + tail * r->elem, head * r->elem, cudaMemcpyDefault, st));
    r->write_pos = head;
    if (r->shadow)
      RCU(cudaMemcpyAsync(r->buf + r->capacity * r->elem, r->buf, std::min(head, r->capacity) * r->elem, cudaMemcpyDefault, st));
  }
  if (sync) RCU(cudaStreamSynchronize(st));
  if (written) *written = n;
  return CRLOT_OK;
}
int crlot_ring_write(crlot_ring_t r, const void* h_src, size_t n, size_t* written) {
  return ring_write_impl(r, h_src, n, written, nullptr, true);
}
int crlot_ring_write_device(crlot_ring_t r, const void* d_src, size_t n, size_t* written, void* stream) {
  return ring_write_impl(r, d_src, n, written, static_cast<cudaStream_t>(stream), false);
}

int crlot_ring_prefetch(crlot_ring_t r, int to_device, void* stream) {
  if (!r) return rfail(CRLOT_ERR_INVALID_ARGUMENT, "null pointer");
  Dev d(r->device);
  const size_t bytes = (r->shadow ? 2 : 1) * r->capacity * r->elem;
#if CUDART_VERSION >= 13000
  cudaMemLocation loc{};
  loc.type = to_device ? cudaMemLocationTypeDevice : cudaMemLocationTypeHost;
  loc.id = to_device ? r->device : 0;
  RCU(cudaMemPrefetchAsync(r->buf, bytes, loc, 0, static_cast<cudaStream_t>(stream)));
#else
  RCU(cudaMemPrefetchAsync(r->buf, bytes, to_device ? r->device : cudaCpuDeviceId, static_cast<cudaStream_t>(stream)));
#endif
  return CRLOT_OK;
}

}  // extern "C"
