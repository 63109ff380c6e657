// aux_kernels.cu -- small helpers around the three hot kernels: the counter-based synthetic
// signal generator (same hash as oracle/crlot_oracle.c, so any slice of a 2^33-sample signal
// can be regenerated on either side), and the device-resident ring operations that back the
// OLAAccumulator facade and the streaming (config 5) path.
#include "common.cuh"
#include "kernels.h"

namespace crlot {

namespace {

__device__ __forceinline__ float synth_one(unsigned long long seed, unsigned long long ch, unsigned long long i) {
  unsigned long long z = seed + ch * 0xD1B54A32D192ED03ull + i * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  const unsigned int m = static_cast<unsigned int>(z >> 40);
  return __fadd_rn(__fmul_rn(static_cast<float>(m), 1.0f / 8388608.0f), -1.0f);
}

__global__ void synth_uniform_kernel(float* out, long long chan_stride, long long n, unsigned long long seed,
                                     unsigned long long channel0, unsigned long long i0) {
  const int c = blockIdx.y;
  float* dst = out + static_cast<long long>(c) * chan_stride;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = synth_one(seed, channel0 + c, i0 + static_cast<unsigned long long>(i));
}

// acc[(start+j) % ring_len] = fma( fl(frame[j]*w[j]) , gain, acc[...] )   (axpy_windowed), or
// fma(frame[j], gain, acc) without a window (axpy).  One thread per (channel, j): no two threads
// touch the same ring cell within a launch (n <= ring_len), so no atomics.
__global__ void ring_axpy_kernel(float* ring, long long ring_len, const float* frames, long long frame_chan_stride,
                                 const float* window, float gain, long long start, int n) {
  const int c = blockIdx.y;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  float* acc = ring + static_cast<long long>(c) * ring_len + (start + j) % ring_len;
  float f = frames[static_cast<long long>(c) * frame_chan_stride + j];
  if (window) f = __fmul_rn(f, window[j]);
  *acc = __fmaf_rn(f, gain, *acc);
}

// out[j] = acc[p] / max(norm[p], eps); acc[p] = 0; p = (read_pos + j) % ring_len   (normalize_and_clear)
__global__ void ring_produce_kernel(float* ring, long long ring_len, const float* norm, float eps, long long read_pos,
                                    int n, float* out, long long out_chan_stride) {
  const int c = blockIdx.y;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const long long pidx = (read_pos + j) % ring_len;
  float* acc = ring + static_cast<long long>(c) * ring_len + pidx;
  out[static_cast<long long>(c) * out_chan_stride + j] = __fdiv_rn(*acc, fmaxf(norm[pidx], eps));
  *acc = 0.0f;
}

__global__ void build_den_rcp_kernel(const float* __restrict__ norm, float eps, long long n, float2* __restrict__ out) {
  const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i < n) {
    const float b = fmaxf(norm[i], eps);
    out[i] = make_float2(b, div_refined_rcp(b));
  }
}

__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// operand pairs: numerator = random sign/mantissa with exponent in [-60, 60] (or exactly 0 once in
// 1024), denominator alternately a table entry and a random value with exponent in [-40, 40]
__global__ void division_selftest_kernel(const float2* __restrict__ den_rcp, long long ring_len, unsigned long long n,
                                         unsigned long long seed, unsigned long long* mismatches) {
  unsigned long long bad = 0;
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * blockDim.x;
  for (unsigned long long i = blockIdx.x * static_cast<unsigned long long>(blockDim.x) + threadIdx.x; i < n; i += stride) {
    const unsigned long long h = mix64(seed ^ (i * 0xD6E8FEB86659FD93ull));
    const unsigned long long h2 = mix64(h);
    const unsigned int ea = 127u - 60u + static_cast<unsigned int>((h >> 32) % 120u);
    float a = __uint_as_float((static_cast<unsigned int>(h) & 0x807fffffu) | (ea << 23));
    if (((h >> 40) & 1023u) == 0u) a = 0.0f;
    float b, r;
    if (ring_len > 0 && (h2 & 1ull)) {
      const float2 t = den_rcp[(h2 >> 8) % static_cast<unsigned long long>(ring_len)];
      b = t.x; r = t.y;
    } else {
      const unsigned int eb = 127u - 40u + static_cast<unsigned int>((h2 >> 32) % 81u);
      b = __uint_as_float((static_cast<unsigned int>(h2 >> 1) & 0x007fffffu) | (eb << 23));
      r = div_refined_rcp(b);
    }
    if (!div_tail_ok(a)) { ++bad; continue; }
    const float want = __fdiv_rn(a, b), got = div_tail(a, b, r);
    if (__float_as_uint(want) != __float_as_uint(got)) ++bad;
  }
  if (bad) atomicAdd(mismatches, bad);
}

// Interleaved (AoS, [n][c]) <-> planar ([c][n]) copies (reference: dsp/ola/aos_to_soa.cc:13-17 and the
// Framer's interleaved push, dsp/frame/framer.cc:133-141).  One thread per time index; the channel loop
// is inside the thread so that every 32-byte sector of the interleaved side is consumed by one warp.
template <int CH>
__global__ void deinterleave_kernel(const float* aos, long long n, int channels, float* planar, long long chan_stride) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (CH == 2) {
      const float2 v = *reinterpret_cast<const float2*>(aos + 2 * i);
      planar[i] = v.x;
      planar[chan_stride + i] = v.y;
    } else if (CH == 4) {
      const float4 v = *reinterpret_cast<const float4*>(aos + 4 * i);
      planar[i] = v.x;
      planar[chan_stride + i] = v.y;
      planar[2 * chan_stride + i] = v.z;
      planar[3 * chan_stride + i] = v.w;
    } else {
      for (int c = 0; c < channels; ++c) planar[c * chan_stride + i] = aos[i * channels + c];
    }
  }
}
template <int CH>
__global__ void interleave_kernel(const float* planar, long long chan_stride, long long n, int channels, float* aos) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (CH == 2) {
      *reinterpret_cast<float2*>(aos + 2 * i) = make_float2(planar[i], planar[chan_stride + i]);
    } else if (CH == 4) {
      *reinterpret_cast<float4*>(aos + 4 * i) =
          make_float4(planar[i], planar[chan_stride + i], planar[2 * chan_stride + i], planar[3 * chan_stride + i]);
    } else {
      for (int c = 0; c < channels; ++c) aos[i * channels + c] = planar[c * chan_stride + i];
    }
  }
}

// FrameQueue::getAllFrames (reference: dsp/frame/FrameQueue.cc:28-46): pure index math, bit-exact
__global__ void framequeue_extract_kernel(const float* sig, long long len, int frame, int hop, int pad, int pad_mode,
                                          long long n_frames, float* frames) {
  const long long total = n_frames * frame;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long k = i / frame;
    const int u = static_cast<int>(i - k * frame);
    frames[i] = pad_value(sig, len, k * hop + u - pad, pad_mode);
  }
}

// spectral-domain hook as its own pass: one thread per bin, rows strided over the grid's y
__global__ void spectral_gain_kernel(cpx* spec, long long rows, int bins, const float* gain, long long row_stride) {
  const long long total = rows * bins;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long r = i / bins;
    const int k = static_cast<int>(i - r * bins);
    const float g = __ldg(gain + r * row_stride + k);
    const cpx x = spec[i];
    spec[i] = mk(__fmul_rn(x.x, g), __fmul_rn(x.y, g));
  }
}

}  // namespace

cudaError_t launch_build_den_rcp(const float* norm, float eps, long long n, float2* den_rcp, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  build_den_rcp_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, st>>>(norm, eps, n, den_rcp);
  return cudaGetLastError();
}

cudaError_t launch_division_selftest(const float2* den_rcp, long long ring_len, unsigned long long n,
                                     unsigned long long seed, unsigned long long* mismatches, cudaStream_t st) {
  division_selftest_kernel<<<device_sm_count() * 8, 256, 0, st>>>(den_rcp, ring_len, n, seed, mismatches);
  return cudaGetLastError();
}

cudaError_t launch_synth_uniform(float* out, long long chan_stride, int channels, long long n, unsigned long long seed,
                                 unsigned long long channel0, unsigned long long i0, cudaStream_t st) {
  if (n <= 0 || channels <= 0) return cudaSuccess;
  long long blocks = (n + 255) / 256;
  const long long cap = static_cast<long long>(device_sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  synth_uniform_kernel<<<dim3(static_cast<unsigned>(blocks), channels), 256, 0, st>>>(out, chan_stride, n, seed,
                                                                                      channel0, i0);
  return cudaGetLastError();
}

cudaError_t launch_ring_axpy(float* ring, long long ring_len, int channels, const float* frames,
                             long long frame_chan_stride, const float* window, float gain, long long start, int n,
                             cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  ring_axpy_kernel<<<dim3((n + 255) / 256, channels), 256, 0, st>>>(ring, ring_len, frames, frame_chan_stride, window,
                                                                   gain, start, n);
  return cudaGetLastError();
}

cudaError_t launch_ring_produce(float* ring, long long ring_len, int channels, const float* norm, float eps,
                                long long read_pos, int n, float* out, long long out_chan_stride, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  ring_produce_kernel<<<dim3((n + 255) / 256, channels), 256, 0, st>>>(ring, ring_len, norm, eps, read_pos, n, out,
                                                                      out_chan_stride);
  return cudaGetLastError();
}

int device_sm_count() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached = n;
    cached_dev = dev;
  }
  return cached;
}

cudaError_t launch_spectral_gain(cpx* spec, long long rows, int bins, const float* gain, long long row_stride,
                                 cudaStream_t st) {
  if (rows <= 0) return cudaSuccess;
  const long long total = rows * bins;
  const long long want = (total + 255) / 256;
  const int grid = static_cast<int>(want < 8ll * device_sm_count() ? want : 8ll * device_sm_count());
  spectral_gain_kernel<<<grid, 256, 0, st>>>(spec, rows, bins, gain, row_stride);
  return cudaGetLastError();
}

cudaError_t launch_framequeue_extract(const float* sig, long long len, int frame, int hop, int pad, int pad_mode,
                                      long long n_frames, float* frames, cudaStream_t st) {
  if (n_frames <= 0) return cudaSuccess;
  const long long want = (n_frames * frame + 255) / 256;
  const int grid = static_cast<int>(want < 8ll * device_sm_count() ? want : 8ll * device_sm_count());
  framequeue_extract_kernel<<<grid, 256, 0, st>>>(sig, len, frame, hop, pad, pad_mode, n_frames, frames);
  return cudaGetLastError();
}

cudaError_t launch_deinterleave(const float* aos, long long n, int channels, float* planar, long long chan_stride,
                                cudaStream_t st) {
  if (n <= 0 || channels <= 0) return cudaSuccess;
  const long long want = (n + 255) / 256;
  const int grid = static_cast<int>(want < 16ll * device_sm_count() ? want : 16ll * device_sm_count());
  const bool a8 = reinterpret_cast<uintptr_t>(aos) % 8 == 0, a16 = reinterpret_cast<uintptr_t>(aos) % 16 == 0;
  if (channels == 2 && a8) deinterleave_kernel<2><<<grid, 256, 0, st>>>(aos, n, channels, planar, chan_stride);
  else if (channels == 4 && a16) deinterleave_kernel<4><<<grid, 256, 0, st>>>(aos, n, channels, planar, chan_stride);
  else deinterleave_kernel<0><<<grid, 256, 0, st>>>(aos, n, channels, planar, chan_stride);
  return cudaGetLastError();
}
cudaError_t launch_interleave(const float* planar, long long chan_stride, long long n, int channels, float* aos,
                              cudaStream_t st) {
  if (n <= 0 || channels <= 0) return cudaSuccess;
  const long long want = (n + 255) / 256;
  const int grid = static_cast<int>(want < 16ll * device_sm_count() ? want : 16ll * device_sm_count());
  const bool a8 = reinterpret_cast<uintptr_t>(aos) % 8 == 0, a16 = reinterpret_cast<uintptr_t>(aos) % 16 == 0;
  if (channels == 2 && a8) interleave_kernel<2><<<grid, 256, 0, st>>>(planar, chan_stride, n, channels, aos);
  else if (channels == 4 && a16) interleave_kernel<4><<<grid, 256, 0, st>>>(planar, chan_stride, n, channels, aos);
  else interleave_kernel<0><<<grid, 256, 0, st>>>(planar, chan_stride, n, channels, aos);
  return cudaGetLastError();
}

}  // namespace crlot
