// k3_ola.cu -- kernel 3: gather-formulated overlap-add + window-sum normalisation (sm_100a).
//
// Replaces OLAAccumulator::add_frame_SoA's accumulate (reference: dsp/ola/OLAAccumulator.cc:
// 86-110 -> axpy_windowed / axpy, dsp/ola/kernels.cc:18-28) followed by produce()'s
// normalize_and_clear (OLAAccumulator.cc:194-209, kernels.cc:30-36), in the streaming drive
// order "push frame k at k*hop, then produce(hop)".  Each output sample sums its contributing
// frames in ascending frame order starting from a cleared accumulator -- the same order and
// the same roundings as the reference's ring (fl(y*w), then fma(.,gain,acc), then a true
// division by max(norm[t mod ring_len], eps)) -- so the result is BIT-IDENTICAL to the
// reference given identical frames, with no atomics and run-to-run deterministic.
//
// Pure streaming: every frame element is read exactly once (16-byte vector loads, L1
// no-allocate), every output element written once.
#include "common.cuh"
#include "kernels.h"

namespace crlot {

namespace {

constexpr int kK3Threads = 256;

template <bool WIN, int VEC>
__global__ void __launch_bounds__(kK3Threads) ola_gather_kernel(const K3Params p) {
  const int c = blockIdx.y;
  const float* __restrict__ frames = p.frames + static_cast<long long>(c) * p.frames_chan_rows * p.nfft;
  float* __restrict__ out = p.out + static_cast<long long>(c) * p.out_chan_stride;
  const long long N = p.nfft, H = p.hop, F = p.frames_per_chan;
  const long long ngroups = (p.n_out + VEC - 1) / VEC;
  const long long tiles = (ngroups + kK3Threads - 1) / kK3Threads;
  const unsigned int Hu = static_cast<unsigned int>(p.hop), Nu = static_cast<unsigned int>(p.nfft);
  // A block walks tiles of 256 groups.  The 64-bit divisions (frame index of the tile's first sample,
  // envelope phase) are done once per tile on block-uniform values; everything per thread is 32-bit
  // arithmetic relative to the tile (the kernel is a pure stream: the index arithmetic is its only
  // instruction cost).
  for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const long long j0 = tile * kK3Threads * VEC;          // first output sample of the tile
    const long long u0 = j0 - p.frame_t0;                  // its offset from the start of local frame 0
    // frame that contains u0 (floor division, also for negative u0) and the offset inside it
    long long kq = u0 >= 0 ? u0 / H : -((-u0 + H - 1) / H);
    const unsigned int rem0 = static_cast<unsigned int>(u0 - kq * H);            // [0, H)
    long long eb = (p.t_offset + j0) % p.ring_len;
    const unsigned int dj = threadIdx.x * VEC;             // sample offset inside the tile
    const long long j = j0 + dj;
    if (j >= p.n_out) continue;
    const unsigned int ru = rem0 + dj;                     // u = kq*H + ru
    const long long k_top = kq + ru / Hu;                  // frame that contains u: floor(u / H)
    const unsigned int i_top = ru % Hu;                    // u - k_top*H, in [0, H)
    float acc[VEC];
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[e] = 0.0f;
    // contributing frames k = k_top - m, m = 0 .. (N - 1 - i_top)/H (VEC == 4: N, H, i_top multiples
    // of 4, so the whole group shares them), clipped to [0, F); summed in ASCENDING k
    unsigned int m_max = (Nu - (VEC == 4 ? 4u : 1u) - i_top) / Hu;
    long long k_lo = k_top - m_max, k_hi = k_top;
    if (k_lo < 0) k_lo = 0;
    if (k_hi > F - 1) k_hi = F - 1;
    if (k_hi >= k_lo) {
      const float* fp = frames + k_lo * N + (static_cast<long long>(k_top - k_lo) * H + i_top);
      unsigned int i = static_cast<unsigned int>(k_top - k_lo) * Hu + i_top;      // offset inside frame k_lo
      const int cnt = static_cast<int>(k_hi - k_lo + 1);
      const long long step = N - H;                        // next frame, same output sample
#pragma unroll 4
      for (int q = 0; q < cnt; ++q) {
        if (VEC == 4) {
          const float4 y = ldg_stream4(reinterpret_cast<const float4*>(fp));
          float4 f = y;
          if (WIN) {
            const float4 w = __ldg(reinterpret_cast<const float4*>(p.window + i));
            f.x = __fmul_rn(y.x, w.x); f.y = __fmul_rn(y.y, w.y);
            f.z = __fmul_rn(y.z, w.z); f.w = __fmul_rn(y.w, w.w);
          }
          acc[0] = __fmaf_rn(f.x, p.gain, acc[0]);
          acc[1] = __fmaf_rn(f.y, p.gain, acc[1]);
          acc[2] = __fmaf_rn(f.z, p.gain, acc[2]);
          acc[3] = __fmaf_rn(f.w, p.gain, acc[3]);
        } else {
          float f = __ldg(fp);
          if (WIN) f = __fmul_rn(f, __ldg(p.window + i));
          acc[0] = __fmaf_rn(f, p.gain, acc[0]);
        }
        fp += step;
        i -= Hu;
      }
    }
    eb += dj;
    while (eb >= p.ring_len) eb -= p.ring_len;
    if (VEC == 4) {
      const float4 nm = __ldg(reinterpret_cast<const float4*>(p.norm + eb));
      float4 o;
      o.x = __fdiv_rn(acc[0], fmaxf(nm.x, p.eps));
      o.y = __fdiv_rn(acc[1], fmaxf(nm.y, p.eps));
      o.z = __fdiv_rn(acc[2], fmaxf(nm.z, p.eps));
      o.w = __fdiv_rn(acc[3], fmaxf(nm.w, p.eps));
      if (j + 3 < p.n_out) {
        stg_stream4(reinterpret_cast<float4*>(out + j), o);
      } else {
        const float t[4] = {o.x, o.y, o.z, o.w};
        for (int e = 0; e < 4 && j + e < p.n_out; ++e) out[j + e] = t[e];
      }
    } else {
      out[j] = __fdiv_rn(acc[0], fmaxf(__ldg(p.norm + eb), p.eps));
    }
  }
}

}  // namespace

cudaError_t launch_ola_gather(const K3Params& p, cudaStream_t st) {
  if (p.n_out <= 0 || p.channels <= 0) return cudaSuccess;
  const bool win = p.window != nullptr;
  // 16-byte path: every quantity that shifts a 4-sample group must be a multiple of 4.
  const bool vec = (p.nfft % 4 == 0) && (p.hop % 4 == 0) && (p.frame_t0 % 4 == 0) && (p.t_offset % 4 == 0) &&
                   (p.ring_len % 4 == 0) && (p.out_chan_stride % 4 == 0) &&
                   (reinterpret_cast<uintptr_t>(p.frames) % 16 == 0) && (reinterpret_cast<uintptr_t>(p.out) % 16 == 0) &&
                   (reinterpret_cast<uintptr_t>(p.norm) % 16 == 0) &&
                   (!win || reinterpret_cast<uintptr_t>(p.window) % 16 == 0);
  const long long groups = vec ? (p.n_out + 3) / 4 : p.n_out;
  if (p.t_offset < 0) return cudaErrorInvalidValue;
  long long blocks = (groups + kK3Threads - 1) / kK3Threads;
  const long long cap = static_cast<long long>(device_sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  const dim3 grid(static_cast<unsigned>(blocks), static_cast<unsigned>(p.channels));
  if (vec) {
    if (win) ola_gather_kernel<true, 4><<<grid, kK3Threads, 0, st>>>(p);
    else     ola_gather_kernel<false, 4><<<grid, kK3Threads, 0, st>>>(p);
  } else {
    if (win) ola_gather_kernel<true, 1><<<grid, kK3Threads, 0, st>>>(p);
    else     ola_gather_kernel<false, 1><<<grid, kK3Threads, 0, st>>>(p);
  }
  return cudaGetLastError();
}

}  // namespace crlot
