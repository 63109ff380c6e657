// k23_istft_ola.cu -- kernels 2+3 fused: inverse real FFT * 1/N + clean-up + synthesis window,
// overlap-added in a per-warp shared-memory accumulator ring and normalised on the way out.
//
// Same arithmetic, in the same order, as kernel 2 followed by kernel 3 (k2_istft.cu, k3_ola.cu):
//   f   = fl( scrub(fl(y * 2/N)) * w[i] )                     (k2, the src*win half of axpy_windowed)
//   acc = fma(f, gain, acc)   frames ascending, from a cleared accumulator   (k3 / axpy)
//   out = acc / max(norm[(t_offset + j) mod ring_len], eps)                  (k3 / normalize_and_clear)
// so the output is BIT-IDENTICAL to the three-kernel path and to the reference ring given identical
// spectra (reference: dsp/fft/backends/kissfft_adapter.cc:124-168, dsp/ola/OLAAccumulator.cc:
// 86-110,194-209, dsp/ola/kernels.cc:18-36) -- but the iFFT frames never travel through HBM:
// per frame the kernel reads one spectrum row (8*bins bytes) and writes one hop (4*hop bytes)
// instead of 8*bins + 4N written/read twice.
//
// Structure (the mirror image of kernel 1): a warp owns a run of consecutive output hop blocks
// of one channel (a "chunk", taken from an atomic counter).  Output block m = samples
// u in [m*hop, (m+1)*hop) relative to the start of local frame 0; it is complete as soon as frame
// m has been accumulated, because frame m+1 starts at (m+1)*hop.  The warp first re-accumulates
// the ceil(N/hop)-1 frames preceding its chunk (contributions at or beyond the chunk start
// only), then for every frame k of the chunk:
//   samples i <  hop      : acc from the ring (+ this frame) -> divided, written to global
//   samples i >= N - hop  : first touch of their ring cells  -> plain store, no read
//   samples in between    : read-modify-write in the ring
// No atomics; every output sample sums its frames in ascending order: deterministic.
#include "common.cuh"
#include "kernels.h"
#include "warpfft.cuh"

namespace crlot {

namespace {

constexpr int kK23Warps = 12;

__host__ __device__ inline int k23_next_pow2(int v) {
  int r = 1;
  while (r < v) r <<= 1;
  return r;
}

// RDIV = nfft / hop when that is 1, 2, 4 or 8 (hop known at compile time: the three register
// classes above are compile-time ranges and the ring loads of a frame issue back to back), else 0
// (any eligible hop; per-register tests at run time).  The synthesis window is always multiplied
// in; "no window" is served by a table of ones (y * 1.0f == y exactly).
template <int NFFT, int RDIV, bool SCRUB>
__global__ void __launch_bounds__(kK23Warps * 32, 1) istft_ola_kernel(const K23Params p, const int acap) {
  using X = GeoX<NFFT>;
  constexpr int C = X::C, G = X::G, A = X::A, B = X::B;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* s_win = reinterpret_cast<float*>(smem_raw);
  cpx* s_tw1 = reinterpret_cast<cpx*>(s_win + NFFT);
  cpx* s_sa = s_tw1 + C;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* wb = reinterpret_cast<unsigned char*>(s_sa + C) + warp * (acap * 4 + X::XCPX * 8);
  float* ring = reinterpret_cast<float*>(wb);
  cpx* ex = reinterpret_cast<cpx*>(ring + acap);
  const int mask = acap - 1;

  for (int i = threadIdx.x; i < NFFT; i += blockDim.x) s_win[i] = p.window[i];
  for (int i = threadIdx.x; i < C; i += blockDim.x) {
    s_tw1[i] = p.tw1[i];
    s_sa[i] = p.sa[i];
  }
  __syncthreads();

  const int H = RDIV > 0 ? NFFT / (RDIV > 0 ? RDIV : 1) : p.hop;
  const int R = (NFFT + H - 1) / H;
  const long long F = p.frames_per_chan;
  const long long nblocks = p.m_end - p.m_begin;
  const long long chunks_per_chan = (nblocks + p.chunk_blocks - 1) / p.chunk_blocks;
  const long long total_chunks = chunks_per_chan * p.channels;
  const long long u_cov = F > 0 ? (F - 1) * H + NFFT : 0;          // u >= u_cov: no frame contributes
  const int lp = mirror_lane<A>(lane);
  const float scale = 2.0f / static_cast<float>(NFFT);             // inv_split hands over halved values
  const int ring_len = static_cast<int>(p.ring_len);
  const float gain = p.gain, eps = p.eps;

  for (;;) {
    unsigned long long q = 0;
    if (lane == 0) q = atomicAdd(p.counter, 1u);
    q = __shfl_sync(0xffffffffu, q, 0);
    if (static_cast<long long>(q) >= total_chunks) break;
    const long long c = q / chunks_per_chan;
    const long long m0 = p.m_begin + (q % chunks_per_chan) * p.chunk_blocks;
    const long long m1 = min(m0 + p.chunk_blocks, p.m_end);
    const cpx* spec = p.spec + c * p.spec_chan_rows * (C + 1);
    float* out = p.out + c * p.out_chan_stride;
    const long long k0 = max(0ll, m0 - R + 1);
    const long long k1 = min(m1, F);

    __syncwarp();
    for (int i = lane * 4; i < acap; i += 128) *reinterpret_cast<float4*>(ring + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();

    // envelope phase of sample 0 of frame k0, advanced by hop per frame (no 64-bit modulo per frame)
    int e_frame = 0;
    {
      long long e = (p.t_offset + p.frame_t0 + k0 * H) % ring_len;
      if (e < 0) e += ring_len;
      e_frame = static_cast<int>(e);
    }
    const int e_step = H % ring_len;

    for (long long kbase = k0; kbase < k1; kbase += G) {
      const long long left = k1 - kbase;
      cpx v[32];
      cpx nyq[X::M1];
#pragma unroll
      for (int j = 0; j < X::M1; ++j) {
        const int i1 = lane + 32 * j, g = i1 / A, a = i1 % A;
        const bool ok = g < left;
        const cpx* row = spec + (kbase + g) * (C + 1);
#pragma unroll
        for (int b = 0; b < B; ++b) v[j * B + b] = ok ? ldg_stream(row + a + A * b) : mk(0.f, 0.f);
        nyq[j] = (ok && a == 0) ? ldg_stream(row + C) : mk(0.f, 0.f);
      }
      inv_split<X>(v, lane, s_sa,
                   [&](int r) {
                     return mk(__shfl_sync(0xffffffffu, v[r].x, lp), __shfl_sync(0xffffffffu, v[r].y, lp));
                   },
                   nyq);
      pass1<X, true>(v);
      twiddle1<X, true>(v, lane, s_tw1);
      xchg_write<X>(v, lane, ex);
      __syncwarp();
      xchg_read<X>(v, lane, ex);
      __syncwarp();
      pass2<X, true>(v);

      // f = fl( scrub(fl(y * 2/N)) * w ), in place, every frame of the step at once
#pragma unroll
      for (int j = 0; j < X::M2; ++j) {
        const int ob = (lane + 32 * j) % B;
#pragma unroll
        for (int oa = 0; oa < A; ++oa) {
          float y0 = v[j * A + oa].x * scale, y1 = v[j * A + oa].y * scale;
          if (SCRUB) {
            y0 = scrub(y0);
            y1 = scrub(y1);
          }
          const float2 w = *reinterpret_cast<const float2*>(s_win + 2 * (ob + B * oa));
          v[j * A + oa] = mk(__fmul_rn(y0, w.x), __fmul_rn(y1, w.y));
        }
      }

      // overlap-add, one frame at a time so that every ring cell sees its frames in ascending order
#pragma unroll
      for (int g = 0; g < G; ++g) {
        if (g < left) {
          const long long k = kbase + g;
          const long long u0 = k * H;
          const int r0 = static_cast<int>(u0 & mask);
          const long long j0 = p.frame_t0 + u0;                           // output index of sample 0
          const int e0 = e_frame;
          e_frame += e_step;
          if (e_frame >= ring_len) e_frame -= ring_len;
          const bool fast = RDIV > 0 && k >= m0 && j0 >= 0 && j0 + H <= p.n_out;
          if (fast) {
            constexpr int AE = RDIV > 0 ? A / RDIV : A;     // registers per hop
#pragma unroll
            for (int j = 0; j < X::M2; ++j) {
              const int i2 = lane + 32 * j, ob = i2 % B;
              if (i2 / B == g) {
                const int rb = r0 + 2 * ob;
#pragma unroll
                for (int oa = 0; oa < A; ++oa) {            // all ring reads of the frame, back to back
                  float2 a = make_float2(0.f, 0.f);
                  if (oa < A - AE) a = *reinterpret_cast<const float2*>(ring + ((rb + 2 * B * oa) & mask));
                  v[j * A + oa] = mk(__fmaf_rn(v[j * A + oa].x, gain, a.x), __fmaf_rn(v[j * A + oa].y, gain, a.y));
                }
#pragma unroll
                for (int oa = AE; oa < A; ++oa)
                  *reinterpret_cast<float2*>(ring + ((rb + 2 * B * oa) & mask)) = make_float2(v[j * A + oa].x, v[j * A + oa].y);
                float2 nm[AE];
#pragma unroll
                for (int oa = 0; oa < AE; ++oa) {
                  int e = e0 + 2 * (ob + B * oa);
                  if (e >= ring_len) e -= ring_len;
                  nm[oa] = __ldg(reinterpret_cast<const float2*>(p.norm + e));
                }
                float* dst = out + j0 + 2 * ob;
#pragma unroll
                for (int oa = 0; oa < AE; ++oa)
                  stg_stream(reinterpret_cast<cpx*>(dst + 2 * B * oa),
                             mk(__fdiv_rn(v[j * A + oa].x, fmaxf(nm[oa].x, eps)), __fdiv_rn(v[j * A + oa].y, fmaxf(nm[oa].y, eps))));
              }
            }
          } else {
            const int thr = k < m0 ? static_cast<int>((m0 - k) * H) : 0;   // warm-up: keep u >= m0*H only
            const bool emit = k >= m0;
#pragma unroll
            for (int j = 0; j < X::M2; ++j) {
              const int i2 = lane + 32 * j, ob = i2 % B;
              if (i2 / B == g) {
#pragma unroll
                for (int oa = 0; oa < A; ++oa) {
                  const int ib = 2 * B * oa;                 // warp-uniform part of the sample index
                  if (ib < thr) continue;
                  const int i = ib + 2 * ob;
                  float* cell = ring + ((r0 + i) & mask);
                  float2 a = make_float2(0.f, 0.f);
                  if (ib < NFFT - H) a = *reinterpret_cast<const float2*>(cell);   // else: first touch
                  a.x = __fmaf_rn(v[j * A + oa].x, gain, a.x);
                  a.y = __fmaf_rn(v[j * A + oa].y, gain, a.y);
                  if (emit && ib < H) {
                    const long long jo = j0 + i;
                    if (jo >= 0 && jo < p.n_out) {
                      int e = e0 + i;
                      if (e >= ring_len) e -= ring_len;
                      const float2 nm = __ldg(reinterpret_cast<const float2*>(p.norm + e));
                      const float o0 = __fdiv_rn(a.x, fmaxf(nm.x, eps));
                      const float o1 = __fdiv_rn(a.y, fmaxf(nm.y, eps));
                      if (jo + 1 < p.n_out) stg_stream(reinterpret_cast<cpx*>(out + jo), mk(o0, o1));
                      else out[jo] = o0;
                    }
                  } else {
                    *reinterpret_cast<float2*>(cell) = a;
                  }
                }
              }
            }
          }
          __syncwarp();
        }
      }
    }

    // blocks of this chunk that have no frame of their own: before frame 0 (zeros) and the
    // tail after the last frame (what is left in the ring)
    for (long long m = m0; m < m1; ++m) {
      if (m >= 0 && m < F) continue;
      const long long u0 = m * H;
      for (int i = 2 * lane; i < H; i += 64) {
        const long long u = u0 + i, jo = p.frame_t0 + u;
        if (jo < 0 || jo >= p.n_out) continue;
        float2 a = make_float2(0.f, 0.f);
        if (u >= 0 && u < u_cov) a = *reinterpret_cast<const float2*>(ring + (u & mask));
        long long e = (p.t_offset + jo) % ring_len;
        if (e < 0) e += ring_len;
        const float2 nm = __ldg(reinterpret_cast<const float2*>(p.norm + e));
        const float o0 = __fdiv_rn(a.x, fmaxf(nm.x, eps));
        const float o1 = __fdiv_rn(a.y, fmaxf(nm.y, eps));
        out[jo] = o0;
        if (jo + 1 < p.n_out) out[jo + 1] = o1;
      }
    }
  }
}

template <int NFFT>
cudaError_t launch_k23(const K23Params& p, cudaStream_t st) {
  using X = GeoX<NFFT>;
  const int S = (X::G - 1) * p.hop + NFFT;
  const int acap = k23_next_pow2(S);
  const int smem = NFFT * 4 + X::C * 16 + kK23Warps * (acap * 4 + X::XCPX * 8);
  if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
  const int rdiv = (NFFT % p.hop == 0) ? NFFT / p.hop : 0;
  void (*kern)(const K23Params, const int) = nullptr;
#define CRLOT_K23_PICK(RD) kern = p.scrub ? istft_ola_kernel<NFFT, RD, true> : istft_ola_kernel<NFFT, RD, false>
  switch (rdiv) {
    case 1: CRLOT_K23_PICK(1); break;
    case 2: CRLOT_K23_PICK(2); break;
    case 4: CRLOT_K23_PICK(4); break;
    case 8: CRLOT_K23_PICK(8); break;
    default: CRLOT_K23_PICK(0); break;
  }
#undef CRLOT_K23_PICK
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return e;
  const long long nblocks = p.m_end - p.m_begin;
  const long long total_chunks = ((nblocks + p.chunk_blocks - 1) / p.chunk_blocks) * p.channels;
  const long long want_blocks = (total_chunks + kK23Warps - 1) / kK23Warps;
  const int grid = static_cast<int>(want_blocks < device_sm_count() ? want_blocks : device_sm_count());
  kern<<<grid > 0 ? grid : 1, kK23Warps * 32, smem, st>>>(p, acap);
  return cudaGetLastError();
}

}  // namespace

bool istft_ola_supported(int nfft, int hop, long long ring_len, long long t_offset, long long frame_t0,
                         long long out_chan_stride, const void* out, const void* norm) {
  int B2 = 0;                                      // 2*B of the geometry: granularity of the per-register tests
  switch (nfft) {
    case 256: B2 = 16; break;
    case 512: case 1024: B2 = 32; break;
    case 2048: B2 = 64; break;
    default: return false;
  }
  if (hop <= 0 || hop > nfft || hop % B2 != 0) return false;
  if (ring_len < hop || ring_len % 2 != 0 || ring_len >= (1ll << 30)) return false;
  if ((t_offset & 1) || (frame_t0 & 1) || (out_chan_stride & 1)) return false;
  if (reinterpret_cast<uintptr_t>(out) % 8 != 0 || reinterpret_cast<uintptr_t>(norm) % 8 != 0) return false;
  return true;
}

cudaError_t launch_istft_ola(K23Params p, cudaStream_t st) {
  if (p.n_out <= 0 || p.channels <= 0) return cudaSuccess;
  const long long H = p.hop;
  // hop blocks that intersect the output range j in [0, n_out), u = j - frame_t0
  const long long ulo = -p.frame_t0, uhi = p.n_out - p.frame_t0;
  p.m_begin = ulo >= 0 ? ulo / H : -((-ulo + H - 1) / H);
  p.m_end = uhi >= 0 ? (uhi + H - 1) / H : -((-uhi) / H);
  const long long nblocks = p.m_end - p.m_begin;
  if (p.chunk_blocks <= 0) {
    // about 6 chunks per resident warp, between 32 and 256 blocks each: the warm-up re-computation
    // (ceil(N/hop)-1 frames per chunk) stays below a few per cent
    const long long warps = static_cast<long long>(device_sm_count()) * kK23Warps;
    long long cb = nblocks * p.channels / (6 * warps);
    if (cb < 32) cb = 32;
    if (cb > 256) cb = 256;
    p.chunk_blocks = static_cast<int>(cb);
  }
  switch (p.nfft) {
    case 256: return launch_k23<256>(p, st);
    case 512: return launch_k23<512>(p, st);
    case 1024: return launch_k23<1024>(p, st);
    case 2048: return launch_k23<2048>(p, st);
    default: return cudaErrorNotSupported;
  }
}

}  // namespace crlot
