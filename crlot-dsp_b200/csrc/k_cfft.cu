// k_cfft.cu -- complex-domain FFT plan: IFftPlan::forward_complex / inverse_complex.
//
// Replaces, per row, KissFftPlan::forward_complex / inverse_complex -> kiss_fft (reference:
// dsp/fft/backends/kissfft_adapter.cc:171-246, third_party/kissfft/kiss_fft.c:236-300): the
// unnormalised forward DFT, and the inverse scaled by 1/nfft with the clean-up rule applied to both
// components (adapter:227-239).
//
// The complex core of kernels 1 and 2 (warpfft.cuh: pass 1, inter-pass twiddle, one shared-memory
// transposition, pass 2) without the real split: a C-point complex transform uses the geometry of
// the 2C-point real transform.  C = 2048 runs on the 64-registers-per-lane machinery of
// warpfft4096.cuh, which exists in the forward direction only: its inverse is conj(FFT(conj x)),
// exact in floating point (a conjugation is a sign flip).
#include "common.cuh"
#include "kernels.h"
#include "warpfft.cuh"
#include "warpfft4096.cuh"

namespace crlot {

namespace {

constexpr int kCfftWarps = 12;

template <int NFFT, bool INV>      // NFFT = 2C: the real-transform geometry that carries a C-point complex core
__global__ void __launch_bounds__(kCfftWarps * 32, 1) cfft_kernel(const CfftParams p) {
  using X = GeoX<NFFT>;
  constexpr int C = X::C, G = X::G;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  cpx* s_tw1 = reinterpret_cast<cpx*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  cpx* ex = s_tw1 + C + warp * X::XCPX;
  for (int i = threadIdx.x; i < C; i += blockDim.x) s_tw1[i] = p.tw1[i];
  __syncthreads();
  const long long total_steps = (p.rows + G - 1) / G;
  const long long stride = static_cast<long long>(gridDim.x) * kCfftWarps;
  const float scale = 1.0f / static_cast<float>(C);
  for (long long step = static_cast<long long>(blockIdx.x) * kCfftWarps + warp; step < total_steps; step += stride) {
    const long long row0 = step * G, left = p.rows - row0;
    cpx v[32];
#pragma unroll
    for (int r = 0; r < 32; ++r) {
      int g, i;
      in_coord<X>(lane, r, g, i);
      v[r] = g < left ? ldg_stream(p.in + (row0 + g) * C + i) : mk(0.f, 0.f);
    }
    pass1<X, INV>(v);
    twiddle1<X, INV>(v, lane, s_tw1);
    __syncwarp();                 // previous step's exchange reads are complete
    xchg_write<X>(v, lane, ex);
    __syncwarp();
    xchg_read<X>(v, lane, ex);
    pass2<X, INV>(v);
#pragma unroll
    for (int r = 0; r < 32; ++r) {
      int g, o;
      out_coord<X>(lane, r, g, o);
      cpx y = v[r];
      if (INV) y = mk(scrub(y.x * scale), scrub(y.y * scale));
      if (g < left) stg_stream(p.out + (row0 + g) * C + o, y);
    }
  }
}

constexpr int kCfftWarps2048 = 8;

template <bool INV>
__global__ void __launch_bounds__(kCfftWarps2048 * 32, 1) cfft_kernel_2048(const CfftParams p) {
  constexpr int C = 2048;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  cpx* s_tw = reinterpret_cast<cpx*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  cpx* ex = s_tw + C + warp * Geo4096::FX_ROUND_CPX;
  for (int i = threadIdx.x; i < C; i += blockDim.x) s_tw[i] = p.tw1[i];
  __syncthreads();
  const long long stride = static_cast<long long>(gridDim.x) * kCfftWarps2048;
  const int ob1 = Geo4096::row1(lane);
  const float scale = 1.0f / static_cast<float>(C);
  const cpx cj = mk(1.f, INV ? -1.f : 1.f);
  for (long long row = static_cast<long long>(blockIdx.x) * kCfftWarps2048 + warp; row < p.rows; row += stride) {
    const cpx* src = p.in + row * C;
    cpx v[64];
#pragma unroll
    for (int b = 0; b < 64; ++b) v[b] = ldg_stream(src + lane + 32 * b);
    if (INV) {
#pragma unroll
      for (int b = 0; b < 64; ++b) v[b] = mul2(v[b], cj);
    }
    fwd4096_pass1(v);
    fwd4096_twiddle(v, lane, s_tw);
    __syncwarp();
    fwd4096_xwrite<0>(v, lane, ex);
    __syncwarp();
    fwd4096_xread<0>(v, lane, ex);
    __syncwarp();
    fwd4096_xwrite<1>(v, lane, ex);
    __syncwarp();
    fwd4096_xread<1>(v, lane, ex);
    fwd4096_pass2(v);
    cpx* dst = p.out + row * C;
#pragma unroll
    for (int r = 0; r < 64; ++r) {
      cpx y = v[r];
      if (INV) y = mk(scrub(y.x * scale), scrub(-y.y * scale));
      stg_stream(dst + (r < 32 ? lane : ob1) + 64 * (r & 31), y);
    }
  }
}

template <int NFFT>
cudaError_t launch_cfft_n(const CfftParams& p, cudaStream_t st) {
  using X = GeoX<NFFT>;
  const int smem = X::C * 8 + kCfftWarps * X::XCPX * 8;
  auto kern = p.inverse ? cfft_kernel<NFFT, true> : cfft_kernel<NFFT, false>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  const long long steps = (p.rows + X::G - 1) / X::G;
  const long long want = (steps + kCfftWarps - 1) / kCfftWarps;
  const int grid = static_cast<int>(want < device_sm_count() ? want : device_sm_count());
  kern<<<grid > 0 ? grid : 1, kCfftWarps * 32, smem, st>>>(p);
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_cfft(const CfftParams& p, cudaStream_t st) {
  if (p.rows <= 0) return cudaSuccess;
  if (p.generic) return launch_cfft_generic(p, st);
  switch (p.n) {
    case 128: return launch_cfft_n<256>(p, st);
    case 256: return launch_cfft_n<512>(p, st);
    case 512: return launch_cfft_n<1024>(p, st);
    case 1024: return launch_cfft_n<2048>(p, st);
    case 2048: {
      const int smem = 2048 * 8 + kCfftWarps2048 * Geo4096::FX_ROUND_CPX * 8;
      auto kern = p.inverse ? cfft_kernel_2048<true> : cfft_kernel_2048<false>;
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      if (e != cudaSuccess) return e;
      const long long want = (p.rows + kCfftWarps2048 - 1) / kCfftWarps2048;
      const int grid = static_cast<int>(want < device_sm_count() ? want : device_sm_count());
      kern<<<grid > 0 ? grid : 1, kCfftWarps2048 * 32, smem, st>>>(p);
      return cudaGetLastError();
    }
    default: return cudaErrorNotSupported;
  }
}

}  // namespace crlot
