// k2_istft.cu -- kernel 2: inverse real FFT, 1/N scale, clean-up, optional synthesis window.
//
// Replaces, per frame, KissFftPlan::inverse -> kiss_fftri (reference:
// dsp/fft/backends/kissfft_adapter.cc:124-168, third_party/kissfft/kiss_fftr.c:119-155) and,
// when a synthesis window is given, the `src*win` half of axpy_windowed
// (dsp/ola/kernels.cc:24-28): the product is rounded to fp32 here exactly as
// std::fma(src, win, 0.0f) rounds it there, so kernel 3 can finish with fma(p, gain, acc).
//
// Same warp-autonomous structure as kernel 1 (warpfft.cuh): a warp-step is 1024 complex
// points; spectrum rows are read straight into registers as contiguous 8-byte-per-lane
// segments, the mirror bins come from one warp shuffle per value, frames are written straight
// from registers.  Warp-steps are interleaved over the grid so neighbouring warps stream
// neighbouring rows.
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "warpfft.cuh"
#include "warpfft4096.cuh"

namespace crlot {

namespace {

constexpr int kK2Warps = 12;

template <int NFFT, bool WIN, bool SCRUB>
__global__ void __launch_bounds__(kK2Warps * 32, 1) istft_inv_kernel(const K2Params p) {
  using X = GeoX<NFFT>;
  constexpr int C = X::C, G = X::G, A = X::A, B = X::B;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* s_win = reinterpret_cast<float*>(smem_raw);
  cpx* s_tw1 = reinterpret_cast<cpx*>(s_win + (WIN ? NFFT : 0));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  cpx* ex = s_tw1 + C + warp * X::XCPX;

  if (WIN)
    for (int i = threadIdx.x; i < NFFT; i += blockDim.x) s_win[i] = p.window[i];
  for (int i = threadIdx.x; i < C; i += blockDim.x) s_tw1[i] = p.tw1[i];
  const SplitLane<X::M1> SL = make_split_inv<X>(lane);   // lane factors of the real split (no table)
  __syncthreads();

  const long long total_steps = (p.rows + G - 1) / G;
  const long long stride = static_cast<long long>(gridDim.x) * kK2Warps;
  const int lp = mirror_lane<A>(lane);
  const float scale = 2.0f / static_cast<float>(NFFT);   // inv_split hands over halved values

  for (long long step = static_cast<long long>(blockIdx.x) * kK2Warps + warp; step < total_steps; step += stride) {
    const long long row0 = step * G;
    const long long left = p.rows - row0;
    cpx v[32];
    cpx nyq[X::M1];
#pragma unroll
    for (int j = 0; j < X::M1; ++j) {
      const int i1 = lane + 32 * j, g = i1 / A, a = i1 % A;
      const bool ok = g < left;
      const cpx* row = p.spec + (row0 + g) * (C + 1);
#pragma unroll
      for (int b = 0; b < B; ++b) v[j * B + b] = ok ? ldg_stream(row + a + A * b) : mk(0.f, 0.f);
      nyq[j] = (ok && a == 0) ? ldg_stream(row + C) : mk(0.f, 0.f);
    }
    if (p.sgain != nullptr) {     // spectral-domain hook: per-bin real gain applied as the row is loaded
#pragma unroll
      for (int j = 0; j < X::M1; ++j) {
        const int i1 = lane + 32 * j, g = i1 / A, a = i1 % A;
        if (g < left) {
          const float* gr = p.sgain + (row0 + g) * p.sgain_row_stride;
#pragma unroll
          for (int b = 0; b < B; ++b) v[j * B + b] = mul2(v[j * B + b], splat(__ldg(gr + a + A * b)));
          if (a == 0) nyq[j] = mul2(nyq[j], splat(__ldg(gr + C)));
        }
      }
    }
    cpx mid[X::M1];
    auto mirror = [&](int r) { return mk(__shfl_sync(0xffffffffu, v[r].x, lp), __shfl_sync(0xffffffffu, v[r].y, lp)); };
    inv_split_pairs<X>(v, lane, SL, mirror, nyq, mid);
    inv_split_finish<X>(v, lane, mirror, mid);
    pass1<X, true>(v);
    twiddle1<X, true>(v, lane, s_tw1);
    __syncwarp();                 // previous step's exchange reads are complete
    xchg_write<X>(v, lane, ex);
    __syncwarp();
    xchg_read<X>(v, lane, ex);
    pass2<X, true>(v);
#pragma unroll
    for (int j = 0; j < X::M2; ++j) {
      const int i2 = lane + 32 * j, g = i2 / B, ob = i2 % B;
      if (g < left) {
        float* dst = p.frames + (row0 + g) * NFFT;
#pragma unroll
        for (int oa = 0; oa < A; ++oa) {
          const int o = ob + B * oa;
          float y0 = v[j * A + oa].x * scale, y1 = v[j * A + oa].y * scale;
          if (SCRUB) {
            y0 = scrub(y0);
            y1 = scrub(y1);
          }
          if (WIN) {
            const float2 w = *reinterpret_cast<const float2*>(s_win + 2 * o);
            y0 = __fmul_rn(y0, w.x);
            y1 = __fmul_rn(y1, w.y);
          }
          stg_stream(reinterpret_cast<cpx*>(dst + 2 * o), mk(y0, y1));
        }
      }
    }
  }
}

// ---- N = 4096: one warp per frame, 64 complex registers per lane (warpfft4096.cuh) --------------
// A lane owns the spectrum columns a0 = lane and a1 = 64 - lane (32 for lane 0): bin a + 64*b, so
// the mirror bin 2048 - k of each of its values sits in the lane's own registers (no shuffles, no
// split table).  Spectrum rows are prefetched one row ahead into a per-warp staging buffer by a
// 1-D bulk async copy (UBLKCP) on the warp's own mbarrier: rows are only 8-byte aligned
// ((2049 bins) x 8 B), so the copy moves the 2048 bins that start at the row's first 16-byte
// boundary and lane 0 fetches the one remaining bin (bin 0 or bin 2048) with a plain load.
constexpr int kK2Warps4096 = 8;
constexpr int kK2Stage4096 = 2048 * 8;                       // bytes staged per row
constexpr int kK2PerWarp4096 = kK2Stage4096 + Geo4096::FX_ROUND_CPX * 8 + 16;

// TM: the per-lane inter-pass twiddles (2 x 32 complex, columns [0, 128)) and synthesis-window samples (64 complex,
// columns [128, 256)) live in tensor memory instead of shared-memory tables (see k1_stft.cu).
template <bool WIN, bool SCRUB, bool TM>
__global__ void __launch_bounds__(kK2Warps4096 * 32, 1) istft_inv_kernel_4096(const K2Params p) {
  constexpr int NFFT = 4096, C = 2048;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint32_t s_tmem_base;
  float* s_win = reinterpret_cast<float*>(smem_raw);
  cpx* s_twi = reinterpret_cast<cpx*>(s_win + (WIN && !TM ? NFFT : 0));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* wb = reinterpret_cast<unsigned char*>(s_twi + (TM ? 0 : C)) + warp * kK2PerWarp4096;
  cpx* stage = reinterpret_cast<cpx*>(wb);
  cpx* ex = stage + C;
  const uint32_t bar = smem_u32(ex + Geo4096::FX_ROUND_CPX);
  const int a1 = Geo4096::row1(lane);
  if (!TM) {
    if (WIN)
      for (int i = threadIdx.x; i < NFFT; i += blockDim.x) s_win[i] = p.window[i];
    for (int i = threadIdx.x; i < C; i += blockDim.x) s_twi[i] = p.pre[i];
  }
  const SplitLane4096 SL = make_split_4096<true>(lane);
  const Coop4096 CO = make_coop_4096<true>(lane);
  cpx dc = mk(0.f, 0.f);
  if (lane == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  uint32_t ttw = 0, twin = 0;
  if (TM) {
    if (warp == 0) tmem_alloc<256>(&s_tmem_base);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    ttw = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    twin = ttw + 128;
    if (warp < 4) {
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        cpx t8[8], w8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int r = 8 * c + i;                         // register r of the pass-1 view: ob = r & 31, column a0 / a1
          t8[i] = p.pre[(r & 31) * 64 + (r < 32 ? lane : a1)];
          w8[i] = WIN ? *reinterpret_cast<const cpx*>(p.window + 2 * (lane + 32 * r)) : mk(1.f, 1.f);
        }
        tmem_st8(ttw + 16 * c, t8);
        tmem_st8(twin + 16 * c, w8);
      }
      tmem_wait_st();
    }
    tmem_fence_before_sync();
  }
  __syncthreads();
  if (TM) tmem_fence_after_sync();
  const long long stride = static_cast<long long>(gridDim.x) * kK2Warps4096;
  const float scale = 2.0f / static_cast<float>(NFFT);    // the split hands over halved values
  uint32_t phase = 0;
  long long rowi = static_cast<long long>(blockIdx.x) * kK2Warps4096 + warp;
  // off = 1 when the row starts 8 bytes past a 16-byte boundary: stage[i] = bin i + off
  auto prefetch = [&](long long r, cpx& extra) {
    const cpx* row = p.spec + r * (C + 1);
    const int off = static_cast<int>((reinterpret_cast<uintptr_t>(row) >> 3) & 1);
    if (lane == 0) {
      fence_proxy_async_smem();
      mbar_arrive_expect_tx(bar, kK2Stage4096);
      bulk_g2s(smem_u32(stage), row + off, kK2Stage4096, bar);
      extra = ldg_stream(off ? row : row + C);
    }
    return off;
  };
  cpx extra = mk(0.f, 0.f);
  int off = 0;
  if (rowi < p.rows) off = prefetch(rowi, extra);
  for (; rowi < p.rows; rowi += stride) {
    mbar_wait(bar, phase);
    phase ^= 1u;
    // lane 0's self-mirrored columns are split in place in the staging buffer, one pair per lane
    // (warpfft4096.cuh "Coop4096"); the general in-register split below then skips lane 0
    cpx* st = stage - off;
    const float* gr = p.sgain ? p.sgain + rowi * p.sgain_row_stride : nullptr;   // spectral-domain hook
    if (lane != 0) {
      cpx xa = st[CO.k], xb = st[C - CO.k];
      if (gr) {
        xa = mul2(xa, splat(__ldg(gr + CO.k)));
        xb = mul2(xb, splat(__ldg(gr + C - CO.k)));
      }
      cpx x1, x2;
      split_pair32<0, true>(xa, xb, CO.q, CO.qr, x1, x2);
      st[CO.k] = x1;
      st[C - CO.k] = x2;
    } else {
      cpx x0 = off ? extra : stage[0];
      cpx xn = off ? stage[C - 1] : extra;
      cpx xm = st[C / 2];
      if (gr) {
        x0 = mul2(x0, splat(__ldg(gr)));
        xn = mul2(xn, splat(__ldg(gr + C)));
        xm = mul2(xm, splat(__ldg(gr + C / 2)));
      }
      dc = mk(0.5f * (x0.x + xn.x), 0.5f * (x0.x - xn.x));     // only the real parts (kiss_fftr.c:131-132)
      st[C / 2] = mul2(xm, mk(1.f, -1.f));
    }
    __syncwarp();
    cpx v[64];
#pragma unroll
    for (int b = 0; b < 32; ++b) v[b] = st[lane + 64 * b];
#pragma unroll
    for (int b = 0; b < 32; ++b) v[32 + b] = st[a1 + 64 * b];
    if (lane == 0) v[0] = dc;
    if (gr && lane != 0) {        // lane 0's values are already scaled (and split) above
#pragma unroll
      for (int b = 0; b < 32; ++b) {
        v[b] = mul2(v[b], splat(__ldg(gr + lane + 64 * b)));
        v[32 + b] = mul2(v[32 + b], splat(__ldg(gr + a1 + 64 * b)));
      }
    }
    __syncwarp();                 // every lane has read the staging buffer
    const long long nxt = rowi + stride;
    if (nxt < p.rows) off = prefetch(nxt, extra);

    if (lane != 0) split4096_general<true>(v, SL.q, SL.qr);
    inv4096_pass1(v);
    if (TM) {
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        cpx t8[8];
        tmem_ld8(ttw + 16 * c, t8);
        tmem_wait_ld8(t8);
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (((8 * c + i) & 31) != 0) v[8 * c + i] = cmulc(v[8 * c + i], t8[i]);
      }
    } else {
      inv4096_twiddle(v, lane, s_twi);
    }
    inv4096_xwrite<0>(v, lane, ex);
    __syncwarp();
    inv4096_xread<0>(v, lane, ex);
    __syncwarp();
    inv4096_xwrite<1>(v, lane, ex);
    __syncwarp();
    inv4096_xread<1>(v, lane, ex);
    __syncwarp();
    inv4096_pass2(v);
    float* dst = p.frames + rowi * NFFT;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      cpx w8[8];
      if (TM && WIN) {
        tmem_ld8(twin + 16 * c, w8);
        tmem_wait_ld8(w8);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int oa = 8 * c + i;
        const int n = lane + 32 * oa;
        float y0 = v[oa].x * scale, y1 = v[oa].y * scale;
        if (SCRUB) {
          y0 = scrub(y0);
          y1 = scrub(y1);
        }
        if (WIN) {
          const float2 w = TM ? make_float2(w8[i].x, w8[i].y) : *reinterpret_cast<const float2*>(s_win + 2 * n);
          y0 = __fmul_rn(y0, w.x);
          y1 = __fmul_rn(y1, w.y);
        }
        stg_stream(reinterpret_cast<cpx*>(dst + 2 * n), mk(y0, y1));
      }
    }
  }
  if (TM) {
    tmem_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<256>(s_tmem_base);
  }
}

static cudaError_t launch_k2_4096(const K2Params& p, cudaStream_t st) {
  // Default: the two-sub-transform kernel (k2_4096v2.cu; needs 16-byte aligned frame rows for its float4 stores and a
  // 16-byte aligned spectrum base for its bulk copies).  CRLOT_K2_4096_V1=1 selects the first version below (64
  // registers per lane, 8 warps) for A/B timing; read per call.
  if (getenv("CRLOT_K2_4096_V1") == nullptr && ((reinterpret_cast<uintptr_t>(p.frames) | reinterpret_cast<uintptr_t>(p.spec)) & 15) == 0)
    return launch_istft_inv_4096v2(p, st);
  const bool win = p.window != nullptr;
  static const bool tm = getenv("CRLOT_K2_NO_TMEM") == nullptr;   // A/B switch: tables in shared memory instead
  const int smem = (tm ? 0 : (win ? 4096 * 4 : 0) + 2048 * 8) + kK2Warps4096 * kK2PerWarp4096;
  void (*kern)(const K2Params) = nullptr;
  if (tm) {
    if (win) kern = p.scrub ? istft_inv_kernel_4096<true, true, true> : istft_inv_kernel_4096<true, false, true>;
    else     kern = p.scrub ? istft_inv_kernel_4096<false, true, true> : istft_inv_kernel_4096<false, false, true>;
  } else {
    if (win) kern = p.scrub ? istft_inv_kernel_4096<true, true, false> : istft_inv_kernel_4096<true, false, false>;
    else     kern = p.scrub ? istft_inv_kernel_4096<false, true, false> : istft_inv_kernel_4096<false, false, false>;
  }
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  const long long want_blocks = (p.rows + kK2Warps4096 - 1) / kK2Warps4096;
  const int grid = static_cast<int>(want_blocks < device_sm_count() ? want_blocks : device_sm_count());
  kern<<<grid > 0 ? grid : 1, kK2Warps4096 * 32, smem, st>>>(p);
  return cudaGetLastError();
}

template <int NFFT>
cudaError_t launch_k2(const K2Params& p, cudaStream_t st) {
  using X = GeoX<NFFT>;
  const bool win = p.window != nullptr;
  const int smem = (win ? NFFT * 4 : 0) + X::C * 8 + kK2Warps * X::XCPX * 8;
  void (*kern)(const K2Params) = nullptr;
  if (win) kern = p.scrub ? istft_inv_kernel<NFFT, true, true> : istft_inv_kernel<NFFT, true, false>;
  else     kern = p.scrub ? istft_inv_kernel<NFFT, false, true> : istft_inv_kernel<NFFT, false, false>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  const long long total_steps = (p.rows + X::G - 1) / X::G;
  const long long want_blocks = (total_steps + kK2Warps - 1) / kK2Warps;
  const int grid = static_cast<int>(want_blocks < device_sm_count() ? want_blocks : device_sm_count());
  kern<<<grid > 0 ? grid : 1, kK2Warps * 32, smem, st>>>(p);
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_istft_inv(const K2Params& p, cudaStream_t st) {
  if (p.rows <= 0) return cudaSuccess;
  if (p.generic) return launch_istft_inv_generic(p, st);
  switch (p.nfft) {
    case 256: return launch_k2<256>(p, st);
    case 512: return launch_k2<512>(p, st);
    case 1024: return launch_k2<1024>(p, st);
    case 2048: return launch_k2<2048>(p, st);
    case 4096: return launch_k2_4096(p, st);
    default: return cudaErrorNotSupported;
  }
}

}  // namespace crlot
