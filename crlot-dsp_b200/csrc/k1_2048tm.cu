// k1_2048tm.cu -- kernel 1 for N = 2048, hop = 512 with the raw samples carried in TENSOR MEMORY: 16 warps per SM.
//
// Replaces, per frame, Framer::pop -> window multiply -> KissFftPlan::forward -> kiss_fftr (reference:
// dsp/frame/framer.cc:128-182, bench/e2e_benchmark.cc:152-156, dsp/fft/backends/kissfft_adapter.cc:83-122,
// third_party/kissfft/kiss_fftr.c:63-117), like stft_fwd_kernel<2048> in k1_stft.cu -- same arithmetic, same bits.
// That kernel is bound by the FP32 pipe with three warps per scheduler (161 registers, 16.6 KB of shared memory per
// warp: the frame ring + the exchange buffer); this one drops the ring: with hop = N/4 = 8 register steps, point
// a + 32*b of frame k is point a + 32*(b-8) of frame k+1 -- it stays in its LANE -- so the three quarters of a frame the
// next frame shares live in 64 tensor-memory columns per warp (frame chunk c in ring chunk (c + s) mod 4), as in the
// one-kernel round trip (k23_istft_ola.cu) and k1_4096v2.cu.  Shared memory per warp: the exchange buffer and a 2 KB
// stage the next hop is fetched into by cp.async a frame ahead.  16 warps need at most 128 registers.
//   TMEM columns: [0,64) window samples of the lane's points, [64,128) inter-pass twiddles, [128,384) four warps x 64 raw.
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "warpfft.cuh"

namespace crlot {

namespace {

constexpr int kTmWarps = 16;
using XT = GeoX<2048>;
constexpr int kTmStage = 2048;                              // bytes: one hop of 512 samples
constexpr int kTmPerWarp = XT::XCPX * 8 + kTmStage;
constexpr int kTmSmem = kTmWarps * kTmPerWarp;

__device__ __noinline__ float4 ld4_guard_tm(const float* sig, long long len, long long idx) {
  return make_float4(idx < len ? __ldg(sig + idx) : 0.f, idx + 1 < len ? __ldg(sig + idx + 1) : 0.f,
                     idx + 2 < len ? __ldg(sig + idx + 2) : 0.f, idx + 3 < len ? __ldg(sig + idx + 3) : 0.f);
}
__device__ __forceinline__ cpx ld_pair_guard_tm(const float* sig, long long idx, long long len) {
  if (idx + 1 < len) return ldg_stream(reinterpret_cast<const cpx*>(sig + idx));
  return mk(idx < len ? __ldg(sig + idx) : 0.0f, 0.0f);
}

template <int SCRUB>
__global__ void __launch_bounds__(kTmWarps * 32, 1) stft_fwd_kernel_2048tm(const K1Params p) {
  constexpr int NFFT = 2048, C = 1024, A = 32, B = 32, H = 512;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint32_t s_tmem_base;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  cpx* ex = reinterpret_cast<cpx*>(smem_raw + warp * kTmPerWarp);
  cpx* stage = ex + XT::XCPX;

  if (warp == 0) tmem_alloc<512>(&s_tmem_base);
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();
  const uint32_t tmb = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  const uint32_t twin = tmb, ttw = tmb + 64, tring = tmb + 128 + 64 * (warp >> 2);
  if (warp < 4) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      cpx w8[8], t8[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int b = 8 * c + i;
        w8[i] = *reinterpret_cast<const cpx*>(p.window + 2 * lane + 2 * A * b);
        t8[i] = p.tw1[b * A + lane];
      }
      tmem_st8(twin + 16 * c, w8);
      tmem_st8(ttw + 16 * c, t8);
    }
    tmem_wait_st();
  }
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();

  const SplitLane<XT::M2> SL = make_split_fwd<XT>(lane);
  const int lp = mirror_lane<B>(lane);
  const long long frames_per_chunk = p.chunk_steps;
  const long long chunks_per_chan = (p.frames + frames_per_chunk - 1) / frames_per_chunk;
  const long long total_chunks = chunks_per_chan * p.channels;

  for (;;) {
    unsigned long long qq = 0;
    if (lane == 0) qq = atomicAdd(p.counter, 1u);
    qq = __shfl_sync(0xffffffffu, qq, 0);
    if (static_cast<long long>(qq) >= total_chunks) break;
    const long long c = qq / chunks_per_chan;
    const long long first = (qq % chunks_per_chan) * frames_per_chunk;
    const int nsteps = static_cast<int>(min(frames_per_chunk, p.frames - first));
    const float* sig = p.sig + c * p.chan_stride;
    cpx* spec = p.spec + (c * p.spec_chan_rows + first) * (C + 1);
    int dirty = 0;

    for (int s = 0; s < nsteps; ++s) {
      const long long f0 = (first + s) * H;                  // first sample of the frame
      unsigned int wmin = 0xffffffffu;
      cpx nf = mk(0.f, 0.f);
      auto screen = [&](const cpx (&x8)[8]) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const unsigned int u0 = __float_as_uint(x8[i].x), u1 = __float_as_uint(x8[i].y);
          wmin = min(wmin, min(u0 + u0 - 2u, u1 + u1 - 2u));
          nf = fma2(x8[i], mk(0.f, 0.f), nf);
        }
      };
      // ---- new samples into the tensor-memory ring -----------------------------------------------------------------
      if (s > 0) {                                           // only the hop that has just arrived, staged by cp.async
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncwarp();
        cpx x8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) x8[i] = stage[lane + 32 * i];
        tmem_st8(tring + 16 * ((3 + s) & 3), x8);
        if (SCRUB == 2) screen(x8);
      } else {
        const bool whole = f0 + NFFT <= p.len;
#pragma unroll 1
        for (int ch = 0; ch < 4; ++ch) {
          cpx x8[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const long long idx = f0 + 2 * lane + 2 * A * (8 * ch + i);
            x8[i] = whole ? ldg_stream(reinterpret_cast<const cpx*>(sig + idx)) : ld_pair_guard_tm(sig, idx, p.len);
          }
          tmem_st8(tring + 16 * ch, x8);
          if (SCRUB == 2) screen(x8);
        }
      }
      tmem_wait_st();
      __syncwarp();                                          // every lane has read the stage
      // ---- request the next frame's new hop now: it lands under the transform ------------------------------------------
      if (s + 1 < nsteps) {
        const long long g0 = f0 + H + 3 * H;                 // the last quarter of the next frame
        const bool in = g0 + H <= p.len;
        float4* st4 = reinterpret_cast<float4*>(stage);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const long long idx = g0 + 4 * (lane + 32 * i);
          if (in)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(st4 + i * 32 + lane)), "l"(sig + idx) : "memory");
          else
            st4[i * 32 + lane] = ld4_guard_tm(sig, p.len, idx);
        }
      }
      bool clean_all = SCRUB == 1;
      if (SCRUB == 2) {
        const bool suspect = (wmin < p.tiny2) || (nf.x != 0.0f) || (nf.y != 0.0f);
        if (__any_sync(0xffffffffu, suspect)) dirty = 5;
        clean_all = dirty > 0 || s == 0;
        if (dirty > 0) --dirty;
      }
      // ---- the frame's points out of the ring, times the window (one rounding per sample, as frame[i]*window[i]) ------
      cpx v[32];
      {
        cpx xb[2][8], wb[2][8];                              // the next chunk of samples and constants in flight
        tmem_ld8(tring + 16 * (s & 3), xb[0]);
        tmem_ld8(twin, wb[0]);
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          tmem_wait_ld8(xb[ch & 1]);
          tmem_wait_ld8(wb[ch & 1]);
          if (ch + 1 < 4) {
            tmem_ld8(tring + 16 * ((ch + 1 + s) & 3), xb[(ch + 1) & 1]);
            tmem_ld8(twin + 16 * (ch + 1), wb[(ch + 1) & 1]);
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) v[8 * ch + i] = mul2(xb[ch & 1][i], wb[ch & 1][i]);
        }
      }
      if (clean_all) {
#pragma unroll
        for (int r = 0; r < 32; ++r) v[r] = mk(scrub(v[r].x), scrub(v[r].y));
      }
      // ---- 1024-point complex transform ---------------------------------------------------------------------------------
      {
        cpx tb[2][8];
        tmem_ld8(ttw, tb[0]);
        pass1<XT, false>(v);
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          tmem_wait_ld8(tb[ch & 1]);
          if (ch + 1 < 4) tmem_ld8(ttw + 16 * (ch + 1), tb[(ch + 1) & 1]);
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (8 * ch + i > 0) v[8 * ch + i] = cmul(v[8 * ch + i], tb[ch & 1][i]);
        }
      }
      xchg_write<XT>(v, lane, ex);
      __syncwarp();
      xchg_read<XT>(v, lane, ex);
      __syncwarp();
      pass2<XT, false>(v);
      // ---- real split (mirror bins by shuffle) and the spectrum row --------------------------------------------------------
      cpx nyq[XT::M2], mid[XT::M2];
      auto mirror = [&](int r) { return mk(__shfl_sync(0xffffffffu, v[r].x, lp), __shfl_sync(0xffffffffu, v[r].y, lp)); };
      fwd_split_pairs<XT>(v, lane, SL, mirror, nyq, mid);
      fwd_split_finish<XT>(v, lane, mirror, mid);
#pragma unroll
      for (int oa = 0; oa < A; ++oa) stg_stream(spec + lane + B * oa, v[oa]);
      if (lane == 0) stg_stream(spec + C, nyq[0]);
      spec += C + 1;
    }
  }
  tmem_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(s_tmem_base);
}

}  // namespace

bool stft_fwd_2048tm_supported(const K1Params& p) {
  return p.nfft == 2048 && p.hop == 512 && p.pad == 0 && p.pad_mode == 0 && !p.generic;   // Framer addressing only (zeros beyond the end)
}

cudaError_t launch_stft_fwd_2048tm(const K1Params& p_in, cudaStream_t st) {
  K1Params p = p_in;
  p.chunk_steps = 64;
  const int mode = !p.scrub ? 0 : p.scrub == 2 ? 2 : 1;
  auto kern = mode == 0 ? stft_fwd_kernel_2048tm<0> : mode == 1 ? stft_fwd_kernel_2048tm<1> : stft_fwd_kernel_2048tm<2>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmSmem);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return e;
  const long long total_chunks = ((p.frames + p.chunk_steps - 1) / p.chunk_steps) * p.channels;
  const long long want_blocks = (total_chunks + kTmWarps - 1) / kTmWarps;
  const int grid = static_cast<int>(want_blocks < device_sm_count() ? want_blocks : device_sm_count());
  kern<<<grid > 0 ? grid : 1, kTmWarps * 32, kTmSmem, st>>>(p);
  return cudaGetLastError();
}

}  // namespace crlot
