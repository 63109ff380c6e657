// k_rt_2048tm.cu -- the whole round trip in one kernel for N = 2048, hop = 512 with BOTH rings in tensor memory: 16 warps per SM.
//
// frame + analysis window + forward real FFT + [spectral hook] + inverse real FFT + synthesis window + overlap-add +
// normalise, for callers that need neither spectra nor iFFT frames (reference: bench/e2e_benchmark.cc:149-176 -- nothing
// happens between forward and inverse but the optional per-bin processing; Framer::pop, KissFftPlan::forward / inverse,
// OLAAccumulator::push_frame_AoS / produce as cited in k1_stft.cu, k2_istft.cu, k3_ola.cu).  Same contract, chunk protocol
// and arithmetic as istft_ola_kernel<2048, 4, TM, FW = true> in k23_istft_ola.cu: output BIT-IDENTICAL to it and to
// K1 -> K2 -> K3 (tests/test_gpu_fused_all.py).  That kernel runs twelve warps (168 registers; the accumulator ring and the
// exchange buffer take 16.4 KB of shared memory per warp) and is bound by FP32 issue and instruction fetch.  This one is
// the forward half of k1_2048tm.cu glued to the inverse half of k23_2048tm.cu:
//
// * the raw samples of the current frame AND the overlap-add accumulator are lane-private (a hop is eight register steps
//   in both the pass-1 and the pass-2 layout): two rings of 64 tensor-memory columns per warp, 4 warps per lane quarter =
//   all 512 columns -- so the per-lane constants (analysis window, twiddles, scaled synthesis window) are shared-memory
//   tables here, 8 KB each, read contiguously;
// * shared memory per warp: the exchange buffer and a 2 KB stage the next hop is fetched into by cp.async a frame ahead;
//   sixteen warps at 128 registers;
// * the forward output layout is the inverse input layout (32 x 32): the spectrum row never leaves the registers; forward
//   and inverse split are both executed (each other's inverse only up to rounding);
// * warm-up frames, clean-up screening, ragged ends and tail blocks are register-granular as in k23_2048tm.cu.
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "warpfft.cuh"

namespace crlot {

namespace {

constexpr int kW = 16;
using X = GeoX<2048>;
constexpr int kStage = 2048;                               // bytes: one hop of 512 samples
constexpr int kTables = 3 * 1024 * 8;                      // analysis window, twiddles, scaled synthesis window
constexpr int kPerWarp = X::XCPX * 8 + kStage;
constexpr int kSmem = kTables + kW * kPerWarp;

__device__ __noinline__ float4 ld4_guard_rt(const float* sig, long long len, long long idx) {
  return make_float4(idx < len ? __ldg(sig + idx) : 0.f, idx + 1 < len ? __ldg(sig + idx + 1) : 0.f,
                     idx + 2 < len ? __ldg(sig + idx + 2) : 0.f, idx + 3 < len ? __ldg(sig + idx + 3) : 0.f);
}
__device__ __forceinline__ cpx ld_pair_guard_rt(const float* sig, long long idx, long long len) {
  if (idx + 1 < len) return ldg_stream(reinterpret_cast<const cpx*>(sig + idx));
  return mk(idx < len ? __ldg(sig + idx) : 0.0f, 0.0f);
}

__global__ void __launch_bounds__(kW * 32, 1) roundtrip_kernel_2048tm(const K23Params p) {
  constexpr int NFFT = 2048, C = 1024, A = 32, B = 32, H = 512;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint32_t s_tmem_base;
  cpx* s_aw = reinterpret_cast<cpx*>(smem_raw);             // analysis window, natural order: (w[2i], w[2i+1]) at i
  cpx* s_tw = s_aw + 1024;                                  // tw[ob*32 + a] = exp(-2 pi i a*ob / 1024)
  cpx* s_ws = s_tw + 1024;                                  // synthesis window times 2/N, natural order
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  cpx* ex = reinterpret_cast<cpx*>(smem_raw + kTables + warp * kPerWarp);
  cpx* stage = ex + X::XCPX;
  const float scale = 2.0f / static_cast<float>(NFFT);     // inv_split hands over halved values
  const float unscale = 0.5f * NFFT;                       // ws * N/2 == w exactly

  for (int i = threadIdx.x; i < 1024; i += blockDim.x) {
    s_aw[i] = *reinterpret_cast<const cpx*>(p.awindow + 2 * i);
    s_tw[i] = p.tw1[i];
    const float2 w = *reinterpret_cast<const float2*>(p.window + 2 * i);
    s_ws[i] = mk(w.x * scale, w.y * scale);
  }
  if (warp == 0) tmem_alloc<512>(&s_tmem_base);
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();
  const uint32_t tmb = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  const uint32_t traw = tmb + 64 * (warp >> 2), tacc = tmb + 256 + 64 * (warp >> 2);

  const SplitLane<X::M2> SLF = make_split_fwd<X>(lane);    // lane factors of the forward and the inverse split
  const SplitLane<X::M1> SLI = make_split_inv<X>(lane);
  const int lpf = mirror_lane<B>(lane), lpi = mirror_lane<A>(lane);
  const long long F = p.frames_per_chan;
  // graded chunks as in k23_2048tm.cu: big ones for the first three quarters of a channel, quarter-size ones after them
  const long long nblocks = p.m_end - p.m_begin;
  const long long cb = p.chunk_blocks, cbs = cb >= 32 ? cb / 4 : cb;
  const long long big_per_chan = (nblocks * 3 / 4) / cb;
  const long long small_per_chan = (nblocks - big_per_chan * cb + cbs - 1) / cbs;
  const long long total_big = big_per_chan * p.channels;
  const long long total_chunks = total_big + small_per_chan * p.channels;
  const long long u_cov = F > 0 ? (F - 1) * H + NFFT : 0;
  const int ring_len = static_cast<int>(p.ring_len);
  const float gain = p.gain, eps = p.eps;
  const cpx gain2 = mk(gain, gain);
  const unsigned int tiny2i = 2u * __float_as_uint(1e-30f * (0.5f * NFFT)) - 2u;

  for (;;) {
    unsigned long long q = 0;
    if (lane == 0) q = atomicAdd(p.counter, 1u);
    q = __shfl_sync(0xffffffffu, q, 0);
    if (static_cast<long long>(q) >= total_chunks) break;
    long long c, m0, m1;
    if (static_cast<long long>(q) < total_big) {
      c = q / big_per_chan;
      m0 = p.m_begin + (q % big_per_chan) * cb;
      m1 = m0 + cb;
    } else {
      const long long qs = q - total_big;
      c = qs / small_per_chan;
      m0 = p.m_begin + big_per_chan * cb + (qs % small_per_chan) * cbs;
      m1 = min(m0 + cbs, p.m_end);
    }
    const float* sig = p.sig + c * p.sig_chan_stride;
    float* out = p.out + c * p.out_chan_stride;
    const float* sgain = p.sgain ? p.sgain + c * p.spec_chan_rows * p.sgain_row_stride : nullptr;
    const long long k0 = max(0ll, m0 - 3);
    const long long k1 = min(m1, F);
    int dirty = 0;                                         // frames still to be cleaned up sample by sample (kernel 1's rule)

    {                                                      // a cleared accumulator
      cpx z8[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) z8[i] = mk(0.f, 0.f);
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) tmem_st8(tacc + 16 * ch, z8);
      tmem_wait_st();
    }
    int e_base = 0;
    {
      long long e = (p.t_offset + p.frame_t0 + k0 * H) % ring_len;
      if (e < 0) e += ring_len;
      e_base = static_cast<int>(e);
    }
    const int e_hop = H % ring_len;

    for (long long k = k0; k < k1; ++k) {
      const int s = static_cast<int>(k - k0);
      const long long f0 = k * H;                          // first sample of the frame
      // ================= forward half: kernel 1's arithmetic (k1_2048tm.cu) ===========================================
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
      if (s > 0) {                                         // only the hop that has just arrived, staged by cp.async
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncwarp();
        cpx x8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) x8[i] = stage[lane + 32 * i];
        tmem_st8(traw + 16 * ((3 + s) & 3), x8);
        if (p.k1_scrub == 2) screen(x8);
      } else {
        const bool whole = f0 + NFFT <= p.sig_len;
#pragma unroll 1
        for (int ch = 0; ch < 4; ++ch) {
          cpx x8[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const long long idx = f0 + 2 * lane + 2 * A * (8 * ch + i);
            x8[i] = whole ? ldg_stream(reinterpret_cast<const cpx*>(sig + idx)) : ld_pair_guard_rt(sig, idx, p.sig_len);
          }
          tmem_st8(traw + 16 * ch, x8);
          if (p.k1_scrub == 2) screen(x8);
        }
      }
      tmem_wait_st();
      __syncwarp();                                        // every lane has read the stage
      if (k + 1 < k1) {                                    // the next frame's new hop: lands under the two transforms
        const long long g0 = f0 + 4 * H;
        const bool in = g0 + H <= p.sig_len;
        float4* st4 = reinterpret_cast<float4*>(stage);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const long long idx = g0 + 4 * (lane + 32 * i);
          if (in)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(st4 + i * 32 + lane)), "l"(sig + idx) : "memory");
          else
            st4[i * 32 + lane] = ld4_guard_rt(sig, p.sig_len, idx);
        }
      }
      bool clean_all = p.k1_scrub == 1;
      if (p.k1_scrub == 2) {
        const bool suspect = (wmin < p.tiny2) || (nf.x != 0.0f) || (nf.y != 0.0f);
        if (__any_sync(0xffffffffu, suspect)) dirty = 5;
        clean_all = dirty > 0 || s == 0;
        if (dirty > 0) --dirty;
      }
      cpx v[32];
      {
        cpx xb[2][8];
        tmem_ld8(traw + 16 * (s & 3), xb[0]);
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          tmem_wait_ld8(xb[ch & 1]);
          if (ch + 1 < 4) tmem_ld8(traw + 16 * ((ch + 1 + s) & 3), xb[(ch + 1) & 1]);
#pragma unroll
          for (int i = 0; i < 8; ++i) v[8 * ch + i] = mul2(xb[ch & 1][i], s_aw[lane + 32 * (8 * ch + i)]);   // one rounding per sample
        }
      }
      if (clean_all) {
#pragma unroll
        for (int r = 0; r < 32; ++r) v[r] = mk(scrub(v[r].x), scrub(v[r].y));
      }
      pass1<X, false>(v);
#pragma unroll
      for (int ob = 1; ob < B; ++ob) v[ob] = cmul(v[ob], s_tw[ob * A + lane]);
      xchg_write<X>(v, lane, ex);
      __syncwarp();
      xchg_read<X>(v, lane, ex);
      __syncwarp();
      pass2<X, false>(v);
      cpx nyq[1];
      {
        cpx nyqf[1], midf[1];
        auto mirror = [&](int r) { return mk(__shfl_sync(0xffffffffu, v[r].x, lpf), __shfl_sync(0xffffffffu, v[r].y, lpf)); };
        fwd_split_pairs<X>(v, lane, SLF, mirror, nyqf, midf);
        fwd_split_finish<X>(v, lane, mirror, midf);
        nyq[0] = lane == 0 ? nyqf[0] : mk(0.f, 0.f);
      }
      // The spectrum row is in registers, register oa of lane ob = bin ob + 32*oa: the layout the inverse half wants.
      if (sgain != nullptr) {                              // spectral hook, as kernel 2 applies it on load
        const float* gr = sgain + k * p.sgain_row_stride;
#pragma unroll
        for (int b = 0; b < B; ++b) v[b] = mul2(v[b], splat(__ldg(gr + lane + A * b)));
        if (lane == 0) nyq[0] = mul2(nyq[0], splat(__ldg(gr + C)));
      }
      // ================= inverse half: kernels 2+3 (k23_2048tm.cu) ===================================================
      {
        cpx mid[1];
        auto mirror = [&](int r) { return mk(__shfl_sync(0xffffffffu, v[r].x, lpi), __shfl_sync(0xffffffffu, v[r].y, lpi)); };
        inv_split_pairs<X>(v, lane, SLI, mirror, nyq, mid);
        inv_split_finish<X>(v, lane, mirror, mid);
      }
      pass1<X, true>(v);
#pragma unroll
      for (int b = 1; b < B; ++b) v[b] = cmulc(v[b], s_tw[b * A + lane]);
      xchg_write<X>(v, lane, ex);
      __syncwarp();
      xchg_read<X>(v, lane, ex);
      __syncwarp();
      pass2<X, true>(v);                                   // v[oa] = samples 2*lane + 64*oa, +1 (raw, times N/2)

      bool clean;
      {
        unsigned int wm[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
        cpx nfs[4] = {mk(0.f, 0.f), mk(0.f, 0.f), mk(0.f, 0.f), mk(0.f, 0.f)};
#pragma unroll
        for (int r = 0; r < 32; ++r) {
          const unsigned int u0 = __float_as_uint(v[r].x), u1 = __float_as_uint(v[r].y);
          wm[r & 3] = min(wm[r & 3], min(u0 + u0 - 2u, u1 + u1 - 2u));
          nfs[r & 3] = fma2(v[r], mk(0.f, 0.f), nfs[r & 3]);
        }
        const unsigned int wmn = min(min(wm[0], wm[1]), min(wm[2], wm[3]));
        const cpx nfv = (nfs[0] + nfs[1]) + (nfs[2] + nfs[3]);
        clean = !__any_sync(0xffffffffu, (wmn < tiny2i) || (nfv.x != 0.0f) || (nfv.y != 0.0f));
      }
      if (clean) {
#pragma unroll
        for (int oa = 0; oa < 32; ++oa) v[oa] = mul2(v[oa], s_ws[lane + 32 * oa]);
      } else {
#pragma unroll
        for (int oa = 0; oa < 32; ++oa) {
          cpx y = mul2(v[oa], splat(scale));
          if (p.scrub) y = mk(scrub(y.x), scrub(y.y));
          v[oa] = mul2(y, mul2(s_ws[lane + 32 * oa], splat(unscale)));
        }
      }
      const int d = k < m0 ? static_cast<int>(m0 - k) : 0;  // warm-up frame: keep the samples at or beyond block m0 only
      const int kc = static_cast<int>(k & 3);
#pragma unroll
      for (int ch = 1; ch < 4; ++ch) {
        if (ch >= d) {
          cpx a8[8];
          if (ch < 3) {
            tmem_ld8(tacc + 16 * ((kc + ch) & 3), a8);
            tmem_wait_ld8(a8);
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) a8[i] = mk(0.f, 0.f);
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) a8[i] = fma2(v[8 * ch + i], gain2, a8[i]);
          tmem_st8(tacc + 16 * ((kc + ch) & 3), a8);
        }
      }
      if (d == 0) {                                        // the hop that leaves: complete, divide, store
        cpx a8[8];
        tmem_ld8(tacc + 16 * kc, a8);
        const long long j_first = p.frame_t0 + k * H;
        const bool inside = j_first >= 0 && j_first + H <= p.n_out;
        float* dst = out + j_first + 2 * lane;
        if (inside && p.den_rcp != nullptr) {
          float4 nm[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            int e = e_base + 2 * (lane + 32 * i);
            if (e >= ring_len) e -= ring_len;
            nm[i] = __ldg(reinterpret_cast<const float4*>(p.den_rcp + e));
          }
          tmem_wait_ld8(a8);
          bool all_ok = true;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            a8[i] = fma2(v[i], gain2, a8[i]);
            all_ok = all_ok & div_tail_ok(a8[i].x) & div_tail_ok(a8[i].y);
            stg_stream(reinterpret_cast<cpx*>(dst + 64 * i), mk(div_tail(a8[i].x, nm[i].x, nm[i].y), div_tail(a8[i].y, nm[i].z, nm[i].w)));
          }
          if (!all_ok) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              stg_stream(reinterpret_cast<cpx*>(dst + 64 * i), mk(__fdiv_rn(a8[i].x, nm[i].x), __fdiv_rn(a8[i].y, nm[i].z)));
          }
        } else {
          tmem_wait_ld8(a8);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const cpx a = fma2(v[i], gain2, a8[i]);
            const long long jo = j_first + 2 * lane + 64 * i;
            if (jo >= 0 && jo < p.n_out) {
              int e = e_base + 2 * (lane + 32 * i);
              if (e >= ring_len) e -= ring_len;
              const float2 nm = __ldg(reinterpret_cast<const float2*>(p.norm + e));
              const float o0 = __fdiv_rn(a.x, fmaxf(nm.x, eps));
              const float o1 = __fdiv_rn(a.y, fmaxf(nm.y, eps));
              if (jo + 1 < p.n_out) stg_stream(reinterpret_cast<cpx*>(out + jo), mk(o0, o1));
              else out[jo] = o0;
            }
          }
        }
      }
      tmem_wait_st();
      e_base = (e_base + e_hop) % ring_len;
    }
    if (k0 < k1) {                                         // a copy may still be in flight into the stage: none is (the last
      asm volatile("cp.async.wait_all;" ::: "memory");     // frame requests nothing), but the next chunk must not depend on it
      __syncwarp();
    }

    for (long long m = m0; m < m1; ++m) {                  // blocks without a frame of their own (see k23_2048tm.cu)
      if (m >= 0 && m < F) continue;
      const long long u0 = m * H;
      cpx a8[8];
      if (m >= 0) {
        tmem_ld8(tacc + 16 * static_cast<int>(m & 3), a8);
        tmem_wait_ld8(a8);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const long long u = u0 + 2 * lane + 64 * i, jo = p.frame_t0 + u;
        if (jo < 0 || jo >= p.n_out) continue;
        cpx a = mk(0.f, 0.f);
        if (u >= 0 && u < u_cov) a = a8[i];
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
  tmem_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(s_tmem_base);
}

}  // namespace

bool roundtrip_2048tm_supported(const K23Params& p) {
  return p.nfft == 2048 && p.hop == 512 && p.sig != nullptr && (reinterpret_cast<uintptr_t>(p.sig) & 15) == 0 &&
         (p.sig_chan_stride & 3) == 0;
}

cudaError_t launch_roundtrip_2048tm(const K23Params& p, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(roundtrip_kernel_2048tm, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return e;
  const long long nblocks = p.m_end - p.m_begin;
  const long long cb = p.chunk_blocks, cbs = cb >= 32 ? cb / 4 : cb;           // as in the kernel
  const long long big_per_chan = (nblocks * 3 / 4) / cb;
  const long long total_chunks = (big_per_chan + (nblocks - big_per_chan * cb + cbs - 1) / cbs) * p.channels;
  const long long want_blocks = (total_chunks + kW - 1) / kW;
  const int grid = static_cast<int>(want_blocks < device_sm_count() ? want_blocks : device_sm_count());
  roundtrip_kernel_2048tm<<<grid > 0 ? grid : 1, kW * 32, kSmem, st>>>(p);
  return cudaGetLastError();
}

}  // namespace crlot
