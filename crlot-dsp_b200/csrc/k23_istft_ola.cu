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
#include <cstdlib>

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

__device__ __forceinline__ void prefetch_l2(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// RDIV = nfft / hop when that is 1, 2, 4 or 8 (hop known at compile time: the three register
// classes above are compile-time ranges, the ring loads of a frame issue back to back and all
// addresses are base + immediate), else 0 (any eligible hop: every frame takes the rolled path).
//
// Shared-memory window table: ws[i] = (2/N) * w[i] (exact: a power-of-two scale; the host refuses
// windows with non-zero values so small that the product would be subnormal).  "No window" is a
// table of ones.  Fast path per frame: f = fl(v * ws) -- equal to fl(scrub(fl(v*2/N)) * w) whenever
// the clean-up rule leaves every sample alone, which a screening pass over the registers decides
// (non-finite, or 0 < |v*2/N| < 1e-30; zeros are fine: their sign cannot reach the output because
// every sum starts from +0).  Frames that fail the screen, warm-up frames and ragged blocks take
// the rolled path, which stages the raw iFFT output in the exchange buffer and applies the
// reference formula sample by sample.
// TMEM columns of a lane (TM): [0, 2B) inter-pass twiddles W_C^(a*ob) of its pass-1 items (a = lane % A),
// [2B, 2B + 2A) the scaled synthesis-window samples of its pass-2 outputs (ob = lane % B, o = ob + B*oa).
// FW (fully fused round trip, nfft 2048 / hop 512): additionally [kAWin, +2B) the analysis-window samples of the lane's
// pass-1 points and, per warp of a lane quarter, 2B columns of RAW signal samples -- the lane's own points of the
// current frame: with hop = nfft/4 = 8 register steps, sample a + A*b of frame k is sample a + A*(b-8) of frame k+1,
// i.e. it stays in the same lane, so the three quarters of a frame that the next frame shares never leave the
// lane's tensor memory and only the new hop is fetched from global memory (8 coalesced 8-byte loads per lane).
template <int NFFT, bool FW = false> struct K23Tm {
  using X = GeoX<NFFT>;
  static constexpr int kTw = 0, kWin = 2 * X::B, kAWin = kWin + 2 * X::A, kRing = kAWin + 2 * X::B;
  static constexpr int kUsed = FW ? kRing + 3 * 2 * X::B : kAWin;
  static constexpr int kCols = kUsed <= 32 ? 32 : kUsed <= 64 ? 64 : kUsed <= 128 ? 128 : kUsed <= 256 ? 256 : 512;
};

// one complex sample pair sig[idx], sig[idx+1] with zeros at and beyond len (Framer's ZERO_PAD tail)
__device__ __forceinline__ cpx ld_pair_guard(const float* sig, long long idx, long long len) {
  if (idx + 1 < len) return ldg_stream(reinterpret_cast<const cpx*>(sig + idx));
  return mk(idx < len ? __ldg(sig + idx) : 0.0f, 0.0f);
}

template <int NFFT, int RDIV, bool TM, bool FW = false>
__global__ void __launch_bounds__(kK23Warps * 32, 1) istft_ola_kernel(const K23Params p, const int acap) {
  using X = GeoX<NFFT>;
  using T = K23Tm<NFFT, FW>;
  static_assert(!FW || (TM && NFFT == 2048 && RDIV == 4), "the fully fused kernel exists for nfft 2048 / hop 512");
  constexpr int C = X::C, G = X::G, A = X::A, B = X::B;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint32_t s_tmem_base;
  float* s_win = reinterpret_cast<float*>(smem_raw);
  cpx* s_tw1 = reinterpret_cast<cpx*>(s_win + NFFT);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* wb = reinterpret_cast<unsigned char*>(s_tw1 + (TM ? 0 : C)) + warp * (acap * 4 + X::XCPX * 8);
  float* ring = reinterpret_cast<float*>(wb);
  cpx* ex = reinterpret_cast<cpx*>(ring + acap);
  const int mask = acap - 1;
  const float scale = 2.0f / static_cast<float>(NFFT);             // inv_split hands over halved values

  // the natural-order window table stays in shared memory for the rolled path (warm-up / ragged / screened frames)
  for (int i = threadIdx.x; i < NFFT; i += blockDim.x) s_win[i] = p.window[i] * scale;
  if (!TM)
    for (int i = threadIdx.x; i < C; i += blockDim.x) s_tw1[i] = p.tw1[i];
  const SplitLane<X::M1> SL = make_split_inv<X>(lane);   // lane factors of the real split (no table)
  uint32_t tmb = 0;
  if (TM) {
    if (warp == 0) tmem_alloc<T::kCols>(&s_tmem_base);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    tmb = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    if (warp < 4) {
      const int a = lane % A, ob = lane % B;
#pragma unroll
      for (int c = 0; c < B / 8; ++c) {
        cpx t8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) t8[i] = p.tw1[(8 * c + i) * A + a];
        tmem_st8(tmb + T::kTw + 16 * c, t8);
      }
#pragma unroll
      for (int c = 0; c < A / 8; ++c) {
        cpx w8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float2 w = *reinterpret_cast<const float2*>(p.window + 2 * (ob + B * (8 * c + i)));
          w8[i] = mk(w.x * scale, w.y * scale);
        }
        tmem_st8(tmb + T::kWin + 16 * c, w8);
      }
      if (FW) {
#pragma unroll
        for (int c = 0; c < B / 8; ++c) {
          cpx w8[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) w8[i] = *reinterpret_cast<const cpx*>(p.awindow + 2 * a + 2 * A * (8 * c + i));
          tmem_st8(tmb + T::kAWin + 16 * c, w8);
        }
      }
      tmem_wait_st();
    }
    tmem_fence_before_sync();
  }
  __syncthreads();
  if (TM) tmem_fence_after_sync();
  const uint32_t tring = tmb + T::kRing + 2 * B * (warp >> 2);     // FW: this warp's raw-sample columns
  const SplitLane<X::M2> SLF = make_split_fwd<X>(lane);            // FW: lane factors of the forward split
  const int lpf = mirror_lane<B>(lane);

  const int H = RDIV > 0 ? NFFT / (RDIV > 0 ? RDIV : 1) : p.hop;
  const int R = (NFFT + H - 1) / H;
  const long long F = p.frames_per_chan;
  const long long nblocks = p.m_end - p.m_begin;
  const long long chunks_per_chan = (nblocks + p.chunk_blocks - 1) / p.chunk_blocks;
  const long long total_chunks = chunks_per_chan * p.channels;
  const long long u_cov = F > 0 ? (F - 1) * H + NFFT : 0;          // u >= u_cov: no frame contributes
  const int lp = mirror_lane<A>(lane);
  const int ring_len = static_cast<int>(p.ring_len);
  const float gain = p.gain, eps = p.eps;
  const cpx gain2 = mk(gain, gain);
  // screen threshold on the raw registers: |v * 2/N| < 1e-30  <=>  |v| < 1e-30 * N/2 (exact scaling);
  // compared as 2*bits - 2 so that +-0 wraps to the top (K1 uses the same trick)
  const unsigned int tiny2 = 2u * __float_as_uint(1e-30f * (0.5f * NFFT)) - 2u;
  const float unscale = 0.5f * NFFT;                                // ws * N/2 == w exactly

  for (;;) {
    unsigned long long q = 0;
    if (lane == 0) q = atomicAdd(p.counter, 1u);
    q = __shfl_sync(0xffffffffu, q, 0);
    if (static_cast<long long>(q) >= total_chunks) break;
    const long long c = q / chunks_per_chan;
    const long long m0 = p.m_begin + (q % chunks_per_chan) * p.chunk_blocks;
    const long long m1 = min(m0 + p.chunk_blocks, p.m_end);
    const cpx* spec = p.spec + c * p.spec_chan_rows * (C + 1);
    const float* sig = FW ? p.sig + c * p.sig_chan_stride : nullptr;
    int dirty = 0;                                        // FW: frames still to be cleaned up sample by sample
    float* out = p.out + c * p.out_chan_stride;
    const float* sgain = p.sgain ? p.sgain + c * p.spec_chan_rows * p.sgain_row_stride : nullptr;
    const long long k0 = max(0ll, m0 - R + 1);
    const long long k1 = min(m1, F);

    __syncwarp();
    for (int i = lane * 4; i < acap; i += 128) *reinterpret_cast<float4*>(ring + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();

    // envelope phase of sample 0 of frame kbase, advanced per step (no 64-bit modulo per frame)
    int e_base = 0;
    {
      long long e = (p.t_offset + p.frame_t0 + k0 * H) % ring_len;
      if (e < 0) e += ring_len;
      e_base = static_cast<int>(e);
    }
    const int e_hop = H % ring_len;

    // spectrum rows of a step, straight into registers (8 bytes per lane, contiguous across the warp)
    auto load_rows = [&](cpx (&dst)[32], cpx (&dnyq)[X::M1], long long kb) {
      const long long lft = k1 - kb;
#pragma unroll
      for (int j = 0; j < X::M1; ++j) {
        const int i1 = lane + 32 * j, g = i1 / A, a = i1 % A;
        const bool ok = g < lft;
        const cpx* row = spec + (kb + g) * (C + 1);
#pragma unroll
        for (int b = 0; b < B; ++b) dst[j * B + b] = ok ? ldg_stream(row + a + A * b) : mk(0.f, 0.f);
        dnyq[j] = (ok && a == 0) ? ldg_stream(row + C) : mk(0.f, 0.f);
      }
      if (sgain != nullptr) {   // spectral-domain hook: per-bin real gain applied as the rows are loaded
#pragma unroll
        for (int j = 0; j < X::M1; ++j) {
          const int i1 = lane + 32 * j, g = i1 / A, a = i1 % A;
          if (g < lft) {
            const float* gr = sgain + (kb + g) * p.sgain_row_stride;
#pragma unroll
            for (int b = 0; b < B; ++b) dst[j * B + b] = mul2(dst[j * B + b], splat(__ldg(gr + a + A * b)));
            if (a == 0) dnyq[j] = mul2(dnyq[j], splat(__ldg(gr + C)));
          }
        }
      }
    };

    for (long long kbase = k0; kbase < k1; kbase += G) {
      const long long left = k1 - kbase;
      cpx v[32];
      cpx nyq[X::M1];
      if constexpr (FW) {
        // ---- forward half: frame kbase from the signal, kernel 1's arithmetic (k1_stft.cu) ------------------
        const int s = static_cast<int>(kbase - k0);
        const long long f0 = kbase * H;                     // first sample of the frame
        const bool whole = f0 + NFFT <= p.sig_len;
        unsigned int wmin = 0xffffffffu;
        cpx nf = mk(0.f, 0.f);
        if (kbase + 1 < k1 && lane < 16 && f0 + NFFT + H <= p.sig_len)   // the hop the NEXT frame brings in: into L2 now
          prefetch_l2(sig + f0 + NFFT + 32 * lane);
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          cpx x8[8], w8[8];
          tmem_ld8(tmb + T::kAWin + 16 * ch, w8);
          const bool fresh = s == 0 || ch == 3;             // from global memory; else carried over in tensor memory
          if (fresh) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const long long idx = f0 + 2 * lane + 2 * A * (8 * ch + i);
              x8[i] = whole ? ldg_stream(reinterpret_cast<const cpx*>(sig + idx)) : ld_pair_guard(sig, idx, p.sig_len);
            }
            // frame chunk ch sits in ring chunk (ch + s) mod 4: the chunk the oldest hop has just left
            if (ch > 0) tmem_st8(tring + 16 * ((ch + s) & 3), x8);
            if (p.k1_scrub == 2) {
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const unsigned int u0 = __float_as_uint(x8[i].x), u1 = __float_as_uint(x8[i].y);
                wmin = min(wmin, min(u0 + u0 - 2u, u1 + u1 - 2u));
                nf = fma2(x8[i], mk(0.f, 0.f), nf);
              }
            }
            tmem_wait_ld8(w8);
          } else {
            tmem_ld8(tring + 16 * ((ch + s) & 3), x8);
            tmem_wait_ld8(w8);
            tmem_wait_ld8(x8);
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            cpx pw = mul2(x8[i], w8[i]);                    // one rounding per sample, as frame[i]*window[i]
            if (p.k1_scrub == 1) pw = mk(scrub(pw.x), scrub(pw.y));
            v[8 * ch + i] = pw;
          }
        }
        tmem_wait_st();                                     // the new hop is in place before a later step reads it back
        if (p.k1_scrub == 2) {
          const bool suspect = (wmin < p.tiny2) || (nf.x != 0.0f) || (nf.y != 0.0f);
          if (__any_sync(0xffffffffu, suspect)) dirty = 5;
          if (dirty > 0 || s == 0) {                        // s == 0: every sample of the frame is new
#pragma unroll
            for (int r = 0; r < 32; ++r) v[r] = mk(scrub(v[r].x), scrub(v[r].y));
            if (dirty > 0) --dirty;
          }
        }
        pass1<X, false>(v);
#pragma unroll
        for (int cc = 0; cc < B / 8; ++cc) {
          cpx t8[8];
          tmem_ld8(tmb + T::kTw + 16 * cc, t8);
          tmem_wait_ld8(t8);
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (8 * cc + i > 0) v[8 * cc + i] = cmul(v[8 * cc + i], t8[i]);
        }
        xchg_write<X>(v, lane, ex);
        __syncwarp();
        xchg_read<X>(v, lane, ex);
        __syncwarp();
        pass2<X, false>(v);
        {
          cpx nyqf[X::M2], midf[X::M2];
          auto mirror = [&](int r) { return mk(__shfl_sync(0xffffffffu, v[r].x, lpf), __shfl_sync(0xffffffffu, v[r].y, lpf)); };
          fwd_split_pairs<X>(v, lane, SLF, mirror, nyqf, midf);
          fwd_split_finish<X>(v, lane, mirror, midf);
          nyq[0] = lane == 0 ? nyqf[0] : mk(0.f, 0.f);
        }
        // The spectrum row is now in registers, register oa of lane ob = bin ob + 32*oa: exactly the layout the
        // inverse half wants (register b of lane a = bin a + 32*b).  The spectral hook, as kernel 2 applies it on load:
        if (sgain != nullptr) {
          const float* gr = sgain + kbase * p.sgain_row_stride;
#pragma unroll
          for (int b = 0; b < B; ++b) v[b] = mul2(v[b], splat(__ldg(gr + lane + A * b)));
          if (lane == 0) nyq[0] = mul2(nyq[0], splat(__ldg(gr + C)));
        }
      } else {
      load_rows(v, nyq, kbase);
      // pull the next step's rows into L2 while this one is transformed (12 warps per SM cannot
      // cover a DRAM round trip by themselves)
      if (kbase + G < k1) {
        const char* nxt = reinterpret_cast<const char*>(spec + (kbase + G) * (C + 1));
        constexpr int kBytes = G * (C + 1) * 8;
#pragma unroll
        for (int o = 0; o < (kBytes + 4095) / 4096; ++o)
          if (o * 4096 + lane * 128 < kBytes) prefetch_l2(nxt + o * 4096 + lane * 128);
      }
      }
      {
        cpx mid[X::M1];
        auto mirror = [&](int r) { return mk(__shfl_sync(0xffffffffu, v[r].x, lp), __shfl_sync(0xffffffffu, v[r].y, lp)); };
        inv_split_pairs<X>(v, lane, SL, mirror, nyq, mid);
        inv_split_finish<X>(v, lane, mirror, mid);
      }
      pass1<X, true>(v);
      if (TM) {
#pragma unroll
        for (int c = 0; c < B / 8; ++c) {
          cpx t8[8];
          tmem_ld8(tmb + T::kTw + 16 * c, t8);
          tmem_wait_ld8(t8);
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (8 * c + i > 0)
#pragma unroll
              for (int j = 0; j < X::M1; ++j) v[j * B + 8 * c + i] = cmulc(v[j * B + 8 * c + i], t8[i]);
        }
      } else {
        twiddle1<X, true>(v, lane, s_tw1);
      }
      xchg_write<X>(v, lane, ex);
      __syncwarp();
      xchg_read<X>(v, lane, ex);
      __syncwarp();
      pass2<X, true>(v);

      // is every frame of this step a plain interior frame of the chunk?
      const long long j_first = p.frame_t0 + kbase * H;               // output index of sample 0 of frame kbase
      bool fast = RDIV > 0 && p.den_rcp != nullptr && left >= G && kbase >= m0 && j_first >= 0 && j_first + static_cast<long long>(G) * H <= p.n_out;
      if (fast) {
        unsigned int wm[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
        cpx nfs[4] = {mk(0.f, 0.f), mk(0.f, 0.f), mk(0.f, 0.f), mk(0.f, 0.f)};
#pragma unroll
        for (int r = 0; r < 32; ++r) {       // four independent chains
          const unsigned int u0 = __float_as_uint(v[r].x), u1 = __float_as_uint(v[r].y);
          wm[r & 3] = min(wm[r & 3], min(u0 + u0 - 2u, u1 + u1 - 2u));
          nfs[r & 3] = fma2(v[r], mk(0.f, 0.f), nfs[r & 3]);
        }
        const unsigned int wmin = min(min(wm[0], wm[1]), min(wm[2], wm[3]));
        const cpx nf = (nfs[0] + nfs[1]) + (nfs[2] + nfs[3]);
        const bool suspect = (wmin < tiny2) || (nf.x != 0.0f) || (nf.y != 0.0f);
        fast = !__any_sync(0xffffffffu, suspect);
      }

      if (fast) {
        constexpr int AE = RDIV > 0 ? A / RDIV : A;     // registers per hop
        // {denominator, reciprocal} pairs of this lane's emit registers, requested before the ring work
        float4 nm[X::M2][AE];
#pragma unroll
        for (int j = 0; j < X::M2; ++j) {
          const int i2 = lane + 32 * j, g = i2 / B, ob = i2 % B;
          int e0 = e_base + g * e_hop;
          e0 %= ring_len;
#pragma unroll
          for (int oa = 0; oa < AE; ++oa) {
            int e = e0 + 2 * (ob + B * oa);
            if (e >= ring_len) e -= ring_len;
            nm[j][oa] = __ldg(reinterpret_cast<const float4*>(p.den_rcp + e));
          }
        }
        // f = fl(v * ws), every frame of the step at once
        if (TM) {
#pragma unroll
          for (int c = 0; c < A / 8; ++c) {
            cpx w8[8];
            tmem_ld8(tmb + T::kWin + 16 * c, w8);
            tmem_wait_ld8(w8);
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
              for (int j = 0; j < X::M2; ++j) v[j * A + 8 * c + i] = mul2(v[j * A + 8 * c + i], w8[i]);
          }
        } else {
#pragma unroll
          for (int j = 0; j < X::M2; ++j) {
            const int ob = (lane + 32 * j) % B;
#pragma unroll
            for (int oa = 0; oa < A; ++oa)
              v[j * A + oa] = mul2(v[j * A + oa], *reinterpret_cast<const cpx*>(s_win + 2 * (ob + B * oa)));
          }
        }
        // overlap-add, one frame at a time so that every ring cell sees its frames in ascending order
#pragma unroll
        for (int g = 0; g < G; ++g) {
          const long long u0 = (kbase + g) * H;
          const int r0 = static_cast<int>(u0 & mask);
#pragma unroll
          for (int j = 0; j < X::M2; ++j) {
            const int i2 = lane + 32 * j, ob = i2 % B;
            if (i2 / B == g) {
              const int rb = r0 + 2 * ob;
#pragma unroll
              for (int oa = 0; oa < A; ++oa) {            // all ring reads of the frame, back to back
                cpx a = mk(0.f, 0.f);
                if (oa < A - AE) a = *reinterpret_cast<const cpx*>(ring + ((rb + 2 * B * oa) & mask));
                v[j * A + oa] = fma2(v[j * A + oa], gain2, a);
              }
#pragma unroll
              for (int oa = AE; oa < A; ++oa) *reinterpret_cast<cpx*>(ring + ((rb + 2 * B * oa) & mask)) = v[j * A + oa];
              float* dst = out + (p.frame_t0 + u0) + 2 * ob;
              bool all_ok = true;
#pragma unroll
              for (int oa = 0; oa < AE; ++oa) {
                const cpx a = v[j * A + oa];
                const float4 t = nm[j][oa];
                all_ok = all_ok & div_tail_ok(a.x) & div_tail_ok(a.y);   // no short-circuit: one predicate chain, no branches
                stg_stream(reinterpret_cast<cpx*>(dst + 2 * B * oa), mk(div_tail(a.x, t.x, t.y), div_tail(a.y, t.z, t.w)));
              }
              if (!all_ok) {                             // operands outside the tail's range: the full division
#pragma unroll
                for (int oa = 0; oa < AE; ++oa) {
                  const cpx a = v[j * A + oa];
                  const float4 t = nm[j][oa];
                  stg_stream(reinterpret_cast<cpx*>(dst + 2 * B * oa), mk(__fdiv_rn(a.x, t.x), __fdiv_rn(a.y, t.z)));
                }
              }
            }
          }
          if (G > 1) __syncwarp();
        }
      } else {
        // rolled path: raw iFFT output -> exchange buffer (natural order), then the reference formula
#pragma unroll
        for (int j = 0; j < X::M2; ++j) {
          const int i2 = lane + 32 * j, g = i2 / B, ob = i2 % B;
#pragma unroll
          for (int oa = 0; oa < A; ++oa) ex[g * C + ob + B * oa] = v[j * A + oa];
        }
        __syncwarp();
        for (int g = 0; g < G && g < left; ++g) {
          const long long k = kbase + g;
          const int thr = k < m0 ? static_cast<int>((m0 - k) * H) : 0;   // warm-up: keep u >= m0*H only
          const bool emit = k >= m0;
          const long long u0 = k * H;
          const int r0 = static_cast<int>(u0 & mask);
          const long long j0 = p.frame_t0 + u0;
          const int e0 = (e_base + g * e_hop) % ring_len;
          for (int o = lane; o < C; o += 32) {
            const int i = 2 * o;
            if (i < thr) continue;                        // thr is a multiple of 64: warp-uniform
            const cpx raw = ex[g * C + o];
            float y0 = raw.x * scale, y1 = raw.y * scale;
            if (p.scrub) {
              y0 = scrub(y0);
              y1 = scrub(y1);
            }
            const float2 w = *reinterpret_cast<const float2*>(s_win + i);
            y0 = __fmul_rn(y0, w.x * unscale);
            y1 = __fmul_rn(y1, w.y * unscale);
            float* cell = ring + ((r0 + i) & mask);
            float2 a = make_float2(0.f, 0.f);
            if (i < NFFT - H) a = *reinterpret_cast<const float2*>(cell);     // else: first touch
            a.x = __fmaf_rn(y0, gain, a.x);
            a.y = __fmaf_rn(y1, gain, a.y);
            if (emit && i < H) {
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
          __syncwarp();
        }
      }
      e_base = (e_base + G * e_hop) % ring_len;
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
  if (TM) {
    tmem_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<T::kCols>(s_tmem_base);
  }
}

template <int NFFT>
cudaError_t launch_k23(const K23Params& p, cudaStream_t st) {
  using X = GeoX<NFFT>;
  const int S = (X::G - 1) * p.hop + NFFT;
  const int acap = k23_next_pow2(S);
  static const bool tm_env = getenv("CRLOT_K23_NO_TMEM") == nullptr;   // A/B switch: twiddle / window tables in shared memory
  const bool tm = tm_env || p.sig != nullptr;
  const int smem = NFFT * 4 + (tm ? 0 : X::C * 8) + kK23Warps * (acap * 4 + X::XCPX * 8);
  if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
  const int rdiv = (NFFT % p.hop == 0) ? NFFT / p.hop : 0;
  // nfft 2048 / hop 512: the 16-warp kernel with the accumulator ring in tensor memory (k23_2048tm.cu; same bits).
  // CRLOT_K23_NO_TM16=1 keeps the kernel below for the A/B (read per call).
  if (NFFT == 2048 && istft_ola_2048tm_supported(p) && getenv("CRLOT_K23_NO_TM16") == nullptr) return launch_istft_ola_2048tm(p, st);
  void (*kern)(const K23Params, const int) = nullptr;
#define CRLOT_K23_PICK(RD) kern = tm ? istft_ola_kernel<NFFT, RD, true> : istft_ola_kernel<NFFT, RD, false>
  switch (rdiv) {
    case 1: CRLOT_K23_PICK(1); break;
    case 2: CRLOT_K23_PICK(2); break;
    case 4: CRLOT_K23_PICK(4); break;
    case 8: CRLOT_K23_PICK(8); break;
    default: CRLOT_K23_PICK(0); break;
  }
#undef CRLOT_K23_PICK
  if (p.sig != nullptr) {                        // fully fused round trip
    if constexpr (NFFT == 2048) {
      if (rdiv != 4) return cudaErrorNotSupported;
      // 16-byte aligned signal rows: the 16-warp kernel with both rings in tensor memory (k_rt_2048tm.cu; same bits).
      // CRLOT_RT_NO_TM16=1 keeps the kernel below for the A/B (read per call).
      if (roundtrip_2048tm_supported(p) && getenv("CRLOT_RT_NO_TM16") == nullptr) return launch_roundtrip_2048tm(p, st);
      kern = istft_ola_kernel<2048, 4, true, true>;
    } else {
      return cudaErrorNotSupported;
    }
  }
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
    case 2048: case 4096: B2 = 64; break;
    default: return false;
  }
  if (hop <= 0 || hop > nfft || hop % B2 != 0) return false;
  if (ring_len < hop || ring_len % 2 != 0 || ring_len >= (1ll << 30)) return false;
  if ((t_offset & 1) || (frame_t0 & 1) || (out_chan_stride & 1)) return false;
  if (reinterpret_cast<uintptr_t>(out) % 8 != 0 || reinterpret_cast<uintptr_t>(norm) % 8 != 0) return false;
  return true;
}

bool roundtrip_fused_supported(int nfft, int hop, long long ring_len, long long t_offset, long long frame_t0,
                               long long out_chan_stride, const void* out, const void* norm, const void* sig,
                               long long sig_chan_stride) {
  if (nfft != 2048 || hop != 512) return false;
  if (reinterpret_cast<uintptr_t>(sig) % 8 != 0 || (sig_chan_stride & 1)) return false;
  return istft_ola_supported(nfft, hop, ring_len, t_offset, frame_t0, out_chan_stride, out, norm);
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
    const long long warps = static_cast<long long>(device_sm_count()) * (p.nfft == 4096 ? 8 : kK23Warps);
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
    case 4096: return launch_istft_ola_4096(p, st);
    default: return cudaErrorNotSupported;
  }
}

}  // namespace crlot
