// k1_stft.cu -- kernel 1: fused frame extraction + analysis window + forward real FFT (sm_100a).
//
// Replaces, per frame, the reference chain Framer::pop -> window multiply loop -> KissFftPlan::
// forward -> kiss_fftr (reference: dsp/frame/framer.cc:128-182, bench/e2e_benchmark.cc:152-156,
// dsp/fft/backends/kissfft_adapter.cc:83-122, third_party/kissfft/kiss_fftr.c:63-117).
//
// Design (see DESIGN.md "Kernel 1"):
//  * warp-autonomous: every warp is its own streaming pipeline over a run ("chunk") of
//    consecutive frames of one channel; there is no block-level barrier after start-up.
//  * the warp keeps the samples of its current G frames in a private shared-memory ring and only
//    the G*hop NEW samples per step are fetched from global memory, by one 1-D bulk async copy
//    (cp.async.bulk -> SASS UBLKCP) completing on the warp's own mbarrier.  The copy for step s+1
//    is issued as soon as step s has been read into registers, so it overlaps the whole FFT.
//    Every input sample therefore crosses L2->SM once per chunk (not nfft/hop times).
//  * 1024 complex points per warp-step live in registers (32 per lane); the transform is two
//    in-register passes with one shared-memory transposition in between (warpfft.cuh), the
//    real split uses warp shuffles, and the spectrum is stored straight from registers as
//    contiguous 8-byte-per-lane rows.
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "warpfft.cuh"
#include "warpfft4096.cuh"

namespace crlot {

namespace {

constexpr int kK1Warps = 12;

__host__ __device__ inline int next_pow2(int v) {
  int r = 1;
  while (r < v) r <<= 1;
  return r;
}

// TM = the per-lane window and inter-pass twiddle constants live in tensor memory (no shared-memory tables)
template <int NFFT, bool TM>
struct K1Smem {
  using X = GeoX<NFFT>;
  static constexpr int kTableBytes = TM ? 0 : NFFT * 4 + X::C * 8;
  // TMEM columns: window (2B) + twiddles (2B), rounded up to a power of two >= 32
  static constexpr int kTmCols = 4 * X::B <= 32 ? 32 : 4 * X::B <= 64 ? 64 : 128;
  __host__ __device__ static int per_warp_bytes(int rcap) { return rcap * 4 + X::XCPX * 8 + 16; }
  __host__ __device__ static int total(int rcap, int nwarps) { return kTableBytes + nwarps * per_warp_bytes(rcap); }
};

// inter-pass twiddle from tensor memory: chunks of 8 factors, the next chunk in flight while this one multiplies
template <class X, bool INV> __device__ __forceinline__ void twiddle1_tmem(cpx (&v)[32], uint32_t ttw, cpx (&t0)[8]) {
  constexpr int NCH = X::B / 8;
  cpx tb[2][8];
#pragma unroll
  for (int i = 0; i < 8; ++i) tb[0][i] = t0[i];
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    tmem_wait_ld8(tb[c & 1]);
    if (c + 1 < NCH) tmem_ld8(ttw + 16 * (c + 1), tb[(c + 1) & 1]);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int ob = 8 * c + i;
      if (ob == 0) continue;
#pragma unroll
      for (int j = 0; j < X::M1; ++j)
        v[j * X::B + ob] = INV ? cmulc(v[j * X::B + ob], tb[c & 1][i]) : cmul(v[j * X::B + ob], tb[c & 1][i]);
    }
  }
}

// Where a frame's samples come from: position q of the (virtually padded) stream of a channel is
// sample q - pad of the signal; positions outside [pad, pad + len) take FrameQueue's padding value
// (pad = 0, mode 0 is the Framer case: zeros beyond the end).
struct SigView {
  const float* sig;
  long long len;
  int pad, pad_mode;
};

// Fill ring positions of stream positions [abs0, abs0+count) of this channel: the 16-byte
// multiple that exists in the signal by bulk copy; padding on either side and a ragged remainder
// by plain loads.  abs0, pad and count are multiples of 4.  All lanes call this after a
// __syncwarp() that follows their last read of the overwritten region.
__device__ __forceinline__ void ring_fill(float* ring, int mask, const SigView& sv, long long abs0, int count,
                                          uint32_t bar, int lane) {
  const long long lo = abs0 - sv.pad;                  // signal index of the first position
  int lead = 0;                                        // positions before the signal starts
  if (lo < 0) lead = static_cast<int>(-lo < count ? -lo : count);
  long long avail = sv.len - (lo + lead);
  if (avail < 0) avail = 0;
  if (avail > count - lead) avail = count - lead;
  const int nb = static_cast<int>(avail) & ~3;
  if (lead > 0 || lead + nb < count) {   // only at the edges of a channel
    for (int i = lane; i < count; i += 32)
      if (i < lead || i >= lead + nb) ring[(abs0 + i) & mask] = pad_value(sv.sig, sv.len, lo + i, sv.pad_mode);
    fence_proxy_async_smem();   // the only generic-proxy WRITES to the ring (the steady state only reads it: no fence there --
                                // a proxy fence is a MEMBAR that also waits for the previous frame's spectrum stores)
    __syncwarp();     // the edge is in place before the barrier phase can complete
  }
  if (elect_one()) {
    mbar_arrive_expect_tx(bar, static_cast<uint32_t>(nb) * 4u);
    if (nb > 0) {
      const int r0 = static_cast<int>((abs0 + lead) & mask);
      const int first = min(nb, mask + 1 - r0);
      const float* src = sv.sig + (lo + lead);
      bulk_g2s(smem_u32(ring + r0), src, static_cast<uint32_t>(first) * 4u, bar);
      if (nb > first) bulk_g2s(smem_u32(ring), src + first, static_cast<uint32_t>(nb - first) * 4u, bar);
    }
  }
}

// SCRUB: 0 = none; 1 = the reference clean-up rule applied to every windowed sample;
//        2 = same results as 1, but samples are SCREENED once, when they first enter the ring
//            (requires nfft == 4*hop): a raw sample can only make a windowed product that the
//            rule would change if it is non-finite or 0 < |x| < tiny_thr (= 1e-30 / smallest
//            non-zero |window|, set by the host).  Only the registers holding the newest hop are
//            screened (2 integer ops + half a packed FMA per sample); when a warp sees a suspect
//            it applies rule 1 to whole frames until the suspect has left the ring.
template <int NFFT, int SCRUB, bool TM>
__global__ void __launch_bounds__(kK1Warps * 32, 1) stft_fwd_kernel(const K1Params p, const int rcap) {
  using X = GeoX<NFFT>;
  using SM = K1Smem<NFFT, TM>;
  constexpr int C = X::C, G = X::G, A = X::A, B = X::B;
  constexpr int NCH = B / 8;                              // chunks of 8 per-lane constants (16 TMEM columns)
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint32_t s_tmem_base;
  float* s_win = reinterpret_cast<float*>(smem_raw);
  cpx* s_tw1 = reinterpret_cast<cpx*>(s_win + NFFT);
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform by construction
  unsigned char* wb = smem_raw + SM::kTableBytes + warp * SM::per_warp_bytes(rcap);
  float* ring = reinterpret_cast<float*>(wb);
  cpx* ex = reinterpret_cast<cpx*>(ring + rcap);
  const uint32_t bar = smem_u32(ex + X::XCPX);
  const int mask = rcap - 1;
  const int a_lane = lane % A;

  if (!TM) {
    for (int i = threadIdx.x; i < NFFT; i += blockDim.x) s_win[i] = p.window[i];
    for (int i = threadIdx.x; i < C; i += blockDim.x) s_tw1[i] = p.tw1[i];
  }
  const SplitLane<X::M2> SL = make_split_fwd<X>(lane);   // lane factors of the real split (no table)
  if (lane == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  // Tensor memory holds what is constant per LANE for the whole kernel: the window samples of the lane's pass-1
  // points (a + A*b, b < B: columns [0, 2B)) and its inter-pass twiddles W_C^(a*ob) (columns [2B, 4B)).  Lane t of
  // every warp has a = t % A, and warps w, w+4, w+8 share TMEM lanes 32*(w%4)..+31, so warps 0..3 write the
  // constants once and all twelve read them: 16 KB per frame that no longer crosses the shared-memory pipe.
  uint32_t twin = 0, ttw = 0;
  if (TM) {
    if (warp == 0) tmem_alloc<SM::kTmCols>(&s_tmem_base);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    twin = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    ttw = twin + 2 * B;
    if (warp < 4) {
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        cpx w8[8], t8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int b = 8 * c + i;
          w8[i] = *reinterpret_cast<const cpx*>(p.window + 2 * a_lane + 2 * A * b);
          t8[i] = p.tw1[b * A + a_lane];
        }
        tmem_st8(twin + 16 * c, w8);
        tmem_st8(ttw + 16 * c, t8);
      }
      tmem_wait_st();
    }
    tmem_fence_before_sync();
  }
  __syncthreads();
  if (TM) tmem_fence_after_sync();

  const int H = p.hop;
  const int S = (G - 1) * H + NFFT;                     // samples spanned by the G frames of one step
  const long long frames_per_chunk = static_cast<long long>(p.chunk_steps) * G;
  const long long chunks_per_chan = (p.frames + frames_per_chunk - 1) / frames_per_chunk;
  const long long total_chunks = chunks_per_chan * p.channels;
  const int lp = mirror_lane<B>(lane);
  uint32_t parity = 0;

  // per-lane constant offsets of the pass-1 items (frame g, point a): float offset g*H + 2a
  int in_off[X::M1];
#pragma unroll
  for (int j = 0; j < X::M1; ++j) {
    const int i1 = lane + 32 * j;
    in_off[j] = (i1 / A) * H + 2 * (i1 % A);
  }

  for (;;) {
    unsigned long long q = 0;
    if (lane == 0) q = atomicAdd(p.counter, 1u);
    q = __shfl_sync(0xffffffffu, q, 0);
    if (static_cast<long long>(q) >= total_chunks) break;
    const long long c = q / chunks_per_chan;
    const long long first = (q % chunks_per_chan) * frames_per_chunk;
    const long long nfr = min(frames_per_chunk, p.frames - first);
    const SigView sig{p.sig + c * p.chan_stride, p.len, p.pad, p.pad_mode};
    cpx* spec = p.spec + (c * p.spec_chan_rows + first) * (C + 1);
    long long pos = first * H;
    const int nsteps = static_cast<int>((nfr + G - 1) / G);

    int dirty = 0;
    __syncwarp();
    ring_fill(ring, mask, sig, pos, S, bar, lane);

    for (int s = 0; s < nsteps; ++s) {
      cpx wb8[2][8];                         // window constants, 8 at a time; the next chunk loads while this one multiplies
      if (TM) tmem_ld8(twin, wb8[0]);        // in flight across the barrier wait
      mbar_wait(bar, parity);
      parity ^= 1u;
      __syncwarp();

      cpx v[32];
      unsigned int wmin = 0xffffffffu;       // min over screened samples of 2*|bits| - 2 (0 maps to max)
      cpx nf = mk(0.f, 0.f);                 // sum of 0*x: becomes NaN iff a screened sample is NaN/Inf
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        cpx x8[X::M1][8];
#pragma unroll
        for (int j = 0; j < X::M1; ++j)
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int ridx = static_cast<int>((pos + in_off[j] + 2 * A * (8 * c + i)) & mask);
            x8[j][i] = *reinterpret_cast<const cpx*>(ring + ridx);
          }
        if (TM) {
          tmem_wait_ld8(wb8[c & 1]);
          if (c + 1 < NCH) tmem_ld8(twin + 16 * (c + 1), wb8[(c + 1) & 1]);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) wb8[c & 1][i] = *reinterpret_cast<const cpx*>(s_win + 2 * a_lane + 2 * A * (8 * c + i));
        }
#pragma unroll
        for (int j = 0; j < X::M1; ++j) {
          // frames of this item across lanes: g in [32j/A, (31+32j)/A]; newest hop of frame g
          // starts at point (3-g)*C/4 (nfft == 4*hop), i.e. register b >= (3-g)*B/4
          const int g_hi = (31 + 32 * j) / A;
          const int newb = g_hi >= 3 ? 0 : (3 - g_hi) * B / 4;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int b = 8 * c + i;
            const cpx x = x8[j][i];
            cpx pw = mul2(x, wb8[c & 1][i]);     // one rounding per sample, as frame[i]*window[i]
            if (SCRUB == 1) pw = mk(scrub(pw.x), scrub(pw.y));
            if (SCRUB == 2 && b >= newb) {
              const unsigned int u0 = __float_as_uint(x.x), u1 = __float_as_uint(x.y);
              wmin = min(wmin, min(u0 + u0 - 2u, u1 + u1 - 2u));   // sign bit leaves by wrap-around
              nf = fma2(x, mk(0.f, 0.f), nf);
            }
            v[j * B + b] = pw;
          }
        }
      }
      if (SCRUB == 2) {
        const bool suspect = (wmin < p.tiny2) || (nf.x != 0.0f) || (nf.y != 0.0f);
        if (__any_sync(0xffffffffu, suspect)) dirty = 3 / G + 2;
        if (dirty > 0 || s == 0) {             // s == 0: the whole ring is new
#pragma unroll
          for (int r = 0; r < 32; ++r) v[r] = mk(scrub(v[r].x), scrub(v[r].y));
          if (dirty > 0) --dirty;
        }
      }
      __syncwarp();
      if (s + 1 < nsteps) ring_fill(ring, mask, sig, pos + S, G * H, bar, lane);

      cpx t0[8];
      if (TM) tmem_ld8(ttw, t0);             // first twiddle chunk in flight during pass 1
      pass1<X, false>(v);
      if (TM) twiddle1_tmem<X, false>(v, ttw, t0);
      else twiddle1<X, false>(v, lane, s_tw1);
      xchg_write<X>(v, lane, ex);
      __syncwarp();
      xchg_read<X>(v, lane, ex);
      __syncwarp();
      pass2<X, false>(v);
      cpx nyq[X::M2], mid[X::M2];
      auto mirror = [&](int r) { return mk(__shfl_sync(0xffffffffu, v[r].x, lp), __shfl_sync(0xffffffffu, v[r].y, lp)); };
      fwd_split_pairs<X>(v, lane, SL, mirror, nyq, mid);
      fwd_split_finish<X>(v, lane, mirror, mid);

      const long long left = nfr - static_cast<long long>(s) * G;
#pragma unroll
      for (int j = 0; j < X::M2; ++j) {
        const int i2 = lane + 32 * j, g = i2 / B, ob = i2 % B;
        if (g < left) {
          cpx* row = spec + static_cast<long long>(g) * (C + 1);
#pragma unroll
          for (int oa = 0; oa < A; ++oa) stg_stream(row + ob + B * oa, v[j * A + oa]);
          if (ob == 0) stg_stream(row + C, nyq[j]);
        }
      }
      pos += static_cast<long long>(G) * H;
      spec += static_cast<long long>(G) * (C + 1);
    }
  }
  if (TM) {
    tmem_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<SM::kTmCols>(s_tmem_base);
  }
}

// ---- N = 4096: one warp per frame, 64 complex registers per lane (warpfft4096.cuh) --------------
// Same pipeline as above with G = 1: private 4096-float ring (the frame itself; the hop that has
// just been read into registers is the one the next bulk copy overwrites), two-round exchange so
// that the warps fit the 227 KB of shared memory, in-lane real split (no shuffles, no table).
// The per-lane window (64 complex = 128 TMEM columns) and inter-pass twiddles (128 columns) live in
// tensor memory as in the kernel above: no shared-memory tables, and one tcgen05.ld per 8 constants
// instead of 8 LDS.64 with their address arithmetic.
constexpr int kK1Warps4096 = 8;

template <int SCRUB, bool TM>
__global__ void __launch_bounds__(kK1Warps4096 * 32, 1) stft_fwd_kernel_4096(const K1Params p) {
  constexpr int NFFT = 4096, C = 2048, RCAP = 4096, mask = RCAP - 1;
  constexpr int kPerWarp = RCAP * 4 + Geo4096::FX_ROUND_CPX * 8 + 16;
  constexpr int kTable = TM ? 0 : NFFT * 4 + C * 8;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint32_t s_tmem_base;
  float* s_win = reinterpret_cast<float*>(smem_raw);
  cpx* s_tw = reinterpret_cast<cpx*>(s_win + NFFT);
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform by construction
  unsigned char* wb = smem_raw + kTable + warp * kPerWarp;
  float* ring = reinterpret_cast<float*>(wb);
  cpx* ex = reinterpret_cast<cpx*>(ring + RCAP);
  const uint32_t bar = smem_u32(ex + Geo4096::FX_ROUND_CPX);

  if (!TM) {
    for (int i = threadIdx.x; i < NFFT; i += blockDim.x) s_win[i] = p.window[i];
    for (int i = threadIdx.x; i < C; i += blockDim.x) s_tw[i] = p.tw1[i];
  }
  const SplitLane4096 SL = make_split_4096<false>(lane);
  const Coop4096 CO = make_coop_4096<false>(lane);
  if (lane == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  uint32_t twin = 0, ttw = 0;
  if (TM) {
    if (warp == 0) tmem_alloc<256>(&s_tmem_base);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    twin = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    ttw = twin + 128;
    if (warp < 4) {
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        cpx w8[8], t8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int b = 8 * c + i;
          w8[i] = *reinterpret_cast<const cpx*>(p.window + 2 * lane + 64 * b);
          t8[i] = p.tw1[b * 32 + lane];
        }
        tmem_st8(twin + 16 * c, w8);
        tmem_st8(ttw + 16 * c, t8);
      }
      tmem_wait_st();
    }
    tmem_fence_before_sync();
  }
  __syncthreads();
  if (TM) tmem_fence_after_sync();

  const int H = p.hop;
  const long long frames_per_chunk = p.chunk_steps;
  const long long chunks_per_chan = (p.frames + frames_per_chunk - 1) / frames_per_chunk;
  const long long total_chunks = chunks_per_chan * p.channels;
  const int ob1 = Geo4096::row1(lane);
  uint32_t parity = 0;

  for (;;) {
    unsigned long long q = 0;
    if (lane == 0) q = atomicAdd(p.counter, 1u);
    q = __shfl_sync(0xffffffffu, q, 0);
    if (static_cast<long long>(q) >= total_chunks) break;
    const long long c = q / chunks_per_chan;
    const long long first = (q % chunks_per_chan) * frames_per_chunk;
    const int nsteps = static_cast<int>(min(frames_per_chunk, p.frames - first));
    const SigView sig{p.sig + c * p.chan_stride, p.len, p.pad, p.pad_mode};
    cpx* spec = p.spec + (c * p.spec_chan_rows + first) * (C + 1);
    long long pos = first * H;

    int dirty = 0;
    __syncwarp();
    ring_fill(ring, mask, sig, pos, NFFT, bar, lane);

    for (int s = 0; s < nsteps; ++s) {
      cpx wb8[8];
      if (TM) tmem_ld8(twin, wb8);           // in flight across the barrier wait
      mbar_wait(bar, parity);
      parity ^= 1u;
      __syncwarp();

      cpx v[64];
      unsigned int wmin = 0xffffffffu;
      cpx nf = mk(0.f, 0.f);
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        cpx x8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int ridx = static_cast<int>((pos + 2 * lane + 64 * (8 * ch + i)) & mask);
          x8[i] = *reinterpret_cast<const cpx*>(ring + ridx);
        }
        if (TM) {
          tmem_wait_ld8(wb8);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) wb8[i] = *reinterpret_cast<const cpx*>(s_win + 2 * lane + 64 * (8 * ch + i));
        }
        cpx pw8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) pw8[i] = mul2(x8[i], wb8[i]);   // one rounding per sample, as frame[i]*window[i]
        if (TM && ch + 1 < 8) tmem_ld8(twin + 16 * (ch + 1), wb8);   // next chunk in flight during the ring reads
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int b = 8 * ch + i;
          const cpx x = x8[i];
          cpx pw = pw8[i];
          if (SCRUB == 1) pw = mk(scrub(pw.x), scrub(pw.y));
          if (SCRUB == 2 && b >= 48) {           // the newest hop (nfft == 4*hop): samples >= 3072
            const unsigned int u0 = __float_as_uint(x.x), u1 = __float_as_uint(x.y);
            wmin = min(wmin, min(u0 + u0 - 2u, u1 + u1 - 2u));
            nf = fma2(x, mk(0.f, 0.f), nf);
          }
          v[b] = pw;
        }
      }
      if (SCRUB == 2) {
        const bool suspect = (wmin < p.tiny2) || (nf.x != 0.0f) || (nf.y != 0.0f);
        if (__any_sync(0xffffffffu, suspect)) dirty = 5;
        if (dirty > 0 || s == 0) {             // s == 0: the whole ring is new
#pragma unroll
          for (int r = 0; r < 64; ++r) v[r] = mk(scrub(v[r].x), scrub(v[r].y));
          if (dirty > 0) --dirty;
        }
      }
      __syncwarp();
      if (s + 1 < nsteps) ring_fill(ring, mask, sig, pos + NFFT, H, bar, lane);

      fwd4096_pass1(v);
      if (TM) {                                // all 64 values are live here: one chunk of 8 factors at a time
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {
          cpx tb[8];
          tmem_ld8(ttw + 16 * ch, tb);
          tmem_wait_ld8(tb);
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (8 * ch + i > 0) v[8 * ch + i] = cmul(v[8 * ch + i], tb[i]);
        }
      } else {
        fwd4096_twiddle(v, lane, s_tw);
      }
      fwd4096_xwrite<0>(v, lane, ex);
      __syncwarp();
      fwd4096_xread<0>(v, lane, ex);
      __syncwarp();
      fwd4096_xwrite<1>(v, lane, ex);
      __syncwarp();
      fwd4096_xread<1>(v, lane, ex);
      __syncwarp();
      fwd4096_pass2(v);
      // lane 0's self-mirrored rows are shared out over the warp (warpfft4096.cuh "Coop4096")
      if (lane == 0) {
#pragma unroll
        for (int r = 0; r < 64; ++r) ex[r] = v[r];
      }
      __syncwarp();
      if (lane != 0) {
        cpx x1, x2;
        split_pair32<0, false>(ex[CO.zi], ex[CO.mi], CO.q, CO.qr, x1, x2);
        stg_stream(spec + CO.k, x1);
        stg_stream(spec + C - CO.k, x2);
        split4096_general<false>(v, SL.q, SL.qr);
#pragma unroll
        for (int oa = 0; oa < 32; ++oa) stg_stream(spec + lane + 64 * oa, v[oa]);
#pragma unroll
        for (int oa = 0; oa < 32; ++oa) stg_stream(spec + ob1 + 64 * oa, v[32 + oa]);
      } else {
        stg_stream(spec, mk(v[0].x + v[0].y, 0.0f));          // exact DC and Nyquist (kiss_fftr.c:93-99)
        stg_stream(spec + C, mk(v[0].x - v[0].y, 0.0f));
        stg_stream(spec + C / 2, mul2(v[16], mk(1.f, -1.f))); // X[1024] = conj(Z[1024])
      }
      pos += H;
      spec += C + 1;
    }
  }
  if (TM) {
    tmem_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<256>(s_tmem_base);
  }
}

static cudaError_t launch_k1_4096(const K1Params& p, cudaStream_t st) {
  static const bool tm = getenv("CRLOT_K1_NO_TMEM") == nullptr;
  const int smem = (tm ? 0 : 4096 * 4 + 2048 * 8) + kK1Warps4096 * (4096 * 4 + Geo4096::FX_ROUND_CPX * 8 + 16);
  const int mode = !p.scrub ? 0 : (p.scrub == 2 && 4096 == 4 * p.hop) ? 2 : 1;
  auto kern = tm ? (mode == 0 ? stft_fwd_kernel_4096<0, true> : mode == 1 ? stft_fwd_kernel_4096<1, true> : stft_fwd_kernel_4096<2, true>)
                 : (mode == 0 ? stft_fwd_kernel_4096<0, false> : mode == 1 ? stft_fwd_kernel_4096<1, false> : stft_fwd_kernel_4096<2, false>);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return e;
  const long long total_chunks = ((p.frames + p.chunk_steps - 1) / p.chunk_steps) * p.channels;
  const long long want_blocks = (total_chunks + kK1Warps4096 - 1) / kK1Warps4096;
  const int grid = static_cast<int>(want_blocks < device_sm_count() ? want_blocks : device_sm_count());
  kern<<<grid > 0 ? grid : 1, kK1Warps4096 * 32, smem, st>>>(p);
  return cudaGetLastError();
}

template <int NFFT>
cudaError_t launch_k1(const K1Params& p, cudaStream_t st) {
  using X = GeoX<NFFT>;
  const int S = (X::G - 1) * p.hop + NFFT;
  const int rcap = next_pow2(S);
  // CRLOT_K1_NO_TMEM=1 keeps the window / twiddle tables in shared memory (A/B timing and the bit-identity test)
  static const bool tm = getenv("CRLOT_K1_NO_TMEM") == nullptr;
  const int smem = tm ? K1Smem<NFFT, true>::total(rcap, kK1Warps) : K1Smem<NFFT, false>::total(rcap, kK1Warps);
  if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
  // screening mode needs nfft == 4*hop and a finite threshold; otherwise the per-sample rule
  const int mode = !p.scrub ? 0 : (p.scrub == 2 && NFFT == 4 * p.hop) ? 2 : 1;
  auto kern = tm ? (mode == 0 ? stft_fwd_kernel<NFFT, 0, true> : mode == 1 ? stft_fwd_kernel<NFFT, 1, true> : stft_fwd_kernel<NFFT, 2, true>)
                 : (mode == 0 ? stft_fwd_kernel<NFFT, 0, false> : mode == 1 ? stft_fwd_kernel<NFFT, 1, false> : stft_fwd_kernel<NFFT, 2, false>);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return e;
  const long long frames_per_chunk = static_cast<long long>(p.chunk_steps) * X::G;
  const long long total_chunks = ((p.frames + frames_per_chunk - 1) / frames_per_chunk) * p.channels;
  const long long want_blocks = (total_chunks + kK1Warps - 1) / kK1Warps;
  const int grid = static_cast<int>(want_blocks < device_sm_count() ? want_blocks : device_sm_count());
  kern<<<grid > 0 ? grid : 1, kK1Warps * 32, smem, st>>>(p, rcap);
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_stft_fwd(const K1Params& p, cudaStream_t st) {
  if (p.frames <= 0 || p.channels <= 0) return cudaSuccess;
  if (p.generic) return launch_stft_fwd_generic(p, st);
  switch (p.nfft) {
    case 256: return launch_k1<256>(p, st);
    case 512: return launch_k1<512>(p, st);
    case 1024:
      if (stft_fwd_2048tm_supported(p) && getenv("CRLOT_K1_NO_TM16") == nullptr) return launch_stft_fwd_2048tm(p, st);   // hop 256: as below
      return launch_k1<1024>(p, st);
    case 2048:
      // hop 512, Framer addressing: the 16-warp kernel that carries the raw samples in tensor memory (k1_2048tm.cu; same
      // bits).  CRLOT_K1_NO_TM16=1 keeps the kernel below for the A/B (read per call).
      if (stft_fwd_2048tm_supported(p) && getenv("CRLOT_K1_NO_TM16") == nullptr) return launch_stft_fwd_2048tm(p, st);
      return launch_k1<2048>(p, st);
    case 4096: {
      // CRLOT_K1_4096_V1=1: the first version (one 64-register transform per frame, 8 warps) for A/B timing
      static const bool v1 = getenv("CRLOT_K1_4096_V1") != nullptr;
      return v1 ? launch_k1_4096(p, st) : launch_stft_fwd_4096v2(p, st);
    }
    default: return cudaErrorNotSupported;
  }
}

}  // namespace crlot
