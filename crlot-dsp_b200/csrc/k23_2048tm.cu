// k23_2048tm.cu -- kernels 2+3 fused for N = 2048, hop = 512 with the ACCUMULATOR RING IN TENSOR MEMORY: 16 warps per SM.
//
// Same contract, chunk protocol and arithmetic as istft_ola_kernel<2048, 4> in k23_istft_ola.cu (read its header first):
// replaces KissFftPlan::inverse -> kiss_fftri, axpy_windowed and OLAAccumulator::produce / normalize_and_clear
// (reference: dsp/fft/backends/kissfft_adapter.cc:124-168, dsp/ola/kernels.cc:18-36, dsp/ola/OLAAccumulator.cc:75-221);
// output BIT-IDENTICAL to that kernel and to K2 -> K3 (tests/test_gpu_fused.py).  That kernel keeps a 2048-float
// accumulator ring per warp in shared memory next to the exchange buffer (16.4 KB per warp: twelve warps, 168 registers,
// half its own roofline).  Here:
//
// * after pass 2 register oa of a lane holds samples 2*lane + 64*oa, +1 and a hop is 8 registers, so a ring cell is
//   only ever touched by ONE lane: the ring is 64 tensor-memory columns per warp, register row (8k + oa) mod 32 of
//   frame k in chunk (k + oa/8) mod 4 -- three tcgen05.ld and three tcgen05.st per frame instead of 24 + 24 shared-memory
//   accesses, compile-time registers, a runtime column offset;
// * the next spectrum row is pulled into the exchange buffer by one bulk async copy (UBLKCP on the warp's mbarrier) as
//   soon as the transposition has been read back (rows of a chunk are consecutive in memory), so a frame starts from
//   shared memory instead of waiting on 33 global loads per lane;
// * what the shared-memory kernel does in a rolled, sample-by-sample path is register-granular here: warm-up frames skip
//   whole chunks of eight registers (the part of the frame before the chunk's first hop block is a multiple of the hop),
//   frames that fail the clean-up screen take the two-step product fl(scrub(fl(v*2/N)) * w) on all registers, ragged
//   ends predicate the stores;
// * shared memory per warp: the exchange buffer; 16 warps at 128 registers.
//   TMEM columns: [0,64) inter-pass twiddles, [64,128) synthesis window times 2/N, [128,384) four warps x 64 accumulators.
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "warpfft.cuh"

namespace crlot {

namespace {

constexpr int kW = 16;
using X = GeoX<2048>;
constexpr int kPerWarp = X::XCPX * 8;
constexpr int kSmem = kW * kPerWarp + kW * 8;

__global__ void __launch_bounds__(kW * 32, 1) istft_ola_kernel_2048tm(const K23Params p) {
  constexpr int NFFT = 2048, C = 1024, A = 32, B = 32, H = 512;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint32_t s_tmem_base;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  cpx* ex = reinterpret_cast<cpx*>(smem_raw + warp * kPerWarp);
  const uint32_t bar = smem_u32(smem_raw + kW * kPerWarp) + 8 * warp;
  const float scale = 2.0f / static_cast<float>(NFFT);   // inv_split hands over halved values
  const float unscale = 0.5f * NFFT;                     // ws * N/2 == w exactly

  if (lane == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&s_tmem_base);
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();
  const uint32_t tmb = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  const uint32_t ttw = tmb, twin = tmb + 64, tring = tmb + 128 + 64 * (warp >> 2);
  if (warp < 4) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      cpx t8[8], w8[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        t8[i] = p.tw1[(8 * c + i) * A + lane];
        const float2 w = *reinterpret_cast<const float2*>(p.window + 2 * (lane + B * (8 * c + i)));
        w8[i] = mk(w.x * scale, w.y * scale);
      }
      tmem_st8(ttw + 16 * c, t8);
      tmem_st8(twin + 16 * c, w8);
    }
    tmem_wait_st();
  }
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();

  const SplitLane<X::M1> SL = make_split_inv<X>(lane);
  const int lp = mirror_lane<A>(lane);
  const long long F = p.frames_per_chan;
  // Chunks are handed out by an atomic counter; a warp that draws its last chunk late leaves the SM waiting at the end
  // (6 % of the stall samples with uniform chunks).  So the first three quarters of every channel go in big chunks (few
  // warm-up frames) and the rest, drawn after all the big ones, in chunks a quarter of the size.
  const long long nblocks = p.m_end - p.m_begin;
  const long long cb = p.chunk_blocks, cbs = cb >= 32 ? cb / 4 : cb;
  const long long big_per_chan = (nblocks * 3 / 4) / cb;
  const long long small_per_chan = (nblocks - big_per_chan * cb + cbs - 1) / cbs;
  const long long total_big = big_per_chan * p.channels;
  const long long total_chunks = total_big + small_per_chan * p.channels;
  const long long u_cov = F > 0 ? (F - 1) * H + NFFT : 0;          // u >= u_cov: no frame contributes
  const int ring_len = static_cast<int>(p.ring_len);
  const float gain = p.gain, eps = p.eps;
  const cpx gain2 = mk(gain, gain);
  const unsigned int tiny2 = 2u * __float_as_uint(1e-30f * (0.5f * NFFT)) - 2u;
  uint32_t parity = 0;

  // spectrum row `row` into the exchange buffer: one copy from the 16-byte boundary at or before it, an odd last bin by
  // a plain load (rows of 1025 bins are only 8-byte aligned)
  auto stage_row = [&](const cpx* row) {
    if (elect_one()) {
      const int off = static_cast<int>((reinterpret_cast<uintptr_t>(row) >> 3) & 1);
      const int total = C + 1 + off;
      const uint32_t nb = static_cast<uint32_t>(total & ~1) * 8u;
      mbar_arrive_expect_tx(bar, nb);
      bulk_g2s(smem_u32(ex), row - off, nb, bar);
      if (total & 1) ex[total - 1] = ldg_stream(row - off + total - 1);
    }
  };

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
    const cpx* spec = p.spec + c * p.spec_chan_rows * (C + 1);
    float* out = p.out + c * p.out_chan_stride;
    const float* sgain = p.sgain ? p.sgain + c * p.spec_chan_rows * p.sgain_row_stride : nullptr;
    const long long k0 = max(0ll, m0 - 3);
    const long long k1 = min(m1, F);

    {                                                    // a cleared accumulator
      cpx z8[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) z8[i] = mk(0.f, 0.f);
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) tmem_st8(tring + 16 * ch, z8);
      tmem_wait_st();
    }
    int e_base = 0;                                      // envelope phase of sample 0 of frame k, advanced per frame
    {
      long long e = (p.t_offset + p.frame_t0 + k0 * H) % ring_len;
      if (e < 0) e += ring_len;
      e_base = static_cast<int>(e);
    }
    const int e_hop = H % ring_len;
    __syncwarp();                                        // the previous chunk's last exchange reads are complete
    if (k0 < k1) stage_row(spec + k0 * (C + 1));

    for (long long k = k0; k < k1; ++k) {
      // ---- the spectrum row out of the exchange buffer (register b of lane a = bin a + 32*b) -----------------------------
      cpx v[32];
      cpx nyq[1];
      mbar_wait(bar, parity);
      parity ^= 1u;
      __syncwarp();                                      // the elected lane's plain load of the last bin
      {
        const cpx* row = spec + k * (C + 1);
        const cpx* st = ex + ((reinterpret_cast<uintptr_t>(row) >> 3) & 1);
#pragma unroll
        for (int b = 0; b < B; ++b) v[b] = st[lane + A * b];
        nyq[0] = lane == 0 ? st[C] : mk(0.f, 0.f);
      }
      if (sgain != nullptr) {                            // spectral-domain hook: per-bin real gain as the row is loaded
        const float* gr = sgain + k * p.sgain_row_stride;
#pragma unroll
        for (int b = 0; b < B; ++b) v[b] = mul2(v[b], splat(__ldg(gr + lane + A * b)));
        if (lane == 0) nyq[0] = mul2(nyq[0], splat(__ldg(gr + C)));
      }
      {
        cpx mid[1];
        auto mirror = [&](int r) { return mk(__shfl_sync(0xffffffffu, v[r].x, lp), __shfl_sync(0xffffffffu, v[r].y, lp)); };
        inv_split_pairs<X>(v, lane, SL, mirror, nyq, mid);
        inv_split_finish<X>(v, lane, mirror, mid);
      }
      {
        cpx tb[2][8];
        tmem_ld8(ttw, tb[0]);
        pass1<X, true>(v);
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          tmem_wait_ld8(tb[ch & 1]);
          if (ch + 1 < 4) tmem_ld8(ttw + 16 * (ch + 1), tb[(ch + 1) & 1]);
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (8 * ch + i > 0) v[8 * ch + i] = cmulc(v[8 * ch + i], tb[ch & 1][i]);
        }
      }
      __syncwarp();                                      // every lane has read its row out of the buffer
      xchg_write<X>(v, lane, ex);
      __syncwarp();
      xchg_read<X>(v, lane, ex);
      fence_proxy_async_smem();                          // this lane's generic-proxy writes to the buffer before the copy's
      __syncwarp();
      if (k + 1 < k1) stage_row(spec + (k + 1) * (C + 1));
      if (k + 2 < k1) {                                  // and the row after it into L2: the copy then takes an L2 round trip
        const char* nx = reinterpret_cast<const char*>(spec + (k + 2) * (C + 1)) + lane * 128;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(nx));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + 4096));
        if (lane == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + 8192));
      }
      pass2<X, true>(v);                                 // v[oa] = samples 2*lane + 64*oa, +1 (raw, times N/2)

      // ---- clean-up screen on the raw registers (non-finite, or 0 < |v * 2/N| < 1e-30) ----------------------------------
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
        const unsigned int wmin = min(min(wm[0], wm[1]), min(wm[2], wm[3]));
        const cpx nf = (nfs[0] + nfs[1]) + (nfs[2] + nfs[3]);
        clean = !__any_sync(0xffffffffu, (wmin < tiny2) || (nf.x != 0.0f) || (nf.y != 0.0f));
      }
      // ---- synthesis window: fl(v * ws), or the reference's two roundings with its clean-up rule in between --------------
      {
        cpx wb[2][8];
        tmem_ld8(twin, wb[0]);
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          tmem_wait_ld8(wb[ch & 1]);
          if (ch + 1 < 4) tmem_ld8(twin + 16 * (ch + 1), wb[(ch + 1) & 1]);
          if (clean) {
#pragma unroll
            for (int i = 0; i < 8; ++i) v[8 * ch + i] = mul2(v[8 * ch + i], wb[ch & 1][i]);
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              cpx y = mul2(v[8 * ch + i], splat(scale));
              if (p.scrub) y = mk(scrub(y.x), scrub(y.y));
              v[8 * ch + i] = mul2(y, mul2(wb[ch & 1][i], splat(unscale)));
            }
          }
        }
      }
      // ---- overlap-add in the tensor-memory ring, eight registers at a time; d = chunks of the frame before the block ----
      const int d = k < m0 ? static_cast<int>(m0 - k) : 0;  // warm-up frame: keep the samples at or beyond block m0 only
      const int kc = static_cast<int>(k & 3);
#pragma unroll
      for (int ch = 1; ch < 4; ++ch) {
        if (ch >= d) {
          cpx a8[8];
          if (ch < 3) {
            tmem_ld8(tring + 16 * ((kc + ch) & 3), a8);
            tmem_wait_ld8(a8);
          } else {                                       // the newest hop: first touch of its cells
#pragma unroll
            for (int i = 0; i < 8; ++i) a8[i] = mk(0.f, 0.f);
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) a8[i] = fma2(v[8 * ch + i], gain2, a8[i]);
          tmem_st8(tring + 16 * ((kc + ch) & 3), a8);
        }
      }
      if (d == 0) {                                      // the hop that leaves: complete, divide, store
        cpx a8[8];
        tmem_ld8(tring + 16 * kc, a8);
        const long long j_first = p.frame_t0 + k * H;    // output index of sample 0 of the frame
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
          if (!all_ok) {                                 // operands outside the tail's range: the full division
#pragma unroll
            for (int i = 0; i < 8; ++i)
              stg_stream(reinterpret_cast<cpx*>(dst + 64 * i), mk(__fdiv_rn(a8[i].x, nm[i].x), __fdiv_rn(a8[i].y, nm[i].z)));
          }
        } else {                                         // ragged ends, or no reciprocal table: the reference formula
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

    // blocks of this chunk that have no frame of their own: before frame 0 (zeros) and the tail after the last frame
    // (what is left in the ring: block m sits in ring chunk m mod 4)
    for (long long m = m0; m < m1; ++m) {
      if (m >= 0 && m < F) continue;
      const long long u0 = m * H;
      cpx a8[8];
      if (m >= 0) {
        tmem_ld8(tring + 16 * static_cast<int>(m & 3), a8);
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

bool istft_ola_2048tm_supported(const K23Params& p) {
  return p.nfft == 2048 && p.hop == 512 && p.sig == nullptr && (reinterpret_cast<uintptr_t>(p.spec) & 15) == 0;
}

cudaError_t launch_istft_ola_2048tm(const K23Params& p, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(istft_ola_kernel_2048tm, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return e;
  const long long nblocks = p.m_end - p.m_begin;
  const long long cb = p.chunk_blocks, cbs = cb >= 32 ? cb / 4 : cb;           // as in the kernel
  const long long big_per_chan = (nblocks * 3 / 4) / cb;
  const long long total_chunks = (big_per_chan + (nblocks - big_per_chan * cb + cbs - 1) / cbs) * p.channels;
  const long long want_blocks = (total_chunks + kW - 1) / kW;
  const int grid = static_cast<int>(want_blocks < device_sm_count() ? want_blocks : device_sm_count());
  istft_ola_kernel_2048tm<<<grid > 0 ? grid : 1, kW * 32, kSmem, st>>>(p);
  return cudaGetLastError();
}

}  // namespace crlot
