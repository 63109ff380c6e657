// k23_4096.cu -- kernels 2+3 fused for N = 4096: inverse real FFT * 1/N + clean-up + synthesis
// window, overlap-added in a per-warp shared-memory accumulator ring and normalised on the way out.
//
// Same contract and chunk logic as k23_istft_ola.cu (read its header first) on the
// 64-registers-per-lane machinery of warpfft4096.cuh: one warp = one frame per step, the lane loads
// the spectrum columns a0 = lane and a1 = 64 - lane straight from global memory (the next row is
// pulled into L2 meanwhile), lane 0's self-mirrored columns are split by the other lanes through the
// exchange buffer (Coop4096), and after pass 2 register oa of a lane holds the samples
// i = 2*(lane + 32*oa), i + 1: registers [0, 64/R) are the hop that leaves, the last 64/R are first
// touches of their ring cells, the rest read-modify-write.  Same roundings in the same order as
// istft_inv_kernel_4096 -> ola_gather_kernel, so the output is BIT-IDENTICAL to that path
// (tests/test_gpu_fused.py).
#include "common.cuh"
#include "kernels.h"
#include "warpfft4096.cuh"

namespace crlot {

namespace {

constexpr int kW = 8;                                   // warps per CTA
constexpr int kNF = 4096, kC = 2048, kACAP = 4096;
constexpr int kPerWarp = kACAP * 4 + Geo4096::FX_ROUND_CPX * 8;

__device__ __forceinline__ void prefetch_l2_4096(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// RDIV = 4096 / hop when that is 1, 2, 4 or 8 (hop known at compile time), else 0: every frame takes
// the rolled path.  Window table in shared memory: ws[i] = (2/N) * w[i] (exact power-of-two scale).
template <int RDIV>
__global__ void __launch_bounds__(kW * 32, 1) istft_ola_kernel_4096(const K23Params p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* s_win = reinterpret_cast<float*>(smem_raw);
  cpx* s_twi = reinterpret_cast<cpx*>(s_win + kNF);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* wb = reinterpret_cast<unsigned char*>(s_twi + kC) + warp * kPerWarp;
  float* ring = reinterpret_cast<float*>(wb);
  cpx* ex = reinterpret_cast<cpx*>(ring + kACAP);
  constexpr int mask = kACAP - 1;
  const float scale = 2.0f / static_cast<float>(kNF);   // the split hands over halved values

  for (int i = threadIdx.x; i < kNF; i += blockDim.x) s_win[i] = p.window[i] * scale;
  for (int i = threadIdx.x; i < kC; i += blockDim.x) s_twi[i] = p.pre[i];
  const SplitLane4096 SL = make_split_4096<true>(lane);
  const Coop4096 CO = make_coop_4096<true>(lane);
  __syncthreads();

  const int H = RDIV > 0 ? kNF / (RDIV > 0 ? RDIV : 1) : p.hop;
  const int R = (kNF + H - 1) / H;
  const long long F = p.frames_per_chan;
  const long long nblocks = p.m_end - p.m_begin;
  const long long chunks_per_chan = (nblocks + p.chunk_blocks - 1) / p.chunk_blocks;
  const long long total_chunks = chunks_per_chan * p.channels;
  const long long u_cov = F > 0 ? (F - 1) * H + kNF : 0;
  const int ring_len = static_cast<int>(p.ring_len);
  const float gain = p.gain, eps = p.eps;
  const cpx gain2 = mk(gain, gain);
  const unsigned int tiny2 = 2u * __float_as_uint(1e-30f * (0.5f * kNF)) - 2u;
  const float unscale = 0.5f * kNF;
  const int a1 = Geo4096::row1(lane);

  for (;;) {
    unsigned long long q = 0;
    if (lane == 0) q = atomicAdd(p.counter, 1u);
    q = __shfl_sync(0xffffffffu, q, 0);
    if (static_cast<long long>(q) >= total_chunks) break;
    const long long c = q / chunks_per_chan;
    const long long m0 = p.m_begin + (q % chunks_per_chan) * p.chunk_blocks;
    const long long m1 = min(m0 + p.chunk_blocks, p.m_end);
    const cpx* spec = p.spec + c * p.spec_chan_rows * (kC + 1);
    float* out = p.out + c * p.out_chan_stride;
    const float* sgain = p.sgain ? p.sgain + c * p.spec_chan_rows * p.sgain_row_stride : nullptr;
    const long long k0 = max(0ll, m0 - R + 1);
    const long long k1 = min(m1, F);

    __syncwarp();
    for (int i = lane * 4; i < kACAP; i += 128) *reinterpret_cast<float4*>(ring + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();

    int e_base = 0;
    {
      long long e = (p.t_offset + p.frame_t0 + k0 * H) % ring_len;
      if (e < 0) e += ring_len;
      e_base = static_cast<int>(e);
    }
    const int e_hop = H % ring_len;

    for (long long k = k0; k < k1; ++k) {
      const cpx* row = spec + k * (kC + 1);
      cpx v[64];
#pragma unroll
      for (int b = 0; b < 32; ++b) v[b] = ldg_stream(row + lane + 64 * b);
#pragma unroll
      for (int b = 0; b < 32; ++b) v[32 + b] = ldg_stream(row + a1 + 64 * b);
      cpx nyq = mk(0.f, 0.f);
      if (lane == 0) nyq = ldg_stream(row + kC);
      if (k + 1 < k1) {                      // pull the next row into L2 while this one is transformed
        const char* nxt = reinterpret_cast<const char*>(row + (kC + 1));
#pragma unroll
        for (int o = 0; o < 5; ++o)
          if (o * 4096 + lane * 128 < (kC + 1) * 8) prefetch_l2_4096(nxt + o * 4096 + lane * 128);
      }
      if (sgain != nullptr) {                // spectral-domain hook: per-bin real gain as the row is loaded
        const float* gr = sgain + k * p.sgain_row_stride;
#pragma unroll
        for (int b = 0; b < 32; ++b) {
          v[b] = mul2(v[b], splat(__ldg(gr + lane + 64 * b)));
          v[32 + b] = mul2(v[32 + b], splat(__ldg(gr + a1 + 64 * b)));
        }
        if (lane == 0) nyq = mul2(nyq, splat(__ldg(gr + kC)));
      }
      // lane 0's self-mirrored columns: shared out over the warp through the exchange buffer
      if (lane == 0) {
#pragma unroll
        for (int r = 0; r < 64; ++r) ex[r] = v[r];
      }
      __syncwarp();
      if (lane != 0) {
        cpx x1, x2;
        split_pair32<0, true>(ex[CO.zi], ex[CO.mi], CO.q, CO.qr, x1, x2);
        ex[CO.zi] = x1;
        ex[CO.mi] = x2;
        split4096_general<true>(v, SL.q, SL.qr);
      }
      __syncwarp();
      if (lane == 0) {
        const cpx x0 = v[0], xm = v[16];
#pragma unroll
        for (int r = 0; r < 64; ++r) v[r] = ex[r];
        v[0] = mk(0.5f * (x0.x + nyq.x), 0.5f * (x0.x - nyq.x));   // only the real parts (kiss_fftr.c:131-132)
        v[16] = mul2(xm, mk(1.f, -1.f));
      }
      __syncwarp();
      inv4096_pass1(v);
      inv4096_twiddle(v, lane, s_twi);
      inv4096_xwrite<0>(v, lane, ex);
      __syncwarp();
      inv4096_xread<0>(v, lane, ex);
      __syncwarp();
      inv4096_xwrite<1>(v, lane, ex);
      __syncwarp();
      inv4096_xread<1>(v, lane, ex);
      __syncwarp();
      inv4096_pass2(v);                       // v[oa] = samples 2*(lane + 32*oa), +1 (raw, times N/2)

      const long long u0 = k * H;
      const int r0 = static_cast<int>(u0 & mask);
      const long long j_first = p.frame_t0 + u0;
      bool fast = RDIV > 0 && p.den_rcp != nullptr && k >= m0 && j_first >= 0 && j_first + H <= p.n_out;
      if (fast) {
        unsigned int wm[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
        cpx nfs[4] = {mk(0.f, 0.f), mk(0.f, 0.f), mk(0.f, 0.f), mk(0.f, 0.f)};
#pragma unroll
        for (int r = 0; r < 64; ++r) {
          const unsigned int b0 = __float_as_uint(v[r].x), b1 = __float_as_uint(v[r].y);
          wm[r & 3] = min(wm[r & 3], min(b0 + b0 - 2u, b1 + b1 - 2u));
          nfs[r & 3] = fma2(v[r], mk(0.f, 0.f), nfs[r & 3]);
        }
        const unsigned int wmin = min(min(wm[0], wm[1]), min(wm[2], wm[3]));
        const cpx nf = (nfs[0] + nfs[1]) + (nfs[2] + nfs[3]);
        fast = !__any_sync(0xffffffffu, (wmin < tiny2) || (nf.x != 0.0f) || (nf.y != 0.0f));
      }

      if (fast) {
        constexpr int AE = RDIV > 0 ? 64 / RDIV : 64;   // registers per hop
#pragma unroll
        for (int oa = 0; oa < 64; ++oa) v[oa] = mul2(v[oa], *reinterpret_cast<const cpx*>(s_win + 2 * (lane + 32 * oa)));
        const int rb = r0 + 2 * lane;
#pragma unroll
        for (int oa = 0; oa < 64; ++oa) {               // all ring reads of the frame, back to back
          cpx a = mk(0.f, 0.f);
          if (oa < 64 - AE) a = *reinterpret_cast<const cpx*>(ring + ((rb + 64 * oa) & mask));
          v[oa] = fma2(v[oa], gain2, a);
        }
#pragma unroll
        for (int oa = AE; oa < 64; ++oa) *reinterpret_cast<cpx*>(ring + ((rb + 64 * oa) & mask)) = v[oa];
        // the hop that leaves: {denominator, reciprocal} pairs, requested once the other registers are free
        float* dst = out + j_first + 2 * lane;
        constexpr int EB = AE < 16 ? AE : 16;          // 16 registers at a time (register budget)
#pragma unroll
        for (int o0 = 0; o0 < AE; o0 += EB) {
          float4 nm[EB];
#pragma unroll
          for (int q = 0; q < EB; ++q) {
            int e = e_base + 2 * (lane + 32 * (o0 + q));
            if (e >= ring_len) e -= ring_len;
            nm[q] = __ldg(reinterpret_cast<const float4*>(p.den_rcp + e));
          }
          bool all_ok = true;
#pragma unroll
          for (int q = 0; q < EB; ++q) {
            const cpx a = v[o0 + q];
            const float4 t = nm[q];
            all_ok = all_ok & div_tail_ok(a.x) & div_tail_ok(a.y);   // no short-circuit: one predicate chain, no branches
            stg_stream(reinterpret_cast<cpx*>(dst + 64 * (o0 + q)), mk(div_tail(a.x, t.x, t.y), div_tail(a.y, t.z, t.w)));
          }
          if (!all_ok) {                                // operands outside the tail's range: the full division
#pragma unroll
            for (int q = 0; q < EB; ++q) {
              const cpx a = v[o0 + q];
              const float4 t = nm[q];
              stg_stream(reinterpret_cast<cpx*>(dst + 64 * (o0 + q)), mk(__fdiv_rn(a.x, t.x), __fdiv_rn(a.y, t.z)));
            }
          }
        }
      } else {
        // rolled path, straight from the registers: the reference formula sample by sample
        const int thr = k < m0 ? static_cast<int>((m0 - k) * H) : 0;   // warm-up: keep u >= m0*H only
        const bool emit = k >= m0;
        // (registers cannot be indexed by a loop variable: each lane parks its own values in the
        // exchange buffer, 32 registers at a time, and the rolled loop reads them back)
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          __syncwarp();
#pragma unroll
          for (int q = 0; q < 32; ++q) ex[q * 32 + lane] = v[32 * half + q];
          __syncwarp();
          for (int q = 0; q < 32; ++q) {
            const int i = 2 * (lane + 32 * (32 * half + q));
            if (i < thr) continue;
            const cpx raw = ex[q * 32 + lane];
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
            if (i < kNF - H) a = *reinterpret_cast<const float2*>(cell);     // else: first touch
            a.x = __fmaf_rn(y0, gain, a.x);
            a.y = __fmaf_rn(y1, gain, a.y);
            if (emit && i < H) {
              const long long jo = j_first + i;
              if (jo >= 0 && jo < p.n_out) {
                int e = e_base + i;
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
        __syncwarp();
      }
      e_base = (e_base + e_hop) % ring_len;
    }

    // blocks of this chunk that have no frame of their own: before frame 0 (zeros) and the
    // tail after the last frame (what is left in the ring)
    __syncwarp();
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

}  // namespace

cudaError_t launch_istft_ola_4096(const K23Params& p, cudaStream_t st) {
  const int smem = kNF * 4 + kC * 8 + kW * kPerWarp;
  const int rdiv = (kNF % p.hop == 0) ? kNF / p.hop : 0;
  void (*kern)(const K23Params) = nullptr;
  switch (rdiv) {
    case 1: kern = istft_ola_kernel_4096<1>; break;
    case 2: kern = istft_ola_kernel_4096<2>; break;
    case 4: kern = istft_ola_kernel_4096<4>; break;
    case 8: kern = istft_ola_kernel_4096<8>; break;
    default: kern = istft_ola_kernel_4096<0>; break;
  }
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return e;
  const long long nblocks = p.m_end - p.m_begin;
  const long long total_chunks = ((nblocks + p.chunk_blocks - 1) / p.chunk_blocks) * p.channels;
  const long long want_blocks = (total_chunks + kW - 1) / kW;
  const int grid = static_cast<int>(want_blocks < device_sm_count() ? want_blocks : device_sm_count());
  kern<<<grid > 0 ? grid : 1, kW * 32, smem, st>>>(p);
  return cudaGetLastError();
}

}  // namespace crlot
