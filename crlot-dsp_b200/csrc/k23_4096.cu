// k23_4096.cu -- kernels 2+3 fused for N = 4096: inverse real FFT * 1/N + clean-up + synthesis
// window, overlap-added in a per-warp shared-memory accumulator ring and normalised on the way out.
//
// Same contract and chunk logic as k23_istft_ola.cu (read its header first).  The transform is kernel 2's for this
// size (k2_4096v2.cu, arithmetic in warpfft4096v2.cuh): quads {k, 1024-k, 1024+k, 2048-k} -> e', o', two 1024-point
// inverse sub-transforms -- here with BOTH planes in registers (64 complex per lane, 8 warps per CTA, the accumulator
// ring takes the shared memory twelve warps would need).  The lane loads X[k] and X[k+1024], k = lane + 32*b, straight
// from global memory (the next row is pulled into L2 meanwhile), partners X[1024-k], X[2048-k] come from the mirror lane
// by shuffle, lane 0's self-paired bins 32*b are computed by lane b through the exchange buffer.  After the transforms
// register oa of a lane holds the four samples i = 4*(lane + 32*oa) .. i+3 (e[oa], o[oa]): registers [0, 32/R) are the
// hop that leaves, the last 32/R are first touches of their ring cells, the rest read-modify-write, all as float4.
// Same roundings in the same order as istft_inv_kernel_4096v2 -> ola_gather_kernel, so the output is BIT-IDENTICAL to
// that path (tests/test_gpu_fused.py).
#include "common.cuh"
#include "kernels.h"
#include "warpfft4096v2.cuh"

namespace crlot {

namespace {

constexpr int kW = 8;                                   // warps per CTA
constexpr int kNF = 4096, kC = 2048, kACAP = 4096;
constexpr int kPerWarp = kACAP * 4 + XV4096::XCPX * 8;

__device__ __forceinline__ void prefetch_l2_4096(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// RDIV = 4096 / hop when that is 1, 2, 4 or 8 (hop known at compile time), else 0: every frame takes
// the rolled path.  Window table in shared memory: ws[i] = (2/N) * w[i] (exact power-of-two scale).
template <int RDIV>
__global__ void __launch_bounds__(kW * 32, 1) istft_ola_kernel_4096(const K23Params p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint32_t s_tmem_base;
  float* s_win = reinterpret_cast<float*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* wb = reinterpret_cast<unsigned char*>(s_win + kNF) + warp * kPerWarp;
  float* ring = reinterpret_cast<float*>(wb);
  cpx* ex = reinterpret_cast<cpx*>(ring + kACAP);
  constexpr int mask = kACAP - 1;
  const float scale = 2.0f / static_cast<float>(kNF);   // the quads hand over halved values

  for (int i = threadIdx.x; i < kNF; i += blockDim.x) s_win[i] = p.window[i] * scale;
  // constants in tensor memory as in k2_4096v2.cu: inter-pass twiddles (64 columns), split factor + combine twiddle
  // per register in the order the quads consume them (128 columns); warps w and w+4 share a lane quarter
  if (warp == 0) tmem_alloc<256>(&s_tmem_base);
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();
  const uint32_t ttw = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16), tqc = ttw + 64;
  if (warp < 4) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      cpx a8[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) a8[i] = p.sa[(8 * c + i) * 32 + lane];
      tmem_st8(ttw + 16 * c, a8);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      cpx c8[8];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int k = lane + 32 * v2_chunk_reg(j, t);
        c8[t] = p.sa[2048 + k];
        c8[4 + t] = p.sa[1024 + k];
      }
      tmem_st8(tqc + 16 * j, c8);
    }
    tmem_wait_st();
  }
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();
  const cpx qc = p.sa[2048 + 32 * lane], tcc = p.sa[1024 + 32 * lane];   // lane 0's quad k = 32*lane, computed by this lane
  const int lp = (32 - lane) & 31;

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
      cpx L[64];                               // L[b] = X[lane + 32 b], L[32 + b] = X[1024 + lane + 32 b]
#pragma unroll
      for (int b = 0; b < 32; ++b) L[b] = ldg_stream(row + lane + 32 * b);
#pragma unroll
      for (int b = 0; b < 32; ++b) L[32 + b] = ldg_stream(row + 1024 + lane + 32 * b);
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
          L[b] = mul2(L[b], splat(__ldg(gr + lane + 32 * b)));
          L[32 + b] = mul2(L[32 + b], splat(__ldg(gr + 1024 + lane + 32 * b)));
        }
        if (lane == 0) nyq = mul2(nyq, splat(__ldg(gr + kC)));
      }
      // lane 0's bins 32*b pair with its own registers: lane b computes that quad through the exchange buffer
      if (lane == 0) {
#pragma unroll
        for (int r = 0; r < 64; ++r) ex[r] = L[r];
      }
      __syncwarp();
      {
        cpx ce, co;
        if (lane != 0) inv4096v2_quad(ex[lane], ex[32 - lane], ex[32 + lane], ex[64 - lane], qc, tcc, ce, co);
        else inv4096v2_quad0(L[0], L[32], nyq, ce, co);   // bin 0: real parts of X[0], X[2048] only (kiss_fftr.c:131-132)
        __syncwarp();
        ex[lane] = ce;
        ex[32 + lane] = co;
      }
      __syncwarp();
      cpx v[64];                               // v[b] = e'[lane + 32 b], v[32 + b] = o'[lane + 32 b]
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        cpx k8[8];
        tmem_ld8(tqc + 16 * j, k8);
        tmem_wait_ld8(k8);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int pp = 2 * j + i, r = 31 - pp;   // own register pp pairs with the mirror lane's register r and vice versa
          const cpx flo_r = mk(__shfl_sync(0xffffffffu, L[r].x, lp), __shfl_sync(0xffffffffu, L[r].y, lp));
          const cpx fhi_r = mk(__shfl_sync(0xffffffffu, L[32 + r].x, lp), __shfl_sync(0xffffffffu, L[32 + r].y, lp));
          const cpx flo_p = mk(__shfl_sync(0xffffffffu, L[pp].x, lp), __shfl_sync(0xffffffffu, L[pp].y, lp));
          const cpx fhi_p = mk(__shfl_sync(0xffffffffu, L[32 + pp].x, lp), __shfl_sync(0xffffffffu, L[32 + pp].y, lp));
          inv4096v2_quad(L[pp], flo_r, L[32 + pp], fhi_r, k8[i], k8[4 + i], v[pp], v[32 + pp]);
          inv4096v2_quad(L[r], flo_p, L[32 + r], fhi_p, k8[2 + i], k8[6 + i], v[r], v[32 + r]);
        }
      }
      if (lane == 0) {
#pragma unroll
        for (int r = 0; r < 64; ++r) v[r] = ex[r];
      }
      __syncwarp();
      {
        cpx(&ve)[32] = *reinterpret_cast<cpx(*)[32]>(&v[0]);
        cpx(&vo)[32] = *reinterpret_cast<cpx(*)[32]>(&v[32]);
        ifft1024(ve, lane, ex, ttw);
        ifft1024(vo, lane, ex, ttw);           // v[oa] = samples 4*(lane + 32*oa), +1; v[32 + oa] = +2, +3 (raw, times N/2)
      }

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
        constexpr int AE = RDIV > 0 ? 32 / RDIV : 32;   // register pairs (e[oa], o[oa]: 128 samples of the frame) per hop
#pragma unroll
        for (int oa = 0; oa < 32; ++oa) {
          const float4 w4 = *reinterpret_cast<const float4*>(s_win + 4 * (lane + 32 * oa));
          v[oa] = mul2(v[oa], mk(w4.x, w4.y));
          v[32 + oa] = mul2(v[32 + oa], mk(w4.z, w4.w));
        }
        const int rb = r0 + 4 * lane;
#pragma unroll
        for (int oa = 0; oa < 32; ++oa) {               // all ring reads of the frame, back to back
          float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
          if (oa < 32 - AE) a = *reinterpret_cast<const float4*>(ring + ((rb + 128 * oa) & mask));
          v[oa] = fma2(v[oa], gain2, mk(a.x, a.y));
          v[32 + oa] = fma2(v[32 + oa], gain2, mk(a.z, a.w));
        }
#pragma unroll
        for (int oa = AE; oa < 32; ++oa)
          *reinterpret_cast<float4*>(ring + ((rb + 128 * oa) & mask)) = make_float4(v[oa].x, v[oa].y, v[32 + oa].x, v[32 + oa].y);
        // the hop that leaves: {denominator, reciprocal} pairs, requested once the other registers are free
        float* dst = out + j_first + 4 * lane;
        constexpr int EB = AE < 8 ? AE : 8;            // 8 register pairs at a time (register budget)
#pragma unroll
        for (int o0 = 0; o0 < AE; o0 += EB) {
          float4 nm[2 * EB];
#pragma unroll
          for (int q = 0; q < EB; ++q) {
            int e = e_base + 4 * (lane + 32 * (o0 + q));
            if (e >= ring_len) e -= ring_len;
            int e2 = e + 2;
            if (e2 >= ring_len) e2 -= ring_len;
            nm[2 * q] = __ldg(reinterpret_cast<const float4*>(p.den_rcp + e));
            nm[2 * q + 1] = __ldg(reinterpret_cast<const float4*>(p.den_rcp + e2));
          }
          bool all_ok = true;
#pragma unroll
          for (int q = 0; q < EB; ++q) {
            const cpx a = v[o0 + q], b = v[32 + o0 + q];
            const float4 t = nm[2 * q], u = nm[2 * q + 1];
            // no short-circuit: one predicate chain, no branches
            all_ok = all_ok & div_tail_ok(a.x) & div_tail_ok(a.y) & div_tail_ok(b.x) & div_tail_ok(b.y);
            stg_stream(reinterpret_cast<cpx*>(dst + 128 * (o0 + q)), mk(div_tail(a.x, t.x, t.y), div_tail(a.y, t.z, t.w)));
            stg_stream(reinterpret_cast<cpx*>(dst + 128 * (o0 + q) + 2), mk(div_tail(b.x, u.x, u.y), div_tail(b.y, u.z, u.w)));
          }
          if (!all_ok) {                                // operands outside the tail's range: the full division
#pragma unroll
            for (int q = 0; q < EB; ++q) {
              const cpx a = v[o0 + q], b = v[32 + o0 + q];
              const float4 t = nm[2 * q], u = nm[2 * q + 1];
              stg_stream(reinterpret_cast<cpx*>(dst + 128 * (o0 + q)), mk(__fdiv_rn(a.x, t.x), __fdiv_rn(a.y, t.z)));
              stg_stream(reinterpret_cast<cpx*>(dst + 128 * (o0 + q) + 2), mk(__fdiv_rn(b.x, u.x), __fdiv_rn(b.y, u.z)));
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
            const int i = 4 * (lane + 32 * q) + 2 * half;   // half 0: the e plane, half 1: the o plane
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
  tmem_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<256>(s_tmem_base);
}

}  // namespace

cudaError_t launch_istft_ola_4096(const K23Params& p, cudaStream_t st) {
  const int smem = kNF * 4 + kW * kPerWarp;
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
