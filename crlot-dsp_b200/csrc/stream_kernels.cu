// stream_kernels.cu -- the per-tick kernels of the streaming (config 5) path: thousands of
// concurrent streams, every stream ingesting one small chunk per tick into a device-resident
// ring and emitting `new_frames * hop` output samples.
//
// Reference semantics reproduced per stream (all `ref:` paths relative to the crlot-dsp tree):
//   Framer in DROP mode fed in chunks      ref: dsp/frame/framer.cc:37-59,88-118,128-182
//     -> frame k = samples [k*hop, k*hop+nfft) becomes available when sample k*hop+nfft-1 arrives
//   window -> IFftPlan::forward -> inverse  ref: bench/e2e_benchmark.cc:152-166, kissfft_adapter.cc
//   OLAAccumulator::push_frame_AoS(frame k at k*hop) then produce(hop)
//                                           ref: dsp/ola/OLAAccumulator.cc:54-221,
//                                                bench/ola_benchmark.cc:150-157 (drive order)
// State lives on the device (StreamCounters + ingest rings + accumulator rings); all streams of a
// state advance in lock step, so one pair of counters serves them all and every launch of a tick
// has tick-independent arguments (=> the tick is CUDA-graph replayable).
//
// The transforms reuse the warp-autonomous register FFT of kernels 1 and 2 (warpfft.cuh); a
// warp-step takes its G frames from G *different streams* when a tick yields one frame per stream.
#include <type_traits>

#include "common.cuh"
#include "kernels.h"
#include "warpfft.cuh"

namespace crlot {

namespace {

constexpr int kSWarps = 4;

// ---- ingest: chunk[s][i] -> ring[s][(w + i) & mask] ----------------------------------------------
// The chunk of this tick starts at chunk[s*chunk_stride + (w - w0)]: w0 = samples ingested when the
// caller's buffer began (run-relative addressing keeps the arguments identical from tick to tick).
__global__ void stream_ingest_kernel(const StreamCounters* __restrict__ ctr, const float* __restrict__ chunk,
                                     long long chunk_stride, int chunk_len, float* __restrict__ ring, int rcap,
                                     int n_streams, long long w0) {
  const long long w = ctr->w;
  chunk += w - w0;
  const int mask = rcap - 1;
  const bool vec = (chunk_len % 4 == 0) && (w % 4 == 0) && (chunk_stride % 4 == 0) &&
                   (reinterpret_cast<uintptr_t>(chunk) % 16 == 0) && (rcap % 4 == 0);
  if (vec) {
    const int per = chunk_len / 4;
    const long long total = static_cast<long long>(n_streams) * per;
    for (long long q = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; q < total;
         q += static_cast<long long>(gridDim.x) * blockDim.x) {
      const int s = static_cast<int>(q / per), i = static_cast<int>(q % per) * 4;
      const float4 v = ldg_stream4(reinterpret_cast<const float4*>(chunk + s * chunk_stride + i));
      *reinterpret_cast<float4*>(ring + static_cast<long long>(s) * rcap + ((w + i) & mask)) = v;
    }
  } else {
    const long long total = static_cast<long long>(n_streams) * chunk_len;
    for (long long q = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; q < total;
         q += static_cast<long long>(gridDim.x) * blockDim.x) {
      const int s = static_cast<int>(q / chunk_len), i = static_cast<int>(q % chunk_len);
      ring[static_cast<long long>(s) * rcap + ((w + i) & mask)] = chunk[s * chunk_stride + i];
    }
  }
}

// ---- gather: frame (k0+i) of stream s out of the ingest ring into row s*nf+i of a linear [rows][nfft] buffer.
// Front end of the sizes the ring-reading forward kernel below does not cover (nfft = 4096 and the any-size
// plans): kernel 1 then runs over the rows as non-overlapping frames (hop = nfft).  nfft % 4 == 0, hop % 4 == 0.
__global__ void stream_gather_kernel(const StreamCounters* __restrict__ ctr, const float* __restrict__ ring, int rcap,
                                     int n_streams, int nf, int nfft, int hop, float* __restrict__ rows) {
  const long long k0 = ctr->k;
  const int mask = rcap - 1, per = nfft / 4;
  const long long total = static_cast<long long>(n_streams) * nf * per;
  for (long long q = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; q < total;
       q += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long f = q / per;
    const int u = static_cast<int>(q % per) * 4;
    const long long s = f / nf;
    const int i = static_cast<int>(f % nf);
    const int pos = static_cast<int>(((k0 + i) * hop + u) & mask);
    *reinterpret_cast<float4*>(rows + f * nfft + u) = *reinterpret_cast<const float4*>(ring + s * rcap + pos);
  }
}

// ---- forward: frames (k0+i) of every stream, straight from the ingest rings ----------------------
// frame list index f = s*nf + i  ->  spectra row f of [n_streams][nf][nfft/2+1]
template <int NFFT, bool SCRUB>
__global__ void __launch_bounds__(kSWarps * 32) stream_fwd_kernel(const StreamFwdParams p) {
  using X = GeoX<NFFT>;
  constexpr int C = X::C, G = X::G, A = X::A, B = X::B;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* s_win = reinterpret_cast<float*>(smem_raw);
  cpx* s_tw1 = reinterpret_cast<cpx*>(s_win + NFFT);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  cpx* ex = s_tw1 + C + warp * X::XCPX;
  for (int i = threadIdx.x; i < NFFT; i += blockDim.x) s_win[i] = p.window[i];
  for (int i = threadIdx.x; i < C; i += blockDim.x) s_tw1[i] = p.tw1[i];
  const SplitLane<X::M2> SL = make_split_fwd<X>(lane);
  __syncthreads();

  const long long k0 = p.ctr->k;
  const int mask = p.rcap - 1;
  const long long total = static_cast<long long>(p.n_streams) * p.nf;
  const long long steps = (total + G - 1) / G;
  const int lp = mirror_lane<B>(lane);
  for (long long step = static_cast<long long>(blockIdx.x) * kSWarps + warp; step < steps;
       step += static_cast<long long>(gridDim.x) * kSWarps) {
    cpx v[32];
#pragma unroll
    for (int j = 0; j < X::M1; ++j) {
      const int i1 = lane + 32 * j, g = i1 / A, a = i1 % A;
      const long long f = step * G + g;
      const bool ok = f < total;
      const long long s = ok ? f / p.nf : 0;
      const int i = ok ? static_cast<int>(f % p.nf) : 0;
      const float* base = p.ring + s * p.rcap;
      const int pos0 = static_cast<int>(((k0 + i) * p.hop) & mask);
#pragma unroll
      for (int b = 0; b < B; ++b) {
        const int off = 2 * a + 2 * A * b;
        cpx x = mk(0.f, 0.f);
        if (ok) x = *reinterpret_cast<const cpx*>(base + ((pos0 + off) & mask));
        cpx pw = mul2(x, *reinterpret_cast<const cpx*>(s_win + off));
        if (SCRUB) pw = mk(scrub(pw.x), scrub(pw.y));
        v[j * B + b] = pw;
      }
    }
    pass1<X, false>(v);
    twiddle1<X, false>(v, lane, s_tw1);
    __syncwarp();
    xchg_write<X>(v, lane, ex);
    __syncwarp();
    xchg_read<X>(v, lane, ex);
    pass2<X, false>(v);
    cpx nyq[X::M2], mid[X::M2];
    auto mirror = [&](int r) { return mk(__shfl_sync(0xffffffffu, v[r].x, lp), __shfl_sync(0xffffffffu, v[r].y, lp)); };
    fwd_split_pairs<X>(v, lane, SL, mirror, nyq, mid);
    fwd_split_finish<X>(v, lane, mirror, mid);
#pragma unroll
    for (int j = 0; j < X::M2; ++j) {
      const int i2 = lane + 32 * j, g = i2 / B, ob = i2 % B;
      const long long f = step * G + g;
      if (f < total) {
        cpx* row = p.spec + f * (C + 1);
#pragma unroll
        for (int oa = 0; oa < A; ++oa) row[ob + B * oa] = v[j * A + oa];
        if (ob == 0) row[C] = nyq[j];
      }
    }
  }
}

// ---- overlap-add into the accumulator rings + produce -------------------------------------------
// One thread owns 4 consecutive ring positions t = k0*hop + j of one stream, j in [0, span):
//   acc[t] += sum over this tick's frames i ascending of fma(frames[s][i][t-(k0+i)*hop], gain, .)
//   (same order as the reference, which pushes frame k0, produces, pushes frame k0+1, ...: a
//   position is only ever produced after every frame that reaches it has been pushed)
//   j < n_prod:  out[s][j] = acc / max(norm[t mod ring_len], eps); acc = 0   (normalize_and_clear)
template <bool WIN>
__global__ void __launch_bounds__(256) stream_ola_kernel(const StreamOlaParams p) {
  const long long k0 = p.ctr->k;
  const int span4 = p.span / 4;
  const long long total = static_cast<long long>(p.n_streams) * span4;
  const int N = p.nfft, H = p.hop;
  for (long long q = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; q < total;
       q += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long s = q / span4;
    const int j = static_cast<int>(q % span4) * 4;
    const long long t = k0 * H + j;
    const int rp = static_cast<int>(t % p.ring_len);
    float* accp = p.acc + s * p.ring_len + rp;
    float4 acc = *reinterpret_cast<const float4*>(accp);
    const float* fr = p.frames + s * p.nf * N;
    // frames i with 0 <= j - i*H < N
    int i_hi = j / H;
    if (i_hi > p.nf - 1) i_hi = p.nf - 1;
    int i_lo = (j - N + H) < 0 ? 0 : (j - N + H) / H;
    for (int i = i_lo; i <= i_hi; ++i) {
      const int u = j - i * H;
      float4 y = *reinterpret_cast<const float4*>(fr + static_cast<long long>(i) * N + u);
      if (WIN) {
        const float4 w = __ldg(reinterpret_cast<const float4*>(p.window + u));
        y.x = __fmul_rn(y.x, w.x); y.y = __fmul_rn(y.y, w.y); y.z = __fmul_rn(y.z, w.z); y.w = __fmul_rn(y.w, w.w);
      }
      acc.x = __fmaf_rn(y.x, p.gain, acc.x);
      acc.y = __fmaf_rn(y.y, p.gain, acc.y);
      acc.z = __fmaf_rn(y.z, p.gain, acc.z);
      acc.w = __fmaf_rn(y.w, p.gain, acc.w);
    }
    if (j < p.n_prod) {
      const float4 nm = __ldg(reinterpret_cast<const float4*>(p.norm + rp));
      float4 o;
      o.x = __fdiv_rn(acc.x, fmaxf(nm.x, p.eps));
      o.y = __fdiv_rn(acc.y, fmaxf(nm.y, p.eps));
      o.z = __fdiv_rn(acc.z, fmaxf(nm.z, p.eps));
      o.w = __fdiv_rn(acc.w, fmaxf(nm.w, p.eps));
      *reinterpret_cast<float4*>(p.out + s * p.out_stride + (k0 - p.k_base) * H + j) = o;
      acc = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    *reinterpret_cast<float4*>(accp) = acc;
  }
}

__global__ void stream_advance_kernel(StreamCounters* ctr, int chunk_len, int nf) {
  ctr->w += chunk_len;
  ctr->k += nf;
}

// ---- the whole tick in ONE kernel, for the steady state of a low-latency stream: exactly one new
// frame per stream per tick (chunk_len == hop once nfft samples are in) -----------------------------
// A warp-step owns G streams from the chunk to the output hop: ingest -> frame from the ring ->
// window + rFFT -> spectra row (kept, crlot_stream_last_spectra) -> irFFT, 1/N, clean-up, synthesis
// window -> accumulate into the OLA ring at k*hop -> produce(hop).  Same arithmetic in the same order
// as the five-launch tick above (bit-identical outputs, tests/test_gpu_stream.py), but one launch
// instead of five: the tick is launch-latency-bound, not bandwidth-bound.  Data a warp wrote and
// reads back within the tick (ring, spectra) goes through L2 (ld.global.cg) after a __syncwarp().
// The last block to finish advances the device counters (every block has read them by then).
__device__ __forceinline__ cpx ldcg2(const void* p) {
  const float2 v = __ldcg(reinterpret_cast<const float2*>(p));
  return mk(v.x, v.y);
}

template <int NFFT, bool SCRUB, bool WIN>
__global__ void __launch_bounds__(kSWarps * 32) stream_tick1_kernel(const StreamTick1Params p) {
  using X = GeoX<NFFT>;
  constexpr int C = X::C, G = X::G, A = X::A, B = X::B;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* s_win = reinterpret_cast<float*>(smem_raw);
  float* s_wsyn = s_win + NFFT;
  cpx* s_tw1 = reinterpret_cast<cpx*>(s_wsyn + (WIN ? NFFT : 0));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  cpx* ex = s_tw1 + C + warp * X::XCPX;
  // Programmatic dependent launch: the next tick's blocks may start as soon as every block of this one is
  // running; they do their prologue (tables into shared memory, per-lane split constants -- nothing a tick
  // writes) while this tick computes, and wait below for this grid to complete and flush before touching
  // the counters, rings or spectra.  Both instructions are no-ops in a launch without the attribute.
  if (p.pdl) asm volatile("griddepcontrol.launch_dependents;");
  for (int i = threadIdx.x; i < NFFT; i += blockDim.x) {
    s_win[i] = p.window[i];
    if (WIN) s_wsyn[i] = p.syn_window[i];
  }
  for (int i = threadIdx.x; i < C; i += blockDim.x) s_tw1[i] = p.tw1[i];
  const SplitLane<X::M2> SLF = make_split_fwd<X>(lane);
  const SplitLane<X::M1> SLI = make_split_inv<X>(lane);
  if (p.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
  const long long w = p.ctr->w, k0 = p.ctr->k;
  __syncthreads();

  const int mask = p.rcap - 1, H = p.hop, S = p.n_streams, CL = p.chunk_len;
  const long long steps = (static_cast<long long>(S) + G - 1) / G;
  const int lpB = mirror_lane<B>(lane), lpA = mirror_lane<A>(lane);
  const int pos0 = static_cast<int>((k0 * H) & mask);
  const int e0 = static_cast<int>((k0 * H) % p.ring_len);
  const float scale = 2.0f / static_cast<float>(NFFT);
  const float* chunk = p.chunk + (w - p.w0);
  for (long long step = static_cast<long long>(blockIdx.x) * kSWarps + warp; step < steps;
       step += static_cast<long long>(gridDim.x) * kSWarps) {
    const long long s_base = step * G;
    // 1. ingest the chunks of this step's streams
    constexpr int IB = 16;                                // sixteen independent loads in flight per lane: with one hop
    const int dg = 32 / CL, di = 32 - dg * CL;            // of 64 samples per stream (G = 8) that is ONE round trip
    for (int q0 = lane; q0 < G * CL; q0 += 32 * IB) {
      float val[IB];
      long long dst[IB];
      int g = q0 / CL, i = q0 - g * CL;                   // (stream, sample) of q0 + 32 u, stepped instead of divided
#pragma unroll
      for (int u = 0; u < IB; ++u) {
        const long long s = s_base + g;
        const bool ok = g < G && s < S;
        dst[u] = ok ? s * p.rcap + ((w + i) & mask) : -1;
        val[u] = ok ? __ldg(chunk + s * p.chunk_stride + i) : 0.0f;
        g += dg;
        i += di;
        if (i >= CL) {
          i -= CL;
          ++g;
        }
      }
#pragma unroll
      for (int u = 0; u < IB; ++u)
        if (dst[u] >= 0) p.ring[dst[u]] = val[u];
    }
    __syncwarp();
    // 2. frame k0 of every stream: window + forward real FFT
    cpx v[32];
#pragma unroll
    for (int j = 0; j < X::M1; ++j) {
      const int i1 = lane + 32 * j, g = i1 / A, a = i1 % A;
      const long long s = s_base + g;
      const bool ok = s < S;
      const float* base = p.ring + (ok ? s : 0) * p.rcap;
#pragma unroll
      for (int b = 0; b < B; ++b) {
        const int off = 2 * a + 2 * A * b;
        cpx x = mk(0.f, 0.f);
        if (ok) x = ldcg2(base + ((pos0 + off) & mask));
        cpx pw = mul2(x, *reinterpret_cast<const cpx*>(s_win + off));
        if (SCRUB) pw = mk(scrub(pw.x), scrub(pw.y));
        v[j * B + b] = pw;
      }
    }
    pass1<X, false>(v);
    twiddle1<X, false>(v, lane, s_tw1);
    __syncwarp();
    xchg_write<X>(v, lane, ex);
    __syncwarp();
    xchg_read<X>(v, lane, ex);
    pass2<X, false>(v);
    {
      cpx nyq[X::M2], mid[X::M2];
      auto mirror = [&](int r) { return mk(__shfl_sync(0xffffffffu, v[r].x, lpB), __shfl_sync(0xffffffffu, v[r].y, lpB)); };
      fwd_split_pairs<X>(v, lane, SLF, mirror, nyq, mid);
      fwd_split_finish<X>(v, lane, mirror, mid);
#pragma unroll
      for (int j = 0; j < X::M2; ++j) {
        const int i2 = lane + 32 * j, g = i2 / B, ob = i2 % B;
        const long long s = s_base + g;
        if (s < S) {
          cpx* row = p.spec + s * (C + 1);
#pragma unroll
          for (int oa = 0; oa < A; ++oa) row[ob + B * oa] = v[j * A + oa];
          if (ob == 0) row[C] = nyq[j];
        }
      }
    }
    __syncwarp();
    // 3. inverse real FFT of the same rows (pass-1 view), 1/N, clean-up, synthesis window
    cpx nyqi[X::M1];
#pragma unroll
    for (int j = 0; j < X::M1; ++j) {
      const int i1 = lane + 32 * j, g = i1 / A, a = i1 % A;
      const long long s = s_base + g;
      const bool ok = s < S;
      const cpx* row = p.spec + (ok ? s : 0) * (C + 1);
#pragma unroll
      for (int b = 0; b < B; ++b) v[j * B + b] = ok ? ldcg2(row + a + A * b) : mk(0.f, 0.f);
      nyqi[j] = (ok && a == 0) ? ldcg2(row + C) : mk(0.f, 0.f);
    }
    {
      cpx mid[X::M1];
      auto mirror = [&](int r) { return mk(__shfl_sync(0xffffffffu, v[r].x, lpA), __shfl_sync(0xffffffffu, v[r].y, lpA)); };
      inv_split_pairs<X>(v, lane, SLI, mirror, nyqi, mid);
      inv_split_finish<X>(v, lane, mirror, mid);
    }
    pass1<X, true>(v);
    twiddle1<X, true>(v, lane, s_tw1);
    __syncwarp();
    xchg_write<X>(v, lane, ex);
    __syncwarp();
    xchg_read<X>(v, lane, ex);
    pass2<X, true>(v);
    // 4. accumulate at k0*hop, produce the first hop (normalize_and_clear), keep the rest.
    //    All accumulator / envelope loads of a register group are issued before the first store, so a
    //    group costs one L2 round trip instead of one per sample pair.  With hop <= nfft/4 (the usual case, tested once
    //    per warp) only the first quarter of a group's registers can fall below the hop: the others are compiled WITHOUT
    //    the normalise-and-clear branch and its two IEEE divisions -- 24 of 32 taken branches and their code go away
    //    (the tick is paced by one warp's instruction stream: `no_instruction` was half of the stall samples of this
    //    step, profiles/r2g_ncu_tick.*).
    auto step4 = [&](auto hop_small) {
      constexpr bool HS = decltype(hop_small)::value;
      constexpr int AQ = A >= 4 ? A / 4 : 1;
#pragma unroll
      for (int j = 0; j < X::M2; ++j) {
        const int i2 = lane + 32 * j, g = i2 / B, ob = i2 % B;
        const long long s = s_base + g;
        if (s < S) {
          float* accb = p.acc + s * p.ring_len;
          float* outb = p.out + s * p.out_stride + (k0 - p.k_base) * H;
          constexpr int AB = A < 16 ? A : 16;                 // loads batched 16 registers at a time
#pragma unroll
          for (int o0 = 0; o0 < A; o0 += AB) {
            float2 a2[AB], nm[AB];
            int rps[AB];
#pragma unroll
            for (int q = 0; q < AB; ++q) {
              const int i = 2 * (ob + B * (o0 + q));
              const bool may = !HS || o0 + q < AQ;            // compile-time
              int rp = e0 + i;
              if (rp >= p.ring_len) rp -= static_cast<int>(p.ring_len);
              rps[q] = rp;
              a2[q] = __ldcg(reinterpret_cast<const float2*>(accb + rp));
              nm[q] = (may && i < H) ? __ldg(reinterpret_cast<const float2*>(p.norm + rp)) : make_float2(1.f, 1.f);
            }
#pragma unroll
            for (int q = 0; q < AB; ++q) {
              const int oa = o0 + q, i = 2 * (ob + B * oa);
              const bool may = !HS || oa < AQ;                // compile-time
              float y0 = v[j * A + oa].x * scale, y1 = v[j * A + oa].y * scale;
              if (SCRUB) {
                y0 = scrub(y0);
                y1 = scrub(y1);
              }
              if (WIN) {
                const float2 ws = *reinterpret_cast<const float2*>(s_wsyn + i);
                y0 = __fmul_rn(y0, ws.x);
                y1 = __fmul_rn(y1, ws.y);
              }
              float2 acc2 = a2[q];
              acc2.x = __fmaf_rn(y0, p.gain, acc2.x);
              acc2.y = __fmaf_rn(y1, p.gain, acc2.y);
              if (may && i < H) {
                *reinterpret_cast<float2*>(outb + i) =
                    make_float2(__fdiv_rn(acc2.x, fmaxf(nm[q].x, p.eps)), __fdiv_rn(acc2.y, fmaxf(nm[q].y, p.eps)));
                acc2 = make_float2(0.f, 0.f);
              }
              a2[q] = acc2;
            }
#pragma unroll
            for (int q = 0; q < AB; ++q) *reinterpret_cast<float2*>(accb + rps[q]) = a2[q];
          }
        }
      }
    };
    if (4 * H <= NFFT) step4(std::true_type{});
    else               step4(std::false_type{});
  }
  // 5. the last block to finish advances the counters
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int ticket = atomicAdd(p.done, 1u);
    if (ticket == gridDim.x - 1) {
      p.ctr->w = w + CL;
      p.ctr->k = k0 + 1;
      *p.done = 0u;
    }
  }
}

template <int NFFT>
cudaError_t launch_tick1_t(const StreamTick1Params& p, cudaStream_t st) {
  using X = GeoX<NFFT>;
  const bool win = p.syn_window != nullptr;
  const int smem = NFFT * 4 * (win ? 2 : 1) + X::C * 8 + kSWarps * X::XCPX * 8;
  void (*kern)(const StreamTick1Params) = nullptr;
  if (win) kern = p.scrub ? stream_tick1_kernel<NFFT, true, true> : stream_tick1_kernel<NFFT, false, true>;
  else     kern = p.scrub ? stream_tick1_kernel<NFFT, true, false> : stream_tick1_kernel<NFFT, false, false>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  const long long steps = (static_cast<long long>(p.n_streams) + X::G - 1) / X::G;
  long long blocks = (steps + kSWarps - 1) / kSWarps;
  const long long cap = static_cast<long long>(device_sm_count()) * 2;
  if (blocks > cap) blocks = cap;
  if (p.pdl) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(static_cast<unsigned>(blocks));
    cfg.blockDim = dim3(kSWarps * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, p);
  }
  kern<<<static_cast<unsigned>(blocks), kSWarps * 32, smem, st>>>(p);
  return cudaGetLastError();
}

template <int NFFT>
cudaError_t launch_fwd_t(const StreamFwdParams& p, cudaStream_t st) {
  using X = GeoX<NFFT>;
  const int smem = NFFT * 4 + X::C * 8 + kSWarps * X::XCPX * 8;
  auto kern = p.scrub ? stream_fwd_kernel<NFFT, true> : stream_fwd_kernel<NFFT, false>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  const long long steps = (static_cast<long long>(p.n_streams) * p.nf + X::G - 1) / X::G;
  long long blocks = (steps + kSWarps - 1) / kSWarps;
  const long long cap = static_cast<long long>(device_sm_count()) * 4;
  if (blocks > cap) blocks = cap;
  kern<<<static_cast<unsigned>(blocks), kSWarps * 32, smem, st>>>(p);
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_stream_ingest(const StreamCounters* ctr, const float* chunk, long long chunk_stride, int chunk_len,
                                 float* ring, int rcap, int n_streams, long long w0, cudaStream_t st) {
  if (chunk_len <= 0 || n_streams <= 0) return cudaSuccess;
  const long long work = static_cast<long long>(n_streams) * ((chunk_len + 3) / 4);
  long long blocks = (work + 255) / 256;
  const long long cap = static_cast<long long>(device_sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  stream_ingest_kernel<<<static_cast<unsigned>(blocks), 256, 0, st>>>(ctr, chunk, chunk_stride, chunk_len, ring, rcap,
                                                                     n_streams, w0);
  return cudaGetLastError();
}

cudaError_t launch_stream_gather(const StreamCounters* ctr, const float* ring, int rcap, int n_streams, int nf, int nfft,
                                 int hop, float* rows, cudaStream_t st) {
  if (nf <= 0 || n_streams <= 0) return cudaSuccess;
  const long long work = static_cast<long long>(n_streams) * nf * (nfft / 4);
  long long blocks = (work + 255) / 256;
  const long long cap = static_cast<long long>(device_sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  stream_gather_kernel<<<static_cast<unsigned>(blocks), 256, 0, st>>>(ctr, ring, rcap, n_streams, nf, nfft, hop, rows);
  return cudaGetLastError();
}

bool stream_fwd_supported(int nfft) { return nfft == 256 || nfft == 512 || nfft == 1024 || nfft == 2048; }

cudaError_t launch_stream_fwd(const StreamFwdParams& p, cudaStream_t st) {
  if (p.nf <= 0 || p.n_streams <= 0) return cudaSuccess;
  switch (p.nfft) {
    case 256: return launch_fwd_t<256>(p, st);
    case 512: return launch_fwd_t<512>(p, st);
    case 1024: return launch_fwd_t<1024>(p, st);
    case 2048: return launch_fwd_t<2048>(p, st);
    default: return cudaErrorNotSupported;
  }
}

bool stream_tick1_supported(int nfft, int hop, int chunk_len, long long ring_len, long long out_stride, const void* out) {
  if (nfft != 256 && nfft != 512 && nfft != 1024 && nfft != 2048) return false;
  if (chunk_len != hop || hop % 2 != 0 || hop > nfft) return false;
  if (ring_len % 2 != 0 || ring_len >= (1ll << 30) || out_stride % 2 != 0) return false;
  return reinterpret_cast<uintptr_t>(out) % 8 == 0;
}

cudaError_t launch_stream_tick1(const StreamTick1Params& p, cudaStream_t st) {
  if (p.n_streams <= 0) return cudaSuccess;
  switch (p.nfft) {
    case 256: return launch_tick1_t<256>(p, st);
    case 512: return launch_tick1_t<512>(p, st);
    case 1024: return launch_tick1_t<1024>(p, st);
    case 2048: return launch_tick1_t<2048>(p, st);
    default: return cudaErrorNotSupported;
  }
}

cudaError_t launch_stream_ola(const StreamOlaParams& p, cudaStream_t st) {
  if (p.span <= 0 || p.n_streams <= 0) return cudaSuccess;
  const long long work = static_cast<long long>(p.n_streams) * (p.span / 4);
  long long blocks = (work + 255) / 256;
  const long long cap = static_cast<long long>(device_sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  if (p.window) stream_ola_kernel<true><<<static_cast<unsigned>(blocks), 256, 0, st>>>(p);
  else          stream_ola_kernel<false><<<static_cast<unsigned>(blocks), 256, 0, st>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_stream_advance(StreamCounters* ctr, int chunk_len, int nf, cudaStream_t st) {
  stream_advance_kernel<<<1, 1, 0, st>>>(ctr, chunk_len, nf);
  return cudaGetLastError();
}

}  // namespace crlot
