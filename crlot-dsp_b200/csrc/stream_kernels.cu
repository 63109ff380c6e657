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
