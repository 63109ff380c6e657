// k1_4096v2.cu -- kernel 1 for N = 4096 as TWO 1024-point complex sub-transforms per frame.
//
// Replaces, per frame, Framer::pop -> window multiply -> KissFftPlan::forward -> kiss_fftr (reference:
// dsp/frame/framer.cc:128-182, bench/e2e_benchmark.cc:152-156, dsp/fft/backends/kissfft_adapter.cc:83-122,
// third_party/kissfft/kiss_fftr.c:63-117), like k1_stft.cu, for the one frame size whose 2048 complex points do not
// fit 32 registers per lane.  The first version (stft_fwd_kernel_4096, k1_stft.cu) spends 64 complex registers per
// lane: 254 registers, 8 warps per SM, two per scheduler -- latency-bound at 0.51-0.54 of the HBM roofline.  This
// version keeps the 32-register machinery of the smaller sizes (12 warps per SM):
//
//   z[n] = x[2n] + i x[2n+1], n < 2048;   e[m] = z[2m],  o[m] = z[2m+1],  m < 1024        (radix-2 decimation in time)
//   E = FFT_1024(e),  O = FFT_1024(o);     Z[k] = E[k] + W_2048^k O[k],  Z[k+1024] = E[k] - W_2048^k O[k]
//   X[q], q = 0..2048: the real split of Z (kiss_fftr.c:93-116 restated)
//
// * a lane's pass-1 points m = lane + 32 b are the float4 groups x[4m .. 4m+3] = (e[m], o[m]): one 16-byte load each;
// * the warp transforms e, parks E in shared memory, transforms o, then combines and splits on the fly: the partner of
//   Z[k] in the split is Z[2048 - k] = Z[(1024 - k) + 1024], which the MIRROR lane computes as E - W O at the mirrored
//   register -- one shuffle per pair, never more than 32 complex values live;
// * raw samples are carried between frames in TENSOR MEMORY when hop = 1024: m advances by 256 = 8 register steps per
//   frame, so a sample stays in its lane; the ring holds the e and o planes (2 x 64 columns per warp), frame chunk c
//   sits in ring chunk (c + s) mod 4, and only the new hop is fetched from global memory.  Other hops fetch every chunk.
//   No shared-memory ring: the 16 KB per warp it cost are what kept the first version at 8 warps.
//   TMEM columns: [0,64) inter-pass twiddles, [64,128) combine twiddles, [128,512) three warps x (e plane, o plane).
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "warpfft.cuh"

namespace crlot {

namespace {

constexpr int kV2Warps = 12;
using XV = GeoX<2048>;                                   // the 1024-point complex sub-transform: 32 x 32
constexpr int kV2PerWarp = XV::XCPX * 8 + 1024 * 8;      // exchange buffer + parked E
constexpr int kV2Tables = 4096 * 4 + 1024 * 8;           // window planes (e, o) + split factors
constexpr int kV2Smem = kV2Tables + kV2Warps * kV2PerWarp;

__device__ __noinline__ float4 ld4_guard(const float* sig, long long len, long long idx, int pad_mode) {
  return make_float4(pad_value(sig, len, idx, pad_mode), pad_value(sig, len, idx + 1, pad_mode),
                     pad_value(sig, len, idx + 2, pad_mode), pad_value(sig, len, idx + 3, pad_mode));
}

// one 1024-point complex transform of the 32 registers (pass-1 view in, pass-2 view out), twiddles from tensor memory
__device__ __forceinline__ void fft1024(cpx (&v)[32], int lane, cpx* ex, uint32_t ttw) {
  cpx t0[8];
  tmem_ld8(ttw, t0);
  pass1<XV, false>(v);
  cpx tb[2][8];
#pragma unroll
  for (int i = 0; i < 8; ++i) tb[0][i] = t0[i];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    tmem_wait_ld8(tb[c & 1]);
    if (c + 1 < 4) tmem_ld8(ttw + 16 * (c + 1), tb[(c + 1) & 1]);
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (8 * c + i > 0) v[8 * c + i] = cmul(v[8 * c + i], tb[c & 1][i]);
  }
  xchg_write<XV>(v, lane, ex);
  __syncwarp();
  xchg_read<XV>(v, lane, ex);
  __syncwarp();
  pass2<XV, false>(v);
}

// predicated streaming store without a compiler memory barrier (nothing in the kernel reads the spectra back)
__device__ __forceinline__ void stg_stream_if(bool on, cpx* p, cpx v) {
  asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %3, 0;\n\t@q st.global.L1::no_allocate.v2.f32 [%0], {%1, %2};\n\t}" ::"l"(p),
               "f"(v.x), "f"(v.y), "r"(static_cast<unsigned int>(on)));
}

// X[q] = s/2 + f, X[2048-q] = conj(s/2 - f) with s = z + conj m, f = q * (z - conj m), q = (-i/2) exp(-i pi q/2048)
__device__ __forceinline__ void split_fwd(cpx z, cpx m, cpx q, cpx& x1, cpx& x2) {
  const cpx s = fma2(m, mk(1.f, -1.f), z);
  const cpx d = fma2(m, mk(-1.f, 1.f), z);
  const cpx f = cmul(d, q);
  x1 = fma2(s, mk(0.5f, 0.5f), f);
  x2 = fma2(s, mk(0.5f, -0.5f), mul2(f, mk(-1.f, 1.f)));
}

template <int SCRUB>
__global__ void __launch_bounds__(kV2Warps * 32, 1) stft_fwd_kernel_4096v2(const K1Params p) {
  constexpr int NFFT = 4096, C = 2048;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint32_t s_tmem_base;
  cpx* s_we = reinterpret_cast<cpx*>(smem_raw);          // s_we[m] = (w[4m], w[4m+1]),  s_wo[m] = (w[4m+2], w[4m+3])
  cpx* s_wo = s_we + 1024;
  cpx* s_qs = s_wo + 1024;                               // split factors, [p*32 + lane]
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform by construction
  cpx* ex = reinterpret_cast<cpx*>(smem_raw + kV2Tables + warp * kV2PerWarp);
  cpx* park = ex + XV::XCPX;

  for (int m = threadIdx.x; m < 1024; m += blockDim.x) {
    const float4 w = *reinterpret_cast<const float4*>(p.window + 4 * m);
    s_we[m] = mk(w.x, w.y);
    s_wo[m] = mk(w.z, w.w);
    s_qs[m] = p.sa[2048 + m];
  }
  if (warp == 0) tmem_alloc<512>(&s_tmem_base);
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();
  const uint32_t tmb = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  const uint32_t ttw = tmb, ttc = tmb + 64;
  const uint32_t tre = tmb + 128 + 128 * (warp >> 2), tro = tre + 64;   // this warp's e and o planes
  if (warp < 4) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      cpx a8[8], b8[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        a8[i] = p.sa[(8 * c + i) * 32 + lane];             // W_1024^(lane*ob)
        b8[i] = p.sa[1024 + (8 * c + i) * 32 + lane];      // W_2048^(lane + 32*oa)
      }
      tmem_st8(ttw + 16 * c, a8);
      tmem_st8(ttc + 16 * c, b8);
    }
    tmem_wait_st();
  }
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();

  // lane l >= 1 finishes lane 0's self-mirrored pair of bins 32*l and 2048 - 32*l (see the end of the frame loop)
  cpx qcoop;
  {
    const double th = 3.14159265358979323846 * (double)(32 * lane) / 2048.0;
    qcoop = mk((float)(-0.5 * sin(th)), (float)(-0.5 * cos(th)));
  }
  const int H = p.hop;
  const bool carry = H == 1024 && p.pad == 0;              // raw samples stay in tensor memory between frames
  // chunk_steps frames per chunk for the first three quarters of a channel, a quarter of that for the rest (drawn after
  // all the big ones): few chunk starts -- each re-reads three hops -- and still a short tail at the end of the kernel
  const long long cb = p.chunk_steps, cbs = cb >= 16 ? cb / 4 : cb;
  const long long big_per_chan = (p.frames * 3 / 4) / cb;
  const long long small_per_chan = (p.frames - big_per_chan * cb + cbs - 1) / cbs;
  const long long total_big = big_per_chan * p.channels;
  const long long total_chunks = total_big + small_per_chan * p.channels;
  const int lp = (32 - lane) & 31;

  for (;;) {
    unsigned long long qq = 0;
    if (lane == 0) qq = atomicAdd(p.counter, 1u);
    qq = __shfl_sync(0xffffffffu, qq, 0);
    if (static_cast<long long>(qq) >= total_chunks) break;
    long long c, first;
    int nsteps;
    if (static_cast<long long>(qq) < total_big) {
      c = qq / big_per_chan;
      first = (qq % big_per_chan) * cb;
      nsteps = static_cast<int>(cb);
    } else {
      const long long qs = qq - total_big;
      c = qs / small_per_chan;
      first = big_per_chan * cb + (qs % small_per_chan) * cbs;
      nsteps = static_cast<int>(min(cbs, p.frames - first));
    }
    const float* sig = p.sig + c * p.chan_stride;
    cpx* spec = p.spec + (c * p.spec_chan_rows + first) * (C + 1);
    int dirty = 0;

    // the next frame's new hop is requested a frame ahead with cp.async into a corner of the exchange buffer that the
    // combine phase leaves alone (cpx [128, 640)): no registers are held across the phase that hides the latency
    float4* stage = reinterpret_cast<float4*>(ex + 128);
    for (int s = 0; s < nsteps; ++s) {
      const long long f0 = (first + s) * H - p.pad;          // signal index of the frame's first sample
      const bool whole = f0 >= 0 && f0 + NFFT <= p.len;
      cpx v[32];
      unsigned int wmin = 0xffffffffu;
      cpx nf = mk(0.f, 0.f);
      auto screen = [&](const cpx (&e8)[8], const cpx (&o8)[8]) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const unsigned int u0 = __float_as_uint(e8[i].x), u1 = __float_as_uint(e8[i].y);
          const unsigned int u2 = __float_as_uint(o8[i].x), u3 = __float_as_uint(o8[i].y);
          wmin = min(min(wmin, min(u0 + u0 - 2u, u1 + u1 - 2u)), min(u2 + u2 - 2u, u3 + u3 - 2u));
          nf = fma2(e8[i], mk(0.f, 0.f), fma2(o8[i], mk(0.f, 0.f), nf));
        }
      };
      // ---- new samples into the two tensor-memory planes (e = x[4m], x[4m+1]; o = x[4m+2], x[4m+3]) ----------------
      if (carry && s > 0) {                                  // only the hop that has just arrived, staged by cp.async
        asm volatile("cp.async.wait_all;" ::: "memory");
        cpx e8[8], o8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 x = stage[i * 32 + lane];
          e8[i] = mk(x.x, x.y);
          o8[i] = mk(x.z, x.w);
        }
        const int rc = (3 + s) & 3;
        tmem_st8(tre + 16 * rc, e8);
        tmem_st8(tro + 16 * rc, o8);
        if (SCRUB == 2) screen(e8, o8);
      } else {
#pragma unroll 1
        for (int ch = 0; ch < 4; ++ch) {
          cpx e8[8], o8[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const long long idx = f0 + 4 * (lane + 32 * (8 * ch + i));
            const float4 x = whole ? ldg_stream4(reinterpret_cast<const float4*>(sig + idx)) : ld4_guard(sig, p.len, idx, p.pad_mode);
            e8[i] = mk(x.x, x.y);
            o8[i] = mk(x.z, x.w);
          }
          const int rc = carry ? ((ch + s) & 3) : ch;
          tmem_st8(tre + 16 * rc, e8);
          tmem_st8(tro + 16 * rc, o8);
          if (SCRUB == 2) screen(e8, o8);
        }
      }
      tmem_wait_st();
      bool clean_all = SCRUB == 1;
      if (SCRUB == 2) {
        const bool suspect = (wmin < p.tiny2) || (nf.x != 0.0f) || (nf.y != 0.0f);
        if (__any_sync(0xffffffffu, suspect)) dirty = 5;
        clean_all = dirty > 0 || s == 0 || !carry;           // without the carry-over every sample of a frame is new
        if (dirty > 0) --dirty;
      }
      // ---- the two 1024-point sub-transforms: ONE copy of the code, run for the e plane and then the o plane --------
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        const uint32_t plane = h ? tro : tre;
        const cpx* win = h ? s_wo : s_we;
        // all four chunks of the plane in flight at once, straight into the registers they are multiplied in (one exposed
        // tensor-memory latency per plane instead of four)
        cpx x8[4][8];
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) tmem_ld8(plane + 16 * (carry ? ((ch + s) & 3) : ch), x8[ch]);
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          tmem_wait_ld8(x8[ch]);
#pragma unroll
          for (int i = 0; i < 8; ++i) v[8 * ch + i] = mul2(x8[ch][i], win[lane + 32 * (8 * ch + i)]);   // one rounding per sample
        }
        if (clean_all) {
#pragma unroll
          for (int r = 0; r < 32; ++r) v[r] = mk(scrub(v[r].x), scrub(v[r].y));
        }
        fft1024(v, lane, ex, ttw);
        if (h == 0) {
#pragma unroll
          for (int oa = 0; oa < 32; ++oa) park[oa * 32 + lane] = v[oa];      // E[lane + 32*oa]
        }
      }
      // ---- request the next frame's new hop now: the loads complete under the combine / split / store below ----------
      if (carry && s + 1 < nsteps) {
        const long long g0 = f0 + H + 3 * H;                 // the last quarter of the next frame
        const bool in = g0 >= 0 && g0 + H <= p.len;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const long long idx = g0 + 4 * (lane + 32 * i);
          if (in)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(stage + i * 32 + lane)), "l"(sig + idx) : "memory");
          else
            stage[i * 32 + lane] = ld4_guard(sig, p.len, idx, p.pad_mode);
        }
      }
      // ---- t = W_2048^k O[k] -------------------------------------------------------------------------------------------
      {
        cpx tb[2][8];                                        // the next chunk of twiddles in flight while this one multiplies
        tmem_ld8(ttc, tb[0]);
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          tmem_wait_ld8(tb[ch & 1]);
          if (ch + 1 < 4) tmem_ld8(ttc + 16 * (ch + 1), tb[(ch + 1) & 1]);
#pragma unroll
          for (int i = 0; i < 8; ++i) v[8 * ch + i] = cmul(v[8 * ch + i], tb[ch & 1][i]);
        }
      }
      // ---- combine + real split + store, a mirrored pair of registers at a time ----------------------------------------
      // own bin k = lane + 32*pp: Z[k] = E + t; its partner Z[2048-k] = (E - t) of the mirror lane at register 31 - pp
      cpx zhi0 = mk(0.f, 0.f);
#pragma unroll
      for (int pp = 0; pp < 16; ++pp) {
        const int r = 31 - pp;
        const cpx Ep = park[pp * 32 + lane], Er = park[r * 32 + lane];
        const cpx zlo_p = Ep + v[pp], zhi_p = Ep - v[pp], zlo_r = Er + v[r], zhi_r = Er - v[r];
        const cpx far_r = mk(__shfl_sync(0xffffffffu, zhi_r.x, lp), __shfl_sync(0xffffffffu, zhi_r.y, lp));
        const cpx far_p = mk(__shfl_sync(0xffffffffu, zhi_p.x, lp), __shfl_sync(0xffffffffu, zhi_p.y, lp));
        // branch-free: every lane runs the split (lane 0's results are garbage and stay unstored: its bins 32*pp pair
        // with its OWN registers and are shared out over the warp below); the stores carry a predicate and no memory
        // clobber, so the loads and shuffles of the next pair can be scheduled above them
        cpx x1, x2, y1, y2;
        split_fwd(zlo_p, far_r, s_qs[pp * 32 + lane], x1, x2);
        split_fwd(zlo_r, far_p, s_qs[r * 32 + lane], y1, y2);
        stg_stream_if(lane != 0, spec + lane + 32 * pp, x1);
        stg_stream_if(lane != 0, spec + C - lane - 32 * pp, x2);
        stg_stream_if(lane != 0, spec + lane + 32 * r, y1);
        stg_stream_if(lane != 0, spec + C - lane - 32 * r, y2);
        if (lane == 0) {
          ex[pp] = zlo_p; ex[r] = zlo_r;
          ex[32 + pp] = zhi_p; ex[32 + r] = zhi_r;
        }
        if (pp == 0) zhi0 = zhi_p;
      }
      __syncwarp();
      if (lane != 0) {                                       // bins 32*l and 2048 - 32*l: Z[32 l] with Z[1024 + 32 (32 - l)]
        cpx x1, x2;
        split_fwd(ex[lane], ex[32 + 32 - lane], qcoop, x1, x2);
        stg_stream(spec + 32 * lane, x1);
        stg_stream(spec + C - 32 * lane, x2);
      } else {
        const cpx z0 = ex[0];
        stg_stream(spec, mk(z0.x + z0.y, 0.0f));             // exact DC and Nyquist (kiss_fftr.c:93-99)
        stg_stream(spec + C, mk(z0.x - z0.y, 0.0f));
        stg_stream(spec + C / 2, mul2(zhi0, mk(1.f, -1.f))); // X[1024] = conj(Z[1024])
      }
      __syncwarp();                                          // ex is the next transform's exchange buffer
      spec += C + 1;
    }
  }
  tmem_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(s_tmem_base);
}

}  // namespace

cudaError_t launch_stft_fwd_4096v2(const K1Params& p_in, cudaStream_t st) {
  K1Params p = p_in;
  // Short chunks: a frame takes ~10 us of a warp, so a 64-frame chunk is a 0.7 ms tail on a 6 ms kernel (32: half); the price of a
  // chunk start is three extra hops read (from L2: the neighbouring chunk has just used them).
  p.chunk_steps = 64;                                      // big chunks; the last quarter of a channel goes in chunks of 16
  const int mode = !p.scrub ? 0 : p.scrub == 2 ? 2 : 1;
  auto kern = mode == 0 ? stft_fwd_kernel_4096v2<0> : mode == 1 ? stft_fwd_kernel_4096v2<1> : stft_fwd_kernel_4096v2<2>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kV2Smem);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(p.counter, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return e;
  const long long cb = p.chunk_steps, cbs = cb >= 16 ? cb / 4 : cb;      // as in the kernel
  const long long big_per_chan = (p.frames * 3 / 4) / cb;
  const long long total_chunks = (big_per_chan + (p.frames - big_per_chan * cb + cbs - 1) / cbs) * p.channels;
  const long long want_blocks = (total_chunks + kV2Warps - 1) / kV2Warps;
  const int grid = static_cast<int>(want_blocks < device_sm_count() ? want_blocks : device_sm_count());
  kern<<<grid > 0 ? grid : 1, kV2Warps * 32, kV2Smem, st>>>(p);
  return cudaGetLastError();
}

}  // namespace crlot
