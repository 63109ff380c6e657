// k2_4096v2.cu -- kernel 2 for N = 4096 as TWO 1024-point complex sub-transforms per frame.
//
// Replaces, per frame, KissFftPlan::inverse -> kiss_fftri (reference: dsp/fft/backends/kissfft_adapter.cc:124-168,
// third_party/kissfft/kiss_fftr.c:119-155) and the `src*win` half of axpy_windowed (dsp/ola/kernels.cc:24-28), like
// k2_istft.cu.  The first version (istft_inv_kernel_4096, k2_istft.cu) spends 64 complex registers per lane: 240
// registers, 8 warps per SM, two per scheduler -- latency-bound at 0.79-0.83 of the HBM roofline.  This version is the
// mirror image of k1_4096v2.cu on the 32-register machinery (12 warps per SM); the arithmetic is in warpfft4096v2.cuh:
//
//   quads {k, 1024-k, 1024+k, 2048-k} -> e'[k] = Z'[k] + Z'[k+1024],  o'[k] = (Z'[k] - Z'[k+1024]) exp(+2 pi i k/2048)
//   e = IFFT_1024(e'), o = IFFT_1024(o');  x[4m], x[4m+1] = e[m];  x[4m+2], x[4m+3] = o[m]   (m = lane + 32*oa)
//
// * a lane loads X[k] and X[k+1024], k = lane + 32*b, straight into registers (contiguous 8 bytes per lane), eight
//   loads at a time, the next eight in flight while the current quads are computed; the next row is pulled into L2
//   a frame ahead.  X[1024-k] and X[2048-k] are the mirror lane's loads at register 31-b: two complex shuffles;
// * lane 0's bins 32*b pair with its own registers: lane l >= 1 computes the quad of k = 32*l (four extra loads)
//   and hands e', o' over through 64 words of the exchange buffer;
// * e' stays in registers, o' is parked in TENSOR MEMORY (64 columns per warp); after the first sub-transform the
//   finished e (scaled, cleaned, windowed) and the parked o' swap places, eight registers at a time; after the second
//   the warp stores float4 = (e[m], o[m]): 16 contiguous bytes per lane;
// * constants in tensor memory as in k1_stft.cu: inter-pass twiddles (64 columns), split factor + combine twiddle per
//   register in the order the main loop consumes them (128), synthesis window planes (2 x 64); 3 warps x 64 parked.
//   No shared memory but the 8.4 KB exchange buffer per warp.
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "warpfft.cuh"
#include "warpfft4096v2.cuh"

namespace crlot {

namespace {

constexpr int kV2Warps = 12;
using XV = XV4096;                                       // the 1024-point complex sub-transform: 32 x 32
constexpr int kV2Ring = 5;                               // staged chunks per warp
constexpr int kV2Seg = 66 * 8;                           // one staged segment: 64 bins + 16 bytes (rows are only 8-byte aligned)
constexpr int kV2Slot = 4 * kV2Seg;                      // a chunk: [lo P0 P1][hi P0 P1][lo R1 R0][hi R1 R0]
// exchange buffer + staging ring + bin 2048 (16 bytes) + one mbarrier per ring slot
constexpr int kV2PerWarp = XV::XCPX * 8 + kV2Ring * kV2Slot + 16 + 48;
static_assert(kV2PerWarp % 16 == 0, "bulk-copy destinations are 16-byte aligned");
constexpr int kV2Smem = kV2Warps * kV2PerWarp;

__device__ __forceinline__ void cp_async8(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// One sub-transform of the frame: h = 0 transforms e' (in the registers), h = 1 first swaps the finished e with the
// parked o'.  Ends with the scale, the clean-up rule and the synthesis window of the plane's samples.
template <bool WIN, bool SCRUB>
__device__ __forceinline__ void sub_transform(cpx (&v)[32], const int h, int lane, cpx* ex, uint32_t ttw, uint32_t tpark,
                                              uint32_t twin, float scale) {
  if (h == 1) {                                        // finished e <-> parked o', eight registers at a time
    cpx tb[2][8];                                      // the next parked chunk in flight while this one swaps
    tmem_ld8(tpark, tb[0]);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      cpx e8[8];
      tmem_wait_ld8(tb[c & 1]);
      if (c + 1 < 4) tmem_ld8(tpark + 16 * (c + 1), tb[(c + 1) & 1]);
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        e8[s] = v[v2_park_reg(c, s)];
        v[v2_park_reg(c, s)] = tb[c & 1][s];
      }
      tmem_st8(tpark + 16 * c, e8);
    }
  }
  ifft1024(v, lane, ex, ttw);
  cpx wb[2][8];                                        // window constants: the next chunk in flight while this one multiplies
  if (WIN) tmem_ld8(twin + 64 * h, wb[0]);
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    if (WIN) {
      tmem_wait_ld8(wb[c & 1]);
      if (c + 1 < 4) tmem_ld8(twin + 64 * h + 16 * (c + 1), wb[(c + 1) & 1]);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int oa = 8 * c + i;
      cpx y = mul2(v[oa], splat(scale));              // exact: a power of two
      if (SCRUB) y = mk(scrub(y.x), scrub(y.y));
      if (WIN) y = mul2(y, wb[c & 1][i]);              // one rounding per sample, as std::fma(src, win, 0.0f) rounds it
      v[oa] = y;
    }
  }
}

// (Two copies of the sub-transform in the instruction stream instead of the two-trip loop were measured: the unrolled
// kernel spills and runs 8.2 ms against 6.85 ms at that stage of the work.)
// GAIN: the spectral-domain hook is compiled in (its loads and multiplies would otherwise ride along predicated off)
template <bool WIN, bool SCRUB, bool GAIN>
__global__ void __launch_bounds__(kV2Warps * 32, 1) istft_inv_kernel_4096v2(const K2Params p) {
  constexpr int NFFT = 4096, C = 2048;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint32_t s_tmem_base;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform by construction
  cpx* ex = reinterpret_cast<cpx*>(smem_raw + warp * kV2PerWarp);
  unsigned char* stage = reinterpret_cast<unsigned char*>(ex + XV::XCPX);
  cpx* nyq = reinterpret_cast<cpx*>(stage + kV2Ring * kV2Slot);
  const uint32_t bar0 = smem_u32(stage + kV2Ring * kV2Slot + 16);          // mbarrier of ring slot s: bar0 + 8 s
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < kV2Ring; ++s) mbar_init(bar0 + 8 * s, 1);
    fence_mbar_init();
  }
  __syncwarp();

  if (warp == 0) tmem_alloc<512>(&s_tmem_base);
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();
  const uint32_t tmb = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  const uint32_t ttw = tmb, tqc = tmb + 64, twin = tmb + 192;
  const uint32_t tpark = tmb + 320 + 64 * (warp >> 2);
  if (warp < 4) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      cpx a8[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) a8[i] = p.sa[(8 * c + i) * 32 + lane];           // W_1024^(lane*ob)
      tmem_st8(ttw + 16 * c, a8);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      cpx c8[8];
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const int k = lane + 32 * v2_chunk_reg(j, s);      // tables are indexed by the bin (tables.h)
        c8[s] = p.sa[2048 + k];                            // (-i/2) exp(-i pi k/2048)
        c8[4 + s] = p.sa[1024 + k];                        // exp(-2 pi i k/2048)
      }
      tmem_st8(tqc + 16 * j, c8);
    }
    if (WIN) {
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          cpx w8[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) w8[i] = *reinterpret_cast<const cpx*>(p.window + 4 * (lane + 32 * (8 * c + i)) + 2 * h);
          tmem_st8(twin + 64 * h + 16 * c, w8);
        }
    }
    tmem_wait_st();
  }
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();

  const int z0 = lane == 0 ? 1 : 0;
  const long long stride = static_cast<long long>(gridDim.x) * kV2Warps;
  const float scale = 2.0f / static_cast<float>(NFFT);    // the quads hand over halved values
  const uint32_t stg = smem_u32(stage);
  uint32_t ph = 0;                                        // phase parity of the ring slots' mbarriers (bit s)

  // Chunk j of a row = the bin-rows (32 bins each) of registers P0 = 2j, P1 = 2j+1, R0 = 31-2j, R1 = 30-2j, lo = bins
  // 32b.., hi = bins 1024+32b..: four segments of 64 contiguous bins, each moved by ONE bulk async copy (UBLKCP) that
  // starts at the 16-byte boundary at or before the segment (off = 1 when the row starts 8 bytes past one: element e of
  // a staged segment is bin first + e - off).  Ring slot j % 5; the copy completes on the slot's mbarrier.
  // One ELECTED lane issues the copies (elect.sync): with the warp index broadcast by a shuffle every operand is
  // warp-uniform, so each copy is a single UBLKCP on uniform registers -- issued under `if (lane == 0)` the compiler
  // wraps every copy in a broadcast loop.  No proxy fence: the ring is only ever READ through the generic proxy, and
  // those reads are ordered before the refill by the __syncwarp() that precedes it.
  auto issue_chunk = [&](const cpx* row, const int j) {
    if (elect_one()) {
      const uint32_t off = static_cast<uint32_t>((reinterpret_cast<uintptr_t>(row) >> 3) & 1);
      const uint32_t bytes = 512 + 16 * off;
      const uint32_t bar = bar0 + 8 * (j % kV2Ring), dst = stg + (j % kV2Ring) * kV2Slot;
      const cpx* src = row - off;
      mbar_arrive_expect_tx(bar, 4 * bytes);
      bulk_g2s(dst, src + 64 * j, bytes, bar);
      bulk_g2s(dst + kV2Seg, src + 1024 + 64 * j, bytes, bar);
      bulk_g2s(dst + 2 * kV2Seg, src + 32 * (30 - 2 * j), bytes, bar);
      bulk_g2s(dst + 3 * kV2Seg, src + 1024 + 32 * (30 - 2 * j), bytes, bar);
      if (j == 0) {
        cp_async8(smem_u32(nyq), row + C);
        cp_async_commit();
      }
    }
  };
  auto wait_chunk = [&](const int j) {
    mbar_wait(bar0 + 8 * (j % kV2Ring), (ph >> (j % kV2Ring)) & 1u);
    ph ^= 1u << (j % kV2Ring);
  };

  long long rowi = static_cast<long long>(blockIdx.x) * kV2Warps + warp;
  if (rowi < p.rows) {
#pragma unroll
    for (int j = 0; j < kV2Ring; ++j) issue_chunk(p.spec + rowi * (C + 1), j);
  }
  for (; rowi < p.rows; rowi += stride) {
    const cpx* row = p.spec + rowi * (C + 1);
    const float* gr = GAIN ? p.sgain + rowi * p.sgain_row_stride : nullptr;      // spectral-domain hook
    const int off = static_cast<int>((reinterpret_cast<uintptr_t>(row) >> 3) & 1);
    const int own8 = (lane + off) * 8;                     // byte offset of this lane's bin in a staged bin-row
    const int mir8 = ((lane ? 32 - lane : 0) + off) * 8;   // ... of the mirror lane's bin (lane 0: its own column)
    cpx v[32];
    cpx o8[8];
    wait_chunk(0);
    cp_async_wait_all();                                   // bin 2048: the elected lane's copy ...
    __syncwarp();                                          // ... made visible to the warp
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (j < 7) wait_chunk(j + 1);                        // lane 0 reads its partners in the neighbouring chunks
      cpx k8[8];
      tmem_ld8(tqc + 16 * j, k8);
      const int jp = (j + kV2Ring - 1) % kV2Ring, jn = (j + 1) % kV2Ring, jc = j % kV2Ring;
      const unsigned char* sb = stage + jc * kV2Slot + own8;
      cpx d[8];                                            // [lo P0, lo P1, hi P0, hi P1, lo R0, lo R1, hi R0, hi R1]
      d[0] = *reinterpret_cast<const cpx*>(sb);
      d[1] = *reinterpret_cast<const cpx*>(sb + 256);
      d[2] = *reinterpret_cast<const cpx*>(sb + kV2Seg);
      d[3] = *reinterpret_cast<const cpx*>(sb + kV2Seg + 256);
      d[4] = *reinterpret_cast<const cpx*>(sb + 2 * kV2Seg + 256);
      d[5] = *reinterpret_cast<const cpx*>(sb + 2 * kV2Seg);
      d[6] = *reinterpret_cast<const cpx*>(sb + 3 * kV2Seg + 256);
      d[7] = *reinterpret_cast<const cpx*>(sb + 3 * kV2Seg);
      // the partner bin-rows: lanes >= 1 find register b's partners in bin-row 31-b of the same chunk at column 32-lane;
      // lane 0's bins 32*b pair with bin-row 32-b at column 0: the previous chunk's R1 for P0, R0 for P1, P1 for R0, the
      // next chunk's P0 for R1 (chunk 7: its own R1, the self-mirrored bin 512)
      constexpr int kR0 = 2 * kV2Seg + 256, kR1 = 2 * kV2Seg, kP0 = 0, kP1 = 256;   // lo bin-rows inside a slot
      const int g_p0 = jc * kV2Slot + kR0, l0_p0 = jp * kV2Slot + kR1;
      const int g_p1 = jc * kV2Slot + kR1, l0_p1 = jc * kV2Slot + kR0;
      const int g_r0 = jc * kV2Slot + kP0, l0_r0 = jc * kV2Slot + kP1;
      const int g_r1 = jc * kV2Slot + kP1, l0_r1 = j < 7 ? jn * kV2Slot + kP0 : jc * kV2Slot + kR1;
      const unsigned char* sm = stage + mir8;
      const unsigned char* a_p0 = sm + g_p0 + z0 * (l0_p0 - g_p0);
      const unsigned char* a_p1 = sm + g_p1 + z0 * (l0_p1 - g_p1);
      const unsigned char* a_r0 = sm + g_r0 + z0 * (l0_r0 - g_r0);
      const unsigned char* a_r1 = sm + g_r1 + z0 * (l0_r1 - g_r1);
      cpx f[8];                                            // partners of [P0, P1, R0, R1]: lo, then hi (the next segment)
      f[0] = *reinterpret_cast<const cpx*>(a_p0); f[4] = *reinterpret_cast<const cpx*>(a_p0 + kV2Seg);
      f[1] = *reinterpret_cast<const cpx*>(a_p1); f[5] = *reinterpret_cast<const cpx*>(a_p1 + kV2Seg);
      f[2] = *reinterpret_cast<const cpx*>(a_r0); f[6] = *reinterpret_cast<const cpx*>(a_r0 + kV2Seg);
      f[3] = *reinterpret_cast<const cpx*>(a_r1); f[7] = *reinterpret_cast<const cpx*>(a_r1 + kV2Seg);
      cpx xn = mk(0.f, 0.f);
      if (j == 0) xn = *nyq;
      if (j >= 1 && j + 4 < 8) {                           // refill the slot of chunk j-1: lane 0 has just read it for the last time
        __syncwarp();
        issue_chunk(row, j + 4);
      }
      if (GAIN) {
        // bins: own lo = lane + 32 b, hi = +1024; partner lo = 1024 - (own lo), hi = 2048 - (own lo)
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int b = v2_chunk_reg(j, t);
          const int own = lane + 32 * b;
          const int s_lo = (t & 1) + ((t & 2) << 1), s_hi = s_lo + 2;
          d[s_lo] = mul2(d[s_lo], splat(__ldg(gr + own)));
          d[s_hi] = mul2(d[s_hi], splat(__ldg(gr + 1024 + own)));
          f[t] = mul2(f[t], splat(__ldg(gr + 1024 - own)));
          f[4 + t] = mul2(f[4 + t], splat(__ldg(gr + 2048 - own)));
        }
        if (j == 0) xn = mul2(xn, splat(__ldg(gr + C)));
      }
      tmem_wait_ld8(k8);
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int pp = 2 * j + i, r = 31 - pp;
        inv4096v2_quad(d[i], f[i], d[2 + i], f[4 + i], k8[i], k8[4 + i], v[pp], o8[4 * (j & 1) + i]);
        inv4096v2_quad(d[4 + i], f[2 + i], d[6 + i], f[6 + i], k8[2 + i], k8[6 + i], v[r], o8[4 * (j & 1) + 2 + i]);
      }
      if (j == 0) {                                        // bin 0: real parts of X[0] and X[2048] only (kiss_fftr.c:131-132)
        cpx e0, o0;
        inv4096v2_quad0(d[0], d[2], xn, e0, o0);
        if (lane == 0) {
          v[0] = e0;
          o8[0] = o0;
        }
      }
      if (j & 1) tmem_st8(tpark + 16 * (j >> 1), o8);
    }
    tmem_wait_st();
    __syncwarp();                                          // every lane has read the last chunks
    if (rowi + stride < p.rows) {                          // the next row's first chunks land while this one is transformed
#pragma unroll
      for (int j = 0; j < kV2Ring; ++j) issue_chunk(p.spec + (rowi + stride) * (C + 1), j);
    }

    // ---- the two sub-transforms: ONE copy of the code, run for e' and then for o' -----------------------------------
#pragma unroll 1
    for (int h = 0; h < 2; ++h) sub_transform<WIN, SCRUB>(v, h, lane, ex, ttw, tpark, twin, scale);
    tmem_wait_st();
    float* dst = p.frames + rowi * NFFT;
    cpx eb[2][8];                                          // the parked e: the next chunk in flight while this one is stored
    tmem_ld8(tpark, eb[0]);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      tmem_wait_ld8(eb[c & 1]);
      if (c + 1 < 4) tmem_ld8(tpark + 16 * (c + 1), eb[(c + 1) & 1]);
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        const int rg = v2_park_reg(c, s);
        stg_stream4(reinterpret_cast<float4*>(dst + 4 * (lane + 32 * rg)),
                    make_float4(eb[c & 1][s].x, eb[c & 1][s].y, v[rg].x, v[rg].y));
      }
    }
  }
  tmem_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(s_tmem_base);
}

}  // namespace

cudaError_t launch_istft_inv_4096v2(const K2Params& p, cudaStream_t st) {
  const bool win = p.window != nullptr;
  void (*kern)(const K2Params) = nullptr;
  const int sel = (win ? 4 : 0) | (p.scrub ? 2 : 0) | (p.sgain != nullptr ? 1 : 0);
  switch (sel) {
#define CRLOT_K2V2_CASE(n) case n: kern = istft_inv_kernel_4096v2<((n) & 4) != 0, ((n) & 2) != 0, ((n) & 1) != 0>; break;
    CRLOT_K2V2_CASE(0) CRLOT_K2V2_CASE(1) CRLOT_K2V2_CASE(2) CRLOT_K2V2_CASE(3)
    CRLOT_K2V2_CASE(4) CRLOT_K2V2_CASE(5) CRLOT_K2V2_CASE(6) CRLOT_K2V2_CASE(7)
#undef CRLOT_K2V2_CASE
  }
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kV2Smem);
  if (e != cudaSuccess) return e;
  const long long want_blocks = (p.rows + kV2Warps - 1) / kV2Warps;
  const int grid = static_cast<int>(want_blocks < device_sm_count() ? want_blocks : device_sm_count());
  kern<<<grid > 0 ? grid : 1, kV2Warps * 32, kV2Smem, st>>>(p);
  return cudaGetLastError();
}

}  // namespace crlot
