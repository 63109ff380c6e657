// warpfft4096v2.cuh -- per-lane pieces of the INVERSE 4096-point real transform as two 1024-point complex
// sub-transforms on the 32-register machinery (kernel 2 for N = 4096, k2_4096v2.cu, and the fused kernels 2+3,
// k23_4096.cu).  Mirror image of k1_4096v2.cu; restates kiss_fftri (reference: third_party/kissfft/kiss_fftr.c:119-155)
// followed by a radix-2 decimation in frequency of the 2048-point inverse complex transform:
//
//   Z'[k]   = inverse real split of X[k], X[2048-k]                      (k < 2048; halved values, see warpfft.cuh)
//   e'[k]   = Z'[k] + Z'[k+1024]                                         (k < 1024)
//   o'[k]   = (Z'[k] - Z'[k+1024]) * exp(+2 pi i k / 2048)
//   e = IFFT_1024(e'),  o = IFFT_1024(o')   (unnormalised);   z[2m] = e[m], z[2m+1] = o[m]
//   x[4m], x[4m+1] = e[m] * 2/N;   x[4m+2], x[4m+3] = o[m] * 2/N
//
// The four bins {k, 1024-k, 1024+k, 2048-k} form a closed group ("quad"): a lane that owns k = lane + 32*b loads
// X[k] and X[k+1024] itself and receives X[1024-k] and X[2048-k] from the MIRROR lane 32-lane, where they are the
// same two loads at register 31-b -- two complex shuffles per register, no table of mirrored bins, never more
// than 32 complex values live.  Lane 0's bins 32*b pair with its own registers; lane l >= 1 computes the quad of
// k = 32*l for it (four extra loads) and hands the result over through shared memory; k = 0 uses only the real
// parts of X[0] and X[2048] (kiss_fftr.c:131-132).
//
// Everything here is a function of (lane, registers): tests/cpu_emu runs it lane by lane on the CPU.
#pragma once
#include "warpfft.cuh"
#if defined(__CUDACC__)
#include "common.cuh"
#endif

namespace crlot {

// e'[k], o'[k] from X0 = X[k], X1 = X[1024-k], X2 = X[1024+k], X3 = X[2048-k];
// qs = (-i/2) exp(-i pi k/2048) (the forward split factor: the inverse one is its conjugate),
// tc = exp(-2 pi i k/2048) (the forward combine twiddle, conjugated here).
CRLOT_HD void inv4096v2_quad(cpx X0, cpx X1, cpx X2, cpx X3, cpx qs, cpx tc, cpx& e, cpx& o) {
  const cpx s0 = fma2(X3, mk(1.f, -1.f), X0);                 // X[k] + conj X[2048-k]
  const cpx d0 = fma2(X3, mk(-1.f, 1.f), X0);                 // X[k] - conj X[2048-k]
  const cpx s1 = fma2(X1, mk(1.f, -1.f), X2);                 // X[k+1024] + conj X[1024-k]
  const cpx d1 = fma2(X1, mk(-1.f, 1.f), X2);
  const cpx z0 = fma2(s0, mk(0.5f, 0.5f), cmulc(d0, qs));     // Z'[k]/2
  // Z'[k+1024]/2: its factor is i times that of k, folded into the operand
  const cpx z1 = fma2(s1, mk(0.5f, 0.5f), cmulc(rot90<true>(d1), qs));
  e = z0 + z1;
  o = cmulc(z0 - z1, tc);
}
// k = 0: X[0] and X[2048] contribute their real parts only; Z'[1024]/2 = conj X[1024]
CRLOT_HD void inv4096v2_quad0(cpx X0, cpx Xm, cpx Xn, cpx& e, cpx& o) {
  const cpx z0 = mk(0.5f * (X0.x + Xn.x), 0.5f * (X0.x - Xn.x));
  const cpx z1 = mk(Xm.x, -Xm.y);
  e = z0 + z1;
  o = z0 - z1;
}

// The main loop walks the 32 registers in 8 chunks of two mirrored pairs: chunk j holds pp = 2j, 2j+1 and
// r = 31-2j, 30-2j.  Slot s < 4 of a chunk: s < 2 -> register 2j+s, else register 31-(2j+s-2).
CRLOT_HD constexpr int v2_chunk_reg(int j, int s) { return s < 2 ? 2 * j + s : 31 - (2 * j + (s - 2)); }
// Parked values (tensor memory, 8 per tcgen05 access): park chunk c < 4, slot s < 8 = chunk 2c + (s >> 2), slot s & 3
CRLOT_HD constexpr int v2_park_reg(int c, int s) { return v2_chunk_reg(2 * c + (s >> 2), s & 3); }

using XV4096 = GeoX<2048>;   // the 1024-point complex sub-transform: 32 x 32 (exchange buffer: XV4096::XCPX values)

#if defined(__CUDACC__)
// inter-pass twiddles W_1024^(lane*ob), ob < 32, in 64 tensor-memory columns at `ttw` (tables.h: sa[ob*32 + lane])
// one inverse 1024-point complex transform of the 32 registers (pass-1 view in, pass-2 view out)
__device__ __forceinline__ void ifft1024(cpx (&v)[32], int lane, cpx* ex, uint32_t ttw) {
  cpx t0[8];
  tmem_ld8(ttw, t0);
  pass1<XV4096, true>(v);
  cpx tb[2][8];
#pragma unroll
  for (int i = 0; i < 8; ++i) tb[0][i] = t0[i];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    tmem_wait_ld8(tb[c & 1]);
    if (c + 1 < 4) tmem_ld8(ttw + 16 * (c + 1), tb[(c + 1) & 1]);
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (8 * c + i > 0) v[8 * c + i] = cmulc(v[8 * c + i], tb[c & 1][i]);
  }
  xchg_write<XV4096>(v, lane, ex);
  __syncwarp();
  xchg_read<XV4096>(v, lane, ex);
  __syncwarp();
  pass2<XV4096, true>(v);
}
#endif

}  // namespace crlot
