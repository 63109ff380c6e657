// tables.h -- host-side (double precision) construction of the constant tables the kernels
// read from shared memory.  Used by capi.cu (uploaded once per plan) and by tests/cpu_emu.
#pragma once
#include <cmath>
#include <vector>

namespace crlot {

struct FftTables {
  int nfft = 0, C = 0, A = 0, B = 0;
  std::vector<float> tw1;   // [B][A] cpx : exp(-2 pi i a*ob / C)      inter-pass twiddle (forward)
  std::vector<float> sa;    // [C]    cpx : 0.5*(1 + exp(-i pi (k/C + 1/2)))  real split (inverse uses conj)
  std::vector<float> pre;   // N=4096 only: [32][64] cpx : exp(-2 pi i a*ob / 2048), the inverse kernel's
                            // transposed inter-pass table twi[ob*64 + a] (warpfft4096.cuh)
};

// C = complex length of the real split (nfft/2) = complex length one warp transforms.
// nfft = 4096: forward uses A=32, B=64 (tw1), inverse A=64, B=32 (pre).
inline FftTables make_fft_tables(int nfft) {
  const double pi = 3.141592653589793238462643383279502884;
  FftTables t;
  t.nfft = nfft;
  t.C = nfft / 2;
  const int Cw = t.C;
  switch (Cw) {
    case 2048: t.A = 32; t.B = 64; break;
    case 1024: t.A = 32; t.B = 32; break;
    case 512:  t.A = 32; t.B = 16; break;
    case 256:  t.A = 16; t.B = 16; break;
    default:   t.A = 16; t.B = 8;  break;
  }
  t.tw1.resize(2 * (size_t)t.A * t.B);
  for (int ob = 0; ob < t.B; ++ob)
    for (int a = 0; a < t.A; ++a) {
      const double ph = -2.0 * pi * (double)a * (double)ob / (double)Cw;
      t.tw1[2 * (ob * t.A + a)] = (float)std::cos(ph);
      t.tw1[2 * (ob * t.A + a) + 1] = (float)std::sin(ph);
    }
  t.sa.resize(2 * (size_t)t.C);
  for (int k = 0; k < t.C; ++k) {
    const double ph = -pi * ((double)k / (double)t.C + 0.5);
    t.sa[2 * k] = (float)(0.5 * (1.0 + std::cos(ph)));
    t.sa[2 * k + 1] = (float)(0.5 * std::sin(ph));
  }
  if (nfft == 4096) {
    // Tables of the two-sub-transform kernels (k1_4096v2.cu): a 4096-point frame = 2048 complex points = two
    // 1024-point complex transforms (even / odd) + a radix-2 combine + the real split.  `sa` is not read by the
    // register kernels of this size, so it carries them: [0,1024) tw[ob*32+a] = exp(-2 pi i a*ob/1024) (inter-pass
    // twiddles of a 1024-point transform), [1024,2048) tc[oa*32+l] = exp(-2 pi i k/2048), k = l + 32*oa (combine),
    // [2048,3072) qs[p*32+l] = (-i/2) exp(-i pi k/2048), k = l + 32*p (forward real split of the 2048-point result).
    t.sa.assign(2 * 3072, 0.0f);
    for (int ob = 0; ob < 32; ++ob)
      for (int a = 0; a < 32; ++a) {
        const double ph = -2.0 * pi * (double)a * (double)ob / 1024.0;
        t.sa[2 * (ob * 32 + a)] = (float)std::cos(ph);
        t.sa[2 * (ob * 32 + a) + 1] = (float)std::sin(ph);
      }
    for (int k = 0; k < 1024; ++k) {       // k = l + 32*oa is stored at index oa*32 + l = k's own digits swapped... both are k
      const int l = k % 32, oa = k / 32;
      const double pc = -2.0 * pi * (double)k / 2048.0, ps = pi * (double)k / 2048.0;
      t.sa[2 * (1024 + oa * 32 + l)] = (float)std::cos(pc);
      t.sa[2 * (1024 + oa * 32 + l) + 1] = (float)std::sin(pc);
      t.sa[2 * (2048 + oa * 32 + l)] = (float)(-0.5 * std::sin(ps));
      t.sa[2 * (2048 + oa * 32 + l) + 1] = (float)(-0.5 * std::cos(ps));
    }
    t.pre.resize(2 * 2048);
    for (int ob = 0; ob < 32; ++ob)
      for (int a = 0; a < 64; ++a) {
        const double ph = -2.0 * pi * (double)a * (double)ob / 2048.0;
        t.pre[2 * (ob * 64 + a)] = (float)std::cos(ph);
        t.pre[2 * (ob * 64 + a) + 1] = (float)std::sin(ph);
      }
  }
  return t;
}

// Tables of the any-size kernels (k_generic.cu) for a real transform of nfft points (complex length
// M = nfft/2): tw1[k] = exp(-2 pi i k / M), k < M, and sa[k] = exp(-i pi (k/M + 1/2)), k < M -- the
// split factor kiss_fftr calls super_twiddles[k-1] (reference: third_party/kissfft/kiss_fftr.c:49-56).
// For a complex plan of n points pass nfft = 2n and use tw1 only.
inline FftTables make_generic_tables(int nfft) {
  const double pi = 3.141592653589793238462643383279502884;
  FftTables t;
  t.nfft = nfft;
  t.C = nfft / 2;
  const int M = t.C > 0 ? t.C : 1;
  t.tw1.resize(2 * (size_t)M);
  t.sa.resize(2 * (size_t)M);
  for (int k = 0; k < M; ++k) {
    const double ph = -2.0 * pi * (double)k / (double)M;
    t.tw1[2 * k] = (float)std::cos(ph);
    t.tw1[2 * k + 1] = (float)std::sin(ph);
    const double ps = -pi * ((double)k / (double)M + 0.5);
    t.sa[2 * k] = (float)std::cos(ps);
    t.sa[2 * k + 1] = (float)std::sin(ps);
  }
  return t;
}

}  // namespace crlot
