// tests/cpu_emu/emu_warpfft.cc -- runs the per-lane code of csrc/warpfft.cuh lane by lane on
// the CPU (a warp = 32 sequential lanes per phase; a shuffle = reading another lane's array)
// and compares with a double-precision DFT.  Checks the index algebra (four-step mapping,
// exchange addressing incl. bank-conflict freedom, mirror lanes) without a GPU.
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <set>
#include <vector>

#include "../../crlot-dsp_b200/csrc/tables.h"
#include "../../crlot-dsp_b200/csrc/warpfft.cuh"
#include "../../crlot-dsp_b200/csrc/warpfft4096.cuh"
#include "../../crlot-dsp_b200/csrc/warpfft4096v2.cuh"

using namespace crlot;
typedef std::complex<double> cd;

static double frand() { return 2.0 * rand() / (double)RAND_MAX - 1.0; }

template <int NFFT> static int run() {
  typedef GeoX<NFFT> X;
  const int C = X::C, G = X::G, N = NFFT;
  FftTables T = make_fft_tables(N);
  const cpx* tw1 = reinterpret_cast<const cpx*>(T.tw1.data());
  std::vector<std::vector<float>> x(G, std::vector<float>(N));
  for (auto& f : x) for (auto& s : f) s = (float)frand();
  // reference spectra in double
  std::vector<std::vector<cd>> Xr(G, std::vector<cd>(C + 1));
  for (int g = 0; g < G; ++g)
    for (int k = 0; k <= C; ++k) {
      cd s = 0;
      for (int n = 0; n < N; ++n) s += (double)x[g][n] * std::polar(1.0, -2.0 * M_PI * (double)n * k / N);
      Xr[g][k] = s;
    }
  // ---------------- forward ----------------
  std::vector<cpx> ex(X::XCPX);
  cpx v[32][32];
  for (int l = 0; l < 32; ++l)
    for (int r = 0; r < 32; ++r) {
      int g, i; in_coord<X>(l, r, g, i);
      v[l][r] = mk(x[g][2 * i], x[g][2 * i + 1]);
    }
  for (int l = 0; l < 32; ++l) { pass1<X, false>(v[l]); twiddle1<X, false>(v[l], l, tw1); }
  // bank-conflict check on the exchange pattern (8-byte accesses: per half-warp 16 distinct pairs)
  int conflicts = 0;
  for (int j = 0; j < X::M1; ++j) for (int ob = 0; ob < X::B; ++ob) for (int h = 0; h < 2; ++h) {
    std::set<int> banks;
    for (int l = 16 * h; l < 16 * h + 16; ++l) { int i1 = l + 32 * j; banks.insert(X::xaddr(i1 / X::A, i1 % X::A, ob) % 16); }
    if (banks.size() != 16) ++conflicts;
  }
  for (int j = 0; j < X::M2; ++j) for (int a = 0; a < X::A; ++a) for (int h = 0; h < 2; ++h) {
    std::set<int> banks;
    for (int l = 16 * h; l < 16 * h + 16; ++l) { int i2 = l + 32 * j; banks.insert(X::xaddr(i2 / X::B, a, i2 % X::B) % 16); }
    if (banks.size() != 16) ++conflicts;
  }
  for (int l = 0; l < 32; ++l) xchg_write<X>(v[l], l, ex.data());
  for (int l = 0; l < 32; ++l) xchg_read<X>(v[l], l, ex.data());
  for (int l = 0; l < 32; ++l) pass2<X, false>(v[l]);
  cpx snap[32][32];
  for (int l = 0; l < 32; ++l) for (int r = 0; r < 32; ++r) snap[l][r] = v[l][r];
  cpx nyq[32][X::M2];
  cpx mid[32][X::M2];
  for (int l = 0; l < 32; ++l) {                       // phase A: mirror lane's registers before the phase
    const int lp = mirror_lane<X::B>(l);
    fwd_split_pairs<X>(v[l], l, make_split_fwd<X>(l), [&](int r) { return snap[lp][r]; }, nyq[l], mid[l]);
  }
  for (int l = 0; l < 32; ++l) for (int r = 0; r < 32; ++r) snap[l][r] = v[l][r];
  for (int l = 0; l < 32; ++l) {                       // phase B: mirror lane's registers after phase A
    const int lp = mirror_lane<X::B>(l);
    fwd_split_finish<X>(v[l], l, [&](int r) { return snap[lp][r]; }, mid[l]);
  }
  double err = 0, nrm = 0;
  std::vector<std::vector<cd>> Xg(G, std::vector<cd>(C + 1));
  for (int l = 0; l < 32; ++l)
    for (int r = 0; r < 32; ++r) {
      int g, o; out_coord<X>(l, r, g, o);
      Xg[g][o] = cd(v[l][r].x, v[l][r].y);
      if (o == 0) Xg[g][C] = cd(nyq[l][r / X::A].x, nyq[l][r / X::A].y);
    }
  for (int g = 0; g < G; ++g) for (int k = 0; k <= C; ++k) { err += std::norm(Xg[g][k] - Xr[g][k]); nrm += std::norm(Xr[g][k]); }
  const double fwd_rel = std::sqrt(err / nrm);
  int exact0 = 1;
  for (int g = 0; g < G; ++g) if (Xg[g][0].imag() != 0.0 || Xg[g][C].imag() != 0.0) exact0 = 0;
  // ---------------- inverse (from the double-precision spectrum rounded to float) ----------------
  cpx nq[32][X::M1];
  for (int l = 0; l < 32; ++l)
    for (int r = 0; r < 32; ++r) {
      int g, i; in_coord<X>(l, r, g, i);
      v[l][r] = mk((float)Xr[g][i].real(), (float)Xr[g][i].imag());
      if (i == 0) { v[l][r].y = 123.0f; nq[l][r / X::B] = mk((float)Xr[g][C].real(), -77.0f); }  // imag of DC/Nyquist must be ignored
    }
  for (int l = 0; l < 32; ++l) for (int r = 0; r < 32; ++r) snap[l][r] = v[l][r];
  cpx midi[32][X::M1];
  for (int l = 0; l < 32; ++l) {
    const int lp = mirror_lane<X::A>(l);
    inv_split_pairs<X>(v[l], l, make_split_inv<X>(l), [&](int r) { return snap[lp][r]; }, nq[l], midi[l]);
  }
  for (int l = 0; l < 32; ++l) for (int r = 0; r < 32; ++r) snap[l][r] = v[l][r];
  for (int l = 0; l < 32; ++l) {
    const int lp = mirror_lane<X::A>(l);
    inv_split_finish<X>(v[l], l, [&](int r) { return snap[lp][r]; }, midi[l]);
  }
  for (int l = 0; l < 32; ++l) { pass1<X, true>(v[l]); twiddle1<X, true>(v[l], l, tw1); }
  for (int l = 0; l < 32; ++l) xchg_write<X>(v[l], l, ex.data());
  for (int l = 0; l < 32; ++l) xchg_read<X>(v[l], l, ex.data());
  for (int l = 0; l < 32; ++l) pass2<X, true>(v[l]);
  err = 0; nrm = 0;
  for (int l = 0; l < 32; ++l)
    for (int r = 0; r < 32; ++r) {
      int g, o; out_coord<X>(l, r, g, o);
      const double y0 = 2.0 * v[l][r].x / (double)N, y1 = 2.0 * v[l][r].y / (double)N;   // inv_split yields halves
      err += (y0 - x[g][2 * o]) * (y0 - x[g][2 * o]) + (y1 - x[g][2 * o + 1]) * (y1 - x[g][2 * o + 1]);
      nrm += (double)x[g][2 * o] * x[g][2 * o] + (double)x[g][2 * o + 1] * x[g][2 * o + 1];
    }
  const double inv_rel = std::sqrt(err / nrm);
  const int ok = fwd_rel < 5e-7 && inv_rel < 5e-7 && conflicts == 0 && exact0;
  printf("N=%4d G=%d A=%2d B=%2d  fwd rel-L2 %.3e  inv rel-L2 %.3e  xchg conflicts %d  exact-zero imag %d  %s\n", N, G,
         X::A, X::B, fwd_rel, inv_rel, conflicts, exact0, ok ? "OK" : "FAIL");
  return ok ? 0 : 1;
}

// N = 4096: one emulated warp per frame, 64 registers per lane (warpfft4096.cuh).
static int run4096() {
  const int N = 4096, C = 2048;
  FftTables T = make_fft_tables(N);
  const cpx* tw = reinterpret_cast<const cpx*>(T.tw1.data());
  const cpx* twi = reinterpret_cast<const cpx*>(T.pre.data());
  std::vector<float> x(N);
  for (auto& s : x) s = (float)frand();
  std::vector<cd> Xr(C + 1);
  for (int k = 0; k <= C; ++k) {
    cd s = 0;
    for (int n = 0; n < N; ++n) s += (double)x[n] * std::polar(1.0, -2.0 * M_PI * (double)n * k / N);
    Xr[k] = s;
  }
  // ---- forward ----
  static cpx v[32][64];
  std::vector<cpx> ex(Geo4096::FX_ROUND_CPX);
  for (int l = 0; l < 32; ++l)
    for (int b = 0; b < 64; ++b) { const int i = l + 32 * b; v[l][b] = mk(x[2 * i], x[2 * i + 1]); }
  for (int l = 0; l < 32; ++l) { fwd4096_pass1(v[l]); fwd4096_twiddle(v[l], l, tw); }
  int conflicts = 0;
  for (int h = 0; h < 2; ++h) for (int a = 0; a < 32; ++a) for (int hw = 0; hw < 2; ++hw) {
    std::set<int> banks;
    for (int l = 16 * hw; l < 16 * hw + 16; ++l) { const int row = h == 0 ? l : (l ? 32 - l : 0); banks.insert((row * Geo4096::FXS + a) % 16); }
    if (banks.size() != 16) ++conflicts;
  }
  for (int l = 0; l < 32; ++l) fwd4096_xwrite<0>(v[l], l, ex.data());
  for (int l = 0; l < 32; ++l) fwd4096_xread<0>(v[l], l, ex.data());
  for (int l = 0; l < 32; ++l) fwd4096_xwrite<1>(v[l], l, ex.data());
  for (int l = 0; l < 32; ++l) fwd4096_xread<1>(v[l], l, ex.data());
  std::vector<cd> Xg(C + 1);
  std::vector<int> seen(C + 1, 0);
  int sink_bad = 0;
  cpx snap0[64];
  for (int l = 0; l < 32; ++l) {
    fwd4096_pass2(v[l]);
    if (l == 0) for (int r = 0; r < 64; ++r) snap0[r] = v[0][r];
    cpx nyq = mk(0, 0);
    fwd4096_split(v[l], l, make_split_4096<false>(l), nyq);
    for (int r = 0; r < 64; ++r) { const int k = fwd4096_bin(l, r); Xg[k] = cd(v[l][r].x, v[l][r].y); ++seen[k]; }
    if (l == 0) { Xg[C] = cd(nyq.x, nyq.y); ++seen[C]; }
  }
  // lane 0's rows shared out over the warp (Coop4096): same bins, single-factor twiddle
  double coop_err = 0;
  for (int l = 1; l < 32; ++l) {
    const Coop4096 co = make_coop_4096<false>(l);
    cpx x1, x2;
    split_pair32<0, false>(snap0[co.zi], snap0[co.mi], co.q, co.qr, x1, x2);
    coop_err = std::max(coop_err, std::abs(cd(x1.x, x1.y) - Xg[co.k]) / std::abs(Xg[co.k]));
    coop_err = std::max(coop_err, std::abs(cd(x2.x, x2.y) - Xg[C - co.k]) / std::abs(Xg[C - co.k]));
  }
  if (coop_err > 2e-6) sink_bad = 1;
  int cover = 1;
  for (int k = 0; k <= C; ++k) if (seen[k] != 1) cover = 0;
  double err = 0, nrm = 0;
  for (int k = 0; k <= C; ++k) { err += std::norm(Xg[k] - Xr[k]); nrm += std::norm(Xr[k]); }
  const double fwd_rel = std::sqrt(err / nrm);
  const int exact0 = Xg[0].imag() == 0.0 && Xg[C].imag() == 0.0;
  // ---- inverse ----
  std::vector<cpx> Xf(C + 1);
  for (int k = 0; k <= C; ++k) Xf[k] = mk((float)Xr[k].real(), (float)Xr[k].imag());
  Xf[0].y = 55.0f; Xf[C].y = -9.0f;    // must be ignored
  for (int l = 0; l < 32; ++l) {
    for (int r = 0; r < 64; ++r) v[l][r] = Xf[inv4096_bin(l, r)];
    inv4096_split(v[l], l, make_split_4096<true>(l), Xf[C]);
    inv4096_pass1(v[l]);
    inv4096_twiddle(v[l], l, twi);
  }
  for (int ob = 0; ob < 32; ++ob) for (int j = 0; j < 2; ++j) for (int hw = 0; hw < 2; ++hw) {
    std::set<int> banks;
    for (int l = 16 * hw; l < 16 * hw + 16; ++l) banks.insert((ob * Geo4096::FXS + (j ? Geo4096::row1(l) - 32 : l)) % 16);
    if (banks.size() != 16) ++conflicts;
  }
  for (int a = 0; a < 32; ++a) for (int hw = 0; hw < 2; ++hw) {
    std::set<int> banks;
    for (int l = 16 * hw; l < 16 * hw + 16; ++l) banks.insert((l * Geo4096::FXS + a) % 16);
    if (banks.size() != 16) ++conflicts;
  }
  for (int l = 0; l < 32; ++l) inv4096_xwrite<0>(v[l], l, ex.data());
  for (int l = 0; l < 32; ++l) inv4096_xread<0>(v[l], l, ex.data());
  for (int l = 0; l < 32; ++l) inv4096_xwrite<1>(v[l], l, ex.data());
  for (int l = 0; l < 32; ++l) inv4096_xread<1>(v[l], l, ex.data());
  std::vector<double> y(N);
  for (int l = 0; l < 32; ++l) {
    inv4096_pass2(v[l]);
    for (int oa = 0; oa < 64; ++oa) {
      const int n = l + 32 * oa;
      y[2 * n] = 2.0 * v[l][oa].x / N;
      y[2 * n + 1] = 2.0 * v[l][oa].y / N;
    }
  }
  err = 0; nrm = 0;
  for (int i = 0; i < N; ++i) { err += (y[i] - x[i]) * (y[i] - x[i]); nrm += (double)x[i] * x[i]; }
  const double inv_rel = std::sqrt(err / nrm);
  const int ok = fwd_rel < 5e-7 && inv_rel < 5e-7 && exact0 && cover && conflicts == 0 && !sink_bad;
  printf("N=4096 (1 warp/frame, 64 regs/lane)  fwd rel-L2 %.3e  inv rel-L2 %.3e  xchg conflicts %d  bins covered once %d  exact-zero imag %d  %s\n",
         fwd_rel, inv_rel, conflicts, cover, exact0, ok ? "OK" : "FAIL");
  return ok ? 0 : 1;
}

// N = 4096 inverse as two 1024-point sub-transforms (warpfft4096v2.cuh; k2_4096v2.cu and k23_4096.cu)
static int run4096v2_inverse() {
  typedef GeoX<2048> XV;
  const int N = 4096, C = 2048;
  FftTables T = make_fft_tables(N);
  const cpx* sa = reinterpret_cast<const cpx*>(T.sa.data());
  const cpx *tw = sa, *tc = sa + 1024, *qs = sa + 2048;
  std::vector<float> x(N);
  for (auto& s : x) s = (float)frand();
  std::vector<cpx> Xf(C + 1);
  for (int k = 0; k <= C; ++k) {
    cd s = 0;
    for (int n = 0; n < N; ++n) s += (double)x[n] * std::polar(1.0, -2.0 * M_PI * (double)n * k / N);
    Xf[k] = mk((float)s.real(), (float)s.imag());
  }
  Xf[0].y = 55.0f; Xf[C].y = -9.0f;    // must be ignored
  // every register is parked exactly once
  int cover = 1;
  { std::vector<int> seen(32, 0);
    for (int c = 0; c < 4; ++c) for (int s = 0; s < 8; ++s) ++seen[v2_park_reg(c, s)];
    for (int r = 0; r < 32; ++r) if (seen[r] != 1) cover = 0; }
  static cpx lo[32][32], hi[32][32], ve[32][32], vo[32][32];
  for (int l = 0; l < 32; ++l) for (int b = 0; b < 32; ++b) { lo[l][b] = Xf[l + 32 * b]; hi[l][b] = Xf[1024 + l + 32 * b]; }
  cpx exE[32], exO[32];
  for (int l = 0; l < 32; ++l) {       // lane l's share of lane 0's quads: four direct loads
    const int kc = 32 * l;
    if (l) inv4096v2_quad(Xf[kc], Xf[1024 - kc], Xf[1024 + kc], Xf[2048 - kc], qs[kc], tc[kc], exE[l], exO[l]);
    else inv4096v2_quad0(Xf[0], Xf[1024], Xf[2048], exE[0], exO[0]);
  }
  for (int l = 0; l < 32; ++l) {
    const int lp = (32 - l) & 31;
    for (int j = 0; j < 8; ++j) for (int s = 0; s < 4; ++s) {
      const int b = v2_chunk_reg(j, s), k = l + 32 * b;
      cpx e, o;
      inv4096v2_quad(lo[l][b], lo[lp][31 - b], hi[l][b], hi[lp][31 - b], qs[k], tc[k], e, o);
      if (l == 0) { e = exE[b]; o = exO[b]; }
      ve[l][b] = e; vo[l][b] = o;
    }
  }
  std::vector<cpx> ex(XV::XCPX);
  std::vector<double> y(N);
  for (int h = 0; h < 2; ++h) {
    cpx (*v)[32] = h ? vo : ve;
    for (int l = 0; l < 32; ++l) { pass1<XV, true>(v[l]); twiddle1<XV, true>(v[l], l, tw); }
    for (int l = 0; l < 32; ++l) xchg_write<XV>(v[l], l, ex.data());
    for (int l = 0; l < 32; ++l) xchg_read<XV>(v[l], l, ex.data());
    for (int l = 0; l < 32; ++l) {
      pass2<XV, true>(v[l]);
      for (int oa = 0; oa < 32; ++oa) {
        const int m = l + 32 * oa;
        y[4 * m + 2 * h] = 2.0 * v[l][oa].x / N;
        y[4 * m + 2 * h + 1] = 2.0 * v[l][oa].y / N;
      }
    }
  }
  double err = 0, nrm = 0;
  for (int i = 0; i < N; ++i) { err += (y[i] - x[i]) * (y[i] - x[i]); nrm += (double)x[i] * x[i]; }
  const double inv_rel = std::sqrt(err / nrm);
  const int ok = inv_rel < 5e-7 && cover;
  printf("N=4096 inverse v2 (two 1024-point sub-transforms)  inv rel-L2 %.3e  registers parked once %d  %s\n", inv_rel, cover, ok ? "OK" : "FAIL");
  return ok ? 0 : 1;
}

int main() {
  srand(1234);
  int bad = 0;
  bad += run<256>();
  bad += run<512>();
  bad += run<1024>();
  bad += run<2048>();
  bad += run4096();
  bad += run4096v2_inverse();
  return bad;
}
