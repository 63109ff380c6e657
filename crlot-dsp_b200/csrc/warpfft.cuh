// warpfft.cuh -- the per-lane pieces of the warp-autonomous real FFT used by kernels 1 and 2.
//
// One warp owns 1024 complex points per step, 32 per lane ("register file" v[32]):
//   N=2048 -> 1 frame  (C=1024 = 32 x 32)      N=1024 -> 2 frames (C=512 = 32 x 16)
//   N= 512 -> 4 frames (C= 256 = 16 x 16)      N= 256 -> 8 frames (C=128 = 16 x  8)
//   N=4096 -> one frame per warp with 64 registers per lane: warpfft4096.cuh
// A C-point complex transform is done "four-step": with input index i = a + A*b and output
// index o = ob + B*oa,
//   pass 1: lane item (g,a) transforms over b (B points, in registers), times W_C^(a*ob)
//   exchange: one transposition through shared memory (the only one)
//   pass 2: lane item (g,ob) transforms over a (A points, in registers)
// so both the input and the output index have the lane as their fastest digit: global
// loads and stores are contiguous across the warp without any further staging.
//
// The real<->complex split needs the mirror bin C-k, which lives in lane (B-ob)%B at register
// A-1-oa: one warp shuffle per value, no shared memory.
//
// Everything here is a function of (lane, registers) so that tests/cpu_emu can run the same
// code lane by lane on the CPU (index algebra check without a GPU).
#pragma once
#include <math.h>

#include "regfft.cuh"

namespace crlot {

template <int NFFT> struct Geo;   // NFFT = real frame length handled with whole frames per warp
template <> struct Geo<2048> { static constexpr int C = 1024, G = 1, A = 32, B = 32; };
template <> struct Geo<1024> { static constexpr int C = 512,  G = 2, A = 32, B = 16; };
template <> struct Geo<512>  { static constexpr int C = 256,  G = 4, A = 16, B = 16; };
template <> struct Geo<256>  { static constexpr int C = 128,  G = 8, A = 16, B = 8;  };

template <int NFFT> struct GeoX : Geo<NFFT> {
  using Geo<NFFT>::C; using Geo<NFFT>::G; using Geo<NFFT>::A; using Geo<NFFT>::B;
  static constexpr int M1 = G * A / 32;       // pass-1 transforms per lane
  static constexpr int M2 = G * B / 32;       // pass-2 transforms per lane
  static_assert(M1 * B == 32 && M2 * A == 32, "32 complex registers per lane");
  // Exchange buffer addressing (units of cpx = 8 bytes).  Rows are padded to an odd length so
  // that both the pass-1 writes (lane = fastest digit of (g,a)) and the pass-2 reads (lane =
  // fastest digit of (g,ob)) touch 16 distinct 8-byte bank pairs per half-warp.
  static constexpr bool kSplitG = (B < 16);   // N=256: two frames share a padded row group
  static constexpr int XS = kSplitG ? (G / 2) * A + 1 : G * A + 1;
  static constexpr int XROWS = kSplitG ? 2 * B : B;
  static constexpr int XCPX = XROWS * XS;     // exchange buffer size per warp in cpx
  CRLOT_HD static int xaddr(int g, int a, int ob) {
    return kSplitG ? (ob + B * (g & 1)) * XS + (g >> 1) * A + a : ob * XS + g * A + a;
  }
};

// ---- pass 1: B-point transforms over b; registers v[j*B + b] -> v[j*B + ob] ------------------
template <class X, bool INV, int J = 0> CRLOT_HD void pass1(cpx (&v)[32]) {
  dft_sub<X::B, INV, J * X::B, 1, 32>(v);
  if constexpr (J + 1 < X::M1) pass1<X, INV, J + 1>(v);
}
// ---- inter-pass twiddle W_C^(a*ob); tw1 is the forward table tw1[ob*A + a] --------------------
template <class X, bool INV> CRLOT_HD void twiddle1(cpx (&v)[32], int lane, const cpx* tw1) {
  const int a = lane % X::A;   // item i1 = lane + 32*j, a = i1 % A, and 32 % A == 0
#pragma unroll
  for (int ob = 1; ob < X::B; ++ob) {
    const cpx w = tw1[ob * X::A + a];
#pragma unroll
    for (int j = 0; j < X::M1; ++j) v[j * X::B + ob] = INV ? cmulc(v[j * X::B + ob], w) : cmul(v[j * X::B + ob], w);
  }
}
// ---- exchange ---------------------------------------------------------------------------------
template <class X> CRLOT_HD void xchg_write(const cpx (&v)[32], int lane, cpx* ex) {
#pragma unroll
  for (int j = 0; j < X::M1; ++j) {
    const int i1 = lane + 32 * j, g = i1 / X::A, a = i1 % X::A;
#pragma unroll
    for (int ob = 0; ob < X::B; ++ob) ex[X::xaddr(g, a, ob)] = v[j * X::B + ob];
  }
}
template <class X> CRLOT_HD void xchg_read(cpx (&v)[32], int lane, const cpx* ex) {
#pragma unroll
  for (int j = 0; j < X::M2; ++j) {
    const int i2 = lane + 32 * j, g = i2 / X::B, ob = i2 % X::B;
#pragma unroll
    for (int a = 0; a < X::A; ++a) v[j * X::A + a] = ex[X::xaddr(g, a, ob)];
  }
}
// ---- pass 2: A-point transforms over a; registers v[j*A + a] -> v[j*A + oa] -------------------
template <class X, bool INV, int J = 0> CRLOT_HD void pass2(cpx (&v)[32]) {
  dft_sub<X::A, INV, J * X::A, 1, 32>(v);
  if constexpr (J + 1 < X::M2) pass2<X, INV, J + 1>(v);
}

// ---- index helpers ---------------------------------------------------------------------------
// pass-1 view: register r = j*B + b of `lane` holds point i = a + A*b of frame g
template <class X> CRLOT_HD void in_coord(int lane, int r, int& g, int& i) {
  const int j = r / X::B, b = r % X::B, i1 = lane + 32 * j;
  g = i1 / X::A;
  i = i1 % X::A + X::A * b;
}
// pass-2 view: register r = j*A + oa of `lane` holds point o = ob + B*oa of frame g
template <class X> CRLOT_HD void out_coord(int lane, int r, int& g, int& o) {
  const int j = r / X::A, oa = r % X::A, i2 = lane + 32 * j;
  g = i2 / X::B;
  o = i2 % X::B + X::B * oa;
}

// ---- real <-> half-length complex split -------------------------------------------------------
// Mirror of point o = ob + B*oa (pass-2 view) is C-o: lane item (g,(B-ob)%B), register
// A-1-oa, except for ob == 0 where it is the lane's own register (A-oa)%A.  Same algebra
// with (A,a,b) for the pass-1 view.
//
// Forward (after pass 2).  For the PAIR k, C-k with z = Z[k], m = Z[C-k], w_k = exp(-i pi k/C):
//   s = z + conj m,  d = z - conj m,  f = (-i/2) w_k d
//   X[k] = s/2 + f            X[C-k] = conj(s/2 - f)
//   (kiss_fftr.c:53-59,101-116 of the reference, restated)
// One lane computes BOTH bins of a pair -- those of its registers oa < A/2 -- and hands X[C-k] to
// the lane that stores it, so the pair costs 9 packed instructions instead of 2 x 5, and no table
// is read: with k = ob + B*oa,  w_k = exp(-i pi ob/C) * exp(-i pi oa/A)  = (lane constant) x
// (compile-time constant).  The lane constants (SplitLane) are computed once per kernel in double.
//   k = 0 gives X[0] = Re+Im, X[C] = Re-Im with exactly zero imaginary parts.
// The split runs in two phases because values travel between lanes twice:
//   phase A (split_pairs):  needs peer(r) = register r of the mirror lane BEFORE the phase
//   phase B (split_finish): needs peer(r) = register r of the mirror lane AFTER phase A
// (on the device both are warp shuffles of the live registers; the CPU emulator snapshots).

// exp(-/+ i pi n/D) for 0 <= n <= D/2 (first quadrant), D in {8,16,32}: the compile-time factor
template <int D, int n, bool CONJ> CRLOT_HD cpx wq() {
  static_assert(D == 8 || D == 16 || D == 32, "split constants: 2D in {16,32,64}");
  static_assert(n >= 0 && 2 * n <= D, "first quadrant only");
  constexpr int m = n * (32 / D);                       // angle = 2 pi m/64
  return mk(c64(m), CONJ ? c64(16 - m) : -c64(16 - m));
}

// Per-lane constants of the split, one pair per register group j: (q, rot(q)) with
//   forward: q = (-i/2) exp(-i pi ob/C)      inverse: q = (+i/2) exp(+i pi a/C)
template <int M> struct SplitLane {
  cpx q[M], qr[M];          // qr = (-q.y, q.x) = i*q, so that e*q = e.x*q + e.y*qr
};
template <class X> CRLOT_HD SplitLane<X::M2> make_split_fwd(int lane) {
  SplitLane<X::M2> L;
#pragma unroll
  for (int j = 0; j < X::M2; ++j) {
    const int ob = (lane + 32 * j) % X::B;
    const double th = 3.14159265358979323846 * (double)ob / (double)X::C;
    L.q[j] = mk((float)(-0.5 * sin(th)), (float)(-0.5 * cos(th)));
    L.qr[j] = mk(-L.q[j].y, L.q[j].x);
  }
  return L;
}
template <class X> CRLOT_HD SplitLane<X::M1> make_split_inv(int lane) {
  SplitLane<X::M1> L;
#pragma unroll
  for (int j = 0; j < X::M1; ++j) {
    const int a = (lane + 32 * j) % X::A;
    const double th = 3.14159265358979323846 * (double)a / (double)X::C;
    L.q[j] = mk((float)(-0.5 * sin(th)), (float)(0.5 * cos(th)));
    L.qr[j] = mk(-L.q[j].y, L.q[j].x);
  }
  return L;
}

// One pair: from z, m and the two factors produce x1 = s/2 + f, x2 = conj(s/2 - f).
template <int D, int n, bool CONJ> CRLOT_HD void split_pair(cpx z, cpx m, cpx q, cpx qr, cpx& x1, cpx& x2) {
  const cpx s = fma2(m, mk(1.f, -1.f), z);               // z + conj(m)
  const cpx d = fma2(m, mk(-1.f, 1.f), z);               // z - conj(m)
  cpx e = d;
  if (n != 0) e = cmul(d, wq<D, n, CONJ>());             // compile-time factor
  const cpx f = fma2(splat(e.y), qr, mul2(splat(e.x), q));   // lane factor
  x1 = fma2(s, mk(0.5f, 0.5f), f);
  x2 = fma2(s, mk(0.5f, -0.5f), mul2(f, mk(-1.f, 1.f)));
}

// Phase A over the registers of group j laid out as v[j*R + r], r < R (R = A forward, B inverse),
// with `self` = this lane's index digit is 0 (mirror inside the lane).  x2 of pair r is parked in
// slot R-1-r.  Descending r, so that the self lane still finds its own partner register R-r intact.
template <int R, int D, bool CONJ, int r, class Peer>
CRLOT_HD void split_pairs_rec(cpx (&v)[32], int base, bool self, cpx q, cpx qr, Peer& peer) {
  const cpx far = peer(base + R - 1 - r);
  const cpx own = v[base + (R - r) % R];
  const cpx m = self ? own : far;
  cpx x1, x2;
  split_pair<D, r, CONJ>(v[base + r], m, q, qr, x1, x2);
  v[base + r] = x1;
  v[base + R - 1 - r] = x2;
  if constexpr (r > 0) split_pairs_rec<R, D, CONJ, r - 1>(v, base, self, q, qr, peer);
}
// Phase B: bins R/2..R-1 of the group come from the mirror lane's parked x2 (self lane: its own,
// one slot lower; its middle bin R/2 is the self-mirrored conj(z) saved before phase A).
template <int R, class Peer> CRLOT_HD void split_finish_group(cpx (&v)[32], int base, bool self, cpx mid, Peer& peer) {
#pragma unroll
  for (int r = R - 1; r > R / 2; --r) {
    const cpx far = peer(base + r);
    const cpx own = v[base + r - 1];
    v[base + r] = self ? own : far;
  }
  const cpx far = peer(base + R / 2);
  v[base + R / 2] = self ? mid : far;
}

// Forward split, pass-2 view.  peer_a / peer_b as described above.
template <class X, class PeerA>
CRLOT_HD void fwd_split_pairs(cpx (&v)[32], int lane, const SplitLane<X::M2>& L, PeerA peer_a, cpx (&nyq)[X::M2],
                              cpx (&mid)[X::M2]) {
#pragma unroll
  for (int j = 0; j < X::M2; ++j) {
    const bool self = (lane + 32 * j) % X::B == 0;
    const cpx z0 = v[j * X::A];
    mid[j] = mul2(v[j * X::A + X::A / 2], mk(1.f, -1.f));          // X[C/2] = conj(Z[C/2]) (self lane only)
    split_pairs_rec<X::A, X::A, false, X::A / 2 - 1>(v, j * X::A, self, L.q[j], L.qr[j], peer_a);
    if (self) {                                                    // k = 0: exact DC and Nyquist
      v[j * X::A] = mk(z0.x + z0.y, 0.0f);
      nyq[j] = mk(z0.x - z0.y, 0.0f);
    }
  }
}
template <class X, class PeerB>
CRLOT_HD void fwd_split_finish(cpx (&v)[32], int lane, PeerB peer_b, const cpx (&mid)[X::M2]) {
#pragma unroll
  for (int j = 0; j < X::M2; ++j) split_finish_group<X::A>(v, j * X::A, (lane + 32 * j) % X::B == 0, mid[j], peer_b);
}

// Inverse (before pass 1), pass-1 view, k = a + A*b.  For the pair k, C-k with x = X[k], m = X[C-k]:
//   s = x + conj m,  d = x - conj m,  f = (+i/2) conj(w_k) d
//   Z'[k]/2 = s/2 + f         Z'[C-k]/2 = conj(s/2 - f)
//   (kiss_fftr.c:119-155 restated).  The HALVED values are produced (exact power-of-two
//   scaling) and the caller folds the 2 into its final 1/N scale.  k = 0 uses only the real
//   parts of X[0] and X[C]:  Z'[0]/2 = ((X0.r + XC.r)/2, (X0.r - XC.r)/2).
template <class X, class PeerA>
CRLOT_HD void inv_split_pairs(cpx (&v)[32], int lane, const SplitLane<X::M1>& L, PeerA peer_a, const cpx (&nyq)[X::M1],
                              cpx (&mid)[X::M1]) {
#pragma unroll
  for (int j = 0; j < X::M1; ++j) {
    const bool self = (lane + 32 * j) % X::A == 0;
    const cpx x0 = v[j * X::B];
    mid[j] = mul2(v[j * X::B + X::B / 2], mk(1.f, -1.f));
    split_pairs_rec<X::B, X::B, true, X::B / 2 - 1>(v, j * X::B, self, L.q[j], L.qr[j], peer_a);
    if (self) v[j * X::B] = mk(0.5f * (x0.x + nyq[j].x), 0.5f * (x0.x - nyq[j].x));
  }
}
template <class X, class PeerB>
CRLOT_HD void inv_split_finish(cpx (&v)[32], int lane, PeerB peer_b, const cpx (&mid)[X::M1]) {
#pragma unroll
  for (int j = 0; j < X::M1; ++j) split_finish_group<X::B>(v, j * X::B, (lane + 32 * j) % X::A == 0, mid[j], peer_b);
}

// partner lane for the mirror: pass-2 view uses B, pass-1 view uses A (lane + 32*j keeps j).
template <int D> CRLOT_HD int mirror_lane(int lane) {
  const int base = lane - lane % D;
  return base + (D - lane % D) % D;
}

}  // namespace crlot
