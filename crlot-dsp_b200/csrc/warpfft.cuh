// warpfft.cuh -- the per-lane pieces of the warp-autonomous real FFT used by kernels 1 and 2.
//
// One warp owns 1024 complex points per step, 32 per lane ("register file" v[32]):
//   N=2048 -> 1 frame  (C=1024 = 32 x 32)      N=1024 -> 2 frames (C=512 = 32 x 16)
//   N= 512 -> 4 frames (C= 256 = 16 x 16)      N= 256 -> 8 frames (C=128 = 16 x  8)
//   N=4096 -> half a frame per warp (radix-2 decimation-in-frequency pre-stage, then C=1024)
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
// Forward (after pass 2), k = o, M = Z[C-k]:
//   X[k] = 0.5*(Z + conj M) + 0.5*t_k*(Z - conj M),  t_k = exp(-i pi (k/C + 1/2))
//        = conj(M) + A_k*(Z - conj M),               A_k = 0.5*(1 + t_k)
//   (kiss_fftr.c:53-59,101-116 of the reference, restated; 5 packed instructions per bin)
//   k = 0 gives X[0] = Re+Im, X[C] = Re-Im with exactly zero imaginary parts.
// sa[] is the table sa[k] = A_k, k in [0,C).
// `peer(r)` returns register r of the partner lane (a warp shuffle on the device).
template <class X, class Peer>
CRLOT_HD void fwd_split(cpx (&v)[32], int lane, const cpx* sa, Peer peer, cpx (&nyq)[X::M2]) {
  cpx out[32];
#pragma unroll
  for (int j = 0; j < X::M2; ++j) {
    const int ob = (lane + 32 * j) % X::B;
#pragma unroll
    for (int oa = 0; oa < X::A; ++oa) {
      const cpx z = v[j * X::A + oa];
      const cpx far = peer(j * X::A + (X::A - 1 - oa));
      const cpx own = v[j * X::A + (X::A - oa) % X::A];
      const cpx m = (ob == 0) ? own : far;                   // Z[C-k]
      const cpx ds = fma2(m, mk(-1.f, 1.f), z);              // Z - conj(M)
      const cpx mc = mul2(m, mk(1.f, -1.f));                 // conj(M)
      out[j * X::A + oa] = cfma(ds, sa[ob + X::B * oa], mc);
    }
  }
#pragma unroll
  for (int j = 0; j < X::M2; ++j) {
    const int ob = (lane + 32 * j) % X::B;
    if (ob == 0) {   // k = 0 of frame g: exact DC and Nyquist (ob = lane % B for every j)
      const cpx z = v[j * X::A];
      out[j * X::A] = mk(z.x + z.y, 0.0f);
      nyq[j] = mk(z.x - z.y, 0.0f);
    }
  }
#pragma unroll
  for (int r = 0; r < 32; ++r) v[r] = out[r];
}

// Inverse (before pass 1), k = i, M = X[C-k]:
//   Z'[k] = (X + conj M) + u_k*(X - conj M),  u_k = exp(+i pi (k/C + 1/2)) = conj(t_k)
//   Z'[k]/2 = conj(M) + conj(A_k)*(X - conj M)
//   (kiss_fftr.c:119-155 restated).  The HALVED values are produced (exact power-of-two
//   scaling) and the caller folds the 2 into its final 1/N scale.  k = 0 uses only the real
//   parts of X[0] and X[C]:  Z'[0]/2 = ((X0.r + XC.r)/2, (X0.r - XC.r)/2).
template <class X, class Peer>
CRLOT_HD void inv_split(cpx (&v)[32], int lane, const cpx* sa, Peer peer, const cpx (&nyq)[X::M1]) {
  cpx out[32];
#pragma unroll
  for (int j = 0; j < X::M1; ++j) {
    const int a = (lane + 32 * j) % X::A;
#pragma unroll
    for (int b = 0; b < X::B; ++b) {
      const cpx x = v[j * X::B + b];
      const cpx far = peer(j * X::B + (X::B - 1 - b));
      const cpx own = v[j * X::B + (X::B - b) % X::B];
      const cpx m = (a == 0) ? own : far;                    // X[C-k]
      const cpx ds = fma2(m, mk(-1.f, 1.f), x);
      const cpx mc = mul2(m, mk(1.f, -1.f));
      out[j * X::B + b] = cfmac(ds, sa[a + X::A * b], mc);
    }
    if (a == 0) out[j * X::B] = mk(0.5f * (v[j * X::B].x + nyq[j].x), 0.5f * (v[j * X::B].x - nyq[j].x));
  }
#pragma unroll
  for (int r = 0; r < 32; ++r) v[r] = out[r];
}

// ---- N = 4096: two warps per frame ------------------------------------------------------------
// The 2048-point complex transform is split by one radix-2 decimation-in-frequency stage done at
// load time:  a[n] = z[n] + z[n+1024],  b[n] = (z[n] - z[n+1024]) * W_2048^n  (n < 1024);
// warp parity 0 transforms a -> bins 2k', parity 1 transforms b -> bins 2k'+1, each with the
// 1024-point machinery above (Geo<2048>).  Mirror pairs keep their parity, so the real split
// stays inside one warp: even bins pair as in fwd_split, odd bins pair (ob,oa) <-> (31-ob,31-oa)
// with no special lane.  sa[] has 2048 entries indexed by the true bin.
template <class Peer>
CRLOT_HD void fwd_split_4096(cpx (&v)[32], int lane, int parity, const cpx* sa, Peer peer, cpx& nyq) {
  cpx out[32];
  const int ob = lane;
  if (parity == 0) {
#pragma unroll
    for (int oa = 0; oa < 32; ++oa) {
      const cpx z = v[oa];
      const cpx far = peer(31 - oa);
      const cpx own = v[(32 - oa) % 32];
      const cpx m = (ob == 0) ? own : far;
      const cpx ds = fma2(m, mk(-1.f, 1.f), z);
      const cpx mc = mul2(m, mk(1.f, -1.f));
      out[oa] = cfma(ds, sa[2 * (ob + 32 * oa)], mc);
    }
    if (ob == 0) {
      out[0] = mk(v[0].x + v[0].y, 0.0f);
      nyq = mk(v[0].x - v[0].y, 0.0f);
    }
  } else {
#pragma unroll
    for (int oa = 0; oa < 32; ++oa) {
      const cpx z = v[oa];
      const cpx m = peer(31 - oa);
      const cpx ds = fma2(m, mk(-1.f, 1.f), z);
      const cpx mc = mul2(m, mk(1.f, -1.f));
      out[oa] = cfma(ds, sa[2 * (ob + 32 * oa) + 1], mc);
    }
  }
#pragma unroll
  for (int r = 0; r < 32; ++r) v[r] = out[r];
}

// Inverse pre-pass for N = 4096, one register: from X[k], X[k+1024] and their mirrors
// X[2048-k], X[1024-k] (all read straight from global memory, k = a + 32 b < 1024) build the
// halved Z'[k], Z'[k+1024] and combine them for this warp's parity:
//   parity 0: Z'[k] + Z'[k+1024]            parity 1: (Z'[k] - Z'[k+1024]) * conj(W_2048^k)
CRLOT_HD cpx inv_pre_4096(int k, int parity, cpx xlo, cpx mlo, cpx xhi, cpx mhi, cpx alo, cpx ahi, cpx pre) {
  cpx zlo = cfmac(fma2(mlo, mk(-1.f, 1.f), xlo), alo, mul2(mlo, mk(1.f, -1.f)));
  if (k == 0) zlo = mk(0.5f * (xlo.x + mlo.x), 0.5f * (xlo.x - mlo.x));   // mlo = X[2048]: real parts only
  const cpx zhi = cfmac(fma2(mhi, mk(-1.f, 1.f), xhi), ahi, mul2(mhi, mk(1.f, -1.f)));
  return parity == 0 ? zlo + zhi : cmulc(zlo - zhi, pre);
}

// partner lane for the mirror: pass-2 view uses B, pass-1 view uses A (lane + 32*j keeps j).
template <int D> CRLOT_HD int mirror_lane(int lane) {
  const int base = lane - lane % D;
  return base + (D - lane % D) % D;
}

}  // namespace crlot
