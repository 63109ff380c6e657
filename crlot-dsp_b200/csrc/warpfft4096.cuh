// warpfft4096.cuh -- N = 4096: one warp per frame, 2048 complex points = 64 registers per lane.
//
// Same four-step scheme as warpfft.cuh (C = A*B, input index i = a + A*b, output o = ob + B*oa)
// with two differences that remove every cross-lane step of the real split:
//
//   forward  (A = 32, B = 64): pass 1 = ONE 64-point transform per lane (a = lane) over b,
//            pass 2 = TWO 32-point transforms per lane over a, for the output rows
//                 ob0 = lane                       (registers v[0..31],  bin lane     + 64*oa)
//                 ob1 = lane ? 64 - lane : 32      (registers v[32..63], bin ob1      + 64*oa)
//            The mirror of bin k = lane + 64*oa is 2048 - k = (64 - lane) + 64*(31 - oa): the SAME
//            lane's register 32 + (31 - oa).  A mirror pair is therefore v[p] / v[63 - p] and the
//            split needs neither shuffles nor a table: w_k = exp(-i pi lane/2048) * exp(-i pi p/32)
//            = (lane constant) x (compile-time constant).  Lane 0 holds the two self-mirrored rows
//            (ob = 0 and ob = 32) and takes its own branch.
//   inverse  (A = 64, B = 32): the transposed assignment -- pass 1 = TWO 32-point transforms per
//            lane over b for a0 = lane and a1 = lane ? 64 - lane : 32 (spectrum bin a + 64*b), so
//            the inverse split is in-lane as well; pass 2 = ONE 64-point transform per lane
//            (ob = lane) whose outputs n = lane + 32*oa are contiguous across the warp.
//
// Both directions keep the lane as the fastest index of their global/ring accesses (rows
// 32..63 are visited in a lane-permuted order, which still covers one contiguous 256-byte
// segment per warp instruction).
//
// Reference arithmetic restated: third_party/kissfft/kiss_fftr.c:63-155 (real split, DC/Nyquist
// handling).  Everything is a function of (lane, registers): tests/cpu_emu runs it on the CPU.
#pragma once
#include "warpfft.cuh"

namespace crlot {

struct Geo4096 {
  static constexpr int C = 2048;
  // exchange buffers in cpx units; rows padded to an odd length (conflict-free 8-byte accesses)
  static constexpr int FXS = 33, FX_ROUND_CPX = 32 * FXS;   // forward: two rounds of 32 rows x 32 (+1)
                                                            // inverse: two rounds of 32 rows x 32 (+1) as well
  CRLOT_HD static int row1(int lane) { return lane ? 64 - lane : 32; }   // second row/column of a lane
};

// exp(-/+ i pi n/32) for n in [0, 32): compile-time factor of the split (second quadrant included)
template <int n, bool CONJ> CRLOT_HD cpx wq32() {
  static_assert(n >= 0 && n < 32, "angle pi*n/32");
  return mk(cos64(n), CONJ ? sin64(n) : -sin64(n));
}

// lane factors: forward q = (-i/2) exp(-i pi k0/2048), inverse q = (+i/2) exp(+i pi k0/2048)
struct SplitLane4096 {
  cpx q, qr;       // rows with k0 = lane (lanes >= 1) / k0 = 0 (lane 0)
  cpx q1, qr1;     // lane 0 only: k0 = 32
};
template <bool INV> CRLOT_HD SplitLane4096 make_split_4096(int lane) {
  SplitLane4096 L;
  const double pi = 3.14159265358979323846;
  const double th = pi * (double)lane / 2048.0, th1 = pi * 32.0 / 2048.0;
  L.q = mk((float)(-0.5 * sin(th)), (float)((INV ? 0.5 : -0.5) * cos(th)));
  L.qr = mk(-L.q.y, L.q.x);
  L.q1 = mk((float)(-0.5 * sin(th1)), (float)((INV ? 0.5 : -0.5) * cos(th1)));
  L.qr1 = mk(-L.q1.y, L.q1.x);
  return L;
}

// one pair with the compile-time factor exp(-/+ i pi n/32):  x1 = s/2 + f,  x2 = conj(s/2 - f)
template <int n, bool CONJ> CRLOT_HD void split_pair32(cpx z, cpx m, cpx q, cpx qr, cpx& x1, cpx& x2) {
  const cpx s = fma2(m, mk(1.f, -1.f), z);               // z + conj(m)
  const cpx d = fma2(m, mk(-1.f, 1.f), z);               // z - conj(m)
  cpx e = d;
  if (n != 0) e = cmul(d, wq32<n, CONJ>());
  const cpx f = fma2(splat(e.y), qr, mul2(splat(e.x), q));
  x1 = fma2(s, mk(0.5f, 0.5f), f);
  x2 = fma2(s, mk(0.5f, -0.5f), mul2(f, mk(-1.f, 1.f)));
}

// lanes >= 1: pairs v[p] <-> v[63 - p], p = 0..31
template <bool CONJ, int p = 0> CRLOT_HD void split4096_general(cpx (&v)[64], cpx q, cpx qr) {
  cpx x1, x2;
  split_pair32<p, CONJ>(v[p], v[63 - p], q, qr, x1, x2);
  v[p] = x1;
  v[63 - p] = x2;
  if constexpr (p + 1 < 32) split4096_general<CONJ, p + 1>(v, q, qr);
}
// lane 0, row 0 (bins 64*p): pairs v[p] <-> v[32 - p], p = 1..15; p = 16 is its own mirror
template <bool CONJ, int p = 1> CRLOT_HD void split4096_row0(cpx (&v)[64], cpx q, cpx qr) {
  cpx x1, x2;
  split_pair32<p, CONJ>(v[p], v[32 - p], q, qr, x1, x2);
  v[p] = x1;
  v[32 - p] = x2;
  if constexpr (p + 1 < 16) split4096_row0<CONJ, p + 1>(v, q, qr);
}
// lane 0, row 32 (bins 32 + 64*p): pairs v[32 + p] <-> v[63 - p], p = 0..15
template <bool CONJ, int p = 0> CRLOT_HD void split4096_row32(cpx (&v)[64], cpx q, cpx qr) {
  cpx x1, x2;
  split_pair32<p, CONJ>(v[32 + p], v[63 - p], q, qr, x1, x2);
  v[32 + p] = x1;
  v[63 - p] = x2;
  if constexpr (p + 1 < 16) split4096_row32<CONJ, p + 1>(v, q, qr);
}

// ---- forward ------------------------------------------------------------------------------------
// pass-1 view: v[b] = point lane + 32*b.
CRLOT_HD void fwd4096_pass1(cpx (&v)[64]) { Dft<64, false>::run(v); }
// tw[ob*32 + a] = exp(-2 pi i a*ob/2048)
CRLOT_HD void fwd4096_twiddle(cpx (&v)[64], int lane, const cpx* tw) {
#pragma unroll
  for (int ob = 1; ob < 64; ++ob) v[ob] = cmul(v[ob], tw[ob * 32 + lane]);
}
// exchange round h moves rows ob in [32h, 32h+32): every lane writes its column a = lane, then
// reads the whole row it transforms in pass 2.  The caller puts __syncwarp() between the two and
// after the read (the buffer is reused by the next round).
template <int h> CRLOT_HD void fwd4096_xwrite(const cpx (&v)[64], int lane, cpx* ex) {
#pragma unroll
  for (int r = 0; r < 32; ++r) ex[r * Geo4096::FXS + lane] = v[32 * h + r];
}
template <int h> CRLOT_HD void fwd4096_xread(cpx (&v)[64], int lane, const cpx* ex) {
  const int row = h == 0 ? lane : (lane ? 32 - lane : 0);
#pragma unroll
  for (int a = 0; a < 32; ++a) v[32 * h + a] = ex[row * Geo4096::FXS + a];
}
CRLOT_HD void fwd4096_pass2(cpx (&v)[64]) {
  dft_sub<32, false, 0, 1, 64>(v);
  dft_sub<32, false, 32, 1, 64>(v);
}
// After pass 2: v[oa] = Z[ob0 + 64*oa], v[32 + oa] = Z[ob1 + 64*oa].  Produces X in the same slots;
// lane 0 also returns the Nyquist bin X[2048].
CRLOT_HD void fwd4096_split(cpx (&v)[64], int lane, const SplitLane4096& L, cpx& nyq) {
  if (lane != 0) {
    split4096_general<false>(v, L.q, L.qr);
  } else {
    const cpx z0 = v[0];
    v[16] = mul2(v[16], mk(1.f, -1.f));                    // X[1024] = conj(Z[1024])
    split4096_row0<false>(v, L.q, L.qr);
    split4096_row32<false>(v, L.q1, L.qr1);
    v[0] = mk(z0.x + z0.y, 0.0f);                          // exact DC and Nyquist (kiss_fftr.c:93-99)
    nyq = mk(z0.x - z0.y, 0.0f);
  }
}
// ---- lane 0's two self-mirrored rows, shared out over the warp ------------------------------------
// Lane 0 holds rows/columns 0 and 32 (bins 64p and 32 + 64p), whose mirrors are in its own
// registers: 31 pairs + DC/Nyquist + bin 1024 that the general pattern above does not cover.  Run
// in a divergent branch they would cost the whole warp 31 more pair computations per frame, so they
// go through 64 words of shared memory instead: lane l >= 1 owns ONE of those pairs for the whole
// kernel (its full twiddle is a lane constant), lane 0 keeps DC, Nyquist and bin 1024.
//   lanes 1..15 : bins k = 64 l and 2048 - k        (slots  l       and 32 - l  of lane 0)
//   lanes 16..31: bins k = 32 + 64 (l-16), 2048 - k (slots 16 + l   and 79 - l  of lane 0)
struct Coop4096 {
  int zi, mi, k;   // slot of z, slot of the mirror, bin of z
  cpx q, qr;       // (-/+ i/2) exp(-/+ i pi k/2048) and i times it
};
template <bool INV> CRLOT_HD Coop4096 make_coop_4096(int lane) {
  Coop4096 c;
  c.zi = lane < 16 ? lane : 16 + lane;
  c.mi = lane < 16 ? 32 - lane : 79 - lane;
  c.k = lane < 16 ? 64 * lane : 32 + 64 * (lane - 16);
  const double th = 3.14159265358979323846 * (double)c.k / 2048.0;
  c.q = mk((float)(-0.5 * sin(th)), (float)((INV ? 0.5 : -0.5) * cos(th)));
  c.qr = mk(-c.q.y, c.q.x);
  return c;
}
// bin held by register r of `lane` after the split
CRLOT_HD int fwd4096_bin(int lane, int r) { return (r < 32 ? lane : Geo4096::row1(lane)) + 64 * (r & 31); }

// ---- inverse ------------------------------------------------------------------------------------
// pass-1 view: v[b] = X[lane + 64*b], v[32 + b] = X[a1 + 64*b], a1 = row1(lane).
CRLOT_HD int inv4096_bin(int lane, int r) { return (r < 32 ? lane : Geo4096::row1(lane)) + 64 * (r & 31); }
// Halved values Z'/2 are produced (the caller folds the 2 into its 1/N scale); only the real parts
// of X[0] and X[2048] are used (kiss_fftr.c:131-132).
CRLOT_HD void inv4096_split(cpx (&v)[64], int lane, const SplitLane4096& L, cpx nyq) {
  if (lane != 0) {
    split4096_general<true>(v, L.q, L.qr);
  } else {
    const cpx x0 = v[0];
    v[16] = mul2(v[16], mk(1.f, -1.f));
    split4096_row0<true>(v, L.q, L.qr);
    split4096_row32<true>(v, L.q1, L.qr1);
    v[0] = mk(0.5f * (x0.x + nyq.x), 0.5f * (x0.x - nyq.x));
  }
}
CRLOT_HD void inv4096_pass1(cpx (&v)[64]) {
  dft_sub<32, true, 0, 1, 64>(v);
  dft_sub<32, true, 32, 1, 64>(v);
}
// twi[ob*64 + a] = exp(-2 pi i a*ob/2048) (forward values; conjugated here)
CRLOT_HD void inv4096_twiddle(cpx (&v)[64], int lane, const cpx* twi) {
  const int a1 = Geo4096::row1(lane);
#pragma unroll
  for (int ob = 1; ob < 32; ++ob) {
    v[ob] = cmulc(v[ob], twi[ob * 64 + lane]);
    v[32 + ob] = cmulc(v[32 + ob], twi[ob * 64 + a1]);
  }
}
// exchange in two rounds of 32 columns (8.4 KB per warp): round h moves the lanes' column a_h.
// The caller puts __syncwarp() between write and read and after the read.
template <int h> CRLOT_HD void inv4096_xwrite(const cpx (&v)[64], int lane, cpx* ex) {
  const int col = h == 0 ? lane : Geo4096::row1(lane) - 32;
#pragma unroll
  for (int ob = 0; ob < 32; ++ob) ex[ob * Geo4096::FXS + col] = v[32 * h + ob];
}
template <int h> CRLOT_HD void inv4096_xread(cpx (&v)[64], int lane, const cpx* ex) {
#pragma unroll
  for (int a = 0; a < 32; ++a) v[32 * h + a] = ex[lane * Geo4096::FXS + a];
}
// pass 2: v[oa] = z[lane + 32*oa]
CRLOT_HD void inv4096_pass2(cpx (&v)[64]) { Dft<64, true>::run(v); }

}  // namespace crlot
