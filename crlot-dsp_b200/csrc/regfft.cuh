// regfft.cuh -- straight-line complex DFTs of 2..32 points held entirely in registers.
//
// Every index below is a compile-time constant after unrolling, so `cpx v[R]` never touches
// local memory: a "permutation" is only a renaming of registers.
//
// Blackwell-native arithmetic: on sm_100a a complex value is one 64-bit register pair and
// every complex add/sub is ONE packed instruction (FADD2), every complex multiply TWO
// (FMUL2 + FFMA2, using the .F32 splat and .LO_HI swap operand modifiers), every
// "+/- i times" rotation is folded into an FFMA2 with a (+1,-1) constant pair.  That halves
// the issue slots of the FFT, which is what bounds kernel 1 (profiles/README.md).
//
// The same header compiles as plain C++ (scalar fallbacks) for the CPU lane-emulator in
// tests/cpu_emu, which checks the index algebra of the warp kernels without a GPU.
//
// Convention: Dft<R,false> computes X[k] = sum_n x[n] exp(-2 pi i n k / R)  (forward),
//             Dft<R,true>  the conjugate kernel (unnormalised inverse).  Natural order in
//             and out.
#pragma once

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#define CRLOT_HD __host__ __device__ __forceinline__
#else
#define CRLOT_HD inline
#endif

namespace crlot {

struct alignas(8) cpx {
  float x, y;
};

CRLOT_HD cpx mk(float a, float b) { cpx r; r.x = a; r.y = b; return r; }

#if defined(__CUDA_ARCH__)
// ---- packed f32x2 backend (sm_100a) ---------------------------------------------------------
__device__ __forceinline__ float2 f2(cpx a) { return make_float2(a.x, a.y); }
__device__ __forceinline__ cpx c2(float2 a) { return mk(a.x, a.y); }
CRLOT_HD cpx operator+(cpx a, cpx b) { return c2(__fadd2_rn(f2(a), f2(b))); }
CRLOT_HD cpx operator-(cpx a, cpx b) { return c2(__fadd2_rn(f2(a), make_float2(-b.x, -b.y))); }
// c + a*(s.x, s.y) element-wise
CRLOT_HD cpx fma2(cpx a, cpx s, cpx c) { return c2(__ffma2_rn(f2(a), f2(s), f2(c))); }
CRLOT_HD cpx mul2(cpx a, cpx s) { return c2(__fmul2_rn(f2(a), f2(s))); }
#else
CRLOT_HD cpx operator+(cpx a, cpx b) { return mk(a.x + b.x, a.y + b.y); }
CRLOT_HD cpx operator-(cpx a, cpx b) { return mk(a.x - b.x, a.y - b.y); }
CRLOT_HD cpx fma2(cpx a, cpx s, cpx c) { return mk(a.x * s.x + c.x, a.y * s.y + c.y); }
CRLOT_HD cpx mul2(cpx a, cpx s) { return mk(a.x * s.x, a.y * s.y); }
#endif

CRLOT_HD cpx swp(cpx a) { return mk(a.y, a.x); }           // .LO_HI operand modifier on the device
CRLOT_HD cpx splat(float s) { return mk(s, s); }           // .F32 operand modifier on the device
CRLOT_HD cpx conj(cpx a) { return mk(a.x, -a.y); }

// a*w = a.x*(w.x, w.y) + a.y*(-w.y, w.x)
CRLOT_HD cpx cmul(cpx a, cpx w) { return fma2(splat(a.y), mk(-w.y, w.x), mul2(splat(a.x), w)); }
// a*conj(w) = a.x*(w.x, -w.y) + a.y*(w.y, w.x)
CRLOT_HD cpx cmulc(cpx a, cpx w) { return fma2(splat(a.y), mk(w.y, w.x), mul2(splat(a.x), mk(w.x, -w.y))); }
// c + a*w  and  c + a*conj(w)
CRLOT_HD cpx cfma(cpx a, cpx w, cpx c) { return fma2(splat(a.y), mk(-w.y, w.x), fma2(splat(a.x), w, c)); }
CRLOT_HD cpx cfmac(cpx a, cpx w, cpx c) { return fma2(splat(a.y), mk(w.y, w.x), fma2(splat(a.x), mk(w.x, -w.y), c)); }
// s + (-i)^(+/-1) * t :  forward uses -i (t.y, -t.x), inverse +i (-t.y, t.x)
template <bool INV> CRLOT_HD cpx add_rot(cpx s, cpx t) { return fma2(swp(t), INV ? mk(-1.f, 1.f) : mk(1.f, -1.f), s); }
template <bool INV> CRLOT_HD cpx sub_rot(cpx s, cpx t) { return fma2(swp(t), INV ? mk(1.f, -1.f) : mk(-1.f, 1.f), s); }
template <bool INV> CRLOT_HD cpx rot90(cpx a) { return mul2(swp(a), INV ? mk(-1.f, 1.f) : mk(1.f, -1.f)); }

// cos(2 pi k/32), k = 0..8; everything else by symmetry.
#define CRLOT_C32_0 1.0f
#define CRLOT_C32_1 0.98078528040323044913f
#define CRLOT_C32_2 0.92387953251128675613f
#define CRLOT_C32_3 0.83146961230254523708f
#define CRLOT_C32_4 0.70710678118654752440f
#define CRLOT_C32_5 0.55557023301960222474f
#define CRLOT_C32_6 0.38268343236508977173f
#define CRLOT_C32_7 0.19509032201612826785f

// exp(-2 pi i k/32) for k in [0,32), as compile-time selected constants.
template <int K> struct W32 {
  static constexpr int k = ((K % 32) + 32) % 32;
  static constexpr int q = k / 8;       // quadrant
  static constexpr int r = k % 8;
  CRLOT_HD static float c8(int j) {     // cos(2 pi j/32), j in [0,8]
    return j == 0 ? CRLOT_C32_0 : j == 1 ? CRLOT_C32_1 : j == 2 ? CRLOT_C32_2 : j == 3 ? CRLOT_C32_3
         : j == 4 ? CRLOT_C32_4 : j == 5 ? CRLOT_C32_5 : j == 6 ? CRLOT_C32_6 : j == 7 ? CRLOT_C32_7 : 0.0f;
  }
  CRLOT_HD static float re() { return q == 0 ? c8(r) : q == 1 ? -c8(8 - r) : q == 2 ? -c8(r) : c8(8 - r); }
  CRLOT_HD static float im() { return q == 0 ? -c8(8 - r) : q == 1 ? -c8(r) : q == 2 ? c8(8 - r) : c8(r); }
  template <bool INV> CRLOT_HD static cpx w() { return mk(re(), INV ? -im() : im()); }
};

// The butterfly "a' = a + w b, b' = a - w b" with w = exp(-/+ 2 pi i K/32), trivial cases folded.
template <int K, bool INV> CRLOT_HD void bfly2w(cpx& a, cpx& b) {
  constexpr int k = ((K % 32) + 32) % 32;
  if (k == 0) {
    const cpx t = a - b;
    a = a + b;
    b = t;
  } else if (k == 8) {
    const cpx t = sub_rot<INV>(a, b);
    a = add_rot<INV>(a, b);
    b = t;
  } else if (k == 16) {
    const cpx t = a + b;
    a = a - b;
    b = t;
  } else if (k == 24) {
    const cpx t = add_rot<INV>(a, b);
    a = sub_rot<INV>(a, b);
    b = t;
  } else {
    const cpx w = W32<k>::template w<INV>();
    const cpx t = cmul(b, w);
    b = a - t;
    a = a + t;
  }
}

// b * exp(-/+ 2 pi i K/32), trivial cases folded (general case: 2 packed instructions)
template <int K, bool INV> CRLOT_HD cpx mulw32(cpx a) {
  constexpr int k = ((K % 32) + 32) % 32;
  if (k == 0) return a;
  if (k == 8) return rot90<INV>(a);
  if (k == 16) return mk(-a.x, -a.y);
  if (k == 24) return rot90<!INV>(a);
  return cmul(a, W32<k>::template w<INV>());
}

// 4-point butterfly on (a,b,c,d) -> natural order outputs (8 packed instructions).
template <bool INV> CRLOT_HD void bfly4(cpx& a, cpx& b, cpx& c, cpx& d) {
  const cpx s0 = a + c, s1 = a - c, s2 = b + d, t = b - d;
  a = s0 + s2;
  c = s0 - s2;
  b = add_rot<INV>(s1, t);
  d = sub_rot<INV>(s1, t);
}

template <int R, bool INV> struct Dft;

template <bool INV> struct Dft<2, INV> {
  CRLOT_HD static void run(cpx (&v)[2]) { bfly2w<0, INV>(v[0], v[1]); }
};
template <bool INV> struct Dft<4, INV> {
  CRLOT_HD static void run(cpx (&v)[4]) { bfly4<INV>(v[0], v[1], v[2], v[3]); }
};
// R = 8: radix 2 on two 4-point transforms, twiddles fused into the butterflies.
template <bool INV> struct Dft<8, INV> {
  CRLOT_HD static void run(cpx (&v)[8]) {
    cpx e[4] = {v[0], v[2], v[4], v[6]};
    cpx o[4] = {v[1], v[3], v[5], v[7]};
    Dft<4, INV>::run(e);
    Dft<4, INV>::run(o);
    bfly2w<0, INV>(e[0], o[0]);
    bfly2w<4, INV>(e[1], o[1]);
    bfly2w<8, INV>(e[2], o[2]);
    bfly2w<12, INV>(e[3], o[3]);
    v[0] = e[0]; v[1] = e[1]; v[2] = e[2]; v[3] = e[3];
    v[4] = o[0]; v[5] = o[1]; v[6] = o[2]; v[7] = o[3];
  }
};
// R = 4*M (16, 32): four M-point transforms of x[4j+q], twiddle, M radix-4 butterflies.
template <int R, bool INV> struct Dft {
  static constexpr int M = R / 4;
  static constexpr int S = 32 / R;   // twiddle stride in units of 2 pi/32
  template <int K> CRLOT_HD static void col(cpx (&v)[R], const cpx (&y0)[M], const cpx (&y1)[M],
                                            const cpx (&y2)[M], const cpx (&y3)[M]) {
    cpx a = y0[K];
    cpx b = mulw32<S * K, INV>(y1[K]);
    cpx c = mulw32<2 * S * K, INV>(y2[K]);
    cpx d = mulw32<3 * S * K, INV>(y3[K]);
    bfly4<INV>(a, b, c, d);
    v[K] = a; v[K + M] = b; v[K + 2 * M] = c; v[K + 3 * M] = d;
    if constexpr (K + 1 < M) col<K + 1>(v, y0, y1, y2, y3);
  }
  CRLOT_HD static void run(cpx (&v)[R]) {
    static_assert(R == 16 || R == 32, "register DFT sizes: 2,4,8,16,32");
    cpx y0[M], y1[M], y2[M], y3[M];
#pragma unroll
    for (int j = 0; j < M; ++j) {
      y0[j] = v[4 * j]; y1[j] = v[4 * j + 1]; y2[j] = v[4 * j + 2]; y3[j] = v[4 * j + 3];
    }
    Dft<M, INV>::run(y0);
    Dft<M, INV>::run(y1);
    Dft<M, INV>::run(y2);
    Dft<M, INV>::run(y3);
    col<0>(v, y0, y1, y2, y3);
  }
};

// ---- 64-point transform (N = 4096 kernels: one 64-point pass per lane) -------------------------
// cos(2 pi n/64), n = 0..16
CRLOT_HD constexpr float c64(int n) {
  return n == 0 ? 1.0f : n == 1 ? 0.99518472667219688624f : n == 2 ? 0.98078528040323044913f
       : n == 3 ? 0.95694033573220886494f : n == 4 ? 0.92387953251128675613f : n == 5 ? 0.88192126434835502971f
       : n == 6 ? 0.83146961230254523708f : n == 7 ? 0.77301045336273696081f : n == 8 ? 0.70710678118654752440f
       : n == 9 ? 0.63439328416364549822f : n == 10 ? 0.55557023301960222474f : n == 11 ? 0.47139673682599764856f
       : n == 12 ? 0.38268343236508977173f : n == 13 ? 0.29028467725446236764f : n == 14 ? 0.19509032201612826785f
       : n == 15 ? 0.09801714032956060199f : 0.0f;
}
// cos / sin of 2 pi m/64 for any m in [0, 64)
CRLOT_HD constexpr float cos64(int m) {
  return m <= 16 ? c64(m) : m <= 32 ? -c64(32 - m) : m <= 48 ? -c64(m - 32) : c64(64 - m);
}
CRLOT_HD constexpr float sin64(int m) {
  return m <= 16 ? c64(16 - m) : m <= 32 ? c64(m - 16) : m <= 48 ? -c64(48 - m) : -c64(m - 48);
}
// a * exp(-/+ 2 pi i K/64), trivial cases folded
template <int K, bool INV> CRLOT_HD cpx mulw64(cpx a) {
  constexpr int k = ((K % 64) + 64) % 64;
  if (k == 0) return a;
  if (k == 16) return rot90<INV>(a);
  if (k == 32) return mk(-a.x, -a.y);
  if (k == 48) return rot90<!INV>(a);
  return cmul(a, mk(cos64(k), INV ? sin64(k) : -sin64(k)));
}
// R = 64: four 16-point transforms of x[4j+q], twiddle W_64^(qK), 16 radix-4 butterflies.
template <bool INV> struct Dft<64, INV> {
  template <int K> CRLOT_HD static void col(cpx (&v)[64], const cpx (&y0)[16], const cpx (&y1)[16],
                                            const cpx (&y2)[16], const cpx (&y3)[16]) {
    cpx a = y0[K];
    cpx b = mulw64<K, INV>(y1[K]);
    cpx c = mulw64<2 * K, INV>(y2[K]);
    cpx d = mulw64<3 * K, INV>(y3[K]);
    bfly4<INV>(a, b, c, d);
    v[K] = a; v[K + 16] = b; v[K + 32] = c; v[K + 48] = d;
    if constexpr (K + 1 < 16) col<K + 1>(v, y0, y1, y2, y3);
  }
  CRLOT_HD static void run(cpx (&v)[64]) {
    cpx y0[16], y1[16], y2[16], y3[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      y0[j] = v[4 * j]; y1[j] = v[4 * j + 1]; y2[j] = v[4 * j + 2]; y3[j] = v[4 * j + 3];
    }
    Dft<16, INV>::run(y0);
    Dft<16, INV>::run(y1);
    Dft<16, INV>::run(y2);
    Dft<16, INV>::run(y3);
    col<0>(v, y0, y1, y2, y3);
  }
};

// DFT over a strided sub-array living inside a larger register file: v[OFF + STRIDE*i], i < R.
template <int R, bool INV, int OFF, int STRIDE, int TOTAL> CRLOT_HD void dft_sub(cpx (&v)[TOTAL]) {
  cpx t[R];
#pragma unroll
  for (int i = 0; i < R; ++i) t[i] = v[OFF + STRIDE * i];
  Dft<R, INV>::run(t);
#pragma unroll
  for (int i = 0; i < R; ++i) v[OFF + STRIDE * i] = t[i];
}

}  // namespace crlot
