// regfft.cuh -- straight-line complex DFTs of 2..32 points held entirely in registers.
//
// Every index below is a compile-time constant after unrolling, so `cpx v[R]` never touches
// local memory: a "permutation" is only a renaming of registers.  The same header compiles
// as plain C++ (CRLOT_HD empty) for the CPU lane-emulator in tests/cpu_emu, which checks
// the index algebra of the warp kernels without a GPU.
//
// Convention: dft<R,false> computes X[k] = sum_n x[n] exp(-2 pi i n k / R)  (forward),
//             dft<R,true>  the conjugate kernel (unnormalised inverse).  Natural order in
//             and out.
#pragma once

#if defined(__CUDACC__)
#define CRLOT_HD __host__ __device__ __forceinline__
#else
#define CRLOT_HD inline
#endif

namespace crlot {

struct alignas(8) cpx {
  float x, y;
};

CRLOT_HD cpx mk(float a, float b) { cpx r; r.x = a; r.y = b; return r; }
CRLOT_HD cpx operator+(cpx a, cpx b) { return mk(a.x + b.x, a.y + b.y); }
CRLOT_HD cpx operator-(cpx a, cpx b) { return mk(a.x - b.x, a.y - b.y); }
CRLOT_HD cpx cmul(cpx a, cpx b) { return mk(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
CRLOT_HD cpx cmulc(cpx a, cpx b) { return mk(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y); }  // a*conj(b)
CRLOT_HD cpx conj(cpx a) { return mk(a.x, -a.y); }
// multiply by -i (forward quarter turn) or +i
template <bool INV> CRLOT_HD cpx rot90(cpx a) { return INV ? mk(-a.y, a.x) : mk(a.y, -a.x); }

// cos(2 pi k/32), sin(2 pi k/32), k = 0..8 (first octant + 1); everything else by symmetry.
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
  // angle = -(q*90 + r*11.25) degrees.  base = (cos r, -sin r); each quadrant multiplies by -i.
  CRLOT_HD static float re() {
    return q == 0 ? c8(r) : q == 1 ? -c8(8 - r) : q == 2 ? -c8(r) : c8(8 - r);
  }
  CRLOT_HD static float im() {
    return q == 0 ? -c8(8 - r) : q == 1 ? -c8(r) : q == 2 ? c8(8 - r) : c8(r);
  }
};

// a * exp(-/+ 2 pi i K/32) with the free cases folded.
template <int K, bool INV> CRLOT_HD cpx mulw32(cpx a) {
  constexpr int k = ((K % 32) + 32) % 32;
  if (k == 0) return a;
  if (k == 8) return rot90<INV>(a);
  if (k == 16) return mk(-a.x, -a.y);
  if (k == 24) return rot90<!INV>(a);
  if (k == 4) {   // (1 -/+ i)/sqrt2
    const float h = CRLOT_C32_4;
    return INV ? mk(h * (a.x - a.y), h * (a.x + a.y)) : mk(h * (a.x + a.y), h * (a.y - a.x));
  }
  if (k == 12) {  // (-1 -/+ i)/sqrt2
    const float h = CRLOT_C32_4;
    return INV ? mk(-h * (a.x + a.y), h * (a.x - a.y)) : mk(h * (a.y - a.x), -h * (a.x + a.y));
  }
  if (k == 20) {
    const float h = CRLOT_C32_4;
    return INV ? mk(h * (a.y - a.x), -h * (a.x + a.y)) : mk(-h * (a.x + a.y), h * (a.x - a.y));
  }
  if (k == 28) {
    const float h = CRLOT_C32_4;
    return INV ? mk(h * (a.x + a.y), h * (a.y - a.x)) : mk(h * (a.x - a.y), h * (a.x + a.y));
  }
  const cpx w = mk(W32<k>::re(), INV ? -W32<k>::im() : W32<k>::im());
  return cmul(a, w);
}

// 4-point butterfly on (a,b,c,d) -> natural order outputs.
template <bool INV> CRLOT_HD void bfly4(cpx& a, cpx& b, cpx& c, cpx& d) {
  const cpx s0 = a + c, s1 = a - c, s2 = b + d, s3 = rot90<INV>(b - d);
  a = s0 + s2;
  c = s0 - s2;
  b = s1 + s3;
  d = s1 - s3;
}

template <int R, bool INV> struct Dft;

template <bool INV> struct Dft<2, INV> {
  CRLOT_HD static void run(cpx (&v)[2]) {
    const cpx t = v[0] - v[1];
    v[0] = v[0] + v[1];
    v[1] = t;
  }
};
template <bool INV> struct Dft<4, INV> {
  CRLOT_HD static void run(cpx (&v)[4]) { bfly4<INV>(v[0], v[1], v[2], v[3]); }
};

// R = 4*M, decimation in time: four M-point transforms of x[4j+q], twiddle, M radix-4 butterflies.
// R = 8 uses radix 2 on two 4-point transforms.
template <bool INV> struct Dft<8, INV> {
  CRLOT_HD static void run(cpx (&v)[8]) {
    cpx e[4] = {v[0], v[2], v[4], v[6]};
    cpx o[4] = {v[1], v[3], v[5], v[7]};
    Dft<4, INV>::run(e);
    Dft<4, INV>::run(o);
    const cpx t0 = o[0], t1 = mulw32<4, INV>(o[1]), t2 = mulw32<8, INV>(o[2]), t3 = mulw32<12, INV>(o[3]);
    v[0] = e[0] + t0; v[4] = e[0] - t0;
    v[1] = e[1] + t1; v[5] = e[1] - t1;
    v[2] = e[2] + t2; v[6] = e[2] - t2;
    v[3] = e[3] + t3; v[7] = e[3] - t3;
  }
};

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

// DFT over a strided sub-array living inside a larger register file: elements
// v[OFF + STRIDE*i], i < R.
template <int R, bool INV, int OFF, int STRIDE, int TOTAL> CRLOT_HD void dft_sub(cpx (&v)[TOTAL]) {
  cpx t[R];
#pragma unroll
  for (int i = 0; i < R; ++i) t[i] = v[OFF + STRIDE * i];
  Dft<R, INV>::run(t);
#pragma unroll
  for (int i = 0; i < R; ++i) v[OFF + STRIDE * i] = t[i];
}

}  // namespace crlot
