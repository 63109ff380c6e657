// k_generic.cu -- the any-size forms of kernels 1 and 2 and of the complex plan (sm_100a).
//
// The register-FFT kernels (k1_stft.cu, k2_istft.cu, k_cfft.cu) take the five power-of-two frame
// sizes the benchmark configurations name.  The reference's plan, however, accepts every even
// real size and every complex size (reference: dsp/fft/backends/kissfft_adapter.cc:13-39; its own
// tests use nfft = 128 real and 513 complex, tests/fft_test.cc:223-248,450-520), because kissfft
// is a mixed-radix transform with a generic butterfly for any prime factor (third_party/kissfft/
// kiss_fft.c:236-328).  These kernels give the drop-in the same domain ON THE GPU (there is no CPU
// path behind any entry point): one thread block per row, the row in shared memory,
// a Stockham autosort transform whose stages are the prime factors of the length (4s first): radix 4, 2,
// 3 and 5 stages as one thread per butterfly (exact +-i rotations, like kissfft's kf_bfly4/2/3/5), any
// other prime factor with each output computed by one thread as an R-term sum (kf_bfly_generic).
// They are the compatibility path, not the benchmark path: a row costs a block-wide barrier per
// stage and the twiddles come from global memory through L1.
//
// Arithmetic follows the reference's definition, not its operation order (the bar for spectra and
// audio is 1e-5 relative L2): twiddles are double-precision cos/sin rounded to float like
// kiss_fft_alloc's (kiss_fft.c:357-363); the real transform is the half-length complex transform
// plus kiss_fftr's split (kiss_fftr.c:63-155), including its exact zeros in the imaginary parts of
// bins 0 and N/2 and its use of only the real parts of those bins in the inverse; clean-up rule and
// 1/N scaling as in the adapter (kissfft_adapter.cc:104-108,150-165,227-239).
#include "common.cuh"
#include "kernels.h"

namespace crlot {

namespace {

constexpr int kGenMaxStages = 16;
constexpr int kGenMaxLen = 8192;      // complex points per row: two ping-pong rows of 64 KB

struct GenStages {
  int len;                       // complex length M
  int count;
  int radix[kGenMaxStages];      // product == len
  // exact division of an index < 2^16 by the stage's stride s and by its radix R without a divide:
  // floor(n / d) == umulhi(n, floor(2^32 / d) + 1) for n, d < 2^16
  unsigned int mul_s[kGenMaxStages], mul_r[kGenMaxStages];
};

GenStages factorize(int len) {
  GenStages g{};
  g.len = len;
  int n = len;
  auto push = [&](int r) { if (g.count < kGenMaxStages) g.radix[g.count] = r; ++g.count; };
  while (n % 4 == 0) { push(4); n /= 4; }
  while (n % 2 == 0) { push(2); n /= 2; }
  for (int r = 3; r * r <= n; r += 2)
    while (n % r == 0) { push(r); n /= r; }
  if (n > 1) push(n);
  unsigned int s = 1;
  for (int i = 0; i < g.count && i < kGenMaxStages; ++i) {
    g.mul_s[i] = static_cast<unsigned int>((1ull << 32) / s + 1);      // s == 1: 2^32 + 1 wraps to 1, handled in the kernel
    g.mul_r[i] = static_cast<unsigned int>((1ull << 32) / static_cast<unsigned int>(g.radix[i]) + 1);
    s *= static_cast<unsigned int>(g.radix[i]);
  }
  return g;
}

// Stockham autosort, decimation in frequency.  A stage of radix R on sub-length n = R*m with stride s
// (n*s == M) maps   y[q + s*(R*p + j)] = W_n^(p*j) * sum_k x[q + s*(p + m*k)] * W_R^(j*k),
// q < s, p < m, j < R; then n <- m, s <- s*R.  W_n^(p*j) = tw[p*j*s], W_R^(j*k) = tw[((j*k) mod R) * (M/R)],
// tw[i] = exp(-2 pi i / M) (conjugated for the inverse).  Returns the buffer that holds the result.
template <bool INV>
__device__ cpx* gen_stages(cpx* x, cpx* y, const GenStages& g, const cpx* __restrict__ tw) {
  const int M = g.len;
  int n = M, s = 1;
  for (int st = 0; st < g.count; ++st) {
    const int R = g.radix[st];
    const int m = n / R, wr = M / R;
    const unsigned int ms = g.mul_s[st], mr = g.mul_r[st];
    if (R == 4) {
      // one thread per butterfly: four inputs, the 4-point DFT with exact +-i rotations, three twiddles
      for (int bt = threadIdx.x; bt < M / 4; bt += blockDim.x) {
        const int p = s == 1 ? bt : static_cast<int>(__umulhi(static_cast<unsigned int>(bt), ms));
        const int q = bt - p * s;
        const cpx* xin = x + q + s * p;
        const cpx a0 = xin[0], a1 = xin[s * m], a2 = xin[2 * s * m], a3 = xin[3 * s * m];
        const cpx e0 = a0 + a2, e1 = a0 - a2, o0 = a1 + a3, o1 = a1 - a3;
        const cpx r = INV ? mk(-o1.y, o1.x) : mk(o1.y, -o1.x);          // -i*o1 forward, +i*o1 inverse
        const cpx b0 = e0 + o0, b1 = e1 + r, b2 = e0 - o0, b3 = e1 - r;
        const int w1 = p * s;
        const cpx t1 = tw[w1], t2 = tw[2 * w1], t3 = tw[3 * w1];
        cpx* yo = y + q + 4 * s * p;
        yo[0] = b0;
        yo[s] = INV ? cmulc(b1, t1) : cmul(b1, t1);
        yo[2 * s] = INV ? cmulc(b2, t2) : cmul(b2, t2);
        yo[3 * s] = INV ? cmulc(b3, t3) : cmul(b3, t3);
      }
    } else if (R == 2) {
      for (int bt = threadIdx.x; bt < M / 2; bt += blockDim.x) {
        const int p = s == 1 ? bt : static_cast<int>(__umulhi(static_cast<unsigned int>(bt), ms));
        const int q = bt - p * s;
        const cpx a0 = x[q + s * p], a1 = x[q + s * p + s * m];
        const cpx t1 = tw[p * s], d = a0 - a1;
        y[q + 2 * s * p] = a0 + a1;
        y[q + 2 * s * p + s] = INV ? cmulc(d, t1) : cmul(d, t1);
      }
    } else if (R == 3) {
      const float c = 0.86602540378443864676f;                          // sin(2 pi / 3)
      for (int bt = threadIdx.x; bt < M / 3; bt += blockDim.x) {
        const int p = s == 1 ? bt : static_cast<int>(__umulhi(static_cast<unsigned int>(bt), ms));
        const int q = bt - p * s;
        const cpx* xin = x + q + s * p;
        const cpx a0 = xin[0], a1 = xin[s * m], a2 = xin[2 * s * m];
        const cpx t = a1 + a2, u = a1 - a2;
        const cpx m1 = fma2(t, splat(-0.5f), a0);
        const cpx m2 = INV ? mk(-c * u.y, c * u.x) : mk(c * u.y, -c * u.x);   // -+ i sin(2 pi/3) u
        const cpx b1 = m1 + m2, b2 = m1 - m2;
        const int w1 = p * s;
        const cpx t1 = tw[w1], t2 = tw[2 * w1];
        cpx* yo = y + q + 3 * s * p;
        yo[0] = a0 + t;
        yo[s] = INV ? cmulc(b1, t1) : cmul(b1, t1);
        yo[2 * s] = INV ? cmulc(b2, t2) : cmul(b2, t2);
      }
    } else if (R == 5) {
      const float c1 = 0.30901699437494742410f, c2 = -0.80901699437494742410f;   // cos(2 pi/5), cos(4 pi/5)
      const float s1 = 0.95105651629515357212f, s2 = 0.58778525229247312917f;    // sin(2 pi/5), sin(4 pi/5)
      for (int bt = threadIdx.x; bt < M / 5; bt += blockDim.x) {
        const int p = s == 1 ? bt : static_cast<int>(__umulhi(static_cast<unsigned int>(bt), ms));
        const int q = bt - p * s;
        const cpx* xin = x + q + s * p;
        const cpx a0 = xin[0], a1 = xin[s * m], a2 = xin[2 * s * m], a3 = xin[3 * s * m], a4 = xin[4 * s * m];
        const cpx t1 = a1 + a4, t2 = a2 + a3, t3 = a1 - a4, t4 = a2 - a3;
        const cpx m1 = fma2(t2, splat(c2), fma2(t1, splat(c1), a0));
        const cpx m2 = fma2(t2, splat(c1), fma2(t1, splat(c2), a0));
        const cpx n1 = fma2(t4, splat(s2), mul2(t3, splat(s1)));
        const cpx n2 = fma2(t4, splat(-s1), mul2(t3, splat(s2)));
        // forward: b1 = m1 - i n1, b4 = m1 + i n1, b2 = m2 - i n2, b3 = m2 + i n2; inverse: signs swapped
        const cpx r1 = INV ? mk(-n1.y, n1.x) : mk(n1.y, -n1.x);
        const cpx r2 = INV ? mk(-n2.y, n2.x) : mk(n2.y, -n2.x);
        const cpx b1 = m1 + r1, b4 = m1 - r1, b2 = m2 + r2, b3 = m2 - r2;
        const int w1 = p * s;
        const cpx u1 = tw[w1], u2 = tw[2 * w1], u3 = tw[3 * w1], u4 = tw[4 * w1];
        cpx* yo = y + q + 5 * s * p;
        yo[0] = a0 + t1 + t2;
        yo[s] = INV ? cmulc(b1, u1) : cmul(b1, u1);
        yo[2 * s] = INV ? cmulc(b2, u2) : cmul(b2, u2);
        yo[3 * s] = INV ? cmulc(b3, u3) : cmul(b3, u3);
        yo[4 * s] = INV ? cmulc(b4, u4) : cmul(b4, u4);
      }
    } else
    for (int idx = threadIdx.x; idx < M; idx += blockDim.x) {
      const int t = s == 1 ? idx : static_cast<int>(__umulhi(static_cast<unsigned int>(idx), ms));   // idx / s
      const int q = idx - t * s;
      const int p = static_cast<int>(__umulhi(static_cast<unsigned int>(t), mr));                     // t / R
      const int j = t - p * R;
      const cpx* xin = x + q + s * p;
      cpx acc = xin[0];
      int jk = 0;
      for (int k = 1; k < R; ++k) {
        jk += j;
        if (jk >= R) jk -= R;
        const cpx w = tw[jk * wr];
        const cpx a = xin[s * m * k];
        acc = INV ? cmulc(a, w) + acc : cfma(a, w, acc);
      }
      const cpx w = tw[p * j * s];
      y[idx] = INV ? cmulc(acc, w) : cmul(acc, w);
    }
    __syncthreads();
    cpx* tmp = x; x = y; y = tmp;
    n = m;
    s *= R;
  }
  return x;
}

// ---- kernel 1, any even N: frame + window + clean-up + forward real transform ---------------------
__global__ void __launch_bounds__(256) stft_fwd_generic_kernel(const K1Params p, const GenStages g) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int M = g.len;
  cpx* bufa = reinterpret_cast<cpx*>(smem_raw);
  cpx* bufb = bufa + M;
  const long long rows = p.frames * p.channels;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const long long c = row / p.frames, k = row - c * p.frames;
    const float* sig = p.sig + c * p.chan_stride;
    const long long base = k * static_cast<long long>(p.hop) - p.pad;     // signed index into the signal
    __syncthreads();                                                      // previous row's reads are complete
    for (int i = threadIdx.x; i < M; i += blockDim.x) {
      float a = __fmul_rn(pad_value(sig, p.len, base + 2 * i, p.pad_mode), __ldg(p.window + 2 * i));
      float b = __fmul_rn(pad_value(sig, p.len, base + 2 * i + 1, p.pad_mode), __ldg(p.window + 2 * i + 1));
      if (p.scrub) { a = scrub(a); b = scrub(b); }
      bufa[i] = mk(a, b);
    }
    __syncthreads();
    const cpx* z = gen_stages<false>(bufa, bufb, g, p.tw1);
    cpx* out = p.spec + (c * p.spec_chan_rows + k) * (M + 1);
    // kiss_fftr's split; pair (k, M-k), the self-mirrored bin M/2 written once with the mirror formula
    for (int b = threadIdx.x; b <= M / 2; b += blockDim.x) {
      if (b == 0) {
        out[0] = mk(z[0].x + z[0].y, 0.0f);
        out[M] = mk(z[0].x - z[0].y, 0.0f);
        continue;
      }
      const cpx zk = z[b], zm = conj(z[M - b]);
      const cpx e = zk + zm, d = zk - zm;
      const cpx t = cmul(d, p.sa[b]);
      if (b != M - b) out[b] = mk(0.5f * (e.x + t.x), 0.5f * (e.y + t.y));
      out[M - b] = mk(0.5f * (e.x - t.x), 0.5f * (t.y - e.y));
    }
  }
}

// ---- kernel 2, any even N: (hook) + inverse real transform * 1/N + clean-up (+ synthesis window) --
__global__ void __launch_bounds__(256) istft_inv_generic_kernel(const K2Params p, const GenStages g) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int M = g.len, N = 2 * M;
  cpx* bufa = reinterpret_cast<cpx*>(smem_raw);
  cpx* bufb = bufa + M;
  const float scale = 1.0f / static_cast<float>(N);
  for (long long row = blockIdx.x; row < p.rows; row += gridDim.x) {
    const cpx* X = p.spec + row * (M + 1);
    const float* gn = p.sgain ? p.sgain + row * p.sgain_row_stride : nullptr;
    __syncthreads();
    for (int b = threadIdx.x; b <= M / 2; b += blockDim.x) {
      if (b == 0) {
        float x0 = X[0].x, xm = X[M].x;                                   // only the real parts are used
        if (gn) { x0 = x0 * gn[0]; xm = xm * gn[M]; }
        bufa[0] = mk(x0 + xm, x0 - xm);
        continue;
      }
      cpx fk = X[b], fm = X[M - b];
      if (gn) { fk = mul2(fk, splat(gn[b])); fm = mul2(fm, splat(gn[M - b])); }
      const cpx fmc = conj(fm);
      const cpx e = fk + fmc, d = fk - fmc;
      const cpx o = cmulc(d, p.sa[b]);
      bufa[b] = e + o;
      if (b != M - b) bufa[M - b] = conj(e - o);
    }
    __syncthreads();
    const cpx* z = gen_stages<true>(bufa, bufb, g, p.tw1);
    float* out = p.frames + row * N;
    for (int i = threadIdx.x; i < M; i += blockDim.x) {
      float a = z[i].x * scale, b = z[i].y * scale;
      if (p.scrub) { a = scrub(a); b = scrub(b); }
      if (p.window) { a = __fmul_rn(a, __ldg(p.window + 2 * i)); b = __fmul_rn(b, __ldg(p.window + 2 * i + 1)); }
      out[2 * i] = a;
      out[2 * i + 1] = b;
    }
  }
}

// ---- complex plan, any n ----------------------------------------------------------------------------
template <bool INV>
__global__ void __launch_bounds__(256) cfft_generic_kernel(const CfftParams p, const GenStages g) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int M = g.len;
  cpx* bufa = reinterpret_cast<cpx*>(smem_raw);
  cpx* bufb = bufa + M;
  const float scale = 1.0f / static_cast<float>(M);
  for (long long row = blockIdx.x; row < p.rows; row += gridDim.x) {
    __syncthreads();
    for (int i = threadIdx.x; i < M; i += blockDim.x) bufa[i] = p.in[row * M + i];
    __syncthreads();
    const cpx* z = gen_stages<INV>(bufa, bufb, g, p.tw1);
    for (int i = threadIdx.x; i < M; i += blockDim.x) {
      cpx y = z[i];
      if (INV) y = mk(scrub(y.x * scale), scrub(y.y * scale));
      p.out[row * M + i] = y;
    }
  }
}

// a radix-4 stage has len/4 butterflies: a quarter of the row length in threads keeps every thread busy in
// the butterfly stages; short rows then leave room for more resident blocks (rows in flight) per SM
int gen_threads(int len) {
  int t = (len / 4 + 31) / 32 * 32;
  return t < 32 ? 32 : (t > 256 ? 256 : t);
}

template <class Kern>
cudaError_t gen_prepare(Kern kern, int len, int* smem) {
  *smem = 2 * len * static_cast<int>(sizeof(cpx));
  if (*smem > 48 * 1024) return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, *smem);
  return cudaSuccess;
}

int gen_grid(long long rows, int len) {
  const int per_sm = 2048 / gen_threads(len) < 32 ? 2048 / gen_threads(len) : 32;
  const long long cap = static_cast<long long>(per_sm) * device_sm_count();
  return static_cast<int>(rows < cap ? (rows > 0 ? rows : 1) : cap);
}

}  // namespace

bool generic_len_supported(int complex_len) { return complex_len >= 1 && complex_len <= kGenMaxLen; }

cudaError_t launch_stft_fwd_generic(const K1Params& p, cudaStream_t st) {
  if (p.frames <= 0 || p.channels <= 0) return cudaSuccess;
  if (p.nfft < 2 || p.nfft % 2 != 0 || !generic_len_supported(p.nfft / 2)) return cudaErrorNotSupported;
  const GenStages g = factorize(p.nfft / 2);
  if (g.count > kGenMaxStages) return cudaErrorNotSupported;
  int smem = 0;
  cudaError_t e = gen_prepare(stft_fwd_generic_kernel, g.len, &smem);
  if (e != cudaSuccess) return e;
  stft_fwd_generic_kernel<<<gen_grid(p.frames * p.channels, g.len), gen_threads(g.len), smem, st>>>(p, g);
  return cudaGetLastError();
}

cudaError_t launch_istft_inv_generic(const K2Params& p, cudaStream_t st) {
  if (p.rows <= 0) return cudaSuccess;
  if (p.nfft < 2 || p.nfft % 2 != 0 || !generic_len_supported(p.nfft / 2)) return cudaErrorNotSupported;
  const GenStages g = factorize(p.nfft / 2);
  if (g.count > kGenMaxStages) return cudaErrorNotSupported;
  int smem = 0;
  cudaError_t e = gen_prepare(istft_inv_generic_kernel, g.len, &smem);
  if (e != cudaSuccess) return e;
  istft_inv_generic_kernel<<<gen_grid(p.rows, g.len), gen_threads(g.len), smem, st>>>(p, g);
  return cudaGetLastError();
}

cudaError_t launch_cfft_generic(const CfftParams& p, cudaStream_t st) {
  if (p.rows <= 0) return cudaSuccess;
  if (!generic_len_supported(p.n)) return cudaErrorNotSupported;
  const GenStages g = factorize(p.n);
  if (g.count > kGenMaxStages) return cudaErrorNotSupported;
  int smem = 0;
  cudaError_t e;
  if (p.inverse) {
    e = gen_prepare(cfft_generic_kernel<true>, g.len, &smem);
    if (e != cudaSuccess) return e;
    cfft_generic_kernel<true><<<gen_grid(p.rows, g.len), gen_threads(g.len), smem, st>>>(p, g);
  } else {
    e = gen_prepare(cfft_generic_kernel<false>, g.len, &smem);
    if (e != cudaSuccess) return e;
    cfft_generic_kernel<false><<<gen_grid(p.rows, g.len), gen_threads(g.len), smem, st>>>(p, g);
  }
  return cudaGetLastError();
}

}  // namespace crlot
