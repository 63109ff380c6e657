/* oracle/crlot_oracle.c -- CPU ORACLE.  TEST INFRASTRUCTURE, NOT PRODUCT.
 * See crlot_oracle.h for the contract and how parity is pinned.
 * Build with -ffp-contract=off (oracle/Makefile) so every rounding below is explicit.
 * Citations are relative to /root/reference.
 */
#define _GNU_SOURCE
#include "crlot_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ------------------------------------------------------------------------------------ */
/* WindowLUT                                                                            */
/* ------------------------------------------------------------------------------------ */

/* Table in double then cast, denominator N (periodic) or N-1 (symmetric), N==1 -> 1.0
 * dsp/window/WindowLUT.cc:256-315.  Normalisations: WindowLUT.cc:317-388 (hop unknown at
 * construction, so OLA_UNITY_GAIN and OLA_SUM_WSQ both reduce to the L2 rule). */
int orc_window(int type, long n, int periodic, int norm, float* out) {
  if (n <= 0 || !out) return -1;
  if (type < ORC_WIN_HANN || type > ORC_WIN_RECT) return -1;
  if (type == ORC_WIN_RECT) {
    for (long i = 0; i < n; ++i) out[i] = 1.0f;
  } else if (n == 1) {
    out[0] = 1.0f;
  } else {
    const double den = periodic ? (double)n : (double)(n - 1);
    const double step = 2.0 * M_PI / den;
    for (long i = 0; i < n; ++i) {
      const double a = step * (double)i;
      double v;
      if (type == ORC_WIN_HANN) {
        v = 0.5 * (1.0 - cos(a));
      } else if (type == ORC_WIN_HAMMING) {
        v = 0.54 - 0.46 * cos(a);
      } else {
        const double c1 = cos(a), c2 = cos(2.0 * a);
        v = 0.42 - 0.5 * c1 + 0.08 * c2;
      }
      out[i] = (float)v;
    }
  }
  if (norm == ORC_NORM_SUM_TO_ONE) {
    double s = 0.0;
    for (long i = 0; i < n; ++i) s += (double)out[i];
    if (s > 0.0) {
      const float k = (float)(1.0 / s);
      for (long i = 0; i < n; ++i) out[i] *= k;
    }
  } else if (norm == ORC_NORM_L2 || norm == ORC_NORM_OLA_UNITY || norm == ORC_NORM_OLA_SUM_WSQ) {
    double s = 0.0;
    for (long i = 0; i < n; ++i) s += (double)out[i] * (double)out[i];
    if (s > 0.0) {
      const float k = (float)(1.0 / sqrt(s));
      for (long i = 0; i < n; ++i) out[i] *= k;
    }
  } else if (norm != ORC_NORM_NONE) {
    return -1;
  }
  return 0;
}

/* ------------------------------------------------------------------------------------ */
/* Framer (one-shot push, then pop until empty)                                         */
/* ------------------------------------------------------------------------------------ */

/* Closed form of what repeated pop() yields after ONE push of `len` sample-frames
 * (dsp/frame/framer.cc:88-118,128-182): read_pos advances by hop per pop (clamped to
 * write_pos); a pop succeeds while at least `frame` samples remain, and in ZERO_PAD mode
 * also while at least one sample remains (zero-tailed frame).  DROP: floor((L-N)/H)+1;
 * ZERO_PAD: ceil(L/H). */
long orc_frame_count(long len, long frame, long hop, int mode) {
  if (len <= 0 || frame <= 0 || hop <= 0) return 0;
  if (mode == ORC_DROP) return len < frame ? 0 : (len - frame) / hop + 1;
  return (len + hop - 1) / hop;
}

long orc_framer(const float* x, long frames_in, long frame, long hop, long channels, int mode,
                float* frames_out, long max_frames) {
  const long total = orc_frame_count(frames_in, frame, hop, mode);
  if (!frames_out) return total;
  for (long k = 0; k < total && k < max_frames; ++k) {
    float* dst = frames_out + k * frame * channels;
    const long start = k * hop;
    long have = frames_in - start;
    if (have > frame) have = frame;
    memcpy(dst, x + start * channels, sizeof(float) * (size_t)(have * channels));
    for (long j = have * channels; j < frame * channels; ++j) dst[j] = 0.0f;
  }
  return total;
}

/* ------------------------------------------------------------------------------------ */
/* FrameQueue (centre / pad front end) restated: dsp/frame/FrameQueue.cc:9-48,68-115,      */
/* dsp/frame/Indexing.h:18-70.  PadMode: 0 CONSTANT (zeros), 1 REFLECT, 2 EDGE.            */
/* ------------------------------------------------------------------------------------ */

/* Indexing.h:18-34 -- note the asymmetric negative branch (-i-1), pinned by                */
/* tests/frame_queue_test.cc:264-366 of the reference                                       */
static long orc_reflect101(long i, long n) {
  if (n <= 1) return 0;
  while (i < 0 || i >= n) {
    if (i < 0) i = -i - 1;
    else i = 2 * n - 2 - i;
  }
  return i;
}

/* Indexing.h:46-70: value of the (virtually) padded signal at signed index idx */
float orc_pad_value(const float* x, long len, long idx, int pad_mode) {
  if (len <= 0) return 0.0f;
  if (idx >= 0 && idx < len) return x[idx];
  switch (pad_mode) {
    case 2: return idx < 0 ? x[0] : x[len - 1];
    case 1: return x[orc_reflect101(idx, len)];
    default: return 0.0f;
  }
}

/* FrameQueue.cc:100-115: padded_len = len + 2*(frame/2) when centred; frames that fit */
long orc_framequeue_count(long len, long frame, long hop, int center) {
  if (frame <= 0 || hop <= 0 || len < 0) return -1;
  const long padded = center ? len + 2 * (frame / 2) : len;
  if (padded < frame) return 0;
  const long tail = frame > hop ? frame - hop : 0;
  if (padded < tail) return 0;
  return (padded - tail) / hop;
}

/* frame k, sample i = padded[k*hop + i], padded[p] = pad_value(p - pad) (FrameQueue.cc:28-46,68-98) */
long orc_framequeue(const float* x, long len, long frame, long hop, int center, int pad_mode, float* frames_out,
                    long max_frames) {
  const long total = orc_framequeue_count(len, frame, hop, center);
  if (total < 0 || !frames_out) return total;
  const long pad = center ? frame / 2 : 0;
  for (long k = 0; k < total && k < max_frames; ++k)
    for (long i = 0; i < frame; ++i) frames_out[k * frame + i] = orc_pad_value(x, len, k * hop + i - pad, pad_mode);
  return total;
}

/* ------------------------------------------------------------------------------------ */
/* kissfft restated: mixed-radix decimation-in-time, float, twiddles from double cos/sin */
/* third_party/kissfft/kiss_fft.c:15-363, kiss_fftr.c:22-155, _kiss_fft_guts.h:87-145     */
/* ------------------------------------------------------------------------------------ */

typedef struct { float r, i; } cpx;

typedef struct {
  int n;          /* complex length                                   */
  int inverse;    /* twiddle sign                                     */
  int fac[64];    /* (radix, remaining) pairs, kiss_fft.c:307-328     */
  cpx* tw;        /* n twiddles exp(-/+ 2 pi i k / n), kiss_fft.c:357-363 */
} cplan;

static inline cpx cmul(cpx a, cpx b) {
  cpx m;
  m.r = a.r * b.r - a.i * b.i;
  m.i = a.r * b.i + a.i * b.r;
  return m;
}
static inline cpx cadd(cpx a, cpx b) { cpx m = {a.r + b.r, a.i + b.i}; return m; }
static inline cpx csub(cpx a, cpx b) { cpx m = {a.r - b.r, a.i - b.i}; return m; }

static int cplan_init(cplan* p, int n, int inverse) {
  p->n = n;
  p->inverse = inverse;
  p->tw = (cpx*)malloc(sizeof(cpx) * (size_t)n);
  if (!p->tw) return -1;
  const double pi = 3.141592653589793238462643383279502884197169399375105820974944;
  for (int k = 0; k < n; ++k) {
    double ph = -2 * pi * k / n;
    if (inverse) ph *= -1;
    p->tw[k].r = (float)cos(ph);
    p->tw[k].i = (float)sin(ph);
  }
  /* factor: all 4s first, then 2s, then odd primes ascending; when the trial radix
   * exceeds floor(sqrt(n0)) the remainder itself is the last radix. */
  int rem = n, radix = 4, nf = 0;
  const double root = floor(sqrt((double)n));
  do {
    while (rem % radix) {
      if (radix == 4) radix = 2;
      else if (radix == 2) radix = 3;
      else radix += 2;
      if (radix > root) radix = rem;
    }
    rem /= radix;
    p->fac[nf++] = radix;
    p->fac[nf++] = rem;
  } while (rem > 1);
  return 0;
}
static void cplan_free(cplan* p) { free(p->tw); p->tw = NULL; }

/* One recombination stage over `m` columns: out[j + q*m], q < radix. `fs` = twiddle stride. */
static void stage2(cpx* F, int fs, const cplan* p, int m) {
  for (int j = 0; j < m; ++j) {
    const cpx t = cmul(F[m + j], p->tw[j * fs]);
    F[m + j] = csub(F[j], t);
    F[j] = cadd(F[j], t);
  }
}
static void stage4(cpx* F, int fs, const cplan* p, int m) {
  for (int j = 0; j < m; ++j) {
    const cpx a = cmul(F[m + j], p->tw[j * fs]);
    const cpx b = cmul(F[2 * m + j], p->tw[2 * j * fs]);
    const cpx c = cmul(F[3 * m + j], p->tw[3 * j * fs]);
    const cpx d = csub(F[j], b);
    F[j] = cadd(F[j], b);
    const cpx s = cadd(a, c);
    const cpx e = csub(a, c);
    F[2 * m + j] = csub(F[j], s);
    F[j] = cadd(F[j], s);
    if (p->inverse) {
      F[m + j].r = d.r - e.i;  F[m + j].i = d.i + e.r;
      F[3 * m + j].r = d.r + e.i;  F[3 * m + j].i = d.i - e.r;
    } else {
      F[m + j].r = d.r + e.i;  F[m + j].i = d.i - e.r;
      F[3 * m + j].r = d.r - e.i;  F[3 * m + j].i = d.i + e.r;
    }
  }
}
static void stage3(cpx* F, int fs, const cplan* p, int m) {
  const float ei = p->tw[fs * m].i;
  for (int j = 0; j < m; ++j) {
    const cpx a = cmul(F[m + j], p->tw[j * fs]);
    const cpx b = cmul(F[2 * m + j], p->tw[2 * j * fs]);
    const cpx s = cadd(a, b);
    cpx d = csub(a, b);
    F[m + j].r = F[j].r - s.r * 0.5f;
    F[m + j].i = F[j].i - s.i * 0.5f;
    d.r *= ei;
    d.i *= ei;
    F[j] = cadd(F[j], s);
    F[2 * m + j].r = F[m + j].r + d.i;
    F[2 * m + j].i = F[m + j].i - d.r;
    F[m + j].r -= d.i;
    F[m + j].i += d.r;
  }
}
static void stage5(cpx* F, int fs, const cplan* p, int m) {
  const cpx ya = p->tw[fs * m], yb = p->tw[fs * 2 * m];
  for (int u = 0; u < m; ++u) {
    cpx* f0 = F + u; cpx* f1 = f0 + m; cpx* f2 = f0 + 2 * m; cpx* f3 = f0 + 3 * m; cpx* f4 = f0 + 4 * m;
    const cpx s0 = *f0;
    const cpx s1 = cmul(*f1, p->tw[u * fs]);
    const cpx s2 = cmul(*f2, p->tw[2 * u * fs]);
    const cpx s3 = cmul(*f3, p->tw[3 * u * fs]);
    const cpx s4 = cmul(*f4, p->tw[4 * u * fs]);
    const cpx s7 = cadd(s1, s4), s10 = csub(s1, s4), s8 = cadd(s2, s3), s9 = csub(s2, s3);
    f0->r += s7.r + s8.r;
    f0->i += s7.i + s8.i;
    cpx s5, s6, s11, s12;
    s5.r = s0.r + s7.r * ya.r + s8.r * yb.r;
    s5.i = s0.i + s7.i * ya.r + s8.i * yb.r;
    s6.r = s10.i * ya.i + s9.i * yb.i;
    s6.i = -(s10.r * ya.i) - s9.r * yb.i;
    *f1 = csub(s5, s6);
    *f4 = cadd(s5, s6);
    s11.r = s0.r + s7.r * yb.r + s8.r * ya.r;
    s11.i = s0.i + s7.i * yb.r + s8.i * ya.r;
    s12.r = -(s10.i * yb.i) + s9.i * ya.i;
    s12.i = s10.r * yb.i - s9.r * ya.i;
    *f2 = cadd(s11, s12);
    *f3 = csub(s11, s12);
  }
}
static void stage_any(cpx* F, int fs, const cplan* p, int m, int radix) {
  cpx* tmp = (cpx*)malloc(sizeof(cpx) * (size_t)radix);
  for (int u = 0; u < m; ++u) {
    for (int q = 0, k = u; q < radix; ++q, k += m) tmp[q] = F[k];
    for (int q1 = 0, k = u; q1 < radix; ++q1, k += m) {
      int idx = 0;
      F[k] = tmp[0];
      for (int q = 1; q < radix; ++q) {
        idx += fs * k;
        if (idx >= p->n) idx -= p->n;
        F[k] = cadd(F[k], cmul(tmp[q], p->tw[idx]));
      }
    }
  }
  free(tmp);
}

/* Recursive decimation in time: split into `radix` sub-transforms of length m taken at
 * input stride fs*radix, then recombine (kiss_fft.c:236-300). */
static void cwork(cpx* out, const cpx* in, int fs, const int* fac, const cplan* p) {
  const int radix = fac[0], m = fac[1];
  if (m == 1) {
    for (int q = 0; q < radix; ++q) out[q] = in[(long)q * fs];
  } else {
    for (int q = 0; q < radix; ++q) cwork(out + (long)q * m, in + (long)q * fs, fs * radix, fac + 2, p);
  }
  switch (radix) {
    case 2: stage2(out, fs, p, m); break;
    case 3: stage3(out, fs, p, m); break;
    case 4: stage4(out, fs, p, m); break;
    case 5: stage5(out, fs, p, m); break;
    default: stage_any(out, fs, p, m, radix); break;
  }
}

typedef struct {
  int nfft;      /* real length (even)              */
  cplan fwd, inv;
  cpx* st_f;     /* split twiddles forward, n/4     */
  cpx* st_i;     /* split twiddles inverse          */
  cpx* tmp;      /* n/2 scratch                     */
  cpx* bins;     /* n/2+1 adapter-side buffer       */
  float* clean;  /* n adapter-side scrubbed input   */
} rplan;

static int rplan_init(rplan* p, int nfft) {
  if (nfft <= 0 || (nfft & 1)) return -1;
  memset(p, 0, sizeof(*p));
  p->nfft = nfft;
  const int h = nfft / 2;
  if (cplan_init(&p->fwd, h, 0) || cplan_init(&p->inv, h, 1)) return -1;
  p->st_f = (cpx*)malloc(sizeof(cpx) * (size_t)(h / 2 + 1));
  p->st_i = (cpx*)malloc(sizeof(cpx) * (size_t)(h / 2 + 1));
  p->tmp = (cpx*)malloc(sizeof(cpx) * (size_t)h);
  p->bins = (cpx*)malloc(sizeof(cpx) * (size_t)(h + 1));
  p->clean = (float*)malloc(sizeof(float) * (size_t)nfft);
  for (int i = 0; i < h / 2; ++i) { /* kiss_fftr.c:53-59 */
    double ph = -3.14159265358979323846264338327 * ((double)(i + 1) / h + .5);
    p->st_f[i].r = (float)cos(ph);
    p->st_f[i].i = (float)sin(ph);
    ph *= -1;
    p->st_i[i].r = (float)cos(ph);
    p->st_i[i].i = (float)sin(ph);
  }
  return 0;
}
static void rplan_free(rplan* p) {
  cplan_free(&p->fwd); cplan_free(&p->inv);
  free(p->st_f); free(p->st_i); free(p->tmp); free(p->bins); free(p->clean);
}

/* kiss_fftr (kiss_fftr.c:63-117): half-length complex transform of (even,odd) pairs, then
 * the split pass.  Bins 0 and n/2 get exactly zero imaginary parts. */
static void rfwd(rplan* p, const float* x, cpx* X) {
  const int h = p->nfft / 2;
  cwork(p->tmp, (const cpx*)x, 1, p->fwd.fac, &p->fwd);
  const cpx dc = p->tmp[0];
  X[0].r = dc.r + dc.i;
  X[h].r = dc.r - dc.i;
  X[0].i = 0.0f;
  X[h].i = 0.0f;
  for (int k = 1; k <= h / 2; ++k) {
    const cpx a = p->tmp[k];
    cpx b; b.r = p->tmp[h - k].r; b.i = -p->tmp[h - k].i;
    const cpx e = cadd(a, b), o = csub(a, b);
    const cpx t = cmul(o, p->st_f[k - 1]);
    X[k].r = (e.r + t.r) * 0.5f;
    X[k].i = (e.i + t.i) * 0.5f;
    X[h - k].r = (e.r - t.r) * 0.5f;
    X[h - k].i = (t.i - e.i) * 0.5f;
  }
}
/* kiss_fftri (kiss_fftr.c:119-155): only the real parts of bins 0 and n/2 are read. */
static void rinv(rplan* p, const cpx* X, float* x) {
  const int h = p->nfft / 2;
  p->tmp[0].r = X[0].r + X[h].r;
  p->tmp[0].i = X[0].r - X[h].r;
  for (int k = 1; k <= h / 2; ++k) {
    const cpx a = X[k];
    cpx b; b.r = X[h - k].r; b.i = -X[h - k].i;
    const cpx e = cadd(a, b), d = csub(a, b);
    const cpx o = cmul(d, p->st_i[k - 1]);
    p->tmp[k] = cadd(e, o);
    p->tmp[h - k] = csub(e, o);
    p->tmp[h - k].i *= -1;
  }
  cwork((cpx*)x, p->tmp, 1, p->inv.fac, &p->inv);
}

static inline float scrub(float v) { /* kissfft_adapter.cc:104-108,159-163 */
  if (isnan(v) || isinf(v)) return 0.0f;
  if (fabsf(v) < 1e-30f) return 0.0f;
  return v;
}

/* KissFftPlan::forward (kissfft_adapter.cc:83-122): batch b reads in + b*si*n with element
 * stride si, writes out + b*so*(n/2+1) with element stride so. */
static void adapter_fwd(rplan* p, int batch, int si, int so, const float* in, cpx* out) {
  const int n = p->nfft, nb = n / 2 + 1;
  for (int b = 0; b < batch; ++b) {
    const float* src = in + (long)b * si * n;
    cpx* dst = out + (long)b * so * nb;
    for (int i = 0; i < n; ++i) p->clean[i] = scrub(src[(long)i * si]);
    rfwd(p, p->clean, p->bins);
    for (int i = 0; i < nb; ++i) dst[(long)i * so] = p->bins[i];
  }
}
/* KissFftPlan::inverse (kissfft_adapter.cc:124-168): multiply by (1.0f/n), then scrub. */
static void adapter_inv(rplan* p, int batch, int si, int so, const cpx* in, float* out) {
  const int n = p->nfft, nb = n / 2 + 1;
  const float scale = 1.0f / (float)n;
  for (int b = 0; b < batch; ++b) {
    const cpx* src = in + (long)b * si * nb;
    float* dst = out + (long)b * so * n;
    for (int i = 0; i < nb; ++i) p->bins[i] = src[(long)i * si];
    rinv(p, p->bins, p->clean);
    for (int i = 0; i < n; ++i) dst[(long)i * so] = scrub(p->clean[i] * scale);
  }
}

static int plan_args_ok(int nfft, int batch, int si, int so) { /* kissfft_adapter.cc:13-39 */
  return !(batch < 1 || batch > 16 || si < 1 || so < 1 || nfft <= 0 || (nfft & 1));
}

int orc_rfft(int nfft, int batch, int si, int so, const float* in, float* out_ri) {
  rplan p;
  if (!plan_args_ok(nfft, batch, si, so) || rplan_init(&p, nfft)) return -1;
  adapter_fwd(&p, batch, si, so, in, (cpx*)out_ri);
  rplan_free(&p);
  return 0;
}
int orc_irfft(int nfft, int batch, int si, int so, const float* in_ri, float* out) {
  rplan p;
  if (!plan_args_ok(nfft, batch, si, so) || rplan_init(&p, nfft)) return -1;
  adapter_inv(&p, batch, si, so, (const cpx*)in_ri, out);
  rplan_free(&p);
  return 0;
}
/* Complex-domain plan: forward_complex / inverse_complex (kissfft_adapter.cc:171-246): strided
 * gather into a scratch buffer, kiss_fft, strided scatter; the inverse scales by 1/nfft and applies
 * the clean-up rule to both components.  Complex plans accept any nfft > 0 (no evenness check). */
int orc_cfft(int nfft, int inverse, int batch, int si, int so, const float* in_ri, float* out_ri) {
  cplan p;
  if (batch < 1 || batch > 16 || si < 1 || so < 1 || nfft <= 0) return -1;
  if (cplan_init(&p, nfft, inverse)) return -1;
  cpx* buf = (cpx*)malloc(sizeof(cpx) * (size_t)nfft);
  cpx* res = (cpx*)malloc(sizeof(cpx) * (size_t)nfft);
  const cpx* in = (const cpx*)in_ri;
  cpx* out = (cpx*)out_ri;
  const float scale = 1.0f / (float)nfft;
  for (int b = 0; b < batch; ++b) {
    const cpx* src = in + (long)b * si * nfft;
    cpx* dst = out + (long)b * so * nfft;
    for (int i = 0; i < nfft; ++i) buf[i] = src[(long)i * si];
    cwork(res, buf, 1, p.fac, &p);
    for (int i = 0; i < nfft; ++i) {
      cpx v = res[i];
      if (inverse) { v.r = scrub(v.r * scale); v.i = scrub(v.i * scale); }
      dst[(long)i * so] = v;
    }
  }
  free(buf); free(res);
  cplan_free(&p);
  return 0;
}

int orc_rfft_many(int nfft, long n_frames, const float* in, float* out_ri) {
  rplan p;
  if (rplan_init(&p, nfft)) return -1;
  const long nb = nfft / 2 + 1;
  for (long k = 0; k < n_frames; ++k) adapter_fwd(&p, 1, 1, 1, in + k * nfft, (cpx*)out_ri + k * nb);
  rplan_free(&p);
  return 0;
}
int orc_irfft_many(int nfft, long n_frames, const float* in_ri, float* out) {
  rplan p;
  if (rplan_init(&p, nfft)) return -1;
  const long nb = nfft / 2 + 1;
  for (long k = 0; k < n_frames; ++k) adapter_inv(&p, 1, 1, 1, (const cpx*)in_ri + k * nb, out + k * nfft);
  rplan_free(&p);
  return 0;
}

/* ------------------------------------------------------------------------------------ */
/* NormBuilder and the three OLA loops                                                  */
/* ------------------------------------------------------------------------------------ */

long orc_ring_len(long frame, long hop) { /* OLAAccumulator.cc:249-258 */
  return ((frame + hop - 1) / hop + 20) * hop;
}

/* norm_builder.cc:8-52.  Frames k = floor(-N/H) .. floor((ring+N-1+H-1)/H) inclusive are
 * folded onto the ring (so the span covered is ring+~2N: the double count near the seam is
 * reference behaviour, pinned by tests/norm_builder_test.cc).  fp32 adds, k then i ascending. */
void orc_build_norm_linear(float* norm, const float* w, long ring, long N, long H) {
  for (long i = 0; i < ring; ++i) norm[i] = 0.0f;
  if (ring <= 0 || N <= 0 || H <= 0) return;
  const long k_lo = -((N + H - 1) / H);                 /* floor(-N/H) */
  const long k_hi = (ring + N - 1 + H - 1) / H;
  for (long k = k_lo; k <= k_hi; ++k) {
    long s = (k * H) % ring;
    if (s < 0) s += ring;
    /* the reference splits the N window taps into at most two spans; when N > ring-s the
     * second span starts at ring position 0 and is NOT wrapped again. */
    long first = ring - s;
    if (first > N) first = N;
    for (long i = 0; i < first; ++i) norm[s + i] += w[i];
    for (long i = first; i < N; ++i) norm[i - first] += w[i];
  }
}

/* OLAAccumulator::initialize_normalization (OLAAccumulator.cc:260-288). */
void orc_ola_envelope(float* norm, const float* w, long ring, long N, long H, int inside, float eps) {
  if (!w || !inside) {
    for (long i = 0; i < ring; ++i) norm[i] = 1.0f;
  } else if (H == N) {
    for (long i = 0; i < ring; ++i) {
      const float v = w[i % N];
      norm[i] = v > eps ? v : eps;      /* std::max(window, eps) */
    }
  } else {
    orc_build_norm_linear(norm, w, ring, N, H);
  }
}

/* dsp/ola/kernels.cc:18-36 */
void orc_axpy(float* d, const float* s, float g, long n) {
  for (long i = 0; i < n; ++i) d[i] = fmaf(s[i], g, d[i]);
}
void orc_axpy_windowed(float* d, const float* s, const float* w, float g, long n) {
  for (long i = 0; i < n; ++i) d[i] = fmaf(fmaf(s[i], w[i], 0.0f), g, d[i]);
}
void orc_normalize_and_clear(float* out, float* acc, const float* norm, float eps, long n) {
  for (long i = 0; i < n; ++i) {
    const float den = norm[i] > eps ? norm[i] : eps;
    out[i] = acc[i] / den;
    acc[i] = 0.0f;
  }
}

/* Ring emulation of add_frame_SoA + produce (OLAAccumulator.cc:54-122,162-221), including
 * the availability rule (produced_ is absolute, read_pos_ wraps). */
typedef struct {
  long N, H, C, ring;
  float eps;
  int inside, has_window;
  float* w;      /* N        */
  float* norm;   /* ring     */
  float* acc;    /* C x ring */
  long read_pos, produced;
  float peak;
} ola_t;

static void ola_init(ola_t* o, long N, long H, long C, float eps, int inside) {
  memset(o, 0, sizeof(*o));
  o->N = N; o->H = H; o->C = C; o->eps = eps; o->inside = inside;
  o->ring = orc_ring_len(N, H);
  o->w = (float*)calloc((size_t)N, sizeof(float));
  o->norm = (float*)malloc(sizeof(float) * (size_t)o->ring);
  o->acc = (float*)calloc((size_t)(C * o->ring), sizeof(float));
  orc_ola_envelope(o->norm, NULL, o->ring, N, H, inside, eps);
}
static void ola_set_window(ola_t* o, const float* w) {
  memcpy(o->w, w, sizeof(float) * (size_t)o->N);
  o->has_window = 1;
  orc_ola_envelope(o->norm, o->w, o->ring, o->N, o->H, o->inside, o->eps);
}
static void ola_free(ola_t* o) { free(o->w); free(o->norm); free(o->acc); }

static void ola_add(ola_t* o, const float* const* ch, const float* win_arg, long start, long off,
                    long size, float g) {
  if (size == 0 || off >= o->N) return;
  if (off + size > o->N) size = o->N - off;
  const int use = o->inside ? o->has_window : (win_arg != NULL);
  const float* w = use ? (o->inside ? o->w : win_arg) : NULL;
  long len = size > o->ring ? o->ring : size;
  const long s = start % o->ring;
  const long first = (s + len <= o->ring) ? len : o->ring - s;
  for (long c = 0; c < o->C; ++c) {
    float* a = o->acc + c * o->ring;
    const float* f = ch[c] + off;
    if (use) {
      orc_axpy_windowed(a + s, f, w + off, g, first);
      orc_axpy_windowed(a, f + first, w + off + first, g, len - first);
    } else {
      orc_axpy(a + s, f, g, first);
      orc_axpy(a, f + first, g, len - first);
    }
  }
  if (start + size > o->produced) o->produced = start + size;
}
static long ola_produce(ola_t* o, float* const* out, long n) {
  if (n == 0) return 0;
  const long avail = o->produced > o->read_pos ? o->produced - o->read_pos : 0;
  if (avail == 0) return 0;
  if (avail < n) n = avail;
  long len = n > o->ring ? o->ring : n;
  const long s = o->read_pos % o->ring;
  const long first = (s + len <= o->ring) ? len : o->ring - s;
  for (long c = 0; c < o->C; ++c) {
    float* a = o->acc + c * o->ring;
    orc_normalize_and_clear(out[c], a + s, o->norm + s, o->eps, first);
    orc_normalize_and_clear(out[c] + first, a, o->norm, o->eps, len - first);
  }
  o->read_pos = (o->read_pos + n) % o->ring;
  for (long i = 0; i < n; ++i) {
    const float v = fabsf(out[0][i]);
    if (v > o->peak) o->peak = v;
  }
  return n;
}

long orc_ola_stream(const float* frames, long F, long N, long H, long C, const float* window,
                    int set_window_flag, int pass_window_flag, int inside, float gain, float eps,
                    float* out, float* peak_out) {
  ola_t o;
  ola_init(&o, N, H, C, eps, inside);
  if (set_window_flag) ola_set_window(&o, window);
  const long out_len = (F - 1) * H + N;
  const float** chp = (const float**)malloc(sizeof(float*) * (size_t)C);
  float** outp = (float**)malloc(sizeof(float*) * (size_t)C);
  long pos = 0;
  for (long k = 0; k < F; ++k) {
    for (long c = 0; c < C; ++c) chp[c] = frames + (k * C + c) * N;
    ola_add(&o, chp, pass_window_flag ? window : NULL, k * H, 0, N, gain);
    for (long c = 0; c < C; ++c) outp[c] = out + c * out_len + pos;
    pos += ola_produce(&o, outp, H);
  }
  if (N > H && F > 0) {
    for (long c = 0; c < C; ++c) outp[c] = out + c * out_len + pos;
    pos += ola_produce(&o, outp, N - H);
  }
  if (peak_out) *peak_out = o.peak;
  free(chp); free(outp);
  ola_free(&o);
  return pos;
}

/* Gather form: out[t] = ( sum over frames k with kH <= t < kH+N, k ascending, of
 * fma( fl(y_k[t-kH]*w[t-kH]), g, acc ) ) / max(norm[t mod ring], eps).  Bit-identical to
 * the streaming drive (each ring cell is cleared by produce before the next lap reaches it). */
void orc_ola_gather(const float* frames, long first_frame, long n_frames, long total_frames,
                    long N, long H, const float* w, float g, const float* norm, long ring,
                    float eps, long t0, long n_out, float* out) {
  (void)total_frames;
  for (long j = 0; j < n_out; ++j) {
    const long t = t0 + j;
    long k_lo = t - N + 1 <= 0 ? 0 : (t - N + 1 + H - 1) / H;
    long k_hi = t / H;
    if (k_lo < first_frame) k_lo = first_frame;
    if (k_hi > first_frame + n_frames - 1) k_hi = first_frame + n_frames - 1;
    float acc = 0.0f;
    for (long k = k_lo; k <= k_hi; ++k) {
      const long i = t - k * H;
      const float y = frames[(k - first_frame) * N + i];
      acc = w ? fmaf(fmaf(y, w[i], 0.0f), g, acc) : fmaf(y, g, acc);
    }
    const float d = norm[t % ring];
    out[j] = acc / (d > eps ? d : eps);
  }
}

/* The composed round trip (drive order of SURVEY.md 3.2; window multiply of
 * bench/e2e_benchmark.cc:154-156; push_frame_AoS(frame,nullptr,k*H,0,N,gain)). */
long orc_roundtrip(const float* x, long len, int nfft, long hop, const float* window, int mode,
                   float gain, float eps, float* spectra, float* iframes, float* out,
                   long* out_len_p) {
  rplan p;
  if (rplan_init(&p, nfft)) return -1;
  const long F = orc_frame_count(len, nfft, hop, mode);
  const long nb = nfft / 2 + 1;
  ola_t o;
  ola_init(&o, nfft, hop, 1, eps, 1);
  ola_set_window(&o, window);
  float* fr = (float*)malloc(sizeof(float) * (size_t)nfft);
  cpx* sp = (cpx*)malloc(sizeof(cpx) * (size_t)nb);
  long pos = 0;
  for (long k = 0; k < F; ++k) {
    const long start = k * hop;
    long have = len - start;
    if (have > nfft) have = nfft;
    for (long i = 0; i < have; ++i) fr[i] = x[start + i] * window[i];
    for (long i = have; i < nfft; ++i) fr[i] = 0.0f * window[i];
    adapter_fwd(&p, 1, 1, 1, fr, sp);
    if (spectra) memcpy(spectra + k * nb * 2, sp, sizeof(cpx) * (size_t)nb);
    adapter_inv(&p, 1, 1, 1, sp, fr);
    if (iframes) memcpy(iframes + k * nfft, fr, sizeof(float) * (size_t)nfft);
    const float* chp[1] = {fr};
    ola_add(&o, chp, NULL, start, 0, nfft, gain);
    float* outp[1] = {out + pos};
    pos += ola_produce(&o, outp, hop);
  }
  if (F > 0 && nfft > hop) {
    float* outp[1] = {out + pos};
    pos += ola_produce(&o, outp, nfft - hop);
  }
  if (out_len_p) *out_len_p = pos;
  free(fr); free(sp);
  ola_free(&o);
  rplan_free(&p);
  return F;
}

/* ------------------------------------------------------------------------------------ */
/* Synthetic inputs (shared definition with the CUDA generator)                         */
/* ------------------------------------------------------------------------------------ */

static inline float synth_one(uint64_t seed, uint64_t ch, uint64_t i) {
  uint64_t z = seed + ch * 0xD1B54A32D192ED03ull + i * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  const uint32_t m = (uint32_t)(z >> 40);            /* 24 random bits */
  return (float)m * (1.0f / 8388608.0f) - 1.0f;      /* exact, in [-1, 1) */
}
void orc_synth_uniform(unsigned long long seed, unsigned long long ch, unsigned long long i0,
                       long n, float* out) {
  for (long j = 0; j < n; ++j) out[j] = synth_one(seed, ch, i0 + (uint64_t)j);
}
void orc_synth_tones(unsigned long long i0, long n, float* out) {
  for (long j = 0; j < n; ++j) {
    const double t = (double)(i0 + (uint64_t)j) / 48000.0;
    out[j] = (float)(0.5f * sin(2.0 * M_PI * 440.0 * t) + 0.3f * sin(2.0 * M_PI * 880.0 * t) +
                     0.2f * sin(2.0 * M_PI * 1320.0 * t));
  }
}

/* ------------------------------------------------------------------------------------ */
/* Timed CPU baseline (kind = "port")                                                   */
/* ------------------------------------------------------------------------------------ */

typedef struct {
  int id, nfft, wtype;
  long len, hop;
  uint64_t seed;
  pthread_barrier_t* bar;
  double t0, t1, sum;
  int err;
} job_t;

static double now_s(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static void* job_main(void* arg) {
  job_t* j = (job_t*)arg;
  float* x = (float*)malloc(sizeof(float) * (size_t)j->len);
  float* w = (float*)malloc(sizeof(float) * (size_t)j->nfft);
  float* y = (float*)malloc(sizeof(float) * (size_t)(j->len + j->nfft));
  j->err = !(x && w && y);
  if (!j->err) {
    orc_synth_uniform(j->seed, (uint64_t)j->id, 0, j->len, x);
    orc_window(j->wtype, j->nfft, 0, ORC_NORM_NONE, w);
  }
  pthread_barrier_wait(j->bar);
  j->t0 = now_s();
  long n_out = 0;
  if (!j->err) orc_roundtrip(x, j->len, j->nfft, j->hop, w, ORC_DROP, 1.0f, 1e-8f, NULL, NULL, y, &n_out);
  j->t1 = now_s();
  pthread_barrier_wait(j->bar);
  double s = 0.0;
  for (long i = 0; i < n_out; ++i) s += (double)y[i];
  j->sum = s;
  free(x); free(w); free(y);
  return NULL;
}

double orc_bench_roundtrip_mt(int threads, long len, int nfft, long hop, int wtype,
                              unsigned long long seed, double* checksum_out) {
  if (threads < 1 || threads > 1024) return -1.0;
  pthread_barrier_t bar;
  pthread_barrier_init(&bar, NULL, (unsigned)threads);
  job_t* jobs = (job_t*)calloc((size_t)threads, sizeof(job_t));
  pthread_t* th = (pthread_t*)calloc((size_t)threads, sizeof(pthread_t));
  for (int i = 0; i < threads; ++i) {
    jobs[i].id = i; jobs[i].nfft = nfft; jobs[i].wtype = wtype; jobs[i].len = len; jobs[i].hop = hop;
    jobs[i].seed = seed; jobs[i].bar = &bar;
    pthread_create(&th[i], NULL, job_main, &jobs[i]);
  }
  double t0 = 1e300, t1 = 0.0, sum = 0.0;
  int err = 0;
  for (int i = 0; i < threads; ++i) {
    pthread_join(th[i], NULL);
    if (jobs[i].t0 < t0) t0 = jobs[i].t0;
    if (jobs[i].t1 > t1) t1 = jobs[i].t1;
    sum += jobs[i].sum;
    err |= jobs[i].err;
  }
  pthread_barrier_destroy(&bar);
  free(jobs); free(th);
  if (checksum_out) *checksum_out = sum;
  return err ? -1.0 : t1 - t0;
}
