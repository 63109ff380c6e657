// facade_test.cc -- exercises the reference-named C++ classes of crlot_dsp.h on a B200, in the
// shape of the reference's own gtest cases (gtest is not in this image, so plain CHECKs):
//   tests/framer_test.cc:100-322, tests/window_lut_test.cc:85-215, tests/fft_test.cc:57-250,
//   tests/ola_accumulator_test.cc:143-330, tests/frame_queue_test.cc:30-420,
//   tests/ring_buffer_test.cc:14-234, tests/norm_builder_test.cc:13-167, tests/kernels_test.cc:67-208 (on
//   device pointers), and the loop of bench/e2e_benchmark.cc:138-186.
// It does not link the oracle.  Arrays it produces are written to <outdir>/*.f32 so that
// tests/test_gpu_facades_cpp.py can compare them with the oracle on the same inputs.
//
//   usage: facade_test <outdir>
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <thread>
#include <tuple>
#include <type_traits>
#include <utility>
#include <vector>

#include "crlot_dsp.h"
#include "crlot_wav.h"

using namespace dsp;
using namespace dsp::fft;

static int g_fail = 0, g_checks = 0;
#define CHECK(cond)                                                               \
  do {                                                                            \
    ++g_checks;                                                                   \
    if (!(cond)) { ++g_fail; std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond); } \
  } while (0)
#define CHECK_THROWS(expr, extype)                          \
  do {                                                      \
    ++g_checks;                                             \
    bool ok__ = false;                                      \
    try { expr; } catch (const extype&) { ok__ = true; } catch (...) {} \
    if (!ok__) { ++g_fail; std::printf("FAIL %s:%d  %s should throw %s\n", __FILE__, __LINE__, #expr, #extype); } \
  } while (0)

#define CHECK_NOTHROW(expr)                                 \
  do {                                                      \
    ++g_checks;                                             \
    try { expr; } catch (...) { ++g_fail; std::printf("FAIL %s:%d  %s threw\n", __FILE__, __LINE__, #expr); } \
  } while (0)

static std::string g_out;
static void dump(const char* name, const float* p, size_t n) {
  const std::string path = g_out + "/" + name + ".f32";
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) { std::printf("cannot write %s\n", path.c_str()); ++g_fail; return; }
  std::fwrite(p, sizeof(float), n, f);
  std::fclose(f);
}

// deterministic test signal (an LCG; the Python side regenerates it from the dumped copy)
static std::vector<float> noise(size_t n, uint32_t seed) {
  std::vector<float> v(n);
  uint32_t s = seed;
  for (size_t i = 0; i < n; ++i) {
    s = s * 1664525u + 1013904223u;
    v[i] = static_cast<float>(static_cast<int32_t>(s >> 8) - (1 << 23)) / static_cast<float>(1 << 23);
  }
  return v;
}

// ---- Framer (tests/framer_test.cc) --------------------------------------------------------------
static void test_framer() {
  std::vector<float> sine(4800);
  for (size_t i = 0; i < sine.size(); ++i) sine[i] = std::sin(2.0 * M_PI * 440.0 * i / 48000.0);
  {  // BasicPushPop :100-112
    Framer f;
    std::vector<float> fr(64);
    CHECK(!f.push(sine.data(), 10));                 // params not set -> false (framer.cc:38-40)
    CHECK(!f.pop(fr.data()));
    f.set_params(64, 16, 1);
    CHECK(f.push(sine.data(), sine.size()));
    CHECK(f.available_frames() > 0);
    CHECK(f.pop(fr.data()));
    bool same = true;
    for (int i = 0; i < 64; ++i) same = same && fr[i] == sine[i];
    CHECK(same);
    CHECK(f.pop(fr.data()));
    same = true;
    for (int i = 0; i < 64; ++i) same = same && fr[i] == sine[16 + i];
    CHECK(same);
  }
  {  // InterleavedChannels :114-134 -- stereo frames stay interleaved [L0,R0,L1,R1,...]
    std::vector<float> st(200);
    for (int i = 0; i < 100; ++i) { st[2 * i] = (float)i; st[2 * i + 1] = -(float)i; }
    Framer f;
    f.set_params(32, 8, 2);
    CHECK(f.push(st.data(), 100));
    std::vector<float> fr(64);
    CHECK(f.pop(fr.data()));
    CHECK(f.pop(fr.data()));
    bool ok = true;
    for (int i = 0; i < 32; ++i) ok = ok && fr[2 * i] == (float)(8 + i) && fr[2 * i + 1] == -(float)(8 + i);
    CHECK(ok);
    CHECK(f.channels() == 2);
  }
  {  // EvenOddLengths :227-241, VeryShortLengths :244-259
    Framer f;
    f.set_params(64, 16, 1);
    CHECK(f.push(sine.data(), 100));
    CHECK(f.available_frames() > 0);
    f.set_params(63, 15, 1);
    f.reset();
    CHECK(f.push(sine.data(), 100));
    CHECK(f.available_frames() == (100 - 63) / 15 + 1);
    f.set_params(32, 8, 1);
    CHECK(f.push(nullptr, 0));
    CHECK(f.available_frames() == 0);
    float one = 1.0f;
    CHECK(f.push(&one, 1));
    CHECK(f.available_frames() == 1);                // ZERO_PAD: a partial frame is available
    std::vector<float> sh(7, 0.5f);
    CHECK(f.push(sh.data(), sh.size()));
    CHECK(!f.push(nullptr, 3));
  }
  {  // BoundaryConditions :262-296
    std::vector<float> sh(20, 1.0f), fr(32, 7.0f);
    Framer pad, drop;
    pad.set_params(32, 16, 1, BoundaryMode::ZERO_PAD);
    drop.set_params(32, 16, 1, BoundaryMode::DROP);
    CHECK(pad.push(sh.data(), 20));
    CHECK(drop.push(sh.data(), 20));
    CHECK(pad.pop(fr.data()));
    bool ok = true;
    for (int i = 0; i < 32; ++i) ok = ok && fr[i] == (i < 20 ? 1.0f : 0.0f);
    CHECK(ok);
    CHECK(pad.available_frames() == 1);              // 4 samples left -> one more padded frame
    CHECK(pad.pop(fr.data()));
    CHECK(fr[3] == 1.0f && fr[4] == 0.0f);
    CHECK(pad.available_frames() == 0);              // cursor clamped at the write position
    CHECK(!pad.pop(fr.data()));
    CHECK(!drop.pop(fr.data()));
    CHECK(drop.available_frames() == 0);
    CHECK_THROWS(pad.set_params(0, 16), std::invalid_argument);
    CHECK_THROWS(pad.set_params(32, 0), std::invalid_argument);
    CHECK_THROWS(pad.set_params(32, 16, 0), std::invalid_argument);
  }
  {  // FramingCompatibilityFormula :299-322 and MemoryEfficiency-style long streaming
    Framer f;
    f.set_params(64, 16, 1);
    std::vector<float> ones(200, 1.0f);
    CHECK(f.push(ones.data(), 200));
    CHECK(f.available_frames() == (200 - 64) / 16 + 1);
    CHECK(f.buffer_size() == 256);                   // max(required 200, 2 * initial 128), framer.cc:120-126
  }
  {  // chunked streaming across compaction/growth: every frame equals the slice of the input
    const size_t N = 256, H = 64, total = 40000;
    std::vector<float> x = noise(total, 7);
    Framer f;
    f.set_params(N, H, 1, BoundaryMode::DROP);
    std::vector<float> fr(N);
    size_t pushed = 0, k = 0;
    bool ok = true;
    const size_t chunks[] = {1, 63, 64, 65, 300, 1024, 17, 4096};
    for (size_t it = 0; pushed < total; ++it) {
      const size_t c = std::min(chunks[it % 8], total - pushed);
      f.push(x.data() + pushed, c);
      pushed += c;
      while (f.pop(fr.data())) {
        for (size_t i = 0; i < N; ++i) ok = ok && fr[i] == x[k * H + i];
        ++k;
      }
    }
    CHECK(ok);
    CHECK(k == (total - N) / H + 1);
  }
}

// ---- WindowLUT (tests/window_lut_test.cc) --------------------------------------------------------
static void test_window() {
  WindowLUT hann(1024, WindowType::HANN);
  CHECK(hann.size() == 1024 && hann.type() == WindowType::HANN && !hann.periodic());
  CHECK(hann.data() != nullptr);
  CHECK(std::fabs(hann.data()[0]) < 1e-7f && std::fabs(hann.data()[1023]) < 1e-6f);
  CHECK(std::fabs(hann.data()[512] - 0.5f * (1.0f - std::cos(2.0 * M_PI * 512 / 1023.0))) < 1e-6f);
  WindowLUT ham(512, WindowType::HAMMING, true);
  CHECK(std::fabs(ham.data()[0] - 0.08f) < 1e-6f);
  WindowLUT bl(512, WindowType::BLACKMAN);
  CHECK(std::fabs(bl.data()[0]) < 1e-6f);
  WindowLUT rect(64, WindowType::RECT);
  CHECK(rect.data()[0] == 1.0f && rect.data()[63] == 1.0f);
  WindowLUT s1(256, WindowType::HANN, false, NormalizationType::SUM_TO_ONE);
  CHECK(std::fabs(WindowLUT::calculateSum(s1.data(), 256) - 1.0) < 1e-5);
  WindowLUT l2(256, WindowType::HANN, false, NormalizationType::L2_NORM);
  CHECK(std::fabs(WindowLUT::calculateSumOfSquares(l2.data(), 256) - 1.0) < 1e-5);
  CHECK_THROWS(WindowLUT(0, WindowType::HANN), std::invalid_argument);
  CHECK_THROWS(WindowLUT(16, WindowType::BLACKMAN_HARRIS), std::invalid_argument);
  auto sp = WindowLUT::getInstance().GetWindowSafe(WindowType::HANN, 2048, true);
  CHECK(sp && sp.get()[1024] == 1.0f);
  dump("win_hann_2048_periodic", sp.get(), 2048);
  dump("win_blackman_512", bl.data(), 512);
}

// ---- IFftPlan (tests/fft_test.cc) -----------------------------------------------------------------
static void test_fft() {
  for (int n : {512, 1024, 2048}) {
    auto plan = MakeFftPlan({FftDomain::Real, n, false, 1, 1, 1});
    CHECK(plan->size() == n && plan->domain() == FftDomain::Real && plan->max_batch_size() == 16);
    std::vector<float> x = noise(n, 100 + n), y(n);
    std::vector<std::complex<float>> X(n / 2 + 1);
    plan->forward(x.data(), X.data());
    plan->inverse(X.data(), y.data());
    double e = 0;
    for (int i = 0; i < n; ++i) e += (double)(y[i] - x[i]) * (y[i] - x[i]);
    CHECK(std::sqrt(e / n) < 1e-5);                                  // :57-129
    CHECK(X[0].imag() == 0.0f && X[n / 2].imag() == 0.0f);           // :344-382
    std::vector<float> tone(n);
    for (int i = 0; i < n; ++i) tone[i] = 2.0f * std::cos(2.0 * M_PI * 10 * i / n);
    plan->forward(tone.data(), X.data());
    CHECK(std::fabs(std::abs(X[10]) - n) < 1e-2 * n / 512);          // :131-197
    if (n == 1024) {
      dump("fft_x_1024", x.data(), n);
      plan->forward(x.data(), X.data());
      dump("fft_X_1024", reinterpret_cast<float*>(X.data()), 2 * (n / 2 + 1));
    }
  }
  {  // NaNAndDenormalHandling :199-221
    auto plan = MakeFftPlan({FftDomain::Real, 512, false, 1, 1, 1});
    std::vector<float> x(512, 0.25f);
    x[3] = NAN; x[9] = INFINITY; x[11] = 1e-38f;
    std::vector<std::complex<float>> X(257);
    plan->forward(x.data(), X.data());
    bool fin = true;
    for (auto& v : X) fin = fin && std::isfinite(v.real()) && std::isfinite(v.imag());
    CHECK(fin);
  }
  // InvalidConfiguration :223-249 (std::runtime_error from the adapter)
  CHECK_THROWS(MakeFftPlan({FftDomain::Real, 511, false, 1, 1, 1}), std::runtime_error);
  CHECK_THROWS(MakeFftPlan({FftDomain::Real, 512, true, 1, 1, 1}), std::runtime_error);
  CHECK_THROWS(MakeFftPlan({FftDomain::Real, 512, false, 0, 1, 1}), std::runtime_error);
  CHECK_THROWS(MakeFftPlan({FftDomain::Real, 512, false, 17, 1, 1}), std::runtime_error);
  CHECK_THROWS(MakeFftPlan({FftDomain::Real, 512, false, 1, 0, 1}), std::runtime_error);
  {  // BatchProcessingReal :384-448
    auto plan = MakeFftPlan({FftDomain::Real, 512, false, 4, 1, 1});
    std::vector<float> x = noise(4 * 512, 5), y(4 * 512);
    std::vector<std::complex<float>> X(4 * 257), X1(257);
    plan->forward(x.data(), X.data(), 4);
    auto one = MakeFftPlan({FftDomain::Real, 512, false, 1, 1, 1});
    bool same = true;
    for (int b = 0; b < 4; ++b) {
      one->forward(x.data() + b * 512, X1.data());
      for (int k = 0; k < 257; ++k) same = same && X1[k] == X[b * 257 + k];
    }
    CHECK(same);
    CHECK_THROWS(plan->forward(x.data(), X.data(), 5), std::runtime_error);
  }
  {  // StrideLayoutVerification :450-520 -- nfft 128 (an any-size plan here), batch 3, strides 2 / 2
    const int n = 128, nb = 65, B = 3;
    auto plan = MakeFftPlan({FftDomain::Real, n, false, B, 2, 2});
    std::vector<float> x(static_cast<size_t>(B) * n * 2, 0.0f);
    for (int b = 0; b < B; ++b)
      for (int i = 0; i < n; ++i) x[(static_cast<size_t>(b) * n + i) * 2] = std::sin(2.0f * (float)M_PI * (b + 1) * i / n);
    std::vector<std::complex<float>> X(static_cast<size_t>(B) * nb * 2);
    plan->forward(x.data(), X.data(), B);
    for (int b = 0; b < B; ++b) {
      int peak = 0;
      float best = 0.0f;
      for (int k = 0; k < nb; ++k) {
        const float m = std::abs(X[(static_cast<size_t>(b) * nb + k) * 2]);
        if (m > best) { best = m; peak = k; }
      }
      CHECK(peak == b + 1);
    }
  }
  {  // InvalidConfiguration :229-237 -- complex plans of even and odd size are accepted; mixed-radix sizes work
    CHECK_NOTHROW(MakeFftPlan({FftDomain::Complex, 512, false, 1, 1, 1}));
    CHECK_NOTHROW(MakeFftPlan({FftDomain::Complex, 513, false, 1, 1, 1}));
    for (int n : {513, 480, 1000}) {
      auto plan = MakeFftPlan({FftDomain::Complex, n, false, 1, 1, 1});
      std::vector<std::complex<float>> x(n), X(n), y(n);
      std::vector<float> r = noise(2 * n, 900 + n);
      for (int i = 0; i < n; ++i) x[i] = {r[2 * i], r[2 * i + 1]};
      plan->forward_complex(x.data(), X.data());
      plan->inverse_complex(X.data(), y.data());
      double e = 0;
      for (int i = 0; i < n; ++i) e += std::norm(y[i] - x[i]);
      CHECK(std::sqrt(e / n) < 1e-5);
    }
    auto rp = MakeFftPlan({FftDomain::Real, 480, false, 1, 1, 1});            // 30 ms at 16 kHz
    std::vector<float> x = noise(480, 77), y(480);
    std::vector<std::complex<float>> X(241);
    rp->forward(x.data(), X.data());
    rp->inverse(X.data(), y.data());
    double e = 0;
    for (int i = 0; i < 480; ++i) e += (double)(y[i] - x[i]) * (y[i] - x[i]);
    CHECK(std::sqrt(e / 480) < 1e-5);
    CHECK(X[0].imag() == 0.0f && X[240].imag() == 0.0f);
  }
}

// ---- OLAAccumulator (tests/ola_accumulator_test.cc) -----------------------------------------------
static void test_ola() {
  OLAConfig cfg{48000, 256, 64, 1, 1e-8f, true};
  {
    OLAAccumulator ola(cfg);
    CHECK(ola.config().frame_size == 256 && ola.ring_size() == (256 / 64 + 20) * 64);     // :143-153
    CHECK(!ola.has_window() && ola.produced_samples() == 0 && ola.read_pos() == 0);
    WindowLUT w(256, WindowType::HANN);
    ola.set_window(w.data(), 256);
    CHECK(ola.has_window());
    CHECK_THROWS(ola.set_window(w.data(), 128), std::invalid_argument);                    // :178-192
    CHECK_THROWS(ola.set_window(nullptr, 256), std::invalid_argument);
    std::vector<float> out(64);
    float* outs[1] = {out.data()};
    CHECK(ola.produce(outs, 0) == 0);                                                      // :279-292
    CHECK(ola.produce(outs, 64) == 0);                                                     // nothing pushed yet
  }
  {
    OLAConfig bad = cfg;
    bad.frame_size = 0;
    CHECK_THROWS(OLAAccumulator o(bad), std::invalid_argument);                            // :155-176
    bad = cfg; bad.hop_size = 0;
    CHECK_THROWS(OLAAccumulator o(bad), std::invalid_argument);
    bad = cfg; bad.channels = 0;
    CHECK_THROWS(OLAAccumulator o(bad), std::invalid_argument);
    bad = cfg; bad.eps = 0.0f;
    CHECK_THROWS(OLAAccumulator o(bad), std::invalid_argument);
  }
  {  // AoS == SoA :392-437, two channels, streaming drive
    OLAConfig c2{48000, 512, 128, 2, 1e-8f, true};
    OLAAccumulator a(c2), b(c2);
    WindowLUT w(512, WindowType::HANN);
    a.set_window(w.data(), 512);
    b.set_window(w.data(), 512);
    const int F = 20;                                             // stays below ring_size: read_pos does not wrap
    std::vector<float> fr = noise((size_t)F * 2 * 512, 11);       // [frame][channel][sample]
    std::vector<float> inter(1024), oa0(128), oa1(128), ob0(128), ob1(128), all_a, all_b;
    bool same = true;
    for (int k = 0; k < F; ++k) {
      const float* p0 = fr.data() + (size_t)k * 1024;
      const float* p1 = p0 + 512;
      const float* chs[2] = {p0, p1};
      for (int i = 0; i < 512; ++i) { inter[2 * i] = p0[i]; inter[2 * i + 1] = p1[i]; }
      a.add_frame_SoA(chs, nullptr, (size_t)k * 128, 0, 512, 0.75f);
      b.push_frame_AoS(inter.data(), nullptr, (size_t)k * 128, 0, 512, 0.75f);
      float* oa[2] = {oa0.data(), oa1.data()};
      float* ob[2] = {ob0.data(), ob1.data()};
      const size_t na = a.produce(oa, 128), nb = b.produce(ob, 128);
      same = same && na == nb && na == 128;
      for (size_t i = 0; i < na; ++i) same = same && oa0[i] == ob0[i] && oa1[i] == ob1[i];
      all_a.insert(all_a.end(), oa0.begin(), oa0.begin() + na);
      all_b.insert(all_b.end(), oa1.begin(), oa1.begin() + na);
    }
    CHECK(same);
    CHECK(a.produced_samples() == (size_t)(F - 1) * 128 + 512 && a.read_pos() == (size_t)F * 128);   // OLAAccumulator.cc:114,215
    dump("ola_frames", fr.data(), fr.size());
    dump("ola_out_ch0", all_a.data(), all_a.size());
    dump("ola_out_ch1", all_b.data(), all_b.size());
    a.reset();
    CHECK(a.produced_samples() == 0 && a.read_pos() == 0);                                  // :307-332
  }
}

// ---- the reference's FullPipeline loop, verbatim shape (bench/e2e_benchmark.cc:138-186) ------------
static void test_e2e_loop() {
  const size_t N = 1024, H = 256, L = 48000;
  std::vector<float> in = noise(L, 2024);
  WindowLUT lut(N, WindowType::HANN);
  const float* window = lut.data();
  Framer framer;
  framer.set_params(N, H, 1, BoundaryMode::DROP);
  OLAConfig cfg{48000, N, H, 1, 1e-8f, true};
  OLAAccumulator ola(cfg);
  ola.set_window(window, (int)N);
  auto plan = MakeFftPlan({FftDomain::Real, (int)N, false, 1, 1, 1});

  framer.push(in.data(), L);
  std::vector<float> frame(N), proc(N), out_loop;
  std::vector<std::complex<float>> spec(N / 2 + 1);
  std::vector<float> hopbuf(H);
  size_t k = 0;
  while (framer.pop(frame.data())) {
    for (size_t i = 0; i < N; ++i) proc[i] = frame[i] * window[i];
    plan->forward(proc.data(), spec.data());
    plan->inverse(spec.data(), proc.data());
    ola.push_frame_AoS(proc.data(), nullptr, k * H, 0, N, 1.0f);
    float* ch[1] = {hopbuf.data()};
    const size_t n = ola.produce(ch, H);
    out_loop.insert(out_loop.end(), hopbuf.begin(), hopbuf.begin() + n);
    ++k;
  }
  CHECK(k == (L - N) / H + 1);
  CHECK(out_loop.size() == k * H);

  // the same signal through the whole-signal call
  StftRoundTrip rt(N, H, window, window, BoundaryMode::DROP);
  CHECK(rt.frames_for(L) == k);
  std::vector<float> out_batch(rt.output_len_for(L));
  const size_t n_out = rt.process(in.data(), L, L, 1, out_batch.data(), out_batch.size());
  CHECK(n_out == (k - 1) * H + N);
  size_t diff = 0;
  double num = 0, den = 0;
  for (size_t i = 0; i < out_loop.size(); ++i) {
    if (out_loop[i] != out_batch[i]) ++diff;
    num += (double)(out_loop[i] - out_batch[i]) * (out_loop[i] - out_batch[i]);
    den += (double)out_batch[i] * out_batch[i];
  }
  std::printf("e2e loop vs whole-signal call: %zu of %zu samples differ in bits, rel-L2 %.3g\n", diff, out_loop.size(),
              std::sqrt(num / den));
  CHECK(std::sqrt(num / den) < 1e-6);
  dump("e2e_in", in.data(), in.size());
  dump("e2e_out_loop", out_loop.data(), out_loop.size());
  dump("e2e_out_batch", out_batch.data(), n_out);
}

// ---- Framer -> kernel 1 without a host hop (B200 extension) -----------------------------------------
static void test_framer_stft() {
  const size_t N = 512, H = 128, C = 2, L = 10000;
  std::vector<float> inter = noise(L * C, 99);
  WindowLUT lut(N, WindowType::HAMMING);
  Framer fr;
  fr.set_params(N, H, C, BoundaryMode::DROP);
  StftRoundTrip rt(N, H, lut.data(), lut.data(), BoundaryMode::DROP);
  fr.push(inter.data(), L / 2);
  fr.push(inter.data() + (L / 2) * C, L - L / 2);
  const size_t F = fr.drainable_frames();
  CHECK(F == (L - N) / H + 1);
  const size_t bins = N / 2 + 1;
  crlot_buf_t spec = nullptr;
  detail::check(crlot_buf_create(C * F * bins * 2 * sizeof(float), 0, &spec));
  long long got = 0;
  detail::check(crlot_framer_stft(fr.handle(), rt.plan(), -1, crlot_buf_device_ptr(spec), &got, nullptr));
  CHECK((size_t)got == F);
  CHECK(fr.available_frames() == 0);
  std::vector<float> S(C * F * bins * 2);
  detail::check(crlot_buf_download(spec, 0, S.data(), S.size() * sizeof(float), nullptr));
  detail::check(crlot_cuda_stream_sync(0, nullptr));
  crlot_buf_destroy(spec);
  dump("fstft_in_interleaved", inter.data(), inter.size());
  dump("fstft_spec", S.data(), S.size());
  // channel 1, frame 3 against the per-frame plan on the de-interleaved, windowed slice
  std::vector<float> x(N);
  for (size_t i = 0; i < N; ++i) x[i] = inter[(3 * H + i) * C + 1] * lut.data()[i];
  auto plan = MakeFftPlan({FftDomain::Real, (int)N, false, 1, 1, 1});
  std::vector<std::complex<float>> X(bins);
  plan->forward(x.data(), X.data());
  bool same = true;
  for (size_t k = 0; k < bins; ++k) {
    const float* s = &S[((1 * F + 3) * bins + k) * 2];
    same = same && s[0] == X[k].real() && s[1] == X[k].imag();
  }
  CHECK(same);
}

// ---- FrameQueue (tests/frame_queue_test.cc) -----------------------------------------------------------
static void test_framequeue() {
  std::vector<float> x(1000);
  for (size_t i = 0; i < x.size(); ++i) x[i] = std::sin(2.0 * M_PI * 440.0 * i / 48000.0);
  {  // BasicConstruction :30-36, hop sizes :69-108
    FrameQueue fq(x.data(), x.size(), 32, 16);
    CHECK(fq.getFrameSize() == 32);
    CHECK(fq.getHopSize() == 16);
    CHECK(fq.getNumFrames() == (1000 + 32 - 16) / 16);
    CHECK(fq.getFrame(1) == fq.getFrame(0) + 32);          // AoS continuity :255-260
    CHECK(fq.getAllFrames().size() == fq.getNumFrames() * 32);
    for (size_t h : {8u, 16u, 24u}) {
      FrameQueue q(x.data(), x.size(), 32, h);
      CHECK(q.getHopSize() == h && q.getNumFrames() > 0);
    }
  }
  {  // odd frame size :39-51, center = false
    FrameQueue fq(x.data(), x.size(), 31, 15);
    CHECK(fq.getNumFrames() == (1000 + 30 - 16) / 15);
    FrameQueue nc(x.data(), x.size(), 32, 16, false);
    CHECK(nc.getNumFrames() == (1000 - 16) / 16);
    bool same = true;
    for (size_t i = 0; i < 32; ++i) same = same && nc.getFrame(3)[i] == x[3 * 16 + i];
    CHECK(same);
  }
  {  // REFLECTIndexMapping :321-364: [1,2,3,4], frame 8, hop 4 -> 4 3 2 1 | 1 2 3 4 ; 1 2 3 4 | 3 2 1 1
    const float pat[4] = {1.f, 2.f, 3.f, 4.f};
    FrameQueue fq(pat, 4, 8, 4, true, PadMode::REFLECT);
    CHECK(fq.getNumFrames() == 2);
    const float want0[8] = {4, 3, 2, 1, 1, 2, 3, 4}, want1[8] = {1, 2, 3, 4, 3, 2, 1, 1};
    bool ok = true;
    for (int i = 0; i < 8; ++i) ok = ok && fq.getFrame(0)[i] == want0[i] && fq.getFrame(1)[i] == want1[i];
    CHECK(ok);
    FrameQueue ed(pat, 4, 8, 4, true, PadMode::EDGE);
    const float e0[8] = {1, 1, 1, 1, 1, 2, 3, 4};
    ok = true;
    for (int i = 0; i < 8; ++i) ok = ok && ed.getFrame(0)[i] == e0[i];
    CHECK(ok);
  }
  {  // extreme cases :264-318,367-: empty, single sample, shorter than a frame
    FrameQueue e(nullptr, 0, 32, 16, true, PadMode::REFLECT);
    CHECK(e.getNumFrames() == 1);
    bool zeros = true;
    for (size_t i = 0; i < 32; ++i) zeros = zeros && e.getFrame(0)[i] == 0.0f;
    CHECK(zeros);
    const float one[1] = {1.0f};
    FrameQueue s1(one, 1, 32, 16, true, PadMode::REFLECT);
    CHECK(s1.getNumFrames() > 0);
    bool found = false;
    for (size_t i = 0; i < 32; ++i) found = found || s1.getFrame(0)[i] == 1.0f;
    CHECK(found);
    const float three[3] = {1.f, 2.f, 3.f};
    FrameQueue s3(three, 3, 32, 16, true, PadMode::REFLECT);
    CHECK(s3.getNumFrames() > 0);
    CHECK_THROWS(FrameQueue(three, 3, 0, 16), std::invalid_argument);
    CHECK_THROWS(FrameQueue(three, 3, 32, 0), std::invalid_argument);
    CHECK_THROWS(FrameQueue(nullptr, 3, 32, 16), std::invalid_argument);
    CHECK_THROWS(s3.getFrame(s3.getNumFrames()), std::out_of_range);
    CHECK_THROWS(s3.copyFrame(0, nullptr), std::invalid_argument);
  }
  {  // dumped for the oracle comparison on the Python side
    std::vector<float> v = noise(777, 99u);
    dump("fq_in", v.data(), v.size());
    FrameQueue a(v.data(), v.size(), 64, 16, true, PadMode::REFLECT);
    dump("fq_reflect_64_16", a.getAllFrames().data(), a.getAllFrames().size());
    FrameQueue b(v.data(), v.size(), 64, 48, true, PadMode::EDGE);
    dump("fq_edge_64_48", b.getAllFrames().data(), b.getAllFrames().size());
    FrameQueue c(v.data(), v.size(), 50, 50, false, PadMode::CONSTANT);
    dump("fq_const_50_50_nc", c.getAllFrames().data(), c.getAllFrames().size());
  }
}

// ---- WAV ingest -> interleaved round trip (main/main.cc:120-160 reads a stereo file) -----
// The Python side writes <outdir>/stereo16.wav before starting the driver; skipped when absent.
static void test_wav_interleaved() {
  WavReader rd;
  if (!rd.open(g_out + "/stereo16.wav")) { std::printf("stereo16.wav absent: WAV test skipped\n"); return; }
  CHECK(rd.is_open() && rd.get_channels() == 2 && rd.get_bits_per_sample() == 16 && rd.get_sample_rate() == 44100);
  const size_t frames = static_cast<size_t>(rd.get_total_frames());
  std::vector<float> x = rd.read_all();                     // interleaved [n][2]
  CHECK(x.size() == frames * 2);
  size_t got = 123;
  CHECK(!rd.read(x.data(), 16, &got) && got == 0);          // at the end: false, nothing read (io/wav.cc:86)
  dump("wav_interleaved", x.data(), x.size());
  WindowLUT lut(1024, WindowType::HANN);
  StftRoundTrip rt(1024, 256, lut.data(), lut.data());
  const size_t n_out = rt.output_len_for(frames);
  std::vector<float> y(n_out * 2, -1.0f);
  long long n = 0;
  detail::check(crlot_roundtrip_host_interleaved(rt.plan(), x.data(), (long long)frames, 2, y.data(), &n));
  CHECK(static_cast<size_t>(n) == n_out);
  dump("wav_roundtrip_interleaved", y.data(), y.size());
  // the same through the planar entry point, channel by channel, must agree bit for bit
  std::vector<float> pl(frames * 2), yp(n_out * 2);
  for (size_t i = 0; i < frames; ++i) { pl[i] = x[2 * i]; pl[frames + i] = x[2 * i + 1]; }
  rt.process(pl.data(), frames, frames, 2, yp.data(), n_out);
  bool same = true;
  for (size_t i = 0; i < n_out; ++i) same = same && y[2 * i] == yp[i] && y[2 * i + 1] == yp[n_out + i];
  CHECK(same);
  WavReader bad;
  CHECK(!bad.open(g_out + "/does_not_exist.wav") && !bad.is_open() && bad.get_channels() == 0);
}


// ---- RingBuffer (tests/ring_buffer_test.cc:14-234; bodies as there, gtest macros replaced) --------
#define EXPECT_EQ(a, b) CHECK((a) == (b))
#define EXPECT_TRUE(a) CHECK(a)
#define EXPECT_THROW(expr, extype) CHECK_THROWS(expr, extype)
static void test_ring_buffer() {
  using dsp::ring::RingBuffer;
  {  // Constructor :14-18, ConstructorWithShadow :20-23
    RingBuffer<float> rb(1024);
    EXPECT_EQ(rb.capacity(), 1024u);
    EXPECT_EQ(rb.write_pos(), 0u);
    RingBuffer<float> rs(512, true);
    EXPECT_EQ(rs.capacity(), 512u);
    EXPECT_EQ(rs.physical_capacity(), 1024u);
    EXPECT_TRUE(rs.has_shadow());
  }
  {  // SplitNoWrapAround :25-43
    RingBuffer<int> rb(10);
    for (size_t i = 0; i < 10; ++i) rb.data()[i] = static_cast<int>(i);
    auto [span1, span2] = rb.split(2, 5);
    EXPECT_EQ(span1.size(), 5u);
    EXPECT_TRUE(span2.empty());
    for (size_t i = 0; i < 5; ++i) EXPECT_EQ(span1[i], static_cast<int>(i + 2));
  }
  {  // SplitWithWrapAround :45-68
    RingBuffer<int> rb(10);
    for (size_t i = 0; i < 10; ++i) rb.data()[i] = static_cast<int>(i);
    auto [span1, span2] = rb.split(7, 5);
    EXPECT_EQ(span1.size(), 3u);
    EXPECT_EQ(span2.size(), 2u);
    EXPECT_EQ(span1[0], 7); EXPECT_EQ(span1[1], 8); EXPECT_EQ(span1[2], 9);
    EXPECT_EQ(span2[0], 0); EXPECT_EQ(span2[1], 1);
  }
  {  // SplitEmpty :70-76
    RingBuffer<int> rb(10);
    auto [span1, span2] = rb.split(5, 0);
    EXPECT_TRUE(span1.empty());
    EXPECT_TRUE(span2.empty());
  }
  {  // SplitFullBuffer :78-95, SplitWrapAroundFullBuffer :97-117, SplitStartBeyondCapacity :119-137
    RingBuffer<int> rb(5);
    for (size_t i = 0; i < 5; ++i) rb.data()[i] = static_cast<int>(i);
    {
      auto [span1, span2] = rb.split(0, 5);
      EXPECT_EQ(span1.size(), 5u);
      EXPECT_TRUE(span2.empty());
      for (size_t i = 0; i < 5; ++i) EXPECT_EQ(span1[i], static_cast<int>(i));
    }
    {
      auto [span1, span2] = rb.split(3, 5);
      EXPECT_EQ(span1.size(), 2u);
      EXPECT_EQ(span2.size(), 3u);
      EXPECT_EQ(span1[0], 3); EXPECT_EQ(span1[1], 4);
      EXPECT_EQ(span2[0], 0); EXPECT_EQ(span2[1], 1); EXPECT_EQ(span2[2], 2);
    }
    {
      auto [span1, span2] = rb.split(7, 3);     // 7 % 5 = 2
      EXPECT_EQ(span1.size(), 3u);
      EXPECT_TRUE(span2.empty());
      EXPECT_EQ(span1[0], 2); EXPECT_EQ(span1[1], 3); EXPECT_EQ(span1[2], 4);
    }
    {  // LargeLenClamping :221-234
      auto [span1, span2] = rb.split(0, 10);
      EXPECT_EQ(span1.size() + span2.size(), 5u);
      EXPECT_EQ(span1.size(), 5u);
      EXPECT_TRUE(span2.empty());
    }
  }
  {  // SplitSpansConcatenation :139-176
    RingBuffer<int> rb(8);
    for (size_t i = 0; i < 8; ++i) rb.data()[i] = static_cast<int>(i * 10);
    std::vector<std::pair<size_t, size_t>> test_cases = {{0, 3}, {5, 4}, {6, 5}, {2, 8}};
    for (const auto& [start, len] : test_cases) {
      auto [span1, span2] = rb.split(start, len);
      std::vector<int> concatenated;
      for (size_t i = 0; i < span1.size(); ++i) concatenated.push_back(span1[i]);
      for (size_t i = 0; i < span2.size(); ++i) concatenated.push_back(span2[i]);
      EXPECT_EQ(concatenated.size(), len);
      for (size_t i = 0; i < len; ++i) EXPECT_EQ(concatenated[i], static_cast<int>(((start + i) % 8) * 10));
    }
  }
  {  // ShadowSyncStub :178-182
    RingBuffer<float> rb(100);
    CHECK_NOTHROW(rb.shadow_sync(50));
  }
  {  // ConstSplitReturnsConstSpans :184-206
    RingBuffer<int> rb(10);
    for (size_t i = 0; i < 10; ++i) rb.data()[i] = static_cast<int>(i);
    const RingBuffer<int>& const_rb = rb;
    auto [span1, span2] = const_rb.split(2, 5);
    static_assert(std::is_same_v<decltype(span1), dsp::base::Span<const int>>);
    static_assert(std::is_same_v<decltype(span2), dsp::base::Span<const int>>);
    EXPECT_EQ(span1.size(), 5u);
    EXPECT_TRUE(span2.empty());
    for (size_t i = 0; i < 5; ++i) EXPECT_EQ(span1[i], static_cast<int>(i + 2));
  }
  {  // OverflowProtection :208-217, ZeroCapacityThrows :219-221, NoexceptMethods :236-244
    RingBuffer<int> rb(10);
    auto [span1, span2] = rb.split(0, std::numeric_limits<size_t>::max());
    EXPECT_EQ(span1.size() + span2.size(), 10u);
    EXPECT_THROW(RingBuffer<int>(0), std::invalid_argument);
    static_assert(noexcept(rb.capacity()));
    static_assert(noexcept(rb.write_pos()));
    static_assert(noexcept(rb.data()));
    static_assert(noexcept(std::as_const(rb).data()));
  }
  {  // write(): wrap + shadow mirror (ring_buffer.cc:166-205), contiguous_read_ptr (:144-163)
    RingBuffer<float> rb(8, true);
    std::vector<float> a = {1, 2, 3, 4, 5, 6}, b = {7, 8, 9, 10, 11};
    EXPECT_EQ(rb.write(a.data(), a.size()), 6u);
    EXPECT_EQ(rb.write_pos(), 6u);
    EXPECT_EQ(rb.write(b.data(), b.size()), 5u);           // 2 in the tail, 3 wrap to the head
    EXPECT_EQ(rb.write_pos(), 3u);
    const float want[8] = {9, 10, 11, 4, 5, 6, 7, 8};
    bool ok = true;
    for (int i = 0; i < 8; ++i) ok = ok && rb.data()[i] == want[i];
    CHECK(ok);
    ok = true;
    for (int i = 0; i < 3; ++i) ok = ok && rb.data()[8 + i] == want[i];    // mirrored head
    CHECK(ok);
    const float* c = rb.contiguous_read_ptr(6);            // window [6, 6+5) is contiguous through the mirror
    CHECK(c == rb.data() + 6 && c[0] == 7 && c[1] == 8 && c[2] == 9 && c[3] == 10 && c[4] == 11);
    EXPECT_EQ(rb.write(nullptr, 4), 0u);
    EXPECT_EQ(rb.write(a.data(), 0), 0u);
    RingBuffer<float> plain(8);
    CHECK(plain.contiguous_read_ptr(13) == plain.data() + 5);
    // the cells are device memory too: the device-pointer seam functions operate on data() directly
    std::vector<float> ones(8, 1.0f);
    plain.write(ones.data(), 8);
    RingBuffer<float> src(8), nrm(8), out(8);
    for (int i = 0; i < 8; ++i) { src.data()[i] = float(i); nrm.data()[i] = i < 4 ? 2.0f : 0.0f; }
    dsp::ola::axpy_device(plain.data(), src.data(), 0.5f, 8);                  // 1 + i/2
    dsp::ola::normalize_and_clear_device(out.data(), plain.data(), nrm.data(), 0.25f, 8);
    CHECK(crlot_cuda_stream_sync(0, nullptr) == CRLOT_OK);
    ok = true;
    for (int i = 0; i < 8; ++i) ok = ok && out.data()[i] == (1.0f + 0.5f * i) / (i < 4 ? 2.0f : 0.25f) && plain.data()[i] == 0.0f;
    CHECK(ok);
  }
}

// ---- build_norm_linear (tests/norm_builder_test.cc:13-167) ----------------------------------------
static void build_norm_scalar(float* norm, const float* window, size_t ring_len, int frame_size, int hop) {   // :17-47
  std::fill(norm, norm + ring_len, 0.0f);
  for (size_t ring_pos = 0; ring_pos < ring_len; ++ring_pos) {
    float overlap_sum = 0.0f;
    int min_offset = -static_cast<int>(std::ceil(static_cast<double>(frame_size) / hop));
    int max_offset = static_cast<int>(std::ceil(static_cast<double>(ring_len + frame_size - 1) / hop));
    for (int frame_offset = min_offset; frame_offset <= max_offset; ++frame_offset) {
      int64_t frame_start = static_cast<int64_t>(frame_offset) * hop;
      for (int t = 0; t < frame_size; ++t) {
        int64_t sample_pos = frame_start + t;
        size_t mapped_pos = static_cast<size_t>(sample_pos % static_cast<int64_t>(ring_len));
        if (sample_pos < 0) {
          mapped_pos = ring_len - (static_cast<size_t>(-sample_pos) % ring_len);
          if (mapped_pos == ring_len) mapped_pos = 0;
        }
        if (mapped_pos == ring_pos) overlap_sum += window[t];
      }
    }
    norm[ring_pos] = std::max(overlap_sum, 1e-8f);
  }
}
static void test_norm_builder() {
  using dsp::ola::build_norm_linear;
  std::vector<std::tuple<int, int, size_t>> cases = {{256, 64, 1024}, {128, 32, 512}, {512, 128, 2048}, {64, 16, 256}, {1024, 256, 4096}};
  for (auto [frame_size, hop, ring_len] : cases) {      // BasicAccuracy :50-74, ParameterCombinations :77-110
    std::vector<float> window(frame_size);
    for (int i = 0; i < frame_size; ++i) window[i] = 0.5f * (1.0f - std::cos(2.0f * M_PI * i / (frame_size - 1)));
    std::vector<float> norm_linear(ring_len), norm_scalar(ring_len);
    build_norm_linear(norm_linear.data(), window.data(), ring_len, frame_size, hop);
    build_norm_scalar(norm_scalar.data(), window.data(), ring_len, frame_size, hop);
    double max_error = 0.0;
    for (size_t i = 0; i < ring_len; ++i) max_error = std::max(max_error, (double)std::abs(norm_linear[i] - norm_scalar[i]));
    CHECK(max_error < 1e-5);
  }
  {  // EdgeCases :113-131 (rectangular window)
    std::vector<float> window(128, 1.0f), a(256), b(256);
    build_norm_linear(a.data(), window.data(), 256, 128, 32);
    build_norm_scalar(b.data(), window.data(), 256, 128, 32);
    bool ok = true;
    for (size_t i = 0; i < 256; ++i) ok = ok && std::abs(a[i] - b[i]) <= 1e-5f;
    CHECK(ok);
  }
  {  // MinimumValueGuarantee :134-152
    std::vector<float> window(256, 1e-10f), norm(1024);
    build_norm_linear(norm.data(), window.data(), 1024, 256, 256);
    bool ok = true;
    for (float v : norm) ok = ok && v >= 1e-10f;
    CHECK(ok);
  }
  // WindowLUT extras the reference's callers use (WindowLUT.h:146-199)
  auto& lut = WindowLUT::getInstance();
  lut.clearCache(true);
  auto p1 = lut.GetWindowSafe(WindowType::HANN, 256);
  auto p2 = lut.GetWindowSafe(WindowType::HANN, 256);
  CHECK(p1.get() == p2.get() && lut.getCacheSize() == 1);
  const uint64_t g0 = lut.getCurrentGeneration();
  lut.clearCache();
  CHECK(lut.getCurrentGeneration() == g0 + 1 && p1.get()[128] > 0.99f);     // outstanding references stay valid
  auto p3 = lut.GetWindowSafe(WindowType::HANN, 256, true);
  CHECK(WindowLUT::calculateRMSError(p1.get(), p1.get(), 256) == 0.0);
  CHECK(WindowLUT::calculateRMSError(p1.get(), p3.get(), 256) > 0.0);
  CHECK(WindowLUT::calculateRMSError(nullptr, p3.get(), 256) == 0.0);
#pragma GCC diagnostic push
#pragma GCC diagnostic ignored "-Wdeprecated-declarations"
  const float* legacy = lut.GetWindow(WindowType::HAMMING, 64);
  CHECK(legacy == lut.GetWindow(WindowType::HAMMING, 64) && std::abs(legacy[0] - 0.08f) < 1e-6f);
  CHECK_THROWS(lut.GetWindow(WindowType::HANN, 0), std::invalid_argument);
#pragma GCC diagnostic pop
  CHECK_THROWS(lut.GetWindowSafe(WindowType::HANN, 0), std::invalid_argument);
  CHECK_THROWS(WindowLUT(0, WindowType::HANN), std::invalid_argument);
  CHECK_THROWS(WindowLUT::getInstance().data(), std::runtime_error);
}

// ---- time-sharded round trip through the C-ABI: world 1, and world 2 (one thread per GPU) when two are visible
static void test_sharded(int ndev) {
  const size_t N = 1024, H = 256;
  const size_t S = 2 * crlot_ola_ring_len(N, H) + 3 * H;
  WindowLUT hann(N, WindowType::HANN);
  for (int world : {1, 2, 4}) {
    const size_t total = world * S;
    const auto x = noise(total, 77u + world);
    StftRoundTrip whole(N, H, hann.data(), hann.data(), BoundaryMode::DROP);
    std::vector<float> want(whole.output_len_for(total));
    whole.process(x.data(), total, total, 1, want.data(), want.size());
    // (a) logical ranks on ONE device, halos filled by slicing: needs no communicator
    std::vector<float> got(want.size(), -1.0f);
    for (int r = 0; r < world; ++r) {
      crlot_time_shard sh{};
      CHECK(crlot_shard_plan((long long)total, (int)N, (int)H, r, world, CRLOT_DROP, &sh) == CRLOT_OK);
      crlot_buf_t xb = nullptr, sp = nullptr, ob = nullptr;
      const size_t L = (size_t)(sh.in_hi - sh.in_lo);
      CHECK(crlot_buf_create((L + 16) * 4, 0, &xb) == CRLOT_OK);
      CHECK(crlot_buf_create((size_t)sh.n_frames * (N / 2 + 1) * 8, 0, &sp) == CRLOT_OK);
      CHECK(crlot_buf_create((size_t)sh.n_out * 4, 0, &ob) == CRLOT_OK);
      CHECK(crlot_buf_upload(xb, 0, x.data() + sh.in_lo, L * 4, nullptr) == CRLOT_OK);
      crlot_time_shard solo = sh;            // same offsets, no neighbours to talk to
      solo.world = 1; solo.rank = 0;
      CHECK(crlot_sharded_roundtrip_execute(whole.plan(), nullptr, &solo, (float*)crlot_buf_device_ptr(xb), crlot_buf_device_ptr(sp),
                                            nullptr, (float*)crlot_buf_device_ptr(ob), nullptr) == CRLOT_OK);
      CHECK(crlot_buf_download(ob, 0, got.data() + sh.t0, (size_t)sh.n_out * 4, nullptr) == CRLOT_OK);
      CHECK(crlot_cuda_stream_sync(0, nullptr) == CRLOT_OK);
      crlot_buf_destroy(xb); crlot_buf_destroy(sp); crlot_buf_destroy(ob);
    }
    CHECK(std::memcmp(got.data(), want.data(), want.size() * 4) == 0);          // bit-identical to the unsharded call
    if (world == 1) dump("shard_in", x.data(), x.size()), dump("shard_out", want.data(), want.size());
    // (b) real ranks: one thread per GPU, NCCL halo exchange (ncclCommInitAll)
    if (world > 1 && ndev >= world) {
      std::vector<int> devs(world);
      for (int i = 0; i < world; ++i) devs[i] = i;
      std::vector<crlot_shard_group_t> groups(world, nullptr);
      const int rc = crlot_shard_group_create_all(world, devs.data(), groups.data());
      CHECK(rc == CRLOT_OK);
      if (rc != CRLOT_OK) { std::printf("  group: %s\n", crlot_cuda_last_error()); continue; }
      std::vector<float> multi(want.size(), -1.0f);
      std::vector<int> bad(world, 0);
      std::vector<std::thread> th;
      for (int r = 0; r < world; ++r)
        th.emplace_back([&, r] {
          try {
            ShardedRoundTrip rt(total, N, H, hann.data(), hann.data(), r, world, groups[r], BoundaryMode::DROP, devs[r]);
            rt.load_owned(x.data() + rt.shard().own_lo);
            crlot_cuda_stream_sync(devs[r], nullptr);
            rt.execute();
            rt.read_output(multi.data() + rt.shard().t0);
          } catch (const std::exception& e) { std::printf("  rank %d: %s\n", r, e.what()); bad[r] = 1; }
        });
      for (auto& t : th) t.join();
      for (int r = 0; r < world; ++r) CHECK(!bad[r]);
      CHECK(std::memcmp(multi.data(), want.data(), want.size() * 4) == 0);
      for (auto g : groups) crlot_shard_group_destroy(g);
      std::printf("sharded world %d over NCCL: ok\n", world);
    }
  }
}

// ---- facade_test --loop <samples> <N> <H>: the reference's per-frame loop, timed (bench.py "facade_loop") ----
// bench/e2e_benchmark.cc:149-176 driven in the streaming order of bench/ola_benchmark.cc:150-157, every call a
// host-pointer call on the drop-in classes.  Prints one JSON line.
#include <chrono>
static int run_loop(size_t len, size_t N, size_t H) {
  const auto x = noise(len, 1234u);
  WindowLUT hann(N, WindowType::HANN);
  const float* w = hann.data();
  auto run = [&](std::vector<float>& y) -> double {
    Framer framer;
    framer.set_params(N, H, 1, BoundaryMode::DROP);
    FftPlanDesc desc{FftDomain::Real, (int)N, false, 1, 1, 1};
    auto plan = MakeFftPlan(desc);
    OLAConfig cfg{};
    cfg.sample_rate = 48000; cfg.frame_size = N; cfg.hop_size = H; cfg.channels = 1; cfg.eps = 1e-8f; cfg.apply_window_inside = true;
    OLAAccumulator ola(cfg);
    ola.set_window(w, (int)N);
    std::vector<float> frame(N), proc(N);
    std::vector<std::complex<float>> spec(N / 2 + 1);
    const auto t0 = std::chrono::steady_clock::now();
    framer.push(x.data(), len);
    size_t k = 0, pos = 0;
    float* outp[1];
    while (framer.pop(frame.data())) {
      for (size_t i = 0; i < N; ++i) proc[i] = frame[i] * w[i];
      plan->forward(proc.data(), spec.data());
      plan->inverse(spec.data(), proc.data());
      ola.push_frame_AoS(proc.data(), nullptr, k * H, 0, N, 1.0f);
      outp[0] = y.data() + pos;
      pos += ola.produce(outp, H);
      ++k;
    }
    if (k > 0 && N > H) { outp[0] = y.data() + pos; pos += ola.produce(outp, N - H); }
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    y.resize(pos);
    return dt;
  };
  std::vector<float> y(len + N), y2(len + N);
  run(y2);                                   // warm-up pass (plans, staging buffers, first-touch)
  y2.assign(len + N, 0.0f);
  const double dt = run(y);
  // whole-signal call on the same input: the per-frame route must give the same samples bit for bit
  StftRoundTrip rt(N, H, w, w, BoundaryMode::DROP);
  std::vector<float> whole(rt.output_len_for(len));
  rt.process(x.data(), len, len, 1, whole.data(), whole.size());
  bool same = whole.size() == y.size();
  size_t diff = 0;
  double num = 0.0, den = 0.0;
  for (size_t i = 0; same && i < y.size(); ++i) {
    if (std::memcmp(&y[i], &whole[i], 4) != 0) ++diff;
    num += (double)(y[i] - whole[i]) * (y[i] - whole[i]); den += (double)whole[i] * whole[i];
  }
  std::printf("{\"value\": %.6f, \"unit\": \"Msamples/s\", \"samples\": %zu, \"frame\": %zu, \"hop\": %zu, \"frames\": %zu, "
              "\"seconds\": %.6f, \"us_per_frame\": %.3f, \"rel_l2_vs_whole_signal_call\": %.3e, \"bit_mismatches_vs_whole_signal_call\": %zu, "
              "\"api\": \"dsp::Framer::pop -> window -> IFftPlan::forward -> IFftPlan::inverse -> OLAAccumulator::push_frame_AoS -> produce, "
              "host pointers at every call (bench/e2e_benchmark.cc:149-176 order)\"}\n",
              len / dt / 1e6, len, N, H, (len - N) / H + 1, dt, dt / ((len - N) / H + 1) * 1e6, den > 0 ? std::sqrt(num / den) : 0.0, diff);
  return same ? 0 : 1;
}

int main(int argc, char** argv) {
  if (argc > 1 && std::string(argv[1]) == "--loop") {
    int nd = 0;
    if (crlot_cuda_device_count(&nd) != CRLOT_OK || nd == 0) { std::printf("no CUDA device\n"); return 2; }
    try {
      return run_loop(argc > 2 ? std::strtoull(argv[2], nullptr, 10) : (1u << 22), argc > 3 ? std::atoi(argv[3]) : 2048,
                      argc > 4 ? std::atoi(argv[4]) : 512);
    } catch (const std::exception& e) { std::printf("unexpected exception: %s\n", e.what()); return 1; }
  }
  g_out = argc > 1 ? argv[1] : ".";
  int ndev = 0;
  if (crlot_cuda_device_count(&ndev) != CRLOT_OK || ndev == 0) {
    std::printf("no CUDA device: %s\n", crlot_cuda_last_error());
    return 2;
  }
  try {
    test_window();
    test_framer();
    test_fft();
    test_ola();
    test_e2e_loop();
    test_framer_stft();
    test_framequeue();
    test_wav_interleaved();
    test_ring_buffer();
    test_norm_builder();
    test_sharded(ndev);
  } catch (const std::exception& e) {
    std::printf("unexpected exception: %s\n", e.what());
    return 1;
  }
  std::printf("%d checks, %d failed\n", g_checks, g_fail);
  return g_fail ? 1 : 0;
}
