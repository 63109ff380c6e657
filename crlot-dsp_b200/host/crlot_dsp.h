// crlot_dsp.h -- the reference's C++ class surface for the STFT -> iSTFT path, re-hosted on the
// B200 C-ABI (include/crlot_cuda.h).  Header-only; link with -lcrlot_cuda.
//
// Same namespaces, class names, method names, argument meaning and exception types as
//   dsp::BoundaryMode / dsp::Framer                 ref: dsp/frame/framer.h:11-131
//   dsp::PadMode / dsp::FrameQueue                  ref: dsp/frame/FrameQueue.h:8-112
//   dsp::WindowType / NormalizationType / WindowLUT ref: dsp/window/WindowLUT.h:13-190
//   dsp::fft::FftDomain / FftPlanDesc / IFftPlan / MakeFftPlan
//                                                   ref: dsp/fft/api/fft_api.h:9-51
//   dsp::OLAConfig / dsp::OLAAccumulator            ref: dsp/ola/OLAAccumulator.h:15-164
//   dsp::ola::build_norm_linear                     ref: dsp/ola/norm_builder.h:27-28
//   dsp::base::Span / dsp::ring::RingBuffer<T>      ref: dsp/base/span.h:11-33, dsp/ring/ring_buffer.h:25-110
// so that code written against the reference (bench/e2e_benchmark.cc:138-186, tests/*.cc) compiles
// against this header unchanged.  Nothing here computes on the CPU: every method forwards to the
// C-ABI, whose compute entry points launch sm_100a kernels or fail (no fallback).
//
// Extensions that the reference does not have are grouped under "B200 extension" comments; they
// are what makes the path fast (whole-signal calls instead of one frame per call).
#ifndef CRLOT_DSP_HOST_FACADES_H_
#define CRLOT_DSP_HOST_FACADES_H_

#include <complex>
#include <cstddef>
#include <cstdint>
#include <limits>
#include <map>
#include <memory>
#include <mutex>
#include <new>
#include <stdexcept>
#include <string>
#include <thread>
#include <tuple>
#include <type_traits>
#include <utility>
#include <vector>
#include <cmath>

#include "../../include/crlot_cuda.h"

namespace dsp {

namespace detail {
// status code -> the exception type the reference throws at the corresponding site
inline void check(int rc) {
  if (rc == CRLOT_OK) return;
  const std::string msg = crlot_cuda_last_error();
  switch (rc) {
    case CRLOT_ERR_INVALID_ARGUMENT: throw std::invalid_argument(msg);
    case CRLOT_ERR_ALLOC: throw std::bad_alloc();
    default: throw std::runtime_error(msg);   // RUNTIME, UNSUPPORTED, CUDA
  }
}
inline int& default_device() {
  static int dev = 0;
  return dev;
}
}  // namespace detail

// B200 extension: device ordinal used by objects constructed afterwards (one process per GPU)
inline void set_default_device(int device) { detail::default_device() = device; }

// ------------------------------------------------------------------------------------------------
// Framer
// ------------------------------------------------------------------------------------------------
enum class BoundaryMode { ZERO_PAD, DROP };

class Framer {
 public:
  Framer() = default;
  ~Framer() { crlot_framer_destroy(h_); }
  Framer(const Framer&) = delete;
  Framer& operator=(const Framer&) = delete;

  void set_params(size_t frame_size, size_t hop_size, size_t channels = 1,
                  BoundaryMode boundary_mode = BoundaryMode::ZERO_PAD) {
    crlot_framer_t nh = nullptr;
    detail::check(crlot_framer_create(frame_size, hop_size, channels,
                                      boundary_mode == BoundaryMode::DROP ? CRLOT_DROP : CRLOT_ZERO_PAD,
                                      detail::default_device(), &nh));
    crlot_framer_destroy(h_);
    h_ = nh;
    frame_size_ = frame_size; hop_size_ = hop_size; channels_ = channels; boundary_mode_ = boundary_mode;
  }
  bool push(const float* interleaved, size_t frames) {
    if (!h_) return false;
    if (interleaved == nullptr && frames > 0) return false;
    detail::check(crlot_framer_push(h_, interleaved, frames));
    return true;
  }
  bool pop(float* out_frame) {
    if (!h_ || out_frame == nullptr) return false;
    int popped = 0;
    detail::check(crlot_framer_pop(h_, out_frame, &popped));
    return popped != 0;
  }
  size_t available_frames() const { return h_ ? crlot_framer_available(h_) : 0; }
  void reset() { if (h_) detail::check(crlot_framer_reset(h_)); }
  size_t frame_size() const { return frame_size_; }
  size_t hop_size() const { return hop_size_; }
  size_t channels() const { return channels_; }
  BoundaryMode boundary_mode() const { return boundary_mode_; }
  size_t buffer_size() const { return h_ ? crlot_framer_get(h_, 4) : 0; }

  // ---- B200 extension --------------------------------------------------------------------------
  crlot_framer_t handle() const { return h_; }
  size_t drainable_frames() const { return h_ ? crlot_framer_get(h_, 8) : 0; }

 private:
  crlot_framer_t h_ = nullptr;
  size_t frame_size_ = 0, hop_size_ = 0, channels_ = 1;
  BoundaryMode boundary_mode_ = BoundaryMode::ZERO_PAD;
};

// ------------------------------------------------------------------------------------------------
// FrameQueue (centre / pad front end)            ref: dsp/frame/FrameQueue.h:8-112
// ------------------------------------------------------------------------------------------------
enum class PadMode { CONSTANT, REFLECT, EDGE };

class FrameQueue {
 public:
  // Same checks, messages and frame layout as the reference constructor (FrameQueue.cc:9-48); the
  // frames are gathered on the device (crlot_framequeue_extract) and kept as the AoS vector.
  FrameQueue(const float* in, size_t len, size_t frame_size, size_t hop_size, bool center = true,
             PadMode pad_mode = PadMode::CONSTANT)
      : frame_size_(frame_size), hop_size_(hop_size), center_(center), pad_mode_(pad_mode) {
    if (frame_size == 0) throw std::invalid_argument("Frame size must be greater than 0");
    if (hop_size == 0) throw std::invalid_argument("Hop size must be greater than 0");
    if (in == nullptr && len > 0) throw std::invalid_argument("Input pointer cannot be null when length > 0");
    const long long nf = crlot_framequeue_count((long long)len, (int)frame_size, (int)hop_size, center ? 1 : 0);
    num_frames_ = nf > 0 ? static_cast<size_t>(nf) : 0;
    frames_.resize(num_frames_ * frame_size_);
    if (num_frames_ == 0) return;
    const int dev = detail::default_device();
    crlot_buf_t din = nullptr, dfr = nullptr;
    detail::check(crlot_buf_create((len ? len : 1) * sizeof(float), dev, &din));
    int rc = crlot_buf_create(frames_.size() * sizeof(float), dev, &dfr);
    if (rc == CRLOT_OK && len) rc = crlot_buf_upload(din, 0, in, len * sizeof(float), nullptr);
    if (rc == CRLOT_OK)
      rc = crlot_framequeue_extract(static_cast<const float*>(crlot_buf_device_ptr(din)), (long long)len, (int)frame_size,
                                    (int)hop_size, center ? 1 : 0, static_cast<int>(pad_mode), nf,
                                    static_cast<float*>(crlot_buf_device_ptr(dfr)), dev, nullptr);
    if (rc == CRLOT_OK) rc = crlot_buf_download(dfr, 0, frames_.data(), frames_.size() * sizeof(float), nullptr);
    if (rc == CRLOT_OK) rc = crlot_cuda_stream_sync(dev, nullptr);
    crlot_buf_destroy(din);
    crlot_buf_destroy(dfr);
    detail::check(rc);
  }
  size_t getNumFrames() const { return num_frames_; }
  size_t getFrameSize() const { return frame_size_; }
  size_t getHopSize() const { return hop_size_; }
  const float* getFrame(size_t frame_idx) const {
    if (frame_idx >= num_frames_) throw std::out_of_range("Frame index out of range");
    return &frames_[frame_idx * frame_size_];
  }
  void copyFrame(size_t frame_idx, float* output) const {
    if (frame_idx >= num_frames_) throw std::out_of_range("Frame index out of range");
    if (output == nullptr) throw std::invalid_argument("Output buffer cannot be null");
    const float* f = &frames_[frame_idx * frame_size_];
    for (size_t i = 0; i < frame_size_; ++i) output[i] = f[i];
  }
  const std::vector<float>& getAllFrames() const { return frames_; }

 private:
  size_t frame_size_, hop_size_, num_frames_ = 0;
  bool center_;
  PadMode pad_mode_;
  std::vector<float> frames_;
};

// ------------------------------------------------------------------------------------------------
// WindowLUT
// ------------------------------------------------------------------------------------------------
enum class WindowType { HANN, HAMMING, BLACKMAN, RECT, BLACKMAN_HARRIS };
enum class NormalizationType { NONE, SUM_TO_ONE, L2_NORM, OLA_UNITY_GAIN, OLA_SUM_WSQ };

class WindowLUT {
 public:
  WindowLUT(size_t nfft, WindowType type, bool periodic = false, NormalizationType norm = NormalizationType::NONE)
      : size_(nfft), type_(type), periodic_(periodic), norm_(norm) {
    if (nfft == 0) throw std::invalid_argument("Window size must be greater than 0");
    data_ = make(type, nfft, periodic, norm);
  }
  WindowLUT() = default;
  WindowLUT(const WindowLUT&) = delete;
  WindowLUT& operator=(const WindowLUT&) = delete;

  const float* data() const {
    if (!data_) throw std::runtime_error("Window data not initialized");      // ref: WindowLUT.cc:63-68
    return data_.get();
  }
  size_t size() const { return data_ ? size_ : 0; }
  WindowType type() const { return data_ ? type_ : WindowType::HANN; }
  bool periodic() const { return data_ ? periodic_ : false; }
  NormalizationType normalization() const { return data_ ? norm_ : NormalizationType::NONE; }

  // Process-wide memo of tables behind a mutex and a generation counter, as the reference's
  // (WindowLUT.cc:60-165): a hit hands out shared ownership of the cached table, clearCache() bumps the
  // generation so that older entries are rebuilt on their next request while outstanding references stay valid.
  std::shared_ptr<const float> GetWindowSafe(WindowType type, size_t N, bool periodic = false,
                                             NormalizationType norm = NormalizationType::NONE) {
    if (N == 0) throw std::invalid_argument("Window size must be greater than 0");
    const Key key{static_cast<int>(type), N, periodic, static_cast<int>(norm)};
    std::lock_guard<std::mutex> lock(cache_mutex_);
    auto it = safe_cache_.find(key);
    if (it != safe_cache_.end() && it->second.second == generation_) return it->second.first;
    auto p = make(type, N, periodic, norm);
    safe_cache_[key] = {p, generation_};
    return p;
  }
  // ref: WindowLUT.h:155-157 (deprecated there as well; valid until clearCache())
  [[deprecated("Use GetWindowSafe() for thread safety")]]
  const float* GetWindow(WindowType type, size_t N, bool periodic = false,
                         NormalizationType norm = NormalizationType::NONE) {
    if (N == 0) throw std::invalid_argument("Window size must be greater than 0");
    const Key key{static_cast<int>(type), N, periodic, static_cast<int>(norm)};
    std::lock_guard<std::mutex> lock(cache_mutex_);
    auto it = legacy_cache_.find(key);
    if (it != legacy_cache_.end()) return it->second.get();
    auto p = make(type, N, periodic, norm);
    legacy_cache_[key] = p;
    return p.get();
  }
  size_t getCacheSize() const {
    std::lock_guard<std::mutex> lock(cache_mutex_);
    return safe_cache_.size() + legacy_cache_.size();
  }
  void clearCache(bool force_immediate = false) {
    std::lock_guard<std::mutex> lock(cache_mutex_);
    ++generation_;
    legacy_cache_.clear();
    if (force_immediate) safe_cache_.clear();
    else
      for (auto it = safe_cache_.begin(); it != safe_cache_.end();)
        it = it->second.second < generation_ ? safe_cache_.erase(it) : std::next(it);
  }
  uint64_t getCurrentGeneration() const {
    std::lock_guard<std::mutex> lock(cache_mutex_);
    return generation_;
  }
  static WindowLUT& getInstance() {
    static WindowLUT inst;
    return inst;
  }
  static double calculateSum(const float* window, size_t N) {
    double s = 0.0;
    for (size_t i = 0; i < N; ++i) s += static_cast<double>(window[i]);
    return s;
  }
  static double calculateSumOfSquares(const float* window, size_t N) {
    double s = 0.0;
    for (size_t i = 0; i < N; ++i) s += static_cast<double>(window[i]) * static_cast<double>(window[i]);
    return s;
  }
  // ref: WindowLUT.cc:195-206
  static double calculateRMSError(const float* window1, const float* window2, size_t N) {
    if (!window1 || !window2 || N == 0) return 0.0;
    double s = 0.0;
    for (size_t i = 0; i < N; ++i) {
      const double d = static_cast<double>(window1[i]) - static_cast<double>(window2[i]);
      s += d * d;
    }
    return std::sqrt(s / static_cast<double>(N));
  }

 private:
  static std::shared_ptr<const float> make(WindowType type, size_t N, bool periodic, NormalizationType norm) {
    std::shared_ptr<float> p(new float[N ? N : 1], std::default_delete<float[]>());
    detail::check(crlot_window_make(static_cast<int>(type), N, periodic ? 1 : 0, static_cast<int>(norm), p.get()));
    return p;
  }
  using Key = std::tuple<int, size_t, bool, int>;
  mutable std::mutex cache_mutex_;
  std::map<Key, std::pair<std::shared_ptr<const float>, uint64_t>> safe_cache_;
  std::map<Key, std::shared_ptr<const float>> legacy_cache_;
  uint64_t generation_ = 0;
  std::shared_ptr<const float> data_;
  size_t size_ = 0;
  WindowType type_ = WindowType::HANN;
  bool periodic_ = false;
  NormalizationType norm_ = NormalizationType::NONE;
};

// ------------------------------------------------------------------------------------------------
// OLAAccumulator
// ------------------------------------------------------------------------------------------------
struct OLAConfig {
  int sample_rate;
  size_t frame_size;
  size_t hop_size;
  size_t channels;
  float eps = 1e-8f;
  bool apply_window_inside;
  bool shadow_ring = false;
  bool isValid() const {
    return sample_rate > 0 && frame_size > 0 && hop_size > 0 && channels > 0 && eps > 0.0f;
  }
};

class OLAAccumulator {
 public:
  explicit OLAAccumulator(const OLAConfig& cfg) : cfg_(cfg) {
    detail::check(crlot_ola_create(cfg.sample_rate, cfg.frame_size, cfg.hop_size, cfg.channels, cfg.eps,
                                   cfg.apply_window_inside ? 1 : 0, cfg.shadow_ring ? 1 : 0,
                                   detail::default_device(), &h_));
  }
  ~OLAAccumulator() { crlot_ola_destroy(h_); }
  OLAAccumulator(const OLAAccumulator&) = delete;
  OLAAccumulator& operator=(const OLAAccumulator&) = delete;
  OLAAccumulator(OLAAccumulator&&) = delete;
  OLAAccumulator& operator=(OLAAccumulator&&) = delete;

  void set_window(const float* w, int wlen) { detail::check(crlot_ola_set_window(h_, w, wlen)); }
  void add_frame_SoA(const float* const* ch_frames, const float* window, size_t start_sample, size_t start_off,
                     size_t size, float gain) {
    detail::check(crlot_ola_add_frame_soa(h_, ch_frames, window, start_sample, start_off, size, gain));
  }
  void push_frame_AoS(const float* interleaved, const float* window, size_t start_sample, size_t start_off,
                      size_t size, float gain) {
    detail::check(crlot_ola_push_frame_aos(h_, interleaved, window, start_sample, start_off, size, gain));
  }
  size_t produce(float* const* ch_out, size_t n) {
    size_t produced = 0;
    detail::check(crlot_ola_produce(h_, ch_out, n, &produced));
    return produced;
  }
  void flush() { detail::check(crlot_ola_flush(h_)); }
  void reset() { detail::check(crlot_ola_reset(h_)); }
  size_t produced_samples() const { return crlot_ola_get(h_, 0); }
  size_t read_pos() const { return crlot_ola_get(h_, 1); }
  float meter_peak() const { return crlot_ola_meter_peak(h_); }
  const OLAConfig& config() const { return cfg_; }
  bool has_window() const { return crlot_ola_get(h_, 3) != 0; }
  size_t ring_size() const { return crlot_ola_get(h_, 2); }

  // ---- B200 extension --------------------------------------------------------------------------
  crlot_ola_t handle() const { return h_; }

 private:
  OLAConfig cfg_;
  crlot_ola_t h_ = nullptr;
};

// ------------------------------------------------------------------------------------------------
// dsp::ola free functions: the envelope builder and the kernel-dispatch seam
// ------------------------------------------------------------------------------------------------
namespace ola {
// ref: dsp/ola/norm_builder.h:27-28, norm_builder.cc:8-52 (fp32 adds in the reference's order; host table)
inline void build_norm_linear(float* norm, const float* window, size_t ring_len, size_t frame_size, size_t hop) {
  detail::check(crlot_build_norm_linear(norm, window, ring_len, frame_size, hop));
}
// B200 extension: the reference's dispatch functions (dsp/ola/kernels.h:28-53) on DEVICE pointers.
// noexcept like the reference's: a failed launch is reported by crlot_cuda_last_error().
inline void axpy_device(float* dst, const float* src, float g, size_t n, void* stream = nullptr) noexcept {
  crlot_axpy(dst, src, g, n, detail::default_device(), stream);
}
inline void axpy_windowed_device(float* dst, const float* src, const float* win, float g, size_t n,
                                 void* stream = nullptr) noexcept {
  crlot_axpy_windowed(dst, src, win, g, n, detail::default_device(), stream);
}
inline void normalize_and_clear_device(float* out, float* acc, const float* norm, float eps, size_t n,
                                       void* stream = nullptr) noexcept {
  crlot_normalize_and_clear(out, acc, norm, eps, n, detail::default_device(), stream);
}
}  // namespace ola

// ------------------------------------------------------------------------------------------------
// base::Span, ring::RingBuffer<T>          ref: dsp/base/span.h:11-33, dsp/ring/ring_buffer.h:25-110
// ------------------------------------------------------------------------------------------------
namespace base {
template <typename T>
class Span {
 public:
  Span() : ptr_(nullptr), len_(0) {}
  Span(T* ptr, size_t len) : ptr_(ptr), len_(len) {}
  T* data() const { return ptr_; }
  size_t size() const { return len_; }
  bool empty() const { return len_ == 0; }
  T& operator[](size_t index) const { return ptr_[index]; }
  T* begin() const { return ptr_; }
  T* end() const { return ptr_ + len_; }

 private:
  T* ptr_;
  size_t len_;
};
}  // namespace base

namespace ring {
// The cells live in CUDA unified memory (crlot_ring_*): data() and the spans are valid on the host, as in the
// reference, and on the device -- kernels launched on data() see the same cells, pages migrating on first touch.
// Index arithmetic (split / write / shadow_sync / contiguous_read_ptr) is the C-ABI's, i.e. the reference's.
template <typename T>
class RingBuffer {
 public:
  explicit RingBuffer(size_t capacity, bool shadow = false) {
    static_assert(std::is_trivially_copyable<T>::value, "Shadow ring requires trivially copyable T");
    detail::check(crlot_ring_create(capacity, sizeof(T), shadow ? 1 : 0, detail::default_device(), &h_));
    buffer_ = static_cast<T*>(crlot_ring_data(h_));
    capacity_ = capacity;
    shadow_ = shadow;
  }
  ~RingBuffer() { crlot_ring_destroy(h_); }
  RingBuffer(const RingBuffer&) = delete;
  RingBuffer& operator=(const RingBuffer&) = delete;
  RingBuffer(RingBuffer&&) = delete;
  RingBuffer& operator=(RingBuffer&&) = delete;

  size_t capacity() const noexcept { return capacity_; }
  size_t physical_capacity() const noexcept { return shadow_ ? capacity_ * 2 : capacity_; }
  bool has_shadow() const noexcept { return shadow_; }
  size_t write_pos() const noexcept { return crlot_ring_write_pos(h_); }

  [[nodiscard]] std::pair<base::Span<T>, base::Span<T>> split(size_t start, size_t len) noexcept {
    crlot_ring_span sp{};
    crlot_ring_split(h_, start, len, &sp);
    return {sp.len1 ? base::Span<T>(buffer_ + sp.off1, sp.len1) : base::Span<T>(),
            sp.len2 ? base::Span<T>(buffer_ + sp.off2, sp.len2) : base::Span<T>()};
  }
  [[nodiscard]] std::pair<base::Span<const T>, base::Span<const T>> split(size_t start, size_t len) const noexcept {
    crlot_ring_span sp{};
    crlot_ring_split(h_, start, len, &sp);
    return {sp.len1 ? base::Span<const T>(buffer_ + sp.off1, sp.len1) : base::Span<const T>(),
            sp.len2 ? base::Span<const T>(buffer_ + sp.off2, sp.len2) : base::Span<const T>()};
  }
  void shadow_sync(size_t bytes) { detail::check(crlot_ring_shadow_sync(h_, bytes)); }
  T* contiguous_read_ptr(size_t read_pos) noexcept { return buffer_ + crlot_ring_contiguous_offset(h_, read_pos); }
  const T* contiguous_read_ptr(size_t read_pos) const noexcept { return buffer_ + crlot_ring_contiguous_offset(h_, read_pos); }
  T* data() noexcept { return buffer_; }
  const T* data() const noexcept { return buffer_; }
  size_t write(const T* src, size_t n) {
    size_t written = 0;
    detail::check(crlot_ring_write(h_, src, n, &written));
    return written;
  }

  // ---- B200 extension --------------------------------------------------------------------------
  crlot_ring_t handle() const { return h_; }
  size_t write_device(const T* d_src, size_t n, void* stream = nullptr) {
    size_t written = 0;
    detail::check(crlot_ring_write_device(h_, d_src, n, &written, stream));
    return written;
  }
  void prefetch(bool to_device, void* stream = nullptr) { detail::check(crlot_ring_prefetch(h_, to_device ? 1 : 0, stream)); }

 private:
  crlot_ring_t h_ = nullptr;
  T* buffer_ = nullptr;
  size_t capacity_ = 0;
  bool shadow_ = false;
};
}  // namespace ring

// ------------------------------------------------------------------------------------------------
// FFT plan API
// ------------------------------------------------------------------------------------------------
namespace fft {

enum class FftDomain { Real, Complex };

struct FftPlanDesc {
  FftDomain domain;
  int nfft;
  bool in_place;
  int batch;
  int stride_in;
  int stride_out;
};

class IFftPlan {
 public:
  virtual ~IFftPlan() = default;
  virtual void forward(const float* in, std::complex<float>* out, int batch = 1) = 0;
  virtual void inverse(const std::complex<float>* in, float* out, int batch = 1) = 0;
  virtual void forward_complex(const std::complex<float>* in, std::complex<float>* out, int batch = 1) = 0;
  virtual void inverse_complex(const std::complex<float>* in, std::complex<float>* out, int batch = 1) = 0;
  virtual FftDomain domain() const = 0;
  virtual int size() const = 0;
  virtual bool supports_batch() const { return true; }
  virtual int max_batch_size() const { return 16; }
};

// IFftPlan over crlot_fft_* (the register-FFT kernels of K1/K2 run without window or framing)
class CudaFftPlan final : public IFftPlan {
 public:
  explicit CudaFftPlan(const FftPlanDesc& d) : desc_(d) {
    if (d.domain != FftDomain::Real && d.domain != FftDomain::Complex)
      throw std::runtime_error("Unsupported FFT domain");                 // ref: kissfft_adapter.cc:14-16
    if (d.domain == FftDomain::Real)
      dsp::detail::check(crlot_fft_plan_create(d.nfft, d.in_place ? 1 : 0, d.batch, d.stride_in, d.stride_out,
                                               dsp::detail::default_device(), &h_));
    else
      dsp::detail::check(crlot_fft_plan_create_complex(d.nfft, d.in_place ? 1 : 0, d.batch, d.stride_in, d.stride_out,
                                                       dsp::detail::default_device(), &h_));
  }
  ~CudaFftPlan() override { crlot_fft_plan_destroy(h_); }
  void forward(const float* in, std::complex<float>* out, int batch = 1) override {
    dsp::detail::check(crlot_fft_forward(h_, in, reinterpret_cast<float*>(out), batch));
  }
  void inverse(const std::complex<float>* in, float* out, int batch = 1) override {
    dsp::detail::check(crlot_fft_inverse(h_, reinterpret_cast<const float*>(in), out, batch));
  }
  void forward_complex(const std::complex<float>* in, std::complex<float>* out, int batch = 1) override {
    dsp::detail::check(crlot_fft_forward_complex(h_, reinterpret_cast<const float*>(in), reinterpret_cast<float*>(out), batch));
  }
  void inverse_complex(const std::complex<float>* in, std::complex<float>* out, int batch = 1) override {
    dsp::detail::check(crlot_fft_inverse_complex(h_, reinterpret_cast<const float*>(in), reinterpret_cast<float*>(out), batch));
  }
  FftDomain domain() const override { return desc_.domain; }
  int size() const override { return desc_.nfft; }

 private:
  FftPlanDesc desc_;
  crlot_fft_t h_ = nullptr;
};

inline std::unique_ptr<IFftPlan> MakeFftPlan(const FftPlanDesc& desc) {
  return std::unique_ptr<IFftPlan>(new CudaFftPlan(desc));
}

}  // namespace fft

// ------------------------------------------------------------------------------------------------
// B200 extension: the whole round trip as one object (what a maintainer would switch the loop of
// bench/e2e_benchmark.cc:149-176 to).  Host pointers in, host pointers out; device-resident
// variants through plan() and the C-ABI.
// ------------------------------------------------------------------------------------------------
class StftRoundTrip {
 public:
  StftRoundTrip(size_t frame_size, size_t hop_size, const float* analysis_window, const float* ola_window,
                BoundaryMode mode = BoundaryMode::DROP, bool apply_window_inside = true, float gain = 1.0f,
                float eps = 1e-8f) {
    crlot_stft_desc d{};
    d.nfft = static_cast<int>(frame_size);
    d.hop = static_cast<int>(hop_size);
    d.boundary_mode = mode == BoundaryMode::DROP ? CRLOT_DROP : CRLOT_ZERO_PAD;
    d.analysis_window = analysis_window;
    d.ola_window = ola_window;
    d.push_window = nullptr;
    d.apply_window_inside = apply_window_inside ? 1 : 0;
    d.gain = gain;
    d.eps = eps;
    d.scrub = 1;
    d.device = detail::default_device();
    detail::check(crlot_stft_plan_create(&d, &plan_));
    n_ = d.nfft; h_ = d.hop; mode_ = d.boundary_mode;
  }
  ~StftRoundTrip() { crlot_stft_plan_destroy(plan_); }
  StftRoundTrip(const StftRoundTrip&) = delete;
  StftRoundTrip& operator=(const StftRoundTrip&) = delete;

  size_t frames_for(size_t len) const { return static_cast<size_t>(crlot_frame_count((long long)len, n_, h_, mode_)); }
  size_t output_len_for(size_t len) const {
    const size_t f = frames_for(len);
    return f ? (f - 1) * static_cast<size_t>(h_) + static_cast<size_t>(n_) : 0;
  }
  // planar channels: in[c*chan_stride + i]; returns samples written per channel
  size_t process(const float* in, size_t chan_stride, size_t len, size_t channels, float* out, size_t out_chan_stride) {
    long long n_out = 0;
    detail::check(crlot_roundtrip_host(plan_, in, (long long)chan_stride, (long long)len, (int)channels, out,
                                       (long long)out_chan_stride, &n_out));
    return static_cast<size_t>(n_out);
  }
  crlot_plan_t plan() const { return plan_; }

 private:
  crlot_plan_t plan_ = nullptr;
  int n_ = 0, h_ = 0, mode_ = CRLOT_DROP;
};

// ------------------------------------------------------------------------------------------------
// B200 extension: one long signal time-sharded over the GPUs of a box (csrc/shard.cu).
// One object per rank (one process per GPU, or one thread per GPU): it owns the rank's plan, its
// [left halo | owned | right halo] input buffer, workspaces and output slice, all device-resident.
//   ShardedRoundTrip rt(total_len, N, H, w, w, rank, world, group);
//   rt.load_owned(host_samples_of_this_rank);   // or fill rt.owned_device_ptr() on the device
//   rt.execute();                               // NCCL halo exchange + K1 + fused K2+K3
//   rt.read_output(host_out);                   // output samples [t0, t1) of the whole run
// The concatenation over ranks is bit-identical to StftRoundTrip on the whole signal.
// ------------------------------------------------------------------------------------------------
class ShardedRoundTrip {
 public:
  ShardedRoundTrip(size_t total_len, size_t frame_size, size_t hop_size, const float* analysis_window,
                   const float* ola_window, int rank, int world, crlot_shard_group_t group,
                   BoundaryMode mode = BoundaryMode::DROP, int device = -1, bool materialise_frames = false)
      : group_(group), device_(device < 0 ? detail::default_device() : device) {
    detail::check(crlot_shard_plan((long long)total_len, (int)frame_size, (int)hop_size, rank, world,
                                   mode == BoundaryMode::DROP ? CRLOT_DROP : CRLOT_ZERO_PAD, &sh_));
    crlot_stft_desc d{};
    d.nfft = (int)frame_size; d.hop = (int)hop_size; d.boundary_mode = sh_.boundary_mode;
    d.analysis_window = analysis_window; d.ola_window = ola_window; d.apply_window_inside = 1;
    d.gain = 1.0f; d.eps = 1e-8f; d.scrub = 1; d.device = device_;
    detail::check(crlot_stft_plan_create(&d, &plan_));
    const size_t bins = frame_size / 2 + 1;
    const size_t own = (size_t)(sh_.own_hi - sh_.own_lo);
    // kernel 1 wants the buffer start 16-byte aligned (cudaMalloc gives 256); 16 floats of slack at the end
    detail::check(crlot_buf_create((pad_ + (size_t)sh_.left_halo + own + (size_t)sh_.right_halo + 16) * sizeof(float), device_, &x_));
    detail::check(crlot_buf_zero(x_, 0, crlot_buf_bytes(x_), nullptr));
    detail::check(crlot_buf_create(std::max<size_t>((size_t)sh_.n_frames * bins * 2 * sizeof(float), 8), device_, &spec_));
    if (materialise_frames)
      detail::check(crlot_buf_create(std::max<size_t>((size_t)sh_.n_frames * frame_size * sizeof(float), 8), device_, &frames_));
    detail::check(crlot_buf_create(std::max<size_t>((size_t)sh_.n_out * sizeof(float), 8), device_, &out_));
  }
  ~ShardedRoundTrip() {
    crlot_buf_destroy(x_); crlot_buf_destroy(spec_); crlot_buf_destroy(frames_); crlot_buf_destroy(out_);
    crlot_stft_plan_destroy(plan_);
  }
  ShardedRoundTrip(const ShardedRoundTrip&) = delete;
  ShardedRoundTrip& operator=(const ShardedRoundTrip&) = delete;

  const crlot_time_shard& shard() const { return sh_; }
  float* xbuf_device_ptr() const { return static_cast<float*>(crlot_buf_device_ptr(x_)) + pad_; }
  float* owned_device_ptr() const { return xbuf_device_ptr() + sh_.left_halo; }
  float* output_device_ptr() const { return static_cast<float*>(crlot_buf_device_ptr(out_)); }
  void* spectra_device_ptr() const { return crlot_buf_device_ptr(spec_); }
  // host samples [own_lo, own_hi) of the signal
  void load_owned(const float* host) {
    detail::check(crlot_buf_upload(x_, (pad_ + (size_t)sh_.left_halo) * sizeof(float), host,
                                   (size_t)(sh_.own_hi - sh_.own_lo) * sizeof(float), nullptr));
  }
  void execute(void* stream = nullptr) {
    detail::check(crlot_sharded_roundtrip_execute(plan_, group_, &sh_, xbuf_device_ptr(), crlot_buf_device_ptr(spec_),
                                                  frames_ ? static_cast<float*>(crlot_buf_device_ptr(frames_)) : nullptr,
                                                  output_device_ptr(), stream));
  }
  // output samples [t0, t1) of the whole run -> host (synchronises)
  size_t read_output(float* host, void* stream = nullptr) {
    detail::check(crlot_buf_download(out_, 0, host, (size_t)sh_.n_out * sizeof(float), stream));
    detail::check(crlot_cuda_stream_sync(device_, stream));
    return (size_t)sh_.n_out;
  }
  crlot_plan_t plan() const { return plan_; }

 private:
  crlot_time_shard sh_{};
  crlot_shard_group_t group_ = nullptr;
  int device_ = 0;
  size_t pad_ = 0;
  crlot_plan_t plan_ = nullptr;
  crlot_buf_t x_ = nullptr, spec_ = nullptr, frames_ = nullptr, out_ = nullptr;
};

}  // namespace dsp

#endif  // CRLOT_DSP_HOST_FACADES_H_
