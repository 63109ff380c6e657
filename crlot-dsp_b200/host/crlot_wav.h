// crlot_wav.h -- WavReader with the reference's interface (ref: io/wav.h:13-77, io/wav.cc:11-141),
// enough to feed real files (interleaved [n][channels] float) into the interleaved entry points
// (crlot_roundtrip_host_interleaved, dsp::Framer::push).  Host-only, header-only, no GPU code.
//
// The reference decodes through dr_wav (third-party single-header library, not vendored under
// /root/reference and absent from this image), so this is a restatement of the RIFF/WAVE subset the
// reference's format guard accepts (io/wav.cc:31-57): 1 or 2 channels; 16/24/32-bit PCM or 32-bit
// IEEE float; scaling as dr_wav's drwav_read_pcm_frames_f32 publishes it:
//   s16 -> x * (1/32768)      s24 -> x * (1/8388608)      s32 -> x / 2147483648      f32 -> as is
// The 16-bit path is pinned by config 1's golden vector (tests/golden/oboe_c1.npz, generated from
// assets/oboe.wav); 24/32-bit are "parity unpinned" (no reference fixture exercises them).
#ifndef CRLOT_WAV_H_
#define CRLOT_WAV_H_

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

class WavReader {
 public:
  WavReader() = default;
  ~WavReader() { close(); }
  WavReader(const WavReader&) = delete;
  WavReader& operator=(const WavReader&) = delete;

  bool open(const std::string& filename) {
    close();
    f_ = std::fopen(filename.c_str(), "rb");
    if (!f_) return false;
    unsigned char hdr[12];
    if (std::fread(hdr, 1, 12, f_) != 12 || std::memcmp(hdr, "RIFF", 4) != 0 || std::memcmp(hdr + 8, "WAVE", 4) != 0)
      return fail();
    bool have_fmt = false;
    for (;;) {                                   // walk the chunks up to "data"
      unsigned char ch[8];
      if (std::fread(ch, 1, 8, f_) != 8) return fail();
      const uint32_t sz = le32(ch + 4);
      if (std::memcmp(ch, "fmt ", 4) == 0) {
        unsigned char fmt[40] = {0};
        const size_t want = sz < sizeof(fmt) ? sz : sizeof(fmt);
        if (sz < 16 || std::fread(fmt, 1, want, f_) != want) return fail();
        if (sz > want) std::fseek(f_, static_cast<long>(sz - want), SEEK_CUR);
        tag_ = le16(fmt);
        channels_ = le16(fmt + 2);
        rate_ = le32(fmt + 4);
        block_ = le16(fmt + 12);
        bits_ = le16(fmt + 14);
        if (tag_ == 0xFFFE && sz >= 26) tag_ = le16(fmt + 24);   // WAVE_FORMAT_EXTENSIBLE: sub-format GUID
        have_fmt = true;
      } else if (std::memcmp(ch, "data", 4) == 0) {
        data_bytes_ = sz;
        break;
      } else {
        std::fseek(f_, static_cast<long>(sz), SEEK_CUR);
      }
      if (sz & 1) std::fseek(f_, 1, SEEK_CUR);   // chunks are word aligned
    }
    // the reference's format guard (io/wav.cc:31-57)
    const bool channels_ok = channels_ == 1 || channels_ == 2;
    const bool is_pcm = tag_ == 1, is_f32 = tag_ == 3 && bits_ == 32;
    const bool bps_ok = bits_ == 16 || bits_ == 24 || bits_ == 32;
    if (!have_fmt || !channels_ok || !bps_ok || !(is_pcm || is_f32) || block_ == 0) return fail();
    total_frames_ = data_bytes_ / block_;
    pos_ = 0;
    open_ = true;
    return true;
  }
  void close() {
    if (f_) std::fclose(f_);
    f_ = nullptr;
    open_ = false;
  }
  // interleaved frames into buffer; same return rule as the reference (io/wav.cc:75-87)
  bool read(float* buffer, size_t frames_to_read, size_t* frames_read = nullptr) {
    if (!open_) return false;
    const uint64_t left = total_frames_ - pos_;
    const size_t n = frames_to_read < left ? frames_to_read : static_cast<size_t>(left);
    std::vector<unsigned char> raw(n * block_);
    const size_t got = raw.empty() ? 0 : std::fread(raw.data(), block_, n, f_);
    const size_t bps = bits_ / 8;
    for (size_t i = 0; i < got * channels_; ++i) {
      const unsigned char* p = raw.data() + i * bps;
      if (tag_ == 3) {
        std::memcpy(buffer + i, p, 4);
      } else if (bits_ == 16) {
        buffer[i] = static_cast<float>(static_cast<int16_t>(le16(p))) * 0.000030517578125f;
      } else if (bits_ == 24) {
        const int32_t v = static_cast<int32_t>((uint32_t(p[0]) << 8) | (uint32_t(p[1]) << 16) | (uint32_t(p[2]) << 24)) >> 8;
        buffer[i] = static_cast<float>(static_cast<double>(v) * 0.00000011920928955078125);
      } else {
        buffer[i] = static_cast<float>(static_cast<int32_t>(le32(p)) / 2147483648.0);
      }
    }
    pos_ += got;
    if (frames_read) *frames_read = got;
    return got > 0 || frames_to_read == 0;
  }
  std::vector<float> read_all() {
    if (!open_) return {};
    std::vector<float> data(static_cast<size_t>(total_frames_ - pos_) * channels_);
    if (data.empty()) return {};
    size_t got = 0;
    read(data.data(), data.size() / channels_, &got);
    data.resize(got * channels_);
    return data;
  }
  uint32_t get_channels() const { return open_ ? channels_ : 0; }
  uint32_t get_sample_rate() const { return open_ ? rate_ : 0; }
  uint64_t get_total_frames() const { return open_ ? total_frames_ : 0; }
  uint32_t get_bits_per_sample() const { return open_ ? bits_ : 0; }
  bool is_open() const { return open_; }

 private:
  static uint16_t le16(const unsigned char* p) { return static_cast<uint16_t>(p[0] | (p[1] << 8)); }
  static uint32_t le32(const unsigned char* p) { return uint32_t(p[0]) | (uint32_t(p[1]) << 8) | (uint32_t(p[2]) << 16) | (uint32_t(p[3]) << 24); }
  bool fail() { close(); return false; }

  std::FILE* f_ = nullptr;
  bool open_ = false;
  uint32_t tag_ = 0, channels_ = 0, rate_ = 0, block_ = 0, bits_ = 0, data_bytes_ = 0;
  uint64_t total_frames_ = 0, pos_ = 0;
};

#endif  // CRLOT_WAV_H_
