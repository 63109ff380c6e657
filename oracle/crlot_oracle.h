/* oracle/crlot_oracle.h -- CPU ORACLE.  TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * A plain-C restatement of the one reference hot path this repository accelerates
 * (crlot-dsp: Framer -> WindowLUT -> real FFT / inverse -> OLAAccumulator + NormBuilder).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker or the reported CPU baseline --
 * never on the product path (the product fails loudly when its CUDA library is missing).
 *
 * Pinning: every function below is compared BIT-FOR-BIT against the unmodified reference
 * compiled from /root/reference (oracle/_ref/libcrlot_ref_strict.so, built by
 * oracle/Makefile) in tests/test_oracle_vs_ref.py, and against committed golden vectors
 * generated from that same build (tests/golden/, script tests/golden/make_golden.py) in
 * tests/test_oracle_golden.py.  Parity is therefore PINNED.
 *
 * All "file:line" citations are relative to /root/reference.
 */
#ifndef CRLOT_ORACLE_H_
#define CRLOT_ORACLE_H_

#ifdef __cplusplus
extern "C" {
#endif

/* enums mirror dsp/window/WindowLUT.h:13-30 and dsp/frame/framer.h:11-14 */
enum { ORC_WIN_HANN = 0, ORC_WIN_HAMMING = 1, ORC_WIN_BLACKMAN = 2, ORC_WIN_RECT = 3 };
enum { ORC_NORM_NONE = 0, ORC_NORM_SUM_TO_ONE = 1, ORC_NORM_L2 = 2, ORC_NORM_OLA_UNITY = 3,
       ORC_NORM_OLA_SUM_WSQ = 4 };
enum { ORC_ZERO_PAD = 0, ORC_DROP = 1 };

/* WindowLUT table generation           dsp/window/WindowLUT.cc:215-388 */
int orc_window(int type, long n, int periodic, int norm, float* out);

/* Framer, one-shot push then pop-all    dsp/frame/framer.cc:37-182 */
long orc_frame_count(long len, long frame, long hop, int boundary_mode);
long orc_framer(const float* interleaved, long frames_in, long frame, long hop, long channels,
                int boundary_mode, float* frames_out, long max_frames);

/* FrameQueue centre/pad front end      dsp/frame/FrameQueue.cc:9-48,68-115, dsp/frame/Indexing.h:18-70 */
enum { ORC_PAD_CONSTANT = 0, ORC_PAD_REFLECT = 1, ORC_PAD_EDGE = 2 };
float orc_pad_value(const float* x, long len, long idx, int pad_mode);
long orc_framequeue_count(long len, long frame, long hop, int center);
long orc_framequeue(const float* x, long len, long frame, long hop, int center, int pad_mode, float* frames_out,
                    long max_frames);

/* kissfft-based real FFT through the adapter (scrub, strides, 1/N)
 *   dsp/fft/backends/kissfft_adapter.cc:83-168, third_party/kissfft/kiss_fft.c, kiss_fftr.c */
int orc_rfft(int nfft, int batch, int stride_in, int stride_out, const float* in, float* out_ri);
int orc_irfft(int nfft, int batch, int stride_in, int stride_out, const float* in_ri, float* out);
/* Complex-domain plan (forward_complex / inverse_complex)   kissfft_adapter.cc:171-246 */
int orc_cfft(int nfft, int inverse, int batch, int stride_in, int stride_out, const float* in_ri, float* out_ri);
int orc_rfft_many(int nfft, long n_frames, const float* in, float* out_ri);
int orc_irfft_many(int nfft, long n_frames, const float* in_ri, float* out);

/* NormBuilder + OLAAccumulator          dsp/ola/norm_builder.cc:8-52, OLAAccumulator.cc */
long orc_ring_len(long frame, long hop);
void orc_build_norm_linear(float* norm, const float* window, long ring_len, long frame, long hop);
void orc_ola_envelope(float* norm, const float* window_or_null, long ring_len, long frame, long hop,
                      int apply_window_inside, float eps);
void orc_axpy(float* dst, const float* src, float g, long n);
void orc_axpy_windowed(float* dst, const float* src, const float* win, float g, long n);
void orc_normalize_and_clear(float* out, float* acc, const float* norm, float eps, long n);

/* Faithful ring emulation of the streaming drive (same contract as ref_ola_stream). */
long orc_ola_stream(const float* frames, long n_frames, long frame, long hop, long channels,
                    const float* window, int set_window_flag, int pass_window_flag,
                    int apply_window_inside, float gain, float eps, float* out, float* peak_out);

/* Closed-form gather OLA (the formulation kernel 3 implements; SURVEY.md 8a "exact
 * output definition").  frames [n_frames][frame] mono; t0 = absolute index of out[0];
 * first_frame = absolute index of frames[0]; total_frames = frames of the whole signal. */
void orc_ola_gather(const float* frames, long first_frame, long n_frames, long total_frames,
                    long frame, long hop, const float* window_or_null, float gain,
                    const float* norm, long ring_len, float eps, long t0, long n_out, float* out);

/* Whole hot path, mono.  Same contract as ref_roundtrip in oracle/ref_shim.cc. */
long orc_roundtrip(const float* x, long len, int nfft, long hop, const float* window,
                   int boundary_mode, float gain, float eps, float* spectra, float* iframes,
                   float* out, long* out_len_p);

/* Counter-based synthetic signal shared with the GPU generator
 * (crlot-dsp_b200/csrc/signal_gen.cuh): x[i] = uniform[-1,1) from hash(seed, channel, i). */
void orc_synth_uniform(unsigned long long seed, unsigned long long channel, unsigned long long i0,
                       long n, float* out);
/* 3-tone test signal of bench/e2e_benchmark.cc:34-39 (440/880/1320 Hz @ 48 kHz). */
void orc_synth_tones(unsigned long long i0, long n, float* out);

/* Timed CPU baseline: `threads` independent mono round trips of `len` samples each
 * (one per thread, distinct channels of the synthetic generator).  Returns wall seconds
 * of the slowest thread region (barrier to barrier), <0 on error. */
double orc_bench_roundtrip_mt(int threads, long len, int nfft, long hop, int window_type,
                              unsigned long long seed, double* checksum_out);

#ifdef __cplusplus
}
#endif
#endif /* CRLOT_ORACLE_H_ */
