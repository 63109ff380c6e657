/* include/crlot_cuda.h -- C-ABI of the B200 (sm_100a) implementation of crlot-dsp's
 * STFT -> iSTFT hot path:  Framer frame extraction, WindowLUT application, batched real
 * FFT / inverse FFT, OLAAccumulator overlap-add with the NormBuilder envelope.
 *
 * This is the ONE boundary between host code (the reference's C++ classes, kept as thin
 * facades in crlot-dsp_b200/host/, or any other language's FFI) and the CUDA side.  Plain
 * pointers and sizes only; every call returns an int status (0 = ok) and never throws;
 * crlot_cuda_last_error() gives the message for the calling thread.  There is no CPU
 * fallback behind any entry point: without a CUDA device every compute call fails.
 *
 * Each entry point names the reference interface it replaces ("ref:" = path relative to
 * the crlot-dsp tree).  INTEGRATION.md shows the reference-side binding.
 */
#ifndef CRLOT_CUDA_H_
#define CRLOT_CUDA_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CRLOT_ABI_VERSION 1

/* status codes; the C++ facades map them back to the reference's exception types */
enum {
  CRLOT_OK = 0,
  CRLOT_ERR_INVALID_ARGUMENT = 1, /* -> std::invalid_argument (ref: framer.cc:17-25, OLAAccumulator.cc:19,40,43,57) */
  CRLOT_ERR_RUNTIME = 2,          /* -> std::runtime_error    (ref: kissfft_adapter.cc:16-45,85-131)             */
  CRLOT_ERR_ALLOC = 3,            /* -> std::bad_alloc        (ref: aligned_alloc.h:21,32)                         */
  CRLOT_ERR_UNSUPPORTED = 4,      /* shape not implemented on the GPU path (no fallback by design)                */
  CRLOT_ERR_CUDA = 5              /* CUDA runtime failure; message carries cudaGetErrorString                     */
};

/* ref: dsp/window/WindowLUT.h:13-30 */
enum { CRLOT_WIN_HANN = 0, CRLOT_WIN_HAMMING = 1, CRLOT_WIN_BLACKMAN = 2, CRLOT_WIN_RECT = 3, CRLOT_WIN_BLACKMAN_HARRIS = 4 };
enum { CRLOT_NORM_NONE = 0, CRLOT_NORM_SUM_TO_ONE = 1, CRLOT_NORM_L2 = 2, CRLOT_NORM_OLA_UNITY_GAIN = 3, CRLOT_NORM_OLA_SUM_WSQ = 4 };
/* ref: dsp/frame/framer.h:11-14 */
enum { CRLOT_ZERO_PAD = 0, CRLOT_DROP = 1 };

/* ---- library ------------------------------------------------------------------------------- */
int crlot_cuda_abi_version(void);
const char* crlot_cuda_last_error(void);
int crlot_cuda_device_count(int* count);
/* cudaStreamSynchronize(stream) on `device` (stream NULL = default stream) */
int crlot_cuda_stream_sync(int device, void* stream);

/* ---- device buffer handles ------------------------------------------------------------------ */
typedef struct crlot_buf_s* crlot_buf_t;
int crlot_buf_create(size_t bytes, int device, crlot_buf_t* out);               /* owns cudaMalloc memory   */
int crlot_buf_wrap(void* device_ptr, size_t bytes, int device, crlot_buf_t* out); /* borrows caller's memory */
int crlot_buf_destroy(crlot_buf_t buf);
int crlot_buf_upload(crlot_buf_t buf, size_t dst_offset, const void* host, size_t bytes, void* stream);
int crlot_buf_download(crlot_buf_t buf, size_t src_offset, void* host, size_t bytes, void* stream);
int crlot_buf_zero(crlot_buf_t buf, size_t offset, size_t bytes, void* stream);
void* crlot_buf_device_ptr(crlot_buf_t buf);
size_t crlot_buf_bytes(crlot_buf_t buf);
/* page-locked host staging memory for the host-buffer entry points */
int crlot_host_alloc(size_t bytes, void** out);
int crlot_host_free(void* p);

/* ---- host-side table builders (tiny, double precision; uploaded once per plan) ---------------- */
/* ref: WindowLUT::createWindow + generators + applyNormalization, dsp/window/WindowLUT.cc:215-388 */
int crlot_window_make(int type, size_t n, int periodic, int norm, float* out);
/* ref: OLAAccumulator::calculate_ring_size, dsp/ola/OLAAccumulator.cc:249-258 */
size_t crlot_ola_ring_len(size_t frame, size_t hop);
/* ref: dsp::ola::build_norm_linear, dsp/ola/norm_builder.cc:8-52 (double count at the seam included) */
int crlot_build_norm_linear(float* norm, const float* window, size_t ring_len, size_t frame, size_t hop);
/* ref: OLAAccumulator::initialize_normalization, dsp/ola/OLAAccumulator.cc:260-288 */
int crlot_ola_envelope(float* norm, const float* window_or_null, size_t ring_len, size_t frame, size_t hop,
                       int apply_window_inside, float eps);
/* ref: Framer::calculate_available_frames after one push of `len` samples, dsp/frame/framer.cc:88-118 */
long long crlot_frame_count(long long len, int frame, int hop, int boundary_mode);

/* ---- STFT plan ------------------------------------------------------------------------------ */
typedef struct crlot_plan_s* crlot_plan_t;

typedef struct crlot_stft_desc {
  int nfft;                      /* frame size N: any even N <= 16384 (ref: kissfft_adapter.cc:36-39).  */
                                 /*   256..4096 (powers of two) with hop % 4 == 0 run the register-FFT  */
                                 /*   kernels; every other shape the any-size kernels (k_generic.cu)    */
  int hop;                       /* hop H: 0 < H <= N                                                  */
  int boundary_mode;             /* CRLOT_ZERO_PAD / CRLOT_DROP (Framer boundary mode)                 */
  const float* analysis_window;  /* host, N floats, or NULL (no multiply before the FFT)              */
                                 /*   ref: bench/e2e_benchmark.cc:154-156                             */
  const float* ola_window;       /* host, N floats, what OLAAccumulator::set_window received, or NULL */
  const float* push_window;      /* host, N floats, the per-call `window` argument of add_frame_SoA   */
                                 /*   (honoured only when !apply_window_inside), or NULL              */
  int apply_window_inside;       /* OLAConfig::apply_window_inside, ref: OLAAccumulator.h:21          */
  float gain;                    /* gain argument of add_frame_SoA / push_frame_AoS                   */
  float eps;                     /* OLAConfig::eps (default 1e-8f)                                    */
  int scrub;                     /* 1: NaN/Inf/|v|<1e-30 -> 0 on FFT input and inverse output         */
                                 /*   ref: kissfft_adapter.cc:104-108,159-163                         */
  const float* envelope;         /* host override of the normalisation envelope, or NULL = reference  */
  size_t envelope_len;           /*   rule (crlot_ola_envelope over crlot_ola_ring_len)               */
  int device;                    /* CUDA device ordinal                                               */
} crlot_stft_desc;

typedef struct crlot_plan_info {
  int nfft, hop, bins;           /* bins = nfft/2 + 1                                                  */
  size_t ring_len;               /* envelope period                                                    */
  int synthesis_window_active;   /* the window policy resolved (OLAAccumulator.cc:82-83)               */
  int device;
} crlot_plan_info;

/* Stream and thread rules.  Every execute entry point enqueues on the `stream` it is given and returns without
 * synchronising (host-buffer entry points synchronise before returning).  One plan may be used from several
 * streams / host threads at once: the dynamically scheduled kernels (K1, fused K2+K3, the one-kernel round trip)
 * take their work counter per stream (up to 1024 streams per plan are tracked; launches on ONE stream are ordered
 * and share a counter).  crlot_stft_plan_set_spectral_gain is plan state: change it only while no launch of that
 * plan is being enqueued.  The stateful objects (crlot_framer_t, crlot_ola_t, crlot_stream_t, crlot_fft_t) are, like
 * the reference's classes, for one host thread at a time; the Framer orders its own operations across streams. */
int crlot_stft_plan_create(const crlot_stft_desc* desc, crlot_plan_t* out);
int crlot_stft_plan_destroy(crlot_plan_t plan);
int crlot_stft_plan_info(crlot_plan_t plan, crlot_plan_info* info);
/* copies the plan's envelope (ring_len floats) to host */
int crlot_stft_plan_envelope(crlot_plan_t plan, float* host_out);

/* Spectral-domain hook between kernels 1 and 2.  The reference leaves an explicit gap
 * between IFftPlan::forward and IFftPlan::inverse for per-bin processing
 *   ref: bench/e2e_benchmark.cc:158-166 (forward; inverse), bench/performance_benchmark.cc:221
 * -- the reason spectra are materialised at all.  The hook is a REAL gain per bin:
 *   X'[row][k] = ( fl(Re X[row][k] * g), fl(Im X[row][k] * g) ),  g = d_gain[row*row_stride + k]
 * row_stride == 0: one table of `bins` gains for every frame (an equaliser);
 * row_stride >= bins: one row of gains per spectrum row, laid out like the spectra
 *                     [channels][frames][row_stride] (a time-frequency mask).
 * crlot_stft_plan_set_spectral_gain attaches the table (device pointer, caller-owned, must stay valid
 * until cleared with NULL) to the plan: crlot_istft_execute, crlot_istft_ola_execute,
 * crlot_roundtrip_execute and crlot_roundtrip_host then apply it as each spectrum row is LOADED by
 * the inverse kernel -- the spectra written by kernel 1 stay untouched and no extra pass over them
 * is made.  crlot_spectral_gain_apply is the same arithmetic as its own in-place pass; the two routes
 * give bit-identical frames/audio. */
int crlot_stft_plan_set_spectral_gain(crlot_plan_t plan, const float* d_gain, long long row_stride);
int crlot_spectral_gain_apply(crlot_plan_t plan, void* d_spec, long long rows, const float* d_gain,
                              long long row_stride, void* stream);

/* Kernel 1.  Frame k of channel c = d_signal[c*chan_stride + k*hop + (0..nfft)), samples at or
 * beyond `len` read as zero.  d_spec: [channels][n_frames][nfft/2+1] interleaved (re,im) fp32.
 * replaces: Framer::pop + analysis-window loop + IFftPlan::forward
 *   ref: dsp/frame/framer.cc:61-67,128-182; bench/e2e_benchmark.cc:152-156;
 *        dsp/fft/backends/kissfft_adapter.cc:83-122 */
int crlot_stft_execute(crlot_plan_t plan, const float* d_signal, long long chan_stride, long long len, int channels,
                       long long n_frames, void* d_spec, void* stream);

/* FrameQueue centre / pad front end as an addressing mode of kernel 1.
 * The stream the frames index is [pad | signal | pad] with pad = nfft/2 when center != 0 (else no
 * padding), frame k = stream[k*hop, k*hop + nfft); padding values follow getPaddingValueSafe:
 * CONSTANT zeros, EDGE clamps, REFLECT is reflect101 INCLUDING its asymmetric negative branch
 * (-1 -> 0, -2 -> 1, ...; n -> n-2) that the reference's tests pin.
 *   ref: dsp/frame/FrameQueue.cc:9-48,68-115; dsp/frame/Indexing.h:18-70;
 *        tests/frame_queue_test.cc:264-366
 * crlot_framequeue_count   = FrameQueue::getNumFrames()  (-1 on zero sizes)
 * crlot_framequeue_extract = FrameQueue::getAllFrames() on the device, [n_frames][frame], bit-exact
 * crlot_stft_execute_framequeue = kernel 1 over those frames without materialising them: the padded
 *   stream is never built, interior frames take the same bulk-copy path as crlot_stft_execute and
 *   only the pad positions are computed by index.  Lengths are limited to < 2^31 (the reference
 *   indexes the padded signal with int). */
enum { CRLOT_PAD_CONSTANT = 0, CRLOT_PAD_REFLECT = 1, CRLOT_PAD_EDGE = 2 };   /* dsp::PadMode */
long long crlot_framequeue_count(long long len, int frame, int hop, int center);
int crlot_framequeue_extract(const float* d_signal, long long len, int frame, int hop, int center, int pad_mode,
                             long long n_frames, float* d_frames, int device, void* stream);
int crlot_stft_execute_framequeue(crlot_plan_t plan, const float* d_signal, long long chan_stride, long long len,
                                  int channels, long long n_frames, int center, int pad_mode, void* d_spec,
                                  void* stream);
/* The same from HOST buffers: h_signal planar [channels][chan_stride], h_spec [channels][n_frames][nfft/2+1] complex,
 * n_frames = crlot_framequeue_count (also returned through n_frames_out).  Slabs of frames go through the plan's three
 * host-path workspace slots (H2D, kernel 1, D2H of the spectra overlap across slabs); interior slabs run the plain
 * path, the slabs that hold a signal end compute the pad positions.  Page-locked buffers (crlot_host_alloc) make the
 * copies asynchronous; blocks until h_spec is complete.
 * replaces: FrameQueue(input, len, frame, hop, center, pad_mode) + per frame window multiply + IFftPlan::forward, the
 * frame loop of the reference's batch-framing pipeline   ref: dsp/frame/FrameQueue.cc:9-115; bench/performance_benchmark.cc:174-246 */
int crlot_stft_host_framequeue(crlot_plan_t plan, const float* h_signal, long long chan_stride, long long len,
                               int channels, int center, int pad_mode, void* h_spec, long long* n_frames_out);

/* Kernel 2.  rows spectra -> rows frames of nfft floats, scaled by 1/nfft and cleaned up;
 * with fuse_synthesis_window != 0 each sample is additionally multiplied (one rounding) by the
 * plan's resolved synthesis window, so that crlot_ola_accumulate can be told frames are
 * pre-windowed.
 * replaces: IFftPlan::inverse   ref: dsp/fft/backends/kissfft_adapter.cc:124-168 */
int crlot_istft_execute(crlot_plan_t plan, const void* d_spec, long long rows, float* d_frames,
                        int fuse_synthesis_window, void* stream);

/* Kernel 3.  Gather overlap-add of local frames [0,frames_per_chan) of each channel into n_out
 * samples per channel.  Local frame k starts at local output index frame_t0 + k*hop; output
 * sample j has absolute index t_offset + j (selects the envelope phase).
 * replaces: OLAAccumulator::add_frame_SoA / push_frame_AoS + produce in the streaming drive
 *   ref: dsp/ola/OLAAccumulator.cc:54-122,162-221; dsp/ola/kernels.cc:18-36 */
int crlot_ola_accumulate(crlot_plan_t plan, const float* d_frames, long long frames_chan_rows,
                         long long frames_per_chan, int channels, int frames_prewindowed, long long t_offset,
                         long long frame_t0, float* d_out, long long out_chan_stride, long long n_out, void* stream);

/* Kernels 2+3 fused.  Same inputs as crlot_istft_execute (spectra [channels][spec_chan_rows][bins],
 * frames_per_chan of them used per channel) and same outputs / t_offset / frame_t0 / n_out meaning
 * as crlot_ola_accumulate, bit-identical to calling the two in sequence with the synthesis window
 * fused -- but the iFFT frames stay in shared memory: per frame 8*bins bytes are read and 4*hop
 * written, instead of 8*bins + 4*nfft + 4*nfft + 4*hop.  Returns CRLOT_ERR_UNSUPPORTED outside its
 * shapes (crlot_istft_ola_supported() == 0): nfft in {256, 512, 1024, 2048, 4096}, hop a multiple of 64 (16 for
 * nfft 256, 32 for 512/1024), even t_offset / frame_t0 / out_chan_stride, 8-byte aligned d_out.
 * replaces: IFftPlan::inverse + OLAAccumulator::push_frame_AoS + produce
 *   ref: kissfft_adapter.cc:124-168; OLAAccumulator.cc:54-122,162-221; dsp/ola/kernels.cc:18-36 */
int crlot_istft_ola_supported(crlot_plan_t plan, long long t_offset, long long frame_t0, const float* d_out,
                              long long out_chan_stride);
int crlot_istft_ola_execute(crlot_plan_t plan, const void* d_spec, long long spec_chan_rows, long long frames_per_chan,
                            int channels, long long t_offset, long long frame_t0, float* d_out,
                            long long out_chan_stride, long long n_out, void* stream);

/* Self-test of the fused kernel's table-driven IEEE division (csrc/common.cuh div_tail): n hashed
 * operand pairs, including the plan's own envelope denominators, compared bit for bit with the
 * hardware-correct __fdiv_rn on the device; *mismatches receives the number that differ (must be 0). */
int crlot_selftest_division(crlot_plan_t plan, unsigned long long n, unsigned long long seed,
                            unsigned long long* mismatches);

/* The composed round trip on device buffers (K1 -> K2 -> K3 on `stream`, or K1 -> fused K2+K3 when
 * d_frames is NULL, or the one-kernel round trip below when d_spec is NULL as well).  d_spec and d_frames
 * are caller-provided materialisation buffers ([channels][n_frames][bins] complex and
 * [channels][n_frames][nfft] fp32).  n_frames = crlot_frame_count(len,...); output length per
 * channel = (n_frames-1)*hop + nfft.
 * replaces the composition in bench/e2e_benchmark.cc:138-186 (driven in the order of
 * bench/ola_benchmark.cc:150-157). */
int crlot_roundtrip_execute(crlot_plan_t plan, const float* d_signal, long long chan_stride, long long len,
                            int channels, void* d_spec, float* d_frames, float* d_out, long long out_chan_stride,
                            void* stream);

/* The whole round trip in ONE kernel, for callers that need neither the spectra nor the iFFT frames (the reference's
 * loop does nothing between forward and inverse but the optional per-bin processing,
 *   ref: bench/e2e_benchmark.cc:158-166): each warp frames, windows and transforms a frame, applies the plan's
 * spectral gain (crlot_stft_plan_set_spectral_gain) in registers, inverse-transforms, overlap-adds and normalises --
 * 8 bytes of HBM traffic per sample instead of 72.  Same roundings in the same order as the three-kernel path:
 * bit-identical output.  Frame k = d_signal[c*chan_stride + k*hop + (0..nfft)), zero at and beyond len;
 * t_offset / frame_t0 / n_out as in crlot_ola_accumulate (time shards).  Exists for nfft 2048 / hop 512 (where the
 * forward transform's register layout is the inverse's and the overlapping three quarters of a frame stay in the
 * lane's tensor memory); crlot_roundtrip_fused_supported() == 0 otherwise and the call returns
 * CRLOT_ERR_UNSUPPORTED.  crlot_roundtrip_execute(d_spec = NULL, d_frames = NULL) forwards here.
 * replaces: Framer::pop + window + IFftPlan::forward + IFftPlan::inverse + OLAAccumulator::push_frame_AoS + produce */
int crlot_roundtrip_fused_supported(crlot_plan_t plan, const float* d_signal, long long chan_stride, long long t_offset,
                                    long long frame_t0, const float* d_out, long long out_chan_stride);
int crlot_roundtrip_fused_execute(crlot_plan_t plan, const float* d_signal, long long chan_stride, long long len,
                                  int channels, long long n_frames, long long t_offset, long long frame_t0, float* d_out,
                                  long long out_chan_stride, long long n_out, void* stream);

/* Same round trip with HOST buffers: host->device and device->host copies are inside the call,
 * processed in time slabs through three internal workspace slots (one stream each) so transfers overlap
 * compute.  h_out receives (n_frames-1)*hop + nfft samples per channel; *n_out_per_chan is set.
 * h_in / h_out should be page-locked (crlot_host_alloc) for full PCIe rate. */
int crlot_roundtrip_host(crlot_plan_t plan, const float* h_in, long long chan_stride, long long len, int channels,
                         float* h_out, long long out_chan_stride, long long* n_out_per_chan);

/* Interleaved (AoS) multichannel I/O.  Framer::push and WavReader hand over [n][channels]
 * samples (ref: dsp/frame/framer.cc:45,133-141; io/wav.cc:90-123) and OLAAccumulator::push_frame_AoS
 * de-interleaves with deinterleave_to_scratch (ref: dsp/ola/aos_to_soa.cc:5-19).
 * crlot_deinterleave / crlot_interleave are those copies on the device (bit-exact);
 * crlot_roundtrip_host_interleaved is crlot_roundtrip_host for interleaved host buffers: h_in holds
 * len*channels floats, h_out receives n_out*channels floats; every time slab carries all channels and
 * is transposed on the device between the PCIe copy and the kernels. */
int crlot_deinterleave(const float* d_interleaved, long long n, int channels, float* d_planar, long long chan_stride,
                       int device, void* stream);
int crlot_interleave(const float* d_planar, long long chan_stride, long long n, int channels, float* d_interleaved,
                     int device, void* stream);
int crlot_roundtrip_host_interleaved(crlot_plan_t plan, const float* h_in, long long len, int channels, float* h_out,
                                     long long* n_out_per_chan);

/* Device generator of the benchmark signal: x[c][i] = uniform[-1,1) from a counter hash of
 * (seed, channel0 + c, i0 + i); identical to oracle/crlot_oracle.c:orc_synth_uniform. */
int crlot_synth_uniform(float* d_out, long long chan_stride, int channels, long long n, unsigned long long seed,
                        unsigned long long channel0, unsigned long long i0, int device, void* stream);

/* ---- streaming state: many concurrent streams, chunked ring-buffer ingest (config 5) ------------ */
/* Device-resident analogue, per stream, of a Framer in DROP mode fed in chunks
 *   (ref: dsp/frame/framer.cc:37-59,88-118,128-182) followed by the window / forward / inverse
 *   / OLAAccumulator::push_frame_AoS(k*hop) / produce(hop) drive of
 *   bench/ola_benchmark.cc:150-157, with the accumulator in a RingBuffer of
 *   crlot_ola_ring_len() cells (ref: dsp/ring/ring_buffer.cc:44-85, OLAAccumulator.cc:249-258).
 * All streams of one state advance in lock step (same chunk length per tick).  The plan must use
 * CRLOT_DROP (chunked ZERO_PAD emits premature frames in the reference, framer.cc:98-103); every plan
 * shape with nfft % 4 == 0 and hop % 4 == 0 runs (256..2048 on the ring-reading kernels, nfft = 4096 and
 * the any-size plans through a gather + kernel 1).  Outputs are bit-identical to
 * crlot_roundtrip_execute on the concatenated input.  After a failed launch the state refuses further
 * ticks until crlot_stream_reset (host mirrors and device counters may disagree). */
typedef struct crlot_stream_s* crlot_stream_t;
int crlot_stream_create(crlot_plan_t plan, int n_streams, int max_chunk, crlot_stream_t* out);
int crlot_stream_destroy(crlot_stream_t st);
int crlot_stream_reset(crlot_stream_t st, void* stream);
/* frames a tick of chunk_len samples would complete, given the samples ingested so far */
int crlot_stream_frames_ready(crlot_stream_t st, int chunk_len);
/* One tick.  d_chunk: [n_streams][chunk_stride] device floats, chunk_len new samples per stream.
 * Every frame completed by them is transformed, inverse-transformed, accumulated and produced:
 * d_out [n_streams][out_stride] receives *n_out = new_frames*hop samples per stream. */
int crlot_stream_tick(crlot_stream_t st, const float* d_chunk, long long chunk_stride, int chunk_len,
                      float* d_out, long long out_stride, int* n_out, void* stream);
/* n_ticks ticks over consecutive chunks of d_in [n_streams][in_stride] (tick i reads samples
 * [i*chunk_len, (i+1)*chunk_len)); outputs are appended per stream in d_out; *n_out = total samples
 * written per stream.  The tick loop runs inside the library (CUDA graph replay in steady state; the
 * instantiated graphs are kept on the state and re-pointed per call).  Asynchronous on `stream`. */
int crlot_stream_run(crlot_stream_t st, const float* d_in, long long in_stride, int chunk_len, long long n_ticks,
                     float* d_out, long long out_stride, long long* n_out, void* stream);
/* One tick with HOST buffers (ideally page-locked): h_chunk [n_streams][chunk_stride] is copied to
 * the device, the tick runs, *n_out samples per stream are copied back into h_out
 * [n_streams][out_stride], and the call returns when h_out is complete (the low-latency use). */
int crlot_stream_tick_host(crlot_stream_t st, const float* h_chunk, long long chunk_stride, int chunk_len,
                           float* h_out, long long out_stride, int* n_out);
/* The same tick without the wait: enqueues it on the state's own stream and returns a ticket;
 * crlot_stream_wait_host(ticket) blocks until that tick's outputs are complete in h_out.  Several ticks may be
 * in flight (each with its own, still-valid buffers); they execute in order.  When both buffers are page-locked
 * (crlot_host_alloc / cudaHostRegister) the tick kernel reads the chunk and writes the hop across PCIe itself
 * (zero-copy: one launch per tick, no copy-engine round trips) -- also in crlot_stream_tick_host. */
int crlot_stream_submit_host(crlot_stream_t st, const float* h_chunk, long long chunk_stride, int chunk_len,
                             float* h_out, long long out_stride, int* n_out, unsigned long long* ticket);
int crlot_stream_wait_host(crlot_stream_t st, unsigned long long ticket);
/* final produce(nfft - hop) of the drive; *n_out = nfft - hop (0 if no frame was ever pushed) */
int crlot_stream_drain(crlot_stream_t st, float* d_out, long long out_stride, int* n_out, void* stream);
/* spectra of the most recent tick, device pointer to [n_streams][frames][nfft/2+1] complex, and
 * the frame count per stream: the reference's "processing goes here" gap (e2e_benchmark.cc:161-162) */
int crlot_stream_last_spectra(crlot_stream_t st, void** d_spec, int* frames_per_stream);
/* what = 0 samples ingested per stream, 1 frames processed per stream, 2 ingest ring capacity */
long long crlot_stream_get(crlot_stream_t st, int what);

/* ---- drop-in Framer state (device-resident sample store) --------------------------------------- */
/* replaces dsp::Framer   ref: dsp/frame/framer.h:25-131, dsp/frame/framer.cc:15-182
 * Samples are de-interleaved once on push into a planar [channel][capacity] device buffer; frame
 * extraction is then pure addressing inside kernel 1 (crlot_framer_stft).  pop() to a host pointer
 * is kept for source compatibility.  available/advance/clamp/ZERO_PAD/DROP rules are the
 * reference's (framer.cc:88-118,162-165). */
typedef struct crlot_framer_s* crlot_framer_t;
int crlot_framer_create(size_t frame_size, size_t hop_size, size_t channels, int boundary_mode, int device,
                        crlot_framer_t* out);                                   /* Framer() + set_params */
int crlot_framer_destroy(crlot_framer_t f);
int crlot_framer_reset(crlot_framer_t f);                                       /* Framer::reset         */
/* Framer::push: h_interleaved holds frames*channels floats ([frame][channel]) */
int crlot_framer_push(crlot_framer_t f, const float* h_interleaved, size_t frames);
/* same from device memory that is already planar: d_planar[c*chan_stride + (0..frames)) */
int crlot_framer_push_device(crlot_framer_t f, const float* d_planar, long long chan_stride, size_t frames,
                             void* stream);
size_t crlot_framer_available(crlot_framer_t f);                                /* available_frames()    */
/* Framer::pop: *popped = 1 and channels*frame_size interleaved floats in h_out_frame, or *popped = 0 */
int crlot_framer_pop(crlot_framer_t f, float* h_out_frame, int* popped);
/* Every frame that repeated pop() would deliver (at most max_frames; < 0 = all) goes through kernel
 * 1 of `plan` straight from the device store: d_spec [channels][*n_frames][nfft/2+1] complex.
 * The read cursor advances exactly as *n_frames pops would.
 * replaces the pop / window / forward loop of bench/e2e_benchmark.cc:152-158 */
int crlot_framer_stft(crlot_framer_t f, crlot_plan_t plan, long long max_frames, void* d_spec, long long* n_frames,
                      void* stream);
/* what = 0 frame_size, 1 hop_size, 2 channels, 3 boundary_mode, 4 buffer_size() as the reference
 * would report it, 5 read cursor, 6 write cursor (per-channel samples), 7 device capacity per
 * channel, 8 frames repeated pop() would deliver */
size_t crlot_framer_get(crlot_framer_t f, int what);
/* device view of the unread samples: returns &buf[0][read], *chan_stride floats between channels,
 * *valid samples per channel */
const float* crlot_framer_device_ptr(crlot_framer_t f, long long* chan_stride, long long* valid);

/* ---- drop-in FFT plan (host pointers, reference batch/stride layout) --------------------------- */
/* replaces dsp::fft::MakeFftPlan / IFftPlan::{forward,inverse} for FftDomain::Real
 *   ref: dsp/fft/api/fft_api.h:16-51; dsp/fft/backends/kissfft_adapter.cc:13-63,83-168
 * Argument rules and their failures (CRLOT_ERR_RUNTIME) mirror the adapter: batch 1..16,
 * strides >= 1, in_place unsupported, even nfft.  Every even nfft <= 16384 runs on the GPU (the
 * reference's own tests use 128 with strides 2/2, tests/fft_test.cc:450-520); larger sizes return
 * CRLOT_ERR_UNSUPPORTED. */
typedef struct crlot_fft_s* crlot_fft_t;
int crlot_fft_plan_create(int nfft, int in_place, int batch, int stride_in, int stride_out, int device,
                          crlot_fft_t* out);
int crlot_fft_plan_destroy(crlot_fft_t plan);
int crlot_fft_forward(crlot_fft_t plan, const float* h_in, float* h_out_ri, int batch);
int crlot_fft_inverse(crlot_fft_t plan, const float* h_in_ri, float* h_out, int batch);
/* Complex-domain plan: FftPlanDesc{Complex,...} + IFftPlan::forward_complex /
 * inverse_complex.  Same constructor checks and messages as the reference (batch 1..16, strides >= 1,
 * no in-place; no evenness requirement: tests/fft_test.cc:229-237 builds a 513-point plan).  Every size
 * 1..8192 runs on the GPU: 128..2048 (powers of two) on the register-FFT core, the others on the
 * mixed-radix any-size kernel; larger sizes return CRLOT_ERR_UNSUPPORTED.  A Real plan refuses the complex calls and vice versa with the
 * reference's messages.  Host arrays are interleaved (re,im); batch/stride addressing as in
 *   ref: dsp/fft/backends/kissfft_adapter.cc:171-246
 * crlot_fft_complex_device is the same transform over `rows` contiguous rows resident on the device. */
int crlot_fft_plan_create_complex(int nfft, int in_place, int batch, int stride_in, int stride_out, int device,
                                  crlot_fft_t* out);
int crlot_fft_forward_complex(crlot_fft_t plan, const float* h_in_ri, float* h_out_ri, int batch);
int crlot_fft_inverse_complex(crlot_fft_t plan, const float* h_in_ri, float* h_out_ri, int batch);
int crlot_fft_complex_device(crlot_fft_t plan, const void* d_in, void* d_out, long long rows, int inverse,
                             void* stream);

/* ---- drop-in OLAAccumulator state (device rings, host pointers at the calls) -------------------- */
/* replaces dsp::OLAAccumulator   ref: dsp/ola/OLAAccumulator.h:15-164, .cc:13-295 */
typedef struct crlot_ola_s* crlot_ola_t;
int crlot_ola_create(int sample_rate, size_t frame_size, size_t hop_size, size_t channels, float eps,
                     int apply_window_inside, int shadow_ring, int device, crlot_ola_t* out);
int crlot_ola_destroy(crlot_ola_t ola);
int crlot_ola_set_window(crlot_ola_t ola, const float* w, int wlen);
int crlot_ola_add_frame_soa(crlot_ola_t ola, const float* const* ch_frames, const float* window,
                            size_t start_sample, size_t start_off, size_t size, float gain);
int crlot_ola_push_frame_aos(crlot_ola_t ola, const float* interleaved, const float* window, size_t start_sample,
                             size_t start_off, size_t size, float gain);
int crlot_ola_produce(crlot_ola_t ola, float* const* ch_out, size_t n, size_t* produced);
int crlot_ola_flush(crlot_ola_t ola);
int crlot_ola_reset(crlot_ola_t ola);
/* what = 0 produced_samples, 1 read_pos, 2 ring_size, 3 has_window */
size_t crlot_ola_get(crlot_ola_t ola, int what);
float crlot_ola_meter_peak(crlot_ola_t ola);

/* ---- kernel-dispatch seam on device pointers (SURVEY 8b seam 2) ----------------------------------- */
/* replaces dsp::ola::axpy / axpy_windowed / normalize_and_clear   ref: dsp/ola/kernels.h:28-53,
 *   dsp/ola/kernels.cc:18-36,42-52 (the functions the reference forwards to kernels_hwy.cc)
 * Same element-wise arithmetic on device arrays of n floats, bit-exact with the scalar reference:
 *   axpy:                dst[i] = fma(src[i], g, dst[i])
 *   axpy_windowed:       dst[i] = fma(fl(src[i]*win[i]), g, dst[i])      (two roundings, kernels.cc:24-28)
 *   normalize_and_clear: out[i] = acc[i] / max(norm[i], eps); acc[i] = 0  (true division, kernels.cc:30-36) */
int crlot_axpy(float* d_dst, const float* d_src, float g, size_t n, int device, void* stream);
int crlot_axpy_windowed(float* d_dst, const float* d_src, const float* d_win, float g, size_t n, int device, void* stream);
int crlot_normalize_and_clear(float* d_out, float* d_acc, const float* d_norm, float eps, size_t n, int device,
                              void* stream);

/* ---- drop-in RingBuffer<T> (unified-memory cells, reference index arithmetic) --------------------- */
/* replaces dsp::ring::RingBuffer<T>   ref: dsp/ring/ring_buffer.h:25-110, dsp/ring/ring_buffer.cc:15-205
 * The (1 + shadow) * capacity cells of elem_bytes each live in CUDA unified memory: crlot_ring_data()
 * is valid on the host (the reference's data()[i] / Span element access keep working) AND on the device
 * (kernels index the same pointer; pages migrate to HBM on first device touch, crlot_ring_prefetch moves
 * them eagerly).  split() is the reference's arithmetic -- len == 0 -> two empty spans, overflow guard,
 * len = min(len, capacity), start %= capacity, one or two spans (ring_buffer.cc:44-85) -- returning
 * element OFFSETS from data().  write() appends at the internal write position, wraps and mirrors the
 * head like the reference (ring_buffer.cc:166-205; n is clamped to capacity, where the reference would
 * overrun).  shadow_sync(n) copies cells [0, min(n, capacity)) to [capacity, ...) (ring_buffer.cc:132-141);
 * contiguous_offset is contiguous_read_ptr's rule: shadow ? read_pos : read_pos % capacity (:144-163).
 * capacity == 0 fails with CRLOT_ERR_INVALID_ARGUMENT and the reference's message (ring_buffer.cc:21-23). */
typedef struct crlot_ring_s* crlot_ring_t;
typedef struct crlot_ring_span { size_t off1, len1, off2, len2; } crlot_ring_span;
int crlot_ring_create(size_t capacity, size_t elem_bytes, int shadow, int device, crlot_ring_t* out);
int crlot_ring_destroy(crlot_ring_t r);
void* crlot_ring_data(crlot_ring_t r);
size_t crlot_ring_capacity(crlot_ring_t r);
size_t crlot_ring_physical_capacity(crlot_ring_t r);
size_t crlot_ring_write_pos(crlot_ring_t r);
int crlot_ring_has_shadow(crlot_ring_t r);
int crlot_ring_split(crlot_ring_t r, size_t start, size_t len, crlot_ring_span* out);
int crlot_ring_shadow_sync(crlot_ring_t r, size_t n_elems);
size_t crlot_ring_contiguous_offset(crlot_ring_t r, size_t read_pos);
/* write(src, n): returns the elements written through *written (0 for a null source or n == 0) */
int crlot_ring_write(crlot_ring_t r, const void* h_src, size_t n, size_t* written);
int crlot_ring_write_device(crlot_ring_t r, const void* d_src, size_t n, size_t* written, void* stream);
/* migrate the cells to the device (to_device != 0) or back to the host ahead of use */
int crlot_ring_prefetch(crlot_ring_t r, int to_device, void* stream);

/* ---- multi-GPU partitioning (csrc/shard.cu) ------------------------------------------ */
/* Channels / streams never interact (ref: dsp/ola/OLAAccumulator.cc:76-111): contiguous blocks, no
 * data-path communication. */
int crlot_channel_shard(long long channels, int rank, int world, long long* first, long long* count);

/* One long signal split by OUTPUT/time range.  Rank r owns outputs [t0, t1) = [r*S, (r+1)*S) clipped to
 * the run's output length, S = total_len / world (a multiple of hop).  Frame k depends only on
 * x[k*hop, k*hop+nfft) (ref: dsp/frame/framer.cc:128-182), out[t] only on frames with k*hop in
 * (t-nfft, t] (ref: dsp/ola/OLAAccumulator.cc:75-111,194-209): the rank needs left_halo samples from
 * rank-1 and right_halo from rank+1 (each about nfft-hop) and recomputes the frames that start in its left
 * halo, so its outputs are bit-identical to the one-GPU run.  Its device buffer is laid out
 * [left_halo | owned S samples | right_halo] = input samples [in_lo, in_hi). */
typedef struct crlot_time_shard {
  int rank, world, nfft, hop, boundary_mode;
  long long total_len;
  long long t0, t1, n_out;            /* owned output range, n_out = t1 - t0                                  */
  long long k_lo, n_frames;           /* global index of local frame 0; frames computed locally              */
  long long n_owned_frames;           /* of those, the frames that start inside the owned input range        */
  long long in_lo, in_hi;             /* input samples needed                                                 */
  long long own_lo, own_hi;           /* input samples held / generated by this rank                         */
  long long left_halo, right_halo;    /* samples received from rank-1 / rank+1                               */
  long long send_left, send_right;    /* samples sent to rank-1 (its right halo) / rank+1 (its left halo)    */
  long long frame_t0;                 /* local output index of local frame 0's first sample (<= 0)           */
} crlot_time_shard;
int crlot_shard_plan(long long total_len, int nfft, int hop, int rank, int world, int boundary_mode,
                     crlot_time_shard* out);

/* One NCCL communicator per rank (one process per GPU, or one thread per GPU of one process).  NCCL is
 * bound at run time (libnccl.so.2; CRLOT_NCCL_LIB overrides the name); world == 1 needs none.
 *   multi-process: rank 0 calls crlot_shard_group_unique_id, ships the 128 bytes to the others by any
 *                  means (MPI, torch.distributed, a file), every rank calls crlot_shard_group_create;
 *   one process:   crlot_shard_group_create_all(n, devices, groups[n])  (ncclCommInitAll);
 *   existing comm: crlot_shard_group_wrap borrows a caller-owned ncclComm_t. */
#define CRLOT_NCCL_UNIQUE_ID_BYTES 128
typedef struct crlot_shard_group_s* crlot_shard_group_t;
int crlot_shard_group_unique_id(void* id_out_128);
int crlot_shard_group_create(int world, int rank, const void* id_128, int device, crlot_shard_group_t* out);
int crlot_shard_group_create_all(int n_dev, const int* devices, crlot_shard_group_t* out_groups);
int crlot_shard_group_wrap(void* nccl_comm, int world, int rank, int device, crlot_shard_group_t* out);
int crlot_shard_group_destroy(crlot_shard_group_t g);
int crlot_shard_group_info(crlot_shard_group_t g, int* world, int* rank, int* device);
/* Fills the halo regions of d_xbuf from the neighbours: ncclSend/ncclRecv pairs of send_* / *_halo floats
 * inside ONE ncclGroupStart/End, enqueued on `stream` (asynchronous; 12 KiB per direction at config 4). */
int crlot_halo_exchange(crlot_shard_group_t g, const crlot_time_shard* shard, float* d_xbuf, void* stream);
/* halo exchange + kernel 1 over the haloed buffer + (kernel 2, kernel 3 | fused 2+3 when d_frames is NULL)
 * at the shard's envelope phase (t0) and frame offset (frame_t0), all on `stream`.  d_spec
 * [n_frames][bins], d_frames [n_frames][nfft] or NULL, d_out [n_out].  group may be NULL when world == 1. */
int crlot_sharded_roundtrip_execute(crlot_plan_t plan, crlot_shard_group_t group, const crlot_time_shard* shard,
                                    float* d_xbuf, void* d_spec, float* d_frames, float* d_out, void* stream);

/* ---- round trip on static scratch: a call that allocates nothing --------------------------------- */
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * crlot_roundtrip_execute (d_frames = NULL semantics: K1 -> fused K2+K3, or K1 -> K2 -> K3 where the fused kernel
 * does not cover the plan) with every intermediate -- the kernels' work counter, the spectra and, on the unfused
 * route, the iFFT frames -- in ONE file-scope `static __device__` array of CRLOT_STATIC_SCRATCH_BYTES inside the
 * library (each GPU that loads the library holds it until exit).  The caller passes only the signal and the output;
 * the call itself performs no cudaMalloc / cudaFree / allocating library routine, and produces the same bits as
 * crlot_roundtrip_execute.  Limit: 256 + round256(F*channels*bins*8) [+ round256(F*channels*nfft*4) unfused]
 * bytes must fit, F = crlot_frame_count(len, ...) -- e.g. 4091 frames x channels at nfft 2048; beyond that the call
 * fails with CRLOT_ERR_INVALID_ARGUMENT and a message that states both figures.  Every call shares the array, which
 * keeps the previous call's contents (the work counter is cleared on `stream` before each kernel; everything else is
 * written before it is read): calls that may OVERLAP on the GPU -- different streams, different plans, different
 * host threads -- race and must be ordered by the caller.  The first call on a device resolves the array's device
 * address (cudaGetSymbolAddress); make that call before a graph capture begins. */
#define CRLOT_STATIC_SCRATCH_BYTES (32u << 20)
int crlot_roundtrip_execute_static(crlot_plan_t plan, const float* d_signal, long long chan_stride, long long len,
                                   int channels, float* d_out, long long out_chan_stride, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CRLOT_CUDA_H_ */
