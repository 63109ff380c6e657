// kernels.h -- internal (C++) launch interface between the C-ABI layer (capi.cu) and the three
// hot-path kernels + helpers.  Not installed; the public boundary is include/crlot_cuda.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "regfft.cuh"

namespace crlot {

// ---- kernel 1: fused frame extraction + analysis window + forward real FFT ---------------------
// Frame k of channel c = samples sig[c*chan_stride + k*hop + (0..nfft)), zero beyond `len`
// (ZERO_PAD tail; DROP callers simply pass a frame count that never reaches past len).
// spec row (c*spec_chan_rows + k) holds nfft/2+1 interleaved (re,im) bins.
struct K1Params {
  const float* sig;
  long long chan_stride;     // floats between channels (multiple of 4)
  long long len;             // valid samples per channel
  int pad, pad_mode;         // FrameQueue front end: frames index the stream [pad | signal | pad] (0, 0 = Framer)
  int channels;
  long long frames;          // frames per channel to compute
  int nfft, hop;
  cpx* spec;
  long long spec_chan_rows;  // rows between channels in spec
  const float* window;       // device, nfft floats (all ones = no window)
  const cpx* tw1;            // device tables (tables.h)
  const cpx* sa;             // real-split table A_k (tables.h)
  const cpx* pre;            // nfft == 4096 only
  unsigned int* counter;     // zeroed work counter (dynamic chunk scheduling)
  int chunk_steps;           // steps (of G frames) per scheduled chunk
  int scrub;                 // 0 none, 1 reference clean-up on every windowed sample, 2 same via screening
  unsigned int tiny2;        // scrub == 2: 2*bits(tiny_thr) - 2 (see k1_stft.cu)
  int generic;               // 1: any-size kernel (k_generic.cu); tw1/sa are then make_generic_tables' (tables.h)
};
cudaError_t launch_stft_fwd(const K1Params& p, cudaStream_t st);
cudaError_t launch_stft_fwd_generic(const K1Params& p, cudaStream_t st);   // k_generic.cu, called by launch_stft_fwd
cudaError_t launch_stft_fwd_4096v2(const K1Params& p, cudaStream_t st);    // k1_4096v2.cu: nfft 4096 as two 1024-point sub-transforms
bool stft_fwd_2048tm_supported(const K1Params& p);                         // k1_2048tm.cu: nfft 2048 / hop 512, raw samples in tensor memory, 16 warps
cudaError_t launch_stft_fwd_2048tm(const K1Params& p, cudaStream_t st);
bool generic_len_supported(int complex_len);

// ---- kernel 2: inverse real FFT * (1/nfft) + clean-up (+ synthesis window) --------------------
struct K2Params {
  const cpx* spec;           // rows of nfft/2+1 bins, `rows` of them, contiguous
  long long rows;
  int nfft;
  float* frames;             // rows of nfft floats
  const float* window;       // device, nfft floats, or nullptr: no synthesis window here
  const cpx* tw1;
  const cpx* sa;
  const cpx* pre;
  int scrub;
  // spectral-domain hook: X[row][k] *= sgain[row*sgain_row_stride + k] (real gain, one
  // rounding per component) as the row is loaded; nullptr = none; stride 0 = the same gains for every row
  const float* sgain;
  long long sgain_row_stride;
  int generic;               // as in K1Params
};
cudaError_t launch_istft_inv(const K2Params& p, cudaStream_t st);
cudaError_t launch_istft_inv_generic(const K2Params& p, cudaStream_t st);
cudaError_t launch_istft_inv_4096v2(const K2Params& p, cudaStream_t st);   // k2_4096v2.cu: nfft 4096 as two 1024-point sub-transforms

// ---- complex-domain plan: rows of n complex points, n in {128, 256, 512, 1024, 2048} --------------
struct CfftParams {
  const cpx* in;             // [rows][n]
  cpx* out;                  // [rows][n]
  long long rows;
  int n;
  int inverse;               // 1: conjugate kernel, times 1/n, clean-up rule on both components
  const cpx* tw1;            // inter-pass table of the 2n-point real geometry (tables.h)
  int generic;               // 1: any n (k_generic.cu); tw1 is then exp(-2 pi i k / n), k < n
};
cudaError_t launch_cfft(const CfftParams& p, cudaStream_t st);
cudaError_t launch_cfft_generic(const CfftParams& p, cudaStream_t st);

// ---- kernel 3: gather overlap-add + envelope normalisation ------------------------------------
// out[c][j] = ( sum_k' fma( f(k',u) , gain, acc ) ) / max(norm[(t_offset + j) % ring_len], eps),
//   u = j - frame_t0 - k'*hop in [0,nfft), k' ascending in [0,frames),
//   f = frames[c][k'][u] * window[u] rounded to fp32 when window != nullptr, else frames[c][k'][u].
struct K3Params {
  const float* frames;
  long long frames_chan_rows;   // rows between channels
  long long frames_per_chan;
  int channels;
  int nfft, hop;
  const float* window;          // device or nullptr (frames already carry the synthesis window)
  float gain, eps;
  const float* norm;            // device envelope, ring_len floats
  long long ring_len;
  long long t_offset;           // absolute sample index of out[.][0] (envelope phase)
  long long frame_t0;           // local output index of the first sample of local frame 0 (may be < 0)
  float* out;
  long long out_chan_stride;
  long long n_out;              // samples per channel to produce
};
cudaError_t launch_ola_gather(const K3Params& p, cudaStream_t st);

// ---- kernels 2+3 fused: inverse real FFT + synthesis window + overlap-add + normalise ------------
// Same inputs as kernel 2 (spectra) and same outputs as kernel 3 (see K3Params for the meaning of
// t_offset / frame_t0 / n_out); the iFFT frames stay in shared memory.  Bit-identical to K2 -> K3.
struct K23Params {
  const cpx* spec;              // [channels][spec_chan_rows][nfft/2+1]
  long long spec_chan_rows;
  long long frames_per_chan;
  int channels;
  int nfft, hop;
  const float* window;          // device synthesis window or nullptr
  const cpx* tw1;
  const cpx* sa;
  const cpx* pre;               // nfft == 4096: transposed inverse inter-pass table (tables.h)
  int scrub;
  float gain, eps;
  const float* norm;
  const float2* den_rcp;        // [ring_len] {max(norm, eps), refined reciprocal} or nullptr (common.cuh)
  long long ring_len;
  long long t_offset, frame_t0;
  float* out;
  long long out_chan_stride;
  long long n_out;
  unsigned int* counter;        // zeroed work counter
  int chunk_blocks;             // output hop blocks per scheduled chunk (0 = choose)
  const float* sgain;           // spectral-domain hook, as in K2Params; rows indexed like `spec`
  long long sgain_row_stride;
  long long m_begin, m_end;     // filled by the launcher: hop blocks [m_begin, m_end) cover [0, n_out)
  // fully fused round trip (sig != nullptr): the spectrum rows are not read from `spec` but computed in the same
  // warp from the signal -- frame k of channel c = sig[c*sig_chan_stride + k*hop + (0..nfft)), zero beyond sig_len,
  // times awindow, forward real FFT with kernel 1's arithmetic -- and never leave the registers.
  const float* sig;
  long long sig_chan_stride, sig_len;
  const float* awindow;         // device analysis window (all ones = none)
  int k1_scrub;                 // kernel 1's clean-up mode (0 none, 1 every sample, 2 screening)
  unsigned int tiny2;           // screening threshold of kernel 1
};
// nfft = 2048 with hop = 512 (the layout of the forward output registers equals the layout the inverse wants, and
// the raw-sample carry-over in tensor memory wants nfft == 4*hop), otherwise the conditions of istft_ola_supported
bool roundtrip_fused_supported(int nfft, int hop, long long ring_len, long long t_offset, long long frame_t0,
                               long long out_chan_stride, const void* out, const void* norm, const void* sig,
                               long long sig_chan_stride);
bool istft_ola_supported(int nfft, int hop, long long ring_len, long long t_offset, long long frame_t0,
                         long long out_chan_stride, const void* out, const void* norm);
cudaError_t launch_istft_ola(K23Params p, cudaStream_t st);
cudaError_t launch_istft_ola_4096(const K23Params& p, cudaStream_t st);   // k23_4096.cu, called by launch_istft_ola
bool istft_ola_2048tm_supported(const K23Params& p);                      // k23_2048tm.cu: nfft 2048 / hop 512, accumulator ring in tensor memory, 16 warps
cudaError_t launch_istft_ola_2048tm(const K23Params& p, cudaStream_t st);
bool roundtrip_2048tm_supported(const K23Params& p);                      // k_rt_2048tm.cu: the one-kernel round trip, both rings in tensor memory, 16 warps
cudaError_t launch_roundtrip_2048tm(const K23Params& p, cudaStream_t st);

// in-place spectral hook as its own pass: spec[row][k] *= gain[row*row_stride + k]
cudaError_t launch_spectral_gain(cpx* spec, long long rows, int bins, const float* gain, long long row_stride,
                                 cudaStream_t st);

// den_rcp[i] = {max(norm[i], eps), div_refined_rcp(max(norm[i], eps))}
cudaError_t launch_build_den_rcp(const float* norm, float eps, long long n, float2* den_rcp, cudaStream_t st);
// compares div_tail with __fdiv_rn on `n` hashed operand pairs inside the tail's operand range plus
// the plan's own denominators; *mismatches (device) is incremented per differing result
cudaError_t launch_division_selftest(const float2* den_rcp, long long ring_len, unsigned long long n,
                                     unsigned long long seed, unsigned long long* mismatches, cudaStream_t st);

// FrameQueue::getAllFrames on the device: frames[k][i] = padded[k*hop + i] (bit-exact gather)
cudaError_t launch_framequeue_extract(const float* sig, long long len, int frame, int hop, int pad, int pad_mode,
                                      long long n_frames, float* frames, cudaStream_t st);

// interleaved [n][channels] <-> planar [channels][chan_stride] copies (bit-exact)
cudaError_t launch_deinterleave(const float* aos, long long n, int channels, float* planar, long long chan_stride,
                                cudaStream_t st);
cudaError_t launch_interleave(const float* planar, long long chan_stride, long long n, int channels, float* aos,
                              cudaStream_t st);

// ---- helpers ----------------------------------------------------------------------------------
// x[c][i] = uniform[-1,1) from hash(seed, channel0 + c, i0 + i); same function as oracle/crlot_oracle.c
cudaError_t launch_synth_uniform(float* out, long long chan_stride, int channels, long long n,
                                 unsigned long long seed, unsigned long long channel0,
                                 unsigned long long i0, cudaStream_t st);
// device-resident OLAAccumulator ring ops (facade + streaming path)
cudaError_t launch_ring_axpy(float* ring, long long ring_len, int channels, const float* frames,
                             long long frame_chan_stride, const float* window, float gain,
                             long long start, int n, cudaStream_t st);
cudaError_t launch_ring_produce(float* ring, long long ring_len, int channels, const float* norm,
                                float eps, long long read_pos, int n, float* out,
                                long long out_chan_stride, cudaStream_t st);

// ---- streaming (config 5): device-resident per-stream state, one tick = 4 launches --------------
struct StreamCounters {
  long long w;   // samples ingested per stream so far
  long long k;   // next frame index (== frames processed); read position of the OLA ring = k*hop
};
cudaError_t launch_stream_ingest(const StreamCounters* ctr, const float* chunk, long long chunk_stride, int chunk_len,
                                 float* ring, int rcap, int n_streams, long long w0, cudaStream_t st);
struct StreamFwdParams {
  const StreamCounters* ctr;
  const float* ring;         // [n_streams][rcap] ingest rings (rcap a power of two)
  int rcap, n_streams, nf;   // nf = frames per stream this tick
  int nfft, hop;
  cpx* spec;                 // [n_streams][nf][nfft/2+1]
  const float* window;
  const cpx* tw1;
  const cpx* sa;
  int scrub;
};
cudaError_t launch_stream_fwd(const StreamFwdParams& p, cudaStream_t st);
bool stream_fwd_supported(int nfft);   // frame sizes the ring-reading forward kernel covers (256..2048)
// the other sizes: frames gathered from the rings into rows [n_streams*nf][nfft], then kernel 1 with hop = nfft
cudaError_t launch_stream_gather(const StreamCounters* ctr, const float* ring, int rcap, int n_streams, int nf, int nfft,
                                 int hop, float* rows, cudaStream_t st);
struct StreamOlaParams {
  const StreamCounters* ctr;
  const float* frames;       // [n_streams][nf][nfft]
  int n_streams, nf, nfft, hop;
  int span;                  // ring positions touched per stream: nfft + (nf-1)*hop, or n_prod when nf == 0
  int n_prod;                // samples produced per stream (nf*hop, or nfft-hop for the final drain)
  const float* window;       // nullptr: frames already carry the synthesis window
  float gain, eps;
  const float* norm;
  long long ring_len;
  float* acc;                // [n_streams][ring_len] accumulator rings
  float* out;                // [n_streams][out_stride]; this tick writes at offset (k - k_base)*hop
  long long out_stride;
  long long k_base;          // frame counter value that maps to out[.][0]
};
cudaError_t launch_stream_ola(const StreamOlaParams& p, cudaStream_t st);
cudaError_t launch_stream_advance(StreamCounters* ctr, int chunk_len, int nf, cudaStream_t st);
// the whole tick in one launch when it yields exactly one frame per stream (chunk_len == hop)
struct StreamTick1Params {
  StreamCounters* ctr;       // read at the start, advanced by the last block to finish
  unsigned int* done;        // zero-initialised block ticket
  const float* chunk;        // chunk of this tick starts at chunk[s*chunk_stride + (w - w0)]
  long long chunk_stride, w0;
  int chunk_len;
  float* ring;               // [n_streams][rcap] ingest rings
  int rcap, n_streams, nfft, hop;
  cpx* spec;                 // [n_streams][nfft/2+1]
  const float* window;       // analysis window
  const float* syn_window;   // synthesis window or nullptr
  const cpx* tw1;
  int scrub;
  float gain, eps;
  const float* norm;
  long long ring_len;
  float* acc;                // [n_streams][ring_len]
  float* out;
  long long out_stride, k_base;
  int pdl;                   // 1: launch with programmatic stream serialization (prologue overlaps the previous tick)
};
bool stream_tick1_supported(int nfft, int hop, int chunk_len, long long ring_len, long long out_stride, const void* out);
cudaError_t launch_stream_tick1(const StreamTick1Params& p, cudaStream_t st);

int device_sm_count();

}  // namespace crlot
