// tmem_probe.cu -- microbenchmark: can tensor memory (TMEM) serve as lane-private scratch for the FFT kernels?
// Measures, per SM, (a) tcgen05.ld.32x32b throughput with 12 warps per CTA, (b) LDS.64 throughput alone,
// (c) both together (do they share a pipe?), (d) tcgen05.st throughput.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_probe tools/tmem_probe.cu && ./tmem_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void tmem_ld16(uint32_t addr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(addr));
}
__device__ __forceinline__ void tmem_st16(uint32_t addr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};\n" ::"r"(addr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]));
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }

// MODE bit 0: tcgen05.ld, bit 1: LDS.64, bit 2: tcgen05.st
template <int MODE>
__global__ void __launch_bounds__(384, 1) probe(int iters, unsigned long long* out, long long* cycles) {
  __shared__ uint32_t s_addr;
  extern __shared__ __align__(16) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"r"(
        (uint32_t)__cvta_generic_to_shared(&s_addr)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = i;
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  const uint32_t base = s_addr;
  // warp w may touch TMEM lanes 32*(w%4) .. +31; give each of the 3 warps of a quarter its own 160 columns
  const uint32_t taddr = base + (static_cast<uint32_t>((warp & 3) * 32) << 16) + static_cast<uint32_t>((warp >> 2) * 160);
  uint32_t r[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) r[i] = lane * 16 + i;
  for (int c = 0; c < 160; c += 16) tmem_st16(taddr + c, r);
  tmem_wait_st();
  __syncthreads();
  unsigned long long acc = 0;
  const uint2* sp = reinterpret_cast<const uint2*>(smem) + warp * 512 + lane;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE & 1) {
#pragma unroll
      for (int c = 0; c < 128; c += 32) {      // two loads in flight per wait (32 registers), as a kernel would
        uint32_t q[16];
        tmem_ld16(taddr + c, r);
        tmem_ld16(taddr + c + 16, q);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) acc += r[i] ^ q[i];
      }
    }
    if (MODE & 2) {
#pragma unroll
      for (int c = 0; c < 64; ++c) {       // 64 x LDS.64 per lane = 16 KB per warp-iteration
        const uint2 v = sp[((c * 8 + it) & 511) * 1];
        acc += v.x ^ v.y;
      }
    }
    if (MODE & 4) {
#pragma unroll
      for (int c = 0; c < 128; c += 16) {
        r[0] += it;
        tmem_st16(taddr + c, r);
      }
      tmem_wait_st();
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" ::"r"(base));
}

template <int MODE> void run(const char* name, int iters) {
  unsigned long long* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 384 * 8);
  cudaMalloc(&cyc, 148 * 8);
  cudaFuncSetAttribute(probe<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  probe<MODE><<<148, 384, 64 * 1024>>>(10, out, cyc);
  cudaDeviceSynchronize();
  probe<MODE><<<148, 384, 64 * 1024>>>(iters, out, cyc);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double c = 0;
  for (int i = 0; i < 148; ++i) c += h[i];
  c /= 148;
  const double tm_bytes = (MODE & 1 ? 12.0 * 128 * 32 * 4 : 0) + (MODE & 4 ? 12.0 * 128 * 32 * 4 : 0);   // per CTA-iteration
  const double ls_bytes = (MODE & 2 ? 12.0 * 64 * 32 * 8 : 0);
  std::printf("%-28s %s  cycles/iter %.1f   TMEM %.1f B/clk/SM   LDS %.1f B/clk/SM\n", name, cudaGetErrorString(e), c / iters,
              tm_bytes * iters / c, ls_bytes * iters / c);
  cudaFree(out);
  cudaFree(cyc);
}

int main() {
  run<1>("tcgen05.ld only", 2000);
  run<2>("LDS.64 only", 2000);
  run<3>("tcgen05.ld + LDS.64", 2000);
  run<4>("tcgen05.st only", 2000);
  run<7>("ld + LDS + st", 2000);
  return 0;
}
