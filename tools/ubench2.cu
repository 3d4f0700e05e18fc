// tools/ubench2.cu -- the compiled 32-row block of the SSV fast path in isolation.
// Includes the kernel source itself and runs RowGroups<0>::run in a bare loop (no barriers, no
// hit checks, no refills), so the number it prints is the ceiling of the block code as ptxas
// scheduled it: cells per clock per SM, against the 128 of the issue roofline.
#include "../havac_b200/csrc/ssv_kernels.cu"

#include <cstdio>

namespace {

template <int MODE>
__global__ void __launch_bounds__(THREADS, 2) core_loop(uint32_t* out, int iters, long long* cycles) {
  extern __shared__ __align__(16) unsigned char smemRaw[];
  FastSmem& sm = *reinterpret_cast<FastSmem*>(smemRaw);
  const int tid = threadIdx.x;
  uint32_t* ringw = reinterpret_cast<uint32_t*>(&sm.ring[0][0]);
  for (int i = tid; i < 8 * NSLOT * 4; i += THREADS) {
    uint32_t h = (uint32_t)i * 2654435761u;
    uint32_t x0 = (h >> 7) & 3, x1 = (h >> 11) & 3;
    ringw[i] = (0x10u + 0x22u * x0) | ((0x10u + 0x22u * x1) << 8);
  }
  for (int i = tid; i < 2 * TC; i += THREADS) {
    // scores around -24 .. +18 times 256 as fp16 pairs (values do not matter for timing)
    sm.tab[0][i] = make_uint2(0xee00ee00u, 0x6c80ee00u);
  }
  __syncthreads();
  uint32_t w[R], e0[R], e1[R];
#pragma unroll
  for (int k = 0; k < R; k++) { w[k] = 0; e0[k] = ringw[(tid + k) & 1023]; e1[k] = ringw[(tid + 17 * k) & 1023]; }
  uint32_t g = (uint32_t)tid % (uint32_t)NSLOT;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
    RowGroups<0>::run(w, e0, e1, sm.tab[0] + (it & 15) * BLOCK_ROWS, &sm.ring[0][g]);
    g = (g + 1 == (uint32_t)NSLOT) ? 0u : g + 1;
    if (MODE == 1 && (it & 15) == 15) __syncthreads();  // one CTA barrier per 512 rows, like the kernel
  }
  long long t1 = clock64();
  uint32_t acc = 0;
#pragma unroll
  for (int k = 0; k < R; k++) acc ^= w[k] ^ e0[k] ^ e1[k];
  out[blockIdx.x * THREADS + tid] = acc;
  if (tid == 0) {
    unsigned long long ns;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    cycles[blockIdx.x] = t1 - t0;
    cycles[gridDim.x + blockIdx.x] = (long long)ns;      // end time
    cycles[2 * gridDim.x + blockIdx.x] = (long long)smid;
  }
}

template <int MODE>
void run(const char* name, int ctasPerSM) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaFuncSetAttribute(core_loop<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FastSmem));
  cudaFuncSetAttribute(core_loop<MODE>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  int grid = sms * ctasPerSM, iters = 4000;
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, sizeof(uint32_t) * grid * THREADS);
  cudaMalloc(&cyc, sizeof(long long) * grid * 3);
  core_loop<MODE><<<grid, THREADS, sizeof(FastSmem)>>>(out, 100, cyc);
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  core_loop<MODE><<<grid, THREADS, sizeof(FastSmem)>>>(out, iters, cyc);
  cudaEventRecord(b);
  cudaError_t e = cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  long long c0 = 0;
  cudaMemcpy(&c0, cyc, sizeof c0, cudaMemcpyDeviceToHost);
  {
    long long* h = new long long[grid * 3];
    cudaMemcpy(h, cyc, sizeof(long long) * grid * 3, cudaMemcpyDeviceToHost);
    long long cmin = h[0], cmax = h[0], emin = h[grid], emax = h[grid];
    int perSm[256] = {0};
    for (int i = 0; i < grid; i++) {
      cmin = h[i] < cmin ? h[i] : cmin; cmax = h[i] > cmax ? h[i] : cmax;
      emin = h[grid + i] < emin ? h[grid + i] : emin; emax = h[grid + i] > emax ? h[grid + i] : emax;
      perSm[h[2 * grid + i] & 255]++;
    }
    int hist[8] = {0};
    for (int i = 0; i < 256; i++) hist[perSm[i] > 7 ? 7 : perSm[i]]++;
    printf("    per-CTA cycles min %lld max %lld; end-time spread %.3f ms; SMs by CTA count: 0:%d 1:%d 2:%d 3:%d 4+:%d\n", cmin, cmax,
           (emax - emin) / 1e6, hist[0], hist[1], hist[2], hist[3], hist[4] + hist[5] + hist[6] + hist[7]);
    delete[] h;
  }
  double cells = (double)iters * BLOCK_ROWS * (2.0 * R) * THREADS * ctasPerSM;  // per SM
  printf("%-22s ctas/SM=%d  %s  cycles=%lld  cells/clk/SM=%.1f (%.1f%% of 128)  time=%.2f ms  -> %.1f GCUPS on %d SMs\n", name,
         ctasPerSM, cudaGetErrorString(e), c0, cells / c0, 100.0 * cells / c0 / 128.0, ms, cells * sms / ms / 1e6, sms);
  cudaFree(out); cudaFree(cyc);
}

}  // namespace

int main() {
  for (int c : {1, 2}) {
    run<0>("block loop", c);
    run<1>("block loop + barrier", c);
  }
  return 0;
}
