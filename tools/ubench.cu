// tools/ubench.cu -- issue-rate microbenchmarks for the instructions the SSV kernel is made of.
// Prints warp-instructions per clock per SM for PRMT-only, HFMA2-only, the PRMT+HFMA2 pair,
// and a few alternatives (IADD3, LOP3, VIADDMNMX, IMAD). This is how the integer-ALU / issue
// roofline denominator in DESIGN.md was measured ("INT32 lane count of
// the B200 SM is not in MEASURED_PEAKS.json").
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define ITER 4096
#define CHAINS 16

template <int MODE>
__global__ void k(uint32_t* out, uint32_t seed, long long* cycles, int reps) {
  uint32_t w[CHAINS], s[CHAINS];
  for (int i = 0; i < CHAINS; i++) { w[i] = seed * (i + 1) + threadIdx.x; s[i] = 0x3210 + ((seed + i) & 0x1111); }
  uint32_t a = seed ^ 0x12345678u, b = seed * 7u;
  long long t0 = clock64();
  for (int rep = 0; rep < reps; rep++)
  for (int it = 0; it < ITER; it++) {
#pragma unroll
    for (int i = 0; i < CHAINS; i++) {
      if (MODE == 0) {  // PRMT only (dependent chain per register)
        asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(w[i]) : "r"(b), "r"(s[i]));
      } else if (MODE == 1) {  // HFMA2.RELU only
        asm volatile("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w[i]) : "r"(a), "r"(0x3c003c00u));
      } else if (MODE == 2) {  // the kernel's pair: PRMT feeding HFMA2.RELU
        uint32_t m;
        asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(a), "r"(b), "r"(s[i]));
        asm volatile("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w[i]) : "r"(m), "r"(0x3c003c00u));
      } else if (MODE == 3) {  // IADD3
        asm volatile("add.u32 %0, %0, %1;" : "+r"(w[i]) : "r"(a));
      } else if (MODE == 4) {  // LOP3
        asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(w[i]) : "r"(a), "r"(b));
      } else if (MODE == 5) {  // VIADDMNMX.S16x2.RELU
        w[i] = __viaddmax_s16x2_relu(w[i], a, b);
      } else if (MODE == 6) {  // IMAD
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(w[i]) : "r"(a), "r"(b));
      } else if (MODE == 7) {  // PRMT + VIADDMNMX pair (all-integer alternative)
        uint32_t m;
        asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(a), "r"(b), "r"(s[i]));
        w[i] = __viaddmax_s16x2_relu(w[i], m, 0u);
      } else if (MODE == 8) {  // HADD2 (no relu) for comparison
        asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(w[i]) : "r"(a));
      }
    }
  }
  long long t1 = clock64();
  uint32_t acc = 0;
  for (int i = 0; i < CHAINS; i++) acc ^= w[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

static int g_reps = 1;
template <int MODE>
void run(const char* name, int instrPerChain, int warpsPerSM) {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int threads = 32 * warpsPerSM;
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, sizeof(uint32_t) * sms * threads);
  cudaMalloc(&cyc, sizeof(long long) * sms);
  k<MODE><<<sms, threads>>>(out, 12345u, cyc, g_reps);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<sms, threads>>>(out, 12345u, cyc, g_reps);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long c0; cudaMemcpy(&c0, cyc, sizeof c0, cudaMemcpyDeviceToHost);
  double warpInstr = (double)g_reps * ITER * CHAINS * instrPerChain * warpsPerSM;
  printf("%-28s warps/SM=%2d  cycles=%9lld  warp-instr/clk/SM=%.3f  (lanes/clk/SM=%.1f)  time=%.3f ms  implied MHz=%.0f\n",
         name, warpsPerSM, c0, warpInstr / c0, 32 * warpInstr / c0, ms, c0 / (ms * 1e3));
  cudaFree(out); cudaFree(cyc);
}

int main(int argc, char** argv) {
  if (argc > 1) g_reps = atoi(argv[1]);  // >1: long runs that show the sustained (power-managed) clock
  for (int w : {4, 8, 16, 32}) {
    run<0>("PRMT", 1, w);
    run<1>("HFMA2.RELU", 1, w);
    run<2>("PRMT+HFMA2.RELU", 2, w);
    run<3>("IADD", 1, w);
    run<4>("LOP3", 1, w);
    run<5>("VIADDMNMX.S16x2.RELU", 1, w);
    run<6>("IMAD", 1, w);
    run<7>("PRMT+VIADDMNMX", 2, w);
    run<8>("HADD2", 1, w);
  }
  return 0;
}
