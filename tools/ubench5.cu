// tools/ubench5.cu -- can PRMT (ALU pipe) + HFMA2.RELU (FMA pipe) really sustain 4 warp-instr/clk/SM?
// ubench.cu / ubench4.cu left the PRMT operands loop-invariant and ptxas hoisted the PRMTs out of
// the timed loop, so their "pair" figures were HFMA2-only rates counted twice. Here the table
// registers a,b change every 16 pairs (two cheap integer updates, counted), so nothing can be
// hoisted; the selector of chain i is a per-thread vector register as in the SSV kernel.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define ITER 1024
#define CH 16
#define UNR 4

// MODE 0: PRMT(a, b, e_i)   MODE 1: PRMT(a, a, e_i)   MODE 2: HFMA2 only   MODE 3: PRMT only (xor sink)
template <int MODE>
__global__ void k(uint32_t* out, uint32_t seed, long long* cycles, int reps) {
  uint32_t w[CH], e[CH];
  const uint32_t t = threadIdx.x * 2654435761u + seed;
#pragma unroll
  for (int i = 0; i < CH; i++) {
    w[i] = (t >> (i & 7)) & 0x03ff03ffu;
    e[i] = 0x3210u ^ ((t >> i) & 0x4444u);
  }
  uint32_t a = t * 3u, b = t * 7u;
  long long t0 = clock64();
  for (int rep = 0; rep < reps; rep++)
    for (int it = 0; it < ITER; it++) {
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        a += t;
        b ^= a;
#pragma unroll
        for (int i = 0; i < CH; i++) {
          uint32_t m;
          if (MODE == 0) {
            asm("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(a), "r"(b), "r"(e[i]));
            asm("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w[i]) : "r"(m), "r"(0x3c003c00u));
          } else if (MODE == 1) {
            asm("prmt.b32 %0, %1, %1, %2;" : "=r"(m) : "r"(a), "r"(e[i]));
            asm("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w[i]) : "r"(m), "r"(0x3c003c00u));
          } else if (MODE == 2) {
            asm("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w[i]) : "r"(a), "r"(0x3c003c00u));
          } else {
            asm("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(a), "r"(b), "r"(e[i]));
            w[i] ^= m;
          }
        }
      }
    }
  long long t1 = clock64();
  uint32_t acc = 0;
  for (int i = 0; i < CH; i++) acc ^= w[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

static int g_reps = 4;
template <int MODE>
void run(const char* name, int perChain, int warpsPerSM) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int threads = 32 * warpsPerSM;
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, sizeof(uint32_t) * sms * threads);
  cudaMalloc(&cyc, sizeof(long long) * sms);
  k<MODE><<<sms, threads>>>(out, 12345u, cyc, 1);
  k<MODE><<<sms, threads>>>(out, 12345u, cyc, g_reps);
  cudaDeviceSynchronize();
  long long c0;
  cudaMemcpy(&c0, cyc, sizeof c0, cudaMemcpyDeviceToHost);
  double warpInstr = (double)g_reps * ITER * UNR * (CH * perChain + 2) * warpsPerSM;
  printf("%-34s warps/SM=%2d  warp-instr/clk/SM=%.3f\n", name, warpsPerSM, warpInstr / c0);
  cudaFree(out);
  cudaFree(cyc);
}

int main(int argc, char** argv) {
  if (argc > 1) g_reps = atoi(argv[1]);
  for (int w : {4, 8, 16, 32}) {
    run<0>("PRMT(a,b,e_i)+HFMA2.RELU", 2, w);
    run<1>("PRMT(a,a,e_i)+HFMA2.RELU", 2, w);
    run<2>("HFMA2.RELU only", 1, w);
    run<3>("PRMT(a,b,e_i)+LOP3 (both ALU)", 2, w);
  }
  return 0;
}
