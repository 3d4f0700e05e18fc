// tools/ubench9.cu -- does a PRMT whose table operand(s) live in UNIFORM registers (no vector
// register-file read) lift the PRMT + HFMA2.RELU pair above ~2.9 of 4 issue slots?
// The uniform operands are functions of a kernel parameter and the loop counter (so ptxas keeps
// them on the uniform datapath) and change every 16 pairs (so nothing is hoisted).
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define ITER 1024
#define CH 16
#define UNR 4

// MODE 0: a vec, b vec    MODE 1: a vec, b uniform    MODE 2: a uniform, b uniform    MODE 3: a uniform, b vec
template <int MODE>
__global__ void k(uint32_t* out, uint32_t seed, long long* cycles, int reps, uint32_t ua, uint32_t ub) {
  uint32_t w[CH], e[CH];
  const uint32_t t = threadIdx.x * 2654435761u + seed;
#pragma unroll
  for (int i = 0; i < CH; i++) {
    w[i] = (t >> (i & 7)) & 0x03ff03ffu;
    e[i] = 0x3210u ^ ((t >> i) & 0x4444u);
  }
  uint32_t av = t * 3u, bv = t * 7u;
  long long t0 = clock64();
  for (int rep = 0; rep < reps; rep++)
    for (int it = 0; it < ITER; it++) {
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        const uint32_t row = (uint32_t)(it * UNR + u);
        const uint32_t au = ua + row * 0x01010101u, bu = ub ^ (row * 0x00010203u);  // uniform
        av += t;
        bv ^= av;
        const uint32_t a = (MODE == 2 || MODE == 3) ? au : av;
        const uint32_t b = (MODE == 1 || MODE == 2) ? bu : bv;
#pragma unroll
        for (int i = 0; i < CH; i++) {
          uint32_t m;
          asm("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(a), "r"(b), "r"(e[i]));
          asm("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w[i]) : "r"(m), "r"(0x3c003c00u));
        }
      }
    }
  long long t1 = clock64();
  uint32_t acc = av ^ bv;
  for (int i = 0; i < CH; i++) acc ^= w[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

static int g_reps = 4;
template <int MODE>
void run(const char* name, int warpsPerSM) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int threads = 32 * warpsPerSM;
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, sizeof(uint32_t) * sms * threads);
  cudaMalloc(&cyc, sizeof(long long) * sms);
  k<MODE><<<sms, threads>>>(out, 12345u, cyc, 1, 0x11223344u, 0x55667788u);
  k<MODE><<<sms, threads>>>(out, 12345u, cyc, g_reps, 0x11223344u, 0x55667788u);
  cudaDeviceSynchronize();
  long long c0;
  cudaMemcpy(&c0, cyc, sizeof c0, cudaMemcpyDeviceToHost);
  double pairs = (double)g_reps * ITER * UNR * CH * warpsPerSM;
  printf("%-34s warps/SM=%2d  pairs/clk/SM=%.3f  (pair instr/clk/SM=%.3f)\n", name, warpsPerSM, pairs / c0, 2 * pairs / c0);
  cudaFree(out);
  cudaFree(cyc);
}

int main(int argc, char** argv) {
  if (argc > 1) g_reps = atoi(argv[1]);
  for (int w : {8, 16}) {
    run<0>("PRMT(a.vec, b.vec)", w);
    run<1>("PRMT(a.vec, b.uniform)", w);
    run<2>("PRMT(a.uniform, b.uniform)", w);
    run<3>("PRMT(a.uniform, b.vec)", w);
  }
  return 0;
}
