// tools/ubench8.cu -- which half of the PRMT + HFMA2.RELU pair holds it at ~2.9 of 4 issue slots?
// X (producer, ALU class) x Y (consumer, FMA class) matrix, 16 independent chains, the two shared
// producer operands change every 16 pairs so nothing is loop-invariant (cf. ubench5).
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define ITER 1024
#define CH 16
#define UNR 4

template <int X>
__device__ __forceinline__ uint32_t produce(uint32_t a, uint32_t b, uint32_t e) {
  uint32_t m;
  if (X == 0) asm("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(a), "r"(b), "r"(e));
  else if (X == 1) asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(m) : "r"(a), "r"(b), "r"(e));
  else if (X == 2) asm("{.reg .u32 q; add.u32 q, %1, %2; add.u32 %0, q, %3;}" : "=r"(m) : "r"(a), "r"(b), "r"(e));
  else asm("max.s32 %0, %1, %2;" : "=r"(m) : "r"(a), "r"(e));
  return m;
}
template <int Y>
__device__ __forceinline__ void consume(uint32_t& w, uint32_t m) {
  if (Y == 0) asm("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w) : "r"(m), "r"(0x3c003c00u));
  else if (Y == 1) asm("fma.rn.f16x2 %0, %1, %2, %0;" : "+r"(w) : "r"(m), "r"(0x3c003c00u));
  else if (Y == 2) asm("add.rn.f16x2 %0, %1, %0;" : "+r"(w) : "r"(m));
  else if (Y == 3) asm("fma.rn.f32 %0, %1, %2, %0;" : "+f"(*(float*)&w) : "f"(__uint_as_float(m)), "f"(1.0f));
  else if (Y == 4) asm("mad.lo.u32 %0, %1, 3, %0;" : "+r"(w) : "r"(m));
  else asm("add.u32 %0, %1, %0;" : "+r"(w) : "r"(m));
}

template <int X, int Y>
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
        for (int i = 0; i < CH; i++) consume<Y>(w[i], produce<X>(a, b, e[i]));
      }
    }
  long long t1 = clock64();
  uint32_t acc = 0;
  for (int i = 0; i < CH; i++) acc ^= w[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

static int g_reps = 4;
static const char* XN[] = {"PRMT", "LOP3", "IADD3", "VIMNMX"};
static const char* YN[] = {"HFMA2.RELU", "HFMA2", "HADD2", "FFMA", "IMAD", "IADD"};
template <int X, int Y>
void run(int warpsPerSM) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int threads = 32 * warpsPerSM;
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, sizeof(uint32_t) * sms * threads);
  cudaMalloc(&cyc, sizeof(long long) * sms);
  k<X, Y><<<sms, threads>>>(out, 12345u, cyc, 1);
  k<X, Y><<<sms, threads>>>(out, 12345u, cyc, g_reps);
  cudaDeviceSynchronize();
  long long c0;
  cudaMemcpy(&c0, cyc, sizeof c0, cudaMemcpyDeviceToHost);
  double warpInstr = (double)g_reps * ITER * UNR * (CH * 2 + 2) * warpsPerSM;
  printf("%-7s + %-10s warps/SM=%2d  warp-instr/clk/SM=%.3f\n", XN[X], YN[Y], warpsPerSM, warpInstr / c0);
  cudaFree(out);
  cudaFree(cyc);
}
template <int X>
void row(int w) {
  run<X, 0>(w); run<X, 1>(w); run<X, 2>(w); run<X, 3>(w); run<X, 4>(w); run<X, 5>(w);
}
int main(int argc, char** argv) {
  if (argc > 1) g_reps = atoi(argv[1]);
  row<0>(16); row<1>(16); row<2>(16); row<3>(16);
  return 0;
}
