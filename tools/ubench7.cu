// tools/ubench7.cu -- would moving a share of the emission lookups from PRMT (ALU pipe) to a
// shared-memory lookup (LSU pipe, LDS.32 of a 16-entry pair table per row) lift the pair rate?
// NL of the 16 chains per row use LDS + HFMA2, the others PRMT + HFMA2.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define ITER 512
#define CH 16
#define UNR 4

template <int NL>
__global__ void k(uint32_t* out, uint32_t seed, long long* cycles, int reps) {
  __shared__ uint32_t lut[64 * 16];  // 64 rows x 16 pair codes
  for (int i = threadIdx.x; i < 64 * 16; i += blockDim.x) lut[i] = (seed * i) & 0x03ff03ffu;
  __syncthreads();
  uint32_t w[CH], e[CH], ad[CH];
  const uint32_t t = threadIdx.x * 2654435761u + seed;
#pragma unroll
  for (int i = 0; i < CH; i++) {
    w[i] = (t >> (i & 7)) & 0x03ff03ffu;
    e[i] = 0x3210u ^ ((t >> i) & 0x4444u);
    ad[i] = ((t >> (2 * i)) & 15u);
  }
  uint32_t a = t * 3u, b = t * 7u;
  long long t0 = clock64();
  for (int rep = 0; rep < reps; rep++)
    for (int it = 0; it < ITER; it++) {
      const uint32_t rowOff = (uint32_t)(it & 15) * 64;
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        a += t;
        b ^= a;
#pragma unroll
        for (int i = 0; i < CH; i++) {
          uint32_t m;
          if (i < NL) m = *(volatile uint32_t*)&lut[ad[i] + rowOff + 16 * u];
          else asm("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(a), "r"(b), "r"(e[i]));
          asm("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w[i]) : "r"(m), "r"(0x3c003c00u));
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
template <int NL>
void run(int warpsPerSM) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int threads = 32 * warpsPerSM;
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, sizeof(uint32_t) * sms * threads);
  cudaMalloc(&cyc, sizeof(long long) * sms);
  k<NL><<<sms, threads>>>(out, 12345u, cyc, 1);
  k<NL><<<sms, threads>>>(out, 12345u, cyc, g_reps);
  cudaDeviceSynchronize();
  long long c0;
  cudaMemcpy(&c0, cyc, sizeof c0, cudaMemcpyDeviceToHost);
  double pairs = (double)g_reps * ITER * UNR * CH * warpsPerSM;
  printf("LDS lookups %2d of 16  warps/SM=%2d  pairs/clk/SM=%.3f  (= %.1f cells/clk/SM; PRMT-only ceiling 128)\n", NL, warpsPerSM,
         pairs / c0, 64 * pairs / c0);
  cudaFree(out);
  cudaFree(cyc);
}

int main(int argc, char** argv) {
  if (argc > 1) g_reps = atoi(argv[1]);
  for (int w : {8, 16}) {
    run<0>(w); run<1>(w); run<2>(w); run<3>(w); run<4>(w); run<5>(w); run<6>(w); run<8>(w);
  }
  return 0;
}
