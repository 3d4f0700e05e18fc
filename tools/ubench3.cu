// tools/ubench3.cu -- register-file read bandwidth probe.
// Every operand is a per-thread vector register (nothing uniform, nothing loop-invariant for the
// compiler to hoist), so the rates below include operand fetch. Question it answers: is the
// PRMT(3 register reads) + HFMA2(2 register reads) pair limited by the 1 instruction/clk/scheduler
// issue rate, or by the register file delivering about two 32-lane operands per clock?
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define ITER 2048
#define CH 16

template <int MODE>
__global__ void k(uint32_t* out, uint32_t seed, long long* cycles, int reps) {
  uint32_t w[CH], s[CH], ta[4], tb[4];
  const uint32_t t = threadIdx.x * 2654435761u + seed;
#pragma unroll
  for (int i = 0; i < CH; i++) { w[i] = (t >> (i & 7)) & 0x03ff03ffu; s[i] = 0x3210u ^ ((t >> i) & 0x1111u); }
#pragma unroll
  for (int i = 0; i < 4; i++) { ta[i] = t * (i + 3); tb[i] = t * (i + 11); }
  float f[CH], fa = __uint_as_float(0x3f800000u | (t & 0xffff)), fb = __uint_as_float(0x3f000000u | (t >> 16));
#pragma unroll
  for (int i = 0; i < CH; i++) f[i] = (float)i + fa;
  long long t0 = clock64();
  for (int rep = 0; rep < reps; rep++)
    for (int it = 0; it < ITER; it++) {
      if (MODE >= 9) {  // grouped: G PRMTs sharing the table registers, then their G HFMA2s
        constexpr int G = MODE == 9 ? 4 : MODE == 10 ? 8 : 16;
#pragma unroll
        for (int g0 = 0; g0 < CH; g0 += G) {
          uint32_t mm[G];
#pragma unroll
          for (int j = 0; j < G; j++)
            asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(mm[j]) : "r"(ta[(g0 / G) & 3]), "r"(tb[(g0 / G) & 3]), "r"(w[g0 + j]));
#pragma unroll
          for (int j = 0; j < G; j++)
            asm volatile("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w[g0 + j]) : "r"(mm[j]), "r"(0x3c003c00u));
        }
        continue;
      }
#pragma unroll
      for (int i = 0; i < CH; i++) {
        uint32_t m;
        if (MODE == 0) {  // PRMT, 3 distinct vector registers
          asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(ta[i & 3]), "r"(tb[i & 3]), "r"(w[i]));
          w[i] ^= m;  // LOP3 consumer keeps it alive (counted below as a second instruction)
        } else if (MODE == 1) {  // the kernel's pair: PRMT (3 reads) feeding HFMA2.RELU (2 reads)
          asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(ta[i & 3]), "r"(tb[i & 3]), "r"(w[i] ^ 0));
          asm volatile("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w[i]) : "r"(m), "r"(0x3c003c00u));
        } else if (MODE == 2) {  // pair with PRMT reading the same register twice (2 distinct reads)
          asm volatile("prmt.b32 %0, %1, %1, %2;" : "=r"(m) : "r"(ta[i & 3]), "r"(w[i]));
          asm volatile("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w[i]) : "r"(m), "r"(0x3c003c00u));
        } else if (MODE == 3) {  // pair with PRMT's third operand RZ (2 reads)
          asm volatile("prmt.b32 %0, %1, %2, 0;" : "=r"(m) : "r"(ta[i & 3]), "r"(w[i]));
          asm volatile("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w[i]) : "r"(m), "r"(0x3c003c00u));
        } else if (MODE == 4) {  // IADD3, 3 distinct registers
          asm volatile("{.reg .u32 q; add.u32 q, %1, %2; add.u32 %0, %0, q;}" : "+r"(w[i]) : "r"(ta[i & 3]), "r"(s[i]));
        } else if (MODE == 5) {  // FFMA, 3 distinct registers
          asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(f[i]) : "f"(fa), "f"(fb));
        } else if (MODE == 6) {  // FFMA with an immediate multiplier (2 reads)
          asm volatile("fma.rn.f32 %0, %1, 0f3F7FF000, %0;" : "+f"(f[i]) : "f"(fa));
        } else if (MODE == 7) {  // LOP3 with 3 distinct registers
          asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(w[i]) : "r"(ta[i & 3]), "r"(s[i]));
        } else if (MODE == 8) {  // PRMT (3 reads) + HFMA2 whose multiplier is a register too (3 reads)
          asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(ta[i & 3]), "r"(tb[i & 3]), "r"(w[i]));
          asm volatile("fma.rn.relu.f16x2 %0, %1, %2, %0;" : "+r"(w[i]) : "r"(m), "r"(tb[(i + 1) & 3]));
        }
      }
    }
  long long t1 = clock64();
  uint32_t acc = 0;
  for (int i = 0; i < CH; i++) acc ^= w[i] ^ __float_as_uint(f[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

static int g_reps = 4;
template <int MODE>
void run(const char* name, int instrPerChain, int warpsPerSM) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int threads = 32 * warpsPerSM;
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, sizeof(uint32_t) * sms * threads);
  cudaMalloc(&cyc, sizeof(long long) * sms);
  k<MODE><<<sms, threads>>>(out, 12345u, cyc, 1);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<sms, threads>>>(out, 12345u, cyc, g_reps);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long c0; cudaMemcpy(&c0, cyc, sizeof c0, cudaMemcpyDeviceToHost);
  double warpInstr = (double)g_reps * ITER * CH * instrPerChain * warpsPerSM;
  printf("%-44s warps/SM=%2d  warp-instr/clk/SM=%.3f  time=%.3f ms  implied MHz=%.0f\n", name, warpsPerSM,
         warpInstr / c0, ms, c0 / (ms * 1e3));
  cudaFree(out); cudaFree(cyc);
}

int main(int argc, char** argv) {
  if (argc > 1) g_reps = atoi(argv[1]);
  for (int w : {8, 16, 32}) {
    run<0>("PRMT(3 regs)+LOP3", 2, w);
    run<1>("PRMT(3 regs)+HFMA2(2 regs)", 2, w);
    run<2>("PRMT(a,a,sel)+HFMA2", 2, w);
    run<3>("PRMT(a,sel,RZ)+HFMA2", 2, w);
    run<4>("IADD3(3 regs)", 1, w);
    run<5>("FFMA(3 regs)", 1, w);
    run<6>("FFMA(2 regs+imm)", 1, w);
    run<7>("LOP3(3 regs)", 1, w);
    run<8>("PRMT(3 regs)+HFMA2(3 regs)", 2, w);
    run<9>("grouped 4xPRMT(shared a,c) + 4xHFMA2", 2, w);
    run<10>("grouped 8xPRMT(shared a,c) + 8xHFMA2", 2, w);
    run<11>("grouped 16xPRMT(shared a,c) + 16xHFMA2", 2, w);
  }
  return 0;
}
