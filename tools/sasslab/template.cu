// tools/sasslab/template.cu -- carrier kernel for SASS-level issue-rate experiments.
// The loop body below is a run of NBODY placeholder instructions over NREG live registers; sasslab.py
// overwrites those instruction words in the compiled cubin with an exact sequence (opcode order, operand
// registers, .reuse flags, stall counts), so the measured stream is the one written down, not the one
// ptxas would schedule. Fixed-latency pipes have no hardware interlock, so any order runs at the rate the
// issue logic allows; the values computed are meaningless and never read back for content.
#include <cstdint>
#ifndef NBODY
#define NBODY 1024
#endif
#ifndef NREG
#define NREG 96
#endif
extern "C" __global__ void __launch_bounds__(512, 1) ub(uint32_t* out, long long* cycles, int iters) {
  uint32_t r[NREG];
  const uint32_t t = threadIdx.x * 2654435761u + blockIdx.x;
#pragma unroll
  for (int i = 0; i < NREG; i++) r[i] = (t >> (i & 7)) * (2 * i + 1);
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < NBODY; i++)
      asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(r[i % NREG]) : "r"(r[(i + 1) % NREG]), "r"(r[(i + 2) % NREG]));
  }
  long long t1 = clock64();
  uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < NREG; i++) acc ^= r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}
// encodings sasslab.py copies opcode bits from
extern "C" __global__ void samples(uint32_t* out) {
  uint32_t a = out[0], b = out[1], c = out[2], d = out[3];
  uint32_t m, w;
  asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(a), "r"(b), "r"(c));
  asm volatile("fma.rn.relu.f16x2 %0, %1, %2, %3;" : "=r"(w) : "r"(m), "r"(0x3c003c00u), "r"(d));
  out[4] = w;
}
