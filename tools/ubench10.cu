// tools/ubench10.cu -- inner loop of a ROW-BAND SSV formulation: a lane owns 32 model rows
// (16 fp16x2 registers, register i = rows i and i+16), the warp sweeps the sequence one column at a
// time, the column's symbol is warp-uniform and selects one of four per-lane register sets holding
// the emission pairs, so a cell pair costs ONE HFMA2.RELU (no PRMT lookup). Per column: 16
// HFMA2.RELU + 1 SHFL (row 31 -> next lane's row 0) + 1 PRMT + the symbol dispatch.
//   MODE 0: fixed symbol (no dispatch): ceiling of the HFMA2 stream
//   MODE 1: switch per column on a 2-bit symbol
//   MODE 2: switch per 2 columns on a 4-bit symbol pair (16 cases)
//   MODE 3: MODE 1 without SHFL/PRMT (dispatch cost alone)
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#ifndef H
#define H 16
#endif
#ifndef CTAS
#define CTAS 2
#endif
#ifndef THREADS
#define THREADS 256
#endif
#define ONE 0x3c003c00u

__device__ __forceinline__ uint32_t frelu(uint32_t a, uint32_t c) {
  uint32_t d;
  asm("fma.rn.relu.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(ONE), "r"(c));
  return d;
}

#define COLUMN(T)                                   \
  {                                                 \
    _Pragma("unroll") for (int i = H - 1; i >= 1; i--) s[i] = frelu(s[i - 1], T[i]); \
    s[0] = frelu(X, T[0]);                          \
  }

#define PAIR(TA, TB)                                                      \
  {                                                                       \
    uint32_t o = s[H - 1], sh = __shfl_up_sync(0xffffffffu, o, 1), X = __byte_perm(sh, o, 0x1032); \
    COLUMN(TA);                                                           \
    o = s[H - 1]; sh = __shfl_up_sync(0xffffffffu, o, 1); X = __byte_perm(sh, o, 0x1032); \
    COLUMN(TB);                                                           \
  }

template <int MODE>
__global__ void __launch_bounds__(THREADS, CTAS) k(uint32_t* out, const uint32_t* __restrict__ tab, const uint32_t* __restrict__ seq,
                                            int words, long long* cycles) {
  uint32_t s[H], tA[H], tC[H], tG[H], tT[H];
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll
  for (int i = 0; i < H; i++) {
    s[i] = 0;
    tA[i] = tab[(tid * 4 + 0) * H + i];
    tC[i] = tab[(tid * 4 + 1) * H + i];
    tG[i] = tab[(tid * 4 + 2) * H + i];
    tT[i] = tab[(tid * 4 + 3) * H + i];
  }
  const int warp = threadIdx.x >> 5;
  long long t0 = clock64();
  for (int w = 0; w < words; w++) {
    uint32_t word = seq[(warp * 131 + w) & 4095];  // same address for the whole warp
    if (MODE == 5 || MODE == 6) {
      // as MODE 4, but the column loop is NOT unrolled (small code): MODE 5 if-tree per column,
      // MODE 6 a 16-way switch per column pair
      const uint32_t w2 = seq[(warp * 131 + w + 1) & 4095];
      const uint32_t lane = threadIdx.x & 31;
      const uint32_t mine = ((lane < 16 ? word : w2) >> (2 * (lane & 15))) & 3u;
      uint32_t b0 = __ballot_sync(0xffffffffu, mine & 1u), b1 = __ballot_sync(0xffffffffu, mine & 2u);
      if (MODE == 5) {
#pragma unroll 1
        for (int c = 0; c < 32; c++) {
          const uint32_t o = s[H - 1], sh = __shfl_up_sync(0xffffffffu, o, 1);
          const uint32_t X = __byte_perm(sh, o, 0x1032);
          if (b1 & 1u) {
            if (b0 & 1u) COLUMN(tT) else COLUMN(tG)
          } else {
            if (b0 & 1u) COLUMN(tC) else COLUMN(tA)
          }
          b0 >>= 1;
          b1 >>= 1;
        }
      } else {
#pragma unroll 1
        for (int c = 0; c < 16; c++) {
          const uint32_t xx = (b0 & 3u) | ((b1 & 3u) << 2);  // bit0: col0.lo, bit1: col1.lo, bit2: col0.hi, bit3: col1.hi
          switch (xx) {
            case 0: PAIR(tA, tA) break;  case 1: PAIR(tC, tA) break;  case 4: PAIR(tG, tA) break;  case 5: PAIR(tT, tA) break;
            case 2: PAIR(tA, tC) break;  case 3: PAIR(tC, tC) break;  case 6: PAIR(tG, tC) break;  case 7: PAIR(tT, tC) break;
            case 8: PAIR(tA, tG) break;  case 9: PAIR(tC, tG) break;  case 12: PAIR(tG, tG) break; case 13: PAIR(tT, tG) break;
            case 10: PAIR(tA, tT) break; case 11: PAIR(tC, tT) break; case 14: PAIR(tG, tT) break; default: PAIR(tT, tT) break;
          }
          b0 >>= 2;
          b1 >>= 2;
        }
      }
      w++;
    } else if (MODE == 4) {
      // two words = 32 columns: lane k holds column k's symbol, two ballots give warp-uniform bit masks
      const uint32_t w2 = seq[(warp * 131 + w + 1) & 4095];
      const uint32_t lane = threadIdx.x & 31;
      const uint32_t mine = ((lane < 16 ? word : w2) >> (2 * (lane & 15))) & 3u;
      const uint32_t b0 = __ballot_sync(0xffffffffu, mine & 1u), b1 = __ballot_sync(0xffffffffu, mine & 2u);
#pragma unroll
      for (int c = 0; c < 32; c++) {
        const uint32_t o = s[H - 1], sh = __shfl_up_sync(0xffffffffu, o, 1);
        const uint32_t X = __byte_perm(sh, o, 0x1032);
        if (b1 & (1u << c)) {
          if (b0 & (1u << c)) COLUMN(tT) else COLUMN(tG)
        } else {
          if (b0 & (1u << c)) COLUMN(tC) else COLUMN(tA)
        }
      }
      w++;
    } else if (MODE == 2) {
#pragma unroll
      for (int c = 0; c < 8; c++) {
        const uint32_t xx = (word >> (4 * c)) & 15u;
        switch (xx) {
          case 0: PAIR(tA, tA) break;  case 1: PAIR(tC, tA) break;  case 2: PAIR(tG, tA) break;  case 3: PAIR(tT, tA) break;
          case 4: PAIR(tA, tC) break;  case 5: PAIR(tC, tC) break;  case 6: PAIR(tG, tC) break;  case 7: PAIR(tT, tC) break;
          case 8: PAIR(tA, tG) break;  case 9: PAIR(tC, tG) break;  case 10: PAIR(tG, tG) break; case 11: PAIR(tT, tG) break;
          case 12: PAIR(tA, tT) break; case 13: PAIR(tC, tT) break; case 14: PAIR(tG, tT) break; default: PAIR(tT, tT) break;
        }
      }
    } else {
#pragma unroll
      for (int c = 0; c < 16; c++) {
        const uint32_t x = (word >> (2 * c)) & 3u;
        uint32_t X;
        if (MODE == 3) X = s[H - 1] ^ word;
        else {
          const uint32_t o = s[H - 1], sh = __shfl_up_sync(0xffffffffu, o, 1);
          X = __byte_perm(sh, o, 0x1032);
        }
        if (MODE == 0) { COLUMN(tA) }
        else {
          switch (x) {
            case 0: COLUMN(tA) break;
            case 1: COLUMN(tC) break;
            case 2: COLUMN(tG) break;
            default: COLUMN(tT) break;
          }
        }
      }
    }
  }
  long long t1 = clock64();
  uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < H; i++) acc ^= s[i];
  out[tid] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, int ctasPerSM, int words) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int threads = THREADS, grid = sms * ctasPerSM, n = grid * threads;
  uint32_t *out, *tab, *seq;
  long long* cyc;
  cudaMalloc(&out, 4 * n);
  cudaMalloc(&tab, 4 * (size_t)n * 4 * H);
  cudaMalloc(&seq, 4 * 4096);
  cudaMalloc(&cyc, 8 * grid);
  uint32_t* h = (uint32_t*)malloc(4 * (size_t)n * 4 * H);
  for (size_t i = 0; i < (size_t)n * 4 * H; i++) {  // emission pairs: small negative / positive halves (x256 scale irrelevant here)
    const int a = (int)(rand() % 7) - 4, b = (int)(rand() % 7) - 4;
    auto half = [](int v) -> uint32_t { if (v == 0) return 0; uint32_t sgn = v < 0 ? 0x8000u : 0; int m = v < 0 ? -v : v; int e = 0; while ((1 << (e + 1)) <= m) e++; return sgn | ((uint32_t)(e + 15) << 10) | (((uint32_t)m << (10 - e)) & 0x3ffu); };
    h[i] = half(a) | (half(b) << 16);
  }
  cudaMemcpy(tab, h, 4 * (size_t)n * 4 * H, cudaMemcpyHostToDevice);
  uint32_t hs[4096];
  for (int i = 0; i < 4096; i++) hs[i] = (uint32_t)rand() * 65537u ^ (uint32_t)rand();
  cudaMemcpy(seq, hs, sizeof hs, cudaMemcpyHostToDevice);
  k<MODE><<<grid, threads>>>(out, tab, seq, 64, cyc);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<grid, threads>>>(out, tab, seq, words, cyc);
  cudaEventRecord(e1);
  cudaError_t e = cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  long long c0 = 0;
  cudaMemcpy(&c0, cyc, 8, cudaMemcpyDeviceToHost);
  const double cols = (double)words * 16, warps = ctasPerSM * (THREADS / 32);
  const double cellsPerClkSM = cols * warps * 32 * 2 * H / c0;
  const double tcups = cols * warps * sms * 32 * 2 * H / (ms * 1e-3) / 1e12;
  if (name[0] != '(') printf("%-44s CTAs/SM=%d  cells/clk/SM=%7.2f (clock64)  %7.2f Tcell/s by events (%.3f ms)  %s\n", name, ctasPerSM, cellsPerClkSM, tcups, ms,
         e == cudaSuccess ? "" : cudaGetErrorString(e));
  free(h);
  cudaFree(out); cudaFree(tab); cudaFree(seq); cudaFree(cyc);
}

int main(int argc, char** argv) {
  const int words = argc > 1 ? atoi(argv[1]) : 4096;
  for (int i = 0; i < 40; i++) run<0>("(warm-up)", CTAS, words);
  for (int c : {CTAS}) {
    run<0>("fixed symbol: 16 HFMA2 + SHFL + PRMT", c, words);
    run<1>("switch per column", c, words);
    run<2>("switch per column pair (16 cases)", c, words);
    run<3>("switch per column, no SHFL/PRMT", c, words);
    run<4>("ballot masks, if-tree per column", c, words);
    run<5>("ballot masks, if-tree, rolled loop", c, words);
    run<6>("ballot masks, 16-way per pair, rolled", c, words);
  }
  return 0;
}
