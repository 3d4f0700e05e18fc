// ssv_kernels.cu -- hand-written sm_100a kernels for HAVAC's SSV filter.
//
// What is computed (bit-exact with test/softSsv/SoftSsv.cpp:31-62 and the FPGA cell of
// device/HavacHls.cpp:370-402):   S(p,j) = S(p-1,j-1) + e[p][x_j];  S<0 -> 0;  S>=256 -> hit, 0.
//
// Fast path design ("f16x2", diagonal-owner):
//  * Each diagonal is an independent chain, so a thread owns 2*R adjacent diagonals and keeps
//    their scores in R registers for the whole model axis; nothing ever moves between threads.
//  * A score s in [0,255] is held as the fp16 value 256*s, two diagonals per register. The cell
//    update is ONE instruction per two cells: HFMA2.RELU  w = relu(m*1 + w)  with m = 256*e.
//    All values are multiples of 256 below 2^16, so fp16 arithmetic is exact; a negative sum is
//    clamped by .RELU (the reset-to-zero of HavacHls.cpp:382); a sum >= 256*256 = 65536
//    overflows fp16 to +inf, and inf is sticky under further adds. A hit is therefore never
//    lost: each thread looks for inf once per 32-row block and, if it finds one, replays that
//    block from a saved copy with exact integer arithmetic (which also performs the reset that
//    follows a hit). Hits are ~1e-7 per cell, so the replay is off the critical path.
//  * The emission lookup e[p][x_j] for two cells is ONE PRMT: the row's four fp16 scores sit
//    in two registers (8 bytes) and the per-symbol selector byte 0x10+0x22*x picks the two
//    bytes of score x. Selectors for the thread's sliding window are kept in registers and
//    reused across rows (register k at row p needs the same window as register k+1 at row
//    p-2), so only one new selector per row per thread is produced.
//  * Per 2 cells: 1 PRMT (ALU pipe) + 1 HFMA2 (FMA pipe). The kernel is issue-bound, not
//    memory-bound: per 8192-diagonal row it reads 8 B of table and 1/4 B of sequence.
//  * Score-table rows are staged into shared memory in 512-row chunks by TMA bulk copies
//    (cp.async.bulk + mbarrier, double buffered). The 2-bit packed sequence is read from HBM
//    with coalesced 128-bit loads, one chunk ahead, and expanded to selector bytes in a
//    shared-memory ring that slides along the diagonal band.
//  * Hits are compacted with warp ballot + prefix popcount and one atomicAdd per warp-row.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>

#include "ssv_kernels.cuh"

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr long long SEG = 12288;  // device/PublicDefines.h:18-22

__device__ __forceinline__ unsigned long long encode_hit(long long col, uint32_t row) {
  // device/HitReporting.cpp:421-430
  unsigned long long seg = (unsigned long long)(col / SEG);
  unsigned long long cell = (unsigned long long)(col % SEG);
  return cell | (seg << 14) | ((unsigned long long)row << 40);
}

// Warp-convergent hit compaction: ballot, one atomic per warp, prefix popcount for the slot.
__device__ __forceinline__ void emit_hits_warp(const SsvParams& prm, bool hit, long long col,
                                               uint32_t row) {
  unsigned m = __ballot_sync(FULL, hit);
  if (m == 0) return;
  unsigned lane = threadIdx.x & 31;
  unsigned long long base = 0;
  if (lane == (unsigned)(__ffs(m) - 1)) base = atomicAdd(prm.hitCount, (unsigned long long)__popc(m));
  base = __shfl_sync(FULL, base, __ffs(m) - 1);
  if (hit) {
    unsigned long long slot = base + __popc(m & ((1u << lane) - 1));
    if (slot < prm.hitCap) prm.hits[slot] = encode_hit(col, row);
    else atomicOr(prm.status, SSV_STATUS_OVERFLOW);
  }
}

// ---------------------------------------------------------------------------------------
// Score-table preparation: int8 [row][4] -> fp16x2 pairs scaled by 256; pad rows get -128.
// ---------------------------------------------------------------------------------------
__global__ void table_prep_kernel(const int8_t* __restrict__ tab8, uint32_t nRows,
                                  uint2* __restrict__ tab16, uint32_t rowsPadded) {
  uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rowsPadded) return;
  int e[4] = {-128, -128, -128, -128};
  if (r < nRows) {
    uint32_t v = reinterpret_cast<const uint32_t*>(tab8)[r];
    for (int i = 0; i < 4; i++) e[i] = (int)(int8_t)(v >> (8 * i));
  }
  uint32_t h[4];
  for (int i = 0; i < 4; i++) h[i] = __half_as_ushort(__int2half_rn(e[i] * 256));
  tab16[r] = make_uint2(h[0] | (h[1] << 16), h[2] | (h[3] << 16));
}

// ---------------------------------------------------------------------------------------
// Plain kernel: one thread per diagonal, integer arithmetic, everything from global memory.
// The correctness-first variant and the cross-check for the fast path in tests.
// ---------------------------------------------------------------------------------------
__global__ void ssv_scalar_kernel(SsvParams prm) {
  long long nDiag = prm.nLoad + (long long)prm.nRows - 1;  // d in [-(P-1), nLoad-1]
  long long warpBase = ((long long)blockIdx.x * blockDim.x + (threadIdx.x & ~31));
  for (; warpBase < nDiag; warpBase += (long long)gridDim.x * blockDim.x) {
    long long d = warpBase + (threadIdx.x & 31) - ((long long)prm.nRows - 1);
    bool live = d < prm.nLoad;
    int s = 0;
    for (uint32_t p = 0; p < prm.nRows; p++) {
      long long j = d + p;
      bool hit = false;
      if (live && j >= 0 && j < prm.nLoad) {
        uint32_t x = (prm.seqPacked[j >> 4] >> ((j & 15) * 2)) & 3;
        s += prm.tab8[4 * (size_t)p + x];
        if (s < 0) s = 0;
        else if (s >= 256) { s = 0; hit = true; }
      }
      long long col = j + prm.loadBegin;
      hit = hit && col >= prm.colBegin && col < prm.colEnd;
      emit_hits_warp(prm, hit, col, p);
    }
  }
}

}  // namespace

// =======================================================================================
// Fast path
// =======================================================================================
namespace {

constexpr int R = SSV_R;
constexpr int THREADS = SSV_THREADS;
constexpr int W = SSV_TILE_DIAGS;
constexpr int TC = SSV_ROW_CHUNK;
constexpr int BLOCK_ROWS = SSV_BLOCK_ROWS;
constexpr int PAD = 64;              // symbols the ring stays ahead of the furthest reader
constexpr int RING = 16384;          // bytes; power of two >= W + 2*TC + PAD
constexpr int RING_MASK = RING - 1;
static_assert(W + 2 * TC + PAD <= RING, "ring too small");
static_assert(TC % BLOCK_ROWS == 0, "row chunk must hold whole blocks");
static_assert(TC / 64 <= 32, "chunk refill is done by one warp");

struct __align__(16) FastSmem {
  uint8_t ring[RING];          // selector bytes 0x10 + 0x22*x, ring[(j - D0) & RING_MASK]
  uint2 tab[2][TC];            // double-buffered score-table chunks
  unsigned long long mbar[2];  // TMA completion barriers
  unsigned int nextTile;
  int abortSeen;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE_%=;\n"
      "bra WAIT_LOOP_%=;\n"
      "WAIT_DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// TMA 1-D bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes,
                                            unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// 16 packed symbols -> 16 selector bytes. Byte for symbol x: low nibble 2x, high nibble 2x+1,
// i.e. "take bytes 2x and 2x+1 of the 8-byte score row" when used as a PRMT selector.
__device__ __forceinline__ uint4 expand16(uint32_t x) {
  uint32_t o[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    uint32_t b = (x >> (8 * i)) & 0xffu;
    uint32_t s = (b & 3u) | ((b & 0xcu) << 6) | ((b & 0x30u) << 12) | ((b & 0xc0u) << 18);
    o[i] = 0x10101010u + 0x22u * s;
  }
  return make_uint4(o[0], o[1], o[2], o[3]);
}

// Packed word holding local symbols [16*wi, 16*wi+16); zero ('A') outside the loaded range.
__device__ __forceinline__ uint32_t load_seq_word(const SsvParams& prm, long long wi) {
  long long nWords = prm.nLoad >> 4;
  return (wi >= 0 && wi < nWords) ? __ldg(prm.seqPacked + wi) : 0u;
}

// PRMT with the raw PTX semantics (CUDA's __byte_perm would first AND the selector with
// 0x7777). Only selector bits [15:0] are read; our nibbles are all <= 7 (plain byte copy).
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}

__device__ __forceinline__ uint32_t hfma2_relu_add(uint32_t m, uint32_t w) {
  uint32_t r;
  const uint32_t one = 0x3c003c00u;
  asm("fma.rn.relu.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(m), "r"(one), "r"(w));
  return r;
}

// true when either fp16 half of v is +inf (0x7c00); all finite states are <= 0x7bf8.
__device__ __forceinline__ bool has_inf(uint32_t v) {
  return ((v + 0x04000400u) & 0x80008000u) != 0u;
}

// Exact integer replay of one 32-row block for the registers of this thread that need it.
// Warp-convergent: every lane runs the same trip counts, lanes without work idle, so the
// ballot-based hit compaction inside is legal.
//   sv   : state at the start of the block (fp16x2)      w : state to overwrite with the result
//   need : per-thread "replay every register" flag (left-edge band); otherwise registers are
//          replayed when their fast-path result holds an inf.
__device__ __noinline__ void exact_block(const SsvParams& prm, const FastSmem& sm, const uint2* tab,
                                         const uint32_t (&sv)[R], uint32_t (&w)[R], bool needAll,
                                         long long d0, long long D0, uint32_t pb) {
#pragma unroll
  for (int k = 0; k < R; k++) {
    bool need = needAll || has_inf(w[k]);
    if (!__any_sync(FULL, need)) continue;
    uint32_t res = 0;
#pragma unroll 1
    for (int h = 0; h < 2; h++) {
      int s = __half2int_rn(__ushort_as_half((unsigned short)(sv[k] >> (16 * h)))) >> 8;
      long long d = d0 + 2 * k + h;
#pragma unroll 1
      for (int r = 0; r < BLOCK_ROWS; r++) {
        uint32_t p = pb + r;
        long long j = d + p;
        bool hit = false;
        if (need) {
          if (j < 0) {
            s = 0;  // column left of the sequence start: prev score is 0 (SoftSsv.cpp:37)
          } else {
            uint32_t x = sm.ring[(uint32_t)(j - D0) & RING_MASK] >> 5;
            uint2 e2 = tab[p & (TC - 1)];
            uint32_t eh = ((x & 2) ? e2.y : e2.x) >> (16 * (x & 1));
            s += __half2int_rn(__ushort_as_half((unsigned short)eh)) >> 8;
            if (s < 0) s = 0;
            else if (s >= 256) { s = 0; hit = true; }
          }
        }
        long long col = j + prm.loadBegin;
        hit = hit && p < prm.nRows && col >= prm.colBegin && col < prm.colEnd;
        emit_hits_warp(prm, hit, col, p);
      }
      res |= (uint32_t)__half_as_ushort(__int2half_rn(s * 256)) << (16 * h);
    }
    if (need) w[k] = res;
  }
}

// One 4-row group of the fast path. u = group index inside the 32-row block (compile time),
// so every register index below is static after unrolling.
template <int U>
__device__ __forceinline__ void four_rows(uint32_t (&w)[R], uint32_t (&e0)[R], uint32_t (&e1)[R],
                                          const uint2* __restrict__ tabRow, uint32_t wn,
                                          uint32_t wn1) {
  // rows 4U+0..3  ->  (q, c) = (2U,0) (2U,1) (2U+1,0) (2U+1,1)
  const uint4 ab01 = *reinterpret_cast<const uint4*>(tabRow + 4 * U);      // rows 4U, 4U+1
  const uint4 ab23 = *reinterpret_cast<const uint4*>(tabRow + 4 * U + 2);  // rows 4U+2, 4U+3
  constexpr int Q0 = 2 * U, Q1 = 2 * U + 1;
#pragma unroll
  for (int k = 0; k < R; k++) w[k] = hfma2_relu_add(prmt(ab01.x, ab01.y, e0[(Q0 + k) % R]), w[k]);
  e0[Q0 % R] = wn;  // window at symbol b        (PRMT reads selector bits [15:0] only)
#pragma unroll
  for (int k = 0; k < R; k++) w[k] = hfma2_relu_add(prmt(ab01.z, ab01.w, e1[(Q0 + k) % R]), w[k]);
  e1[Q0 % R] = wn >> 8;  // window at b+1
#pragma unroll
  for (int k = 0; k < R; k++) w[k] = hfma2_relu_add(prmt(ab23.x, ab23.y, e0[(Q1 + k) % R]), w[k]);
  e0[Q1 % R] = wn >> 16;  // window at b+2
#pragma unroll
  for (int k = 0; k < R; k++) w[k] = hfma2_relu_add(prmt(ab23.z, ab23.w, e1[(Q1 + k) % R]), w[k]);
  e1[Q1 % R] = __funnelshift_r(wn, wn1, 24);  // window at b+3 (crosses into the next word)
}

template <int U>
struct RowGroups {
  static __device__ __forceinline__ void run(uint32_t (&w)[R], uint32_t (&e0)[R], uint32_t (&e1)[R],
                                             const uint2* tabRow, const uint8_t* ring,
                                             uint32_t pos, uint32_t& wn) {
    uint32_t wn1 = *reinterpret_cast<const uint32_t*>(ring + ((pos + 4 * (U + 1)) & RING_MASK));
    four_rows<U>(w, e0, e1, tabRow, wn, wn1);
    wn = wn1;
    RowGroups<U + 1>::run(w, e0, e1, tabRow, ring, pos, wn);
  }
};
template <>
struct RowGroups<BLOCK_ROWS / 4> {
  static __device__ __forceinline__ void run(uint32_t (&)[R], uint32_t (&)[R], uint32_t (&)[R],
                                             const uint2*, const uint8_t*, uint32_t, uint32_t&) {}
};

__global__ void __launch_bounds__(THREADS, 2) ssv_f16x2_kernel(const SsvParams prm) {
  __shared__ FastSmem sm;
  const int tid = threadIdx.x;
  uint32_t mbarPhase = 0;  // bit b = parity to wait for on sm.mbar[b]

  if (tid == 0) {
    mbar_init(&sm.mbar[0], 1);
    mbar_init(&sm.mbar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    sm.abortSeen = 0;
    sm.nextTile = atomicAdd(prm.tileCounter, 1u);
  }
  __syncthreads();

  while (true) {
    const unsigned tile = sm.nextTile;
    const bool aborted = sm.abortSeen != 0;
    __syncthreads();  // everyone has read nextTile before thread 0 overwrites it
    if (tile >= prm.nTiles || aborted) break;

    const long long D0 = prm.diagBegin + (long long)tile * W;  // first diagonal of this tile
    const long long d0 = D0 + (long long)tid * (2 * R);        // first diagonal of this thread
    // rows this tile has real cells in: leftmost diagonal D0 lives longest on the right side,
    // rightmost diagonal D0+W-1 enters first on the left side.
    long long rowEnd = prm.nLoad - D0;
    if (rowEnd > (long long)prm.nRows) rowEnd = prm.nRows;
    long long rowBegin = -(D0 + W - 1);
    if (rowBegin < 0) rowBegin = 0;
    const int c0 = (int)(rowBegin / TC);
    const int c1 = (int)((rowEnd + TC - 1) / TC);  // exclusive
    if (rowEnd <= rowBegin) {                       // nothing to do (defensive)
      if (tid == 0) sm.nextTile = atomicAdd(prm.tileCounter, 1u);
      __syncthreads();
      continue;
    }

    // --- prologue: queue the first two table chunks, fill the ring, load selectors ---------
    if (tid == 0) {
      sm.nextTile = atomicAdd(prm.tileCounter, 1u);
      if (*prm.abortFlag) sm.abortSeen = 1;
      for (int i = 0; i < 2 && c0 + i < c1; i++) {
        mbar_expect_tx(&sm.mbar[i], TC * sizeof(uint2));
        tma_load_1d(sm.tab[i], prm.tab16 + (size_t)(c0 + i) * TC, TC * sizeof(uint2), &sm.mbar[i]);
      }
    }
    const long long jFill0 = D0 + (long long)c0 * TC;  // first symbol the tile needs
    for (int i = tid; i < (W + 2 * TC + PAD) / 16; i += THREADS) {
      long long j = jFill0 + 16ll * i;
      uint4 v = expand16(load_seq_word(prm, j >> 4));
      *reinterpret_cast<uint4*>(sm.ring + ((uint32_t)(j - D0) & RING_MASK)) = v;
    }
    __syncthreads();

    uint32_t w[R], sv[R], e0[R], e1[R];
    {
      const uint32_t p0 = (uint32_t)c0 * TC;
      const uint32_t base = (uint32_t)(tid * 2 * R) + p0;  // ring position of symbol d0 + p0
#pragma unroll
      for (int n = 0; n < R; n += 2) {
        uint32_t a = *reinterpret_cast<const uint32_t*>(sm.ring + ((base + 2 * n) & RING_MASK));
        uint32_t b = *reinterpret_cast<const uint32_t*>(sm.ring + ((base + 2 * n + 4) & RING_MASK));
        e0[n] = a;
        e1[n] = a >> 8;
        e0[n + 1] = a >> 16;
        e1[n + 1] = __funnelshift_r(a, b, 24);
      }
#pragma unroll
      for (int k = 0; k < R; k++) w[k] = 0u;
    }

    // --- main loop over row chunks -----------------------------------------------------------
    for (int c = c0; c < c1; c++) {
      const int buf = (c - c0) & 1;
      // prefetch the packed symbols that extend the ring for chunk c+2 (consumed after compute)
      uint4 refill = make_uint4(0, 0, 0, 0);
      const long long jRefill = D0 + W + (long long)(c + 2) * TC + PAD;
      if (tid < TC / 64 && c + 1 < c1) {
        long long wi = (jRefill >> 4) + 4 * tid;
        refill.x = load_seq_word(prm, wi);
        refill.y = load_seq_word(prm, wi + 1);
        refill.z = load_seq_word(prm, wi + 2);
        refill.w = load_seq_word(prm, wi + 3);
      }
      mbar_wait(&sm.mbar[buf], (mbarPhase >> buf) & 1u);
      mbarPhase ^= 1u << buf;

      long long chunkEnd = (long long)(c + 1) * TC;
      if (chunkEnd > rowEnd) chunkEnd = rowEnd;
      for (uint32_t pb = (uint32_t)c * TC; (long long)pb < chunkEnd; pb += BLOCK_ROWS) {
#pragma unroll
        for (int k = 0; k < R; k++) sv[k] = w[k];
        const uint32_t pos = (uint32_t)(tid * 2 * R) + pb + 2 * R;  // ring position of d0+pb+2R
        uint32_t wn = *reinterpret_cast<const uint32_t*>(sm.ring + (pos & RING_MASK));
        RowGroups<0>::run(w, e0, e1, sm.tab[buf] + (pb & (TC - 1)), sm.ring, pos, wn);

        // post-pass: left-edge band and hit replay
        uint32_t mx = 0;
#pragma unroll
        for (int k = 0; k < R; k += 2) mx = __vimax3_s16x2(mx, w[k], w[k + 1]);
        bool needAll = false;
        if (d0 + (long long)pb < 0) {  // some of my cells in this block lie left of column 0
          if (d0 + 2 * R - 1 + (long long)pb + BLOCK_ROWS - 1 < 0) {
#pragma unroll
            for (int k = 0; k < R; k++) w[k] = 0u;  // all of them: nothing real yet
            mx = 0;
          } else {
            needAll = true;
          }
        }
        if (__any_sync(FULL, needAll || has_inf(mx))) {
          // cold path: hand copies to the out-of-line replay so w/sv stay in registers above
          uint32_t svc[R], wc[R];
#pragma unroll
          for (int k = 0; k < R; k++) { svc[k] = sv[k]; wc[k] = w[k]; }
          exact_block(prm, sm, sm.tab[buf], svc, wc, needAll, d0, D0, pb);
#pragma unroll
          for (int k = 0; k < R; k++) w[k] = wc[k];
        }
      }

      __syncthreads();  // all reads of tab[buf] and of the ring tail about to be recycled are done
      if (tid == 0) {
        if (c + 2 < c1) {
          mbar_expect_tx(&sm.mbar[buf], TC * sizeof(uint2));
          tma_load_1d(sm.tab[buf], prm.tab16 + (size_t)(c + 2) * TC, TC * sizeof(uint2), &sm.mbar[buf]);
        }
        if (*prm.abortFlag) sm.abortSeen = 1;
      }
      if (tid < TC / 64 && c + 1 < c1) {
        uint8_t* dst = sm.ring + ((uint32_t)(jRefill - D0 + 64 * tid) & RING_MASK);
        reinterpret_cast<uint4*>(dst)[0] = expand16(refill.x);
        reinterpret_cast<uint4*>(dst)[1] = expand16(refill.y);
        reinterpret_cast<uint4*>(dst)[2] = expand16(refill.z);
        reinterpret_cast<uint4*>(dst)[3] = expand16(refill.w);
      }
      if (sm.abortSeen) break;  // uniform: written before the barrier of the previous chunk
    }
    __syncthreads();
    if (sm.abortSeen) break;
  }
  if (tid == 0 && sm.abortSeen) atomicOr(prm.status, SSV_STATUS_ABORTED);
}

}  // namespace

// =======================================================================================
// launchers
// =======================================================================================
void ssv_launch_table_prep(const int8_t* tab8, uint32_t nRows, uint2* tab16, uint32_t rowsPadded,
                           void* stream) {
  table_prep_kernel<<<(rowsPadded + 255) / 256, 256, 0, (cudaStream_t)stream>>>(tab8, nRows, tab16,
                                                                              rowsPadded);
}

int ssv_f16x2_ctas_per_sm() {
  int n = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, ssv_f16x2_kernel, THREADS, 0);
  return n > 0 ? n : 1;
}

void ssv_launch_f16x2(const SsvParams& prm, int numSMs, void* stream) {
  int grid = numSMs * ssv_f16x2_ctas_per_sm();
  if ((unsigned)grid > prm.nTiles) grid = (int)prm.nTiles;
  if (grid < 1) grid = 1;
  ssv_f16x2_kernel<<<grid, THREADS, 0, (cudaStream_t)stream>>>(prm);
}

void ssv_launch_scalar(const SsvParams& prm, void* stream) {
  long long nDiag = prm.nLoad + (long long)prm.nRows - 1;
  long long blocks = (nDiag + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  ssv_scalar_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(prm);
}
