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
//    lost: each thread looks for inf once per superblock and, if it finds one, replays that
//    superblock from a checkpoint with exact integer arithmetic (which also performs the reset that
//    follows a hit). Hits are ~1e-7 per cell, so the replay is off the critical path.
//  * The emission lookup e[p][x_j] for two cells is ONE PRMT: the row's four fp16 scores sit
//    in two registers (8 bytes) and the per-symbol selector byte 0x10+0x22*x picks the two
//    bytes of score x. Selectors for the thread's sliding window are kept in registers and
//    reused across rows (register k at row p needs the same window as register k+1 at row
//    p-2), so only one new selector per row per thread is produced.
//  * Per 2 cells: 1 PRMT (ALU pipe) + 1 HFMA2 (FMA pipe). The kernel is issue-bound, not
//    memory-bound: per 8192-diagonal row it reads 8 B of table and 1/4 B of sequence.
//  * Score-table rows are staged into shared memory in 2048-row chunks by TMA bulk copies
//    (cp.async.bulk + mbarrier, double buffered). The 2-bit packed sequence is read from HBM
//    with 128-bit loads and expanded into a shared-memory ring of ready-made selector windows
//    (one 32-bit entry per symbol) that slides along the diagonal band; a thread fetches the
//    windows of four rows as two 64-bit halves of one conflict-free ring entry, so the hot loop
//    spends no ALU instruction on selectors. The warps of a CTA do not meet between chunks: the
//    warp that is last through a chunk extends the ring and queues the table chunk after next.
//  * The state at the start of every superblock (one 32-row block: a replay costs the rows it
//    covers, and real searches have dense runs of hits; four blocks would be 0.7 % faster on
//    hit-free input and up to 37 % slower on hit-dense input) is checkpointed in shared memory;
//    hits (inf) are looked for once per superblock. The replay runs sixteen rows at a time
//    without looking and walks only the sixteen that end in inf row by row.
//  * A hit found in the replay takes a place in its warp's stage in shared memory; the warp moves the stage to
//    the device's hit list with one atomicAdd and coalesced stores (the plain kernel compacts with warp ballot +
//    prefix popcount and one atomicAdd per warp-row).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <atomic>
#include <cstdint>
#include <cstdlib>

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
// Score-table projection on the device: negative-log match emissions ->
// int8 scores at threshold 256, bit for bit the loop of p7HmmProjectForThreshold256
// (PhmmReprojection/PhmmReprojection.cpp:117-144): float alpha - (e * beta) with the product
// and the difference rounded separately (no FMA contraction), round() half away from zero,
// saturate to [-128, 127]. One thread per model position; alpha = 2*scale and
// beta = log2(e)*scale come per model from the host, which owns the scale-factor arithmetic.
// ---------------------------------------------------------------------------------------
__global__ void project_table_kernel(const float4* __restrict__ emissions, uint32_t nRows,
                                     const uint32_t* __restrict__ modelFirstRow, const float2* __restrict__ alphaBeta,
                                     uint32_t nModels, uint32_t* __restrict__ tab8) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nRows) return;
  uint32_t a = 0, b = nModels;  // last model whose first row is <= r
  while (b - a > 1) {
    const uint32_t mid = a + (b - a) / 2;
    if (modelFirstRow[mid] <= r) a = mid;
    else b = mid;
  }
  const float2 ab = alphaBeta[a];
  const float4 e = emissions[r];
  const float in[4] = {e.x, e.y, e.z, e.w};
  uint32_t packed = 0;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    float v = __fsub_rn(ab.x, __fmul_rn(in[i], ab.y));
    v = roundf(v);
    if (v < -128.f) v = -128.f;
    if (v > 127.f) v = 127.f;
    packed |= ((uint32_t)(int)v & 0xffu) << (8 * i);
  }
  tab8[r] = packed;
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

// Per-cell trace (the reference's HAVAC_PER_CELL_DATA_TESTING mode, device/PublicDefines.h:11): one thread per
// diagonal that crosses the window, run from its first row with the plain kernel's arithmetic on what the
// DEVICE holds (packed symbols, int8 table), writing a record for every cell inside the window.
__global__ void ssv_trace_kernel(SsvParams prm, uint32_t rowBegin, uint32_t rowEnd, long long colBegin, long long colEnd,
                                 SsvCellRecord* __restrict__ out) {
  // diagonals (local column - row) that have a cell in the window: [colBegin - (rowEnd - 1), colEnd - rowBegin)
  const long long dFirst = colBegin - (long long)(rowEnd - 1);
  const long long nDiag = (colEnd - (long long)rowBegin) - dFirst;
  const long long width = colEnd - colBegin;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < nDiag; t += (long long)gridDim.x * blockDim.x) {
    const long long d = dFirst + t;  // in columns of the loaded range
    int s = 0;
    for (uint32_t p = d < 0 ? (uint32_t)(-d) : 0u; p < rowEnd; p++) {
      const long long j = d + p;
      if (j >= prm.nLoad) break;
      const uint32_t x = (prm.seqPacked[j >> 4] >> ((j & 15) * 2)) & 3;
      const int e = prm.tab8[4 * (size_t)p + x];
      const int prev = s;
      bool hit = false;
      s += e;
      if (s < 0) s = 0;
      else if (s >= 256) { s = 0; hit = true; }
      if (p >= rowBegin && j >= colBegin && j < colEnd) {
        SsvCellRecord r;
        r.prevValue = (uint8_t)prev;
        r.matchScore = (int8_t)e;
        r.cellValue = (uint8_t)s;
        for (int i = 0; i < 4; i++) r.phmmVector[i] = prm.tab8[4 * (size_t)p + i];
        r.symbol = (uint8_t)x;
        r.passesThreshold = hit ? 1 : 0;
        r.pad[0] = r.pad[1] = r.pad[2] = 0;
        out[(size_t)(p - rowBegin) * width + (j - colBegin)] = r;
      }
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
constexpr int BLOCK_ROWS = SSV_BLOCK_ROWS;          // rows unrolled in one block (= 2R = 32)
// One 32-row block between two checkpoints / hit checks: a replay costs the rows it covers, and two blocks per check
// measured slower even on hit-free input (profiles/r1_ubench10_row_band.txt; again in round 2 with the re-scheduled block).
constexpr int NSLOT = (W + 2 * TC) / 32;            // 32-symbol groups the window ring holds
static_assert(BLOCK_ROWS == 32 && R == 16, "the ring layout assumes 32 diagonals and 32 rows per thread-block");
static_assert(TC % BLOCK_ROWS == 0, "row chunk must hold whole blocks");
static_assert(TC % 64 == 0, "the ring is extended 64 symbols per lane and pass");
static_assert((W + 2 * TC) % 32 == 0, "ring must hold whole 32-symbol groups");

constexpr int WARPS = THREADS / 32;
// Hits on their way out. A hit is found by ONE lane deep inside the replay; taking its place in the device's hit
// list there costs an atomic on one global counter whose answer the lane has to wait for (a microsecond, with the
// other 31 lanes idle). Instead the lane takes a place in its warp's stage with a shared-memory atomic, and the
// warp moves the stage out -- one global atomic, coalesced stores -- when it is half full or the tile ends.
constexpr int HSTAGE = 240;  // hit words per warp (what two CTAs per SM leave of the shared memory)
struct __align__(16) HitStage {
  unsigned long long hit[WARPS][HSTAGE];
  unsigned int count[WARPS];  // places taken; may run past HSTAGE (those hits went straight to the list)
};

// Shared memory of one CTA (dynamic; ~112 KB at 2048-row chunks, two CTAs per SM).
//  ring : selector WINDOWS, one 32-bit entry per symbol. The entry of symbol j is
//         sel(x_j) | sel(x_j+1) << 8 | ..., sel(x) = 0x10 + 0x22 x; PRMT reads bits [15:0], which
//         pick the fp16 scores of x_j and x_j+1 out of the 8-byte table row. Symbol offset
//         s = j - D0 lives at ring[(s & 31) >> 2][(s >> 5) % NSLOT] component s & 3, so the four
//         windows a thread needs for four consecutive rows are one LDS.128, and the 32 lanes of a
//         warp (32 symbols apart) read 32 consecutive uint4 -- conflict-free.
//  tab  : double-buffered score-table chunks (TMA bulk copies).
//  ckpt : the thread's 16 state registers at the start of the current superblock, for the exact
//         replay (kept in shared memory so the hot loop carries no copies of them).
//  done : warps that are through with the chunk in tab[b] (and with the TC oldest symbols of the ring). The warp that
//         makes it WARPS refills both for the chunk after next; nobody waits for it before that chunk begins.
//  stage: hit words waiting for their place in the device's hit list, per warp (see HitStage)
struct __align__(16) FastSmem {
  uint4 ring[8][NSLOT];
  uint2 tab[2][TC];
  uint4 ckpt[R / 4][THREADS];
  HitStage stage;
  unsigned long long mbar[2];  // "chunk staged": the TMA's bytes, armed by the warp that also extended the ring
  unsigned int nextTile;
  unsigned int done[2];
  int abortSeen;  // set (never cleared) by whoever polls the device's abort flag: tile prologue, chunk refill
};

// geometry of the short-table kernel (its description is at the kernel)
constexpr int SHORT_ROWS = SSV_SHORT_ROWS;                 // rows the short-table kernel accepts (multiple of 32)
constexpr int WARP_DIAGS = 32 * 2 * R;                     // 1024 diagonals per warp tile
constexpr int NS_SHORT = (WARP_DIAGS + SHORT_ROWS + 64) / 32;  // 32-symbol groups per plane of a warp's ring (never wraps)

struct __align__(16) ShortSmem {
  uint4 ring[WARPS][8][NS_SHORT];
  uint2 tab[SHORT_ROWS];
  uint4 ckpt[R / 4][THREADS];
  HitStage stage;
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

// Packed word holding local symbols [16*wi, 16*wi+16); zero ('A') outside the loaded range.
__device__ __forceinline__ uint32_t load_seq_word(const SsvParams& prm, long long wi) {
  long long nWords = prm.nLoad >> 4;
  return (wi >= 0 && wi < nWords) ? __ldg(prm.seqPacked + wi) : 0u;
}

// Four packed words (64 symbols) with one 128-bit load; wi is a multiple of 4 and so is the number
// of resident words, so the vector is either wholly inside the loaded range or wholly outside.
__device__ __forceinline__ uint4 load_seq_vec(const SsvParams& prm, long long wi) {
  const long long nWords = prm.nLoad >> 4;
  if (wi >= 0 && wi + 4 <= nWords) return __ldg(reinterpret_cast<const uint4*>(prm.seqPacked + wi));
  return make_uint4(0u, 0u, 0u, 0u);
}

// PRMT with the raw PTX semantics (CUDA's __byte_perm would first AND the selector with
// 0x7777). Only selector bits [15:0] are read; our nibbles are all <= 7 (plain byte copy).
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}

// Write the 16 window entries of the symbols in packed word `x` (symbol offsets s .. s+15 from
// the tile's first diagonal, s a multiple of 16); `xNext` is the following packed word, whose
// first symbol completes the last window.
// The selector byte of symbol value v is byte v of 0x76543210, so PRMT does the expansion itself: with the packed
// word masked to 0x3333 its selector nibbles ARE the even symbols (and shifted right by two, the odd ones): four
// PRMTs turn 16 symbols into their selector bytes, even and odd positions apart, and one more PRMT per window puts
// two neighbours side by side. Only bits [15:0] of a window are ever read (by the lookups' PRMT and by the
// replay); the upper half holds whatever the same PRMT brings along.
// `ring` points at plane 0 of a ring of `nslot` 32-symbol groups per plane (the CTA-wide ring of the streaming
// kernels, or one warp's private ring in the short-table kernel).
__device__ __forceinline__ void ring_fill16(uint4* ring, uint32_t nslot, uint32_t s, uint32_t x, uint32_t xNext) {
  constexpr uint32_t K = 0x76543210u;
  const uint32_t e0 = prmt(K, 0u, x & 0x3333u);           // selectors of symbols 0 2 4 6
  const uint32_t o0 = prmt(K, 0u, (x >> 2) & 0x3333u);    //                      1 3 5 7
  const uint32_t e1 = prmt(K, 0u, (x >> 16) & 0x3333u);   //                      8 10 12 14
  const uint32_t o1 = prmt(K, 0u, (x >> 18) & 0x3333u);   //                      9 11 13 15
  const uint32_t nx = prmt(K, 0u, xNext & 0x3u);          // first symbol of the next word (byte 0)
  const uint32_t slot = (s >> 5) % nslot;
  const uint32_t plane0 = (s & 31u) >> 2;  // 0 or 4
  uint4 v;
  // windows of symbols 0..3: (e0.0 o0.0) (o0.0 e0.1) (e0.1 o0.1) (o0.1 e0.2)
  v.x = prmt(e0, o0, 0x5140u); v.y = prmt(e0, o0, 0x2514u); v.z = prmt(e0, o0, 0x6251u); v.w = prmt(e0, o0, 0x3625u);
  ring[(plane0 + 0) * nslot + slot] = v;
  // 4..7: (e0.2 o0.2) (o0.2 e0.3) (e0.3 o0.3) (o0.3 e1.0)
  v.x = prmt(e0, o0, 0x7362u); v.y = prmt(e0, o0, 0x7736u); v.z = prmt(e0, o0, 0x7773u); v.w = prmt(o0, e1, 0x4443u);
  ring[(plane0 + 1) * nslot + slot] = v;
  // 8..11
  v.x = prmt(e1, o1, 0x5140u); v.y = prmt(e1, o1, 0x2514u); v.z = prmt(e1, o1, 0x6251u); v.w = prmt(e1, o1, 0x3625u);
  ring[(plane0 + 2) * nslot + slot] = v;
  // 12..15: the last window ends with the next word's first symbol
  v.x = prmt(e1, o1, 0x7362u); v.y = prmt(e1, o1, 0x7736u); v.z = prmt(e1, o1, 0x7773u); v.w = prmt(o1, nx, 0x4443u);
  ring[(plane0 + 3) * nslot + slot] = v;
}

// The same for 64 symbols (one 128-bit vector of the packed sequence plus the word after it).
__device__ __forceinline__ void ring_fill64(uint4* ring, uint32_t nslot, uint32_t s, uint4 v, uint32_t next) {
  ring_fill16(ring, nslot, s, v.x, v.y);
  ring_fill16(ring, nslot, s + 16u, v.y, v.z);
  ring_fill16(ring, nslot, s + 32u, v.z, v.w);
  ring_fill16(ring, nslot, s + 48u, v.w, next);
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

// Hit emission from divergent code: the lanes that are here together aggregate into one atomic
// (ballot of the active lanes + prefix popcount for the slot).
__device__ __forceinline__ void emit_hit_active(const SsvParams& prm, long long col, uint32_t row) {
  const unsigned m = __activemask();
  const unsigned lane = threadIdx.x & 31;
  const int leader = __ffs(m) - 1;
  unsigned long long base = 0;
  if ((int)lane == leader) base = atomicAdd(prm.hitCount, (unsigned long long)__popc(m));
  base = __shfl_sync(m, base, leader);
  const unsigned long long slot = base + __popc(m & ((1u << lane) - 1));
  if (slot < prm.hitCap) prm.hits[slot] = encode_hit(col, row);
  else atomicOr(prm.status, SSV_STATUS_OVERFLOW);
}

// Exact replay of one superblock (nRowsReplay rows from row pb) for the registers of this
// thread that need it, starting from the checkpoint in shared memory. It is the same fp16x2
// cell update as the fast path, one row at a time, with the overflow looked at after every row:
// 256 (s + e) rounds to +inf exactly when s + e >= 256, so "inf" IS the hit test of
// SoftSsv.cpp:44 / HavacHls.cpp:386, and the half that hit is reset to 0 as the reference does.
// Lanes and registers with nothing to do skip; no warp-synchronous call is made in here.
//   infMask : bit k set = the caller's register k holds an inf after the fast path
//   edge : the tile still has diagonals left of column 0 in this superblock. Registers wholly
//          left of it are zeroed (nothing real yet: prev score is 0 at column 0, SoftSsv.cpp:37),
//          registers crossing it are replayed with the cells left of column 0 held at 0.
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint2 lds_v2(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}

// A register whose fp16 halves just reached inf at row p: report the hit(s), reset those halves
// to 0 as the reference does (SoftSsv.cpp:44-49), return the register. `stage` = shared-space address of the warp's
// hit stage (HitStage::hit[warp]); its counter sits at `stageCount`.
__device__ __noinline__ uint32_t on_hit(const SsvParams& prm, uint32_t v, long long dlo, uint32_t p, uint32_t stage, uint32_t stageCount) {
#pragma unroll 1
  for (int h = 0; h < 2; h++) {
    if (((v >> (16 * h)) & 0xffffu) != 0x7c00u) continue;
    v &= h ? 0x0000ffffu : 0xffff0000u;
    const long long col = dlo + h + (long long)p + prm.loadBegin;
    if (p < prm.nRows && col >= prm.colBegin && col < prm.colEnd) {
      uint32_t place;
      asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(place) : "r"(stageCount) : "memory");
      if (place < (uint32_t)HSTAGE) {
        const unsigned long long word = encode_hit(col, p);
        asm volatile("st.shared.u64 [%0], %1;" ::"r"(stage + 8u * place), "l"(word) : "memory");
      } else {
        emit_hit_active(prm, col, p);  // stage full (a warp-block with more than 120 hits): straight to the list
      }
    }
  }
  return v;
}

// Moves the warp's staged hits into the device's hit list: one atomic for all of them, coalesced stores. Called by
// the whole warp (converged).
__device__ __noinline__ void flush_hits(const SsvParams& prm, HitStage& st) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncwarp();
  unsigned n = *(volatile unsigned*)&st.count[warp];
  if (n == 0) return;
  if (n > (unsigned)HSTAGE) n = HSTAGE;
  unsigned long long base = 0;
  if (lane == 0) base = atomicAdd(prm.hitCount, (unsigned long long)n);
  base = __shfl_sync(FULL, base, 0);
  for (unsigned i = lane; i < n; i += 32) {
    if (base + i < prm.hitCap) prm.hits[base + i] = st.hit[warp][i];
    else atomicOr(prm.status, SSV_STATUS_OVERFLOW);
  }
  __syncwarp();
  if (lane == 0) *(volatile unsigned*)&st.count[warp] = 0;
  __syncwarp();
}

// SHORT: called from the short-table kernel (the warp's private ring, the whole table at `tab`) or from the streaming
// kernels (the CTA's ring, `tab` = the staged chunk). The shared-memory addresses are found from the kernel family's
// layout in here, so that the call costs the hot loop's register allocation no more than its six arguments.
extern __shared__ __align__(16) unsigned char smemRaw[];
template <bool SHORT>
__device__ __noinline__ uint32_t exact_rows(const SsvParams& prm, const uint2* tab, uint32_t infMask, bool edge,
                                            long long d0, long long D0, uint32_t pb, uint32_t nRowsReplay) {
  struct Geom { uint32_t ckBase, ringBase, nslot, tabBase, tabMask, stage, stageCount; } geo;
  if constexpr (SHORT) {
    ShortSmem& sm = *reinterpret_cast<ShortSmem*>(smemRaw);
    geo = {smem_u32(&sm.ckpt[0][0]), smem_u32(&sm.ring[threadIdx.x >> 5][0][0]), (uint32_t)NS_SHORT, smem_u32(tab), 0xffffffffu,
           smem_u32(&sm.stage.hit[threadIdx.x >> 5][0]), smem_u32(&sm.stage.count[threadIdx.x >> 5])};
  } else {
    FastSmem& sm = *reinterpret_cast<FastSmem*>(smemRaw);
    geo = {smem_u32(&sm.ckpt[0][0]), smem_u32(&sm.ring[0][0]), (uint32_t)NSLOT, smem_u32(tab), (uint32_t)TC - 1u,
           smem_u32(&sm.stage.hit[threadIdx.x >> 5][0]), smem_u32(&sm.stage.count[threadIdx.x >> 5])};
  }
  // Registers do not cross this call: the caller says which of its registers hold an inf
  // (infMask), the exact end state of every replayed register is written over its checkpoint
  // slot, and the returned mask tells the caller which slots to read back. Shared memory is read
  // with shared-space addresses (a generic pointer would cost an address translation per load).
  const uint32_t ckBase = geo.ckBase, ringBase = geo.ringBase, tabBase = geo.tabBase, NSL = geo.nslot;
  uint32_t changed = 0;
#pragma unroll 1
  for (uint32_t todo = edge ? 0xffffu : infMask; todo; todo &= todo - 1) {
    const int k = __ffs(todo) - 1;
    const uint32_t slot = ckBase + 4u * (uint32_t)(((k >> 2) * THREADS + threadIdx.x) * 4 + (k & 3));
    const long long dlo = d0 + 2 * k;  // diagonals dlo (low half) and dlo + 1 (high half)
    bool crossing = false;
    if (edge) {
      if (dlo + 1 + (long long)pb + nRowsReplay - 1 < 0) {  // wholly left of column 0
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(slot), "r"(0u) : "memory");
        changed |= 1u << k;
        continue;
      }
      crossing = dlo + (long long)pb < 0;
    }
    if (!crossing && !((infMask >> k) & 1u)) continue;
    uint32_t v = lds_u32(slot);
    const uint32_t s0 = (uint32_t)(dlo + (long long)pb - D0);  // ring offset of the low half's symbol at row pb
    if (crossing) {
      // first tiles only: row by row, with the cells left of column 0 held at 0. Nothing is real before the high
      // half's diagonal (dlo + 1) reaches column 0, at row -(dlo + 1): the walk starts there, from 0.
      v = 0u;
#pragma unroll 4
      for (uint32_t r = (uint32_t)(-(dlo + 1) - (long long)pb); r < nRowsReplay; r++) {
        const uint32_t sy = s0 + r, p = pb + r;
        const uint32_t sel = lds_u32(ringBase + (((sy & 31u) >> 2) * NSL + (sy >> 5) % NSL) * 16u + (sy & 3u) * 4u);
        const uint2 e2 = lds_v2(tabBase + (p & geo.tabMask) * 8u);
        v = hfma2_relu_add(prmt(e2.x, e2.y, sel), v);
        const long long j = dlo + (long long)p;
        if (j < 0) v &= 0xffff0000u;
        if (j + 1 < 0) v = 0u;
        if (has_inf(v)) v = on_hit(prm, v, dlo, p, geo.stage, geo.stageCount);
      }
    } else {
      // Sixteen rows at a time: the lookups do not depend on the state, so their loads pipeline,
      // and the chain is one HFMA2 per row. The sixteen are first run without looking; only if
      // they end in inf are they walked again, from the same looked-up values, one row at a time.
      // nRowsReplay is a multiple of 32.
#pragma unroll 1
      for (uint32_t rb = 0; rb < nRowsReplay; rb += 16) {
        // s0 is even (32 per thread, 2 per register, 32 per block) and so is the row: two consecutive symbols' windows
        // are one 64-bit half of a ring entry, two consecutive table rows one 128-bit load
        const uint32_t sA = s0 + rb, qA = sA >> 5;
        const uint32_t slotA = qA % NSL, slotB = slotA + 1 == NSL ? 0u : slotA + 1;
        uint32_t m[16];
#pragma unroll
        for (uint32_t i = 0; i < 16; i += 2) {
          const uint32_t sy = sA + i;
          const uint32_t sl = (sy >> 5) == qA ? slotA : slotB;
          const uint2 sel2 = lds_v2(ringBase + (((sy & 31u) >> 2) * NSL + sl) * 16u + (sy & 2u) * 4u);
          const uint4 e4 = lds_v4(tabBase + ((pb + rb + i) & geo.tabMask) * 8u);
          m[i] = prmt(e4.x, e4.y, sel2.x);
          m[i + 1] = prmt(e4.z, e4.w, sel2.y);
        }
        uint32_t vv = v;
#pragma unroll
        for (uint32_t i = 0; i < 16; i++) vv = hfma2_relu_add(m[i], vv);
        if (!has_inf(vv)) {
          v = vv;
          continue;
        }
#pragma unroll
        for (uint32_t i = 0; i < 16; i++) {
          v = hfma2_relu_add(m[i], v);
          if (has_inf(v)) v = on_hit(prm, v, dlo, pb + rb + i, geo.stage, geo.stageCount);
        }
      }
    }
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(slot), "r"(v) : "memory");
    changed |= 1u << k;
  }
  return changed;
}

// One 4-row group of the fast path. U = group index inside the 32-row block (compile time),
// so every register index below is static after unrolling.
//   rows 4U+0..3 -> (q, c) = (2U,0) (2U,1) (2U+1,0) (2U+1,1)
//   register k at row r needs the window of symbol d0 + 2k + r: even rows read e0, odd rows e1,
//   and the slot a row just finished with is refilled with the window 32 symbols further on.
// NS = 32-symbol groups per plane of the ring the windows come from (compile time: it is an address stride)
template <int U, int NS>
__device__ __forceinline__ void four_rows(uint32_t (&w)[R], uint32_t (&e0)[R], uint32_t (&e1)[R],
                                          const uint2* __restrict__ tabRow,
                                          const uint4* __restrict__ ringSlot) {
  // The four new windows are fetched as two 64-bit halves: a half can be loaded IN PLACE over the two windows that
  // die in rows 4U and 4U+1 (resp. 4U+2 and 4U+3) two rows before its first use, which is what lets the block
  // scheduler (havac_b200/sass/reblock.py) keep every window in one register for the whole block.
  // (ld.volatile: ptxas fuses two adjacent plain 64-bit loads, inline asm or not, back into one LDS.128.)
  const uint32_t ringAddr = smem_u32(ringSlot + U * NS);
  uint2 win01, win23;  // windows of the next block: rows 4U, 4U+1 and rows 4U+2, 4U+3
  asm volatile("ld.volatile.shared.v2.u32 {%0, %1}, [%2];" : "=r"(win01.x), "=r"(win01.y) : "r"(ringAddr));
  asm volatile("ld.volatile.shared.v2.u32 {%0, %1}, [%2+8];" : "=r"(win23.x), "=r"(win23.y) : "r"(ringAddr));
  const uint4 ab01 = *reinterpret_cast<const uint4*>(tabRow + 4 * U);      // rows 4U, 4U+1
  const uint4 ab23 = *reinterpret_cast<const uint4*>(tabRow + 4 * U + 2);  // rows 4U+2, 4U+3
  constexpr int Q0 = 2 * U, Q1 = 2 * U + 1;
#pragma unroll
  for (int k = 0; k < R; k++) w[k] = hfma2_relu_add(prmt(ab01.x, ab01.y, e0[(Q0 + k) % R]), w[k]);
  e0[Q0 % R] = win01.x;
#pragma unroll
  for (int k = 0; k < R; k++) w[k] = hfma2_relu_add(prmt(ab01.z, ab01.w, e1[(Q0 + k) % R]), w[k]);
  e1[Q0 % R] = win01.y;
#pragma unroll
  for (int k = 0; k < R; k++) w[k] = hfma2_relu_add(prmt(ab23.x, ab23.y, e0[(Q1 + k) % R]), w[k]);
  e0[Q1 % R] = win23.x;
#pragma unroll
  for (int k = 0; k < R; k++) w[k] = hfma2_relu_add(prmt(ab23.z, ab23.w, e1[(Q1 + k) % R]), w[k]);
  e1[Q1 % R] = win23.y;
}

template <int U, int NS>
struct RowGroups {
  static __device__ __forceinline__ void run(uint32_t (&w)[R], uint32_t (&e0)[R], uint32_t (&e1)[R],
                                             const uint2* tabRow, const uint4* ringSlot) {
    four_rows<U, NS>(w, e0, e1, tabRow, ringSlot);
    RowGroups<U + 1, NS>::run(w, e0, e1, tabRow, ringSlot);
  }
};
template <int NS>
struct RowGroups<BLOCK_ROWS / 4, NS> {
  static __device__ __forceinline__ void run(uint32_t (&)[R], uint32_t (&)[R], uint32_t (&)[R],
                                             const uint2*, const uint4*) {}
};

// The first END groups of a block only: the last block of a tile's rows when the table does not
// end on a 32-row boundary (a 200-row model would otherwise pay for 224 rows). Nothing runs after
// it in the tile, so the window registers it leaves half rotated are never used again.
template <int U, int END, int NS>
struct RowGroupsTo {
  static __device__ __forceinline__ void run(uint32_t (&w)[R], uint32_t (&e0)[R], uint32_t (&e1)[R],
                                             const uint2* tabRow, const uint4* ringSlot) {
    four_rows<U, NS>(w, e0, e1, tabRow, ringSlot);
    RowGroupsTo<U + 1, END, NS>::run(w, e0, e1, tabRow, ringSlot);
  }
};
template <int END, int NS>
struct RowGroupsTo<END, END, NS> {
  static __device__ __forceinline__ void run(uint32_t (&)[R], uint32_t (&)[R], uint32_t (&)[R],
                                             const uint2*, const uint4*) {}
};

// TAIL = the build for short tables, where a partial last block is worth a branch per block; the
// plain build keeps the hot loop exactly as it was measured (the branch cost 2 % at 35 k rows).
template <bool TAIL>
__device__ __forceinline__ void ssv_f16x2_body(const SsvParams& prm) {
  FastSmem& sm = *reinterpret_cast<FastSmem*>(smemRaw);
  const int tid = threadIdx.x;
  uint32_t mbarPhase = 0;  // bit b = parity to wait for on sm.mbar[b]
  const long long clk0 = clock64();

  if (tid == 0) {
    mbar_init(&sm.mbar[0], 1);
    mbar_init(&sm.mbar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    sm.abortSeen = 0;
    sm.nextTile = atomicAdd(prm.tileCounter, 1u);
  }
  if ((tid & 31) == 0) sm.stage.count[tid >> 5] = 0;
  __syncthreads();
  bool aborted = false;

  while (true) {
    const unsigned ticket = sm.nextTile;
    __syncthreads();  // everyone has read nextTile before thread 0 overwrites it
    if (ticket >= prm.nTiles) break;
    // Tiles are handed out left to right: the tiles that start left of column 0 (an unsharded run has rows/W of
    // them) go first. They have fewer rows than the others but the warp the column-0 boundary runs through walks
    // its crossing registers row by row there, so they are the SLOW ones, and a slow tile handed out last is the
    // kernel's tail (profiles/r2_ncu_full_f16x2.txt before this: SMs active 218.0 M of 226.9 M elapsed cycles).
    const unsigned tile = ticket;

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
    // rows (from row 0) in which the tile still has diagonals left of column 0: D0 + row < 0
    // ... seen from this WARP: its first diagonal is D0 + 1024 warp, so it is clear of column 0 from row -(that) on.
    // (The warps to the right of the one the boundary runs through never take the edge path.)
    const long long dWarp = D0 + (long long)(tid & ~31) * (2 * R);
    const uint32_t edgeRows = dWarp < 0 ? (uint32_t)(-dWarp < 0x7fffffffll ? -dWarp : 0x7fffffffll) : 0u;
    if (rowEnd <= rowBegin) {                       // nothing to do (defensive)
      if (tid == 0) sm.nextTile = atomicAdd(prm.tileCounter, 1u);
      __syncthreads();
      continue;
    }

    // --- prologue: queue the first two table chunks, fill the ring, load the windows --------
    if (tid == 0) {
      sm.nextTile = atomicAdd(prm.tileCounter, 1u);
      if (*prm.abortFlag) sm.abortSeen = 1;
      sm.done[0] = sm.done[1] = 0;
      for (int i = 0; i < 2 && c0 + i < c1; i++) {
        mbar_expect_tx(&sm.mbar[i], TC * sizeof(uint2));
        tma_load_1d(sm.tab[i], prm.tab16 + (size_t)(c0 + i) * TC, TC * sizeof(uint2), &sm.mbar[i]);
      }
    }
    const uint32_t p0 = (uint32_t)c0 * TC;  // first row; ring holds offsets [p0, p0 + W + 2 TC)
    {
      const long long wFill0 = (D0 + (long long)p0) >> 4;  // D0 and p0 are multiples of 64
      // a short table needs only the symbols its rows reach (+ one block of window prefetch)
      long long need = (long long)W + ((rowEnd - (long long)p0 + 63) & ~63ll) + 64;
      if (need > W + 2 * TC) need = W + 2 * TC;
      for (int i = tid; i < (int)(need / 64); i += THREADS)
        ring_fill64(&sm.ring[0][0], NSLOT, p0 + 64u * (uint32_t)i, load_seq_vec(prm, wFill0 + 4 * i), load_seq_word(prm, wFill0 + 4 * i + 4));
    }
    __syncthreads();
    aborted = *(volatile int*)&sm.abortSeen != 0;  // CTA-uniform: written before the barrier above
    if (aborted) {  // drain the table chunks queued above, then leave
      for (int i = 0; i < 2 && c0 + i < c1; i++) mbar_wait(&sm.mbar[i], (mbarPhase >> i) & 1u);
      break;
    }

    uint32_t w[R], e0[R], e1[R];
    uint32_t g = ((uint32_t)tid + p0 / 32u) % (uint32_t)NSLOT;  // slot of symbol offset 32 tid + p0
#pragma unroll
    for (int u = 0; u < R / 2; u++) {
      const uint4 v = sm.ring[u][g];
      e0[2 * u] = v.x;
      e1[2 * u] = v.y;
      e0[2 * u + 1] = v.z;
      e1[2 * u + 1] = v.w;
    }
#pragma unroll
    for (int k = 0; k < R; k++) w[k] = 0u;
    g = (g + 1 == (uint32_t)NSLOT) ? 0u : g + 1;  // slot the first block loads its new windows from

    // --- main loop over row chunks -----------------------------------------------------------
    // The warps of a CTA do NOT meet between chunks. A warp waits for "chunk staged" (the TMA's bytes and the ring
    // extension that goes with them), runs the chunk's blocks, and counts itself off the chunk; the warp that counts
    // last -- the one everybody would have waited for at a barrier -- extends the ring and queues the table for the
    // chunk after next, which re-use what the finished chunk occupied. So a warp can be up to two chunks ahead of the
    // slowest one before it has to wait, and what a warp loses to replays in one chunk it can make up in the next.
    for (int c = c0; c < c1; c++) {
      const int buf = (c - c0) & 1;
      mbar_wait(&sm.mbar[buf], (mbarPhase >> buf) & 1u);
      mbarPhase ^= 1u << buf;
      if (*(volatile int*)&sm.abortSeen != 0) {  // a refill found the abort flag up: this chunk was never staged
        aborted = true;
        break;
      }

      // Per block of 32 rows: checkpoint, the block, one look at the states. Everything around the block is kept in
      // 32-bit registers and as short as it can be: these instructions queue for the same (saturated) ALU pipe as the
      // block's lookups, so each one of them costs two of its clocks.
      const uint32_t pbEnd = (long long)(c + 1) * TC < rowEnd ? (uint32_t)(c + 1) * TC : (uint32_t)rowEnd;
      const uint2* tabRow = sm.tab[buf];
      for (uint32_t pb = (uint32_t)c * TC; pb < pbEnd; pb += BLOCK_ROWS) {
#pragma unroll
        for (int q = 0; q < R / 4; q++)
          sm.ckpt[q][tid] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
#ifdef SSV_DEBUG_SPLIT_HEAD
        __syncwarp();
#endif
        if constexpr (TAIL) {
          const uint32_t left = pbEnd - pb;  // rows of the tile from this block on
          if (left > 24) RowGroups<0, NSLOT>::run(w, e0, e1, tabRow, &sm.ring[0][g]);
          else if (left > 16) RowGroupsTo<0, 6, NSLOT>::run(w, e0, e1, tabRow, &sm.ring[0][g]);
          else if (left > 8) RowGroupsTo<0, 4, NSLOT>::run(w, e0, e1, tabRow, &sm.ring[0][g]);
          else RowGroupsTo<0, 2, NSLOT>::run(w, e0, e1, tabRow, &sm.ring[0][g]);
        } else {
          RowGroups<0, NSLOT>::run(w, e0, e1, tabRow, &sm.ring[0][g]);
        }
        tabRow += BLOCK_ROWS;
        g = (g + 1 == (uint32_t)NSLOT) ? 0u : g + 1;
#ifdef SSV_DEBUG_SPLIT_TAIL
        __syncwarp();
#endif

        // post-pass: hit replay, and the band left of column 0 in the first tiles
        uint32_t mx = 0;
#pragma unroll
        for (int k = 0; k < R; k += 2) mx = __vimax3_s16x2(mx, w[k], w[k + 1]);
        const bool edge = pb < edgeRows;  // warp-uniform: this warp still has diagonals left of column 0
        if (edge || __any_sync(FULL, has_inf(mx))) {
          // cold path, out of line: the replay leaves the exact states in the checkpoint slots
          uint32_t infMask = 0;
#pragma unroll
          for (int k = 0; k < R; k++) infMask |= has_inf(w[k]) ? (1u << k) : 0u;
          const uint32_t changed = exact_rows<false>(prm, sm.tab[buf], infMask, edge, d0, D0, pb, BLOCK_ROWS);
          const uint32_t* ck = reinterpret_cast<const uint32_t*>(&sm.ckpt[0][0]);
#pragma unroll
          for (int k = 0; k < R; k++)
            if ((changed >> k) & 1u) w[k] = ck[((k >> 2) * THREADS + tid) * 4 + (k & 3)];
          __syncwarp();
          if (*(volatile unsigned*)&sm.stage.count[tid >> 5] >= (unsigned)(HSTAGE / 2)) flush_hits(prm, sm.stage);
        }
      }

      if (c + 2 < c1) {
        // through with tab[buf] and with ring offsets [c TC, (c + 1) TC): count this warp off
        __syncwarp();
        unsigned last = 0;
        if ((tid & 31) == 0) {
          __threadfence_block();  // this warp's reads, before the count that lets another warp overwrite them
          last = atomicAdd(&sm.done[buf], 1u) == (unsigned)(THREADS / 32 - 1) ? 1u : 0u;
          __threadfence_block();
        }
        last = __shfl_sync(FULL, last, 0);
        if (last) {
          // the ring's windows of offsets [W + (c + 2) TC, W + (c + 3) TC), straight from the packed sequence ...
          const uint32_t sRefill = (uint32_t)W + (uint32_t)(c + 2) * TC;
          const long long wFill = (D0 + (long long)sRefill) >> 4;
          for (int i = tid & 31; i < TC / 64; i += 32)
            ring_fill64(&sm.ring[0][0], NSLOT, sRefill + 64u * (uint32_t)i, load_seq_vec(prm, wFill + 4 * i), load_seq_word(prm, wFill + 4 * i + 4));
          __syncwarp();
          // ... then the table chunk; its arrival on the barrier publishes the windows with it (release / acquire).
          // With the abort flag up nothing is queued -- no TMA may be in flight when the CTA exits -- and the waiters
          // are let through to find abortSeen.
          if ((tid & 31) == 0) {
            sm.done[buf] = 0;
            if (*prm.abortFlag) {
              *(volatile int*)&sm.abortSeen = 1;
              asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&sm.mbar[buf])) : "memory");
            } else {
              mbar_expect_tx(&sm.mbar[buf], TC * sizeof(uint2));
              tma_load_1d(sm.tab[buf], prm.tab16 + (size_t)(c + 2) * TC, TC * sizeof(uint2), &sm.mbar[buf]);
            }
          }
        }
      }
    }
    __syncthreads();  // the next tile's prologue rewrites the ring and both table buffers
    aborted = *(volatile int*)&sm.abortSeen != 0;  // CTA-uniform after the barrier (a warp may have finished the tile without seeing it)
    if (aborted) break;
  }
  flush_hits(prm, sm.stage);
  if (tid == 0 && aborted) atomicOr(prm.status, SSV_STATUS_ABORTED);
  if (tid == 0 && blockIdx.x == 0) *prm.smCycles = (unsigned long long)(clock64() - clk0);
}

// =======================================================================================
// Short tables (at most SHORT_ROWS rows): every WARP is on its own
// =======================================================================================
// A streaming tile costs two CTA-wide barriers and a dependent chain global load -> window expansion -> barrier ->
// window loads before its first block, whatever the number of rows: at 72 rows the warps of the streaming kernel
// spend 22 % of their time at those barriers and another 9 % waiting for the loads (profiles/r2_ncu_rows72.txt).
// A table of at most 512 rows fits in shared memory whole (4 KB), so nothing has to be streamed and nothing
// has to be shared between warps but that table: here a warp takes 1024 diagonals of its own from the tile counter,
// fills a PRIVATE ring of selector windows (1024 + rows + 64 symbols), runs the same 32-row blocks over all rows and
// fetches the next 1024 diagonals -- no barrier after the table has been loaded, so the warps of an SM drift apart
// and one warp's load latency is another warp's compute.
__global__ void __launch_bounds__(THREADS, 512 / THREADS) ssv_f16x2_short_kernel(const SsvParams prm) {
  ShortSmem& sm = *reinterpret_cast<ShortSmem*>(smemRaw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long clk0 = clock64();
  // the whole table, once per CTA (pad rows of the device table are -128: they can never hit)
  const uint32_t rowsPad = (prm.nRows + BLOCK_ROWS - 1) / BLOCK_ROWS * BLOCK_ROWS;
  for (uint32_t r = tid; r < rowsPad; r += THREADS) sm.tab[r] = __ldg(prm.tab16 + r);
  if (lane == 0) sm.stage.count[warp] = 0;
  __syncthreads();

  uint4* ring = &sm.ring[warp][0][0];
  const uint32_t nTilesW = prm.nTiles * (uint32_t)(W / WARP_DIAGS);
  bool aborted = false;
  while (true) {
    unsigned ticket = 0;
    int stop = 0;
    if (lane == 0) {
      ticket = atomicAdd(prm.tileCounter, 1u);
      stop = *prm.abortFlag;
    }
    ticket = __shfl_sync(FULL, ticket, 0);
    aborted = __shfl_sync(FULL, stop, 0) != 0;
    if (ticket >= nTilesW || aborted) break;
    const unsigned tile = ticket;  // left to right, as in the streaming kernels: the column-0 tiles first
    const long long D0 = prm.diagBegin + (long long)tile * WARP_DIAGS;
    const long long d0 = D0 + (long long)lane * (2 * R);
    long long rowEnd = prm.nLoad - D0;
    if (rowEnd > (long long)prm.nRows) rowEnd = prm.nRows;
    if (rowEnd <= 0 || D0 + WARP_DIAGS - 1 + (long long)prm.nRows - 1 < 0) continue;  // nothing real in this band
    const uint32_t edgeRows = D0 < 0 ? (uint32_t)(-D0 < 0x7fffffffll ? -D0 : 0x7fffffffll) : 0u;
    const uint32_t pbEnd = (uint32_t)rowEnd;

    // ring: the windows of symbols [D0, D0 + 1024 + rows + 64), 64 per lane and pass
    {
      const long long wFill0 = D0 >> 4;  // D0 is a multiple of 64
      const int n64 = (WARP_DIAGS + (int)((pbEnd + 63u) & ~63u) + 64) / 64;
      for (int i = lane; i < n64; i += 32)
        ring_fill64(ring, NS_SHORT, 64u * (uint32_t)i, load_seq_vec(prm, wFill0 + 4 * i), load_seq_word(prm, wFill0 + 4 * i + 4));
    }
    __syncwarp();

    uint32_t w[R], e0[R], e1[R];
    uint32_t g = (uint32_t)lane;  // slot of symbol offset 32 lane
#pragma unroll
    for (int u = 0; u < R / 2; u++) {
      const uint4 v = ring[u * NS_SHORT + g];
      e0[2 * u] = v.x;
      e1[2 * u] = v.y;
      e0[2 * u + 1] = v.z;
      e1[2 * u + 1] = v.w;
    }
#pragma unroll
    for (int k = 0; k < R; k++) w[k] = 0u;
    g += 1;

    const uint2* tabRow = sm.tab;
    for (uint32_t pb = 0; pb < pbEnd; pb += BLOCK_ROWS) {
#pragma unroll
      for (int q = 0; q < R / 4; q++)
        sm.ckpt[q][tid] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
      const uint32_t left = pbEnd - pb;  // rows of the tile from this block on
      if (left > 24) RowGroups<0, NS_SHORT>::run(w, e0, e1, tabRow, ring + g);
      else if (left > 16) RowGroupsTo<0, 6, NS_SHORT>::run(w, e0, e1, tabRow, ring + g);
      else if (left > 8) RowGroupsTo<0, 4, NS_SHORT>::run(w, e0, e1, tabRow, ring + g);
      else RowGroupsTo<0, 2, NS_SHORT>::run(w, e0, e1, tabRow, ring + g);
      tabRow += BLOCK_ROWS;
      g += 1;  // 32 + rows / 32 + 1 groups at most: the private ring never wraps

      uint32_t mx = 0;
#pragma unroll
      for (int k = 0; k < R; k += 2) mx = __vimax3_s16x2(mx, w[k], w[k + 1]);
      const bool edge = pb < edgeRows;
      if (edge || __any_sync(FULL, has_inf(mx))) {
        uint32_t infMask = 0;
#pragma unroll
        for (int k = 0; k < R; k++) infMask |= has_inf(w[k]) ? (1u << k) : 0u;
        const uint32_t changed = exact_rows<true>(prm, sm.tab, infMask, edge, d0, D0, pb, BLOCK_ROWS);
        const uint32_t* ck = reinterpret_cast<const uint32_t*>(&sm.ckpt[0][0]);
#pragma unroll
        for (int k = 0; k < R; k++)
          if ((changed >> k) & 1u) w[k] = ck[((k >> 2) * THREADS + tid) * 4 + (k & 3)];
        __syncwarp();
        if (*(volatile unsigned*)&sm.stage.count[warp] >= (unsigned)(HSTAGE / 2)) flush_hits(prm, sm.stage);
      }
    }
    __syncwarp();  // every lane has read its last windows before the ring is refilled
  }
  flush_hits(prm, sm.stage);
  if (lane == 0 && aborted) atomicOr(prm.status, SSV_STATUS_ABORTED);
  if (tid == 0 && blockIdx.x == 0) *prm.smCycles = (unsigned long long)(clock64() - clk0);
}

__global__ void __launch_bounds__(THREADS, 512 / THREADS) ssv_f16x2_kernel(const SsvParams prm) {
  ssv_f16x2_body<false>(prm);
}
// Short tables that end 1..24 rows into their last block (chosen in ssv_launch_f16x2).
__global__ void __launch_bounds__(THREADS, 512 / THREADS) ssv_f16x2_tail_kernel(const SsvParams prm) {
  ssv_f16x2_body<true>(prm);
}

// Every (previous score, emission) pair through the fast path's own cell update, for the exhaustive parity
// test against the FPGA cell (device/HavacHls.cpp:370-402): thread i takes pair i in the low fp16 half and pair
// 65535 - i in the high half, so both halves and their independence are covered in one launch.
__global__ void cell_selftest_kernel(uint8_t* __restrict__ outScore, uint8_t* __restrict__ outHit) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 65536u) return;
  const uint32_t j = 65535u - i;
  auto enc = [](int v) { return (uint32_t)__half_as_ushort(__int2half_rn(v * 256)); };
  const uint32_t w = enc((int)(i >> 8)) | (enc((int)(j >> 8)) << 16);
  const uint32_t m = enc((int)(i & 255u) - 128) | (enc((int)(j & 255u) - 128) << 16);
  const uint32_t r = hfma2_relu_add(m, w);
  const uint32_t lo = r & 0xffffu, hi = r >> 16;
  const bool hitLo = lo == 0x7c00u, hitHi = hi == 0x7c00u;
  // the kernel's own test sees an inf exactly where a half is 0x7c00
  const bool seen = has_inf(r);
  const int sLo = hitLo ? 0 : __half2int_rn(__ushort_as_half((unsigned short)lo)) / 256;
  const int sHi = hitHi ? 0 : __half2int_rn(__ushort_as_half((unsigned short)hi)) / 256;
  // low half reports pair i; the high half's result must equal what thread j's low half reports: written to
  // a second plane so the host can compare the two
  outScore[i] = (uint8_t)sLo;
  outHit[i] = (uint8_t)((hitLo ? 1 : 0) | ((seen == (hitLo || hitHi)) ? 0 : 0x80));
  outScore[65536u + j] = (uint8_t)sHi;
  outHit[65536u + j] = (uint8_t)(hitHi ? 1 : 0);
}

}  // namespace

// =======================================================================================
// launchers
// =======================================================================================
void ssv_launch_trace(const SsvParams& prm, uint32_t rowBegin, uint32_t rowEnd, long long colBegin, long long colEnd,
                      SsvCellRecord* out, void* stream) {
  const long long nDiag = (colEnd - colBegin) + (long long)(rowEnd - rowBegin) - 1;
  long long blocks = (nDiag + 127) / 128;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  ssv_trace_kernel<<<(int)blocks, 128, 0, (cudaStream_t)stream>>>(prm, rowBegin, rowEnd, colBegin, colEnd, out);
}

void ssv_launch_cell_selftest(uint8_t* outScore, uint8_t* outHit, void* stream) {
  cell_selftest_kernel<<<256, 256, 0, (cudaStream_t)stream>>>(outScore, outHit);
}

void ssv_launch_table_prep(const int8_t* tab8, uint32_t nRows, uint2* tab16, uint32_t rowsPadded,
                           void* stream) {
  table_prep_kernel<<<(rowsPadded + 255) / 256, 256, 0, (cudaStream_t)stream>>>(tab8, nRows, tab16,
                                                                              rowsPadded);
}

void ssv_launch_project_table(const float* emissions, uint32_t nRows, const uint32_t* modelFirstRow, const float* alphaBeta,
                              uint32_t nModels, int8_t* tab8, void* stream) {
  project_table_kernel<<<(nRows + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4*>(emissions), nRows, modelFirstRow, reinterpret_cast<const float2*>(alphaBeta), nModels,
      reinterpret_cast<uint32_t*>(tab8));
}

// The kernels need more than the 48 KB of shared memory a launch gets by default. Function attributes are per
// DEVICE, so they are set (and the occupancy asked for) once per device ordinal of this process and remembered:
// a search that lasts microseconds must not pay five runtime calls per launch.
int ssv_f16x2_ctas_per_sm() {
  constexpr int MAX_DEV = 64;
  static std::atomic<int> cached[MAX_DEV];  // 0 = not asked yet
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < MAX_DEV) {
    const int have = cached[dev].load(std::memory_order_acquire);
    if (have > 0) return have;
  }
  cudaError_t e = cudaSuccess;
  if (e == cudaSuccess) e = cudaFuncSetAttribute(ssv_f16x2_short_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ShortSmem));
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(ssv_f16x2_short_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  for (auto kernel : {ssv_f16x2_kernel, ssv_f16x2_tail_kernel}) {
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FastSmem));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  }
  int n = 0;
  if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, ssv_f16x2_kernel, THREADS, sizeof(FastSmem));
  if (e != cudaSuccess || n < 1) return 1;  // not remembered: the launch that follows reports the error
  if (dev >= 0 && dev < MAX_DEV) cached[dev].store(n, std::memory_order_release);
  return n;
}

bool ssv_takes_short_kernel(uint32_t nRows, bool streamingOnly) { return nRows <= (uint32_t)SHORT_ROWS && !streamingOnly; }

void ssv_launch_f16x2(const SsvParams& prm, int numSMs, void* stream, bool streamingOnly) {
  int grid = numSMs * ssv_f16x2_ctas_per_sm();
  if (ssv_takes_short_kernel(prm.nRows, streamingOnly)) {
    // whole table in shared memory, warps on their own (the grid never exceeds the 8192-diagonal tiles there are,
    // each of which is eight warp tiles)
    if ((unsigned)grid > prm.nTiles) grid = (int)prm.nTiles;
    if (grid < 1) grid = 1;
    ssv_f16x2_short_kernel<<<grid, THREADS, sizeof(ShortSmem), (cudaStream_t)stream>>>(prm);
    return;
  }
  if ((unsigned)grid > prm.nTiles) grid = (int)prm.nTiles;
  if (grid < 1) grid = 1;
  // A table that ends 1..24 rows into its last 32-row block can skip the 8, 16 or 24 rows it does not
  // have, at the price of the tail build's hot loop (a branch per block: 2 % slower, measured). It is
  // taken where the rows saved are at least 1/32 of the table: 72 rows +20 %, 200 rows +7.6 %,
  // 1000 rows break-even (plain), 2040 rows would lose 1.4 % (plain).
  const uint32_t tailRows = prm.nRows % BLOCK_ROWS;
  const uint32_t saved = tailRows ? BLOCK_ROWS - ((tailRows + 7u) & ~7u) : 0u;
  const uint32_t padded = (prm.nRows + BLOCK_ROWS - 1) / BLOCK_ROWS * BLOCK_ROWS;
  if (saved != 0 && saved * 32u >= padded)
    ssv_f16x2_tail_kernel<<<grid, THREADS, sizeof(FastSmem), (cudaStream_t)stream>>>(prm);
  else
    ssv_f16x2_kernel<<<grid, THREADS, sizeof(FastSmem), (cudaStream_t)stream>>>(prm);
}

void ssv_launch_scalar(const SsvParams& prm, void* stream) {
  long long nDiag = prm.nLoad + (long long)prm.nRows - 1;
  long long blocks = (nDiag + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  ssv_scalar_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(prm);
}
