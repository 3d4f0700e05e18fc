// ssv_kernels.cuh -- launch parameters shared by the SSV kernels and the C-ABI client.
//
// Geometry (all "local" to one device shard):
//   symbols  j in [0, nLoad)   = global columns [loadBegin, loadBegin + nLoad)
//   rows     p in [0, nRows)   = concatenated model positions (no resets between models,
//                                test/softSsv/SoftSsv.cpp:31-62, device/HavacHls.cpp:321-360)
//   diagonal d = j - p         : S(p, j) depends only on S(p-1, j-1), so every diagonal is an
//                                independent byte-wide scan (device/HavacHls.cpp:370-402).
// A tile is TILE_DIAGS consecutive diagonals run over all rows; tiles are handed out to
// persistent CTAs through an atomic counter.
#pragma once
#include <cstdint>

struct SsvParams {
  const uint32_t* seqPacked;  // 2-bit packed symbols, 16 per word, nLoad/16 words (+ slack)
  long long nLoad;            // symbols resident on this device (multiple of 64)
  long long loadBegin;        // global column of local symbol 0 (multiple of 64)
  long long colBegin;         // hits are reported for global columns [colBegin, colEnd)
  long long colEnd;
  const uint2* tab16;         // fp16x2 rows {(256*eA,256*eC),(256*eG,256*eT)}, rowsPadded rows
  const int8_t* tab8;         // the int8 table as written by the host (4 B / row)
  uint32_t nRows;             // P_total
  uint32_t rowsPadded;        // multiple of the row chunk; pad rows hold -128 everywhere
  long long diagBegin;        // first diagonal of tile 0 (multiple of 64, may be negative)
  uint32_t nTiles;
  unsigned long long* hits;   // u64 hit words: cell[13:0] | segment[39:14] | row[63:40]
  unsigned long long hitCap;
  unsigned long long* hitCount;  // total hits found (may exceed hitCap -> overflow)
  unsigned int* tileCounter;
  const volatile int* abortFlag;  // device memory; non-zero = stop (written by the host on a side stream)
  int* status;                    // bit 0: hit buffer overflow, bit 1: aborted
  unsigned long long* smCycles;   // SM clock cycles CTA 0 spent in the kernel (effective-clock probe)
};

enum : int { SSV_STATUS_OVERFLOW = 1, SSV_STATUS_ABORTED = 2 };

// One cell of the per-cell trace: the fields of test/byCellComparator/byCellComparator.hpp:23-31 (12 bytes).
struct SsvCellRecord {
  uint8_t prevValue;
  int8_t matchScore;
  uint8_t cellValue;
  int8_t phmmVector[4];
  uint8_t symbol;
  uint8_t passesThreshold;
  uint8_t pad[3];
};

// Kernel geometry of the fast path (see ssv_kernels.cu).
constexpr int SSV_R = 16;            // fp16x2 state registers per thread (2 diagonals each)
#ifndef SSV_THREADS_V
#define SSV_THREADS_V 256
#endif
constexpr int SSV_THREADS = SSV_THREADS_V;
constexpr int SSV_TILE_DIAGS = SSV_THREADS * 2 * SSV_R;  // 8192
#ifndef SSV_ROW_CHUNK_V
#define SSV_ROW_CHUNK_V 2048  // 512 -> 2048 measured +1.9 % (fewer CTA-wide barriers per tile)
#endif
constexpr int SSV_ROW_CHUNK = SSV_ROW_CHUNK_V;   // rows of the score table staged per TMA bulk copy
constexpr int SSV_BLOCK_ROWS = 2 * SSV_R;
#ifndef SSV_SHORT_ROWS_V
#define SSV_SHORT_ROWS_V 512
#endif
constexpr int SSV_SHORT_ROWS = SSV_SHORT_ROWS_V;  // tables up to this many rows take the short-table kernel (warps on their own)
constexpr int SSV_WARP_TILE_DIAGS = 32 * 2 * SSV_R;  // diagonals of a warp tile there

// host-callable launchers (defined in ssv_kernels.cu); all asynchronous on `stream`.
void ssv_launch_table_prep(const int8_t* tab8, uint32_t nRows, uint2* tab16, uint32_t rowsPadded,
                           void* stream);
void ssv_launch_project_table(const float* emissions, uint32_t nRows, const uint32_t* modelFirstRow, const float* alphaBeta,
                              uint32_t nModels, int8_t* tab8, void* stream);
// streamingOnly: tables short enough for the short-table kernel take the streaming kernels all the same (parity tests
// of the streaming kernels' partial last block, A/B measurements)
void ssv_launch_f16x2(const SsvParams& prm, int numSMs, void* stream, bool streamingOnly = false);
bool ssv_takes_short_kernel(uint32_t nRows, bool streamingOnly);
void ssv_launch_scalar(const SsvParams& prm, void* stream);
void ssv_launch_trace(const SsvParams& prm, uint32_t rowBegin, uint32_t rowEnd, long long colBegin, long long colEnd,
                      SsvCellRecord* out, void* stream);
int ssv_f16x2_ctas_per_sm();
// 2 x 65536 bytes each: plane 0 = the low fp16 half's result for pair i = prev * 256 + (e + 128), plane 1 = the high half's
void ssv_launch_cell_selftest(uint8_t* outScore, uint8_t* outHit, void* stream);
