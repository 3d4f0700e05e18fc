// client_internal.cuh -- private state of the C-ABI device client, shared by its translation
// units (havac_cuda.cu: buffers / launch / state; fasta_ingest.cu: FASTA text -> packed sequence).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/havac_cuda.h"

constexpr long long SEG = HAVAC_CUDA_SEGMENT_SYMBOLS;
constexpr unsigned long long MAX_ROWS = 1ull << 24;  // phmmPos_t is 24 bits (device/HavacHls.hpp:19)
constexpr long long HALO_RESIDENT = 1ll << 24;       // symbols kept left of a shard (>= MAX_ROWS - 1)

enum KernelKind { KERNEL_AUTO = 0, KERNEL_F16X2 = 1, KERNEL_SCALAR = 2, KERNEL_F16X2_STREAMING = 3 };

struct DeviceResultBlock {
  unsigned long long hitCount;  // total hits found (may exceed the capacity -> overflow)
  unsigned long long cycles;    // SM cycles of CTA 0 (effective-clock probe)
  int status;                   // SSV_STATUS_* bits
  unsigned int tileCounter;     // next tile to hand out
  int abort;                    // non-zero = stop; written from the control stream
  int pad;
};
constexpr size_t RESULT_BYTES = 24;  // hitCount, cycles, status, tileCounter: what the host reads back

struct DeviceCtx {
  int dev = 0;
  int numSMs = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t evStart = nullptr, evStop = nullptr;
  cudaEvent_t evDone = nullptr;  // after the result block of the run has landed in hostRes
  // sequence shard
  uint32_t* dSeq = nullptr;
  size_t seqCap = 0;                   // bytes allocated at dSeq (kept across writes of the same size)
  long long resBegin = 0, resEnd = 0;  // global columns resident in dSeq
  long long colBegin = 0, colEnd = 0;  // global columns whose hits this device reports
  // score table
  int8_t* dTab8 = nullptr;
  uint2* dTab16 = nullptr;
  uint32_t rowsPadded = 0;
  size_t tabCap = 0;  // bytes allocated at dTab8
  // outputs
  unsigned long long* dHits = nullptr;
  // One 32-byte control block per device (DeviceResultBlock): cleared by ONE memset per launch, copied
  // to pinned host memory by ONE asynchronous copy behind the stop event. The pointers below point into it.
  DeviceResultBlock* dCtl = nullptr;
  DeviceResultBlock* hostRes = nullptr;  // pinned
  unsigned long long* dHitCount = nullptr;
  unsigned int* dTileCounter = nullptr;
  int* dStatus = nullptr;
  unsigned long long* dCycles = nullptr;  // SM cycles of CTA 0 in the last run
  unsigned long long cycles = 0;
  int* dAbort = nullptr;          // device-resident abort flag the kernels poll (L2 hit, no PCIe)
  cudaStream_t ctlStream = nullptr;  // carries the abort write while the run stream is busy
  // results of the last finished run
  unsigned long long hitCount = 0;
  int status = 0;
  bool hitsSorted = false;  // dHits already put into emission order (hit_postprocess.cu)
  // scratch of the emission-order sort (second key buffer, digit counts, OR/AND word): kept between runs and grown
  // when a run has more hits, so that reading the hit list costs no cudaMalloc / cudaFree in the steady state
  void* sortScratch = nullptr;
  size_t sortScratchCap = 0;
  // how the last launch cut its work (havac_cuda_tile_geometry)
  long long lastLoadBegin = 0, lastDiagBegin = 0;
  uint32_t lastTiles = 0, lastTileDiags = 0;
};

struct havac_cuda_client {
  std::vector<DeviceCtx> devs;
  std::string err;
  unsigned long long hitCap = 16ull << 20;  // per device
  bool hitCapGrows = true;                  // on overflow the buffers are enlarged and the search repeated
  int regrowths = 0;                        // how often that happened in the last run
  uint64_t segBegin = 0, segEnd = 0;        // 0,0 = whole buffer
  uint64_t seqSegments = 0;                 // segments in the buffer given to write_sequence
  uint64_t mirrorSymbols = 0;               // N when the resident sequence is the reverse complement over N symbols (strand.cu)
  uint32_t nRows = 0;
  bool launched = false, finished = false, aborted = false;
  bool abortRequested = false;  // havac_cuda_abort/destroy asked for a stop: no repeat after a hit-buffer overflow
  int finalState = HAVAC_CUDA_STATE_NEW;
  int* hostOne = nullptr;    // pinned source word for the asynchronous abort write
  int kernel = KERNEL_AUTO;
  uint32_t launches = 0;
  float lastMs = 0.f;
  uint64_t h2dBytes = 0, d2hBytes = 0;  // running totals of what crossed PCIe
  // record table of the last havac_cuda_ingest_fasta
  std::vector<uint64_t> ingestEnd, ingestHeaderByte;
};

#define CK(call)                                                                         \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) {                                                             \
      c->err = std::string(#call) + ": " + cudaGetErrorString(e_);                       \
      return e_ == cudaErrorMemoryAllocation ? HAVAC_CUDA_E_NOMEM : HAVAC_CUDA_E_RUNTIME; \
    }                                                                                    \
  } while (0)

// sets the client's (or, for c == nullptr, this thread's create-time) error text; returns `code`
int havac_fail(havac_cuda_client* c, int code, const std::string& msg);
void havac_free_sequence(DeviceCtx& d);
int havac_place_sequence(havac_cuda_client* c, const uint8_t* packed, int srcDevice, uint64_t nBytes);
// hit_postprocess.cu: sorts d.dHits[0, d.hitCount) into the FPGA's emission order on the device
int havac_sort_hits_on_device(havac_cuda_client* c, DeviceCtx& d);
void havac_postprocess_preload();  // loads the sort / resolve kernels onto the current device
