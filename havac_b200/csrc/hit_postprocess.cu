// hit_postprocess.cu -- what happens to the raw hit words after the search, on the device
// (SURVEY.md 8f-2):
//   * emission order: the FPGA writes hits by (segment, row, cell) ascending (device/HavacHls.cpp:
//     131-162: outer loop over segments, inner over rows, sieve in cell order). The SSV kernel
//     finds them by diagonal band, so they are sorted here: the key seg<<38 | row<<14 | cell is a
//     bit permutation of the hit word; an LSD radix sort (8-bit digits, constant digits skipped)
//     orders the keys, which are then turned back into hit words.
//   * resolution: Havac::getHitsFromFinishedRun (host/Havac.cpp:145-187) maps every hit to
//     (sequenceIndex, positionInSequence) with fastaVectorGetLocalSequencePositionFromGlobal
//     (FastaVector/src/FastaVector.c:345-373: linear scan, INCLUSIVE end test) and to
//     (phmmIndex, phmmPosition) with a prefix-sum search (host/Havac.cpp:104-142); here both are
//     binary searches, one thread per hit, with the same answers.
#include <cuda_runtime.h>

#include <algorithm>
#include <string>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <vector>

#include "client_internal.cuh"

namespace {

constexpr int STHREADS = 256;
constexpr int ITEMS = 16;
constexpr int TILE = STHREADS * ITEMS;
constexpr unsigned FULLMASK = 0xffffffffu;

__host__ __device__ __forceinline__ unsigned long long hit_to_key(unsigned long long h) {
  return ((h >> 14) & 0x3ffffffull) << 38 | (h >> 40) << 14 | (h & 0x3fffull);
}
__host__ __device__ __forceinline__ unsigned long long key_to_hit(unsigned long long k) {
  return (k & 0x3fffull) | (k >> 38) << 14 | ((k >> 14) & 0xffffffull) << 40;
}

// hits -> keys in place; OR and AND of all keys tell which digits vary at all
__global__ void hits_to_keys(unsigned long long* __restrict__ v, unsigned long long n, unsigned long long* __restrict__ orAnd) {
  unsigned long long o = 0, a = ~0ull;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
    const unsigned long long k = hit_to_key(v[i]);
    v[i] = k;
    o |= k;
    a &= k;
  }
#pragma unroll
  for (int off = 16; off; off >>= 1) {
    o |= __shfl_xor_sync(FULLMASK, o, off);
    a &= __shfl_xor_sync(FULLMASK, a, off);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicOr(orAnd, o);
    atomicAnd(orAnd + 1, a);
  }
}

__global__ void keys_to_hits(unsigned long long* __restrict__ v, unsigned long long n) {
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x)
    v[i] = key_to_hit(v[i]);
}

// digit histogram of one tile; counts are stored digit-major ([digit][block]) so that one exclusive
// scan over the whole array yields, for every (digit, block), where that block's keys of that digit go
__global__ void __launch_bounds__(STHREADS) radix_count(const unsigned long long* __restrict__ keys, unsigned long long n, int shift,
                                                        unsigned* __restrict__ counts, unsigned nBlocks) {
  __shared__ unsigned hist[256];
  hist[threadIdx.x] = 0;
  __syncthreads();
  const unsigned long long base = (unsigned long long)blockIdx.x * TILE;
#pragma unroll 4
  for (int r = 0; r < ITEMS; r++) {
    const unsigned long long i = base + (unsigned long long)r * STHREADS + threadIdx.x;
    if (i < n) atomicAdd(&hist[(keys[i] >> shift) & 255u], 1u);
  }
  __syncthreads();
  counts[(size_t)threadIdx.x * nBlocks + blockIdx.x] = hist[threadIdx.x];
}

// one CTA: exclusive scan of `total` 32-bit counters in place (total is a multiple of 4: 256 digits x tiles, and the
// array is 16-byte aligned). 4096 counters per round: one coalesced 128-bit load per thread, a shuffle scan inside
// each warp, the 32 warp totals scanned by warp 0, a running carry across rounds.
__global__ void __launch_bounds__(1024) scan_exclusive(unsigned* __restrict__ v, unsigned long long total) {
  __shared__ unsigned warpSum[32];
  __shared__ unsigned carry;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  uint4* v4 = reinterpret_cast<uint4*>(v);
  const unsigned long long quads = total / 4;
  for (unsigned long long base = 0; base < quads; base += 1024) {
    const unsigned long long i = base + threadIdx.x;
    uint4 x = i < quads ? v4[i] : make_uint4(0u, 0u, 0u, 0u);
    const unsigned mine = x.x + x.y + x.z + x.w;
    unsigned incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const unsigned up = __shfl_up_sync(FULLMASK, incl, off);
      if (lane >= (unsigned)off) incl += up;
    }
    if (lane == 31) warpSum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const unsigned t = warpSum[lane];
      unsigned ti = t;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const unsigned up = __shfl_up_sync(FULLMASK, ti, off);
        if (lane >= (unsigned)off) ti += up;
      }
      warpSum[lane] = ti - t;  // exclusive offset of each warp inside the round
    }
    __syncthreads();
    unsigned at = carry + warpSum[warp] + (incl - mine);
    if (i < quads) {
      uint4 o;
      o.x = at;
      o.y = at + x.x;
      o.z = o.y + x.y;
      o.w = o.z + x.z;
      v4[i] = o;
    }
    __syncthreads();  // everyone has read carry and warpSum
    if (threadIdx.x == 1023) carry = at + mine;  // thread 1023's inclusive end = the round's total on top of the carry
    __syncthreads();
  }
}

// stable scatter of one tile: keys keep their order within a digit (tile order = round, warp, lane)
__global__ void __launch_bounds__(STHREADS) radix_scatter(const unsigned long long* __restrict__ in, unsigned long long* __restrict__ out,
                                                          unsigned long long n, int shift, const unsigned* __restrict__ offsets,
                                                          unsigned nBlocks) {
  __shared__ unsigned warpCnt[STHREADS / 32][256];
  __shared__ unsigned running[256];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  running[threadIdx.x] = offsets[(size_t)threadIdx.x * nBlocks + blockIdx.x];
  const unsigned long long base = (unsigned long long)blockIdx.x * TILE;
  for (int r = 0; r < ITEMS; r++) {
#pragma unroll
    for (int w = 0; w < STHREADS / 32; w++) warpCnt[w][threadIdx.x] = 0;
    __syncthreads();
    const unsigned long long i = base + (unsigned long long)r * STHREADS + threadIdx.x;
    const bool valid = i < n;
    const unsigned long long key = valid ? in[i] : 0ull;
    const unsigned digit = valid ? (unsigned)((key >> shift) & 255u) : 256u;
    const unsigned same = __match_any_sync(FULLMASK, digit);
    const unsigned rank = __popc(same & ((1u << lane) - 1u));
    if (valid && rank == 0) warpCnt[warp][digit] = __popc(same);
    __syncthreads();
    if (valid) {
      unsigned at = running[digit] + rank;
      for (unsigned w = 0; w < warp; w++) at += warpCnt[w][digit];
      out[at] = key;
    }
    __syncthreads();
    unsigned add = 0;
#pragma unroll
    for (int w = 0; w < STHREADS / 32; w++) add += warpCnt[w][threadIdx.x];
    running[threadIdx.x] += add;
    __syncthreads();
  }
}

// one thread per (sorted) hit; sequenceIndex = 0xffffffff marks a hit in the zero padding, which the
// reference drops (host/Havac.cpp:169-173)
__global__ void resolve_hits(const unsigned long long* __restrict__ hits, unsigned long long n,
                             const unsigned long long* __restrict__ seqEnd, unsigned nSeq, const unsigned* __restrict__ prefix,
                             unsigned nModels, unsigned long long mirrorSymbols, unsigned long long* __restrict__ seqPos,
                             unsigned* __restrict__ seqIdx, unsigned* __restrict__ phmmPos, unsigned* __restrict__ phmmIdx) {
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
    const unsigned long long h = hits[i];
    unsigned long long pos = ((h >> 14) & 0x3ffffffull) * (unsigned long long)SEG + (h & 0x3fffull);
    const unsigned row = (unsigned)(h >> 40);
    // reverse-complemented sequence (strand.cu): report the forward-strand coordinate of the symbol;
    // anything at or beyond the mirrored length lies in the padding
    const bool beyond = mirrorSymbols != 0 && pos >= mirrorSymbols;
    if (mirrorSymbols != 0 && !beyond) pos = mirrorSymbols - 1 - pos;
    // first record whose end offset is >= pos (the reference's `<=` test, FastaVector.c:361)
    unsigned lo = 0, hi = nSeq;
    while (lo < hi) {
      const unsigned mid = lo + (hi - lo) / 2;
      if (seqEnd[mid] >= pos) hi = mid;
      else lo = mid + 1;
    }
    // last model whose first row is <= row (prefix[0] = 0, so there always is one)
    unsigned a = 0, b = nModels;
    while (b - a > 1) {
      const unsigned mid = a + (b - a) / 2;
      if (prefix[mid] <= row) a = mid;
      else b = mid;
    }
    if (beyond) lo = nSeq;
    seqIdx[i] = lo == nSeq ? 0xffffffffu : lo;
    seqPos[i] = pos - (lo == 0 || lo == nSeq ? 0ull : seqEnd[lo - 1]);
    phmmIdx[i] = a;
    phmmPos[i] = row - prefix[a];
  }
}

}  // namespace

// The device context's post-processing scratch (DeviceCtx::sortScratch), at least `need` bytes: kept between calls and
// grown by half again when a run needs more, so that the steady state makes no cudaMalloc / cudaFree. The sort and the
// resolution use it one after the other (each drains the stream before it returns). The device must be current.
static int ensure_scratch(havac_cuda_client* c, DeviceCtx& d, size_t need) {
  if (d.sortScratchCap >= need) return HAVAC_CUDA_OK;
  if (d.sortScratch) cudaFree(d.sortScratch);
  d.sortScratch = nullptr;
  d.sortScratchCap = 0;
  const size_t want = need + need / 2;
  if (cudaMalloc(&d.sortScratch, want) == cudaSuccess) {
    d.sortScratchCap = want;
    return HAVAC_CUDA_OK;
  }
  cudaGetLastError();
  if (cudaMalloc(&d.sortScratch, need) == cudaSuccess) {
    d.sortScratchCap = need;
    return HAVAC_CUDA_OK;
  }
  cudaGetLastError();
  d.sortScratch = nullptr;
  return havac_fail(c, HAVAC_CUDA_E_NOMEM, "cudaMalloc(post-processing scratch): out of device memory");
}

// Loads this file's kernels onto the current device (a query of their attributes does it), so that the first hit
// list of a client is not the one that pays for it. Called while the device is being opened.
void havac_postprocess_preload() {
  cudaFuncAttributes a;
  cudaFuncGetAttributes(&a, hits_to_keys);
  cudaFuncGetAttributes(&a, keys_to_hits);
  cudaFuncGetAttributes(&a, radix_count);
  cudaFuncGetAttributes(&a, scan_exclusive);
  cudaFuncGetAttributes(&a, radix_scatter);
  cudaFuncGetAttributes(&a, resolve_hits);
  cudaGetLastError();
}

// Sorts the hitCount raw hits of device d into emission order, in place in d.dHits.
int havac_sort_hits_on_device(havac_cuda_client* c, DeviceCtx& d) {
  const unsigned long long n = d.hitCount;
  if (d.hitsSorted || n < 2) { d.hitsSorted = true; return HAVAC_CUDA_OK; }
  if (n >= (1ull << 32)) return havac_fail(c, HAVAC_CUDA_E_LENGTH, "more than 2^32 hits on one device");
  CK(cudaSetDevice(d.dev));
  cudaStream_t st = d.stream;
  const unsigned nBlocks = (unsigned)((n + TILE - 1) / TILE);
  // one allocation, kept on the device context: [keys n x 8 | counts 256 x tiles x 4 | OR/AND 16], 256-byte aligned parts
  const size_t keysBytes = ((size_t)n * 8 + 255) & ~(size_t)255;
  const size_t countsBytes = ((size_t)256 * nBlocks * sizeof(unsigned) + 255) & ~(size_t)255;
  const size_t need = keysBytes + countsBytes + 256;
  {
    const int rc = ensure_scratch(c, d, need);
    if (rc) return rc;
  }
  unsigned char* const base = static_cast<unsigned char*>(d.sortScratch);
  unsigned long long* const dTmp = reinterpret_cast<unsigned long long*>(base);
  unsigned* const dCounts = reinterpret_cast<unsigned*>(base + keysBytes);
  unsigned long long* const dOrAnd = reinterpret_cast<unsigned long long*>(base + keysBytes + countsBytes);
  const unsigned long long init[2] = {0ull, ~0ull};
  CK(cudaMemcpyAsync(dOrAnd, init, sizeof init, cudaMemcpyHostToDevice, st));
  const unsigned grid = (unsigned)std::min<unsigned long long>((n + 255) / 256, 148ull * 16);
  hits_to_keys<<<grid, 256, 0, st>>>(d.dHits, n, dOrAnd);
  unsigned long long orAnd[2];
  CK(cudaMemcpyAsync(orAnd, dOrAnd, sizeof orAnd, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  const unsigned long long varying = orAnd[0] ^ orAnd[1];
  unsigned long long* src = d.dHits;
  unsigned long long* dst = dTmp;
  for (int shift = 0; shift < 64; shift += 8) {
    if (((varying >> shift) & 255ull) == 0) continue;
    radix_count<<<nBlocks, STHREADS, 0, st>>>(src, n, shift, dCounts, nBlocks);
    scan_exclusive<<<1, 1024, 0, st>>>(dCounts, (unsigned long long)256 * nBlocks);
    radix_scatter<<<nBlocks, STHREADS, 0, st>>>(src, dst, n, shift, dCounts, nBlocks);
    std::swap(src, dst);
  }
  keys_to_hits<<<grid, 256, 0, st>>>(src, n);
  if (src != d.dHits) CK(cudaMemcpyAsync(d.dHits, src, n * 8, cudaMemcpyDeviceToDevice, st));
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
  d.hitsSorted = true;
  return HAVAC_CUDA_OK;
}

extern "C" int havac_cuda_resolve_hits(havac_cuda_client* c, const uint64_t* sequenceEnd, uint64_t nSequences,
                                       const uint32_t* modelLength, uint32_t nModels, uint64_t* sequencePosition,
                                       uint32_t* sequenceIndex, uint32_t* phmmPosition, uint32_t* phmmIndex, uint64_t cap,
                                       uint64_t* nOut) {
  if (!c || !nOut || (!sequenceEnd && nSequences) || !modelLength || nModels == 0) return HAVAC_CUDA_E_INVALID;
  uint64_t total = 0;
  int rc = havac_cuda_num_hits(c, &total);  // E_LOGIC before a run, E_OVERFLOW after an overflowed one
  if (rc) return rc;
  *nOut = total;  // upper bound until the padding hits are dropped below
  if (cap == 0 || total == 0) return HAVAC_CUDA_OK;
  if (cap < total || !sequencePosition || !sequenceIndex || !phmmPosition || !phmmIndex)
    return havac_fail(c, HAVAC_CUDA_E_LENGTH, "hit arrays too small: " + std::to_string(total) + " hits");
  if (nSequences >= (1ull << 32)) return havac_fail(c, HAVAC_CUDA_E_LENGTH, "more than 2^32 records");
  std::vector<unsigned> prefix(nModels);
  for (uint32_t i = 0, acc = 0; i < nModels; i++) {  // generatePhmmLenPrefixSums, host/Havac.cpp:104-116
    prefix[i] = acc;
    acc += modelLength[i];
  }
  uint64_t kept = 0;
  const bool trace = std::getenv("HAVAC_TRACE") != nullptr;
  auto now = [] { return std::chrono::steady_clock::now(); };
  auto secs = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double>(b - a).count(); };
  for (DeviceCtx& d : c->devs) {  // devices hold increasing segment ranges: device order is emission order
    const unsigned long long n = d.hitCount;
    if (!n) continue;
    const auto t0 = now();
    rc = havac_sort_hits_on_device(c, d);
    if (rc) return rc;
    const auto t1 = now();
    CK(cudaSetDevice(d.dev));
    // record ends, model prefix sums and the four result arrays, carved out of the context's scratch (256-byte parts)
    auto part = [](size_t bytes) { return (bytes + 255) & ~(size_t)255; };
    const size_t oEnd = 0, oPrefix = oEnd + part(nSequences * 8), oPos = oPrefix + part((size_t)nModels * 4), oSeqIdx = oPos + part(n * 8),
                 oPhmmPos = oSeqIdx + part(n * 4), oPhmmIdx = oPhmmPos + part(n * 4), total = oPhmmIdx + part(n * 4);
    rc = ensure_scratch(c, d, total);
    if (rc) return rc;
    unsigned char* const sb = static_cast<unsigned char*>(d.sortScratch);
    unsigned long long* const dEnd = reinterpret_cast<unsigned long long*>(sb + oEnd);
    unsigned* const dPrefix = reinterpret_cast<unsigned*>(sb + oPrefix);
    unsigned long long* const dPos = reinterpret_cast<unsigned long long*>(sb + oPos);
    unsigned* const dSeqIdx = reinterpret_cast<unsigned*>(sb + oSeqIdx);
    unsigned* const dPhmmPos = reinterpret_cast<unsigned*>(sb + oPhmmPos);
    unsigned* const dPhmmIdx = reinterpret_cast<unsigned*>(sb + oPhmmIdx);
    CK(cudaMemcpyAsync(dEnd, sequenceEnd, nSequences * 8, cudaMemcpyHostToDevice, d.stream));
    CK(cudaMemcpyAsync(dPrefix, prefix.data(), (size_t)nModels * 4, cudaMemcpyHostToDevice, d.stream));
    const unsigned grid = (unsigned)std::min<unsigned long long>((n + 255) / 256, 148ull * 16);
    resolve_hits<<<grid, 256, 0, d.stream>>>(d.dHits, n, dEnd, (unsigned)nSequences, dPrefix, nModels, c->mirrorSymbols, dPos, dSeqIdx, dPhmmPos,
                                              dPhmmIdx);
    CK(cudaGetLastError());
    // straight into the caller's arrays at the running offset, then squeeze out the padding hits
    CK(cudaMemcpyAsync(sequencePosition + kept, dPos, n * 8, cudaMemcpyDeviceToHost, d.stream));
    CK(cudaMemcpyAsync(sequenceIndex + kept, dSeqIdx, n * 4, cudaMemcpyDeviceToHost, d.stream));
    CK(cudaMemcpyAsync(phmmPosition + kept, dPhmmPos, n * 4, cudaMemcpyDeviceToHost, d.stream));
    CK(cudaMemcpyAsync(phmmIndex + kept, dPhmmIdx, n * 4, cudaMemcpyDeviceToHost, d.stream));
    CK(cudaStreamSynchronize(d.stream));
    const auto t2 = now();
    c->h2dBytes += nSequences * 8 + (uint64_t)nModels * 4;
    c->d2hBytes += n * 20;
    uint64_t w = kept;
    for (uint64_t r = kept; r < kept + n; r++) {
      if (sequenceIndex[r] == 0xffffffffu) continue;
      if (w != r) {
        sequencePosition[w] = sequencePosition[r];
        sequenceIndex[w] = sequenceIndex[r];
        phmmPosition[w] = phmmPosition[r];
        phmmIndex[w] = phmmIndex[r];
      }
      w++;
    }
    kept = w;
    if (trace)
      std::fprintf(stderr, "[havac trace] device %d: %llu hits, sort %.3f s, resolve + copy back %.3f s, squeeze %.3f s\n", d.dev, n,
                   secs(t0, t1), secs(t1, t2), secs(t2, now()));
  }
  *nOut = kept;
  return HAVAC_CUDA_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Emission-order sort with NO allocation (havac_cuda_sort_hits_static, include/havac_cuda.h). No counterpart in the
// reference; added for callers that cannot call cudaMalloc (code under CUDA
// graph capture, tight serving loops, harnesses that forbid it). The launch sequence is the one of
// havac_sort_hits_on_device above, with every buffer that function takes from cudaMalloc laid out in ONE file-scope
// device array instead. Every GPU that loads the library holds the array (16.25 MiB) until the process exits.
//
// Scratch layout (bytes), for the maximum of HAVAC_CUDA_SORT_STATIC_MAX_HITS = 2^20 hits:
//   [0, 8 Mi)            keys, buffer A (the input is copied here and turned into keys in place)
//   [8 Mi, 16 Mi)        keys, buffer B (the radix passes go A -> B -> A ...)
//   [16 Mi, 16 Mi+256 Ki) digit counts, 256 digits x (2^20 / TILE = 256 tiles) x 4
//   [.., +256)           OR / AND of all keys (which digits vary)
// All calls share the array and it keeps the last call's contents: the OR/AND pair is the only part a call reads
// before writing, and it is re-initialised by every call; keys and counts are fully overwritten over [0, n) /
// [0, 256 * tiles) before they are read. Calls that overlap ON THE GPU (two streams, two host threads) would race
// on the array: the entry point holds a per-process mutex and returns only after its stream has drained, so calls
// through it cannot overlap; nothing else may use the array.
namespace {

constexpr unsigned long long STATIC_MAX_HITS = HAVAC_CUDA_SORT_STATIC_MAX_HITS;
constexpr size_t STATIC_KEYS_BYTES = (size_t)STATIC_MAX_HITS * 8;
constexpr size_t STATIC_MAX_TILES = (STATIC_MAX_HITS + TILE - 1) / TILE;
constexpr size_t STATIC_COUNTS_BYTES = 256 * STATIC_MAX_TILES * sizeof(unsigned);
constexpr size_t STATIC_OFF_B = STATIC_KEYS_BYTES;
constexpr size_t STATIC_OFF_COUNTS = 2 * STATIC_KEYS_BYTES;
constexpr size_t STATIC_OFF_ORAND = STATIC_OFF_COUNTS + STATIC_COUNTS_BYTES;
constexpr size_t STATIC_SCRATCH_BYTES = STATIC_OFF_ORAND + 256;
static_assert(STATIC_OFF_B % 256 == 0 && STATIC_OFF_COUNTS % 256 == 0 && STATIC_OFF_ORAND % 256 == 0, "sub-buffers stay 256-byte aligned");

// In host code this NAME is a host-side stand-in for the symbol, not a device address: kernels get the pointer
// that cudaGetSymbolAddress returns (static_scratch_on below), never `g_sortScratch` itself.
static __device__ __align__(256) unsigned char g_sortScratch[STATIC_SCRATCH_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
std::mutex g_staticMutex;                           // one call at a time per process: the array is shared
unsigned char* g_staticAddr[STATIC_MAX_DEVICES];    // device address of g_sortScratch per device ordinal, looked up once

// The device address of the scratch array on `dev`; `dev` must be the current device (the symbol has one
// instance per device, and the lookup answers for the current one).
cudaError_t static_scratch_on(int dev, unsigned char** out) {
  if (!g_staticAddr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_sortScratch);
    if (e != cudaSuccess) return e;
    g_staticAddr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_staticAddr[dev];
  return cudaSuccess;
}

}  // namespace

// CK for an entry point without a client: the text goes where havac_cuda_create's goes (havac_cuda_last_error(NULL))
#define CKS(call)                                                                                              \
  do {                                                                                                         \
    cudaError_t e_ = (call);                                                                                   \
    if (e_ != cudaSuccess)                                                                                     \
      return havac_fail(nullptr, HAVAC_CUDA_E_RUNTIME, std::string("havac_cuda_sort_hits_static: " #call ": ") + cudaGetErrorString(e_)); \
  } while (0)

extern "C" int havac_cuda_sort_hits_static(uint32_t deviceIndex, const uint64_t* hitsIn, uint64_t n, uint64_t* hitsOut, void* stream) {
  havac_cuda_client* const c = nullptr;  // no client: error texts go to the calling thread's slot, havac_cuda_last_error(NULL)
  if (n > STATIC_MAX_HITS)
    return havac_fail(c, HAVAC_CUDA_E_LENGTH, "havac_cuda_sort_hits_static: " + std::to_string(n) + " hits exceed the static scratch's maximum of " +
                                                  std::to_string(STATIC_MAX_HITS) + " (HAVAC_CUDA_SORT_STATIC_MAX_HITS); sort through a client instead");
  if (n == 0) return HAVAC_CUDA_OK;
  if (!hitsIn || !hitsOut) return havac_fail(c, HAVAC_CUDA_E_INVALID, "havac_cuda_sort_hits_static: null hit array");
  if (deviceIndex >= (uint32_t)STATIC_MAX_DEVICES) return havac_fail(c, HAVAC_CUDA_E_INVALID, "havac_cuda_sort_hits_static: device index out of range");
  std::lock_guard<std::mutex> lock(g_staticMutex);
  CKS(cudaSetDevice((int)deviceIndex));
  unsigned char* base = nullptr;
  CKS(static_scratch_on((int)deviceIndex, &base));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned long long* const bufA = reinterpret_cast<unsigned long long*>(base);
  unsigned long long* const bufB = reinterpret_cast<unsigned long long*>(base + STATIC_OFF_B);
  unsigned* const dCounts = reinterpret_cast<unsigned*>(base + STATIC_OFF_COUNTS);
  unsigned long long* const dOrAnd = reinterpret_cast<unsigned long long*>(base + STATIC_OFF_ORAND);
  const unsigned nBlocks = (unsigned)((n + TILE - 1) / TILE);
  // cudaMemcpyDefault: hitsIn / hitsOut may be device or host memory (unified addressing tells which)
  CKS(cudaMemcpyAsync(bufA, hitsIn, n * 8, cudaMemcpyDefault, st));
  const unsigned long long init[2] = {0ull, ~0ull};  // the one part of the scratch that must not carry over
  CKS(cudaMemcpyAsync(dOrAnd, init, sizeof init, cudaMemcpyHostToDevice, st));
  const unsigned grid = (unsigned)std::min<unsigned long long>((n + 255) / 256, 148ull * 16);
  hits_to_keys<<<grid, 256, 0, st>>>(bufA, n, dOrAnd);
  unsigned long long orAnd[2];
  CKS(cudaMemcpyAsync(orAnd, dOrAnd, sizeof orAnd, cudaMemcpyDeviceToHost, st));
  CKS(cudaStreamSynchronize(st));
  const unsigned long long varying = orAnd[0] ^ orAnd[1];
  unsigned long long* src = bufA;
  unsigned long long* dst = bufB;
  for (int shift = 0; shift < 64; shift += 8) {
    if (((varying >> shift) & 255ull) == 0) continue;
    radix_count<<<nBlocks, STHREADS, 0, st>>>(src, n, shift, dCounts, nBlocks);
    scan_exclusive<<<1, 1024, 0, st>>>(dCounts, (unsigned long long)256 * nBlocks);
    radix_scatter<<<nBlocks, STHREADS, 0, st>>>(src, dst, n, shift, dCounts, nBlocks);
    std::swap(src, dst);
  }
  keys_to_hits<<<grid, 256, 0, st>>>(src, n);
  CKS(cudaGetLastError());
  CKS(cudaMemcpyAsync(hitsOut, src, n * 8, cudaMemcpyDefault, st));
  CKS(cudaStreamSynchronize(st));
  return HAVAC_CUDA_OK;
}
