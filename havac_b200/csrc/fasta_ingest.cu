// fasta_ingest.cu -- FASTA text -> packed 2-bit sequence, on the device.
//
// Replaces, for a client that holds no sequence yet, the three host passes of the reference:
// fastaVectorReadFasta (FastaVector/src/FastaVector.c:54-168, a per-character fgetc state
// machine), SequencePreprocessor (host/sequence/SequencePreprocessor.cpp:9-85, a scalar
// read-modify-write packer) and HavacHwClient::writeSequence (host/HavacHwClient.cpp:78-110).
// The result is bit-identical: same symbols, same record end offsets, same padding.
//
// The reader is a 4-state automaton (line start / header / comment / sequence). A byte is a
// function state -> state, functions compose associatively, so the state in front of every
// 64-byte slice comes out of a prefix scan of 8-bit transition functions; with the states known
// every slice counts and then writes its symbols independently:
//   K1 fasta_scan_states   per-slice transition function, block-wide scan, block aggregates
//   K2 fasta_chain_blocks  start state of every block (one CTA over the block aggregates)
//   K3 fasta_count         per-slice symbol / record / non-ACGT counts, block-wide scan
//   K4 fasta_chain_counts  64-bit bases per block, grand totals
//   K5 fasta_emit          2-bit symbols OR-ed into the packed buffer; record table; fix-up list
//   K6 fasta_apply_fixups  non-ACGT symbols (and record terminators) resolved by the caller
// Symbols that are not A/C/G/T -- including the '\0' the reference appends to every record --
// become rand()-chosen bases in the reference (SequencePreprocessor.cpp:62-85). rand() is
// sequential host state, so the device only lists those positions in sequence order; the caller's
// callback resolves them in that order and K6 writes the answers.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "client_internal.cuh"

namespace {

constexpr int SLICE = 64;             // text bytes per thread
constexpr int FTHREADS = 256;         // threads per CTA
constexpr int BLOCK_BYTES = SLICE * FTHREADS;
constexpr unsigned FULLMASK = 0xffffffffu;

enum : unsigned { ST_NL = 0, ST_HEADER = 1, ST_COMMENT = 2, ST_SEQ = 3 };
constexpr unsigned F_IDENTITY = 0xE4;  // state s -> s, two bits per state

// Transition function of one byte (next state for each of the four current states, 2 bits each),
// following FastaVector.c:66-157: a printable character is one in [' ', 0x7f]; bytes >= 0x80
// are negative chars there and never printable.
__device__ __forceinline__ unsigned byte_fn(unsigned c) {
  if (c == '\n') return 0x00;                  // every state -> line start
  if (c == ';') return 0xAA;                   // every state -> comment
  if (c == '>') return 0xE5;                   // line start -> header; others stay
  if (c >= ' ' && c < 0x80) return 0xE7;       // line start -> sequence; others stay
  return F_IDENTITY;
}
__device__ __forceinline__ unsigned fn_apply(unsigned f, unsigned s) { return (f >> (2 * s)) & 3u; }
// "a, then b"
__device__ __forceinline__ unsigned fn_compose(unsigned a, unsigned b) {
  return fn_apply(b, a & 3u) | fn_apply(b, (a >> 2) & 3u) << 2 | fn_apply(b, (a >> 4) & 3u) << 4 |
         fn_apply(b, (a >> 6) & 3u) << 6;
}
// what the automaton does with byte c in state s
__device__ __forceinline__ bool emits(unsigned s, unsigned c) {
  const bool printable = c >= ' ' && c < 0x80;
  return printable && c != ';' && ((s == ST_NL && c != '>') || s == ST_SEQ);
}
__device__ __forceinline__ bool starts_record(unsigned s, unsigned c) { return s == ST_NL && c == '>'; }
// 0..3 for A/C/G/T in either case, 4 otherwise (SequencePreprocessor.cpp:62-71)
__device__ __forceinline__ unsigned base_code(unsigned c) {
  const unsigned u = c & 0xdfu;  // fold case
  return u == 'A' ? 0u : u == 'C' ? 1u : u == 'G' ? 2u : u == 'T' ? 3u : 4u;
}

struct Slice {
  uint4 v[SLICE / 16];
  __device__ __forceinline__ void load(const uint8_t* text, size_t slice) {
    const uint4* p = reinterpret_cast<const uint4*>(text + slice * SLICE);
#pragma unroll
    for (int i = 0; i < SLICE / 16; i++) v[i] = __ldg(p + i);
  }
  __device__ __forceinline__ unsigned byte(int i) const {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(v);
    return (w[i >> 2] >> (8 * (i & 3))) & 0xffu;
  }
};

// Inclusive scan of transition functions across the CTA (order matters: lower threads first).
// Returns this thread's EXCLUSIVE prefix; *blockTotal receives the whole CTA's function.
__device__ unsigned block_scan_fn(unsigned f, unsigned* blockTotal) {
  __shared__ unsigned warpTotal[FTHREADS / 32];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned incl = f;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const unsigned prev = __shfl_up_sync(FULLMASK, incl, off);
    if (lane >= (unsigned)off) incl = fn_compose(prev, incl);
  }
  if (lane == 31) warpTotal[warp] = incl;
  __syncthreads();
  unsigned before = F_IDENTITY;
  for (unsigned w = 0; w < warp; w++) before = fn_compose(before, warpTotal[w]);
  unsigned exclInWarp = __shfl_up_sync(FULLMASK, incl, 1);
  if (lane == 0) exclInWarp = F_IDENTITY;
  if (threadIdx.x == FTHREADS - 1) *blockTotal = fn_compose(before, incl);
  return fn_compose(before, exclInWarp);
}

__global__ void __launch_bounds__(FTHREADS) fasta_scan_states(const uint8_t* __restrict__ text, uint8_t* __restrict__ sliceFn,
                                                              uint8_t* __restrict__ blockFn) {
  __shared__ unsigned total;
  const size_t slice = (size_t)blockIdx.x * FTHREADS + threadIdx.x;
  Slice s;
  s.load(text, slice);
  unsigned f = F_IDENTITY;
#pragma unroll 8
  for (int i = 0; i < SLICE; i++) f = fn_compose(f, byte_fn(s.byte(i)));
  const unsigned excl = block_scan_fn(f, &total);
  sliceFn[slice] = (uint8_t)excl;
  __syncthreads();
  if (threadIdx.x == 0) blockFn[blockIdx.x] = (uint8_t)total;
}

// One CTA: state in front of every block, from the file's initial state (line start).
__global__ void __launch_bounds__(1024) fasta_chain_blocks(const uint8_t* __restrict__ blockFn, uint8_t* __restrict__ blockState,
                                                           unsigned nBlocks) {
  __shared__ unsigned part[1024];
  const unsigned per = (nBlocks + 1023) / 1024;
  const unsigned b0 = min(threadIdx.x * per, nBlocks), b1 = min(b0 + per, nBlocks);
  unsigned f = F_IDENTITY;
  for (unsigned b = b0; b < b1; b++) f = fn_compose(f, blockFn[b]);
  part[threadIdx.x] = f;
  __syncthreads();
  if (threadIdx.x == 0) {  // 1024 eight-bit compositions: not worth a second scan level
    unsigned acc = F_IDENTITY;
    for (int t = 0; t < 1024; t++) {
      const unsigned mine = part[t];
      part[t] = acc;
      acc = fn_compose(acc, mine);
    }
  }
  __syncthreads();
  unsigned state = fn_apply(part[threadIdx.x], ST_NL);
  for (unsigned b = b0; b < b1; b++) {
    blockState[b] = (uint8_t)state;
    state = fn_apply(blockFn[b], state);
  }
}

// counts packed in one 64-bit word: symbols emitted | record starts << 20 | non-ACGT symbols << 40
constexpr int R_SHIFT = 20, N_SHIFT = 40;
constexpr unsigned long long FIELD = (1ull << 20) - 1;

__device__ unsigned long long block_scan_counts(unsigned long long x, unsigned long long* blockTotal) {
  __shared__ unsigned long long warpTotal[FTHREADS / 32];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long incl = x;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const unsigned long long prev = __shfl_up_sync(FULLMASK, incl, off);
    if (lane >= (unsigned)off) incl += prev;
  }
  if (lane == 31) warpTotal[warp] = incl;
  __syncthreads();
  unsigned long long before = 0;
  for (unsigned w = 0; w < warp; w++) before += warpTotal[w];
  if (threadIdx.x == FTHREADS - 1) *blockTotal = before + incl;
  return before + incl - x;
}

__global__ void __launch_bounds__(FTHREADS) fasta_count(const uint8_t* __restrict__ text, uint8_t* __restrict__ sliceFn,
                                                        const uint8_t* __restrict__ blockState,
                                                        unsigned long long* __restrict__ slicePrefix,
                                                        unsigned long long* __restrict__ blockCount) {
  __shared__ unsigned long long total;
  const size_t slice = (size_t)blockIdx.x * FTHREADS + threadIdx.x;
  Slice s;
  s.load(text, slice);
  const unsigned start = fn_apply(sliceFn[slice], blockState[blockIdx.x]);
  unsigned state = start, nE = 0, nR = 0, nN = 0;
#pragma unroll 8
  for (int i = 0; i < SLICE; i++) {
    const unsigned c = s.byte(i);
    if (emits(state, c)) {
      nE++;
      nN += base_code(c) == 4u;
    }
    nR += starts_record(state, c);
    state = fn_apply(byte_fn(c), state);
  }
  const unsigned long long mine = (unsigned long long)nE | (unsigned long long)nR << R_SHIFT | (unsigned long long)nN << N_SHIFT;
  const unsigned long long excl = block_scan_counts(mine, &total);
  slicePrefix[slice] = excl;
  sliceFn[slice] = (uint8_t)start;  // from here on this array holds the slice's start STATE
  __syncthreads();
  if (threadIdx.x == 0) blockCount[blockIdx.x] = total;
}

// One CTA: exclusive 64-bit bases (symbols, records, non-ACGT) per block + grand totals.
__global__ void __launch_bounds__(1024) fasta_chain_counts(const unsigned long long* __restrict__ blockCount,
                                                           unsigned long long* __restrict__ baseE, unsigned long long* __restrict__ baseR,
                                                           unsigned long long* __restrict__ baseN, unsigned nBlocks,
                                                           unsigned long long* __restrict__ totals) {
  __shared__ unsigned long long pe[1024], pr[1024], pn[1024];
  const unsigned per = (nBlocks + 1023) / 1024;
  const unsigned b0 = min(threadIdx.x * per, nBlocks), b1 = min(b0 + per, nBlocks);
  unsigned long long e = 0, r = 0, n = 0;
  for (unsigned b = b0; b < b1; b++) {
    const unsigned long long v = blockCount[b];
    e += v & FIELD;
    r += (v >> R_SHIFT) & FIELD;
    n += v >> N_SHIFT;
  }
  pe[threadIdx.x] = e;
  pr[threadIdx.x] = r;
  pn[threadIdx.x] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long ae = 0, ar = 0, an = 0;
    for (int t = 0; t < 1024; t++) {
      const unsigned long long te = pe[t], tr = pr[t], tn = pn[t];
      pe[t] = ae; pr[t] = ar; pn[t] = an;
      ae += te; ar += tr; an += tn;
    }
    totals[0] = ae;
    totals[1] = ar;
    totals[2] = an;
  }
  __syncthreads();
  e = pe[threadIdx.x];
  r = pr[threadIdx.x];
  n = pn[threadIdx.x];
  for (unsigned b = b0; b < b1; b++) {
    baseE[b] = e;
    baseR[b] = r;
    baseN[b] = n;
    const unsigned long long v = blockCount[b];
    e += v & FIELD;
    r += (v >> R_SHIFT) & FIELD;
    n += v >> N_SHIFT;
  }
}

// Symbol index of something the automaton produces = characters emitted before it + record
// starts before it - 1: every record start but the first puts the previous record's terminator
// in front of the new record (FastaVector.c:78-86), and the last record's terminator comes at
// the very end (:159-164).
__global__ void __launch_bounds__(FTHREADS) fasta_emit(const uint8_t* __restrict__ text, const uint8_t* __restrict__ sliceState,
                                                       const unsigned long long* __restrict__ slicePrefix,
                                                       const unsigned long long* __restrict__ baseE,
                                                       const unsigned long long* __restrict__ baseR,
                                                       const unsigned long long* __restrict__ baseN, uint32_t* __restrict__ packed,
                                                       unsigned long long* __restrict__ recByte, unsigned long long* __restrict__ recTerm,
                                                       unsigned long long* __restrict__ fix, unsigned long long totalE,
                                                       unsigned long long totalR, unsigned long long totalN, int* __restrict__ orphan) {
  const size_t slice = (size_t)blockIdx.x * FTHREADS + threadIdx.x;
  Slice s;
  s.load(text, slice);
  const unsigned long long pre = slicePrefix[slice];
  unsigned long long E = baseE[blockIdx.x] + (pre & FIELD);
  unsigned long long Rn = baseR[blockIdx.x] + ((pre >> R_SHIFT) & FIELD);
  unsigned long long N = baseN[blockIdx.x] + (pre >> N_SHIFT);
  unsigned state = sliceState[slice];
  unsigned long long curWord = ~0ull;
  uint32_t acc = 0;
  auto put = [&](unsigned long long idx, unsigned v) {
    const unsigned long long w = idx >> 4;
    if (w != curWord) {
      if (acc) atomicOr(packed + curWord, acc);
      curWord = w;
      acc = 0;
    }
    acc |= v << (2 * (unsigned)(idx & 15));
  };
#pragma unroll 4
  for (int i = 0; i < SLICE; i++) {
    const unsigned c = s.byte(i);
    if (starts_record(state, c)) {
      recByte[Rn] = (unsigned long long)slice * SLICE + i;
      if (Rn >= 1) {
        const unsigned long long idx = E + Rn - 1;
        recTerm[Rn] = idx;
        fix[N + Rn - 1] = idx << 8;  // '\0'
      } else {
        recTerm[0] = 0;
      }
      Rn++;
    } else if (emits(state, c)) {
      if (Rn == 0) {
        *orphan = 1;  // sequence data before any '>' header
      } else {
        const unsigned long long idx = E + Rn - 1;
        const unsigned v = base_code(c);
        if (v < 4u) {
          put(idx, v);
        } else {
          fix[N + Rn - 1] = idx << 8 | c;
          N++;
        }
      }
      E++;
    }
    state = fn_apply(byte_fn(c), state);
  }
  if (acc) atomicOr(packed + curWord, acc);
  if (slice == 0 && totalR >= 1) fix[totalN + totalR - 1] = (totalE + totalR - 1) << 8;  // the last record's terminator
}

// entries: symbol index << 8 | resolved 2-bit value
__global__ void fasta_apply_fixups(const unsigned long long* __restrict__ fix, unsigned long long n, uint32_t* __restrict__ packed) {
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
    const unsigned long long e = fix[i];
    const unsigned long long idx = e >> 8;
    const uint32_t v = (uint32_t)(e & 3u);
    if (v) atomicOr(packed + (idx >> 4), v << (2 * (unsigned)(idx & 15)));
  }
}

struct Scratch {  // device buffers and staging of one ingest, released on every exit path
  struct Stage {
    cudaStream_t stream = nullptr;
    void* pinned[2] = {nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    cudaEvent_t done = nullptr;
  };
  std::vector<void*> ptrs;
  std::vector<Stage> stage;
  ~Scratch() {
    for (Stage& w : stage) {
      if (w.stream) cudaStreamSynchronize(w.stream);
      for (void* p : w.pinned) if (p) cudaFreeHost(p);
      for (cudaEvent_t e : w.ev) if (e) cudaEventDestroy(e);
      if (w.done) cudaEventDestroy(w.done);
      if (w.stream) cudaStreamDestroy(w.stream);
    }
    for (void* p : ptrs) cudaFree(p);
  }
};

}  // namespace

extern "C" {

int havac_cuda_ingest_fasta(havac_cuda_client* c, const uint8_t* text, uint64_t nBytes, havac_cuda_symbol_fn resolve,
                            void* user, uint64_t* nSymbols, uint64_t* nRecords) {
  if (!c || (!text && nBytes) || !resolve) return HAVAC_CUDA_E_INVALID;
  if (c->launched && !c->finished) return havac_fail(c, HAVAC_CUDA_E_LOGIC, "a run is in flight");
  if (nBytes >= (1ull << 36)) return havac_fail(c, HAVAC_CUDA_E_LENGTH, "FASTA text of 64 GiB or more");
  DeviceCtx& d = c->devs[0];
  CK(cudaSetDevice(d.dev));
  cudaStream_t st = d.stream;
  Scratch scratch;
  auto dalloc = [&](void** p, size_t bytes) {
    cudaError_t e = cudaMalloc(p, bytes ? bytes : 16);
    if (e == cudaSuccess) scratch.ptrs.push_back(*p);
    return e;
  };

  const uint64_t nBlocks64 = (nBytes + BLOCK_BYTES - 1) / BLOCK_BYTES;
  if (nBlocks64 == 0) return havac_fail(c, HAVAC_CUDA_E_RUNTIME, "Error while reading from the opened fasta file: no '>' record header before sequence data.");
  const unsigned nBlocks = (unsigned)nBlocks64;
  const size_t padded = (size_t)nBlocks * BLOCK_BYTES, nSlices = (size_t)nBlocks * FTHREADS;

  // ---- text to the device through two pinned staging buffers (copy-in overlaps the DMA) ----------
  uint8_t* dText = nullptr;
  CK(dalloc((void**)&dText, padded));
  if (padded > nBytes) CK(cudaMemsetAsync(dText + nBytes, 0, padded - nBytes, st));  // NUL: no state change, no symbol
  {
    // Worker threads take 4 MB pieces round-robin, each with its own stream and pair of pinned
    // buffers: one thread's copy out of the page cache (~2.5 GB/s, page faults included) no longer
    // bounds the upload. The client's stream then waits for every worker's last copy.
    const size_t PIECE = 4u << 20;
    const uint64_t nPieces = (nBytes + PIECE - 1) / PIECE;
    unsigned workers = 1 + (unsigned)std::min<uint64_t>(3, nBytes >> 28);  // setting a worker up costs a few ms: one per 256 MB, 4 at most
    if (const char* env = std::getenv("HAVAC_INGEST_THREADS")) workers = (unsigned)std::max(1, std::atoi(env));
    workers = (unsigned)std::min<uint64_t>(std::min(workers, 16u), nPieces);
    scratch.stage.resize(workers);
    std::vector<cudaError_t> result(workers, cudaSuccess);
    auto work = [&](unsigned t) {
      Scratch::Stage& w = scratch.stage[t];
      cudaError_t e = cudaSetDevice(d.dev);
      if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking);
      for (int i = 0; i < 2 && e == cudaSuccess; i++) {
        e = cudaHostAlloc(&w.pinned[i], PIECE, cudaHostAllocDefault);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&w.ev[i], cudaEventDisableTiming);
      }
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&w.done, cudaEventDisableTiming);
      int k = 0;
      for (uint64_t piece = t; piece < nPieces && e == cudaSuccess; piece += workers, k ^= 1) {
        const uint64_t off = piece * PIECE;
        const size_t n = (size_t)std::min<uint64_t>(PIECE, nBytes - off);
        e = cudaEventSynchronize(w.ev[k]);  // the copy that last used this buffer is done
        if (e != cudaSuccess) break;
        std::memcpy(w.pinned[k], text + off, n);
        e = cudaMemcpyAsync(dText + off, w.pinned[k], n, cudaMemcpyHostToDevice, w.stream);
        if (e == cudaSuccess) e = cudaEventRecord(w.ev[k], w.stream);
      }
      if (e == cudaSuccess) e = cudaEventRecord(w.done, w.stream);
      result[t] = e;
    };
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < workers; t++) pool.emplace_back(work, t);
    work(0);
    for (std::thread& th : pool) th.join();
    for (unsigned t = 0; t < workers; t++) CK(result[t]);
    for (unsigned t = 0; t < workers; t++) CK(cudaStreamWaitEvent(st, scratch.stage[t].done, 0));
    c->h2dBytes += nBytes;
  }

  // ---- states, counts ---------------------------------------------------------------------------
  uint8_t *dSliceFn = nullptr, *dBlockFn = nullptr, *dBlockState = nullptr;
  unsigned long long *dSlicePrefix = nullptr, *dBlockCount = nullptr, *dBaseE = nullptr, *dBaseR = nullptr, *dBaseN = nullptr,
                     *dTotals = nullptr;
  int* dOrphan = nullptr;
  CK(dalloc((void**)&dSliceFn, nSlices));
  CK(dalloc((void**)&dBlockFn, nBlocks));
  CK(dalloc((void**)&dBlockState, nBlocks));
  CK(dalloc((void**)&dSlicePrefix, nSlices * 8));
  CK(dalloc((void**)&dBlockCount, (size_t)nBlocks * 8));
  CK(dalloc((void**)&dBaseE, (size_t)nBlocks * 8));
  CK(dalloc((void**)&dBaseR, (size_t)nBlocks * 8));
  CK(dalloc((void**)&dBaseN, (size_t)nBlocks * 8));
  CK(dalloc((void**)&dTotals, 3 * 8));
  CK(dalloc((void**)&dOrphan, sizeof(int)));
  CK(cudaMemsetAsync(dOrphan, 0, sizeof(int), st));
  fasta_scan_states<<<nBlocks, FTHREADS, 0, st>>>(dText, dSliceFn, dBlockFn);
  fasta_chain_blocks<<<1, 1024, 0, st>>>(dBlockFn, dBlockState, nBlocks);
  fasta_count<<<nBlocks, FTHREADS, 0, st>>>(dText, dSliceFn, dBlockState, dSlicePrefix, dBlockCount);
  fasta_chain_counts<<<1, 1024, 0, st>>>(dBlockCount, dBaseE, dBaseR, dBaseN, nBlocks, dTotals);
  CK(cudaGetLastError());
  unsigned long long totals[3] = {0, 0, 0};
  CK(cudaMemcpyAsync(totals, dTotals, sizeof totals, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  const uint64_t totalE = totals[0], totalR = totals[1], totalN = totals[2];
  if (totalR == 0)
    return havac_fail(c, HAVAC_CUDA_E_RUNTIME, "Error while reading from the opened fasta file: no '>' record header before sequence data.");
  const uint64_t symbols = totalE + totalR;  // one terminator per record (FastaVector.c:78-86,159-164)
  const uint64_t segments = (symbols + SEG - 1) / SEG;
  const uint64_t packedBytes = segments * SEG / 4;
  if (packedBytes >= 4ull * 1024 * 1024 * 1024)  // host/HavacHwClient.cpp:92-97
    return havac_fail(c, HAVAC_CUDA_E_LENGTH,
                      "compressed sequence size must be less than 4GiB. length requested: " + std::to_string(packedBytes) + " bytes.");

  // ---- symbols ------------------------------------------------------------------------------------
  uint32_t* dPacked = nullptr;
  unsigned long long *dRecByte = nullptr, *dRecTerm = nullptr, *dFix = nullptr;
  const uint64_t nFix = totalN + totalR;
  CK(cudaMalloc((void**)&dPacked, packedBytes));  // becomes the client's sequence buffer on success
  auto dropPacked = [&](int code) { cudaFree(dPacked); return code; };
  auto ck2 = [&](cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    c->err = std::string(what) + ": " + cudaGetErrorString(e);
    return false;
  };
  if (!ck2(dalloc((void**)&dRecByte, totalR * 8), "cudaMalloc") || !ck2(dalloc((void**)&dRecTerm, totalR * 8), "cudaMalloc") ||
      !ck2(dalloc((void**)&dFix, nFix * 8), "cudaMalloc"))
    return dropPacked(HAVAC_CUDA_E_NOMEM);
  if (!ck2(cudaMemsetAsync(dPacked, 0, packedBytes, st), "cudaMemsetAsync")) return dropPacked(HAVAC_CUDA_E_RUNTIME);
  fasta_emit<<<nBlocks, FTHREADS, 0, st>>>(dText, dSliceFn, dSlicePrefix, dBaseE, dBaseR, dBaseN, dPacked, dRecByte, dRecTerm, dFix,
                                           totalE, totalR, totalN, dOrphan);
  std::vector<unsigned long long> fix(nFix), recByte(totalR), recTerm(totalR);
  int orphan = 0;
  if (!ck2(cudaGetLastError(), "fasta_emit") ||
      !ck2(cudaMemcpyAsync(fix.data(), dFix, nFix * 8, cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync") ||
      !ck2(cudaMemcpyAsync(recByte.data(), dRecByte, totalR * 8, cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync") ||
      !ck2(cudaMemcpyAsync(recTerm.data(), dRecTerm, totalR * 8, cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync") ||
      !ck2(cudaMemcpyAsync(&orphan, dOrphan, sizeof orphan, cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync") ||
      !ck2(cudaStreamSynchronize(st), "cudaStreamSynchronize"))
    return dropPacked(HAVAC_CUDA_E_RUNTIME);
  c->d2hBytes += (nFix + 2 * totalR) * 8;
  if (orphan) {
    cudaFree(dPacked);
    return havac_fail(c, HAVAC_CUDA_E_RUNTIME, "Error while reading from the opened fasta file: no '>' record header before sequence data.");
  }
  // non-ACGT symbols and terminators, resolved by the caller IN SEQUENCE ORDER (it owns the rand() state)
  for (uint64_t i = 0; i < nFix; i++) fix[i] = (fix[i] & ~0xffull) | (resolve(user, (unsigned char)(fix[i] & 0xff)) & 3u);
  if (!ck2(cudaMemcpyAsync(dFix, fix.data(), nFix * 8, cudaMemcpyHostToDevice, st), "cudaMemcpyAsync")) return dropPacked(HAVAC_CUDA_E_RUNTIME);
  c->h2dBytes += nFix * 8;
  {
    const unsigned grid = (unsigned)std::min<uint64_t>((nFix + 255) / 256, 148 * 8);
    fasta_apply_fixups<<<grid ? grid : 1, 256, 0, st>>>(dFix, nFix, dPacked);
  }
  if (!ck2(cudaGetLastError(), "fasta_apply_fixups") || !ck2(cudaStreamSynchronize(st), "cudaStreamSynchronize"))
    return dropPacked(HAVAC_CUDA_E_RUNTIME);

  // ---- record table ---------------------------------------------------------------------------------
  c->ingestEnd.assign(totalR, 0);
  c->ingestHeaderByte.assign(recByte.begin(), recByte.end());
  for (uint64_t r = 1; r < totalR; r++) c->ingestEnd[r - 1] = recTerm[r] + 1;
  c->ingestEnd[totalR - 1] = symbols;
  if (nSymbols) *nSymbols = symbols;
  if (nRecords) *nRecords = totalR;

  // ---- install as the client's sequence ----------------------------------------------------------------
  const bool whole = c->devs.size() == 1 && c->segBegin == 0 && (c->segEnd == 0 || c->segEnd >= segments);
  if (whole) {
    havac_free_sequence(d);
    d.dSeq = dPacked;
    d.seqCap = packedBytes;
    d.colBegin = d.resBegin = 0;
    d.colEnd = d.resEnd = (long long)(segments * SEG);
    c->seqSegments = segments;
    c->mirrorSymbols = 0;
    return HAVAC_CUDA_OK;
  }
  // sharded clients: ranges and halos as write_sequence lays them out, copied device to device
  const int rc = havac_place_sequence(c, (const uint8_t*)dPacked, d.dev, packedBytes);
  cudaSetDevice(d.dev);
  cudaFree(dPacked);
  return rc;
}

int havac_cuda_ingest_records(havac_cuda_client* c, uint64_t* sequenceEnd, uint64_t* headerByteOffset, uint64_t cap, uint64_t* n) {
  if (!c) return HAVAC_CUDA_E_INVALID;
  if (n) *n = c->ingestEnd.size();
  if (cap < c->ingestEnd.size()) return cap == 0 ? HAVAC_CUDA_OK : havac_fail(c, HAVAC_CUDA_E_LENGTH, "record table larger than the buffers given");
  for (size_t i = 0; i < c->ingestEnd.size(); i++) {
    if (sequenceEnd) sequenceEnd[i] = c->ingestEnd[i];
    if (headerByteOffset) headerByteOffset[i] = c->ingestHeaderByte[i];
  }
  return HAVAC_CUDA_OK;
}

int havac_cuda_read_sequence(havac_cuda_client* c, uint8_t* packedOut, uint64_t capBytes, uint64_t* nBytes) {
  if (!c) return HAVAC_CUDA_E_INVALID;
  if (c->devs.size() != 1) return havac_fail(c, HAVAC_CUDA_E_LOGIC, "read_sequence is for single-device clients");
  DeviceCtx& d = c->devs[0];
  const uint64_t bytes = (uint64_t)(d.resEnd - d.resBegin) / 4;
  if (nBytes) *nBytes = bytes;
  if (!packedOut || capBytes == 0) return HAVAC_CUDA_OK;
  if (capBytes < bytes) return havac_fail(c, HAVAC_CUDA_E_LENGTH, "buffer smaller than the resident sequence");
  if (bytes == 0) return HAVAC_CUDA_OK;
  CK(cudaSetDevice(d.dev));
  CK(cudaMemcpy(packedOut, d.dSeq, bytes, cudaMemcpyDeviceToHost));
  c->d2hBytes += bytes;
  return HAVAC_CUDA_OK;
}

}  // extern "C"
