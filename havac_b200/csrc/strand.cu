// strand.cu -- the other strand: the resident 2-bit sequence is replaced, on the
// device, by its reverse complement, so that the same SSV kernel searches the Crick strand the
// reference leaves out (its benchmark pins nhmmer to --watson, benchmark/readme.txt:62-64).
//
// With A,C,G,T packed as 0..3 (host/sequence/SequencePreprocessor.cpp:62-85) the complement is the
// bitwise NOT of a symbol, so for a buffer of N real symbols (terminators included) followed by
// zero padding up to whole segments:   out[k] = ~in[N-1-k] for k < N,   out[k] = in[k] for k >= N.
// One thread builds one 32-bit word (16 symbols) from the two input words that hold its source
// symbols. Calling it twice restores the buffer. A client that spreads the sequence over several
// GPUs of this process gathers the shards on its first device over peer copies, reverses there,
// and scatters the ranges (halo included) back: nothing goes through the host.
#include <cuda_runtime.h>

#include <algorithm>
#include <string>
#include <vector>

#include "client_internal.cuh"

namespace {

__device__ __forceinline__ uint32_t reverse_symbols(uint32_t x) {  // symbol j <-> symbol 15-j
  x = __brev(x);                                                    // reverses the two bits of each symbol too
  return ((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1);
}

__global__ void reverse_complement_words(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, unsigned long long nSymbols,
                                         unsigned long long nWords) {
  for (unsigned long long w = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; w < nWords;
       w += (unsigned long long)gridDim.x * blockDim.x) {
    const unsigned long long first = 16 * w;
    const uint32_t keep = in[w];
    if (first >= nSymbols) { out[w] = keep; continue; }  // padding stays where it is
    const unsigned long long hi = nSymbols - 1 - first;  // source of this word's symbol 0; symbol j comes from hi - j
    const unsigned long long b = hi >> 4;
    const uint64_t window = ((uint64_t)in[b] << 32) | (b ? in[b - 1] : 0u);
    const uint32_t x = (uint32_t)(window >> (2 * ((unsigned)(hi & 15) + 1)));  // symbol 15 of x is source hi
    uint32_t r = ~reverse_symbols(x);
    const unsigned long long m = nSymbols - first;  // symbols of this word that are real
    if (m < 16) {
      const uint32_t mask = (1u << (2 * (unsigned)m)) - 1u;
      r = (r & mask) | (keep & ~mask);
    }
    out[w] = r;
  }
}

}  // namespace

extern "C" int havac_cuda_reverse_complement(havac_cuda_client* c, uint64_t nSymbols) {
  if (!c) return HAVAC_CUDA_E_INVALID;
  if (c->launched && !c->finished) return havac_fail(c, HAVAC_CUDA_E_LOGIC, "a run is in flight");
  const uint64_t total = c->seqSegments * SEG;
  if (total == 0) return havac_fail(c, HAVAC_CUDA_E_LENGTH, "no sequence is resident");
  if (nSymbols > total) return havac_fail(c, HAVAC_CUDA_E_LENGTH, "more symbols than the resident sequence holds");
  if (c->segBegin != 0 || (c->segEnd != 0 && c->segEnd < c->seqSegments))
    return havac_fail(c, HAVAC_CUDA_E_LOGIC,
                      "this client holds a segment range only: reverse the whole packed buffer first and write that");
  if (c->mirrorSymbols != 0 && c->mirrorSymbols != nSymbols)
    return havac_fail(c, HAVAC_CUDA_E_LOGIC, "the resident sequence was reversed over a different symbol count");

  DeviceCtx& d0 = c->devs[0];
  const size_t bytes = (size_t)(total / 4);
  const unsigned long long nWords = total / 16;
  CK(cudaSetDevice(d0.dev));
  uint32_t* rev = nullptr;
  CK(cudaMalloc((void**)&rev, bytes));
  uint8_t* fwd = nullptr;  // the whole forward buffer on the first device
  bool gathered = false;
  auto drop = [&](int code) {
    cudaSetDevice(d0.dev);
    cudaFree(rev);
    if (gathered) cudaFree(fwd);
    return code;
  };
  auto bad = [&](cudaError_t e, const char* what) { return drop(havac_fail(c, HAVAC_CUDA_E_RUNTIME, std::string(what) + ": " + cudaGetErrorString(e))); };

  if (c->devs.size() == 1) {
    fwd = (uint8_t*)d0.dSeq;
  } else {
    cudaError_t e = cudaMalloc((void**)&fwd, bytes);
    if (e != cudaSuccess) return drop(havac_fail(c, HAVAC_CUDA_E_NOMEM, "cudaMalloc: out of device memory"));
    gathered = true;
    for (DeviceCtx& d : c->devs) {
      if (d.colEnd <= d.colBegin) continue;
      e = cudaMemcpyPeerAsync(fwd + d.colBegin / 4, d0.dev, (const uint8_t*)d.dSeq + (d.colBegin - d.resBegin) / 4, d.dev,
                              (size_t)(d.colEnd - d.colBegin) / 4, d0.stream);
      if (e != cudaSuccess) return bad(e, "cudaMemcpyPeerAsync");
    }
  }
  {
    const unsigned grid = (unsigned)std::min<unsigned long long>((nWords + 255) / 256, 148ull * 32);
    reverse_complement_words<<<grid, 256, 0, d0.stream>>>((const uint32_t*)fwd, rev, nSymbols, nWords);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return bad(e, "reverse_complement_words");
  }
  // everything -- gather, reversal, scatter -- is ordered on the first device's stream and finished
  // before this call returns, so no other stream of any device can see a half-written buffer
  if (c->devs.size() > 1)
    for (DeviceCtx& d : c->devs) {
      if (d.colEnd <= d.colBegin) continue;
      cudaError_t e = cudaMemcpyPeerAsync(d.dSeq, d.dev, (const uint8_t*)rev + d.resBegin / 4, d0.dev,
                                          (size_t)(d.resEnd - d.resBegin) / 4, d0.stream);
      if (e != cudaSuccess) return bad(e, "cudaMemcpyPeerAsync");
    }
  {
    cudaError_t e = cudaStreamSynchronize(d0.stream);
    if (e != cudaSuccess) return bad(e, "reverse complement");
  }
  if (c->devs.size() == 1) {
    cudaFree(d0.dSeq);
    d0.dSeq = rev;
    rev = nullptr;
  } else {
    drop(0);
  }
  c->mirrorSymbols = c->mirrorSymbols ? 0 : nSymbols;
  return HAVAC_CUDA_OK;
}

extern "C" int havac_cuda_strand_state(havac_cuda_client* c, uint64_t* mirroredSymbols) {
  if (!c || !mirroredSymbols) return HAVAC_CUDA_E_INVALID;
  *mirroredSymbols = c->mirrorSymbols;
  return HAVAC_CUDA_OK;
}
