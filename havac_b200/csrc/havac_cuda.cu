// havac_cuda.cu -- implementation of the C-ABI device client declared in include/havac_cuda.h.
// It stands where the reference's HavacHwClient talks to XRT (host/HavacHwClient.cpp): device
// buffers, asynchronous launch, state/wait/abort, hit read-back. One client may drive several
// GPUs of this process; the sequence axis is split on segment boundaries with a left halo of
// (rows - 1) symbols per shard, which makes every shard's hits exact without any exchange
// between devices (each diagonal is an independent chain).
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "client_internal.cuh"
#include "ssv_kernels.cuh"

namespace {

thread_local std::string g_createError;

}  // namespace

int havac_fail(havac_cuda_client* c, int code, const std::string& msg) {
  if (c) c->err = msg;
  else g_createError = msg;
  return code;
}

namespace {

inline int fail(havac_cuda_client* c, int code, const std::string& msg) { return havac_fail(c, code, msg); }

bool parse_spec(const char* spec, uint32_t deviceIndex, int nDev, std::vector<int>& out,
                unsigned long long& hitCap, bool& hitCapGrows, std::string& err) {
  std::string s = spec ? spec : "";
  std::string devPart = s;
  size_t semi = s.find(';');
  if (semi != std::string::npos) {
    devPart = s.substr(0, semi);
    std::string opt = s.substr(semi + 1);
    if (opt.rfind("hitcap=", 0) == 0) {  // "hitcap=N": fixed, overflow is an error; "hitcap=N+": first size, grows
      hitCap = strtoull(opt.c_str() + 7, nullptr, 10);
      hitCapGrows = !opt.empty() && opt.back() == '+';
    }
    if (hitCap == 0) { err = "hitcap must be positive"; return false; }
  }
  if (devPart.rfind("cuda:", 0) != 0) {  // an xclbin path or empty: single device
    out.push_back((int)deviceIndex);
  } else {
    std::string list = devPart.substr(5);
    if (list == "all") {
      for (int i = 0; i < nDev; i++) out.push_back(i);
    } else {
      size_t pos = 0;
      while (pos <= list.size()) {
        size_t comma = list.find(',', pos);
        std::string tok = list.substr(pos, comma == std::string::npos ? std::string::npos : comma - pos);
        if (tok.empty() || tok.find_first_not_of("0123456789") != std::string::npos) {
          err = "bad device list '" + list + "'";
          return false;
        }
        out.push_back(atoi(tok.c_str()));
        if (comma == std::string::npos) break;
        pos = comma + 1;
      }
    }
  }
  for (int d : out)
    if (d < 0 || d >= nDev) {
      err = "CUDA device " + std::to_string(d) + " not present (" + std::to_string(nDev) + " visible)";
      return false;
    }
  return !out.empty();
}

inline void free_sequence(DeviceCtx& d) { havac_free_sequence(d); }
void free_table(DeviceCtx& d) {
  cudaSetDevice(d.dev);
  if (d.dTab8) { cudaFree(d.dTab8); d.dTab8 = nullptr; }
  if (d.dTab16) { cudaFree(d.dTab16); d.dTab16 = nullptr; }
  d.tabCap = 0;
}

// Queues one search on every device: counters cleared, start event, kernel, stop event.
int launch_all(havac_cuda_client* c) {
  c->launches = 0;
  for (DeviceCtx& d : c->devs) {
    CK(cudaSetDevice(d.dev));
    d.hitsSorted = false;
    CK(cudaMemsetAsync(d.dCtl, 0, sizeof(DeviceResultBlock), d.stream));  // counters, status and the abort flag
    CK(cudaEventRecord(d.evStart, d.stream));
    if (d.colEnd > d.colBegin) {
      SsvParams p{};
      const long long P = c->nRows;
      long long loadBegin = std::max(d.resBegin, (d.colBegin - (P - 1)) & ~63ll);
      if (loadBegin < 0) loadBegin = 0;
      p.seqPacked = d.dSeq + (loadBegin - d.resBegin) / 16;
      p.nLoad = d.resEnd - loadBegin;
      p.loadBegin = loadBegin;
      p.colBegin = d.colBegin;
      p.colEnd = d.colEnd;
      p.tab16 = d.dTab16;
      p.tab8 = d.dTab8;
      p.nRows = c->nRows;
      p.rowsPadded = d.rowsPadded;
      long long dLo = d.colBegin - loadBegin - (P - 1);  // first diagonal that can reach a reported column
      p.diagBegin = dLo >= 0 ? (dLo & ~63ll) : -((-dLo + 63) & ~63ll);
      p.nTiles = (uint32_t)((p.nLoad - p.diagBegin + SSV_TILE_DIAGS - 1) / SSV_TILE_DIAGS);
      p.hits = d.dHits;
      p.hitCap = c->hitCap;
      p.hitCount = d.dHitCount;
      p.tileCounter = d.dTileCounter;
      p.abortFlag = d.dAbort;
      p.status = d.dStatus;
      p.smCycles = d.dCycles;
      d.lastLoadBegin = loadBegin;
      d.lastDiagBegin = p.diagBegin;
      const bool streamingOnly = c->kernel == KERNEL_F16X2_STREAMING;
      const bool warpTiles = c->kernel != KERNEL_SCALAR && ssv_takes_short_kernel(p.nRows, streamingOnly);
      d.lastTiles = warpTiles ? p.nTiles * (SSV_TILE_DIAGS / SSV_WARP_TILE_DIAGS) : p.nTiles;
      d.lastTileDiags = warpTiles ? SSV_WARP_TILE_DIAGS : SSV_TILE_DIAGS;
      if (c->kernel == KERNEL_SCALAR) ssv_launch_scalar(p, d.stream);
      else ssv_launch_f16x2(p, d.numSMs, d.stream, streamingOnly);
      CK(cudaGetLastError());
      c->launches++;
    }
    CK(cudaEventRecord(d.evStop, d.stream));
    // the results follow the stop event to pinned memory; nobody blocks in a cudaMemcpy later
    CK(cudaMemcpyAsync(d.hostRes, d.dCtl, RESULT_BYTES, cudaMemcpyDeviceToHost, d.stream));
    CK(cudaEventRecord(d.evDone, d.stream));
  }
  return HAVAC_CUDA_OK;
}

// Enlarges every device's hit buffer to hold `needed` hits (plus slack) if device memory allows.
bool grow_hit_buffers(havac_cuda_client* c, unsigned long long needed) {
  const unsigned long long newCap = needed + needed / 8 + 1024;
  for (DeviceCtx& d : c->devs) {
    if (cudaSetDevice(d.dev) != cudaSuccess) return false;
    size_t freeB = 0, totalB = 0;
    if (cudaMemGetInfo(&freeB, &totalB) != cudaSuccess) return false;
    // the sort of the hits needs a second buffer of the same size later on
    if ((double)newCap * 8.0 * 2.2 > (double)freeB) return false;
  }
  // the larger buffers are allocated BEFORE the old ones are freed: a failure leaves every device as it was
  std::vector<unsigned long long*> fresh(c->devs.size(), nullptr);
  bool ok = true;
  for (size_t i = 0; i < c->devs.size() && ok; i++) {
    cudaSetDevice(c->devs[i].dev);
    if (cudaMalloc(&fresh[i], newCap * sizeof(unsigned long long)) != cudaSuccess) { fresh[i] = nullptr; cudaGetLastError(); ok = false; }
  }
  for (size_t i = 0; i < c->devs.size(); i++) {
    cudaSetDevice(c->devs[i].dev);
    if (ok) { cudaFree(c->devs[i].dHits); c->devs[i].dHits = fresh[i]; }
    else if (fresh[i]) cudaFree(fresh[i]);
  }
  if (ok) c->hitCap = newCap;
  return ok;
}

// Pull the per-device results after the stop events have fired; decides the final state. A hit
// buffer that overflowed is enlarged to the count the kernel reported and the search repeated
// (unless the capacity was fixed with ";hitcap=N"): with mayBlock the repeat is waited for here,
// otherwise the client goes back to "running".
int collect(havac_cuda_client* c, bool mayBlock = true) {
  while (!c->finished) {
    float worst = 0.f;
    bool overflow = false, aborted = false;
    unsigned long long most = 0;
    for (DeviceCtx& d : c->devs) {
      CK(cudaSetDevice(d.dev));
      CK(cudaEventSynchronize(d.evDone));
      d.hitCount = d.hostRes->hitCount;
      d.status = d.hostRes->status;
      d.cycles = d.hostRes->cycles;
      c->d2hBytes += RESULT_BYTES;
      float ms = 0.f;
      CK(cudaEventElapsedTime(&ms, d.evStart, d.evStop));
      worst = std::max(worst, ms);
      overflow |= (d.status & SSV_STATUS_OVERFLOW) != 0 || d.hitCount > c->hitCap;
      aborted |= (d.status & SSV_STATUS_ABORTED) != 0;
      most = std::max<unsigned long long>(most, d.hitCount);
    }
    c->lastMs += worst;  // a repeated search counts in full
    // a stop that was asked for wins over the repeat: launch_all would clear the abort flag again
    if (overflow && c->abortRequested) aborted = true;
    if (overflow && !aborted && c->hitCapGrows && c->regrowths < 2) {
      if (grow_hit_buffers(c, most)) {
        c->regrowths++;
        int rc = launch_all(c);
        if (rc) return rc;
        if (!mayBlock) { c->finalState = HAVAC_CUDA_STATE_RUNNING; return HAVAC_CUDA_OK; }
        continue;
      }
    }
    c->finished = true;
    c->aborted = aborted;
    if (aborted) c->finalState = HAVAC_CUDA_STATE_ABORT;
    else if (overflow) {
      c->finalState = HAVAC_CUDA_STATE_ERROR;
      c->err = "device hit buffer overflowed (capacity " + std::to_string(c->hitCap) + " hits per device, " + std::to_string(most) +
               " needed)";
    } else c->finalState = HAVAC_CUDA_STATE_COMPLETED;
  }
  return HAVAC_CUDA_OK;
}

}  // namespace

void havac_free_sequence(DeviceCtx& d) {
  if (d.dSeq) { cudaSetDevice(d.dev); cudaFree(d.dSeq); d.dSeq = nullptr; d.seqCap = 0; }
}

// The packed sequence goes to the devices: every device gets its column range plus the resident
// halo in front of it. The source is a host buffer (srcDevice < 0: write_sequence, blocking like
// the reference's bo.sync, HavacHwClient.cpp:104) or a buffer on one of this process's GPUs (the
// device-side FASTA ingest: shards travel by peer copies, nothing through the host).
int havac_place_sequence(havac_cuda_client* c, const uint8_t* packed, int srcDevice, uint64_t nBytes) {
  // host/HavacHwClient.cpp:81-97: whole segments only, and strictly less than 4 GiB packed
  uint64_t nSymbols = nBytes * 4;
  if (nSymbols % SEG != 0)
    return fail(c, HAVAC_CUDA_E_LENGTH, "sequence length must be a multiple of the sequence segment length " +
                                            std::to_string(SEG) + " but sequence of length " + std::to_string(nSymbols) + " was not");
  if (nBytes >= 4ull * 1024 * 1024 * 1024)
    return fail(c, HAVAC_CUDA_E_LENGTH, "compressed sequence size must be less than 4GiB. length requested: " +
                                            std::to_string(nBytes) + " bytes.");
  if (c->launched && !c->finished) return fail(c, HAVAC_CUDA_E_LOGIC, "a run is in flight");
  c->seqSegments = nSymbols / SEG;
  c->mirrorSymbols = 0;
  uint64_t sb = c->segBegin, se = c->segEnd ? c->segEnd : c->seqSegments;
  if (se > c->seqSegments) se = c->seqSegments;
  if (sb > se) sb = se;
  const uint64_t nSeg = se - sb;
  const size_t nd = c->devs.size();
  for (size_t i = 0; i < nd; i++) {
    DeviceCtx& d = c->devs[i];
    d.colBegin = (long long)((sb + nSeg * i / nd) * SEG);
    d.colEnd = (long long)((sb + nSeg * (i + 1) / nd) * SEG);
    d.resBegin = std::max(0ll, d.colBegin - HALO_RESIDENT) & ~63ll;
    d.resEnd = d.colEnd;
    if (d.colEnd <= d.colBegin) { d.resBegin = d.resEnd = d.colBegin; free_sequence(d); continue; }
    size_t bytes = (size_t)(d.resEnd - d.resBegin) / 4;
    CK(cudaSetDevice(d.dev));
    if (bytes != d.seqCap) {  // a reload of the same size (the common case) reuses the allocation
      free_sequence(d);
      CK(cudaMalloc(&d.dSeq, bytes));
      d.seqCap = bytes;
    }
    if (srcDevice < 0) {
      CK(cudaMemcpyAsync(d.dSeq, packed + d.resBegin / 4, bytes, cudaMemcpyHostToDevice, d.stream));
      c->h2dBytes += bytes;
    } else {
      CK(cudaMemcpyPeerAsync(d.dSeq, d.dev, packed + d.resBegin / 4, srcDevice, bytes, d.stream));
    }
  }
  for (DeviceCtx& d : c->devs) {
    CK(cudaSetDevice(d.dev));
    CK(cudaStreamSynchronize(d.stream));
  }
  return HAVAC_CUDA_OK;
}

extern "C" {

int havac_cuda_create(const char* deviceSpec, uint32_t deviceIndex, havac_cuda_client** out) {
  if (!out) return fail(nullptr, HAVAC_CUDA_E_INVALID, "null output handle");
  *out = nullptr;
  int nDev = 0;
  cudaError_t e = cudaGetDeviceCount(&nDev);
  if (e != cudaSuccess || nDev == 0)
    return fail(nullptr, HAVAC_CUDA_E_NODEVICE,
                std::string("no CUDA device available: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "count is 0"));
  std::vector<int> ids;
  unsigned long long hitCap = 16ull << 20;
  std::string perr;
  bool hitCapGrows = true;
  if (!parse_spec(deviceSpec, deviceIndex, nDev, ids, hitCap, hitCapGrows, perr)) return fail(nullptr, HAVAC_CUDA_E_NODEVICE, perr);

  havac_cuda_client* c = new (std::nothrow) havac_cuda_client();
  if (!c) return fail(nullptr, HAVAC_CUDA_E_NOMEM, "out of host memory");
  c->hitCap = hitCap;
  c->hitCapGrows = hitCapGrows;
  auto bail = [&](int code) { g_createError = c->err; havac_cuda_destroy(c); return code; };
  {
    cudaError_t ee = cudaHostAlloc((void**)&c->hostOne, sizeof(int), cudaHostAllocPortable);
    if (ee != cudaSuccess) { c->err = std::string("cudaHostAlloc: ") + cudaGetErrorString(ee); return bail(HAVAC_CUDA_E_NOMEM); }
    *c->hostOne = 1;
  }
  for (int id : ids) {
    DeviceCtx d;
    d.dev = id;
    c->devs.push_back(d);
  }
  for (DeviceCtx& d : c->devs) {
    auto ck = [&](cudaError_t ee, const char* what) {
      if (ee == cudaSuccess) return true;
      c->err = std::string(what) + ": " + cudaGetErrorString(ee);
      return false;
    };
    cudaDeviceProp prop;
    if (!ck(cudaSetDevice(d.dev), "cudaSetDevice") || !ck(cudaGetDeviceProperties(&prop, d.dev), "cudaGetDeviceProperties"))
      return bail(HAVAC_CUDA_E_RUNTIME);
    if (prop.major != 10) {  // sm_100a SASS only: a later major would fail at the first launch with 'no kernel image'
      c->err = "device " + std::to_string(d.dev) + " (" + prop.name + ") is sm_" + std::to_string(prop.major) +
               std::to_string(prop.minor) + "; this library carries sm_100a code only";
      return bail(HAVAC_CUDA_E_NODEVICE);
    }
    d.numSMs = prop.multiProcessorCount;
    if (!ck(cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking), "cudaStreamCreate") ||
        !ck(cudaStreamCreateWithFlags(&d.ctlStream, cudaStreamNonBlocking), "cudaStreamCreate") ||
        !ck(cudaEventCreate(&d.evStart), "cudaEventCreate") || !ck(cudaEventCreate(&d.evStop), "cudaEventCreate") ||
        !ck(cudaEventCreateWithFlags(&d.evDone, cudaEventDisableTiming), "cudaEventCreate"))
      return bail(HAVAC_CUDA_E_RUNTIME);
    // the hit buffers are allocated once, like HavacHwClient::allocateOutputBuffers (HavacHwClient.cpp:73-76)
    if (!ck(cudaMalloc(&d.dHits, c->hitCap * sizeof(unsigned long long)), "cudaMalloc(hits)") ||
        !ck(cudaMalloc(&d.dCtl, sizeof(DeviceResultBlock)), "cudaMalloc") ||
        !ck(cudaHostAlloc((void**)&d.hostRes, sizeof(DeviceResultBlock), cudaHostAllocPortable), "cudaHostAlloc"))
      return bail(HAVAC_CUDA_E_NOMEM);
    d.dHitCount = &d.dCtl->hitCount;
    d.dCycles = &d.dCtl->cycles;
    d.dStatus = &d.dCtl->status;
    d.dTileCounter = &d.dCtl->tileCounter;
    d.dAbort = &d.dCtl->abort;
    // the SSV kernels' attributes are set (which also loads their code onto this device) while the device is being
    // opened, like the xclbin load of HavacHwClient::generateKernel: the first search does not pay for it
    ssv_f16x2_ctas_per_sm();
    havac_postprocess_preload();
  }
  *out = c;
  return HAVAC_CUDA_OK;
}

int havac_cuda_destroy(havac_cuda_client* c) {
  if (!c) return HAVAC_CUDA_E_INVALID;
  for (DeviceCtx& d : c->devs) {
    cudaSetDevice(d.dev);
    if (c->launched && !c->finished && d.dAbort && d.ctlStream)  // let a still-running kernel drain quickly
      c->abortRequested = true, cudaMemcpyAsync(d.dAbort, c->hostOne, sizeof(int), cudaMemcpyHostToDevice, d.ctlStream);
    if (d.stream) cudaStreamSynchronize(d.stream);
    free_sequence(d);
    free_table(d);
    if (d.dHits) cudaFree(d.dHits);
    if (d.sortScratch) cudaFree(d.sortScratch);
    if (d.dCtl) cudaFree(d.dCtl);
    if (d.hostRes) cudaFreeHost(d.hostRes);
    if (d.evDone) cudaEventDestroy(d.evDone);
    if (d.ctlStream) cudaStreamDestroy(d.ctlStream);
    if (d.evStart) cudaEventDestroy(d.evStart);
    if (d.evStop) cudaEventDestroy(d.evStop);
    if (d.stream) cudaStreamDestroy(d.stream);
  }
  if (c->hostOne) cudaFreeHost(c->hostOne);
  delete c;
  return HAVAC_CUDA_OK;
}

int havac_cuda_set_segment_range(havac_cuda_client* c, uint64_t segBegin, uint64_t segEnd) {
  if (!c) return HAVAC_CUDA_E_INVALID;
  if (segEnd != 0 && segEnd <= segBegin) return fail(c, HAVAC_CUDA_E_LENGTH, "empty segment range");
  c->segBegin = segBegin;
  c->segEnd = segEnd;
  return HAVAC_CUDA_OK;
}

int havac_cuda_write_sequence(havac_cuda_client* c, const uint8_t* packed, uint64_t nBytes) {
  if (!c || (!packed && nBytes)) return HAVAC_CUDA_E_INVALID;
  return havac_place_sequence(c, packed, -1, nBytes);
}

int havac_cuda_write_phmm(havac_cuda_client* c, const int8_t* flat, uint64_t nBytes) {
  if (!c || (!flat && nBytes)) return HAVAC_CUDA_E_INVALID;
  // host/HavacHwClient.cpp:112-125
  if (nBytes % 4 != 0)
    return fail(c, HAVAC_CUDA_E_LENGTH, "phmm of length " + std::to_string(nBytes) +
                                            " was given, but was not divisible by 4. only nucleotide phmms with 4 scores/position is supported.");
  if (nBytes >= 1024ull * 1024 * 1024)
    return fail(c, HAVAC_CUDA_E_LENGTH, "model length must be less than 1GiB. length requested: " + std::to_string(nBytes) + " bytes.");
  if (nBytes / 4 > MAX_ROWS)
    return fail(c, HAVAC_CUDA_E_LENGTH, "more than 2^24 model positions: the hit word holds the row in 24 bits");
  if (c->launched && !c->finished) return fail(c, HAVAC_CUDA_E_LOGIC, "a run is in flight");
  c->nRows = (uint32_t)(nBytes / 4);
  for (DeviceCtx& d : c->devs) {
    if (c->nRows == 0) { free_table(d); continue; }
    CK(cudaSetDevice(d.dev));
    if (nBytes != d.tabCap) {
      free_table(d);
      d.rowsPadded = (c->nRows + SSV_ROW_CHUNK - 1) / SSV_ROW_CHUNK * SSV_ROW_CHUNK;
      CK(cudaMalloc(&d.dTab8, nBytes));
      CK(cudaMalloc(&d.dTab16, (size_t)d.rowsPadded * sizeof(uint2)));
      d.tabCap = nBytes;
    }
    CK(cudaMemcpyAsync(d.dTab8, flat, nBytes, cudaMemcpyHostToDevice, d.stream));
    c->h2dBytes += nBytes;
    ssv_launch_table_prep(d.dTab8, c->nRows, d.dTab16, d.rowsPadded, d.stream);
    CK(cudaGetLastError());
  }
  for (DeviceCtx& d : c->devs) {
    CK(cudaSetDevice(d.dev));
    CK(cudaStreamSynchronize(d.stream));
  }
  return HAVAC_CUDA_OK;
}

int havac_cuda_project_phmm(havac_cuda_client* c, const float* negLogEmissions, uint64_t nRows, const uint32_t* modelLength,
                            const float* scaleFactor, uint32_t nModels) {
  if (!c || (!negLogEmissions && nRows) || (nModels && (!modelLength || !scaleFactor))) return HAVAC_CUDA_E_INVALID;
  const uint64_t nBytes = nRows * 4;
  if (nBytes >= 1024ull * 1024 * 1024)  // host/HavacHwClient.cpp:121-125
    return fail(c, HAVAC_CUDA_E_LENGTH, "model length must be less than 1GiB. length requested: " + std::to_string(nBytes) + " bytes.");
  if (nRows > MAX_ROWS) return fail(c, HAVAC_CUDA_E_LENGTH, "more than 2^24 model positions: the hit word holds the row in 24 bits");
  if (c->launched && !c->finished) return fail(c, HAVAC_CUDA_E_LOGIC, "a run is in flight");
  std::vector<uint32_t> firstRow(nModels);
  std::vector<float> alphaBeta(2 * (size_t)nModels);
  uint64_t at = 0;
  const float LOG2_E = 1.44269504089;  // PhmmReprojection.cpp:113-115
  for (uint32_t m = 0; m < nModels; m++) {
    firstRow[m] = (uint32_t)at;
    at += modelLength[m];
    alphaBeta[2 * m] = 2 * scaleFactor[m];
    alphaBeta[2 * m + 1] = LOG2_E * scaleFactor[m];
  }
  if (at != nRows) return fail(c, HAVAC_CUDA_E_LENGTH, "model lengths do not add up to the number of emission rows");
  c->nRows = (uint32_t)nRows;
  for (DeviceCtx& d : c->devs) {
    if (c->nRows == 0) { free_table(d); continue; }
    CK(cudaSetDevice(d.dev));
    if (nBytes != d.tabCap) {
      free_table(d);
      d.rowsPadded = (c->nRows + SSV_ROW_CHUNK - 1) / SSV_ROW_CHUNK * SSV_ROW_CHUNK;
      CK(cudaMalloc(&d.dTab8, nBytes));
      CK(cudaMalloc(&d.dTab16, (size_t)d.rowsPadded * sizeof(uint2)));
      d.tabCap = nBytes;
    }
    float *dEm = nullptr, *dAb = nullptr;
    uint32_t* dFirst = nullptr;
    CK(cudaMalloc(&dEm, nRows * 16));
    cudaError_t e1 = cudaMalloc(&dAb, alphaBeta.size() * 4), e2 = cudaMalloc(&dFirst, firstRow.size() * 4);
    auto drop = [&] { cudaFree(dEm); cudaFree(dAb); cudaFree(dFirst); };
    if (e1 != cudaSuccess || e2 != cudaSuccess) { drop(); return fail(c, HAVAC_CUDA_E_NOMEM, "cudaMalloc: out of device memory"); }
    cudaError_t e = cudaMemcpyAsync(dEm, negLogEmissions, nRows * 16, cudaMemcpyHostToDevice, d.stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dAb, alphaBeta.data(), alphaBeta.size() * 4, cudaMemcpyHostToDevice, d.stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dFirst, firstRow.data(), firstRow.size() * 4, cudaMemcpyHostToDevice, d.stream);
    if (e == cudaSuccess) {
      ssv_launch_project_table(dEm, c->nRows, dFirst, dAb, nModels, d.dTab8, d.stream);
      ssv_launch_table_prep(d.dTab8, c->nRows, d.dTab16, d.rowsPadded, d.stream);
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(d.stream);
    drop();
    if (e != cudaSuccess) return fail(c, HAVAC_CUDA_E_RUNTIME, std::string("project_phmm: ") + cudaGetErrorString(e));
    c->h2dBytes += nRows * 16 + alphaBeta.size() * 4 + firstRow.size() * 4;
  }
  return HAVAC_CUDA_OK;
}

int havac_cuda_read_phmm(havac_cuda_client* c, int8_t* flatScores, uint64_t capBytes, uint64_t* nBytes) {
  if (!c) return HAVAC_CUDA_E_INVALID;
  const uint64_t bytes = (uint64_t)c->nRows * 4;
  if (nBytes) *nBytes = bytes;
  if (!flatScores || capBytes == 0 || bytes == 0) return HAVAC_CUDA_OK;
  if (capBytes < bytes) return fail(c, HAVAC_CUDA_E_LENGTH, "buffer smaller than the resident score table");
  DeviceCtx& d = c->devs[0];
  CK(cudaSetDevice(d.dev));
  CK(cudaMemcpy(flatScores, d.dTab8, bytes, cudaMemcpyDeviceToHost));
  c->d2hBytes += bytes;
  return HAVAC_CUDA_OK;
}

int havac_cuda_invoke_async(havac_cuda_client* c) {
  if (!c) return HAVAC_CUDA_E_INVALID;
  // host/HavacHwClient.cpp:142-147
  if (c->seqSegments == 0) return fail(c, HAVAC_CUDA_E_LENGTH, "sequence length in segments cannot be 0, but 0 was given to the client.");
  if (c->nRows == 0) return fail(c, HAVAC_CUDA_E_LENGTH, "phmm length in vectors cannot be 0, but 0 was given to the client.");
  if (c->launched && !c->finished) return fail(c, HAVAC_CUDA_E_LOGIC, "a run is already in flight");
  c->regrowths = 0;
  c->lastMs = 0.f;
  c->abortRequested = false;
  int rc = launch_all(c);
  if (rc) return rc;
  c->launched = true;
  c->finished = false;
  c->aborted = false;
  c->finalState = HAVAC_CUDA_STATE_QUEUED;
  return HAVAC_CUDA_OK;
}

int havac_cuda_state(havac_cuda_client* c, int* state) {
  if (!c || !state) return HAVAC_CUDA_E_INVALID;
  if (!c->launched)  // host/HavacHwClient.cpp:163-170
    return fail(c, HAVAC_CUDA_E_LOGIC, "run object was not initialized. run function invokeHavacSsvAsync to initialize this object.");
  if (c->finished) { *state = c->finalState; return HAVAC_CUDA_OK; }
  bool allDone = true, anyStarted = false;
  for (DeviceCtx& d : c->devs) {
    CK(cudaSetDevice(d.dev));
    cudaError_t e = cudaEventQuery(d.evDone);
    if (e == cudaErrorNotReady) allDone = false;
    else if (e != cudaSuccess) { c->err = cudaGetErrorString(e); *state = HAVAC_CUDA_STATE_ERROR; return HAVAC_CUDA_E_RUNTIME; }
    if (cudaEventQuery(d.evStart) == cudaSuccess) anyStarted = true;
  }
  if (allDone) {
    int rc = collect(c, false);
    if (rc) return rc;
    *state = c->finalState;
  } else {
    *state = anyStarted ? HAVAC_CUDA_STATE_RUNNING : HAVAC_CUDA_STATE_QUEUED;
  }
  return HAVAC_CUDA_OK;
}

int havac_cuda_wait(havac_cuda_client* c, uint64_t timeoutMs, int* state) {
  if (!c) return HAVAC_CUDA_E_INVALID;
  if (!c->launched) return fail(c, HAVAC_CUDA_E_LOGIC, "nothing was launched");
  int st = 0;
  if (timeoutMs == 0) {
    int rc = collect(c);  // blocks on the stop events
    if (rc) return rc;
    st = c->finalState;
  } else {
    auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeoutMs);
    while (true) {
      int rc = havac_cuda_state(c, &st);
      if (rc) return rc;
      if (c->finished) break;
      if (std::chrono::steady_clock::now() >= deadline) { st = HAVAC_CUDA_STATE_TIMEOUT; break; }
      std::this_thread::sleep_for(std::chrono::microseconds(200));
    }
  }
  if (state) *state = st;
  return HAVAC_CUDA_OK;
}

int havac_cuda_abort(havac_cuda_client* c, int* state) {
  if (!c) return HAVAC_CUDA_E_INVALID;
  if (!c->launched) return fail(c, HAVAC_CUDA_E_LOGIC, "nothing was launched");
  if (!c->finished) {
    c->abortRequested = true;
    for (DeviceCtx& d : c->devs) {
      CK(cudaSetDevice(d.dev));
      // wait until the run has actually started, or the memset that opens it would clear the flag
      CK(cudaEventSynchronize(d.evStart));
      CK(cudaMemcpyAsync(d.dAbort, c->hostOne, sizeof(int), cudaMemcpyHostToDevice, d.ctlStream));
    }
    int rc = collect(c);
    if (rc) return rc;
  }
  if (state) *state = c->finalState;
  return HAVAC_CUDA_OK;
}

int havac_cuda_num_hits(havac_cuda_client* c, uint64_t* n) {
  if (!c || !n) return HAVAC_CUDA_E_INVALID;
  if (!c->launched) return fail(c, HAVAC_CUDA_E_LOGIC, "no run to read hits from");
  int rc = collect(c);
  if (rc) return rc;
  if (c->finalState == HAVAC_CUDA_STATE_ERROR) return HAVAC_CUDA_E_OVERFLOW;
  uint64_t total = 0;
  for (DeviceCtx& d : c->devs) total += d.hitCount;
  *n = total;
  return HAVAC_CUDA_OK;
}

int havac_cuda_read_hits(havac_cuda_client* c, uint64_t* out, uint64_t cap, uint64_t* n) {
  uint64_t total = 0;
  int rc = havac_cuda_num_hits(c, &total);
  if (rc) return rc;
  if (n) *n = total;
  if (total == 0) return HAVAC_CUDA_OK;
  if (!out || cap < total) return fail(c, HAVAC_CUDA_E_LENGTH, "hit buffer too small: " + std::to_string(total) + " hits");
  // each device sorts its own hits into the FPGA's emission order (hit_postprocess.cu); devices hold
  // increasing segment ranges, so laying the lists end to end in device order IS the global order
  uint64_t off = 0;
  for (DeviceCtx& d : c->devs) {
    if (!d.hitCount) continue;
    rc = havac_sort_hits_on_device(c, d);
    if (rc) return rc;
    CK(cudaSetDevice(d.dev));
    CK(cudaMemcpy(out + off, d.dHits, d.hitCount * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    c->d2hBytes += d.hitCount * sizeof(uint64_t);
    off += d.hitCount;
  }
  return HAVAC_CUDA_OK;
}

const char* havac_cuda_last_error(havac_cuda_client* c) { return c ? c->err.c_str() : g_createError.c_str(); }

int havac_cuda_device_time_ms(havac_cuda_client* c, float* ms) {
  if (!c || !ms) return HAVAC_CUDA_E_INVALID;
  if (!c->launched) return fail(c, HAVAC_CUDA_E_LOGIC, "nothing was launched");
  int rc = collect(c);
  if (rc) return rc;
  *ms = c->lastMs;
  return HAVAC_CUDA_OK;
}

int havac_cuda_device_count(havac_cuda_client* c, int* n) {
  if (!c || !n) return HAVAC_CUDA_E_INVALID;
  *n = (int)c->devs.size();
  return HAVAC_CUDA_OK;
}

int havac_cuda_run_stats(havac_cuda_client* c, uint64_t* cells, uint32_t* launches) {
  if (!c) return HAVAC_CUDA_E_INVALID;
  uint64_t cols = 0;
  for (DeviceCtx& d : c->devs) cols += (uint64_t)(d.colEnd - d.colBegin);
  if (cells) *cells = cols * c->nRows;
  if (launches) *launches = c->launches;
  return HAVAC_CUDA_OK;
}

int havac_cuda_sm_cycles(havac_cuda_client* c, uint64_t* cycles) {
  if (!c || !cycles) return HAVAC_CUDA_E_INVALID;
  if (!c->launched) return fail(c, HAVAC_CUDA_E_LOGIC, "nothing was launched");
  int rc = collect(c);
  if (rc) return rc;
  uint64_t worst = 0;
  for (DeviceCtx& d : c->devs) worst = std::max<uint64_t>(worst, d.cycles);
  *cycles = worst;
  return HAVAC_CUDA_OK;
}

int havac_cuda_transfer_stats(havac_cuda_client* c, uint64_t* h2d, uint64_t* d2h) {
  if (!c) return HAVAC_CUDA_E_INVALID;
  if (h2d) *h2d = c->h2dBytes;
  if (d2h) *d2h = c->d2hBytes;
  return HAVAC_CUDA_OK;
}

int havac_cuda_trace_cells(havac_cuda_client* c, uint32_t rowBegin, uint32_t rowEnd, uint64_t colBegin, uint64_t colEnd,
                           havac_cuda_cell_record* out, uint64_t capRecords, uint64_t* nRecords) {
  static_assert(sizeof(havac_cuda_cell_record) == sizeof(SsvCellRecord), "record layouts must agree");
  if (!c) return HAVAC_CUDA_E_INVALID;
  if (c->devs.size() != 1) return fail(c, HAVAC_CUDA_E_LOGIC, "the per-cell trace runs on a single-device client");
  if (c->launched && !c->finished) return fail(c, HAVAC_CUDA_E_LOGIC, "a run is in flight");
  DeviceCtx& d = c->devs[0];
  if (!d.dSeq || !d.dTab8) return fail(c, HAVAC_CUDA_E_LOGIC, "write the sequence and the score table first");
  if (rowEnd > c->nRows || rowBegin >= rowEnd || colBegin >= colEnd || (long long)colEnd > d.resEnd || (long long)colBegin < d.resBegin)
    return fail(c, HAVAC_CUDA_E_LENGTH, "trace window outside the resident matrix");
  const uint64_t n = (uint64_t)(rowEnd - rowBegin) * (colEnd - colBegin);
  if (nRecords) *nRecords = n;
  if (!out || capRecords < n) return fail(c, HAVAC_CUDA_E_LENGTH, "record buffer too small: " + std::to_string(n) + " cells");
  // a diagonal of the window starts rowEnd - 1 columns to its left at most; those symbols must be resident
  if ((long long)colBegin - (long long)(rowEnd - 1) < d.resBegin && d.resBegin > 0)
    return fail(c, HAVAC_CUDA_E_LENGTH, "trace window needs symbols left of the resident range");
  CK(cudaSetDevice(d.dev));
  SsvParams p{};
  p.seqPacked = d.dSeq;
  p.nLoad = d.resEnd - d.resBegin;
  p.loadBegin = d.resBegin;
  p.tab8 = d.dTab8;
  p.nRows = c->nRows;
  SsvCellRecord* dOut = nullptr;
  CK(cudaMalloc(&dOut, n * sizeof(SsvCellRecord)));
  ssv_launch_trace(p, rowBegin, rowEnd, (long long)colBegin - d.resBegin, (long long)colEnd - d.resBegin, dOut, d.stream);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, dOut, n * sizeof(SsvCellRecord), cudaMemcpyDeviceToHost, d.stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(d.stream);
  cudaFree(dOut);
  if (e != cudaSuccess) return fail(c, HAVAC_CUDA_E_RUNTIME, std::string("trace_cells: ") + cudaGetErrorString(e));
  c->d2hBytes += n * sizeof(SsvCellRecord);
  return HAVAC_CUDA_OK;
}

int havac_cuda_debug_poke_phmm(havac_cuda_client* c, uint64_t byteOffset, int8_t value) {
  if (!c) return HAVAC_CUDA_E_INVALID;
  if (c->launched && !c->finished) return fail(c, HAVAC_CUDA_E_LOGIC, "a run is in flight");
  if (byteOffset >= (uint64_t)c->nRows * 4) return fail(c, HAVAC_CUDA_E_LENGTH, "offset beyond the resident score table");
  for (DeviceCtx& d : c->devs) {
    CK(cudaSetDevice(d.dev));
    CK(cudaMemcpyAsync(d.dTab8 + byteOffset, &value, 1, cudaMemcpyHostToDevice, d.stream));
    ssv_launch_table_prep(d.dTab8, c->nRows, d.dTab16, d.rowsPadded, d.stream);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(d.stream));
  }
  return HAVAC_CUDA_OK;
}

int havac_cuda_cell_selftest(havac_cuda_client* c, uint8_t* score, uint8_t* hit) {
  if (!c || !score || !hit) return HAVAC_CUDA_E_INVALID;
  DeviceCtx& d = c->devs[0];
  CK(cudaSetDevice(d.dev));
  uint8_t* buf = nullptr;
  CK(cudaMalloc(&buf, 4 * 65536));
  ssv_launch_cell_selftest(buf, buf + 2 * 65536, d.stream);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(score, buf, 2 * 65536, cudaMemcpyDeviceToHost, d.stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(hit, buf + 2 * 65536, 2 * 65536, cudaMemcpyDeviceToHost, d.stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(d.stream);
  cudaFree(buf);
  if (e != cudaSuccess) return fail(c, HAVAC_CUDA_E_RUNTIME, std::string("cell_selftest: ") + cudaGetErrorString(e));
  return HAVAC_CUDA_OK;
}

int havac_cuda_tile_geometry(havac_cuda_client* c, uint32_t device, int64_t* loadBegin, int64_t* diagBegin, uint32_t* nTiles,
                             uint32_t* tileDiagonals, uint32_t* rowChunk, uint32_t* blockRows) {
  if (!c || device >= c->devs.size()) return HAVAC_CUDA_E_INVALID;
  if (!c->launched) return fail(c, HAVAC_CUDA_E_LOGIC, "nothing was launched");
  const DeviceCtx& d = c->devs[device];
  if (loadBegin) *loadBegin = d.lastLoadBegin;
  if (diagBegin) *diagBegin = d.lastDiagBegin;
  if (nTiles) *nTiles = d.lastTiles;
  if (tileDiagonals) *tileDiagonals = d.lastTileDiags;
  if (rowChunk) *rowChunk = SSV_ROW_CHUNK;
  if (blockRows) *blockRows = SSV_BLOCK_ROWS;
  return HAVAC_CUDA_OK;
}

int havac_cuda_set_kernel(havac_cuda_client* c, const char* name) {
  if (!c || !name) return HAVAC_CUDA_E_INVALID;
  std::string s = name;
  if (s == "auto") c->kernel = KERNEL_AUTO;
  else if (s == "f16x2") c->kernel = KERNEL_F16X2;
  else if (s == "scalar") c->kernel = KERNEL_SCALAR;
  else if (s == "f16x2-streaming") c->kernel = KERNEL_F16X2_STREAMING;
  else return fail(c, HAVAC_CUDA_E_INVALID, "unknown kernel '" + s + "'");
  return HAVAC_CUDA_OK;
}

}  // extern "C"
