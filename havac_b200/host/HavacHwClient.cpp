// HavacHwClient.cpp -- see HavacHwClient.hpp. Error mapping follows the reference:
//   length contract broken -> std::length_error   (HavacHwClient.cpp:83-96,112-125,142-147)
//   call out of order      -> std::logic_error    (HavacHwClient.cpp:168)
//   device failure         -> std::runtime_error  (HavacHwClient.cpp:179-183)
//   allocation failure     -> std::bad_alloc
#include "HavacHwClient.hpp"

#include <utility>
#include <initializer_list>
#include <thread>
#include <algorithm>
#include <cstdio>
#include <new>
#include <stdexcept>

HavacHwClient::HavacHwClient(const std::string& xclbinFileSrc, const std::string& havacKernelName, const uint32_t deviceIndex)
    : kernelName(havacKernelName) {
  int rc = havac_cuda_create(xclbinFileSrc.c_str(), deviceIndex, &handle);
  if (rc != HAVAC_CUDA_OK) {
    std::string msg = havac_cuda_last_error(nullptr);
    handle = nullptr;
    if (rc == HAVAC_CUDA_E_NOMEM) throw std::bad_alloc();
    throw std::runtime_error("could not open the CUDA device client: " + msg);
  }
}

HavacHwClient::~HavacHwClient() {
  if (handle) havac_cuda_destroy(handle);
}

void HavacHwClient::raise(int code, const char* where) {
  std::string msg = std::string(where) + ": " + havac_cuda_last_error(handle);
  switch (code) {
    case HAVAC_CUDA_E_LENGTH: throw std::length_error(msg);
    case HAVAC_CUDA_E_LOGIC: throw std::logic_error(msg);
    case HAVAC_CUDA_E_NOMEM: throw std::bad_alloc();
    case HAVAC_CUDA_E_INVALID: throw std::invalid_argument(msg);
    default: throw std::runtime_error(msg);  // E_RUNTIME, E_NODEVICE, E_OVERFLOW
  }
}

void HavacHwClient::writeSequence(const std::vector<uint8_t>& compressedSequence) {
  check(havac_cuda_write_sequence(handle, compressedSequence.data(), compressedSequence.size()), "writeSequence");
}

void HavacHwClient::writePhmm(std::shared_ptr<std::vector<int8_t>> phmmAsFlattenedArray) {
  if (!phmmAsFlattenedArray) throw std::invalid_argument("writePhmm: null score table");
  check(havac_cuda_write_phmm(handle, phmmAsFlattenedArray->data(), phmmAsFlattenedArray->size()), "writePhmm");
}

void HavacHwClient::invokeHavacSsvAsync() { check(havac_cuda_invoke_async(handle), "invokeHavacSsvAsync"); }

havac_hw_cmd_state HavacHwClient::getHwState() {
  int st = 0;
  check(havac_cuda_state(handle, &st), "getHwState");
  return (havac_hw_cmd_state)st;
}

havac_hw_cmd_state HavacHwClient::waitForHavacSsvAsync(const std::chrono::milliseconds& timeout) {
  int st = 0;
  check(havac_cuda_wait(handle, (uint64_t)timeout.count(), &st), "waitForHavacSsvAsync");
  return (havac_hw_cmd_state)st;
}

havac_hw_cmd_state HavacHwClient::abort() {
  int st = 0;
  check(havac_cuda_abort(handle, &st), "abort");
  return (havac_hw_cmd_state)st;
}

uint32_t HavacHwClient::getNumHits() {
  uint64_t n = 0;
  check(havac_cuda_num_hits(handle, &n), "getNumHits");
  return (uint32_t)n;
}

std::vector<uint64_t> HavacHwClient::getHitList() {
  uint64_t n = 0;
  check(havac_cuda_num_hits(handle, &n), "getHitList");
  std::printf("client reported %llu hits\n", (unsigned long long)n);  // HavacHwClient.cpp:189
  std::vector<uint64_t> hits(n);
  uint64_t got = 0;
  check(havac_cuda_read_hits(handle, hits.data(), hits.size(), &got), "getHitList");
  hits.resize(got);
  return hits;
}

void HavacHwClient::ingestFasta(const uint8_t* text, uint64_t nBytes, uint8_t (*resolve)(void*, unsigned char), void* user,
                                std::vector<uint64_t>& sequenceEndPositions, std::vector<uint64_t>& headerByteOffsets) {
  uint64_t nSymbols = 0, nRecords = 0;
  check(havac_cuda_ingest_fasta(handle, text, nBytes, resolve, user, &nSymbols, &nRecords), "ingestFasta");
  sequenceEndPositions.assign(nRecords, 0);
  headerByteOffsets.assign(nRecords, 0);
  uint64_t n = 0;
  check(havac_cuda_ingest_records(handle, sequenceEndPositions.data(), headerByteOffsets.data(), nRecords, &n), "ingestFasta");
}

std::vector<uint8_t> HavacHwClient::readSequence() {
  uint64_t n = 0;
  check(havac_cuda_read_sequence(handle, nullptr, 0, &n), "readSequence");
  std::vector<uint8_t> out(n);
  check(havac_cuda_read_sequence(handle, out.data(), out.size(), &n), "readSequence");
  return out;
}

void HavacHwClient::projectPhmm(const std::vector<float>& negLogEmissions, const std::vector<uint32_t>& modelLengths,
                                const std::vector<float>& scaleFactors) {
  if (modelLengths.size() != scaleFactors.size()) throw std::invalid_argument("projectPhmm: one scale factor per model");
  check(havac_cuda_project_phmm(handle, negLogEmissions.data(), negLogEmissions.size() / 4, modelLengths.data(),
                                scaleFactors.data(), (uint32_t)modelLengths.size()),
        "writePhmm");
}

std::vector<int8_t> HavacHwClient::readPhmm() {
  uint64_t n = 0;
  check(havac_cuda_read_phmm(handle, nullptr, 0, &n), "readPhmm");
  std::vector<int8_t> out(n);
  check(havac_cuda_read_phmm(handle, out.data(), out.size(), &n), "readPhmm");
  return out;
}

void HavacHwClient::reverseComplement(uint64_t nSymbols) {
  check(havac_cuda_reverse_complement(handle, nSymbols), "reverseComplement");
}

namespace {
// first write to every page of [p, p + bytes), spread over threads
void touchPages(std::initializer_list<std::pair<void*, size_t>> ranges) {
  const unsigned threads = std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
  for (const auto& r : ranges) {
    char* p = (char*)r.first;
    const size_t bytes = r.second;
    if (bytes < (64u << 20) || threads == 1) continue;  // small arrays: the copy itself touches them
    std::vector<std::thread> pool;
    const size_t slice = (bytes / threads + 4095) & ~(size_t)4095;
    for (unsigned t = 0; t < threads; t++)
      pool.emplace_back([=] {
        const size_t a = (size_t)t * slice, b = std::min(bytes, a + slice);
        for (size_t k = a; k < b; k += 4096) p[k] = 0;
      });
    for (std::thread& th : pool) th.join();
  }
}
}  // namespace

HavacHwClient::ResolvedHits HavacHwClient::resolveHits(const std::vector<uint64_t>& sequenceEndPositions,
                                                       const std::vector<uint32_t>& modelLengths) {
  uint64_t n = 0;
  check(havac_cuda_num_hits(handle, &n), "getHitList");
  std::printf("client reported %llu hits\n", (unsigned long long)n);  // HavacHwClient.cpp:189
  ResolvedHits out;
  if (n == 0) return out;
  out.sequencePosition.reset(new uint64_t[n]);
  out.sequenceIndex.reset(new uint32_t[n]);
  out.phmmPosition.reset(new uint32_t[n]);
  out.phmmIndex.reset(new uint32_t[n]);
  touchPages({{out.sequencePosition.get(), n * 8}, {out.sequenceIndex.get(), n * 4}, {out.phmmPosition.get(), n * 4},
              {out.phmmIndex.get(), n * 4}});
  check(havac_cuda_resolve_hits(handle, sequenceEndPositions.data(), sequenceEndPositions.size(), modelLengths.data(),
                                (uint32_t)modelLengths.size(), out.sequencePosition.get(), out.sequenceIndex.get(),
                                out.phmmPosition.get(), out.phmmIndex.get(), n, &out.count),
        "getHitsFromFinishedRun");
  return out;
}

void HavacHwClient::setSegmentRange(uint64_t segBegin, uint64_t segEnd) {
  check(havac_cuda_set_segment_range(handle, segBegin, segEnd), "setSegmentRange");
}

float HavacHwClient::deviceTimeMs() {
  float ms = 0.f;
  check(havac_cuda_device_time_ms(handle, &ms), "deviceTimeMs");
  return ms;
}

int HavacHwClient::deviceCount() {
  int n = 0;
  check(havac_cuda_device_count(handle, &n), "deviceCount");
  return n;
}
