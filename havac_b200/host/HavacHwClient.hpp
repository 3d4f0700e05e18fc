// HavacHwClient.hpp -- the device client of the Havac driver, B200 edition.
//
// Same class name, method names, argument meaning and exception behaviour as the reference's
// XRT client (host/HavacHwClient.hpp:21-115 and HavacHwClient.cpp), but every XRT call is
// replaced by one call into the C-ABI of libhavac_cuda (include/havac_cuda.h). Nothing here
// computes: it only maps error codes back onto the exception types the reference throws.
#ifndef HAVAC_HW_CLIENT_HPP
#define HAVAC_HW_CLIENT_HPP

#include <chrono>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "havac_cuda.h"

// numeric values of XRT's ert_cmd_state, which the reference returns (HavacHwClient.cpp:163-170)
enum havac_hw_cmd_state {
  HAVAC_HW_STATE_NEW = 1,
  HAVAC_HW_STATE_QUEUED = 2,
  HAVAC_HW_STATE_RUNNING = 3,
  HAVAC_HW_STATE_COMPLETED = 4,
  HAVAC_HW_STATE_ERROR = 5,
  HAVAC_HW_STATE_ABORT = 6,
  HAVAC_HW_STATE_SUBMITTED = 7,
  HAVAC_HW_STATE_TIMEOUT = 8,
  HAVAC_HW_STATE_NORESPONSE = 9
};

class HavacHwClient {
public:
  // `xclbinFileSrc` no longer names a bitstream: it is handed to the C-ABI as the device spec
  // ("cuda:all", "cuda:0,1", optional ";hitcap=N"); anything else selects GPU `deviceIndex`.
  HavacHwClient(const std::string& xclbinFileSrc, const std::string& havacKernelName, const uint32_t deviceIndex = 0);
  HavacHwClient(HavacHwClient&& hc) = delete;
  HavacHwClient(HavacHwClient& hc) = delete;
  ~HavacHwClient();

  void writeSequence(const std::vector<uint8_t>& compressedSequence);
  void writePhmm(std::shared_ptr<std::vector<int8_t>> phmmAsFlattenedArray);
  void invokeHavacSsvAsync();
  havac_hw_cmd_state getHwState();
  havac_hw_cmd_state waitForHavacSsvAsync(const std::chrono::milliseconds& timeout = std::chrono::milliseconds{0});
  havac_hw_cmd_state abort();
  std::vector<uint64_t> getHitList();

  // FASTA text -> packed sequence on the device (include/havac_cuda.h: havac_cuda_ingest_fasta);
  // stands for FastaVector + SequencePreprocessor + writeSequence. `resolve` is asked, in sequence
  // order, for every symbol that is not A/C/G/T and for every record terminator (c == 0).
  void ingestFasta(const uint8_t* text, uint64_t nBytes, uint8_t (*resolve)(void*, unsigned char), void* user,
                   std::vector<uint64_t>& sequenceEndPositions, std::vector<uint64_t>& headerByteOffsets);
  std::vector<uint8_t> readSequence();
  // Negative-log match emissions -> int8 score table on the device (havac_cuda_project_phmm);
  // stands for p7HmmProjectForThreshold256's loop + PhmmPreprocessor + writePhmm.
  void projectPhmm(const std::vector<float>& negLogEmissions, const std::vector<uint32_t>& modelLengths,
                   const std::vector<float>& scaleFactors);
  std::vector<int8_t> readPhmm();
  // resident sequence <-> its reverse complement over the first nSymbols symbols (havac_cuda_reverse_complement)
  void reverseComplement(uint64_t nSymbols);
  // Hits of the finished run, sorted and resolved on the device (havac_cuda_resolve_hits): four
  // parallel arrays in emission order, padding hits dropped. The arrays are allocated without
  // being zeroed and their pages are first touched by several threads: at 10^8 hits the page
  // faults of a plain vector resize cost more than the device work.
  struct ResolvedHits {
    std::unique_ptr<uint64_t[]> sequencePosition;
    std::unique_ptr<uint32_t[]> sequenceIndex, phmmPosition, phmmIndex;
    uint64_t count = 0;
  };
  ResolvedHits resolveHits(const std::vector<uint64_t>& sequenceEndPositions, const std::vector<uint32_t>& modelLengths);

  // harness helpers (no reference counterpart)
  void setSegmentRange(uint64_t segBegin, uint64_t segEnd);
  float deviceTimeMs();
  int deviceCount();

protected:
  havac_cuda_client* handle = nullptr;
  std::string kernelName;
  uint32_t getNumHits();
  [[noreturn]] void raise(int code, const char* where);
  void check(int code, const char* where) { if (code != HAVAC_CUDA_OK) raise(code, where); }
};

#endif
