// Havac.hpp -- the public driver API, unchanged from the reference (host/Havac.hpp:16-107):
// same class, same method names and argument meaning, same exceptions. Behind it the XRT/U50
// path is replaced by the CUDA device client (HavacHwClient -> include/havac_cuda.h).
#ifndef HAVAC_HOST_HPP
#define HAVAC_HOST_HPP

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "FastaSequences.hpp"
#include "HavacHwClient.hpp"
#include "PhmmModels.hpp"

using std::shared_ptr;
using std::vector;

enum havac_cmd_state {
  HAVAC_CMD_STATE_NEW = 1,
  HAVAC_CMD_STATE_QUEUED = 2,
  HAVAC_CMD_STATE_RUNNING = 3,
  HAVAC_CMD_STATE_COMPLETED = 4,
  HAVAC_CMD_STATE_ERROR = 5,
  HAVAC_CMD_STATE_ABORT = 6,
  HAVAC_CMD_STATE_SUBMITTED = 7,
  HAVAC_CMD_STATEIMEOUT = 8,  // spelled as in the reference (host/Havac.hpp:24)
  HAVAC_CMD_STATE_NORESPONSE = 9
};

// The reference's default names an FPGA bitstream. The string is kept as the third constructor
// argument and now carries the device spec of include/havac_cuda.h: a path like this one (or
// "") means "the single GPU deviceIndex"; "cuda:all" / "cuda:0,1,2,3" shard the sequence.
const std::string xclbinSrcDefault = "../device/bin/havac.xclbin";

class HavacHit {
public:
  HavacHit(const uint64_t sequencePosition, const uint32_t sequenceIndex, const uint32_t phmmPosition, const uint32_t phmmIndex);
  HavacHit() {}  // an addition: leaves the fields unset, so that a vector of hits can be sized first and filled by several threads
  uint64_t sequencePosition;
  uint32_t sequenceIndex;
  uint32_t phmmPosition;
  uint32_t phmmIndex;
  std::string toString();
};

class Havac {
public:
  Havac(const uint32_t deviceIndex = 0, const float requiredPValue = 0.02f, const std::string xclbinSrc = xclbinSrcDefault);
  Havac(Havac&& havac) = delete;
  Havac(Havac& havac) = delete;
  ~Havac();
  void loadSequence(const std::string fastaSrc);
  void loadPhmm(const std::string phmmSrc);
  void runHardwareClient();
  void runHardwareClientAsync();
  void waitHardwareClientAsync();
  void abortHardwareClient();
  vector<HavacHit> getHitsFromFinishedRun();
  enum havac_cmd_state currentHardwareState();

  // The other strand (no reference counterpart: the reference searches the strand in the file and
  // leaves the reverse complement to a second run on a reversed file). true: the sequence resident
  // on the device is replaced, there, by its reverse complement; the next run searches the Crick
  // strand and getHitsFromFinishedRun reports each hit at the FORWARD-strand coordinate of the
  // symbol that crossed the threshold (record index and position as usual, model positions as
  // usual). false: back to the strand in the file. loadSequence starts on the file's strand.
  void useReverseComplement(const bool crick);
  bool usingReverseComplement() const { return crickStrand; }

  // additions for harnesses (no reference counterpart)
  float lastDeviceTimeMs();
  std::vector<uint8_t> devicePackedSequence() { return hwClient->readSequence(); }  // what the device holds, copied back
  std::vector<int8_t> devicePhmmScores() { return hwClient->readPhmm(); }  // the int8 score table as the device holds it
  const FastaSequences& sequences() const { return fasta; }
  const PhmmModels& phmms() const { return models; }

private:
  shared_ptr<HavacHwClient> hwClient;
  FastaSequences fasta;
  PhmmModels models;
  shared_ptr<vector<int8_t>> compressedPhmmScores;
  uint32_t deviceIndex;
  float requiredPValue;
  bool phmmLoadedToDevice = false;
  bool sequenceLoadedToDevice = false;
  bool crickStrand = false;
  std::string havacXclbinFileSrc;
  const std::string havacKernelName = "HavacKernel";
};

// Raw device hits -> HavacHit, exactly as Havac::getHitsFromFinishedRun does it
// (host/Havac.cpp:145-187): decode cell[13:0] | segment[39:14] | row[63:40], resolve the record
// with the inclusive end test, drop hits in the padding, resolve the model by prefix sums.
vector<HavacHit> havacResolveHits(const vector<uint64_t>& rawHits, const FastaSequences& fasta,
                                  const vector<uint32_t>& phmmLengthPrefixSums);

#endif
