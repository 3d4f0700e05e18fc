// tests/cpp/reference_comparison_test.cpp -- the shape of the reference's on-card test
// (host/test/RefernceComparisonTest/ReferenceComparisonTest.cpp:52-128) against this repo's
// driver: build a Havac object, load a model file and a FASTA file, run the device, and compare its
// hits two ways with a software SSV on the same inputs (:132-176: every hardware hit must be a
// software hit and every software hit a hardware hit, on the full 4-tuple). "abort" as the second
// argument exercises abortHardwareClient as :66-86 does, on a run that is in flight.
//
// The software side is the oracle (oracle/soft_ssv.c, the restatement of the reference's
// softSsvThreshold256) on the buffers the device holds, resolved by the driver's own
// havacResolveHits -- the reference's test uses HitsFromSsv, which resets the state per
// (model, sequence) pair and is therefore not an exact mirror of the device.
//
//   reference_comparison_test <device spec> <fasta> <hmm> [p-value]
//   reference_comparison_test <device spec> abort <fasta> <hmm>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <memory>
#include <string>
#include <tuple>
#include <vector>

#include "Havac.hpp"

extern "C" {
int64_t oracle_soft_ssv(const uint8_t* symbols, uint64_t n_symbols, const int8_t* table, uint64_t n_rows, uint64_t* hit_pos,
                        uint32_t* hit_row, uint64_t cap);
void oracle_unpack_2bit(const uint8_t* packed, uint64_t n_symbols, uint8_t* symbols);
uint64_t oracle_encode_hit(uint64_t seq_pos, uint32_t row);
}

using Tuple = std::tuple<uint64_t, uint32_t, uint32_t, uint32_t>;  // sequencePosition, sequenceIndex, phmmPosition, phmmIndex

static std::vector<Tuple> asTuples(const std::vector<HavacHit>& hits) {
  std::vector<Tuple> t;
  t.reserve(hits.size());
  for (const HavacHit& h : hits) t.emplace_back(h.sequencePosition, h.sequenceIndex, h.phmmPosition, h.phmmIndex);
  std::sort(t.begin(), t.end());
  return t;
}

int main(int argc, char** argv) {
  if (argc < 4) {
    std::cout << "error: program requires a device spec, a fasta file and a phmm file (or: spec abort fasta phmm)" << std::endl;
    return 2;
  }
  const bool abortRequested = std::string(argv[2]) == "abort";
  if (abortRequested && argc < 5) return 2;
  const std::string src = argv[1], seqFileSrc = argv[abortRequested ? 3 : 2], phmmFileSrc = argv[abortRequested ? 4 : 3];
  const float desiredPValue = (!abortRequested && argc > 4) ? (float)std::atof(argv[4]) : 0.02f;
  srand(1);
  try {
    std::cout << "generating havac object" << std::endl;
    std::shared_ptr<Havac> havac = std::make_shared<Havac>(0, desiredPValue, src);
    std::cout << "loading phmm to device" << std::endl;
    havac->loadPhmm(phmmFileSrc);
    std::cout << "loading sequence to device" << std::endl;
    havac->loadSequence(seqFileSrc);

    if (abortRequested) {
      havac->runHardwareClientAsync();
      std::cout << "abort requested, current hw state: " << havac->currentHardwareState() << "." << std::endl;
      havac->abortHardwareClient();
      const int state = havac->currentHardwareState();
      std::cout << "aborted. hw state is: " << state << "." << std::endl;
      // a run that had already finished reports COMPLETED; one caught in flight reports ABORT
      return (state == HAVAC_CMD_STATE_ABORT || state == HAVAC_CMD_STATE_COMPLETED) ? 0 : 1;
    }

    std::cout << "beginning hardware run..." << std::endl;
    havac->runHardwareClient();
    std::vector<HavacHit> hardwareHits = havac->getHitsFromFinishedRun();
    std::cout << "hardware run generated " << hardwareHits.size() << " hits." << std::endl;

    // the software implementation on the same inputs
    const std::vector<uint8_t> packed = havac->devicePackedSequence();
    const std::vector<int8_t> table = havac->devicePhmmScores();
    std::vector<uint8_t> symbols(packed.size() * 4);
    oracle_unpack_2bit(packed.data(), symbols.size(), symbols.data());
    const uint64_t cap = 1ull << 24;
    std::vector<uint64_t> pos(cap);
    std::vector<uint32_t> row(cap);
    const int64_t n = oracle_soft_ssv(symbols.data(), symbols.size(), table.data(), (uint64_t)(table.size() / 4), pos.data(), row.data(), cap);
    if (n < 0 || (uint64_t)n > cap) {
      std::cout << "error: software ssv failed or overflowed: " << n << std::endl;
      return 3;
    }
    std::vector<uint64_t> raw((size_t)n);
    for (int64_t i = 0; i < n; i++) raw[(size_t)i] = oracle_encode_hit(pos[(size_t)i], row[(size_t)i]);
    const std::vector<HavacHit> softwareHits = havacResolveHits(raw, havac->sequences(), havac->phmms().lengthPrefixSums());
    std::cout << "soft ssv finished, found " << softwareHits.size() << " hits. comparing hits..." << std::endl;

    const std::vector<Tuple> hw = asTuples(hardwareHits), sw = asTuples(softwareHits);
    bool foundAllHwHits = true, foundAllSwHits = true;
    for (const Tuple& h : hw)
      if (!std::binary_search(sw.begin(), sw.end(), h)) {
        foundAllHwHits = false;
        std::cout << "\tERROR: hw hit not found in software hits: " << HavacHit(std::get<0>(h), std::get<1>(h), std::get<2>(h), std::get<3>(h)).toString() << std::endl;
      }
    for (const Tuple& s : sw)
      if (!std::binary_search(hw.begin(), hw.end(), s)) {
        foundAllSwHits = false;
        std::cout << "\terror: sw hit not found in hardware hits: " << HavacHit(std::get<0>(s), std::get<1>(s), std::get<2>(s), std::get<3>(s)).toString() << std::endl;
      }
    if (foundAllHwHits) std::cout << "found all " << hw.size() << " hardware hits" << std::endl;
    if (foundAllSwHits) std::cout << "found all " << sw.size() << " software hits" << std::endl;
    std::cout << "finished comparison" << std::endl;
    return (foundAllHwHits && foundAllSwHits && hw.size() == sw.size()) ? 0 : 1;
  } catch (const std::exception& e) {
    std::cout << "error: " << e.what() << std::endl;
    return 4;
  }
}
