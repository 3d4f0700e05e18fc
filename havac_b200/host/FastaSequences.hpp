// FastaSequences.hpp -- FASTA ingest straight into the device wire format.
//
// The reference reads a FASTA into one big char vector with a '\0' after every record
// (FastaVector/src/FastaVector.c:54-168) and then packs that vector two bits per symbol
// (host/sequence/SequencePreprocessor.cpp:9-85). Only two things survive into the search: the
// packed buffer and each record's end offset. This reader produces exactly those, in one pass
// over the file, without materialising the char vector; the record bookkeeping and the symbol
// mapping (including the terminators that become rand()-chosen bases) are kept bit-for-bit.
#ifndef HAVAC_FASTA_SEQUENCES_HPP
#define HAVAC_FASTA_SEQUENCES_HPP

#include <cstdint>
#include <string>
#include <vector>

constexpr uint64_t HAVAC_SEGMENT_SYMBOLS = 12288;  // device/PublicDefines.h:18-22 (16 groups x 768 cells)

struct FastaLocalPosition {
  uint64_t sequenceIndex;
  uint64_t positionInSequence;
};

class FastaSequences {
public:
  // Parses `fastaSrc` and APPENDS its records (the reference's FastaVector appends too when
  // loadSequence is called twice). Throws std::runtime_error when the file cannot be opened or
  // holds no record header before its first sequence line; std::bad_alloc on allocation failure.
  void readFasta(const std::string& fastaSrc);

  // sequenceEndPosition of every record, counting one terminator symbol per record
  // (FastaVector.c:19-24,78-86,159-164).
  const std::vector<uint64_t>& sequenceEndPositions() const { return ends; }
  const std::vector<std::string>& headers() const { return names; }
  uint64_t numSequences() const { return ends.size(); }
  // FastaVector's sequence.count: all symbols of all records plus one terminator each.
  uint64_t symbolCount() const { return ends.empty() ? 0 : ends.back(); }

  // The packed device buffer: symbol k in byte k/4 at bit (k%4)*2, zero-filled up to a whole
  // number of 12288-symbol segments (SequencePreprocessor.cpp:15-19,37-59).
  const std::vector<uint8_t>& compressedSequenceBuffer() const { return packed; }
  uint64_t lengthInSegments() const { return packed.size() * 4 / HAVAC_SEGMENT_SYMBOLS; }

  // fastaVectorGetLocalSequencePositionFromGlobal (FastaVector.c:345-373), including its
  // inclusive upper bound: a global position equal to a record's end offset still resolves to
  // that record. Binary search instead of the reference's linear scan; same answers.
  bool localPositionFromGlobal(uint64_t globalPosition, FastaLocalPosition& out) const;

  // Record table of a FASTA that was parsed and packed on the device (HavacHwClient::ingestFasta):
  // the packed buffer stays in device memory, so compressedSequenceBuffer() is empty afterwards.
  void adoptDeviceIngest(const std::vector<uint64_t>& sequenceEndPositions, std::vector<std::string> recordHeaders);
  bool packedOnDeviceOnly() const { return deviceOnly; }
  // Gives the object its packed buffer back (copied from the device) so that a later readFasta can append.
  void restorePacked(std::vector<uint8_t> devicePacked) { packed = std::move(devicePacked); deviceOnly = false; }

  // A record table without sequence data (for resolving hits of a search whose FASTA was read elsewhere).
  static FastaSequences fromEndPositions(const std::vector<uint64_t>& sequenceEndPositions);

private:
  std::vector<uint64_t> ends;
  std::vector<std::string> names;
  std::vector<uint8_t> packed;
  uint64_t nSymbols = 0;  // symbols written so far (terminators included)
  bool deviceOnly = false;
  void pushSymbol(unsigned char c);
  void finishPadding();
};

// SequencePreprocessor::getCompressedSymbol (SequencePreprocessor.cpp:62-85): A/C/G/T in either
// case map to 0..3; everything else draws from rand() in the reference's order (one draw for the
// two-way ambiguity codes, a second one for anything else). 'Y' reproduces the reference's
// operator-precedence slip (`randMod2 << 1 + 1`), whose two low bits are always 0.
uint8_t havacCompressedSymbol(unsigned char c);

// The packed buffer's reverse complement over its first nSymbols symbols, in place (the padding
// behind them stays): symbol k becomes 3 - symbol(nSymbols-1-k). The host form of
// havac_cuda_reverse_complement, for clients that hold a segment range only.
void havacReverseComplementPacked(std::vector<uint8_t>& packed, uint64_t nSymbols);

#endif
