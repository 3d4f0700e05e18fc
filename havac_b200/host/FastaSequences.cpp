#include "FastaSequences.hpp"

#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>

uint8_t havacCompressedSymbol(unsigned char c) {
  switch (c) {
    case 'a': case 'A': return 0;
    case 'c': case 'C': return 1;
    case 'g': case 'G': return 2;
    case 't': case 'T': return 3;
    default: break;
  }
  const unsigned r = (unsigned)(rand() % 2);
  switch (c) {
    case 'r': case 'R': return (uint8_t)(r * 2);  // A or G
    case 'y': case 'Y': return 0;                 // r << 2 in the reference: the symbol's own two bits are 0
    case 's': case 'S': return (uint8_t)(r + 1);  // C or G
    case 'w': case 'W': return (uint8_t)(r * 3);  // A or T
    case 'k': case 'K': return (uint8_t)(r + 2);  // G or T
    case 'm': case 'M': return (uint8_t)r;        // A or C
    default: return (uint8_t)(rand() % 4);
  }
}

void FastaSequences::pushSymbol(unsigned char c) {
  const uint8_t sym = havacCompressedSymbol(c);
  const uint64_t byte = nSymbols >> 2;
  if (byte >= packed.size()) packed.resize(packed.size() < 4096 ? 4096 : packed.size() * 2, 0);
  packed[byte] |= (uint8_t)(sym << ((nSymbols & 3) * 2));
  nSymbols++;
  ends.back()++;
}

void FastaSequences::finishPadding() {
  const uint64_t segments = (nSymbols + HAVAC_SEGMENT_SYMBOLS - 1) / HAVAC_SEGMENT_SYMBOLS;
  packed.resize(segments * HAVAC_SEGMENT_SYMBOLS / 4, 0);
  // a resize that shrinks a doubled buffer must not leave stale bits: everything past the last
  // symbol was never written, so it is already zero ('A').
}

void FastaSequences::readFasta(const std::string& fastaSrc) {
  FILE* f = std::fopen(fastaSrc.c_str(), "rb");
  if (!f) throw std::runtime_error("Could not open fasta file for reading.");
  // continue after previously loaded records: drop the zero padding, keep the symbols
  packed.resize((nSymbols + 3) / 4);

  enum { ON_NEWLINE, HEADER, COMMENT, SEQUENCE } state = ON_NEWLINE;
  std::vector<unsigned char> buf(1 << 22);
  const size_t firstNewRecord = ends.size();
  std::string header;
  bool readFailed = false, orphanSequence = false;
  while (true) {
    size_t n = std::fread(buf.data(), 1, buf.size(), f);
    if (n == 0) {
      readFailed = std::ferror(f) != 0;
      break;
    }
    for (size_t i = 0; i < n && !orphanSequence; i++) {
      const unsigned char c = buf[i];
      // the reference compares a (signed) char with ' ': bytes >= 0x80 are negative there and
      // therefore never "printable" (FastaVector.c:97,144)
      const bool printable = c >= ' ' && c < 0x80;
      switch (state) {
        case COMMENT:
          if (c == '\n') state = ON_NEWLINE;
          break;
        case ON_NEWLINE:
          if (c == '>') {
            state = HEADER;
            // terminate the previous record; like the reference (FastaVector.c:78-86) this also fires
            // for the last record of an earlier file, which therefore ends in two terminators
            if (!ends.empty()) pushSymbol('\0');
            ends.push_back(ends.empty() ? 0 : ends.back());
            header.clear();
          } else if (c == ';') {
            state = COMMENT;
          } else if (printable) {
            if (ends.empty()) { orphanSequence = true; break; }
            state = SEQUENCE;
            pushSymbol(c);
          }
          break;
        case HEADER:
          if (c == ';' || c == '\n') {
            state = c == ';' ? COMMENT : ON_NEWLINE;
            names.push_back(header);
          } else if (printable || c == '\t') {
            header.push_back((char)c);
          }
          break;
        case SEQUENCE:
          if (c == ';') state = COMMENT;
          else if (c == '\n') state = ON_NEWLINE;
          else if (printable) pushSymbol(c);
          break;
      }
    }
    if (orphanSequence) break;
  }
  std::fclose(f);
  if (state == HEADER) names.push_back(header);  // file ended inside a header line
  if (orphanSequence || ends.size() == firstNewRecord) {
    // the reference indexes metadata[-1] here (FastaVector.c:19-24 with count == 0); refuse instead
    ends.resize(firstNewRecord);
    names.resize(firstNewRecord);
    nSymbols = symbolCount();
    packed.resize((nSymbols + 3) / 4);
    finishPadding();
    throw std::runtime_error("Error while reading from the opened fasta file: no '>' record header before sequence data.");
  }
  if (readFailed) throw std::runtime_error("Error while reading from the opened fasta file.");
  pushSymbol('\0');  // terminate the last record (FastaVector.c:159-164)
  while (names.size() < ends.size()) names.emplace_back();
  finishPadding();
}

bool FastaSequences::localPositionFromGlobal(uint64_t globalPosition, FastaLocalPosition& out) const {
  // first record whose end offset is >= the position (the reference's `<=` test, FastaVector.c:361)
  size_t lo = 0, hi = ends.size();
  while (lo < hi) {
    size_t mid = lo + (hi - lo) / 2;
    if (ends[mid] >= globalPosition) hi = mid;
    else lo = mid + 1;
  }
  if (lo == ends.size()) return false;
  out.sequenceIndex = lo;
  out.positionInSequence = globalPosition - (lo == 0 ? 0 : ends[lo - 1]);
  return true;
}

void FastaSequences::adoptDeviceIngest(const std::vector<uint64_t>& sequenceEndPositions, std::vector<std::string> recordHeaders) {
  ends = sequenceEndPositions;
  names = std::move(recordHeaders);
  names.resize(ends.size());
  nSymbols = symbolCount();
  packed.clear();
  deviceOnly = true;
}

FastaSequences FastaSequences::fromEndPositions(const std::vector<uint64_t>& sequenceEndPositions) {
  FastaSequences f;
  f.ends = sequenceEndPositions;
  f.names.resize(sequenceEndPositions.size());
  f.nSymbols = f.symbolCount();
  return f;
}

void havacReverseComplementPacked(std::vector<uint8_t>& packed, uint64_t nSymbols) {
  if (nSymbols > packed.size() * 4) throw std::length_error("reverse complement over more symbols than the buffer holds");
  // 16 symbols (one little-endian 32-bit word) at a time, out of place: word w takes its symbols
  // from the two input words around symbol nSymbols-1-16w, as the device kernel does (strand.cu)
  const uint64_t nWords = (packed.size() + 3) / 4;
  std::vector<uint32_t> in(nWords + 1, 0u), out(nWords, 0u);
  std::memcpy(in.data(), packed.data(), packed.size());
  auto reverseSymbols = [](uint32_t x) -> uint32_t {
    x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
    x = ((x >> 4) & 0x0f0f0f0fu) | ((x & 0x0f0f0f0fu) << 4);
    return __builtin_bswap32(x);
  };
  for (uint64_t w = 0; w < nWords; w++) {
    const uint64_t first = 16 * w;
    if (first >= nSymbols) { out[w] = in[w]; continue; }
    const uint64_t hi = nSymbols - 1 - first, b = hi >> 4;
    const uint64_t window = ((uint64_t)in[b] << 32) | (b ? in[b - 1] : 0u);
    uint32_t r = ~reverseSymbols((uint32_t)(window >> (2 * ((unsigned)(hi & 15) + 1))));
    const uint64_t m = nSymbols - first;
    if (m < 16) {
      const uint32_t mask = (1u << (2 * (unsigned)m)) - 1u;
      r = (r & mask) | (in[w] & ~mask);
    }
    out[w] = r;
  }
  std::memcpy(packed.data(), out.data(), packed.size());
}
