#include "Havac.hpp"

#include <thread>
#include <cstdio>
#include <chrono>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <stdexcept>

Havac::Havac(const uint32_t deviceIndex, const float requiredPValue, const std::string xclbinSrc)
    : deviceIndex(deviceIndex), requiredPValue(requiredPValue), havacXclbinFileSrc(xclbinSrc) {
  // opens the device and allocates the hit buffers (host/Havac.cpp:24, HavacHwClient.cpp:24-31);
  // throws when no sm_100 GPU is present -- there is no host fallback for the search
  this->hwClient = std::make_shared<HavacHwClient>(this->havacXclbinFileSrc, this->havacKernelName, deviceIndex);
}

Havac::~Havac() {}

void Havac::loadPhmm(const std::string phmmSrc) {
  this->models.readP7Hmm(phmmSrc);  // std::runtime_error on a malformed file (host/Havac.cpp:48-50)
  if (std::getenv("HAVAC_HOST_PROJECT")) {  // the reference's route: project on the host, write the int8 table
    this->compressedPhmmScores = this->models.projectForThreshold256(this->requiredPValue);
    this->hwClient->writePhmm(this->compressedPhmmScores);
  } else {  // scale factors here, the per-position arithmetic and the table layout on the device
    this->compressedPhmmScores.reset();
    this->hwClient->projectPhmm(this->models.concatenatedMatchEmissions(), this->models.lengths(),
                                this->models.threshold256ScaleFactors(this->requiredPValue));
  }
  this->phmmLoadedToDevice = true;
}

namespace {

// The whole file, mapped read-only (the text is only read once, by the staging copy to the device).
struct MappedFile {
  const uint8_t* data = nullptr;
  size_t size = 0;
  int fd = -1;
  explicit MappedFile(const std::string& path) {
    fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) throw std::runtime_error("Could not open fasta file for reading.");
    struct stat st;
    if (::fstat(fd, &st) != 0) { ::close(fd); throw std::runtime_error("Could not open fasta file for reading."); }
    size = (size_t)st.st_size;
    if (size) {
      void* p = ::mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
      if (p == MAP_FAILED) { ::close(fd); throw std::runtime_error("Error while reading from the opened fasta file."); }
      ::madvise(p, size, MADV_SEQUENTIAL);
      data = (const uint8_t*)p;
    }
  }
  ~MappedFile() {
    if (data) ::munmap((void*)data, size);
    if (fd >= 0) ::close(fd);
  }
};

uint8_t resolveSymbol(void*, unsigned char c) { return havacCompressedSymbol(c); }

// header text of the record whose '>' sits at `at` (FastaVector.c:88-112: up to ';' or newline,
// keeping printable characters and tabs)
std::string headerAt(const MappedFile& f, uint64_t at) {
  std::string h;
  for (uint64_t i = at + 1; i < f.size; i++) {
    const unsigned char c = f.data[i];
    if (c == ';' || c == '\n') break;
    if ((c >= ' ' && c < 0x80) || c == '\t') h.push_back((char)c);
  }
  return h;
}

}  // namespace

void Havac::loadSequence(const std::string fastaSrc) {
  // First load: the file is parsed and packed on the GPU. A second load appends
  // to the records already held, as the reference's FastaVector does; that rare case (the packed
  // buffer is first copied back from the device) and HAVAC_HOST_FASTA=1 take the host reader, which
  // produces the same bytes.
  if (this->crickStrand) this->useReverseComplement(false);  // an append goes onto the file's strand
  const char* forceHost = std::getenv("HAVAC_HOST_FASTA");
  if (this->fasta.numSequences() == 0 && !(forceHost && forceHost[0] == '1')) {
    MappedFile file(fastaSrc);
    std::vector<uint64_t> ends, headerAtByte;
    this->hwClient->ingestFasta(file.data, file.size, resolveSymbol, nullptr, ends, headerAtByte);
    std::vector<std::string> headers;
    headers.reserve(ends.size());
    for (uint64_t at : headerAtByte) headers.push_back(headerAt(file, at));
    this->fasta.adoptDeviceIngest(ends, std::move(headers));
  } else {
    if (this->fasta.packedOnDeviceOnly())  // appending to a device-packed sequence: take the buffer back first
      this->fasta.restorePacked(this->hwClient->readSequence());
    this->fasta.readFasta(fastaSrc);  // runtime_error / bad_alloc as host/Havac.cpp:59-67
    this->hwClient->writeSequence(this->fasta.compressedSequenceBuffer());
  }
  this->sequenceLoadedToDevice = true;
}

void Havac::runHardwareClient() {
  this->runHardwareClientAsync();
  this->waitHardwareClientAsync();
}

void Havac::runHardwareClientAsync() {
  if (!this->phmmLoadedToDevice)
    throw std::logic_error("Phmm was not loaded to device before hardware was requested to run.");
  if (!this->sequenceLoadedToDevice)
    throw std::logic_error("Sequence was not loaded to device before hardware was requested to run.");
  this->hwClient->invokeHavacSsvAsync();
}

void Havac::waitHardwareClientAsync() { this->hwClient->waitForHavacSsvAsync(); }

void Havac::abortHardwareClient() { this->hwClient->abort(); }

enum havac_cmd_state Havac::currentHardwareState() { return (havac_cmd_state)hwClient->getHwState(); }

float Havac::lastDeviceTimeMs() { return hwClient->deviceTimeMs(); }

vector<HavacHit> havacResolveHits(const vector<uint64_t>& rawHits, const FastaSequences& fasta,
                                  const vector<uint32_t>& prefixSums) {
  vector<HavacHit> resolved;
  resolved.reserve(rawHits.size());
  for (uint64_t i = 0; i < rawHits.size(); i++) {
    const uint64_t raw = rawHits[i];
    const uint64_t cell = raw & ((1ULL << 14) - 1);
    const uint64_t segment = (raw & ((1ULL << 40) - 1)) >> 14;
    const uint64_t globalSequencePosition = segment * HAVAC_SEGMENT_SYMBOLS + cell;
    const uint32_t globalPhmmPosition = (uint32_t)(raw >> 40);

    FastaLocalPosition where;
    if (!fasta.localPositionFromGlobal(globalSequencePosition, where)) continue;  // a hit in the zero padding

    // last prefix sum <= the global row (phmmPrefixSumsBinarySearch, host/Havac.cpp:119-142)
    auto it = std::upper_bound(prefixSums.begin(), prefixSums.end(), globalPhmmPosition);
    if (it == prefixSums.begin()) {
      std::cerr << "ERROR: could not resolve phmm position for raw hit report #" << i << "\n" << std::endl;
      continue;
    }
    const uint32_t phmmIndex = (uint32_t)(it - prefixSums.begin()) - 1;
    resolved.push_back(HavacHit(where.positionInSequence, (uint32_t)where.sequenceIndex,
                                globalPhmmPosition - prefixSums[phmmIndex], phmmIndex));
  }
  return resolved;
}

void Havac::useReverseComplement(const bool crick) {
  if (!this->sequenceLoadedToDevice) throw std::logic_error("useReverseComplement: no sequence has been loaded");
  if (crick == this->crickStrand) return;
  this->hwClient->reverseComplement(this->fasta.symbolCount());
  this->crickStrand = crick;
}

vector<HavacHit> Havac::getHitsFromFinishedRun() {
  // sorted into emission order and resolved on the device; havacResolveHits above is
  // the same mapping on the host (HAVAC_HOST_RESOLVE=1 selects it; tests compare the two)
  const char* forceHost = std::getenv("HAVAC_HOST_RESOLVE");
  if (forceHost && forceHost[0] == '1') {
    vector<uint64_t> rawHits = this->hwClient->getHitList();
    if (this->crickStrand) {  // reversed coordinates -> forward-strand hit words; padding hits dropped
      const uint64_t n = this->fasta.symbolCount();
      vector<uint64_t> forward;
      forward.reserve(rawHits.size());
      for (uint64_t h : rawHits) {
        const uint64_t pos = ((h >> 14) & 0x3ffffffull) * HAVAC_SEGMENT_SYMBOLS + (h & 0x3fffull);
        if (pos >= n) continue;
        const uint64_t f = n - 1 - pos;
        forward.push_back((h & ~0xffffffffffull) | ((f / HAVAC_SEGMENT_SYMBOLS) << 14) | (f % HAVAC_SEGMENT_SYMBOLS));
      }
      rawHits.swap(forward);
    }
    return havacResolveHits(rawHits, this->fasta, this->models.lengthPrefixSums());
  }
  vector<uint32_t> lengths;
  lengths.reserve(this->models.count());
  for (const PhmmModel& m : this->models.models()) lengths.push_back(m.modelLength);
  const auto t0 = std::chrono::steady_clock::now();
  const HavacHwClient::ResolvedHits r = this->hwClient->resolveHits(this->fasta.sequenceEndPositions(), lengths);
  const auto t1 = std::chrono::steady_clock::now();
  // the objects are built in place by several threads (HavacHit's empty default constructor leaves
  // the new elements untouched, so the threads also take the page faults of the 24 bytes per hit)
  vector<HavacHit> resolved(r.count);
  auto fill = [&](size_t a, size_t b) {
    for (size_t i = a; i < b; i++) resolved[i] = HavacHit(r.sequencePosition[i], r.sequenceIndex[i], r.phmmPosition[i], r.phmmIndex[i]);
  };
  const unsigned threads = r.count < (1u << 20) ? 1u : std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
  if (threads == 1) {
    fill(0, r.count);
  } else {
    std::vector<std::thread> pool;
    const size_t slice = (r.count + threads - 1) / threads;
    for (unsigned t = 0; t < threads; t++) pool.emplace_back(fill, std::min<size_t>(r.count, t * slice), std::min<size_t>(r.count, (t + 1) * slice));
    for (std::thread& th : pool) th.join();
  }
  if (std::getenv("HAVAC_TRACE"))
    std::fprintf(stderr, "[havac trace] getHitsFromFinishedRun: client %.3f s, %zu HavacHit objects %.3f s\n",
                 std::chrono::duration<double>(t1 - t0).count(), resolved.size(),
                 std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count());
  return resolved;
}

HavacHit::HavacHit(const uint64_t sequencePosition, const uint32_t sequenceIndex, const uint32_t phmmPosition,
                   const uint32_t phmmIndex)
    : sequencePosition(sequencePosition), sequenceIndex(sequenceIndex), phmmPosition(phmmPosition), phmmIndex(phmmIndex) {}

std::string HavacHit::toString() {
  std::stringstream out;  // same text as host/Havac.cpp:202-207
  out << "sequence $" << this->sequenceIndex << ", position " << this->sequencePosition << "; phmm #" << this->phmmIndex
      << " position " << this->phmmPosition;
  return out.str();
}
