// havac_host_c.cpp -- extern "C" doors declared in include/havac_host.h. Thin: one Havac member
// per function, exceptions mapped onto codes.
#include "havac_host.h"

#include <cstdlib>
#include <cstring>
#include <new>
#include <stdexcept>
#include <string>

#include "Havac.hpp"

struct havac_host {
  Havac* havac = nullptr;
  std::string err;
  vector<HavacHit> hits;
  bool hitsFetched = false;
};

namespace {
thread_local std::string g_err;

template <class F>
int guarded(havac_host* h, F&& body) {
  std::string& err = h ? h->err : g_err;
  try {
    body();
    return 0;
  } catch (const std::length_error& e) { err = e.what(); return -2;
  } catch (const std::invalid_argument& e) { err = e.what(); return -1;
  } catch (const std::logic_error& e) { err = e.what(); return -3;
  } catch (const std::bad_alloc& e) { err = e.what(); return -5;
  } catch (const std::exception& e) { err = e.what(); return -4; }
}
}  // namespace

extern "C" {

int havac_host_create(uint32_t deviceIndex, float requiredPValue, const char* src, havac_host** out) {
  if (!out) return -1;
  *out = nullptr;
  havac_host* h = new (std::nothrow) havac_host();
  if (!h) return -5;
  int rc = guarded(nullptr, [&] { h->havac = new Havac(deviceIndex, requiredPValue, src ? std::string(src) : xclbinSrcDefault); });
  if (rc) { delete h; return rc; }
  *out = h;
  return 0;
}

void havac_host_destroy(havac_host* h) {
  if (!h) return;
  delete h->havac;
  delete h;
}

int havac_host_load_sequence(havac_host* h, const char* fastaSrc) {
  if (!h || !fastaSrc) return -1;
  return guarded(h, [&] { h->havac->loadSequence(fastaSrc); });
}

void havac_host_srand(unsigned seed) { srand(seed); }

int havac_host_sequence_table(havac_host* h, uint64_t* sequenceEnd, uint64_t cap, uint64_t* n) {
  if (!h || !n) return -1;
  return guarded(h, [&] {
    const std::vector<uint64_t>& e = h->havac->sequences().sequenceEndPositions();
    *n = e.size();
    if (sequenceEnd && cap >= e.size()) std::copy(e.begin(), e.end(), sequenceEnd);
  });
}

int havac_host_read_sequence(havac_host* h, uint8_t* packed, uint64_t cap, uint64_t* n) {
  if (!h || !n) return -1;
  return guarded(h, [&] {
    std::vector<uint8_t> buf = h->havac->devicePackedSequence();
    *n = buf.size();
    if (packed && cap >= buf.size()) std::copy(buf.begin(), buf.end(), packed);
  });
}

int havac_host_read_phmm(havac_host* h, int8_t* flatScores, uint64_t cap, uint64_t* n) {
  if (!h || !n) return -1;
  return guarded(h, [&] {
    std::vector<int8_t> buf = h->havac->devicePhmmScores();
    *n = buf.size();
    if (flatScores && cap >= buf.size()) std::copy(buf.begin(), buf.end(), flatScores);
  });
}

int havac_host_use_reverse_complement(havac_host* h, int crick) {
  if (!h) return -1;
  return guarded(h, [&] { h->havac->useReverseComplement(crick != 0); });
}

int havac_host_reverse_complement_packed(uint8_t* packed, uint64_t nBytes, uint64_t nSymbols) {
  if (!packed && nBytes) return -1;
  return guarded(nullptr, [&] {
    std::vector<uint8_t> buf(packed, packed + nBytes);
    havacReverseComplementPacked(buf, nSymbols);
    std::copy(buf.begin(), buf.end(), packed);
  });
}

int havac_host_load_phmm(havac_host* h, const char* phmmSrc) {
  if (!h || !phmmSrc) return -1;
  return guarded(h, [&] { h->havac->loadPhmm(phmmSrc); });
}

int havac_host_run(havac_host* h) {
  if (!h) return -1;
  h->hitsFetched = false;
  return guarded(h, [&] { h->havac->runHardwareClient(); });
}

int havac_host_run_async(havac_host* h) {
  if (!h) return -1;
  h->hitsFetched = false;
  return guarded(h, [&] { h->havac->runHardwareClientAsync(); });
}

int havac_host_wait(havac_host* h) {
  if (!h) return -1;
  return guarded(h, [&] { h->havac->waitHardwareClientAsync(); });
}

int havac_host_abort(havac_host* h) {
  if (!h) return -1;
  return guarded(h, [&] { h->havac->abortHardwareClient(); });
}

int havac_host_state(havac_host* h, int* state) {
  if (!h || !state) return -1;
  return guarded(h, [&] { *state = (int)h->havac->currentHardwareState(); });
}

int havac_host_get_hits(havac_host* h, uint64_t* sequencePosition, uint32_t* sequenceIndex, uint32_t* phmmPosition,
                        uint32_t* phmmIndex, uint64_t cap, uint64_t* n) {
  if (!h || !n) return -1;
  int rc = 0;
  if (!h->hitsFetched) {
    rc = guarded(h, [&] { h->hits = h->havac->getHitsFromFinishedRun(); });
    if (rc) return rc;
    h->hitsFetched = true;
  }
  *n = h->hits.size();
  if (cap == 0) return 0;
  if (cap < h->hits.size() || !sequencePosition || !sequenceIndex || !phmmPosition || !phmmIndex) {
    h->err = "hit arrays too small";
    return -2;
  }
  for (size_t i = 0; i < h->hits.size(); i++) {
    sequencePosition[i] = h->hits[i].sequencePosition;
    sequenceIndex[i] = h->hits[i].sequenceIndex;
    phmmPosition[i] = h->hits[i].phmmPosition;
    phmmIndex[i] = h->hits[i].phmmIndex;
  }
  return 0;
}

int havac_host_device_time_ms(havac_host* h, float* ms) {
  if (!h || !ms) return -1;
  return guarded(h, [&] { *ms = h->havac->lastDeviceTimeMs(); });
}

const char* havac_host_last_error(havac_host* h) { return h ? h->err.c_str() : g_err.c_str(); }

int havac_host_pack_fasta(const char* fastaSrc, unsigned seed, uint8_t* packed, uint64_t packedCap, uint64_t* packedBytes,
                          uint64_t* sequenceEnd, uint64_t endCap, uint64_t* nSequences) {
  if (!fastaSrc) return -1;
  return guarded(nullptr, [&] {
    srand(seed);
    FastaSequences f;
    f.readFasta(fastaSrc);
    const std::vector<uint8_t>& buf = f.compressedSequenceBuffer();
    if (packedBytes) *packedBytes = buf.size();
    if (nSequences) *nSequences = f.numSequences();
    if (packed && packedCap >= buf.size()) std::memcpy(packed, buf.data(), buf.size());
    if (sequenceEnd)
      for (uint64_t i = 0; i < f.numSequences() && i < endCap; i++) sequenceEnd[i] = f.sequenceEndPositions()[i];
  });
}

int havac_host_project_phmm(const char* phmmSrc, float pValue, int8_t* table, uint64_t tableCap, uint64_t* tableBytes,
                            uint32_t* modelLength, uint32_t lengthCap, uint32_t* nModels) {
  if (!phmmSrc) return -1;
  return guarded(nullptr, [&] {
    PhmmModels m;
    m.readP7Hmm(phmmSrc);
    std::shared_ptr<std::vector<int8_t>> t = m.projectForThreshold256(pValue);
    if (tableBytes) *tableBytes = t->size();
    if (nModels) *nModels = m.count();
    if (table && tableCap >= t->size()) std::memcpy(table, t->data(), t->size());
    if (modelLength)
      for (uint32_t i = 0; i < m.count() && i < lengthCap; i++) modelLength[i] = m.models()[i].modelLength;
  });
}

int havac_host_parse_phmm(const char* phmmSrc, float* matchEmissions, uint64_t cap, uint64_t* nFloats) {
  if (!phmmSrc || !nFloats) return -1;
  return guarded(nullptr, [&] {
    PhmmModels m;
    m.readP7Hmm(phmmSrc);
    std::vector<float> e = m.concatenatedMatchEmissions();
    *nFloats = e.size();
    if (matchEmissions && cap >= e.size()) std::copy(e.begin(), e.end(), matchEmissions);
  });
}

int havac_host_model_names(const char* phmmSrc, char* text, uint64_t cap, uint64_t* nBytes) {
  if (!phmmSrc || !nBytes) return -1;
  return guarded(nullptr, [&] {
    PhmmModels m;
    m.readP7Hmm(phmmSrc);
    std::string all;
    for (const PhmmModel& x : m.models()) all += x.name + "\t" + x.accession + "\n";
    *nBytes = all.size();
    if (text && cap >= all.size()) std::memcpy(text, all.data(), all.size());
  });
}

float havac_host_scale_factor(float mu, float lambda, uint32_t maxLength, uint32_t modelLength, float pValue) {
  return havacThreshold256ScaleFactor(mu, lambda, maxLength, modelLength, pValue);
}

int havac_host_resolve_hits(const uint64_t* raw, uint64_t nRaw, const uint64_t* sequenceEnd, uint64_t nSequences,
                            const uint32_t* modelLength, uint32_t nModels, uint64_t* sequencePosition,
                            uint32_t* sequenceIndex, uint32_t* phmmPosition, uint32_t* phmmIndex, uint64_t* nOut) {
  if ((!raw && nRaw) || !nOut) return -1;
  return guarded(nullptr, [&] {
    FastaSequences f = FastaSequences::fromEndPositions(std::vector<uint64_t>(sequenceEnd, sequenceEnd + nSequences));
    std::vector<uint32_t> sums(1, 0);
    for (uint32_t i = 0; i < nModels; i++) sums.push_back(sums.back() + modelLength[i]);
    vector<HavacHit> hits = havacResolveHits(std::vector<uint64_t>(raw, raw + nRaw), f, sums);
    *nOut = hits.size();
    for (size_t i = 0; i < hits.size(); i++) {
      sequencePosition[i] = hits[i].sequencePosition;
      sequenceIndex[i] = hits[i].sequenceIndex;
      phmmPosition[i] = hits[i].phmmPosition;
      phmmIndex[i] = hits[i].phmmIndex;
    }
  });
}

}  // extern "C"
