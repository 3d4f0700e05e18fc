/*
 * oracle/ref_shim.cpp -- TEST INFRASTRUCTURE ONLY.
 *
 * extern "C" doorways into the UNMODIFIED reference objects that oracle/Makefile
 * compiles straight from /root/reference into oracle/_ref/libhavac_ref.so.
 * No reference source is copied here; this file only calls it. Used to pin
 * oracle/*.c, to generate tests/golden/, and as bench.py's "reference" CPU arm.
 */
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <vector>
#include <memory>

extern "C" {
#include "FastaVector.h"
#include "p7HmmReader.h"
}
#include "PhmmReprojection.h"
#include "SoftSsv.h"
#include "PhmmPreprocessor.hpp"
#include "SequencePreprocessor.hpp"

extern "C" {

/* test/softSsv/SoftSsv.cpp:15 */
int64_t ref_soft_ssv(const uint8_t* symbols, uint64_t nSymbols, const int8_t* table, uint64_t nRows,
                     uint64_t* hitSeqPos, uint32_t* hitRow, uint64_t cap) {
  int8_t err = 0;
  std::vector<SoftSsvHit> hits = softSsvThreshold256(symbols, nSymbols, table, nRows, err);
  if (err) return -1;
  for (uint64_t i = 0; i < hits.size() && i < cap; i++) {
    hitSeqPos[i] = hits[i].sequencePosition;
    hitRow[i] = hits[i].phmmPosition;
  }
  return (int64_t)hits.size();
}

/* PhmmReprojection/PhmmReprojection.cpp:36 */
float ref_scale_factor(float mu, float lambda, uint32_t maxLength, uint32_t modelLength, float p) {
  P7Hmm h;
  std::memset(&h, 0, sizeof h);
  h.stats.msvGumbelMu = mu;
  h.stats.msvGumbelLambda = lambda;
  h.header.maxLength = maxLength;
  h.header.modelLength = modelLength;
  return findThreshold256ScalingFactor(&h, p);
}

/* readP7Hmm + host/phmm/PhmmPreprocessor.cpp:9. Returns total rows, or <0.
 * modelLen/out may be NULL to query sizes. */
int64_t ref_load_phmm(const char* path, float p, int8_t* out, uint64_t outCap, uint32_t* modelLen,
                      uint32_t modelCap, uint32_t* nModels) {
  P7HmmList list;
  if (readP7Hmm(path, &list) != p7HmmSuccess) return -2;
  PhmmPreprocessor pre(&list, p);
  std::shared_ptr<std::vector<int8_t>> data = pre.getProcessedPhmmData();
  if (nModels) *nModels = list.count;
  for (uint32_t i = 0; i < list.count && modelLen && i < modelCap; i++)
    modelLen[i] = list.phmms[i].header.modelLength;
  if (out && outCap >= data->size()) std::memcpy(out, data->data(), data->size());
  int64_t rows = (int64_t)(data->size() / 4);
  p7HmmListDealloc(&list);
  return rows;
}

/* fastaVectorReadFasta + host/sequence/SequencePreprocessor.cpp:9. Returns the
 * packed size in bytes; fills record end offsets and FastaVector sequence.count. */
int64_t ref_load_fasta(const char* path, unsigned seed, uint8_t* packed, uint64_t packedCap,
                       uint64_t* seqEnd, uint64_t seqCap, uint64_t* nSeqs, uint64_t* count) {
  FastaVector fv;
  if (fastaVectorInit(&fv) != FASTA_VECTOR_OK) return -1;
  if (fastaVectorReadFasta(path, &fv) != FASTA_VECTOR_OK) { fastaVectorDealloc(&fv); return -2; }
  srand(seed);
  SequencePreprocessor pre(&fv);
  std::vector<uint8_t>& buf = pre.getCompressedSequenceBuffer();
  if (nSeqs) *nSeqs = fv.metadata.count;
  if (count) *count = fv.sequence.count;
  for (uint64_t i = 0; i < fv.metadata.count && seqEnd && i < seqCap; i++)
    seqEnd[i] = fv.metadata.data[i].sequenceEndPosition;
  if (packed && packedCap >= buf.size()) std::memcpy(packed, buf.data(), buf.size());
  int64_t n = (int64_t)buf.size();
  fastaVectorDealloc(&fv);
  return n;
}

/* FastaVector/src/FastaVector.c:345 on a synthetic metadata table. */
int ref_local_position(const uint64_t* seqEnd, uint64_t nSeqs, uint64_t globalPos, uint64_t* seqIdx,
                       uint64_t* pos) {
  FastaVector fv;
  std::memset(&fv, 0, sizeof fv);
  std::vector<FastaVectorMetadata> md(nSeqs);
  for (uint64_t i = 0; i < nSeqs; i++) { md[i].headerEndPosition = 0; md[i].sequenceEndPosition = seqEnd[i]; }
  fv.metadata.data = md.data();
  fv.metadata.count = nSeqs;
  fv.metadata.capacity = nSeqs;
  FastaVectorLocalPosition lp;
  bool ok = fastaVectorGetLocalSequencePositionFromGlobal(&fv, globalPos, &lp);
  if (ok) { *seqIdx = lp.sequenceIndex; *pos = lp.positionInSequence; }
  return ok ? 1 : 0;
}

}  // extern "C"
