// PhmmModels.hpp -- HMMER3 profile file -> int8 score table at threshold 256.
//
// Stands where the reference uses P7HmmReader (readP7Hmm, P7HmmReader/src/p7HmmReader.c:45),
// PhmmReprojection (PhmmReprojection/PhmmReprojection.cpp:15-145) and PhmmPreprocessor
// (host/phmm/PhmmPreprocessor.cpp:9-31). Only the fields the SSV path consumes are kept: LENG,
// MAXL, ALPH, the MSV Gumbel parameters and the match-emission rows. The projection arithmetic
// mirrors the reference's float/double mix operation by operation, because that mix decides
// every score byte; build with -ffp-contract=off.
#ifndef HAVAC_PHMM_MODELS_HPP
#define HAVAC_PHMM_MODELS_HPP

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

struct PhmmModel {
  std::string name;       // NAME
  std::string accession;  // ACC (optional)
  uint32_t modelLength = 0;  // LENG
  uint32_t maxLength = 0;    // MAXL
  uint32_t alphabetCardinality = 0;
  float msvGumbelMu = 0.f, msvGumbelLambda = 0.f;
  bool hasMsvStats = false;
  std::vector<float> matchEmissionScores;  // negative natural logs, [position][symbol]
};

class PhmmModels {
public:
  // Parses every model of `phmmSrc` and APPENDS them. std::runtime_error("Phmm file was not
  // formatted correctly.") on a malformed file (host/Havac.cpp:48-50), or when it cannot be opened.
  void readP7Hmm(const std::string& phmmSrc);

  const std::vector<PhmmModel>& models() const { return list; }
  uint32_t count() const { return (uint32_t)list.size(); }
  // 0, L0, L0+L1, ... (Havac::generatePhmmLenPrefixSums, host/Havac.cpp:104-116)
  std::vector<uint32_t> lengthPrefixSums() const;

  // All models projected for `pValue` and concatenated with no separator rows: 4 bytes per
  // model position, [A,C,G,T] (PhmmPreprocessor.cpp:9-31). Throws std::runtime_error for a
  // model whose alphabet is not 4 symbols wide or that lacks MAXL / STATS LOCAL MSV.
  std::shared_ptr<std::vector<int8_t>> projectForThreshold256(float pValue) const;
  // The inputs of the same projection done on the device (havac_cuda_project_phmm): one scale
  // factor per model (same checks and exceptions as above), the model lengths, and the match
  // emissions of all models concatenated.
  std::vector<float> threshold256ScaleFactors(float pValue) const;
  std::vector<uint32_t> lengths() const;
  std::vector<float> concatenatedMatchEmissions() const;

private:
  std::vector<PhmmModel> list;
};

// esl_gumbel_invsurv as adapted by the reference (PhmmReprojection.cpp:15-31).
double havacGumbelInverseSurvival(double p, double mu, double lambda);
// findThreshold256ScalingFactor (PhmmReprojection.cpp:36-66).
float havacThreshold256ScaleFactor(float mu, float lambda, uint32_t maxLength, uint32_t modelLength, float pValue);
// p7HmmProjectForThreshold256's inner loop (PhmmReprojection.cpp:109-145).
void havacProjectEmissions(const float* matchEmissions, size_t count, float scaleFactor, int8_t* out);

#endif
