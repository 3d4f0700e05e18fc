// PhmmProjection.cpp -- p-value -> per-model scale factor -> int8 scores (threshold == 256).
// Types below are chosen to reproduce PhmmReprojection.cpp:36-66,109-145 bit for bit: the
// reference is C++ with <math.h>, so log(float) and round(float) bind to the float overloads
// while log(1.0 - x) is evaluated in double.
#include <cmath>
#include <stdexcept>

#include "PhmmModels.hpp"

double havacGumbelInverseSurvival(double p, double mu, double lambda) {
  // small p: log(-log(1-p)) ~ log(p) ~ (p^p - 1)/p   (PhmmReprojection.cpp:27-30)
  const double logPart = (p < 5e-9) ? (std::pow(p, p) - 1) / p : std::log(-1. * std::log(1 - p));
  return mu - (logPart / lambda);
}

float havacThreshold256ScaleFactor(float mu, float lambda, uint32_t maxLengthU, uint32_t modelLengthU, float pValue) {
  const double LN2 = 0.69314718055994529;
  const float maxLength = maxLengthU;
  const float modelLength = modelLengthU;
  const double fullModelScore = havacGumbelInverseSurvival(pValue, mu, lambda);

  // single-hit core model: N and C loops over MAXL residues, two escapes, B->Mk entry, E->C
  const float nLoop = std::log(maxLength / (maxLength + 3));
  const float nLoopTotal = nLoop * maxLength;
  const float nEscape = std::log(3.0f / (maxLength + 3));
  const float bToMk = std::log(2.0f / (modelLength * (modelLength + 1)));
  const float eToC = std::log(1.0f / 2);
  const float coreAdjustment = (nEscape + nLoopTotal + nEscape + bToMk + eToC);

  // null model: MAXL loops and one move
  const float bgLoop = maxLength / (maxLength + 1);
  const float bgLoopTotal = maxLength * std::log(bgLoop);
  const float bgMove = std::log(1.0 - bgLoop);  // double log, rounded to float on assignment
  const float bgScore = bgLoopTotal + bgMove;

  const float thresholdNats = (fullModelScore * LN2) + bgScore - coreAdjustment;
  const float thresholdBits = thresholdNats / LN2;
  return 256.0f / thresholdBits;
}

void havacProjectEmissions(const float* matchEmissions, size_t count, float scaleFactor, int8_t* out) {
  const float LOG2_E = 1.44269504089;
  const float alpha = 2 * scaleFactor;      // log2(4) * scale: the uniform background 1/4
  const float beta = LOG2_E * scaleFactor;  // nats -> scaled bits
  for (size_t i = 0; i < count; i++) {
    float v = alpha - (matchEmissions[i] * beta);
    v = std::round(v);
    if (v < -128) v = -128;
    if (v > 127) v = 127;
    out[i] = (int8_t)v;
  }
}

std::vector<uint32_t> PhmmModels::lengthPrefixSums() const {
  std::vector<uint32_t> sums;
  sums.reserve(list.size() + 1);
  uint32_t at = 0;
  sums.push_back(at);
  for (const PhmmModel& m : list) sums.push_back(at += m.modelLength);
  return sums;
}

std::vector<float> PhmmModels::threshold256ScaleFactors(float pValue) const {
  std::vector<float> scale;
  scale.reserve(list.size());
  for (const PhmmModel& m : list) {
    if (m.alphabetCardinality != 4)
      throw std::runtime_error("model '" + m.name + "' is not a nucleotide model: only 4 scores/position are supported.");
    if (m.maxLength == 0 || !m.hasMsvStats)
      throw std::runtime_error("model '" + m.name + "' lacks MAXL or STATS LOCAL MSV, which the threshold projection needs.");
    scale.push_back(havacThreshold256ScaleFactor(m.msvGumbelMu, m.msvGumbelLambda, m.maxLength, m.modelLength, pValue));
  }
  return scale;
}

std::vector<uint32_t> PhmmModels::lengths() const {
  std::vector<uint32_t> len;
  len.reserve(list.size());
  for (const PhmmModel& m : list) len.push_back(m.modelLength);
  return len;
}

std::vector<float> PhmmModels::concatenatedMatchEmissions() const {
  size_t rows = 0;
  for (const PhmmModel& m : list) rows += m.modelLength;
  std::vector<float> all;
  all.reserve(rows * 4);
  for (const PhmmModel& m : list) all.insert(all.end(), m.matchEmissionScores.begin(), m.matchEmissionScores.begin() + (size_t)m.modelLength * 4);
  return all;
}

std::shared_ptr<std::vector<int8_t>> PhmmModels::projectForThreshold256(float pValue) const {
  const std::vector<float> scale = threshold256ScaleFactors(pValue);
  size_t rows = 0;
  for (const PhmmModel& m : list) rows += m.modelLength;
  auto table = std::make_shared<std::vector<int8_t>>(rows * 4);
  size_t at = 0, i = 0;
  for (const PhmmModel& m : list) {
    havacProjectEmissions(m.matchEmissionScores.data(), (size_t)m.modelLength * 4, scale[i++], table->data() + at);
    at += (size_t)m.modelLength * 4;
  }
  return table;
}
