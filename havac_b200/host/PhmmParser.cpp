// PhmmParser.cpp -- HMMER3 ASCII profile reader (the subset of P7HmmReader the SSV path needs).
// Line grammar follows the HMMER3 save-file format as the reference parser consumes it
// (P7HmmReader/src/p7HmmReader.c: format tag :99, LENG :191-206, MAXL :207-222, ALPH :224-252,
// STATS :495-534, HMM header + one skipped line :535-568, COMPO / insert-0 / begin transitions
// :572-682, then three lines per node :694-1064, "//" ends a model :687-692).
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>

#include "PhmmModels.hpp"

namespace {

const char* const FORMAT_ERROR = "Phmm file was not formatted correctly.";

inline bool isBlank(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }

// Splits `line` in place into whitespace-separated tokens (reusing `toks`' storage: three lines
// per model position make this the hot loop of a large model database).
void splitTokens(char* line, std::vector<char*>& toks) {
  toks.clear();
  char* p = line;
  while (true) {
    while (isBlank(*p)) p++;
    if (!*p) break;
    toks.push_back(p);
    while (*p && !isBlank(*p)) p++;
    if (!*p) break;
    *p++ = 0;
  }
}

// number of whitespace-separated tokens, without touching the line (insert-emission and
// transition lines are only checked for their shape: nothing on them reaches the score table)
size_t countTokens(const char* p) {
  size_t n = 0;
  while (true) {
    while (isBlank(*p)) p++;
    if (!*p) return n;
    n++;
    while (*p && !isBlank(*p)) p++;
  }
}

// "%f" of the reference's sscanf: strtof gives the same correctly rounded float. '*' stands for
// a zero probability (infinite negative log); the reference leaves such a match emission
// unset, here it is +inf, which projects to the most negative score.
bool parseFloat(const char* tok, float& out) {
  if (tok[0] == '*' && tok[1] == 0) { out = HUGE_VALF; return true; }
  char* end = nullptr;
  out = std::strtof(tok, &end);
  return end != tok;
}

uint32_t cardinalityOf(const char* alph) {
  std::string a(alph);
  for (char& c : a) c = (char)std::tolower((unsigned char)c);
  if (a == "dna" || a == "rna") return 4;
  if (a == "amino") return 20;
  if (a == "coins") return 2;
  if (a == "dice") return 6;
  return 0;
}

}  // namespace

void PhmmModels::readP7Hmm(const std::string& phmmSrc) {
  FILE* f = std::fopen(phmmSrc.c_str(), "r");
  if (!f) throw std::runtime_error("could not open phmm file '" + phmmSrc + "' for reading.");
  struct Closer { FILE* f; ~Closer() { std::fclose(f); } } closer{f};

  enum { IDLE, HEADER, SKIP_TRANSITION_LABELS, COMPO_OR_INSERT0, INSERT0, BEGIN_TRANSITIONS, MATCH_LINE, INSERT_LINE,
         TRANSITION_LINE } state = IDLE;
  std::vector<PhmmModel> parsed;
  PhmmModel* cur = nullptr;
  uint32_t expectedNode = 1;
  char* line = nullptr;
  size_t cap = 0;
  struct LineFree { char** p; ~LineFree() { std::free(*p); } } lineFree{&line};

  auto fail = [&]() -> void { throw std::runtime_error(FORMAT_ERROR); };

  std::vector<char> ioBuffer(1 << 20);
  std::setvbuf(f, ioBuffer.data(), _IOFBF, ioBuffer.size());
  std::vector<char*> t;
  while (getline(&line, &cap, f) != -1) {
    if (state == INSERT_LINE || state == TRANSITION_LINE) {
      const size_t n = countTokens(line);
      if (n == 0) continue;  // blank line
      if (state == INSERT_LINE) {
        if (n < cur->alphabetCardinality) fail();
        state = TRANSITION_LINE;
      } else {
        if (n < 7) fail();
        expectedNode++;
        state = MATCH_LINE;
      }
      continue;
    }
    splitTokens(line, t);
    if (t.empty()) continue;  // blank line
    switch (state) {
      case IDLE:
        if (std::strncmp(t[0], "HMMER3", 6) == 0) {
          parsed.emplace_back();
          cur = &parsed.back();
          expectedNode = 1;
          state = HEADER;
        }
        break;
      case HEADER: {
        const std::string tag = t[0];
        if (tag == "NAME") { if (t.size() > 1) cur->name = t[1]; }
        else if (tag == "LENG") {
          if (t.size() < 2 || std::sscanf(t[1], "%u", &cur->modelLength) != 1 || cur->modelLength == 0) fail();
        } else if (tag == "MAXL") {
          if (t.size() < 2 || std::sscanf(t[1], "%u", &cur->maxLength) != 1 || cur->maxLength == 0) fail();
        } else if (tag == "ALPH") {
          if (t.size() < 2 || (cur->alphabetCardinality = cardinalityOf(t[1])) == 0) fail();
        } else if (tag == "STATS") {
          // STATS LOCAL <MSV|VITERBI|FORWARD> <mu|tau> <lambda>
          if (t.size() < 5) fail();
          float a = 0, b = 0;
          if (!parseFloat(t[3], a) || !parseFloat(t[4], b)) fail();
          if (std::strcmp(t[1], "LOCAL") == 0 && std::strcmp(t[2], "MSV") == 0) {
            cur->msvGumbelMu = a;
            cur->msvGumbelLambda = b;
            cur->hasMsvStats = true;
          }
        } else if (tag == "HMM") {
          if (cur->modelLength == 0 || cur->alphabetCardinality == 0) fail();  // LENG and ALPH are mandatory
          cur->matchEmissionScores.assign((size_t)cur->modelLength * cur->alphabetCardinality, 0.f);
          state = SKIP_TRANSITION_LABELS;
        }
        // every other header tag (ACC, DESC, RF, MM, CONS, CS, MAP, DATE, COM, NSEQ, EFFN, CKSUM, GA, TC, NC, BM, SM)
        // carries nothing the score table depends on
        break;
      }
      case SKIP_TRANSITION_LABELS:  // "m->m m->i m->d i->m i->i d->m d->d"
        state = COMPO_OR_INSERT0;
        break;
      case COMPO_OR_INSERT0:
        if (std::strcmp(t[0], "COMPO") == 0) { state = INSERT0; break; }
        state = BEGIN_TRANSITIONS;  // no COMPO line: this was the insert-0 emission line
        break;
      case INSERT0:
        state = BEGIN_TRANSITIONS;
        break;
      case BEGIN_TRANSITIONS:
        if (t.size() < 7) fail();
        state = MATCH_LINE;
        break;
      case MATCH_LINE: {
        if (std::strcmp(t[0], "//") == 0) {
          if (expectedNode != cur->modelLength + 1) fail();
          state = IDLE;
          cur = nullptr;
          break;
        }
        char* endOfNode = nullptr;
        const unsigned long node = std::strtoul(t[0], &endOfNode, 10);
        if (endOfNode == t[0] || node != expectedNode || node > cur->modelLength) fail();
        const uint32_t K = cur->alphabetCardinality;
        if (t.size() < 1 + K) fail();
        for (uint32_t a = 0; a < K; a++)
          if (!parseFloat(t[1 + a], cur->matchEmissionScores[(size_t)(node - 1) * K + a])) fail();
        state = INSERT_LINE;
        break;
      }
      case INSERT_LINE:
      case TRANSITION_LINE:
        break;  // handled above, without tokenising
    }
  }
  if (state != IDLE || parsed.empty()) fail();  // truncated model, or no model at all
  for (PhmmModel& m : parsed) list.push_back(std::move(m));
}
