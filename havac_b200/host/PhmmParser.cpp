// PhmmParser.cpp -- HMMER3 ASCII profile reader (the subset of P7HmmReader the SSV path needs).
// Line grammar follows the HMMER3 save-file format as the reference parser consumes it
// (P7HmmReader/src/p7HmmReader.c: format tag :99, LENG :191-206, MAXL :207-222, ALPH :224-252,
// STATS :495-534, HMM header + one skipped line :535-568, COMPO / insert-0 / begin transitions
// :572-682, then three lines per node :694-1064, "//" ends a model :687-692).
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <stdexcept>
#include <thread>

#include "PhmmModels.hpp"

// The file is mapped and parsed in place, without copying lines or writing terminators: a token
// is a (pointer, length) pair and every line the parser sees ends in '\n' (an unterminated last
// line is copied once), so strtof/strtoul stop inside the mapping. A large database is cut into
// byte ranges at model ends ("//" lines) and the ranges are parsed by separate threads; see
// readP7Hmm for why that gives the serial parser's result or falls back to it.
namespace {

const char* const FORMAT_ERROR = "Phmm file was not formatted correctly.";

inline bool isBlank(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }

struct Token {
  const char* p;
  size_t n;
  bool is(const char* word) const { return std::strlen(word) == n && std::memcmp(p, word, n) == 0; }
  bool startsWith(const char* word) const { const size_t k = std::strlen(word); return n >= k && std::memcmp(p, word, k) == 0; }
};

// Whitespace-separated tokens of [p, eol); a NUL byte ends the line as it would for a C string.
void splitTokens(const char* p, const char* eol, std::vector<Token>& toks) {
  toks.clear();
  while (true) {
    while (p < eol && isBlank(*p)) p++;
    if (p >= eol || !*p) break;
    const char* b = p;
    while (p < eol && *p && !isBlank(*p)) p++;
    toks.push_back({b, (size_t)(p - b)});
  }
}

// number of tokens only (insert-emission and transition lines are checked for their shape:
// nothing on them reaches the score table)
size_t countTokens(const char* p, const char* eol) {
  size_t n = 0;
  while (true) {
    while (p < eol && isBlank(*p)) p++;
    if (p >= eol || !*p) return n;
    n++;
    while (p < eol && *p && !isBlank(*p)) p++;
  }
}

// "%f" of the reference's sscanf == strtof: the correctly rounded float. Plain decimals
// ("1.38629", what HMMER writes) take a short route: the digits as an integer M < 2^53 and the
// power of ten 10^k <= 10^22 are exact doubles, so M / 10^k is the correctly rounded double; its
// conversion to float equals the correctly rounded float unless that double sits within one
// double ulp of the midpoint between two floats, and then (or for any other spelling) strtof decides.
bool parseFloat(const Token& t, float& out) {
  if (t.n == 1 && t.p[0] == '*') { out = HUGE_VALF; return true; }  // zero probability: +inf negative log
  static const double POW10[] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15};
  const char* p = t.p;
  const char* e = t.p + t.n;
  bool neg = false;
  if (p < e && (*p == '-' || *p == '+')) neg = (*p++ == '-');
  uint64_t m = 0;
  int digits = 0, frac = 0;
  bool dot = false, plain = p < e;
  for (; p < e; p++) {
    const unsigned d = (unsigned)(*p - '0');
    if (d <= 9) { m = m * 10 + d; digits++; frac += dot; }
    else if (*p == '.' && !dot) dot = true;
    else { plain = false; break; }
  }
  if (plain && digits >= 1 && digits <= 15) {
    if (m == 0) { out = neg ? -0.0f : 0.0f; return true; }
    const double q = (double)m / POW10[frac];
    if (q > 1e-30 && q < 1e30) {
      uint64_t bits;
      std::memcpy(&bits, &q, 8);
      const uint64_t low = bits & 0x1fffffffull;  // the 29 mantissa bits a float drops
      if (low < 0x0fffffffull || low > 0x10000001ull) { out = neg ? -(float)q : (float)q; return true; }
    }
  }
  char* end = nullptr;
  out = std::strtof(t.p, &end);
  return end != t.p;
}

bool parseUnsigned(const Token& t, uint32_t& out) {  // "%u"
  char* end = nullptr;
  const char* p = t.p;
  if (*p == '-') return false;
  const unsigned long v = std::strtoul(p, &end, 10);
  if (end == p) return false;
  out = (uint32_t)v;
  return true;
}

uint32_t cardinalityOf(const Token& t) {
  std::string a(t.p, t.n);
  for (char& c : a) c = (char)std::tolower((unsigned char)c);
  if (a == "dna" || a == "rna") return 4;
  if (a == "amino") return 20;
  if (a == "coins") return 2;
  if (a == "dice") return 6;
  return 0;
}

struct MappedText {
  const char* data = nullptr;
  size_t size = 0;
  int fd = -1;
  explicit MappedText(const std::string& path) {
    fd = ::open(path.c_str(), O_RDONLY);
    struct stat st;
    if (fd < 0 || ::fstat(fd, &st) != 0 || !S_ISREG(st.st_mode)) {
      if (fd >= 0) ::close(fd);
      throw std::runtime_error("could not open phmm file '" + path + "' for reading.");
    }
    size = (size_t)st.st_size;
    if (size) {
      void* p = ::mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
      if (p == MAP_FAILED) { ::close(fd); throw std::runtime_error("could not read phmm file '" + path + "'."); }
      data = (const char*)p;
    }
  }
  ~MappedText() {
    if (data) ::munmap((void*)data, size);
    if (fd >= 0) ::close(fd);
  }
};

// The line-driven state machine over [begin, end), every line '\n'-terminated. Starts between
// models; `endedBetweenModels` says whether it also stopped there.
struct RangeParser {
  enum State { IDLE, HEADER, SKIP_TRANSITION_LABELS, COMPO_OR_INSERT0, INSERT0, BEGIN_TRANSITIONS, MATCH_LINE, INSERT_LINE,
               TRANSITION_LINE };
  State state = IDLE;
  std::vector<PhmmModel> parsed;
  uint32_t expectedNode = 1;
  std::vector<Token> t;

  [[noreturn]] static void fail() { throw std::runtime_error(FORMAT_ERROR); }
  bool endedBetweenModels() const { return state == IDLE; }

  void run(const char* begin, const char* end) {
    for (const char* p = begin; p < end;) {
      const char* eol = (const char*)std::memchr(p, '\n', (size_t)(end - p));
      if (!eol) eol = end;  // callers pass '\n'-terminated ranges; kept for safety
      line(p, eol);
      p = eol + 1;
    }
  }

  void line(const char* p, const char* eol) {
    PhmmModel* cur = parsed.empty() ? nullptr : &parsed.back();
    if (state == INSERT_LINE || state == TRANSITION_LINE) {
      const size_t n = countTokens(p, eol);
      if (n == 0) return;  // blank line
      if (state == INSERT_LINE) {
        if (n < cur->alphabetCardinality) fail();
        state = TRANSITION_LINE;
      } else {
        if (n < 7) fail();
        expectedNode++;
        state = MATCH_LINE;
      }
      return;
    }
    splitTokens(p, eol, t);
    if (t.empty()) return;  // blank line
    switch (state) {
      case IDLE:
        if (t[0].startsWith("HMMER3")) {
          parsed.emplace_back();
          expectedNode = 1;
          state = HEADER;
        }
        break;
      case HEADER: {
        const Token& tag = t[0];
        if (tag.is("NAME")) { if (t.size() > 1) cur->name.assign(t[1].p, t[1].n); }
        else if (tag.is("ACC")) { if (t.size() > 1) cur->accession.assign(t[1].p, t[1].n); }
        else if (tag.is("LENG")) {
          if (t.size() < 2 || !parseUnsigned(t[1], cur->modelLength) || cur->modelLength == 0) fail();
        } else if (tag.is("MAXL")) {
          if (t.size() < 2 || !parseUnsigned(t[1], cur->maxLength) || cur->maxLength == 0) fail();
        } else if (tag.is("ALPH")) {
          if (t.size() < 2 || (cur->alphabetCardinality = cardinalityOf(t[1])) == 0) fail();
        } else if (tag.is("STATS")) {
          // STATS LOCAL <MSV|VITERBI|FORWARD> <mu|tau> <lambda>
          if (t.size() < 5) fail();
          float a = 0, b = 0;
          if (!parseFloat(t[3], a) || !parseFloat(t[4], b)) fail();
          if (t[1].is("LOCAL") && t[2].is("MSV")) {
            cur->msvGumbelMu = a;
            cur->msvGumbelLambda = b;
            cur->hasMsvStats = true;
          }
        } else if (tag.is("HMM")) {
          if (cur->modelLength == 0 || cur->alphabetCardinality == 0) fail();  // LENG and ALPH are mandatory
          cur->matchEmissionScores.assign((size_t)cur->modelLength * cur->alphabetCardinality, 0.f);
          state = SKIP_TRANSITION_LABELS;
        }
        // every other header tag (DESC, RF, MM, CONS, CS, MAP, DATE, COM, NSEQ, EFFN, CKSUM, GA, TC, NC, BM, SM)
        // carries nothing the score table depends on
        break;
      }
      case SKIP_TRANSITION_LABELS:  // "m->m m->i m->d i->m i->i d->m d->d"
        state = COMPO_OR_INSERT0;
        break;
      case COMPO_OR_INSERT0:
        state = t[0].is("COMPO") ? INSERT0 : BEGIN_TRANSITIONS;  // no COMPO line: this was the insert-0 emission line
        break;
      case INSERT0:
        state = BEGIN_TRANSITIONS;
        break;
      case BEGIN_TRANSITIONS:
        if (t.size() < 7) fail();
        state = MATCH_LINE;
        break;
      case MATCH_LINE: {
        if (t[0].is("//")) {
          if (expectedNode != cur->modelLength + 1) fail();
          state = IDLE;
          break;
        }
        char* endOfNode = nullptr;
        const unsigned long node = std::strtoul(t[0].p, &endOfNode, 10);
        if (endOfNode == t[0].p || node != expectedNode || node > cur->modelLength) fail();
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
};

// Byte offsets that cut [0, size) into up to `parts` ranges, each cut just after a line that
// begins with "//".
std::vector<size_t> cutAtModelEnds(const char* data, size_t size, unsigned parts) {
  std::vector<size_t> cuts{0};
  for (unsigned k = 1; k < parts; k++) {
    size_t at = std::max(cuts.back(), size / parts * k);
    const char* hit = nullptr;
    while (at < size) {
      const char* nl = (const char*)std::memchr(data + at, '\n', size - at);
      if (!nl || (size_t)(nl - data) + 3 > size) break;
      if (nl[1] == '/' && nl[2] == '/') { hit = nl + 1; break; }
      at = (size_t)(nl - data) + 1;
    }
    if (!hit) break;
    const char* eol = (const char*)std::memchr(hit, '\n', size - (size_t)(hit - data));
    if (!eol) break;
    const size_t cut = (size_t)(eol - data) + 1;
    if (cut > cuts.back() && cut < size) cuts.push_back(cut);
  }
  cuts.push_back(size);
  return cuts;
}

}  // namespace

void PhmmModels::readP7Hmm(const std::string& phmmSrc) {
  MappedText file(phmmSrc);
  // everything up to and including the last '\n', in place; an unterminated last line from a copy
  size_t body = file.size;
  while (body > 0 && file.data[body - 1] != '\n') body--;
  std::string tail(file.data + body, file.size - body);
  tail.push_back('\n');

  unsigned threads = std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
  if (const char* env = std::getenv("HAVAC_PARSE_THREADS")) threads = (unsigned)std::max(1, std::atoi(env));
  threads = (unsigned)std::min<size_t>(threads, body / (4u << 20) + 1);  // a thread per 4 MB at least

  std::vector<PhmmModel> parsed;
  bool done = false;
  if (threads > 1) {
    // Ranges cut after "//" lines, each parsed from the between-models state. If every range but
    // the last also ENDS between models, then by induction over the ranges the serial parser is in
    // that same state at every cut, and the concatenation is exactly its result. Anything else
    // (a "//" line inside a header, a malformed model, ...) is re-read serially so that errors
    // and odd files behave as they always did.
    const std::vector<size_t> cuts = cutAtModelEnds(file.data, body, threads);
    const size_t n = cuts.size() - 1;
    std::vector<RangeParser> part(n);
    std::vector<char> ok(n, 0);
    std::vector<std::thread> pool;
    auto work = [&](size_t i) {
      try {
        part[i].run(file.data + cuts[i], file.data + cuts[i + 1]);
        ok[i] = (i + 1 == n) || part[i].endedBetweenModels();
      } catch (const std::exception&) { ok[i] = 0; }
    };
    for (size_t i = 1; i < n; i++) pool.emplace_back(work, i);
    work(0);
    for (std::thread& th : pool) th.join();
    if (std::all_of(ok.begin(), ok.end(), [](char c) { return c != 0; })) {
      RangeParser& last = part[n - 1];
      last.run(tail.data(), tail.data() + tail.size());
      if (!last.endedBetweenModels()) RangeParser::fail();  // truncated model
      for (RangeParser& r : part)
        for (PhmmModel& m : r.parsed) parsed.push_back(std::move(m));
      done = true;
    }
  }
  if (!done) {
    RangeParser all;
    all.run(file.data, file.data + body);
    all.run(tail.data(), tail.data() + tail.size());
    if (!all.endedBetweenModels()) RangeParser::fail();  // truncated model
    parsed = std::move(all.parsed);
  }
  if (parsed.empty()) RangeParser::fail();  // no model at all
  for (PhmmModel& m : parsed) list.push_back(std::move(m));
}
