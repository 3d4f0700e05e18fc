// examples/havac_benchmark.cpp -- a user program against the Havac driver API, timing the same
// five phases the reference's own driver times (benchmark/benchmark.cpp:43-78: build, load,
// run, hit read-back, total). It compiles unchanged against the reference's host/Havac.hpp; here
// it links libhavac_host + libhavac_cuda (B200). One JSON line on stdout.
//
//   havac_benchmark <device spec | xclbin path> <fasta> <hmm> [p-value] [both]
// "both" also searches the reverse complement (Havac::useReverseComplement, this repo's addition;
// leave it out to compile against the reference's header) and adds crick_* fields.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <string>
#include <vector>

#include "Havac.hpp"

static double seconds(std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
  return std::chrono::duration<double>(b - a).count();
}

int main(int argc, char** argv) {
  if (argc < 4) {
    std::fprintf(stderr, "usage: %s <device spec> <fasta> <hmm> [p-value]\n", argv[0]);
    return 2;
  }
  const std::string src = argv[1], fastaSrc = argv[2], phmmSrc = argv[3];
  const float pValue = argc > 4 ? (float)std::atof(argv[4]) : 0.02f;
  srand(1);
  try {
    auto now = [] { return std::chrono::steady_clock::now(); };
    const auto t0 = now();
    std::shared_ptr<Havac> havac = std::make_shared<Havac>(0, pValue, src);
    const auto t1 = now();
    havac->loadPhmm(phmmSrc);
    const auto t2 = now();
    havac->loadSequence(fastaSrc);
    const auto t3 = now();
    havac->runHardwareClient();
    const auto t4 = now();
    std::vector<HavacHit> hits = havac->getHitsFromFinishedRun();
    const auto t5 = now();
    const bool both = argc > 5 && std::string(argv[5]) == "both";
    size_t crickHits = 0;
    double crickReverse = 0, crickRun = 0, crickGet = 0;
    if (both) {
      const auto c0 = now();
      havac->useReverseComplement(true);  // reversed on the device, nothing re-read or re-sent
      const auto c1 = now();
      havac->runHardwareClient();
      const auto c2 = now();
      crickHits = havac->getHitsFromFinishedRun().size();
      const auto c3 = now();
      crickReverse = seconds(c0, c1);
      crickRun = seconds(c1, c2);
      crickGet = seconds(c2, c3);
    }

    const double symbols = (double)havac->sequences().symbolCount();
    uint64_t rows = 0;
    for (const PhmmModel& m : havac->phmms().models()) rows += m.modelLength;
    const double deviceMs = havac->lastDeviceTimeMs();
    std::printf(
        "{\"fasta\": \"%s\", \"hmm\": \"%s\", \"p_value\": %g, \"symbols\": %.0f, \"records\": %llu, \"models\": %zu, "
        "\"rows\": %llu, \"hits\": %zu, \"build_s\": %.4f, \"load_phmm_s\": %.4f, \"load_sequence_s\": %.4f, "
        "\"run_s\": %.4f, \"get_hits_s\": %.4f, \"total_s\": %.4f, \"device_ms\": %.3f, \"gcups_device\": %.1f, "
        "\"gcups_whole_program\": %.1f, \"run_us\": %.1f, \"get_hits_us\": %.1f",
        fastaSrc.c_str(), phmmSrc.c_str(), pValue, symbols, (unsigned long long)havac->sequences().numSequences(),
        (size_t)havac->phmms().count(), (unsigned long long)rows, hits.size(), seconds(t0, t1), seconds(t1, t2),
        seconds(t2, t3), seconds(t3, t4), seconds(t4, t5), seconds(t0, t5), deviceMs,
        symbols * (double)rows / (deviceMs * 1e-3) / 1e9, symbols * (double)rows / seconds(t0, t5) / 1e9, seconds(t3, t4) * 1e6,
        seconds(t4, t5) * 1e6);
    if (both)
      std::printf(", \"crick_hits\": %zu, \"crick_reverse_s\": %.4f, \"crick_run_s\": %.4f, \"crick_get_hits_s\": %.4f", crickHits,
                  crickReverse, crickRun, crickGet);
    std::printf("}\n");
    if (!hits.empty()) std::fprintf(stderr, "first hit: %s\n", hits.front().toString().c_str());
  } catch (const std::exception& e) {
    std::fprintf(stderr, "havac_benchmark: %s\n", e.what());
    return 1;
  }
  return 0;
}
