/*
 * havac_host.h -- C doors into the host-side C++ driver (libhavac_host.so), for bindings
 * that cannot hold a C++ object (the ctypes mirror in havac_b200/havac.py, test harnesses).
 * Each function forwards to one member of `class Havac` (havac_b200/host/Havac.hpp, which keeps
 * the reference's host/Havac.hpp:42-107 API) and turns its exception into a code:
 *   0 ok | -1 invalid argument | -2 std::length_error | -3 std::logic_error |
 *  -4 std::runtime_error | -5 std::bad_alloc
 * The last four functions need no GPU: they expose the input/output contracts either side of
 * the kernel (FASTA -> packed buffer, .hmm -> int8 table, raw hit -> HavacHit).
 */
#ifndef HAVAC_HOST_H
#define HAVAC_HOST_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct havac_host havac_host;

/* Havac::Havac(deviceIndex, requiredPValue, src)            host/Havac.cpp:20-31 */
int havac_host_create(uint32_t deviceIndex, float requiredPValue, const char* src, havac_host** out);
void havac_host_destroy(havac_host* h);
/* Havac::loadSequence / loadPhmm                            host/Havac.cpp:42-77 */
int havac_host_load_sequence(havac_host* h, const char* fastaSrc);
int havac_host_load_phmm(havac_host* h, const char* phmmSrc);
/* Harness doors around loadSequence: seed rand() (non-ACGT symbols and record terminators are
 * rand()-chosen, SequencePreprocessor.cpp:62-85), the record end offsets the driver holds, and the
 * packed buffer as the DEVICE holds it (copied back), to check the device-side FASTA ingest. */
void havac_host_srand(unsigned seed);
int havac_host_sequence_table(havac_host* h, uint64_t* sequenceEnd, uint64_t cap, uint64_t* n);
int havac_host_read_sequence(havac_host* h, uint8_t* packed, uint64_t cap, uint64_t* n);
/* The int8 score table as the DEVICE holds it (copied back), to check the device-side projection. */
int havac_host_read_phmm(havac_host* h, int8_t* flatScores, uint64_t cap, uint64_t* n);
/* Havac::useReverseComplement (no reference counterpart): 1 = the next runs search the reverse
 * complement of the loaded sequence (reversed on the device), hits reported at forward-strand
 * coordinates; 0 = back to the strand in the file. And the same reversal of a packed buffer on the
 * host, in place, for clients that hold a segment range only. */
int havac_host_use_reverse_complement(havac_host* h, int crick);
int havac_host_reverse_complement_packed(uint8_t* packed, uint64_t nBytes, uint64_t nSymbols);
/* Havac::runHardwareClient / Async / wait / abort / state   host/Havac.cpp:80-102,190-192 */
int havac_host_run(havac_host* h);
int havac_host_run_async(havac_host* h);
int havac_host_wait(havac_host* h);
int havac_host_abort(havac_host* h);
int havac_host_state(havac_host* h, int* state);
/* Havac::getHitsFromFinishedRun                             host/Havac.cpp:145-187
 * First call with cap = 0 to learn *n, then again with arrays of that size. */
int havac_host_get_hits(havac_host* h, uint64_t* sequencePosition, uint32_t* sequenceIndex, uint32_t* phmmPosition,
                        uint32_t* phmmIndex, uint64_t cap, uint64_t* n);
int havac_host_device_time_ms(havac_host* h, float* ms);
const char* havac_host_last_error(havac_host* h); /* h may be NULL: last error of this thread */

/* FASTA -> 2-bit packed device buffer + record end offsets (FastaVector + SequencePreprocessor).
 * srand(seed) is called first so the rand()-resolved symbols are reproducible. */
int havac_host_pack_fasta(const char* fastaSrc, unsigned seed, uint8_t* packed, uint64_t packedCap, uint64_t* packedBytes,
                          uint64_t* sequenceEnd, uint64_t endCap, uint64_t* nSequences);
/* .hmm -> flat int8 table at threshold 256 + model lengths (P7HmmReader + PhmmPreprocessor). */
int havac_host_project_phmm(const char* phmmSrc, float pValue, int8_t* table, uint64_t tableCap, uint64_t* tableBytes,
                            uint32_t* modelLength, uint32_t lengthCap, uint32_t* nModels);
/* .hmm -> the match emissions as parsed (negative natural logs, models concatenated, 4 per
 * position), to check the reader's number parsing on its own. */
int havac_host_parse_phmm(const char* phmmSrc, float* matchEmissions, uint64_t cap, uint64_t* nFloats);
/* .hmm -> one "NAME<tab>ACC<newline>" line per model, in file order (ACC may be empty). */
int havac_host_model_names(const char* phmmSrc, char* text, uint64_t cap, uint64_t* nBytes);
/* findThreshold256ScalingFactor                             PhmmReprojection.cpp:36-66 */
float havac_host_scale_factor(float mu, float lambda, uint32_t maxLength, uint32_t modelLength, float pValue);
/* raw u64 hits -> (sequencePosition, sequenceIndex, phmmPosition, phmmIndex); padding hits dropped. */
int havac_host_resolve_hits(const uint64_t* raw, uint64_t nRaw, const uint64_t* sequenceEnd, uint64_t nSequences,
                            const uint32_t* modelLength, uint32_t nModels, uint64_t* sequencePosition,
                            uint32_t* sequenceIndex, uint32_t* phmmPosition, uint32_t* phmmIndex, uint64_t* nOut);

#ifdef __cplusplus
}
#endif
#endif /* HAVAC_HOST_H */
