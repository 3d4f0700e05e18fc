/*
 * havac_cuda.h -- the drop-in boundary: a C-ABI device client for the HAVAC SSV
 * filter on NVIDIA B200 (sm_100a).
 *
 * It serves exactly the calls the reference's `class HavacHwClient` makes into
 * XRT (host/HavacHwClient.hpp:21-115); the `Havac` driver above it keeps its
 * public API unchanged (host/Havac.hpp:42-107). Plain pointers and sizes only:
 * no C++, CUDA or torch types cross this boundary. Every function returns 0 on
 * success or a negative HAVAC_CUDA_E_* code and never throws.
 *
 * Wire formats (bit-identical to the FPGA design):
 *   sequence : 2 bits/symbol, symbol k in byte k/4 at bit (k%4)*2, a whole number
 *              of 12288-symbol segments (device/PublicDefines.h:18-22,
 *              host/sequence/SequencePreprocessor.cpp:37-59).
 *   table    : int8 [row][A,C,G,T], models concatenated with no separator, hit
 *              threshold projected to exactly 256 (host/phmm/PhmmPreprocessor.cpp:9-31).
 *   hit      : u64 = cell[13:0] | segment[39:14] | row[63:40]
 *              (device/HitReporting.cpp:421-430).
 */
#ifndef HAVAC_CUDA_H
#define HAVAC_CUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HAVAC_CUDA_SEGMENT_SYMBOLS 12288u

/* error codes */
#define HAVAC_CUDA_OK 0
#define HAVAC_CUDA_E_INVALID (-1)   /* bad argument / null handle                       */
#define HAVAC_CUDA_E_LENGTH (-2)    /* size contract broken  -> std::length_error        */
#define HAVAC_CUDA_E_LOGIC (-3)     /* call out of order     -> std::logic_error         */
#define HAVAC_CUDA_E_RUNTIME (-4)   /* CUDA failure          -> std::runtime_error       */
#define HAVAC_CUDA_E_NOMEM (-5)     /* allocation failure    -> std::bad_alloc           */
#define HAVAC_CUDA_E_NODEVICE (-6)  /* no usable CUDA device -> std::runtime_error       */
#define HAVAC_CUDA_E_OVERFLOW (-7)  /* device hit buffer overflowed (the FPGA design does
                                       not check this: device/HitReporting.cpp:412)     */

/* Command states: numerically identical to XRT's ert_cmd_state, which
 * HavacHwClient::getHwState returns and host/Havac.hpp:16-26 mirrors. */
#define HAVAC_CUDA_STATE_NEW 1
#define HAVAC_CUDA_STATE_QUEUED 2
#define HAVAC_CUDA_STATE_RUNNING 3
#define HAVAC_CUDA_STATE_COMPLETED 4
#define HAVAC_CUDA_STATE_ERROR 5
#define HAVAC_CUDA_STATE_ABORT 6
#define HAVAC_CUDA_STATE_SUBMITTED 7
#define HAVAC_CUDA_STATE_TIMEOUT 8
#define HAVAC_CUDA_STATE_NORESPONSE 9

typedef struct havac_cuda_client havac_cuda_client; /* opaque, one per Havac object */

/* Replaces HavacHwClient::HavacHwClient (host/HavacHwClient.cpp:24-31: open device,
 * load xclbin, allocate the hit buffers). `deviceSpec` takes the place of the xclbin
 * path: NULL, "" or anything not starting with "cuda:" selects the single GPU
 * `deviceIndex`; "cuda:all" or "cuda:0,1,2" shards the sequence over those GPUs of
 * this process. The hit buffers start at 16 Mi hits per device and GROW: a run that overflows
 * them is repeated with buffers sized from the count the kernel reported, so callers see one
 * completed run (the FPGA design writes past its fixed 3.5 GiB buffer instead,
 * device/HitReporting.cpp:412). ";hitcap=<n>+" sets the first size; ";hitcap=<n>" (no plus) fixes
 * the capacity, and an overflow is then reported as E_OVERFLOW. */
int havac_cuda_create(const char* deviceSpec, uint32_t deviceIndex, havac_cuda_client** out);
int havac_cuda_destroy(havac_cuda_client* c);

/* Replaces HavacHwClient::writeSequence (host/HavacHwClient.cpp:78-110).
 * nBytes*4 must be a multiple of 12288 and nBytes < 4 GiB, else E_LENGTH. The data
 * is copied; the caller keeps ownership of `packed2bit`. */
int havac_cuda_write_sequence(havac_cuda_client* c, const uint8_t* packed2bit, uint64_t nBytes);

/* Replaces HavacHwClient::writePhmm (host/HavacHwClient.cpp:111-138).
 * nBytes must be a multiple of 4 and < 1 GiB, else E_LENGTH. */
int havac_cuda_write_phmm(havac_cuda_client* c, const int8_t* flatScores, uint64_t nBytes);

/* Restrict this client to the sequence segments [segBegin, segEnd) of the buffer
 * given to write_sequence (it keeps a left halo of rows-1 symbols so the hits in
 * that range are exact). Used by one-process-per-GPU launchers to shard the
 * database; call before write_sequence. segEnd = 0 means "to the end". No
 * reference counterpart (the FPGA is a single card). */
int havac_cuda_set_segment_range(havac_cuda_client* c, uint64_t segBegin, uint64_t segEnd);

/* Replaces HavacHwClient::invokeHavacSsvAsync (host/HavacHwClient.cpp:141-151):
 * E_LENGTH when either input has zero length; returns as soon as the work is queued. */
int havac_cuda_invoke_async(havac_cuda_client* c);

/* Replaces HavacHwClient::getHwState (host/HavacHwClient.cpp:163-170): E_LOGIC when
 * nothing was ever launched. */
int havac_cuda_state(havac_cuda_client* c, int* state);

/* Replaces HavacHwClient::waitForHavacSsvAsync (host/HavacHwClient.cpp:153-157):
 * timeoutMs = 0 waits forever; otherwise *state is HAVAC_CUDA_STATE_TIMEOUT when the
 * run is still going after timeoutMs. */
int havac_cuda_wait(havac_cuda_client* c, uint64_t timeoutMs, int* state);

/* Replaces HavacHwClient::abort (host/HavacHwClient.cpp:159-161): raises a flag the
 * persistent kernel polls between tiles, waits for it to drain, reports ABORT. */
int havac_cuda_abort(havac_cuda_client* c, int* state);

/* Replace HavacHwClient::getNumHits / getHitList (host/HavacHwClient.cpp:172-203).
 * Hits come back in the FPGA's emission order (segment, row, cell ascending).
 * E_LOGIC before a completed run, E_OVERFLOW when the device buffer overflowed. */
int havac_cuda_num_hits(havac_cuda_client* c, uint64_t* n);
int havac_cuda_read_hits(havac_cuda_client* c, uint64_t* out, uint64_t cap, uint64_t* n);

/* Score table built ON THE DEVICE from the models' negative-log match emissions ([row][A,C,G,T]
 * floats as P7HmmReader holds them, models concatenated): stands for the loop of
 * p7HmmProjectForThreshold256 (PhmmReprojection/PhmmReprojection.cpp:109-145) over all models,
 * PhmmPreprocessor's concatenation (host/phmm/PhmmPreprocessor.cpp:9-31) and writePhmm, with
 * byte-identical int8 scores. `scaleFactor[m]` is findThreshold256ScalingFactor of model m
 * (PhmmReprojection.cpp:36-66), which stays on the host. havac_cuda_read_phmm copies the int8
 * table back (parity checks). */
int havac_cuda_project_phmm(havac_cuda_client* c, const float* negLogEmissions, uint64_t nRows, const uint32_t* modelLength,
                            const float* scaleFactor, uint32_t nModels);
int havac_cuda_read_phmm(havac_cuda_client* c, int8_t* flatScores, uint64_t capBytes, uint64_t* nBytes);

/* The other strand. Replaces the resident sequence, on the device, by its reverse complement over
 * the first `nSymbols` symbols (the real ones: records and their terminators; the zero padding up
 * to whole segments stays at the end), so that the next invoke searches the Crick strand, which the
 * reference leaves to a second run on a reversed file (benchmark/readme.txt:62-64 pins nhmmer to
 * --watson for the comparison). A second call with the same count restores the forward sequence;
 * writing or ingesting a sequence resets the state. While reversed, raw hits (read_hits) are in
 * reversed coordinates and havac_cuda_resolve_hits reports forward-strand positions
 * (nSymbols - 1 - position; hits in the padding are dropped). Clients limited to a segment range
 * (set_segment_range) answer E_LOGIC: reverse the packed buffer on the host and write it.
 * havac_cuda_strand_state: 0 = forward, else the count the sequence is reversed over. */
int havac_cuda_reverse_complement(havac_cuda_client* c, uint64_t nSymbols);
int havac_cuda_strand_state(havac_cuda_client* c, uint64_t* mirroredSymbols);

/* Hits of the finished run resolved ON THE DEVICE to the four numbers of a HavacHit, in emission
 * order, hits in the zero padding dropped: stands for the loop of Havac::getHitsFromFinishedRun
 * (host/Havac.cpp:145-187) with fastaVectorGetLocalSequencePositionFromGlobal (FastaVector/src/
 * FastaVector.c:345-373, inclusive end test kept) and the model prefix-sum search
 * (host/Havac.cpp:104-142). `sequenceEnd` = FastaVector's sequenceEndPosition per record,
 * `modelLength` = LENG per model. Call with cap = 0 to learn an upper bound of *n. */
int havac_cuda_resolve_hits(havac_cuda_client* c, const uint64_t* sequenceEnd, uint64_t nSequences,
                            const uint32_t* modelLength, uint32_t nModels, uint64_t* sequencePosition,
                            uint32_t* sequenceIndex, uint32_t* phmmPosition, uint32_t* phmmIndex, uint64_t cap,
                            uint64_t* n);

/* FASTA text -> device-resident packed sequence, parsed and packed ON THE DEVICE. For a client
 * that holds no sequence yet this stands for fastaVectorReadFasta (FastaVector/src/FastaVector.c:
 * 54-168) + SequencePreprocessor (host/sequence/SequencePreprocessor.cpp:9-85) + writeSequence
 * (host/HavacHwClient.cpp:78-110) in one call, with bit-identical symbols, record end offsets
 * and padding. `text` is the whole file in host memory. Symbols that are not A/C/G/T, and the
 * terminator of every record, are chosen with rand() by the reference (SequencePreprocessor.cpp:
 * 62-85); `resolve(user, c)` is called once for each of them IN SEQUENCE ORDER (c == 0 for a
 * terminator) and returns the 2-bit symbol, so the caller keeps that state. E_RUNTIME when the
 * text has no '>' header before its first sequence line, E_LENGTH at >= 4 GiB packed. */
typedef uint8_t (*havac_cuda_symbol_fn)(void* user, unsigned char c);
int havac_cuda_ingest_fasta(havac_cuda_client* c, const uint8_t* text, uint64_t nBytes, havac_cuda_symbol_fn resolve,
                            void* user, uint64_t* nSymbols, uint64_t* nRecords);
/* Record table of the last ingest: FastaVector's sequenceEndPosition per record (terminators
 * counted) and the byte offset of each record's '>' in the text. Call with cap = 0 to learn *n. */
int havac_cuda_ingest_records(havac_cuda_client* c, uint64_t* sequenceEnd, uint64_t* headerByteOffset, uint64_t cap,
                              uint64_t* n);
/* The packed sequence resident on a single-device client, copied back (parity checks). */
int havac_cuda_read_sequence(havac_cuda_client* c, uint8_t* packedOut, uint64_t capBytes, uint64_t* nBytes);

/* Harness helpers (no reference counterpart). */
const char* havac_cuda_last_error(havac_cuda_client* c); /* c may be NULL: last create error */
int havac_cuda_device_time_ms(havac_cuda_client* c, float* ms); /* CUDA-event time of the last
                                                                   run, max over devices       */
int havac_cuda_device_count(havac_cuda_client* c, int* n);
/* Per-run facts for roofline accounting: cells = real cells computed (rows x columns
 * of this client's range, halo and padding excluded); launches = kernels launched. */
int havac_cuda_run_stats(havac_cuda_client* c, uint64_t* cells, uint32_t* launches);
/* SM clock cycles the first persistent CTA spent in the last run's kernel (clock64), max over
 * devices; cycles / device_time = the SM clock times the share of the launch that CTA was alive for
 * (a persistent CTA leaves when the tiles run out, so the quotient falls short of the clock by the
 * kernel's tail). */
int havac_cuda_sm_cycles(havac_cuda_client* c, uint64_t* cycles);
/* Running totals of the bytes this client copied host->device and device->host. */
int havac_cuda_transfer_stats(havac_cuda_client* c, uint64_t* h2dBytes, uint64_t* d2hBytes);
/* Per-cell debug trace: the reference's HAVAC_PER_CELL_DATA_TESTING mode (device/PublicDefines.h:11; record
 * schema of test/byCellComparator/byCellComparator.hpp:23-31) for a window of the matrix. Runs the plain integer
 * recurrence on what the (single) device holds -- its packed symbols, its int8 table -- and writes one record per
 * cell with row in [rowBegin, rowEnd) and global column in [colBegin, colEnd), row-major. For triage of a parity
 * failure: tests/oracle_lib.py walks these against the CPU oracle's records and names the first cell that differs. */
typedef struct havac_cuda_cell_record {
  uint8_t prevValue;
  int8_t matchScore;
  uint8_t cellValue;
  int8_t phmmVector[4];
  uint8_t symbol;
  uint8_t passesThreshold;
  uint8_t pad[3];
} havac_cuda_cell_record;
int havac_cuda_trace_cells(havac_cuda_client* c, uint32_t rowBegin, uint32_t rowEnd, uint64_t colBegin, uint64_t colEnd,
                           havac_cuda_cell_record* out, uint64_t capRecords, uint64_t* nRecords);
/* Debug: overwrite one byte of the score table AS THE DEVICE HOLDS IT (every device of the client) and rebuild the
 * fast path's fp16 rows from it. The host's copy is untouched -- this is how the tests stage "device and host
 * disagree" for the trace above. */
int havac_cuda_debug_poke_phmm(havac_cuda_client* c, uint64_t byteOffset, int8_t value);
/* The fast kernel's cell update on all 65,536 (previous score 0..255, emission -128..127) pairs, one launch:
 * score[i] / hit[i] for pair i = prev * 256 + (emission + 128) computed in the low fp16 half, score[65536 + i] /
 * hit[65536 + i] the same pair computed in the high half. hit bit 7 set = the kernel's own overflow test disagreed
 * with the halves. Both buffers hold 131,072 bytes. Checked against device/HavacHls.cpp:370-402 in the tests. */
int havac_cuda_cell_selftest(havac_cuda_client* c, uint8_t* score, uint8_t* hit);
/* How the last launch on device `device` of this client cut its work: global column of the first resident
 * symbol the kernel was given, first diagonal (column - row, relative to that symbol) of tile 0, number of
 * tiles, diagonals per tile, rows per staged score-table chunk, rows per unrolled block. Parity tests use
 * it to aim their diagonal replays at tile / thread / chunk / block edges. Any pointer may be NULL. */
int havac_cuda_tile_geometry(havac_cuda_client* c, uint32_t device, int64_t* loadBegin, int64_t* diagBegin, uint32_t* nTiles,
                             uint32_t* tileDiagonals, uint32_t* rowChunk, uint32_t* blockRows);
/* Select a kernel variant by name ("auto" | "f16x2" | "f16x2-streaming" | "scalar"); for tests that cross-check the
 * fast path against the plain one. "f16x2-streaming" keeps tables of at most 512 rows away from the short-table
 * kernel (every warp on its own) so that the streaming kernels can be tested with short tables too. */
int havac_cuda_set_kernel(havac_cuda_client* c, const char* name);

/* Emission-order sort of raw hit words that ALLOCATES NOTHING. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it). Same kernels and same result as the sort behind havac_cuda_read_hits
 * (FPGA emission order: segment, row, cell ascending -- device/HavacHls.cpp:131-162), but its scratch (two key
 * buffers, digit counts) lives in a file-scope device array of the library instead of cudaMalloc'ed buffers, so the
 * call makes no cudaMalloc / cudaFree / allocating library call. Every GPU that loads the library holds that array
 * (16.25 MiB) until the process exits, and it bounds the size: n > HAVAC_CUDA_SORT_STATIC_MAX_HITS -> E_LENGTH.
 *   deviceIndex : CUDA device ordinal; becomes the current device.
 *   hitsIn      : n raw hit words, device OR host memory (unified addressing decides); not modified.
 *   hitsOut     : n words, device or host memory; may equal hitsIn.
 *   stream      : a cudaStream_t of that device, or NULL for the default stream. The call returns after the stream
 *                 has drained (it reads one word back between its kernels), so hitsOut is complete on return.
 * No client handle: the text of a failure is read with havac_cuda_last_error(NULL) on the calling thread.
 * All calls share the one array per device and it keeps the last call's contents (what must start clean is
 * re-initialised by every call). Calls through this entry point are serialised by a mutex inside the library; work
 * that overlaps ON THE GPU with a call and touches the array (there is no other door to it) would race. */
#define HAVAC_CUDA_SORT_STATIC_MAX_HITS (1u << 20)
int havac_cuda_sort_hits_static(uint32_t deviceIndex, const uint64_t* hitsIn, uint64_t n, uint64_t* hitsOut, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HAVAC_CUDA_H */
