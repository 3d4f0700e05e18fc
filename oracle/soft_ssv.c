/*
 * oracle/soft_ssv.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the HAVAC SSV hot path, used as the checker by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg. Nothing under
 * havac_b200/ links, imports or executes this file; the product path fails
 * loudly when its CUDA library is missing.
 *
 * Parity status: PINNED. Every function here is checked against the real
 * reference objects compiled from /root/reference into oracle/_ref/
 * (oracle/Makefile, tests/test_oracle_vs_ref.py) and against the committed
 * vectors under tests/golden/ that those objects produced
 * (tests/golden/make_golden.py).
 *
 * Citations are file:line under the reference tree.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_SEGMENT_SYMBOLS 12288u /* device/PublicDefines.h:18-22 (16 x 768) */

/* ---------------------------------------------------------------------------
 * The SSV recurrence itself. Follows test/softSsv/SoftSsv.cpp:15-67:
 * one persistent row of uint8 scores; for each model row walk the sequence
 * from the right so row[i-1] still holds the previous model row; the cell is
 * prev + e[row][symbol]; negative -> 0; >= 256 -> report and reset to 0.
 * Nothing resets between concatenated models or FASTA records.
 *
 * Output: hits as (seqPos, row) pairs in (row ascending, seqPos descending)
 * order, the order the reference emits. Returns the number of hits found
 * (may exceed cap; only the first cap are stored), or -1 on allocation
 * failure (SoftSsv.cpp:25-29).
 * ------------------------------------------------------------------------- */
int64_t oracle_soft_ssv(const uint8_t *symbols, uint64_t n_symbols,
                        const int8_t *table, uint64_t n_rows,
                        uint64_t *hit_seq_pos, uint32_t *hit_row, uint64_t cap)
{
    uint8_t *row = (uint8_t *)calloc(n_symbols ? n_symbols : 1, 1);
    if (!row) return -1;
    int64_t found = 0;
    for (uint64_t p = 0; p < n_rows; p++) {
        const int8_t *e = table + 4 * p;
        for (int64_t i = (int64_t)n_symbols - 1; i >= 0; i--) {
            int32_t prev = (i == 0) ? 0 : row[i - 1];
            int32_t v = prev + (int32_t)e[symbols[i]];
            if (v < 0) {
                v = 0;
            } else if (v >= 256) {
                v = 0;
                if ((uint64_t)found < cap) {
                    hit_seq_pos[found] = (uint64_t)i;
                    hit_row[found] = (uint32_t)p;
                }
                found++;
            }
            row[i] = (uint8_t)v;
        }
    }
    free(row);
    return found;
}

/* Per-cell record of the recurrence over a window of the matrix: the fields the reference's
 * per-cell debug mode stores for every cell (test/byCellComparator/byCellComparator.hpp:23-31,
 * filled at test/softSsv/SoftSsv.cpp:52-58 under HAVAC_PER_CELL_DATA_TESTING). The whole matrix
 * is walked as in oracle_soft_ssv (the window's cells depend on everything above-left of it);
 * cells with row in [row_begin, row_end) and column in [col_begin, col_end) are written row-major. */
typedef struct {
    uint8_t prevValue;
    int8_t matchScore;
    uint8_t cellValue;
    int8_t phmmVector[4];
    uint8_t symbol;
    uint8_t passesThreshold;
    uint8_t pad[3];
} oracle_cell_record;

int64_t oracle_trace_cells(const uint8_t *symbols, uint64_t n_symbols, const int8_t *table, uint64_t n_rows,
                           uint64_t row_begin, uint64_t row_end, uint64_t col_begin, uint64_t col_end,
                           oracle_cell_record *out)
{
    if (row_end > n_rows || col_end > n_symbols || row_begin > row_end || col_begin > col_end) return -2;
    uint8_t *row = (uint8_t *)calloc(n_symbols ? n_symbols : 1, 1);
    if (!row) return -1;
    const uint64_t width = col_end - col_begin;
    for (uint64_t p = 0; p < row_end; p++) {
        const int8_t *e = table + 4 * p;
        for (int64_t i = (int64_t)n_symbols - 1; i >= 0; i--) {
            int32_t prev = (i == 0) ? 0 : row[i - 1];
            int32_t v = prev + (int32_t)e[symbols[i]];
            int hit = 0;
            if (v < 0) v = 0;
            else if (v >= 256) { v = 0; hit = 1; }
            if (p >= row_begin && (uint64_t)i >= col_begin && (uint64_t)i < col_end) {
                oracle_cell_record *r = out + (p - row_begin) * width + ((uint64_t)i - col_begin);
                r->prevValue = (uint8_t)prev;
                r->matchScore = e[symbols[i]];
                r->cellValue = (uint8_t)v;
                memcpy(r->phmmVector, e, 4);
                r->symbol = symbols[i];
                r->passesThreshold = (uint8_t)hit;
                memset(r->pad, 0, 3);
            }
            row[i] = (uint8_t)v;
        }
    }
    free(row);
    return (int64_t)((row_end - row_begin) * width);
}

/* The same recurrence evaluated the way the FPGA cell does it, with a 9-bit
 * sum and the carry/sign trick (device/HavacHls.cpp:370-402). Used by the
 * oracle self-test to show the two formulations agree on all 65536 inputs. */
void oracle_cell_processor(uint8_t prev, int8_t match, uint8_t *score_out, int *hit_out)
{
    uint32_t sum9 = (uint32_t)prev + (uint32_t)(uint8_t)match;
    int carry = (sum9 >> 8) & 1;
    int sign = ((uint8_t)match >> 7) & 1;
    int reset = carry != sign;
    *score_out = reset ? 0 : (uint8_t)(sum9 & 0xff);
    *hit_out = carry && !sign;
}

/* Replay one diagonal d = seqPos - row over rows [0, n_rows) and write the
 * rows at which it reports. Mirrors the role of attemptTraceback
 * (test/hmmerSsvRef/hmmerSsvRef.cpp:504-586): a bounded, per-diagonal check
 * usable where the full matrix is too large to brute-force. Returns hits. */
int64_t oracle_replay_diagonal(const uint8_t *symbols, uint64_t n_symbols,
                               const int8_t *table, uint64_t n_rows, int64_t diagonal,
                               uint32_t *hit_rows, uint64_t cap)
{
    int64_t found = 0;
    int32_t s = 0;
    uint64_t p0 = diagonal < 0 ? (uint64_t)(-diagonal) : 0;
    for (uint64_t p = p0; p < n_rows; p++) {
        int64_t i = diagonal + (int64_t)p;
        if (i < 0) continue;
        if ((uint64_t)i >= n_symbols) break;
        s += table[4 * p + symbols[i]];
        if (s < 0) s = 0;
        else if (s >= 256) {
            s = 0;
            if ((uint64_t)found < cap) hit_rows[found] = (uint32_t)p;
            found++;
        }
    }
    return found;
}

/* ---------------------------------------------------------------------------
 * Wire formats either side of the kernel.
 * ------------------------------------------------------------------------- */

/* 2-bit unpack: symbol k lives in byte k/4 at bit (k%4)*2, LSB first
 * (host/sequence/SequencePreprocessor.cpp:43-57). */
void oracle_unpack_2bit(const uint8_t *packed, uint64_t n_symbols, uint8_t *symbols)
{
    for (uint64_t k = 0; k < n_symbols; k++)
        symbols[k] = (packed[k >> 2] >> ((k & 3) * 2)) & 3;
}

void oracle_pack_2bit(const uint8_t *symbols, uint64_t n_symbols, uint8_t *packed)
{
    memset(packed, 0, (n_symbols + 3) / 4);
    for (uint64_t k = 0; k < n_symbols; k++)
        packed[k >> 2] |= (uint8_t)((symbols[k] & 3) << ((k & 3) * 2));
}

/* Raw hit word: cell[13:0] | segment[39:14] | row[63:40]
 * (device/HitReporting.cpp:421-430; decode at host/Havac.cpp:154-163). */
uint64_t oracle_encode_hit(uint64_t seq_pos, uint32_t row)
{
    uint64_t seg = seq_pos / ORACLE_SEGMENT_SYMBOLS;
    uint64_t cell = seq_pos % ORACLE_SEGMENT_SYMBOLS;
    return (cell & 0x3fffu) | ((seg & 0x3ffffffu) << 14) | ((uint64_t)(row & 0xffffffu) << 40);
}

/* Resolve a raw hit into (sequenceIndex, positionInSequence, phmmIndex,
 * phmmPosition). seq_end[k] is FastaVector's sequenceEndPosition of record k
 * (counts the record's '\0' terminator, FastaVector.c:19-24,78-86,159-164);
 * the record lookup is the inclusive linear scan of FastaVector.c:345-373,
 * the model lookup the "last prefix sum <= row" search of
 * host/Havac.cpp:119-142. Returns 1 when resolved, 0 when the hit is dropped
 * (padding hit, host/Havac.cpp:169-173, or unresolvable row, :177-180). */
int oracle_resolve_hit(uint64_t raw, const uint64_t *seq_end, uint64_t n_seqs,
                       const uint32_t *model_len, uint32_t n_models,
                       uint64_t *seq_pos_out, uint32_t *seq_idx_out,
                       uint32_t *phmm_pos_out, uint32_t *phmm_idx_out)
{
    uint64_t cell = raw & 0x3fffu;
    uint64_t seg = (raw & ((1ull << 40) - 1)) >> 14;
    uint64_t gpos = seg * ORACLE_SEGMENT_SYMBOLS + cell;
    uint32_t grow = (uint32_t)(raw >> 40);

    uint64_t k, start = 0;
    int found = 0;
    for (k = 0; k < n_seqs; k++) {
        start = k ? seq_end[k - 1] : 0;
        if (gpos <= seq_end[k]) { found = 1; break; }
    }
    if (!found) return 0;

    int64_t idx = -1;
    uint64_t acc = 0, base = 0;
    for (uint32_t m = 0; m <= n_models; m++) { /* prefix sums 0, L0, L0+L1, ... */
        if (acc <= grow) { idx = m; base = acc; }
        if (m < n_models) acc += model_len[m];
    }
    if (idx < 0) return 0;
    *seq_pos_out = gpos - start;
    *seq_idx_out = (uint32_t)k;
    *phmm_pos_out = (uint32_t)(grow - base);
    *phmm_idx_out = (uint32_t)idx;
    return 1;
}
