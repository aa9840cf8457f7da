/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement (plain C11) of the reference's bzip2 decode path.
 *
 * Nothing under indexed_bzip2_b200/ (the product) may include, link or call this.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it, as the checker.
 *
 * Parity status: PINNED.  tests/test_oracle_vs_reference.py checks every function here against
 *   (a) the reference's own known-answer vectors (testBitStringFinder.cpp:115-146, testRunAB.cpp:12-77,
 *       testBZ2Reader.cpp:284-285, testParallelGzipReader.cpp:1133,1150-1162, the 11 src/tests/data/ *.bz2), and
 *   (b) golden outputs of the compiled, unmodified reference (oracle/_ref, see oracle/Makefile) committed
 *       under tests/golden/ by tests/golden/make_golden.py.
 *
 * Every function cites the reference file:line it restates (paths relative to /root/reference/src).
 */
#ifndef BZ2_ORACLE_H
#define BZ2_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAGIC_BLOCK 0x314159265359ULL /* indexed_bzip2/bzip2.hpp:103 */
#define ORC_MAGIC_EOS   0x177245385090ULL /* indexed_bzip2/bzip2.hpp:104 */
#define ORC_DBUF_SIZE   900000U           /* indexed_bzip2/bzip2.hpp:414-416 */
#define ORC_MAX_GROUPS  6U                /* indexed_bzip2/bzip2.hpp:95 */
#define ORC_GROUP_SIZE  50U               /* indexed_bzip2/bzip2.hpp:96 */
#define ORC_MAX_CODE_LEN 20U              /* indexed_bzip2/bzip2.hpp:97 */
#define ORC_MAX_SYMBOLS 258U              /* indexed_bzip2/bzip2.hpp:98 */
#define ORC_MAX_SELECTORS 32768U          /* indexed_bzip2/bzip2.hpp:451 */
/* Upper bound on Huffman symbols per block that the stage-level buffers hold (selectors * 50). */
#define ORC_SYMBOL_CAPACITY ( 32767U * 50U + 1U )

/* Per-block status. Values are shared with include/ibz2_cuda.h (IBZ2_ST_*); tests assert they agree.
 * Each error names the reference throw site it stands for. */
enum orc_status {
    ORC_OK = 0,
    ORC_ERR_BAD_MAGIC = 1,        /* domain_error  bzip2.hpp:502-507 */
    ORC_ERR_RANDOMISED = 2,       /* domain_error  bzip2.hpp:509-512 */
    ORC_ERR_ORIGPTR_RANGE = 3,    /* logic_error   bzip2.hpp:514-519 */
    ORC_ERR_GROUP_COUNT = 4,      /* logic_error   bzip2.hpp:578-583 */
    ORC_ERR_SELECTOR_COUNT = 5,   /* logic_error   bzip2.hpp:592-597 */
    ORC_ERR_SELECTOR_UNARY = 6,   /* domain_error  bzip2.hpp:625-629 */
    ORC_ERR_CODE_LENGTH = 7,      /* logic_error   bzip2.hpp:655-662 */
    ORC_ERR_CODE_LENGTHS = 8,     /* domain_error  bzip2.hpp:680-683 (over-subscribed, HuffmanCodingBase.hpp:95-102) */
    ORC_ERR_SELECTOR_OVERRUN = 9, /* domain_error  bzip2.hpp:714-718 */
    ORC_ERR_DBUF_OVERFLOW = 10,   /* domain_error  bzip2.hpp:751-756, 776-780 */
    ORC_ERR_INVALID_CODE = 11,    /* bad_optional_access bzip2.hpp:723 */
    ORC_ERR_ORIGPTR = 12,         /* domain_error  bzip2.hpp:794-798 */
    ORC_ERR_CRC = 13,             /* runtime_error bzip2.hpp:900-907 */
    ORC_ERR_EOF = 14,             /* BitReader::EndOfFileReached, core/BitReader.hpp:82 */
    ORC_ERR_OUTPUT_CAPACITY = 15, /* ours: caller's output buffer too small */
    ORC_ERR_STREAM_HEADER = 16,   /* domain_error  bzip2.hpp:114-142 */
    ORC_ERR_STREAM_CRC = 17,      /* runtime_error BZ2Reader.hpp:406-416 */
    ORC_ERR_BLOCK_ORDER = 18      /* invalid_argument core/BlockMap.hpp:101-103 */
};

typedef struct orc_block_info {
    int32_t  status;
    int32_t  is_eos;          /* bzip2.hpp:489 */
    int32_t  at_eof;          /* bzip2.hpp:498, only meaningful for EOS blocks */
    uint64_t enc_offset_bits; /* bzip2.hpp:484 */
    uint64_t enc_size_bits;   /* bzip2.hpp:497 (EOS, incl. padding), :806 (data block) */
    uint32_t header_crc;      /* bzip2.hpp:488 */
    uint32_t calc_crc;        /* bzip2.hpp:901 */
    uint32_t orig_ptr;        /* bzip2.hpp:514 */
    uint32_t symbol_count;    /* bzip2.hpp:560-570 */
    uint8_t  symbol_to_byte[256];
    uint32_t group_count;     /* bzip2.hpp:578 */
    uint32_t selectors_count; /* bzip2.hpp:592 */
    uint8_t  code_lengths[6][258]; /* bzip2.hpp:652-678 */
    uint32_t nsymbols;        /* Huffman symbols decoded, incl. the end-of-block symbol */
    uint32_t n;               /* dbufCount, bzip2.hpp:793 */
    uint64_t out_size;        /* decoded bytes of the block after RLE1 undo */
} orc_block_info;

/* ---- CRC-32/BZIP2 (bzip2.hpp:59-91) ---- */
uint32_t orc_crc32_update( uint32_t crc, const uint8_t* data, size_t n ); /* raw register update, no init/xorout */
uint32_t orc_crc32( const uint8_t* data, size_t n );                      /* init 0xFFFFFFFF, xorout 0xFFFFFFFF */
uint32_t orc_stream_crc_fold( uint32_t streamCrc, uint32_t blockCrc );    /* BZ2Reader.hpp:481-484 */

/* ---- 48-bit magic finder (core/BitStringFinder.hpp:158-238) ----
 * All bit offsets (MSB-first numbering, ascending) at which the 48-bit pattern lies fully inside buf.
 * Returns the total number of matches; at most cap are stored. */
size_t orc_find_magic( const uint8_t* buf, size_t nbytes, uint64_t magic48, uint64_t* out, size_t cap );

/* ---- RUNA/RUNB arithmetic (bzip2.hpp:726-742; KAT table tests/indexed_bzip2/testRunAB.cpp:12-77) ----
 * symbols[i] in {0 = RUNA, 1 = RUNB}; returns the run length the sequence encodes. */
uint32_t orc_run_length( const uint8_t* runSymbols, size_t n );

/* ---- stream header (bzip2.hpp:114-142). Returns level 1..9, or 0 on error / EOF. *pos advances by 32 bits. */
int orc_read_stream_header( const uint8_t* buf, size_t nbytes, uint64_t* bitpos );

/* ---- whole-block decode with stage-level outputs (bzip2.hpp:479-910) ----
 * Any output pointer may be NULL. Capacities: selectors ORC_MAX_SELECTORS, symbols ORC_SYMBOL_CAPACITY (u16),
 * lbytes ORC_DBUF_SIZE, hist 256 (bzip2.hpp:698,759,788: counts BEFORE the prefix sum), tvec ORC_DBUF_SIZE
 * (the reference's packed dbuf after prepare(): (T[k] << 8) | L[k], bzip2.hpp:827-831), prerle ORC_DBUF_SIZE
 * (the N bytes the traversal yields, bzip2.hpp:876-879, before RLE1 undo), out out_cap bytes.
 * Returns info->status. */
int orc_decode_block( const uint8_t* buf, size_t nbytes, uint64_t bit_offset, orc_block_info* info,
                      uint8_t* selectors, uint16_t* symbols, uint8_t* lbytes, uint32_t* hist,
                      uint32_t* tvec, uint8_t* prerle, uint8_t* out, uint64_t out_cap );

/* ---- reader-level semantics ----
 * ParallelBZ2Reader::read to exhaustion + blockOffsets() (ParallelBZ2Reader.hpp:167-269,339-350;
 * core/BlockMap.hpp:69-119,174-198). offsets = (bits, bytes) pairs. *warnTrailing is set when the
 * "[Warning] Trailing garbage after EOF ignored!" path (ParallelBZ2Reader.hpp:224-235) is taken.
 * out may be NULL (size-only). Returns ORC_OK or the status of the first failing block. */
int orc_parallel_reader( const uint8_t* buf, size_t nbytes, uint8_t* out, uint64_t out_cap, uint64_t* out_size,
                         uint64_t* offsets, size_t offsets_cap_pairs, size_t* offsets_count, int* warnTrailing );

/* BZ2Reader::read to exhaustion + blockOffsets() (BZ2Reader.hpp:233-264,395-424,447-491) incl. the
 * stream-CRC check. *stream_crc receives the last calculated stream CRC (BZ2Reader.hpp:172-176). */
int orc_serial_reader( const uint8_t* buf, size_t nbytes, uint8_t* out, uint64_t out_cap, uint64_t* out_size,
                       uint64_t* offsets, size_t offsets_cap_pairs, size_t* offsets_count, int* warnTrailing,
                       uint32_t* stream_crc );

#ifdef __cplusplus
}
#endif
#endif /* BZ2_ORACLE_H */
