/*
 * ibz2_cuda.h -- the thin C ABI between the host-side readers (C++ / Python / any FFI) and the sm_100a CUDA
 * kernels that replace the bzip2 hot path of indexed_bzip2.  Plain pointers and sizes only; no C++ types, no
 * exceptions, no torch types.  Every call returns an ibz2 status (0 = ok); per-block problems are reported in
 * ibz2_block_result.status so that the host side can rethrow the exception the reference would have thrown.
 *
 * Reference interfaces each entry point stands in for (paths relative to the reference's src/):
 *
 *   ibz2_find_magics      core/BitStringFinder.hpp:158-320 (BitStringFinder<48>::find / findBitStrings),
 *                         core/ParallelBitStringFinder.hpp:159-265, as driven by core/BlockFinder.hpp:164-197
 *                         and tools/ibzip2.cpp:96-143 (findCompressedBlocks, both magics, sorted)
 *   ibz2_decode_blocks    indexed_bzip2/BZ2BlockFetcher.hpp:85-138 (decodeBlock: Block ctor + readBlockData + read
 *                         until eob) i.e. indexed_bzip2/bzip2.hpp:479-910, for a whole batch of block offsets
 *   ibz2_decode_buffer    indexed_bzip2/ParallelBZ2Reader.hpp:167-269 (finder + fetcher over an in-memory file)
 *   ibz2_read_block_header indexed_bzip2/BZ2BlockFetcher.hpp:64-82 (readBlockHeader: EOS probe), host-side, no GPU
 *   ibz2_read_stream_header indexed_bzip2/bzip2.hpp:114-142 (readBzip2Header), host-side, no GPU
 *   ibz2_stream_crc_fold  indexed_bzip2/BZ2Reader.hpp:481-484 (fold of the per-block CRCs the GPU computed)
 *   ibz2_fold_stream_crcs indexed_bzip2/BZ2Reader.hpp:406-416 (stream CRC check at the end-of-stream block), for a
 *                         whole array of block results
 *
 * There is deliberately NO CPU decode behind this interface: if no CUDA device is usable, ibz2_create fails.
 */
#ifndef IBZ2_CUDA_H
#define IBZ2_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IBZ2_ABI_VERSION 3

#define IBZ2_MAGIC_BLOCK 0x314159265359ULL /* indexed_bzip2/bzip2.hpp:103 */
#define IBZ2_MAGIC_EOS   0x177245385090ULL /* indexed_bzip2/bzip2.hpp:104 */
#define IBZ2_MAX_BLOCK_SYMBOLS 900000U     /* dbuf capacity, indexed_bzip2/bzip2.hpp:414-416 */

/* which magics ibz2_find_magics reports */
#define IBZ2_FIND_BLOCK 1U
#define IBZ2_FIND_EOS   2U
/* bit 63 of a reported offset marks an end-of-stream magic */
#define IBZ2_OFFSET_EOS_FLAG ( 1ULL << 63 )

/* ibz2_block_result.stream_flags */
#define IBZ2_STREAM_FIRST_BLOCK 1U /* a stream header ("BZh1".."BZh9", bzip2.hpp:114-142) stands directly in front of the block */
#define IBZ2_STREAM_LAST_BLOCK  2U /* an end-of-stream block follows the block directly; stream_crc_expected is its CRC field */
#define IBZ2_STREAM_CRC_CHECKED 4U /* stream_crc_calculated holds the fold of the block CRCs of the whole stream */

/* where a buffer passed through the ABI lives */
#define IBZ2_MEM_HOST   0 /* pageable or pinned host memory; the call copies through the PCIe link */
#define IBZ2_MEM_DEVICE 1 /* memory of the context's CUDA device (16-byte aligned) */

/* Call-level return codes */
enum ibz2_rc {
    IBZ2_OK = 0,
    IBZ2_E_INVALID_ARGUMENT = 1,
    IBZ2_E_CUDA = 2,          /* a CUDA runtime call failed; see ibz2_last_error */
    IBZ2_E_NO_DEVICE = 3,     /* no usable CUDA device: there is no CPU fallback */
    IBZ2_E_CAPACITY = 4,      /* caller's output array too small; *count / *total hold the needed size */
    IBZ2_E_OUT_OF_MEMORY = 5
};

/* Per-block status.  Same numbering as oracle/bz2_oracle.h (tests assert the two agree).  Each value names the
 * reference throw site it stands for. */
enum ibz2_block_status {
    IBZ2_ST_OK = 0,
    IBZ2_ST_BAD_MAGIC = 1,        /* domain_error  bzip2.hpp:502-507 */
    IBZ2_ST_RANDOMISED = 2,       /* domain_error  bzip2.hpp:509-512 */
    IBZ2_ST_ORIGPTR_RANGE = 3,    /* logic_error   bzip2.hpp:514-519 */
    IBZ2_ST_GROUP_COUNT = 4,      /* logic_error   bzip2.hpp:578-583 */
    IBZ2_ST_SELECTOR_COUNT = 5,   /* logic_error   bzip2.hpp:592-597 */
    IBZ2_ST_SELECTOR_UNARY = 6,   /* domain_error  bzip2.hpp:625-629 */
    IBZ2_ST_CODE_LENGTH = 7,      /* logic_error   bzip2.hpp:655-662 */
    IBZ2_ST_CODE_LENGTHS = 8,     /* domain_error  bzip2.hpp:680-683 */
    IBZ2_ST_SELECTOR_OVERRUN = 9, /* domain_error  bzip2.hpp:714-718 */
    IBZ2_ST_DBUF_OVERFLOW = 10,   /* domain_error  bzip2.hpp:751-756, 776-780 */
    IBZ2_ST_INVALID_CODE = 11,    /* bad_optional_access bzip2.hpp:723 */
    IBZ2_ST_ORIGPTR = 12,         /* domain_error  bzip2.hpp:794-798 */
    IBZ2_ST_CRC = 13,             /* runtime_error bzip2.hpp:900-907 */
    IBZ2_ST_EOF = 14,             /* BitReader::EndOfFileReached, core/BitReader.hpp:82 */
    IBZ2_ST_OUTPUT_CAPACITY = 15, /* ours: the caller's output buffer cannot hold this block */
    IBZ2_ST_STREAM_HEADER = 16,   /* domain_error  bzip2.hpp:114-142 */
    IBZ2_ST_STREAM_CRC = 17,      /* runtime_error BZ2Reader.hpp:406-416 */
    IBZ2_ST_BLOCK_ORDER = 18      /* invalid_argument core/BlockMap.hpp:101-103 */
};

/* What BZ2BlockFetcher::BlockData carries (BZ2BlockFetcher.hpp:18-34) minus the bytes, which land in the
 * caller's output buffer at out_offset. */
typedef struct ibz2_block_result {
    uint64_t enc_offset_bits; /* BlockHeaderData::encodedOffsetInBits */
    uint64_t enc_size_bits;   /* BlockHeaderData::encodedSizeInBits (EOS: incl. padding to the byte boundary) */
    uint64_t out_offset;      /* where this block's bytes start inside the output buffer of the call */
    uint64_t out_size;        /* decoded bytes of the block (after RLE1 undo) */
    uint32_t expected_crc;    /* BlockHeaderData::expectedCRC */
    uint32_t calculated_crc;  /* BlockData::calculatedCRC */
    uint32_t status;          /* enum ibz2_block_status */
    uint32_t is_eos;          /* BlockHeaderData::isEndOfStreamBlock */
    uint32_t is_eof;          /* BlockHeaderData::isEndOfFile (EOS blocks only) */
    uint32_t n_symbols;       /* entries of the BWT vector (dbufCount, bzip2.hpp:793); for a rejected header field (IBZ2_ST_GROUP_COUNT,
                               * _SELECTOR_COUNT, _SELECTOR_UNARY, _CODE_LENGTH) the value the reference's message quotes
                               * (bzip2.hpp:581, 595, 627, 659); for IBZ2_ST_DBUF_OVERFLOW the count that did not fit
                               * (bzip2.hpp:751-756, 776-780); for IBZ2_ST_SELECTOR_OVERRUN the
                               * block's selector count, which the reference's message quotes (bzip2.hpp:714-718) */
    uint32_t orig_ptr;        /* bzip2.hpp:514; for IBZ2_ST_BAD_MAGIC the lower 24 of the 48 bits found where the magic
                               * should be (the upper 24 in n_symbols): the reference's message quotes them */
    uint32_t stream_flags;    /* IBZ2_STREAM_*: where the streams of the file begin and end among the blocks of the call */
    uint32_t stream_crc_expected;   /* last block of a stream: BlockHeader::headerCRC() of the EOS block behind it */
    uint32_t stream_crc_calculated; /* last block of a stream whose first block is part of the same call and whose blocks
                                     * were given in file order: the fold of BZ2Reader.hpp:481-484 over their CRCs */
    uint32_t n_coded;         /* Huffman-coded symbols of the block without the end-of-block symbol (bzip2.hpp:709-723) */
    uint32_t reserved;
} ibz2_block_result;

/* GPU time of the stages of the last ibz2_find_magics / ibz2_decode_blocks call, in milliseconds (CUDA events on
 * the context's stream).  The stage names mirror Block::Statistics::Durations (bzip2.hpp:161-199). */
typedef struct ibz2_timings {
    float h2d_ms;          /* host -> device copy of the compressed bytes (0 for device input) */
    float find_ms;         /* K1 magic finder */
    float huffman_ms;      /* K2 header + tables + Huffman/RLE2/MTF symbol loop (readBlockHeader + readBlockData) */
    float bwt_prepare_ms;  /* K4a counting-sort scatter (BurrowsWheelerTransformData::prepare) */
    float bwt_traverse_ms; /* K4b permutation walk (first half of decodeBlock) */
    float rle_scan_ms;     /* K5a RLE1 state machine + output offsets */
    float rle_write_ms;    /* K5b RLE1 expansion into the output buffer */
    float crc_ms;          /* K6 CRC32 of the output + verification */
    float d2h_ms;          /* device -> host copy of the decoded bytes (0 for device output) */
    float total_ms;        /* first event to last event of the call */
    uint64_t launches;     /* kernels launched by the call */
    float huffman_bits_ms; /* K2b alone (the bit-serial symbol loop inside huffman_ms) */
    float bwt_walk_ms;     /* the walk kernel alone (inside bwt_traverse_ms) */
    float mtf_chunk_ms;    /* K2c alone (the chunk-parallel move-to-front inside huffman_ms) */
} ibz2_timings;

typedef struct ibz2_ctx ibz2_ctx;

/* ---- context: one per GPU; calls on one context must be serialised by the caller ---- */
int ibz2_abi_version( void );
int ibz2_device_count( void ); /* 0 when no CUDA device is usable */
int ibz2_create( int device, ibz2_ctx** ctx );
void ibz2_destroy( ibz2_ctx* ctx );
const char* ibz2_last_error( const ibz2_ctx* ctx ); /* ctx may be NULL: error of a failed ibz2_create */
int ibz2_get_timings( const ibz2_ctx* ctx, ibz2_timings* timings );
/* Upper bound on the data blocks one ibz2_decode_blocks launch wave holds in flight (scratch is sized for it). */
int ibz2_set_batch_blocks( ibz2_ctx* ctx, uint32_t blocks );
/* Stream CRC (BZ2Reader.hpp:406-416, 481-484): ibz2_decode_blocks / ibz2_decode_buffer fold the block CRCs of every stream
 * that lies completely inside the call and, when verify != 0 (the default), give the LAST block of a stream whose fold
 * differs from the CRC field of its end-of-stream block the status IBZ2_ST_STREAM_CRC (its bytes are still delivered).
 * With verify == 0 only stream_flags / stream_crc_expected / stream_crc_calculated are filled in. */
int ibz2_set_verify_stream_crc( ibz2_ctx* ctx, int verify );
/* The same fold over results of several calls laid end to end (blocks of one file in file order, e.g. the shares of
 * several GPUs): fills stream_crc_calculated / IBZ2_STREAM_CRC_CHECKED and, when verify != 0, sets IBZ2_ST_STREAM_CRC.
 * Returns the number of streams whose CRC did not match.  Host-side, no GPU work. */
uint64_t ibz2_fold_stream_crcs( ibz2_block_result* results, uint64_t nblocks, int verify );

/* ---- device memory helpers (so that FFI callers can keep inputs/outputs resident in HBM) ---- */
int ibz2_device_alloc( ibz2_ctx* ctx, uint64_t nbytes, void** dptr );
int ibz2_device_free( ibz2_ctx* ctx, void* dptr );
int ibz2_host_alloc_pinned( ibz2_ctx* ctx, uint64_t nbytes, void** hptr );
int ibz2_host_free_pinned( ibz2_ctx* ctx, void* hptr );
int ibz2_memcpy_h2d( ibz2_ctx* ctx, void* dptr, const void* hptr, uint64_t nbytes );
int ibz2_memcpy_d2h( ibz2_ctx* ctx, void* hptr, const void* dptr, uint64_t nbytes );
/* the same as ibz2_memcpy_h2d for PAGEABLE host memory (a memory-mapped file): host threads copy into two pinned staging
 * buffers that the copy engine empties in turn.  Stands where the reference reads the file through its SharedFileReader
 * (core/filereader/Shared.hpp:267-300): the compressed file goes to the GPU once, when a reader opens it. */
int ibz2_upload_pageable( ibz2_ctx* ctx, void* dptr, const void* hptr, uint64_t nbytes );
int ibz2_synchronize( ibz2_ctx* ctx );
/* device -> PINNED host without waiting, on a stream of its own; ibz2_copy_synchronize waits for the copies queued so
 * far.  For consumers that keep decoded blocks in HBM and hand out the bytes a read asks for (host/BlockSource.hpp). */
int ibz2_memcpy_d2h_async( ibz2_ctx* ctx, void* hptr, const void* dptr, uint64_t nbytes );
int ibz2_copy_synchronize( ibz2_ctx* ctx );

/* ---- K1: magic finder ----
 * Reports every bit offset p (MSB-first numbering, p + 48 <= 8 * nbytes) at which one of the selected 48-bit magics
 * starts, ascending, as base_bit_offset + p, with IBZ2_OFFSET_EOS_FLAG or-ed in for EOS magics.
 * offsets is a HOST array of `capacity` entries; *count receives the number found (IBZ2_E_CAPACITY if larger). */
int ibz2_find_magics( ibz2_ctx* ctx, const void* data, uint64_t nbytes, int data_mem, uint64_t base_bit_offset,
                      uint32_t which, uint64_t* offsets, uint64_t capacity, uint64_t* count );

/* ---- K1 without allocation ----
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same kernel and the
 * same result as ibz2_find_magics, without a context: the match list and its counter live in a device array of fixed
 * size that is part of the library image (512.25 KB on every GPU that loads the library, until the process exits), so the
 * call allocates nothing, neither on the device nor through a library.
 *   device        CUDA device that holds device_data (made current for the call, the caller's device is restored)
 *   stream        a cudaStream_t of that device, or NULL for the default stream; the call waits for it before returning
 *   device_data   DEVICE memory only (the call stages nothing); nbytes bytes
 *   offsets       HOST array of `capacity` entries, capacity <= IBZ2_NOALLOC_MAX_MATCHES: a larger capacity is refused
 *                 with IBZ2_E_CAPACITY before anything runs (an 8 GB `bzip2 -9` file has ~9100 magics)
 *   *count        the number of magics found; IBZ2_E_CAPACITY if it is larger than capacity (offsets then holds
 *                 `capacity` of them in ascending order, not necessarily the first ones)
 * All calls on one device share the array: calls that overlap on the GPU (two streams, two threads) race and must be
 * serialised by the caller.  The first call on a device resolves the array's address (cudaGetSymbolAddress); make it
 * outside a stream capture.  ibz2_last_error( NULL ) describes a failure on the calling thread. */
#define IBZ2_NOALLOC_MAX_MATCHES 65536U
int ibz2_find_magics_noalloc( int device, void* stream, const void* device_data, uint64_t nbytes, uint64_t base_bit_offset,
                              uint32_t which, uint64_t* offsets, uint64_t capacity, uint64_t* count );

/* ---- K2..K6: decode a batch of blocks ----
 * data/nbytes: the compressed file (or a window of it); block_bit_offsets: HOST array of nblocks bit offsets into
 * data at which a block (or EOS) magic is expected.  Decoded bytes of block i go to out + results[i].out_offset;
 * blocks are laid out back to back in the order given.  out may be NULL (sizes, CRCs and statuses only).
 * results: HOST array of nblocks entries.  *total_out receives the sum of out_size.  Blocks that do not fit get
 * IBZ2_ST_OUTPUT_CAPACITY.  Block-level failures do not fail the call. */
int ibz2_decode_blocks( ibz2_ctx* ctx, const void* data, uint64_t nbytes, int data_mem,
                        const uint64_t* block_bit_offsets, uint64_t nblocks,
                        void* out, uint64_t out_capacity, int out_mem,
                        ibz2_block_result* results, uint64_t* total_out );

/* ---- K1 + K2..K6 in one call: what ParallelBZ2Reader does for a file that is in memory ----
 * Copies `data` to the GPU once (host input), finds every data-block magic in it, decodes the blocks whose magic
 * starts in [first_bit, last_bit) (pass 0 and UINT64_MAX for all; a rank of a multi-GPU job passes its share) in
 * file order and returns their bytes back to back in `out`.  results: HOST array of results_capacity entries;
 * *nblocks receives the number of blocks decoded (IBZ2_E_CAPACITY if results is too small).
 * Stands in for the BlockFinder thread + BlockFetcher pool of ParallelBZ2Reader::read
 * (indexed_bzip2/ParallelBZ2Reader.hpp:167-269, core/BlockFinder.hpp:164-197, core/BlockFetcher.hpp:244-317). */
int ibz2_decode_buffer( ibz2_ctx* ctx, const void* data, uint64_t nbytes, int data_mem,
                        uint64_t first_bit, uint64_t last_bit,
                        void* out, uint64_t out_capacity, int out_mem,
                        ibz2_block_result* results, uint64_t results_capacity, uint64_t* nblocks,
                        uint64_t* total_out );

/* ---- one file on several GPUs of one machine, in one process ----
 * Stands in for the pool of decode threads inside ONE reference reader object (indexed_bzip2/ParallelBZ2Reader.hpp:50-69,
 * core/BlockFetcher.hpp:176-190), with devices where the reference has threads: the compressed stream is cut into equal
 * byte ranges, device i finds and decodes the blocks whose magic starts in range i (one host thread and one ibz2_ctx per
 * device), the byte counts of the devices are scanned on the host and every device copies its bytes to its place in the
 * caller's ONE output buffer.  No collective: blocks are independent.  Block results come back in file order with
 * file-relative enc_offset_bits / out_offset, and the stream CRCs are folded over all of them (ibz2_fold_stream_crcs,
 * verdict included).  `devices` may name one device several times (two contexts on one GPU: how the tests run it). */
typedef struct ibz2_multi ibz2_multi;
typedef struct ibz2_multi_timings {
    float decode_ms;  /* wall clock: until the slowest device has decoded its share into its own memory */
    float copy_ms;    /* wall clock: the device -> host copies to the caller's buffer, all devices at once */
    float total_ms;   /* wall clock of the call */
    float reserved;
    uint64_t stream_crc_mismatches;
} ibz2_multi_timings;
int ibz2_multi_create( const int* devices, int ndevices, ibz2_multi** multi );
void ibz2_multi_destroy( ibz2_multi* multi );
const char* ibz2_multi_last_error( const ibz2_multi* multi );
int ibz2_multi_device_count( const ibz2_multi* multi );
/* pinned host memory that every device of the set can copy into (cudaHostAllocPortable) */
int ibz2_multi_host_alloc_pinned( ibz2_multi* multi, uint64_t nbytes, void** hptr );
int ibz2_multi_host_free_pinned( ibz2_multi* multi, void* hptr );
/* data, out, results: HOST memory (pinned for full PCIe speed).  Otherwise like ibz2_decode_buffer( 0, UINT64_MAX ). */
int ibz2_multi_decode_buffer( ibz2_multi* multi, const void* data, uint64_t nbytes, void* out, uint64_t out_capacity,
                              ibz2_block_result* results, uint64_t results_capacity, uint64_t* nblocks,
                              uint64_t* total_out );
/* of the last call; per_device: `capacity` entries, one ibz2_timings per device in the order given to create */
int ibz2_multi_get_timings( const ibz2_multi* multi, ibz2_multi_timings* timings, ibz2_timings* per_device, int capacity );

/* CUDA-event stopwatch on the context's stream: brackets any number of calls (device time, not wall clock). */
int ibz2_timer_start( ibz2_ctx* ctx );
int ibz2_timer_stop( ibz2_ctx* ctx, float* milliseconds );

/* ---- test/diagnostic hook: stage-level intermediates of ONE block, for parity tests against the oracle ----
 * Any of the host output pointers may be NULL.  lbytes: n_symbols bytes (the BWT last column, bzip2.hpp:759,788);
 * tvec: n_symbols u32, the packed (T << 8) | L vector after prepare() (bzip2.hpp:827-831); prerle: n_symbols bytes,
 * the traversal output before RLE1 undo (bzip2.hpp:876-879); out: out_capacity bytes. */
int ibz2_debug_decode_block( ibz2_ctx* ctx, const void* data, uint64_t nbytes, uint64_t bit_offset,
                             ibz2_block_result* result, uint8_t* lbytes, uint32_t* tvec, uint8_t* prerle,
                             uint8_t* out, uint64_t out_capacity );

/* ---- host-side header probes (no GPU work; a few dozen bits each) ---- */
/* Returns the level 1..9, 0 for a bad header (domain_error), -1 when the file ends first. *bit_pos advances. */
int ibz2_read_stream_header( const uint8_t* data, uint64_t nbytes, uint64_t* bit_pos );
/* Parses magic + CRC (+ EOS padding). Fills enc_offset_bits, enc_size_bits (EOS only), expected_crc, is_eos, is_eof,
 * status (IBZ2_ST_OK, IBZ2_ST_BAD_MAGIC or IBZ2_ST_EOF). */
int ibz2_read_block_header( const uint8_t* data, uint64_t nbytes, uint64_t bit_offset, ibz2_block_result* result );

uint32_t ibz2_stream_crc_fold( uint32_t stream_crc, uint32_t block_crc );

#ifdef __cplusplus
}
#endif
#endif /* IBZ2_CUDA_H */
