/*
 * common.cuh -- shared device-side definitions for the sm_100a bzip2 decode kernels.
 *
 * Data layout in HBM (per launch wave of B blocks, see DESIGN.md "Data layout"):
 *   meta[B]                      BlockMeta, one per block (inputs: bit offset; outputs of every stage)
 *   lbytes[B][LBYTES_STRIDE]     u8   BWT last column L written by K2; re-used as the pre-RLE1 byte string by K4b
 *   dbuf[B][DBUF_STRIDE]         u32  the reference's packed vector (T[k] << 8) | L[k] (bzip2.hpp:827-831) written by
 *                                     K4a; re-used by K5a for the per-element output offsets
 *   selectors[B][SEL_STRIDE]     u8   Huffman table selectors of the block (bzip2.hpp:574-637)
 *   warpbase[B][32]              u32  K5a: output offset of each of the 32 element ranges of a block
 */
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/ibz2_cuda.h"

namespace ibz2
{
constexpr uint64_t MAGIC_BLOCK = IBZ2_MAGIC_BLOCK;
constexpr uint64_t MAGIC_EOS = IBZ2_MAGIC_EOS;
constexpr uint32_t DBUF_SIZE = IBZ2_MAX_BLOCK_SYMBOLS;  /* bzip2.hpp:414-416 */
constexpr uint32_t MAX_GROUPS = 6;                      /* bzip2.hpp:95 */
constexpr uint32_t GROUP_SIZE = 50;                     /* bzip2.hpp:96 */
constexpr uint32_t MAX_CODE_LEN = 20;                   /* bzip2.hpp:97 */
constexpr uint32_t MAX_SYMBOLS = 258;                   /* bzip2.hpp:98 */
constexpr uint32_t MAX_SELECTORS = 32768;               /* bzip2.hpp:451 */

constexpr uint32_t LBYTES_STRIDE = 900096;              /* >= DBUF_SIZE, multiple of 128 */
constexpr uint32_t DBUF_STRIDE = 900096;                /* u32 entries */
constexpr uint32_t SEL_STRIDE = MAX_SELECTORS + 64;
constexpr uint32_t RLE_RANGES = 32;                     /* element ranges per block in K5a/K5b (one per warp) */

/* Internal block status between the K2 kernels: "decode this block again with the exact state machine". */
constexpr uint32_t ST_INTERNAL_CAREFUL = 255;

/* One per block, device-resident; the host mirrors it to fill ibz2_block_result. */
struct BlockMeta
{
    uint64_t bitOffset;   /* in: where the 48-bit magic starts */
    uint64_t endBit;      /* K2: first bit after the block (after the EOB symbol / after the EOS padding) */
    uint64_t outOffset;   /* scan: byte offset of the block's output inside the call's output buffer */
    uint32_t status;      /* ibz2_block_status */
    uint32_t headerCrc;
    uint32_t origPtr;
    uint32_t n;           /* dbufCount */
    uint32_t outSize;     /* K5a: decoded bytes */
    uint32_t crcAcc;      /* K6: xor-accumulated raw CRC contributions */
    uint32_t calcCrc;     /* K6 finalize */
    uint32_t isEos;
    uint32_t isEof;
    uint32_t cycleLen;    /* K4b: length of the permutation cycle through origPtr (diagnostic) */
    uint32_t writeOutput; /* scan: 1 when the block's bytes fit the output buffer */
    uint32_t nCoded;      /* K2b: Huffman symbols stored for K2c..K2e (run symbols and MTF symbols, without EOB) */
    int32_t travStar;     /* K4b walk: the segment that passes through origPtr (-1: none), set to -1 by K2a */
    uint32_t travK;       /* K4b walk: number of steps from that segment's start to origPtr */
    uint32_t streamFlags; /* stream probe: STREAM_FIRST_BLOCK | STREAM_LAST_BLOCK */
    uint32_t eosCrc;      /* stream probe: CRC field of the end-of-stream block that follows the block directly */
};
constexpr uint32_t STREAM_FIRST_BLOCK = IBZ2_STREAM_FIRST_BLOCK;   /* a stream header stands directly in front of the block */
constexpr uint32_t STREAM_LAST_BLOCK = IBZ2_STREAM_LAST_BLOCK;     /* an end-of-stream block follows the block directly */

/* ---- K2 intermediates (see huffman_header.cu / huffman_bits.cu / mtf.cu) ----
 * K2a writes one BlockTables per block; K2b reads it (the LUTs go to shared memory, the rest stays in L1/L2). */
constexpr int HUFF_LUT_BITS = 9;                       /* direct-lookup width of K2b */
constexpr int HUFF_LUT_SIZE = 1 << HUFF_LUT_BITS;
constexpr uint32_t HUFF_LUT_WORDS = MAX_GROUPS * HUFF_LUT_SIZE / 2;   /* u32 words per block: 1536 */

struct alignas( 16 ) BlockTables
{
    uint16_t lut[MAX_GROUPS][HUFF_LUT_SIZE];    /* { symbol:11, length:5 }, 0 = code longer than HUFF_LUT_BITS or none */
    uint16_t perm[MAX_GROUPS][260];             /* symbols sorted by (length, symbol) */
    uint32_t first[MAX_GROUPS][MAX_CODE_LEN + 2];
    uint16_t count[MAX_GROUPS][MAX_CODE_LEN + 2];
    uint16_t offset[MAX_GROUPS][MAX_CODE_LEN + 2];
    uint32_t maxLen[MAX_GROUPS + 2];
    uint64_t limit[MAX_GROUPS][MAX_CODE_LEN + 2];   /* ( first + count ) << ( 32 - len ): codes of <= len bits are below it */
    int32_t base[MAX_GROUPS][MAX_CODE_LEN + 2];     /* offset - first: perm index of a code of len bits = code + base */
    uint8_t s2b[256];                           /* symbolToByte (bzip2.hpp:526-571) */
    uint64_t dataStartBit;                      /* first bit of the Huffman-coded symbols */
    uint32_t symbolCount;
    uint32_t selectorsCount;
    uint32_t groupCount;
    uint32_t pad;
    uint16_t eobPerm[MAX_GROUPS + 2];           /* perm index of the end-of-block symbol (the last one of its length) */
};

/* The symbol stream of a block is cut into chunks of MTF_CHUNK coded symbols; K2b leaves one checkpoint per chunk. */
constexpr uint32_t MTF_CHUNK = 1024;
constexpr uint32_t MTF_CHUNK_SHIFT = 10;
constexpr uint32_t MAX_CHUNKS = ( DBUF_SIZE + MTF_CHUNK - 1 ) / MTF_CHUNK;   /* 879 */
struct ChunkCheckpoint
{
    uint32_t outOffset;   /* index in L of the first byte the chunk's first symbol produces */
    uint32_t runShift;    /* number of run symbols directly in front of the chunk (position inside a RUNA/RUNB number) */
};

/* K2 scratch aliases the block's dbuf slot (3.6 MB), which K4a only overwrites after K2e has consumed it:
 *   u16 symbols[900096] | u8 symbolic[900096] | u8 lists[MAX_CHUNKS + 1][256] | ChunkCheckpoint[MAX_CHUNKS + 1] */
constexpr uint64_t K2_SYM16_OFFSET = 0;
constexpr uint64_t K2_SYM8_OFFSET = static_cast<uint64_t>( LBYTES_STRIDE ) * 2U;
constexpr uint64_t K2_LISTS_OFFSET = K2_SYM8_OFFSET + LBYTES_STRIDE;
constexpr uint64_t K2_CKPT_OFFSET = K2_LISTS_OFFSET + static_cast<uint64_t>( MAX_CHUNKS + 1U ) * 256U;
static_assert( sizeof( BlockTables ) % 16U == 0, "K2b copies the LUTs with 16-byte loads" );
static_assert( K2_CKPT_OFFSET + ( MAX_CHUNKS + 1U ) * sizeof( ChunkCheckpoint ) <= static_cast<uint64_t>( DBUF_STRIDE ) * 4U,
               "K2 scratch must fit the dbuf slot" );

__device__ __forceinline__ uint32_t
laneId()
{
    return threadIdx.x & 31U;
}

/* Big-endian 32-bit word i of the byte stream (bytes past the end read as zero). */
__device__ __forceinline__ uint32_t
loadWordBE( const uint8_t* __restrict__ data, uint64_t nbytes, uint64_t wordIndex )
{
    const uint64_t byte = wordIndex * 4U;
    uint32_t w = 0;
    if ( byte + 4U <= nbytes ) {
        w = __ldg( reinterpret_cast<const uint32_t*>( data ) + wordIndex );
    } else {
        for ( uint32_t k = 0; k < 4U; ++k ) {
            if ( byte + k < nbytes ) {
                w |= static_cast<uint32_t>( data[byte + k] ) << ( 8U * k );
            }
        }
    }
    return __byte_perm( w, 0, 0x0123 );
}

/* ---- CRC-32/BZIP2 arithmetic in GF(2)[x] / 0x104C11DB7 (bzip2.hpp:59-91) ---- */
constexpr uint32_t CRC_POLY = 0x04C11DB7U;

/* a * b mod P, polynomials with bit 31 = x^31 */
__host__ __device__ __forceinline__ uint32_t
crcMulMod( uint32_t a, uint32_t b )
{
    uint32_t r = 0;
#pragma unroll 4
    for ( int i = 31; i >= 0; --i ) {
        r = ( r << 1 ) ^ ( ( r & 0x80000000U ) ? CRC_POLY : 0U );
        if ( ( b >> i ) & 1U ) {
            r ^= a;
        }
    }
    return r;
}

/* launch wrappers (one per .cu file); all asynchronous on `stream` */
void launchFindMagics( const uint8_t* data, uint64_t nbytes, uint32_t which, uint64_t* found, uint32_t* foundCount,
                       uint32_t capacity, cudaStream_t stream );

/* K2a..K2e; tables: BlockTables[nblocks]; scratch: the dbuf slots; counter: device u32[16 + nblocks]: [0] = K2b's work
 * counter, [16 + i] = the block K2b's i-th request gets (longest compressed block first) when `ordered` is set. */
constexpr uint32_t K2B_ORDER_OFFSET = 16;
void launchHuffmanDecode( const uint8_t* data, uint64_t nbytes, BlockMeta* meta, uint32_t nblocks,
                          uint8_t* lbytes, uint8_t* selectors, BlockTables* tables, uint32_t* scratch,
                          uint32_t* counter, bool ordered, cudaStream_t stream,
                          cudaEvent_t bitsStart = nullptr, cudaEvent_t bitsStop = nullptr,   /* recorded around K2b + resolve */
                          cudaEvent_t chunkStop = nullptr );                                  /* ... and behind K2c */
constexpr uint32_t HUFFMAN_LAUNCHES = 8;   /* K2a, K2b, resolve, K2c, K2d, K2e, K2x, stream probe */

void launchHuffmanHeader( const uint8_t* data, uint64_t nbytes, BlockMeta* meta, uint32_t nblocks,
                          uint8_t* selectors, BlockTables* tables, cudaStream_t stream );
void launchHuffmanBits( const uint8_t* data, uint64_t nbytes, BlockMeta* meta, uint32_t nblocks,
                        const uint8_t* selectors, const BlockTables* tables, uint32_t* scratch, uint32_t* counter,
                        bool ordered, cudaStream_t stream );
void launchHuffmanCareful( const uint8_t* data, uint64_t nbytes, BlockMeta* meta, uint32_t nblocks, uint8_t* lbytes,
                           uint8_t* selectors, cudaStream_t stream );
void launchMtf( BlockMeta* meta, uint32_t nblocks, const BlockTables* tables, uint32_t* scratch, uint8_t* lbytes,
                cudaStream_t stream, cudaEvent_t chunkStop = nullptr );

void launchBwtPrepare( const uint8_t* lbytes, uint32_t* dbuf, BlockMeta* meta, uint32_t nblocks,
                       cudaStream_t stream );

/* K4b staging, per block: one slot of TRAV_SLOT_BYTES per segment (the first bytes the segment's walk met), then per
 * segment its length and successor segment (one uint2: the ranking follows chains of them, one load per hop; kept
 * apart from the slots so that a block's 56 250 links are 450 KB that stay in L2), the position at which a walk longer
 * than the slot goes on (u32) and, after ranking, its output offset (u32).  The buffer starts with a TRAV_STAGE_HEADER holding the walk kernel's work counter. */
#ifndef IBZ2_TRAV_SPACING_SHIFT
#define IBZ2_TRAV_SPACING_SHIFT 4
#endif
#ifndef IBZ2_TRAV_SLOT_SHIFT
#define IBZ2_TRAV_SLOT_SHIFT 6
#endif
constexpr uint32_t TRAV_SPACING_SHIFT = IBZ2_TRAV_SPACING_SHIFT;
constexpr uint32_t TRAV_SPACING = 1U << TRAV_SPACING_SHIFT;                    /* a splitter every 16 positions */
constexpr uint32_t TRAV_COARSE_SHIFT = 6;                                      /* every 64: the ones that get ranked */
constexpr uint32_t TRAV_MAX_SEGMENTS = ( DBUF_SIZE + TRAV_SPACING - 1U ) / TRAV_SPACING;   /* 56250 */
constexpr uint32_t TRAV_SEG_CAP = ( TRAV_MAX_SEGMENTS + 63U ) & ~63U;          /* 56256 */
constexpr uint32_t TRAV_SLOT_SHIFT = IBZ2_TRAV_SLOT_SHIFT;
constexpr uint32_t TRAV_SLOT_BYTES = 1U << TRAV_SLOT_SHIFT;                    /* 4 x the mean segment length: two slots per line */
constexpr uint32_t TRAV_SLOT2_SHIFT = 7;                                       /* where the 1.8 % of longer segments go on */
constexpr uint32_t TRAV_SLOT2_BYTES = 1U << TRAV_SLOT2_SHIFT;
constexpr uint64_t TRAV_SLOT2_OFFSET = static_cast<uint64_t>( TRAV_SEG_CAP ) * TRAV_SLOT_BYTES;
constexpr uint64_t TRAV_LINK_OFFSET = TRAV_SLOT2_OFFSET + static_cast<uint64_t>( TRAV_SEG_CAP ) * TRAV_SLOT2_BYTES;   /* uint2: length, successor */
constexpr uint64_t TRAV_OVF_OFFSET = TRAV_LINK_OFFSET + TRAV_SEG_CAP * 8ULL;
constexpr uint64_t TRAV_BASE_OFFSET = TRAV_OVF_OFFSET + TRAV_SEG_CAP * 4ULL;    /* u32: output offset of the segment's first byte */
constexpr uint64_t TRAV_STAGE_STRIDE = ( TRAV_BASE_OFFSET + TRAV_SEG_CAP * 4ULL + 255ULL ) & ~255ULL;
constexpr uint64_t TRAV_STAGE_HEADER = 256;
/* stage: device u8[TRAV_STAGE_HEADER + nblocks * TRAV_STAGE_STRIDE] */
void launchBwtTraverse( const uint32_t* dbuf, uint8_t* prerle, BlockMeta* meta, uint32_t nblocks, uint8_t* stage,
                        cudaStream_t stream, cudaEvent_t walkStop = nullptr );   /* recorded after the walk kernel */

void launchRleScan( const uint8_t* prerle, uint32_t* offsets, uint32_t* warpBase, BlockMeta* meta,
                    uint32_t nblocks, cudaStream_t stream );

/* running: device u64[1] = next output offset, carried across the waves of one call.
 * pieceStart: device u32[nblocks + 1], first CRC window of each block (windows of pieceBytes bytes). */
void launchOutputLayout( BlockMeta* meta, uint32_t nblocks, uint64_t* running, uint64_t outCapacity,
                         int haveOutput, uint32_t* pieceStart, uint32_t pieceBytes, cudaStream_t stream );

void launchRleWrite( const uint8_t* prerle, const uint32_t* offsets, const uint32_t* warpBase,
                     const BlockMeta* meta, uint32_t nblocks, uint8_t* out, cudaStream_t stream );

void launchCrc( const uint8_t* out, BlockMeta* meta, uint32_t nblocks, const uint32_t* pieceStart,
                uint32_t pieceBytes, cudaStream_t stream );

void launchCrcFinalize( const uint8_t* out, BlockMeta* meta, uint32_t nblocks, cudaStream_t stream );

/* Where the streams of the file begin and end among the decoded blocks (BZ2Reader.hpp:406-416 needs the CRC field of the
 * end-of-stream block behind a stream's last block); runs once K2 has set every block's endBit. */
void launchStreamProbe( const uint8_t* data, uint64_t nbytes, BlockMeta* meta, uint32_t nblocks, cudaStream_t stream );

constexpr uint32_t CRC_PIECE_BYTES = 131072;

void initCrcConstants();

int smCount();
void setContextFreeError( const char* message );
}  // namespace ibz2
