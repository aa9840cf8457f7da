/*
 * K2b -- the Huffman symbol loop of Block::readBlockData (indexed_bzip2/bzip2.hpp:709-723 and :766-768, with
 * huffman/HuffmanCodingShortBitsCached.hpp:98-150), one LANE per bzip2 block.
 *
 * The bit stream of a block is a dependent chain (window -> table -> code length -> shift); nothing inside one block can
 * shorten it, so (a) the chip holds as many independent chains as possible and (b) the chain carries nothing but the
 * code LENGTH:
 *   - a decoder keeps six 9-bit tables of u8 code lengths in shared memory (3 KB, contiguous per decoder, so the address
 *     of an entry is "table base + top nine bits": one LEA.HI away from the window) plus the 0.4 KB of per-length limits
 *     that codes of 10..20 bits need: 3.5 KB per decoder, 32 decoders per CTA, two CTAs per SM = 64 chains per SM,
 *     9472 per chip, i.e. every block of the 8 GB workload decodes in ONE round;
 *   - what a decoder stores per symbol is the table entry it hit, ( table << 9 ) | index (or 0x8000 | table << 9 |
 *     rank for a long code); the symbol itself is not needed to go on, and resolveSymbolsKernel turns the entries into
 *     symbols afterwards at memory speed;
 *   - the window is two big-endian words A:B and a bit position; three symbols are decoded per refill test (a symbol of
 *     the table takes at most 9 bits and the position is below 32 when a triple starts), each one a shift of the first
 *     symbol's 32 window bits.  A table entry of 0 (code longer than 9 bits, end-of-block symbol, no code) has length 0:
 *     the window stays where it is, the rest of the triple misses too, and one test per triple finds it;
 *   - decoders pull block indices from a global counter (longest compressed block first); the unit of lockstep inside a
 *     warp is one GROUP of 50 symbols (the table switch of bzip2.hpp:710-721).
 * The dbufCount / RUNA-RUNB bookkeeping of bzip2.hpp:726-790 is chunk-parallel work and happens in K2c..K2e (mtf.cu).
 * A block leaves this kernel as "clean" (status OK, nCoded symbols stored, endBit set) only when its symbol stream ends
 * with the end-of-block symbol inside the file.  Anything else -- a bit pattern without code, more groups than
 * selectors, the end of the file, more symbols than a block can hold, an empty symbol map -- marks the block
 * ST_INTERNAL_CAREFUL and the exact state machine in huffman_careful.cu decides what the reference would have thrown.
 *
 * Algorithmic HBM traffic per block: C compressed bytes read, one entry per coded symbol written.
 */
#include "common.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <type_traits>

namespace ibz2
{
namespace
{
constexpr int BITS_SLOTS = 32;          /* decoders per CTA */
constexpr int BITS_CTAS_PER_SM = 2;
constexpr int BITS_LANES_DEFAULT = 4;   /* active lanes per warp; BITS_SLOTS / lanes warps per CTA */
/* experiment switches (the defaults are what the measurements in DESIGN.md chose) */
#ifndef IBZ2_K2B_TOUCH
#define IBZ2_K2B_TOUCH 0   /* measured: 63.2 ms with the touches, 61.9 without */
#endif
#ifndef IBZ2_K2B_PREFETCH_WORDS
#define IBZ2_K2B_PREFETCH_WORDS 40
#endif
constexpr uint32_t SYM_LIMIT = DBUF_SIZE;                         /* a clean block has at most this many symbols */
constexpr uint32_t GROUP_MAX_WORDS = ( GROUP_SIZE * MAX_CODE_LEN + 31U ) / 32U + 4U + 96U;   /* + prefetch distance */

/* ---- shared memory of one decoder ---- */
constexpr uint32_t LONG_FIRST_LEN = HUFF_LUT_BITS;                          /* limits are kept for lengths 9..20 */
constexpr uint32_t LONG_LIMITS = MAX_CODE_LEN - LONG_FIRST_LEN + 1U;        /* 12 */
constexpr uint32_t LONG_OFFSETS = MAX_CODE_LEN - LONG_FIRST_LEN;            /* lengths 10..20: 11 */
constexpr uint32_t DEC_LUT_BYTES = MAX_GROUPS * HUFF_LUT_SIZE;              /* u8 [6][512] code lengths */
constexpr uint32_t DEC_LIM_OFFSET = DEC_LUT_BYTES;                          /* u32 [6][12], 20-bit left-aligned code limits */
constexpr uint32_t DEC_LIM_BYTES = MAX_GROUPS * LONG_LIMITS * 4U;
constexpr uint32_t DEC_OFF_OFFSET = DEC_LIM_OFFSET + DEC_LIM_BYTES;         /* u16 [6][11], perm offset of each length */
constexpr uint32_t DEC_OFF_BYTES = MAX_GROUPS * LONG_OFFSETS * 2U;
constexpr uint32_t DEC_EOB_OFFSET = DEC_OFF_OFFSET + DEC_OFF_BYTES;         /* u16 [6], perm index of the end-of-block symbol */
/* an odd number of words: the same entry of neighbouring decoders sits in different banks */
constexpr uint32_t DEC_STRIDE = ( ( DEC_EOB_OFFSET + MAX_GROUPS * 2U + 3U ) / 4U | 1U ) * 4U;
constexpr uint32_t SH_TOTAL_BYTES = BITS_SLOTS * DEC_STRIDE;
static_assert( BITS_CTAS_PER_SM * ( SH_TOTAL_BYTES + 1024U ) <= 228U * 1024U, "two CTAs per SM" );
static_assert( DEC_LUT_BYTES <= 0x8000U, "stored entries keep bit 15 for long codes" );

#ifndef IBZ2_K2B_PREFETCH_KIND
#define IBZ2_K2B_PREFETCH_KIND 1   /* 0 = none, 1 = into L1, 2 = into L2 */
#endif
__device__ __forceinline__ void
prefetchStream( const void* p )
{
#if IBZ2_K2B_PREFETCH_KIND == 1
    asm volatile( "prefetch.global.L1 [%0];" ::"l"( p ) );
#elif IBZ2_K2B_PREFETCH_KIND == 2
    asm volatile( "prefetch.global.L2 [%0];" ::"l"( p ) );
#else
    (void) p;
#endif
}

__device__ __forceinline__ uint32_t
ldsU8( uint32_t addr )
{
    uint32_t v;   /* a wider destination is zero-extended: no separate mask on the dependent chain */
    asm volatile( "ld.shared.u8 %0, [%1];" : "=r"( v ) : "r"( addr ) );
    return v;
}

__device__ __forceinline__ uint32_t
ldsU16( uint32_t addr )
{
    uint32_t v;
    asm volatile( "ld.shared.u16 %0, [%1];" : "=r"( v ) : "r"( addr ) );
    return v;
}

__device__ __forceinline__ uint32_t
ldsU32( uint32_t addr )
{
    uint32_t v;
    asm volatile( "ld.shared.u32 %0, [%1];" : "=r"( v ) : "r"( addr ) );
    return v;
}

/* 32-bit word i of the byte stream in memory order (bytes past the end read as zero) */
__device__ __forceinline__ uint32_t
loadWordRaw( const uint8_t* __restrict__ data, uint64_t nbytes, uint64_t wordIndex )
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
    return w;
}

/* A load whose only purpose is to bring its sector into L1 (prefetch.global.L1 does not: with it 1 load in 8, the first
 * of every sector, still missed).  The value is kept alive by the caller until long after it has arrived, so that the
 * register is not handed to something that would then wait for the load. */
__device__ __forceinline__ uint32_t
touch( const uint32_t* p )
{
    uint32_t v;
    asm volatile( "ld.global.nc.u32 %0, [%1];" : "=r"( v ) : "l"( p ) );
    return v;
}

__device__ __forceinline__ uint32_t
bswap( uint32_t raw )
{
    return __byte_perm( raw, 0, 0x0123 );
}

/* A code of more than 9 bits (HuffmanCodingShortBitsCached.hpp:117-150), entirely from shared memory and without the
 * symbol itself: canonical codes grow with their length when left-aligned, so the length is 10 + the number of per-length
 * limits the window has reached, and the code's rank among all codes (its index into perm) follows from the limit below
 * it.  Needs 20 valid bits in hi.  Returns ( permIndex << 5 ) | length, 0 when no code matches. */
__device__ __noinline__ uint32_t
decodeLongShared( uint32_t limAddr, uint32_t offAddr, uint32_t hi )
{
    const uint32_t h20 = hi >> 12;
    uint32_t lim[LONG_LIMITS];
#pragma unroll
    for ( uint32_t i = 0; i < LONG_LIMITS; ++i ) {
        lim[i] = ldsU32( limAddr + i * 4U );
    }
    uint32_t len = LONG_FIRST_LEN + 1U;
    uint32_t lower = lim[0];
#pragma unroll
    for ( uint32_t i = 1; i + 1U < LONG_LIMITS; ++i ) {
        const bool ge = h20 >= lim[i];
        len += ge ? 1U : 0U;
        lower = ge ? lim[i] : lower;
    }
    if ( h20 >= lim[LONG_LIMITS - 1U] ) {
        return 0;
    }
    const uint32_t rel = ( h20 - lower ) >> ( MAX_CODE_LEN - len );
    const uint32_t idx = ldsU16( offAddr + ( len - LONG_FIRST_LEN - 1U ) * 2U ) + rel;
    return ( idx << 5 ) | len;
}

/* The bit position of a decoder: A and B are the big-endian words wordIdx and wordIdx + 1 of the file, craw is word
 * wordIdx + 2 in memory order (swapped when it moves up, so that nothing waits for its load), pos < 32 between symbols. */
struct Decoder
{
    uint32_t A;
    uint32_t B;
    uint32_t craw;
    uint32_t pos;
    uint64_t wordIdx;
};

enum : uint32_t { SYM_PLAIN = 0, SYM_EOB = 1, SYM_FAIL = 2 };

/* One symbol the careful way: any code length, the end-of-block symbol, no code at all.  The window must have been
 * refilled (pos < 32).  *entry = what K2b stores for the symbol. */
__device__ __forceinline__ uint32_t
decodeOneSymbol( uint32_t A, uint32_t B, uint32_t& pos, uint32_t lutAddr, uint32_t table, const BlockTables* __restrict__ bt,
                 uint32_t* entry )
{
    const uint32_t tableAddr = lutAddr + table * HUFF_LUT_SIZE;
    const uint32_t hi = __funnelshift_l( B, A, pos );
    const uint32_t idx = hi >> ( 32 - HUFF_LUT_BITS );
    const uint32_t len = ldsU8( tableAddr + idx );
    if ( len != 0 ) {
        *entry = ( table << HUFF_LUT_BITS ) | idx;
        pos += len;
        return SYM_PLAIN;
    }
    const uint32_t limAddr = lutAddr + DEC_LIM_OFFSET + table * ( LONG_LIMITS * 4U );
    if ( ( hi >> 12 ) < ldsU32( limAddr ) ) {
        /* a code of at most 9 bits that is not in the shared-memory table: the end-of-block symbol (bzip2.hpp:766-768) */
        pos += bt->lut[table][idx] & 31U;
        return SYM_EOB;
    }
    const uint32_t r = decodeLongShared( limAddr, lutAddr + DEC_OFF_OFFSET + table * ( LONG_OFFSETS * 2U ), hi );
    if ( r == 0 ) {
        return SYM_FAIL;
    }
    pos += r & 31U;
    *entry = 0x8000U | ( table << HUFF_LUT_BITS ) | ( r >> 5 );
    return ( r >> 5 ) == ldsU16( lutAddr + DEC_EOB_OFFSET + table * 2U ) ? SYM_EOB : SYM_PLAIN;
}

template<int BITS_LANES>
__global__ void __launch_bounds__( ( BITS_SLOTS / BITS_LANES ) * 32, BITS_CTAS_PER_SM )
huffmanBitsKernel( const uint8_t* __restrict__ data, uint64_t nbytes, BlockMeta* __restrict__ meta, uint32_t nblocks,
                   const uint8_t* __restrict__ selectorsAll, const BlockTables* __restrict__ tablesAll,
                   uint32_t* __restrict__ scratchAll, uint32_t* __restrict__ counter, const uint32_t* __restrict__ order,
                   uint32_t slotsPerCta )
{
    extern __shared__ __align__( 16 ) uint8_t decShared[];
    constexpr int BITS_WARPS = BITS_SLOTS / BITS_LANES;
    const uint32_t lane = threadIdx.x & 31U;
    if ( lane >= static_cast<uint32_t>( BITS_LANES ) ) {
        return;
    }
    /* a launch with few blocks (slotsPerCta < 32) leaves the higher lanes of every warp idle rather than whole SMs */
    const uint32_t slot = lane * BITS_WARPS + ( threadIdx.x >> 5 );
    if ( slot >= slotsPerCta ) {
        return;
    }
    uint8_t* const myDec = decShared + slot * DEC_STRIDE;
    const uint32_t lutAddr = static_cast<uint32_t>( __cvta_generic_to_shared( decShared ) ) + slot * DEC_STRIDE;
    const uint64_t nbits = nbytes * 8U;
    const uint64_t fullWords = nbytes >> 2;   /* words that can be loaded without looking at the file end */
    const uint32_t* const dataWords = reinterpret_cast<const uint32_t*>( data );

    bool haveBlock = false;
    uint32_t block = 0;
    const BlockTables* bt = nullptr;
    const uint8_t* selectors = nullptr;
    uint16_t* symOut = nullptr;
    uint32_t selectorsCount = 0;
    Decoder d{};
    uint32_t nsym = 0;
    uint32_t selIdx = 0;
    uint32_t selWord = 0;
    uint32_t selNext = 0;   /* the four selectors behind selWord's, loaded when selWord's come into use */
    uint32_t sink = 0;   /* keeps the touch() loads alive */

    while ( true ) {
        if ( !haveBlock ) {
            block = atomicAdd( counter, 1U );
            if ( block >= nblocks ) {
                break;
            }
            if ( order != nullptr ) {
                block = order[block];   /* longest block first: the decoders then finish at about the same time */
            }
            BlockMeta& bm = meta[block];
            if ( ( bm.status != IBZ2_ST_OK ) || ( bm.isEos != 0 ) ) {
                continue;
            }
            bt = tablesAll + block;
            const uint32_t symbolCount = bt->symbolCount;
            if ( symbolCount == 0 ) {
                /* the end-of-block symbol would be 1 = RUNB, which the reference takes for a run symbol first
                 * (bzip2.hpp:723): such a block never ends cleanly */
                bm.status = ST_INTERNAL_CAREFUL;
                continue;
            }
            selectors = selectorsAll + static_cast<uint64_t>( block ) * SEL_STRIDE;
            symOut = reinterpret_cast<uint16_t*>( reinterpret_cast<uint8_t*>( scratchAll + static_cast<uint64_t>( block ) * DBUF_STRIDE )
                                                  + K2_SYM16_OFFSET );
            selectorsCount = bt->selectorsCount;
            {
                /* code lengths -> this decoder's tables.  The end-of-block symbol is left out: the hot loop then has ONE
                 * exceptional entry value (0 = "not a plain symbol of at most 9 bits") to test for. */
                const uint4* src = reinterpret_cast<const uint4*>( &bt->lut[0][0] );
                const uint32_t groupCount = bt->groupCount;
                const uint32_t quads = groupCount * ( HUFF_LUT_SIZE / 8U );
                uint32_t* dst = reinterpret_cast<uint32_t*>( myDec );
                for ( uint32_t q = 0; q < quads; ++q ) {
                    const uint4 v = __ldg( src + q );
                    const uint32_t w[4] = { v.x, v.y, v.z, v.w };
                    uint32_t lens[2] = { 0, 0 };
#pragma unroll
                    for ( uint32_t i = 0; i < 4U; ++i ) {
                        const uint32_t a = ( ( w[i] & 0xFFFFU ) >> 5 ) > symbolCount ? 0U : ( w[i] & 31U );
                        const uint32_t b = ( w[i] >> 21 ) > symbolCount ? 0U : ( ( w[i] >> 16 ) & 31U );
                        lens[i >> 1] |= ( a | ( b << 8 ) ) << ( 16U * ( i & 1U ) );
                    }
                    dst[2U * q] = lens[0];
                    dst[2U * q + 1U] = lens[1];
                }
                /* what the codes of more than 9 bits need: limits as 20-bit left-aligned values, perm offsets, EOB index */
                for ( uint32_t g = 0; g < groupCount; ++g ) {
                    for ( uint32_t i = 0; i < LONG_LIMITS; ++i ) {
                        const uint64_t limit = bt->limit[g][LONG_FIRST_LEN + i];   /* <= 2^32, a multiple of 2^12 */
                        *reinterpret_cast<uint32_t*>( myDec + DEC_LIM_OFFSET + ( g * LONG_LIMITS + i ) * 4U )
                            = static_cast<uint32_t>( limit >> 12 );
                    }
                    for ( uint32_t i = 0; i < LONG_OFFSETS; ++i ) {
                        *reinterpret_cast<uint16_t*>( myDec + DEC_OFF_OFFSET + ( g * LONG_OFFSETS + i ) * 2U )
                            = bt->offset[g][LONG_FIRST_LEN + 1U + i];
                    }
                    *reinterpret_cast<uint16_t*>( myDec + DEC_EOB_OFFSET + g * 2U ) = bt->eobPerm[g];
                }
                asm volatile( "" ::: "memory" );   /* the asm loads below read what was just stored */
            }
            const uint64_t startBit = bt->dataStartBit;
            d.wordIdx = startBit >> 5;
            d.A = bswap( loadWordRaw( data, nbytes, d.wordIdx ) );
            d.B = bswap( loadWordRaw( data, nbytes, d.wordIdx + 1U ) );
            d.craw = loadWordRaw( data, nbytes, d.wordIdx + 2U );
            d.pos = static_cast<uint32_t>( startBit & 31U );
            nsym = 0;
            selIdx = 0;
            selNext = *reinterpret_cast<const uint32_t*>( selectors );
            haveBlock = true;
        }

        /* 0 = group done, 1 = end-of-block symbol seen, 2 = hand the block to the careful decoder */
        uint32_t outcome = 0;

        /* next group of 50 symbols: switch the table (bzip2.hpp:710-721) */
        if ( ( selIdx >= selectorsCount ) || ( nsym >= SYM_LIMIT ) || ( d.wordIdx * 32U + d.pos > nbits ) ) {
            outcome = 2;
        } else {
            if ( ( selIdx & 3U ) == 0 ) {
                /* SEL_STRIDE leaves room behind the last selector: the load needs no bounds check */
                selWord = selNext;
                selNext = *reinterpret_cast<const uint32_t*>( selectors + selIdx + 4U );
            }
            const uint32_t table = ( selWord >> ( 8U * ( selIdx & 3U ) ) ) & 0xFFU;
            ++selIdx;

            if ( d.wordIdx + GROUP_MAX_WORDS < fullWords ) {
                /* ---- hot loop: no file-end checks, three symbols per refill test ----
                 * Six symbols per iteration; when all six were plain symbols of at most 9 bits their entries leave as three
                 * 32-bit stores back to back (stores to one line that follow one another directly merge in L1; the same
                 * stores spread over the iteration cost 16 ms more on the 8 GB workload).  Otherwise the careful path
                 * stores what the fast path had decoded and goes on one symbol at a time; nothing it computes flows back
                 * into the registers of the fast path except the bit position. */
                const uint32_t tableAddr = lutAddr + table * HUFF_LUT_SIZE;
                const uint32_t* const wbase = dataWords + d.wordIdx;
                uint16_t* const groupOut = symOut + nsym;   /* nsym is a multiple of 50: 4-byte aligned */
                const uint32_t packBias = lutAddr * 0x00010001U;
                prefetchStream( wbase + IBZ2_K2B_PREFETCH_WORDS );   /* stays inside the file: see GROUP_MAX_WORDS */
#if IBZ2_K2B_TOUCH
                /* the sectors the groups after this one will read (a group takes 25 to 60 bytes): every first load of a
                 * sector went to L2 or HBM otherwise, and the word queue hides one L1 hit, not that */
                const uint32_t touched0 = touch( wbase + 24 );
                const uint32_t touched1 = touch( wbase + 32 );
#endif
                uint32_t A = d.A;
                uint32_t B = d.B;
                uint32_t craw = d.craw;
                uint32_t pos = d.pos;
                uint32_t wofs = 0;          /* words consumed in this group */
                uint32_t groupDone = GROUP_SIZE;   /* symbols of the group in front of the end-of-block symbol */

                /* branch-free: a select per register, a predicated load */
#define IBZ2_REFILL()                                                     \
    {                                                                     \
        const bool rf = pos >= 32U;                                       \
        A = rf ? B : A;                                                   \
        B = rf ? bswap( craw ) : B;                                       \
        wofs += rf ? 1U : 0U;                                             \
        if ( rf ) {                                                       \
            craw = __ldg( wbase + wofs + 2U );                            \
        }                                                                 \
        pos &= 31U;                                                       \
    }
                /* COUNT symbols (2 or 3) from a refilled window; a0.. = shared-memory addresses of their entries.  A symbol
                 * that is not a plain one of at most 9 bits has length 0: the position stays in front of it and the rest
                 * of the triple reads the same entry again, so `last` == 0 tells, and e0, e1 say how many were decoded. */
#define IBZ2_FAST( COUNT, a0, a1, a2, e0, e1, last )                                      \
    {                                                                                     \
        const uint32_t h0 = __funnelshift_l( B, A, pos );                                 \
        a0 = tableAddr + ( h0 >> ( 32 - HUFF_LUT_BITS ) );                                \
        e0 = ldsU8( a0 );                                                                 \
        const uint32_t h1 = h0 << e0;                                                     \
        a1 = tableAddr + ( h1 >> ( 32 - HUFF_LUT_BITS ) );                                \
        e1 = ldsU8( a1 );                                                                 \
        last = e1;                                                                        \
        pos += e0 + e1;                                                                   \
        if ( COUNT == 3 ) {                                                               \
            const uint32_t h2 = h1 << e1;                                                 \
            a2 = tableAddr + ( h2 >> ( 32 - HUFF_LUT_BITS ) );                            \
            last = ldsU8( a2 );                                                           \
            pos += last;                                                                  \
        }                                                                                 \
    }
                /* the careful path: symbols [valid, COUNT) of the iteration that starts at symbol FIRST of the group */
#define IBZ2_CAREFUL( COUNT, FIRST, valid )                                                                  \
    {                                                                                                        \
        for ( uint32_t kk = valid; kk < COUNT; ++kk ) {                                                      \
            IBZ2_REFILL();                                                                                   \
            uint32_t entry = 0;                                                                              \
            const uint32_t what = decodeOneSymbol( A, B, pos, lutAddr, table, bt, &entry );                  \
            if ( what != SYM_PLAIN ) {                                                                       \
                outcome = what == SYM_EOB ? 1U : 2U;                                                         \
                groupDone = ( FIRST ) + kk;                                                                  \
                break;                                                                                       \
            }                                                                                                \
            groupOut[( FIRST ) + kk] = static_cast<uint16_t>( entry );                                       \
        }                                                                                                    \
        IBZ2_REFILL();                                                                                       \
    }

#pragma unroll 1
                for ( uint32_t first = 0; first < GROUP_SIZE - 2U; first += 6U ) {
                    uint32_t a0, a1, a2, a3 = 0, a4 = 0, a5 = 0;
                    uint32_t e0, e1, last;
                    uint32_t valid = 6;
                    IBZ2_FAST( 3, a0, a1, a2, e0, e1, last );
                    if ( last == 0 ) {
                        valid = ( e0 != 0 ? 1U : 0U ) + ( e1 != 0 ? 1U : 0U );
                    } else {
                        IBZ2_REFILL();
                        IBZ2_FAST( 3, a3, a4, a5, e0, e1, last );
                        if ( last == 0 ) {
                            valid = 3U + ( e0 != 0 ? 1U : 0U ) + ( e1 != 0 ? 1U : 0U );
                        }
                    }
                    if ( valid == 6U ) {
                        IBZ2_REFILL();
                        /* entry = address - lutAddr < 2^15: both halves at once, modulo 2^32 */
                        uint32_t* const out = reinterpret_cast<uint32_t*>( groupOut + first );
                        out[0] = a0 + ( a1 << 16 ) - packBias;
                        out[1] = a2 + ( a3 << 16 ) - packBias;
                        out[2] = a4 + ( a5 << 16 ) - packBias;
                        continue;
                    }
                    {
                        const uint32_t a[6] = { a0, a1, a2, a3, a4, a5 };
#pragma unroll
                        for ( uint32_t k = 0; k < 5U; ++k ) {
                            if ( k < valid ) {
                                groupOut[first + k] = static_cast<uint16_t>( a[k] - lutAddr );
                            }
                        }
                    }
                    IBZ2_CAREFUL( 6U, first, valid );
                    if ( outcome != 0 ) {
                        break;
                    }
                }
                if ( outcome == 0 ) {
                    /* symbols 48 and 49 */
                    uint32_t a0, a1, a2 = 0;
                    uint32_t e0, e1, last;
                    IBZ2_FAST( 2, a0, a1, a2, e0, e1, last );
                    if ( last != 0 ) {
                        IBZ2_REFILL();
                        *reinterpret_cast<uint32_t*>( groupOut + GROUP_SIZE - 2U ) = a0 + ( a1 << 16 ) - packBias;
                    } else {
                        const uint32_t valid = e0 != 0 ? 1U : 0U;
                        if ( valid != 0 ) {
                            groupOut[GROUP_SIZE - 2U] = static_cast<uint16_t>( a0 - lutAddr );
                        }
                        IBZ2_CAREFUL( 2U, GROUP_SIZE - 2U, valid );
                    }
                }
#undef IBZ2_REFILL
#undef IBZ2_FAST
#undef IBZ2_CAREFUL
                d.A = A;
                d.B = B;
                d.craw = craw;
                d.pos = pos;
                d.wordIdx += wofs;
#if IBZ2_K2B_TOUCH
                sink ^= touched0 ^ touched1;
#endif
                nsym += outcome == 0 ? GROUP_SIZE : groupDone;
            } else {
                /* ---- the last groups before the end of the file: one symbol at a time with checked loads ---- */
                uint32_t g = 0;
                for ( ; g < GROUP_SIZE; ++g ) {
                    uint32_t entry = 0;
                    const uint32_t what = decodeOneSymbol( d.A, d.B, d.pos, lutAddr, table, bt, &entry );
                    if ( d.pos >= 32U ) {
                        d.A = d.B;
                        d.B = bswap( d.craw );
                        ++d.wordIdx;
                        d.craw = loadWordRaw( data, nbytes, d.wordIdx + 2U );
                        d.pos -= 32U;
                    }
                    if ( what != SYM_PLAIN ) {
                        outcome = what == SYM_EOB ? 1U : 2U;
                        break;
                    }
                    symOut[nsym + g] = static_cast<uint16_t>( entry );
                }
                nsym += g;
            }
        }

        if ( outcome != 0 ) {
            const uint64_t endBit = d.wordIdx * 32U + d.pos;
            BlockMeta& bm = meta[block];
            bm.cycleLen = sink;   /* a diagnostic field that K4b overwrites: gives the touch() loads a use */
            if ( ( outcome == 1 ) && ( endBit <= nbits ) ) {
                bm.nCoded = nsym;
                bm.endBit = endBit;
            } else {
                bm.status = ST_INTERNAL_CAREFUL;
            }
            haveBlock = false;
        }
    }
}

/*
 * K2b for launches with few blocks -- one CTA per bzip2 block, the bit positions of a whole group of 50 symbols looked up at
 * once.
 *
 * One lane per block is as fast as a single dependent chain of ~12 instructions per symbol gets (71 cycles per symbol on
 * an otherwise idle SM: 33 ms for a block of 900 k symbols however few blocks there are to decode -- the latency of a random
 * read, and all of a 100 MB file).  With SMs to spare the chain itself is shortened instead:
 *   A  thread r of 512 looks at bit offset r behind the (word-aligned) start of the group AS IF a symbol started there:
 *      nine bits -> code length -> ring1[r] = { offset of the symbol behind it, 1 symbol }; a bit pattern that is not a plain
 *      symbol of at most nine bits (longer code, end-of-block symbol, no code) points at itself with 0 symbols;
 *   B..D  pointer doubling: ring2[r] = ring1 o ring1, ring4 = ring2 o ring2, ring8 = ring4 o ring4, counts added;
 *   the chain: one warp follows ring8 six times and ring2 once from the group's first bit: 50 symbols in 7 dependent
 *      shared-memory loads instead of 50 x ( load + shift + address ).  The counts tell whether all hops were whole; if not
 *      (an exceptional symbol, the end of the block) the group is walked one symbol at a time through ring1 and
 *      decodeOneSymbol, which is also what a table switch in the middle of the ring amounts to;
 *   output: thread s < 50 walks from the chain's stop before symbol s & ~7 to symbol s through ring4/ring2/ring1 and stores
 *      the entry of the offset it arrives at, while the chain warp is already busy with the next group (the rings are
 *      double buffered).
 * The group is the unit because the table switches every 50 symbols at a bit position only the chain knows
 * (bzip2.hpp:710-721): 512 offsets cover 31 + 50 x 9 bits.  Same entries, same exits to the careful decoder as the
 * lane-per-block kernel above; ~20 instructions per symbol instead of 12, so it only pays while SMs would otherwise idle.
 */
constexpr int GRP_THREADS = 512;
constexpr uint32_t GRP_RING = GRP_THREADS;       /* bit offsets per group */
constexpr uint32_t GRP_LEVELS = 4;               /* hops of 1, 2, 4, 8 symbols */
constexpr uint32_t GRP_WORDS = 256;              /* staged big-endian words of the file (a ring) */
constexpr uint32_t GRP_REFILL = 64;
constexpr uint32_t GRP_POS_MASK = 2U * GRP_RING - 1U;   /* an entry = byte offset of the target's entry | symbols << 10 */
constexpr uint32_t GRP_COUNT_SHIFT = 10;
constexpr uint32_t GRP_CAREFUL_STOP = GRP_RING - 32U;   /* the careful walk re-bases the ring behind this offset */
/* launches with up to this many blocks per SM take this kernel.  Measured on config 3's blocks: 1133 blocks (7.7 per SM)
 * 74.0 ms here against 37.4 ms a lane per block, 2266 blocks 134 against 43.9; 112 blocks (config 1) 12.3 against 33.0 */
constexpr uint32_t GRP_BLOCKS_PER_SM = 2;
static_assert( 31U + GROUP_SIZE * HUFF_LUT_BITS < GRP_RING, "a group of plain symbols must fit the ring" );
static_assert( GRP_CAREFUL_STOP + HUFF_LUT_BITS <= GRP_RING, "no plain symbol in front of the stop leaves the ring" );

struct GroupShared
{
    alignas( 16 ) uint8_t dec[DEC_STRIDE];
    uint32_t words[GRP_WORDS];
    uint16_t ring[2][GRP_LEVELS][GRP_RING];
    uint16_t idx[2][GRP_RING];
    uint16_t coarse[2][8];
    unsigned long long nextBit;
    uint32_t left;       /* symbols the current selector still covers */
    uint32_t outcome;    /* 0 = go on, 1 = end-of-block symbol seen, 2 = careful decoder */
    uint32_t nsym;
    uint32_t fast;       /* the group just finished took the fast path: its entries are still to be stored */
    uint32_t block;
};

__device__ __forceinline__ uint32_t
ringLoad( const uint16_t* ring, uint32_t byteOffset )
{
    return *reinterpret_cast<const uint16_t*>( reinterpret_cast<const uint8_t*>( ring ) + byteOffset );
}

/* entries of a group that took the fast path; called by 64 threads with s = 0..63 */
__device__ __forceinline__ void
storeGroupEntries( const GroupShared& sh, uint32_t set, uint32_t s, uint32_t table, uint16_t* __restrict__ out )
{
    if ( s >= GROUP_SIZE ) {
        return;
    }
    const uint32_t hop = s < 48U ? ( s >> 3 ) : 6U;
    const uint32_t within = s < 48U ? ( s & 7U ) : s - 48U;
    uint32_t p2 = sh.coarse[set][hop];
    if ( ( within & 4U ) != 0 ) {
        p2 = ringLoad( sh.ring[set][2], p2 ) & GRP_POS_MASK;
    }
    if ( ( within & 2U ) != 0 ) {
        p2 = ringLoad( sh.ring[set][1], p2 ) & GRP_POS_MASK;
    }
    if ( ( within & 1U ) != 0 ) {
        p2 = ringLoad( sh.ring[set][0], p2 ) & GRP_POS_MASK;
    }
    out[s] = static_cast<uint16_t>( ( table << HUFF_LUT_BITS ) | ringLoad( sh.idx[set], p2 ) );
}

__global__ void __launch_bounds__( GRP_THREADS, 2 )
huffmanGroupsKernel( const uint8_t* __restrict__ data, uint64_t nbytes, BlockMeta* __restrict__ meta, uint32_t nblocks,
                     const uint8_t* __restrict__ selectorsAll, const BlockTables* __restrict__ tablesAll,
                     uint32_t* __restrict__ scratchAll, uint32_t* __restrict__ counter, const uint32_t* __restrict__ order )
{
    __shared__ GroupShared sh;
    const uint32_t t = threadIdx.x;
    const uint32_t decAddr = static_cast<uint32_t>( __cvta_generic_to_shared( sh.dec ) );
    const uint64_t nbits = nbytes * 8U;

    while ( true ) {
        __syncthreads();   /* everybody is done with the previous block's shared state */
        if ( t == 0 ) {
            sh.block = atomicAdd( counter, 1U );
        }
        __syncthreads();
        uint32_t block = sh.block;
        if ( block >= nblocks ) {
            break;
        }
        if ( order != nullptr ) {
            block = order[block];
        }
        BlockMeta& bm = meta[block];
        const BlockTables* const bt = tablesAll + block;
        {
            /* one thread looks and tells: what it writes below must not split the CTA */
            if ( t == 0 ) {
                uint32_t skip = ( ( bm.status != IBZ2_ST_OK ) || ( bm.isEos != 0 ) ) ? 1U : 0U;
                if ( ( skip == 0 ) && ( bt->symbolCount == 0 ) ) {
                    bm.status = ST_INTERNAL_CAREFUL;   /* see huffmanBitsKernel */
                    skip = 1;
                }
                sh.outcome = skip;
            }
            __syncthreads();
            if ( sh.outcome != 0 ) {
                continue;
            }
        }
        const uint32_t symbolCount = bt->symbolCount;
        const uint32_t groupCount = bt->groupCount;
        const uint32_t selectorsCount = bt->selectorsCount;
        const uint8_t* const selectors = selectorsAll + static_cast<uint64_t>( block ) * SEL_STRIDE;
        uint16_t* const symOut = reinterpret_cast<uint16_t*>( reinterpret_cast<uint8_t*>( scratchAll + static_cast<uint64_t>( block ) * DBUF_STRIDE )
                                                              + K2_SYM16_OFFSET );
        {
            /* the same tables as one decoder of huffmanBitsKernel keeps, built by all threads */
            const uint4* const src = reinterpret_cast<const uint4*>( &bt->lut[0][0] );
            uint32_t* const dst = reinterpret_cast<uint32_t*>( sh.dec );
            for ( uint32_t q = t; q < groupCount * ( HUFF_LUT_SIZE / 8U ); q += GRP_THREADS ) {
                const uint4 v = __ldg( src + q );
                const uint32_t w[4] = { v.x, v.y, v.z, v.w };
                uint32_t lens[2] = { 0, 0 };
#pragma unroll
                for ( uint32_t i = 0; i < 4U; ++i ) {
                    const uint32_t a = ( ( w[i] & 0xFFFFU ) >> 5 ) > symbolCount ? 0U : ( w[i] & 31U );
                    const uint32_t b = ( w[i] >> 21 ) > symbolCount ? 0U : ( ( w[i] >> 16 ) & 31U );
                    lens[i >> 1] |= ( a | ( b << 8 ) ) << ( 16U * ( i & 1U ) );
                }
                dst[2U * q] = lens[0];
                dst[2U * q + 1U] = lens[1];
            }
            for ( uint32_t i = t; i < groupCount * LONG_LIMITS; i += GRP_THREADS ) {
                const uint32_t g = i / LONG_LIMITS;
                const uint64_t limit = bt->limit[g][LONG_FIRST_LEN + i % LONG_LIMITS];
                *reinterpret_cast<uint32_t*>( sh.dec + DEC_LIM_OFFSET + i * 4U ) = static_cast<uint32_t>( limit >> 12 );
            }
            for ( uint32_t i = t; i < groupCount * LONG_OFFSETS; i += GRP_THREADS ) {
                const uint32_t g = i / LONG_OFFSETS;
                *reinterpret_cast<uint16_t*>( sh.dec + DEC_OFF_OFFSET + i * 2U ) = bt->offset[g][LONG_FIRST_LEN + 1U + i % LONG_OFFSETS];
            }
            if ( t < groupCount ) {
                *reinterpret_cast<uint16_t*>( sh.dec + DEC_EOB_OFFSET + t * 2U ) = bt->eobPerm[t];
            }
        }
        const uint64_t startBit = bt->dataStartBit;
        uint64_t loadedUpTo = ( startBit >> 5 ) + GRP_WORDS;   /* words in front of this one are staged */
        for ( uint32_t i = t; i < GRP_WORDS; i += GRP_THREADS ) {
            const uint64_t w = ( startBit >> 5 ) + i;
            sh.words[w & ( GRP_WORDS - 1U )] = bswap( loadWordRaw( data, nbytes, w ) );
        }
        if ( t == 0 ) {
            sh.nextBit = startBit;
            sh.left = GROUP_SIZE;
            sh.outcome = 0;
            sh.nsym = 0;
            sh.fast = 0;
        }
        uint32_t selIdx = 0;
        uint32_t tableAhead = selectorsCount > 0 ? selectors[0] : 0U;   /* the selector of the next fresh group, loaded early */
        uint32_t table = 0;
        uint32_t set = 0;
        uint32_t prevFast = 0;
        uint32_t prevTable = 0;
        uint32_t prevNsym = 0;
        uint32_t outcome = 0;
        uint32_t nsym = 0;
        uint64_t X = startBit;
        bool pending = false;
        uint32_t pendingRaw = 0;
        __syncthreads();

        while ( true ) {
            X = sh.nextBit;
            const uint32_t left = sh.left;
            outcome = sh.outcome;
            nsym = sh.nsym;
            if ( outcome != 0 ) {
                break;
            }
            if ( left == GROUP_SIZE ) {
                /* next group of 50 symbols: switch the table (bzip2.hpp:710-721) */
                if ( ( selIdx >= selectorsCount ) || ( nsym >= SYM_LIMIT ) || ( X > nbits ) ) {
                    outcome = 2;
                    break;
                }
                table = tableAhead;
                ++selIdx;
                tableAhead = selIdx < selectorsCount ? selectors[selIdx] : 0U;
            }
            const uint64_t baseWord = X >> 5;
            const uint32_t r0 = static_cast<uint32_t>( X & 31U );

            /* the staged words run GRP_REFILL at a time, loaded in one iteration and stored in the next; the slots they
             * overwrite hold words in front of baseWord (fewer than 160 are staged ahead when a load is issued) */
            if ( pending ) {
                if ( t < GRP_REFILL ) {
                    sh.words[( loadedUpTo + t ) & ( GRP_WORDS - 1U )] = bswap( pendingRaw );
                }
                loadedUpTo += GRP_REFILL;
                pending = false;
            }
            if ( loadedUpTo - baseWord < 160U ) {
                if ( t < GRP_REFILL ) {
                    pendingRaw = loadWordRaw( data, nbytes, loadedUpTo + t );
                } else if ( ( t < GRP_REFILL + 2U ) && ( ( loadedUpTo + 2U * GRP_REFILL ) * 4U < nbytes ) ) {
                    asm volatile( "prefetch.global.L2 [%0];" ::"l"( data + ( loadedUpTo + GRP_REFILL + ( t - GRP_REFILL ) * 32U ) * 4U ) );
                }
                pending = true;
            }

            /* A: every bit offset of the ring as a symbol start */
            uint32_t e;
            {
                const uint32_t w = t >> 5;
                const uint32_t wa = sh.words[( baseWord + w ) & ( GRP_WORDS - 1U )];
                const uint32_t wb = sh.words[( baseWord + w + 1U ) & ( GRP_WORDS - 1U )];
                const uint32_t hi = __funnelshift_l( wb, wa, t & 31U );
                const uint32_t idx = hi >> ( 32 - HUFF_LUT_BITS );
                const uint32_t len = sh.dec[table * HUFF_LUT_SIZE + idx];
                const uint32_t target = t + len;
                e = ( ( len != 0 ) && ( target < GRP_RING ) ) ? ( ( target * 2U ) | ( 1U << GRP_COUNT_SHIFT ) ) : t * 2U;
                sh.ring[set][0][t] = static_cast<uint16_t>( e );
                sh.idx[set][t] = static_cast<uint16_t>( idx );
            }
            __syncthreads();
            /* B..D: hops of 2, 4, 8 symbols */
#pragma unroll
            for ( uint32_t level = 1; level < GRP_LEVELS; ++level ) {
                const uint32_t f = ringLoad( sh.ring[set][level - 1U], e & GRP_POS_MASK );
                e = f + ( e & ~GRP_POS_MASK );
                sh.ring[set][level][t] = static_cast<uint16_t>( e );
                __syncthreads();
            }

            if ( t < 32U ) {
                /* the chain */
                uint32_t p2 = r0 * 2U;
                uint32_t count = 0;
                if ( left == GROUP_SIZE ) {
#pragma unroll
                    for ( uint32_t hop = 0; hop < 6U; ++hop ) {
                        if ( t == 0 ) {
                            sh.coarse[set][hop] = static_cast<uint16_t>( p2 );
                        }
                        const uint32_t f = ringLoad( sh.ring[set][3], p2 );
                        count += f >> GRP_COUNT_SHIFT;
                        p2 = f & GRP_POS_MASK;
                    }
                    if ( t == 0 ) {
                        sh.coarse[set][6] = static_cast<uint16_t>( p2 );
                    }
                    const uint32_t f = ringLoad( sh.ring[set][1], p2 );
                    count += f >> GRP_COUNT_SHIFT;
                    p2 = f & GRP_POS_MASK;
                }
                if ( count == GROUP_SIZE ) {
                    if ( t == 0 ) {
                        sh.nextBit = baseWord * 32U + ( p2 >> 1 );
                        sh.nsym = nsym + GROUP_SIZE;
                        sh.fast = 1;
                    }
                } else {
                    /* one symbol at a time: plain ones through ring1, the others through decodeOneSymbol */
                    uint32_t p = r0;
                    uint32_t done = 0;
                    uint32_t what = SYM_PLAIN;
                    uint16_t* const out = symOut + nsym;
                    while ( ( done < left ) && ( p < GRP_CAREFUL_STOP ) ) {
                        const uint32_t f = sh.ring[set][0][p];
                        if ( ( f >> GRP_COUNT_SHIFT ) != 0 ) {
                            if ( t == 0 ) {
                                out[done] = static_cast<uint16_t>( ( table << HUFF_LUT_BITS ) | sh.idx[set][p] );
                            }
                            p = ( f & GRP_POS_MASK ) >> 1;
                            ++done;
                            continue;
                        }
                        const uint32_t w = p >> 5;
                        uint32_t pos = p & 31U;
                        uint32_t entry = 0;
                        what = decodeOneSymbol( sh.words[( baseWord + w ) & ( GRP_WORDS - 1U )],
                                                sh.words[( baseWord + w + 1U ) & ( GRP_WORDS - 1U )], pos, decAddr, table, bt, &entry );
                        p = ( p & ~31U ) + pos;
                        if ( what != SYM_PLAIN ) {
                            break;
                        }
                        if ( t == 0 ) {
                            out[done] = static_cast<uint16_t>( entry );
                        }
                        ++done;
                    }
                    if ( t == 0 ) {
                        sh.nextBit = baseWord * 32U + p;
                        sh.nsym = nsym + done;
                        sh.fast = 0;
                        sh.outcome = what == SYM_PLAIN ? 0U : ( what == SYM_EOB ? 1U : 2U );
                        sh.left = done < left ? left - done : GROUP_SIZE;
                    }
                }
            } else if ( ( t < 96U ) && ( prevFast != 0 ) ) {
                storeGroupEntries( sh, set ^ 1U, t - 32U, prevTable, symOut + prevNsym );
            }
            __syncthreads();
            prevFast = sh.fast;
            prevTable = table;
            prevNsym = nsym;
            set ^= 1U;
        }

        if ( ( prevFast != 0 ) && ( t >= 32U ) && ( t < 96U ) ) {
            storeGroupEntries( sh, set ^ 1U, t - 32U, prevTable, symOut + prevNsym );
        }
        if ( t == 0 ) {
            if ( ( outcome == 1 ) && ( X <= nbits ) ) {
                bm.nCoded = nsym;
                bm.endBit = X;
            } else {
                bm.status = ST_INTERNAL_CAREFUL;
            }
        }
    }
}

#include "huffman_pipeline.inl"

/* What K2b stored per symbol -> the symbol itself: ( table << 9 ) | index goes through K2a's 9-bit table, 0x8000 | table <<
 * 9 | rank through perm (HuffmanCodingSymbolsPerLength.hpp:105-124).  One pass over the symbol arrays with 16-byte
 * accesses; the 9 KB of tables of a block stay in L1. */
constexpr int RESOLVE_THREADS = 256;
constexpr uint32_t RESOLVE_PER_CTA = RESOLVE_THREADS * 8U * 8U;   /* symbols per CTA */

__device__ __forceinline__ uint32_t
resolveEntry( const BlockTables& bt, uint32_t v )
{
    const uint16_t* const flat = &bt.lut[0][0];
    return ( v & 0x8000U ) != 0 ? static_cast<uint32_t>( __ldg( &bt.perm[( v >> 9 ) & 7U][v & 0x1FFU] ) )
                                : static_cast<uint32_t>( __ldg( flat + ( v & 0x7FFFU ) ) ) >> 5;
}

__global__ void __launch_bounds__( RESOLVE_THREADS )
resolveSymbolsKernel( const BlockMeta* __restrict__ meta, const BlockTables* __restrict__ tablesAll,
                      uint32_t* __restrict__ scratchAll )
{
    const uint32_t block = blockIdx.y;
    const BlockMeta& bm = meta[block];
    if ( ( bm.status != IBZ2_ST_OK ) || ( bm.isEos != 0 ) ) {
        return;
    }
    const uint32_t nCoded = bm.nCoded;
    const uint32_t ctaFirst = blockIdx.x * RESOLVE_PER_CTA;
    if ( ctaFirst >= nCoded ) {
        return;
    }
    const BlockTables& bt = tablesAll[block];
    uint4* const sym = reinterpret_cast<uint4*>( reinterpret_cast<uint8_t*>( scratchAll + static_cast<uint64_t>( block ) * DBUF_STRIDE )
                                                 + K2_SYM16_OFFSET );
    const uint32_t quads = ( nCoded + 7U ) / 8U;   /* the slot is larger than any clean block: whole quads are ours */
#pragma unroll 2
    for ( uint32_t i = 0; i < 8U; ++i ) {
        const uint32_t q = ctaFirst / 8U + i * RESOLVE_THREADS + threadIdx.x;
        if ( q >= quads ) {
            break;
        }
        const uint4 v = sym[q];
        uint32_t w[4] = { v.x, v.y, v.z, v.w };
        const uint32_t remaining = nCoded - q * 8U;   /* entries behind the block's last symbol are not entries */
#pragma unroll
        for ( uint32_t k = 0; k < 4U; ++k ) {
            const uint32_t a = ( 2U * k < remaining ) ? resolveEntry( bt, w[k] & 0xFFFFU ) : 0U;
            const uint32_t b = ( 2U * k + 1U < remaining ) ? resolveEntry( bt, w[k] >> 16 ) : 0U;
            w[k] = a | ( b << 16 );
        }
        sym[q] = make_uint4( w[0], w[1], w[2], w[3] );
    }
}
}  // namespace

void
launchHuffmanBits( const uint8_t* data, uint64_t nbytes, BlockMeta* meta, uint32_t nblocks, const uint8_t* selectors,
                   const BlockTables* tables, uint32_t* scratch, uint32_t* counter, bool ordered, cudaStream_t stream )
{
    /* few blocks per SM: a CTA per block (huffmanGroupsKernel); many: a lane per block.  IBZ2_K2B_KERNEL = lanes | groups
     * forces one of them (the tests run both), IBZ2_K2B_GROUPS_PER_SM moves the switch. */
    {
        const char* const which = std::getenv( "IBZ2_K2B_KERNEL" );
        const char* const perSm = std::getenv( "IBZ2_K2B_GROUPS_PER_SM" );
        const uint32_t limit = static_cast<uint32_t>( smCount() ) * ( perSm != nullptr ? std::strtoul( perSm, nullptr, 10 ) : GRP_BLOCKS_PER_SM );
        const bool groups = ( which != nullptr ) && ( ( which[0] == 'g' ) || ( which[0] == 'p' ) ) ? true
                            : ( which != nullptr ) && ( which[0] == 'l' ) ? false : nblocks <= limit;
        if ( groups && ( ( which == nullptr ) || ( which[0] != 'g' ) ) ) {
            /* the pipelined version (config 1: 11.5 ms against 12.3, random reads 22.5 ms against 24.3);
             * IBZ2_K2B_KERNEL=groups runs the first one */
            cudaMemsetAsync( counter, 0, sizeof( uint32_t ), stream );
            const uint32_t ctas = std::max( 1U, std::min( static_cast<uint32_t>( smCount() * 2 ), nblocks ) );
            cudaFuncSetAttribute( huffmanPipelineKernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  static_cast<int>( sizeof( PipeShared ) ) );
            huffmanPipelineKernel<<<ctas, PIPE_THREADS, sizeof( PipeShared ), stream>>>( data, nbytes, meta, nblocks, selectors, tables, scratch, counter,
                                                                     ordered ? counter + K2B_ORDER_OFFSET : nullptr );
            const dim3 grid( ( DBUF_SIZE + RESOLVE_PER_CTA - 1U ) / RESOLVE_PER_CTA, nblocks );
            resolveSymbolsKernel<<<grid, RESOLVE_THREADS, 0, stream>>>( meta, tables, scratch );
            return;
        }
        if ( groups ) {
            cudaMemsetAsync( counter, 0, sizeof( uint32_t ), stream );
            const uint32_t ctas = std::max( 1U, std::min( static_cast<uint32_t>( smCount() * 2 ), nblocks ) );
            huffmanGroupsKernel<<<ctas, GRP_THREADS, 0, stream>>>( data, nbytes, meta, nblocks, selectors, tables, scratch, counter,
                                                                 ordered ? counter + K2B_ORDER_OFFSET : nullptr );
            const dim3 grid( ( DBUF_SIZE + RESOLVE_PER_CTA - 1U ) / RESOLVE_PER_CTA, nblocks );
            resolveSymbolsKernel<<<grid, RESOLVE_THREADS, 0, stream>>>( meta, tables, scratch );
            return;
        }
    }
    const size_t shared = SH_TOTAL_BYTES;
    static const int lanes = [] () {
        const char* const env = std::getenv( "IBZ2_K2B_LANES" );
        const int v = env != nullptr ? std::atoi( env ) : BITS_LANES_DEFAULT;
        return ( ( v == 2 ) || ( v == 4 ) || ( v == 8 ) || ( v == 16 ) ) ? v : BITS_LANES_DEFAULT;
    }();
    /* one lane per warp would need 64 warps per SM: 32 registers per thread, not enough for this loop */
    auto* const kernel = lanes == 2 ? huffmanBitsKernel<2> : lanes == 4 ? huffmanBitsKernel<4>
                         : lanes == 8 ? huffmanBitsKernel<8> : huffmanBitsKernel<16>;
    cudaFuncSetAttribute( kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>( shared ) );
    cudaFuncSetAttribute( kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared );
    cudaMemsetAsync( counter, 0, sizeof( uint32_t ), stream );
    const uint32_t ctas = std::max( 1U, std::min( static_cast<uint32_t>( smCount() * BITS_CTAS_PER_SM ), nblocks ) );
    const uint32_t slotsPerCta = std::min<uint32_t>( BITS_SLOTS, ( nblocks + ctas - 1U ) / ctas );
    kernel<<<ctas, ( BITS_SLOTS / lanes ) * 32, shared, stream>>>( data, nbytes, meta, nblocks, selectors, tables,
                                                                 scratch, counter, ordered ? counter + K2B_ORDER_OFFSET : nullptr, slotsPerCta );
    const dim3 grid( ( DBUF_SIZE + RESOLVE_PER_CTA - 1U ) / RESOLVE_PER_CTA, nblocks );
    resolveSymbolsKernel<<<grid, RESOLVE_THREADS, 0, stream>>>( meta, tables, scratch );
}
}  // namespace ibz2
