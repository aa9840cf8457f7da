/*
 * K2b -- the Huffman symbol loop of Block::readBlockData (indexed_bzip2/bzip2.hpp:709-723 and :766-768, with
 * huffman/HuffmanCodingShortBitsCached.hpp:98-150), one LANE per bzip2 block.
 *
 * The bit stream of a block is a dependent chain (window -> LUT -> code length -> shift); nothing inside one block can
 * shorten it, so the chip is filled with as many independent chains as shared memory allows:
 *   - every SM runs one persistent CTA of 32 / BITS_LANES warps with BITS_LANES active lanes each = 32 decoders per SM;
 *   - the six 9-bit LUTs of a decoder's block (K2a's BlockTables::lut, 6 KB) sit in shared memory as
 *     u16 lut[index][table][slot]: the decoders of one warp are neighbouring slots, i.e. they read different
 *     banks whatever they look up, and an entry's address is one shift and one multiply-add away from the window;
 *   - decoders pull block indices from a global counter; the unit of lockstep inside a warp is one GROUP of 50 symbols
 *     (the table switch of bzip2.hpp:710-721), so a decoder that finishes its block waits at most 49 symbols for its
 *     warp before it fetches the next block;
 *   - the loop does nothing but decode: refill (every second symbol), look up, shift, store the symbol, test for the
 *     end-of-block symbol.  The dbufCount / RUNA-RUNB bookkeeping of bzip2.hpp:726-790 is chunk-parallel work and
 *     happens in K2c..K2e (mtf.cu).
 * A block leaves this kernel as "clean" (status OK, nCoded symbols stored, endBit set) only when its symbol stream ends
 * with the end-of-block symbol inside the file.  Anything else -- a bit pattern without code, more groups than
 * selectors, the end of the file, more symbols than a block can hold -- marks the block ST_INTERNAL_CAREFUL and the
 * exact state machine in huffman_careful.cu decides what the reference would have thrown.
 *
 * Algorithmic HBM traffic per block: C compressed bytes read, 2 bytes per coded symbol written.
 */
#include "common.cuh"

#include <algorithm>
#include <cstdlib>

namespace ibz2
{
namespace
{
constexpr int BITS_SLOTS = 32;   /* decoders per SM: the LUT rows are 32 decoders wide */
constexpr int BITS_LANES_DEFAULT = 4;   /* active lanes per warp; BITS_SLOTS / lanes warps per CTA.  Measured on the 8 GB workload
                                         * once loads no longer stall the loop: 95.7 / 85.1 / 93.7 ms at 8 / 4 / 2 lanes */
constexpr uint32_t LUT_ROW_BYTES = BITS_SLOTS * 2U;               /* one u16 per decoder */
constexpr uint32_t LUT_IDX_STRIDE = MAX_GROUPS * LUT_ROW_BYTES;   /* u16 lut[index][table][decoder] */
constexpr uint32_t LUT_ENTRIES = MAX_GROUPS * HUFF_LUT_SIZE;      /* 3072 rows */
constexpr uint32_t SYM_LIMIT = DBUF_SIZE;                         /* a clean block has at most this many symbols */
constexpr uint32_t GROUP_MAX_WORDS = ( GROUP_SIZE * MAX_CODE_LEN + 31U ) / 32U + 4U + 96U;   /* + prefetch distance */

__device__ __forceinline__ void
prefetchL1( const void* p )
{
    asm volatile( "prefetch.global.L1 [%0];" ::"l"( p ) );
}

struct Decoder
{
    uint32_t hi;        /* window: next bits, left aligned, 64 bits in hi:lo */
    uint32_t lo;
    int nb;             /* valid bits in the window */
    uint32_t nextW;     /* big-endian word wordIdx of the file, already loaded */
    uint64_t wordIdx;
};

__device__ __forceinline__ void
refillChecked( Decoder& d, const uint8_t* __restrict__ data, uint64_t nbytes )
{
    if ( d.nb <= 32 ) {
        /* d.lo is empty here */
        d.hi |= d.nb == 32 ? 0U : ( d.nextW >> d.nb );
        d.lo = d.nb == 0 ? 0U : ( d.nextW << ( 32 - d.nb ) );
        d.nb += 32;
        ++d.wordIdx;
        d.nextW = loadWordBE( data, nbytes, d.wordIdx );
    }
}

__device__ __forceinline__ void
consume( Decoder& d, uint32_t len )
{
    d.hi = __funnelshift_l( d.lo, d.hi, len );
    d.lo <<= len;
    d.nb -= static_cast<int>( len );
}

/* Code longer than the LUT width, or no code at all (HuffmanCodingShortBitsCached.hpp:117-150).  Needs 20 valid bits.
 * The caller has established that no code of at most HUFF_LUT_BITS bits matches.  Canonical codes grow with their
 * length when left-aligned, so the length is the number of per-length limits the window has reached; the loads are
 * independent of each other (one L1 line).  Returns ( symbol << 5 ) | length like a LUT entry, 0 for "no code". */
__device__ __noinline__ uint32_t
decodeLong( const BlockTables* __restrict__ bt, uint32_t table, uint32_t hi )
{
    const uint64_t* const limit = bt->limit[table];
    uint32_t len = HUFF_LUT_BITS + 1;
#pragma unroll
    for ( uint32_t l = HUFF_LUT_BITS + 1; l <= MAX_CODE_LEN; ++l ) {
        len += ( static_cast<uint64_t>( hi ) >= limit[l] ) ? 1U : 0U;
    }
    if ( len > MAX_CODE_LEN ) {
        return 0;
    }
    const uint32_t idx = ( hi >> ( 32U - len ) ) + static_cast<uint32_t>( bt->base[table][len] );
    return ( static_cast<uint32_t>( bt->perm[table][idx] ) << 5 ) | len;
}

/* ---- shared memory of one CTA: everything a decoder looks at per symbol, one bank column per decoder ---- */
constexpr uint32_t LONG_FIRST_LEN = HUFF_LUT_BITS;                          /* limits are kept for lengths 9..20 */
constexpr uint32_t LONG_LIMITS = MAX_CODE_LEN - LONG_FIRST_LEN + 1U;        /* 12 */
constexpr uint32_t LONG_OFFSETS = MAX_CODE_LEN - LONG_FIRST_LEN;            /* lengths 10..20: 11 */
constexpr uint32_t SH_LUT_BYTES = LUT_ENTRIES * LUT_ROW_BYTES;              /* u16 [6 * 512][32] */
constexpr uint32_t SH_LIM_OFFSET = SH_LUT_BYTES;                            /* u32 [6][12][32], 20-bit left-aligned code limits */
constexpr uint32_t SH_LIM_BYTES = MAX_GROUPS * LONG_LIMITS * BITS_SLOTS * 4U;
constexpr uint32_t SH_OFF_OFFSET = SH_LIM_OFFSET + SH_LIM_BYTES;            /* u16 [6][11][32], perm offset of each length */
constexpr uint32_t SH_OFF_BYTES = MAX_GROUPS * LONG_OFFSETS * BITS_SLOTS * 2U;
constexpr uint32_t SH_EOB_OFFSET = SH_OFF_OFFSET + SH_OFF_BYTES;            /* u16 [8][32], perm index of the end-of-block symbol */
constexpr uint32_t SH_TOTAL_BYTES = SH_EOB_OFFSET + 8U * BITS_SLOTS * 2U;
static_assert( SH_TOTAL_BYTES <= 227U * 1024U, "one CTA per SM" );

__device__ __forceinline__ uint32_t
ldsU16( uint32_t addr )
{
    uint32_t v;   /* a wider destination is zero-extended: no separate mask on the dependent chain */
    asm volatile( "ld.volatile.shared.u16 %0, [%1];" : "=r"( v ) : "r"( addr ) );
    return v;
}

__device__ __forceinline__ uint32_t
ldsU32( uint32_t addr )
{
    uint32_t v;
    asm volatile( "ld.shared.u32 %0, [%1];" : "=r"( v ) : "r"( addr ) );
    return v;
}

/* LUT entry for the 9 bits on top of the window; tableAddr = shared address of lut[0][table][slot] */
__device__ __forceinline__ uint32_t
lookup( uint32_t tableAddr, uint32_t hi )
{
    return ldsU16( ( hi >> ( 32 - HUFF_LUT_BITS ) ) * LUT_IDX_STRIDE + tableAddr );   /* shift + multiply-add */
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
        lim[i] = ldsU32( limAddr + i * ( BITS_SLOTS * 4U ) );
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
    const uint32_t idx = ldsU16( offAddr + ( len - LONG_FIRST_LEN - 1U ) * ( BITS_SLOTS * 2U ) ) + rel;
    return ( idx << 5 ) | len;
}

template<int BITS_LANES>
__global__ void __launch_bounds__( ( BITS_SLOTS / BITS_LANES ) * 32, 1 )
huffmanBitsKernel( const uint8_t* __restrict__ data, uint64_t nbytes, BlockMeta* __restrict__ meta, uint32_t nblocks,
                   const uint8_t* __restrict__ selectorsAll, const BlockTables* __restrict__ tablesAll,
                   uint32_t* __restrict__ scratchAll, uint32_t* __restrict__ counter, const uint32_t* __restrict__ order,
                   uint32_t slotsPerCta )
{
    extern __shared__ __align__( 16 ) uint8_t lutShared[];
    constexpr int BITS_WARPS = BITS_SLOTS / BITS_LANES;
    const uint32_t lane = threadIdx.x & 31U;
    if ( lane >= static_cast<uint32_t>( BITS_LANES ) ) {
        return;
    }
    /* the lanes of one warp are BITS_WARPS slots apart: different banks whatever they look up; a launch with few blocks
     * (slotsPerCta < 32) leaves the second lane of every warp idle rather than whole SMs */
    const uint32_t slot = lane * BITS_WARPS + ( threadIdx.x >> 5 );
    if ( slot >= slotsPerCta ) {
        return;
    }
    uint8_t* const myLut = lutShared + slot * 2U;
    const uint32_t sharedBase = static_cast<uint32_t>( __cvta_generic_to_shared( lutShared ) );
    const uint32_t myLutAddr = sharedBase + slot * 2U;
    const uint32_t myLimAddr = sharedBase + SH_LIM_OFFSET + slot * 4U;
    const uint32_t myOffAddr = sharedBase + SH_OFF_OFFSET + slot * 2U;
    const uint32_t myEobAddr = sharedBase + SH_EOB_OFFSET + slot * 2U;
    const uint64_t nbits = nbytes * 8U;
    const uint64_t fullWords = nbytes >> 2;   /* words that can be loaded without looking at the file end */
    const uint32_t* const dataWords = reinterpret_cast<const uint32_t*>( data );

    bool haveBlock = false;
    uint32_t block = 0;
    const BlockTables* bt = nullptr;
    const uint8_t* selectors = nullptr;
    uint16_t* symOut = nullptr;
    uint32_t symbolCount = 0;
    uint32_t selectorsCount = 0;
    Decoder d{};
    uint32_t nsym = 0;
    uint32_t selIdx = 0;
    uint32_t selWord = 0;

    while ( true ) {
        if ( !haveBlock ) {
            block = atomicAdd( counter, 1U );
            if ( block >= nblocks ) {
                break;
            }
            if ( order != nullptr ) {
                block = order[block];   /* longest block first: the decoders then finish at about the same time */
            }
            const BlockMeta& bm = meta[block];
            if ( ( bm.status != IBZ2_ST_OK ) || ( bm.isEos != 0 ) ) {
                continue;
            }
            bt = tablesAll + block;
            selectors = selectorsAll + static_cast<uint64_t>( block ) * SEL_STRIDE;
            symOut = reinterpret_cast<uint16_t*>( reinterpret_cast<uint8_t*>( scratchAll + static_cast<uint64_t>( block ) * DBUF_STRIDE )
                                                  + K2_SYM16_OFFSET );
            symbolCount = bt->symbolCount;
            selectorsCount = bt->selectorsCount;
            {
                /* LUTs -> this decoder's column */
                const uint4* src = reinterpret_cast<const uint4*>( &bt->lut[0][0] );
                const uint32_t groupCount = bt->groupCount;
                const uint32_t quads = groupCount * ( HUFF_LUT_SIZE / 8U );
                for ( uint32_t q = 0; q < quads; ++q ) {
                    const uint4 v = __ldg( src + q );
                    uint32_t w[4] = { v.x, v.y, v.z, v.w };
                    /* The end-of-block symbol is taken out of the shared-memory copy: the hot loop then has ONE
                     * exceptional entry value (0 = "not a plain symbol of at most 9 bits") to test for. */
#pragma unroll
                    for ( uint32_t i = 0; i < 4U; ++i ) {
                        const uint32_t a = ( ( w[i] & 0xFFFFU ) >> 5 ) > symbolCount ? 0U : ( w[i] & 0xFFFFU );
                        const uint32_t b = ( w[i] >> 21 ) > symbolCount ? 0U : ( w[i] >> 16 );
                        w[i] = a | ( b << 16 );
                    }
                    /* quad q holds entries 8 * ( q % 64 ) .. + 7 of table q / 64 */
                    uint8_t* dst = myLut + ( q >> 6 ) * LUT_ROW_BYTES + ( ( q & 63U ) * 8U ) * LUT_IDX_STRIDE;
#pragma unroll
                    for ( uint32_t i = 0; i < 4U; ++i ) {
                        *reinterpret_cast<uint16_t*>( dst ) = static_cast<uint16_t>( w[i] );
                        *reinterpret_cast<uint16_t*>( dst + LUT_IDX_STRIDE ) = static_cast<uint16_t>( w[i] >> 16 );
                        dst += 2U * LUT_IDX_STRIDE;
                    }
                }
                /* what the codes of more than 9 bits need: limits as 20-bit left-aligned values, perm offsets, EOB index */
                for ( uint32_t g = 0; g < groupCount; ++g ) {
                    for ( uint32_t i = 0; i < LONG_LIMITS; ++i ) {
                        const uint64_t limit = bt->limit[g][LONG_FIRST_LEN + i];   /* <= 2^32, a multiple of 2^12 */
                        *reinterpret_cast<uint32_t*>( lutShared + SH_LIM_OFFSET + ( ( g * LONG_LIMITS + i ) * BITS_SLOTS + slot ) * 4U )
                            = static_cast<uint32_t>( limit >> 12 );
                    }
                    for ( uint32_t i = 0; i < LONG_OFFSETS; ++i ) {
                        *reinterpret_cast<uint16_t*>( lutShared + SH_OFF_OFFSET + ( ( g * LONG_OFFSETS + i ) * BITS_SLOTS + slot ) * 2U )
                            = bt->offset[g][LONG_FIRST_LEN + 1U + i];
                    }
                    *reinterpret_cast<uint16_t*>( lutShared + SH_EOB_OFFSET + ( g * BITS_SLOTS + slot ) * 2U ) = bt->eobPerm[g];
                }
                asm volatile( "" ::: "memory" );   /* the asm loads below read what was just stored */
            }
            const uint64_t startBit = bt->dataStartBit;
            d.wordIdx = startBit >> 5;
            const uint32_t w0 = loadWordBE( data, nbytes, d.wordIdx );
            const uint32_t w1 = loadWordBE( data, nbytes, d.wordIdx + 1U );
            d.wordIdx += 2U;
            d.nextW = loadWordBE( data, nbytes, d.wordIdx );
            const uint32_t skipBits = static_cast<uint32_t>( startBit & 31U );
            d.hi = __funnelshift_l( w1, w0, skipBits );
            d.lo = w1 << skipBits;
            d.nb = 64 - static_cast<int>( skipBits );
            nsym = 0;
            selIdx = 0;
            haveBlock = true;
        }

        /* 0 = group done, 1 = end-of-block symbol seen, 2 = hand the block to the careful decoder */
        uint32_t outcome = 0;

        /* next group of 50 symbols: switch the table (bzip2.hpp:710-721) */
        if ( ( selIdx >= selectorsCount ) || ( nsym >= SYM_LIMIT )
             || ( d.wordIdx * 32U - static_cast<uint64_t>( d.nb ) > nbits ) ) {
            outcome = 2;
        } else {
            if ( ( selIdx & 3U ) == 0 ) {
                selWord = *reinterpret_cast<const uint32_t*>( selectors + selIdx );
            }
            const uint32_t table = ( selWord >> ( 8U * ( selIdx & 3U ) ) ) & 0xFFU;
            const uint8_t* const tableLut = myLut + table * LUT_ROW_BYTES;
            ++selIdx;
            uint32_t* const pairOut = reinterpret_cast<uint32_t*>( symOut + nsym );   /* nsym is a multiple of 50 */

            if ( d.wordIdx + GROUP_MAX_WORDS < fullWords ) {
                /* ---- hot loop: two symbols per iteration, no file-end checks ----
                 * A lane's instruction stream is what bounds a block (one warp instruction every ~5 cycles, nothing to
                 * overlap it with inside the block), so the pair is ONE basic block with as few instructions as the
                 * format allows: look up, shift the window by the entry's length, look up again.  An entry of 0 (code
                 * longer than 9 bits, the end-of-block symbol, or no code) has length 0: the window stays where it
                 * is and every later lookup misses too, so one test at the end of the pair finds it, with the window
                 * still in front of the first symbol that needs the slow path.
                 * Three raw words are kept behind the window and a word is loaded only when one is consumed, so that a load has
                 * the time between two refills (about two pairs) to land before anything waits for its register. */
                const uint32_t tableAddr = myLutAddr + table * LUT_ROW_BYTES;
                const uint32_t limAddr = myLimAddr + table * ( LONG_LIMITS * BITS_SLOTS * 4U );
                const uint32_t offAddr = myOffAddr + table * ( LONG_OFFSETS * BITS_SLOTS * 2U );
                const uint32_t longTag = 0x8000U | ( table << 9 );
                const uint32_t* const wbase = dataWords + d.wordIdx;
                uint32_t wofs = 0;                                   /* words consumed in this group */
                prefetchL1( wbase + 64 );                            /* stays inside the file: see GROUP_MAX_WORDS */
                uint32_t raw0 = __byte_perm( d.nextW, 0, 0x0123 );   /* back to memory order: swapped when used */
                uint32_t raw1 = __ldg( wbase + 1 );
                uint32_t raw2 = __ldg( wbase + 2 );
                uint32_t hi = d.hi;
                uint32_t lo = d.lo;
                int nb = d.nb;
                uint32_t g = 0;
                uint32_t e = lookup( tableAddr, hi );   /* at least 15 valid bits here */
#pragma unroll 5
                for ( ; g < GROUP_SIZE / 2U; ++g ) {
                    {
                        /* refill as selects (no branch); lo is empty whenever rf holds */
                        const bool rf = nb <= 32;
                        const uint32_t x = rf ? __byte_perm( raw0, 0, 0x0123 ) : 0U;
                        hi |= __funnelshift_rc( x, 0U, static_cast<uint32_t>( nb ) );
                        lo |= __funnelshift_lc( 0U, x, static_cast<uint32_t>( 32 - nb ) );
                        nb += rf ? 32 : 0;
                        raw0 = rf ? raw1 : raw0;
                        raw1 = rf ? raw2 : raw1;
                        wofs += rf ? 1U : 0U;
                        if ( rf ) {
                            raw2 = __ldg( wbase + wofs + 2U );   /* a predicated load: nothing is fetched twice */
                        }
                    }
                    /* at least 33 valid bits; a plain symbol takes at most 9 */
                    const uint32_t e0 = e;
                    hi = __funnelshift_l( lo, hi, e0 );   /* shifts by e0 & 31 = the length */
                    lo <<= ( e0 & 31U );
                    nb -= static_cast<int>( e0 & 31U );
                    const uint32_t e1 = lookup( tableAddr, hi );
                    hi = __funnelshift_l( lo, hi, e1 );
                    lo <<= ( e1 & 31U );
                    nb -= static_cast<int>( e1 & 31U );
                    e = lookup( tableAddr, hi );          /* first symbol of the next pair: at least 15 valid bits */
                    pairOut[g] = ( e0 >> 5 ) | ( ( e1 >> 5 ) << 16 );
                    if ( e1 == 0 ) {
                        /* ---- slow path: the window stands in front of symbol `k` of the pair ---- */
                        uint32_t outFirst = e0 >> 5;
                        uint32_t outSecond = 0;
                        const uint32_t lim0 = ldsU32( limAddr );   /* codes of at most 9 bits are below it */
                        for ( uint32_t k = ( e0 == 0 ) ? 0U : 1U; k < 2U; ++k ) {
                            if ( nb <= 32 ) {
                                const uint32_t x = __byte_perm( raw0, 0, 0x0123 );
                                hi |= __funnelshift_rc( x, 0U, static_cast<uint32_t>( nb ) );
                                lo = __funnelshift_lc( 0U, x, static_cast<uint32_t>( 32 - nb ) );
                                nb += 32;
                                raw0 = raw1;
                                raw1 = raw2;
                                ++wofs;
                                raw2 = __ldg( wbase + wofs + 2U );
                            }
                            uint32_t ee = lookup( tableAddr, hi );
                            uint32_t out = ee >> 5;
                            bool isEob = false;
                            if ( ee == 0 ) {
                                if ( ( hi >> 12 ) < lim0 ) {
                                    /* a code of at most 9 bits that is not in the shared-memory LUT: end of block */
                                    ee = bt->lut[table][hi >> ( 32 - HUFF_LUT_BITS )];
                                    isEob = true;
                                } else {
                                    const uint32_t r = decodeLongShared( limAddr, offAddr, hi );
                                    if ( r == 0 ) {
                                        outcome = 2;
                                        break;
                                    }
                                    ee = r;
                                    out = longTag | ( r >> 5 );
                                    isEob = ( r >> 5 ) == ldsU16( myEobAddr + table * ( BITS_SLOTS * 2U ) );
                                }
                            }
                            hi = __funnelshift_l( lo, hi, ee );
                            lo <<= ( ee & 31U );
                            nb -= static_cast<int>( ee & 31U );
                            if ( isEob ) {
                                /* end of block (bzip2.hpp:766-768) */
                                outcome = 1;
                                nsym += 2U * g + k;
                                break;
                            }
                            if ( k == 0 ) {
                                outFirst = out;
                            } else {
                                outSecond = out;
                            }
                        }
                        pairOut[g] = outFirst | ( outSecond << 16 );
                        if ( outcome != 0 ) {
                            break;
                        }
                        if ( nb <= 32 ) {
                            const uint32_t x = __byte_perm( raw0, 0, 0x0123 );
                            hi |= __funnelshift_rc( x, 0U, static_cast<uint32_t>( nb ) );
                            lo = __funnelshift_lc( 0U, x, static_cast<uint32_t>( 32 - nb ) );
                            nb += 32;
                            raw0 = raw1;
                            raw1 = raw2;
                            ++wofs;
                            raw2 = __ldg( wbase + wofs + 2U );
                        }
                        e = lookup( tableAddr, hi );
                    }
                }
                /* the window holds words [-2] and [-1] of wbase + wofs, raw0 = word [0]: hand the position back in file terms */
                d.hi = hi;
                d.lo = lo;
                d.nb = nb;
                d.wordIdx += wofs;
                d.nextW = __byte_perm( raw0, 0, 0x0123 );
                if ( outcome == 0 ) {
                    nsym += GROUP_SIZE;
                }
            } else {
                /* ---- the last groups before the end of the file: same loop with checked loads ---- */
                uint32_t g = 0;
                for ( ; g < GROUP_SIZE; ++g ) {
                    refillChecked( d, data, nbytes );
                    uint32_t e = *reinterpret_cast<const uint16_t*>( tableLut + ( d.hi >> ( 32 - HUFF_LUT_BITS ) ) * LUT_IDX_STRIDE );
                    if ( e == 0 ) {
                        e = bt->lut[table][d.hi >> ( 32 - HUFF_LUT_BITS )];   /* the end-of-block symbol, if short */
                    }
                    if ( e == 0 ) {
                        e = decodeLong( bt, table, d.hi );
                        if ( e == 0 ) {
                            outcome = 2;
                            break;
                        }
                    }
                    consume( d, e & 31U );
                    const uint32_t sym = e >> 5;
                    if ( sym > symbolCount ) {
                        outcome = 1;
                        break;
                    }
                    symOut[nsym + g] = static_cast<uint16_t>( sym );
                }
                nsym += g;
            }
        }

        if ( outcome != 0 ) {
            const uint64_t endBit = d.wordIdx * 32U - static_cast<uint64_t>( d.nb );
            BlockMeta& bm = meta[block];
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

/* Symbols that K2b left as ( 0x8000 | table << 9 | permIndex ) -> the symbol itself (HuffmanCodingSymbolsPerLength.hpp:
 * 105-124).  One pass over the symbol arrays with 16-byte accesses; about one symbol in a hundred is rewritten. */
constexpr int RESOLVE_THREADS = 256;
constexpr uint32_t RESOLVE_PER_CTA = RESOLVE_THREADS * 8U * 8U;   /* symbols per CTA */

__global__ void __launch_bounds__( RESOLVE_THREADS )
resolveLongSymbolsKernel( const BlockMeta* __restrict__ meta, const BlockTables* __restrict__ tablesAll,
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
#pragma unroll
    for ( uint32_t i = 0; i < 8U; ++i ) {
        const uint32_t q = ctaFirst / 8U + i * RESOLVE_THREADS + threadIdx.x;
        if ( q >= quads ) {
            break;
        }
        uint4 v = sym[q];
        if ( ( ( v.x | v.y | v.z | v.w ) & 0x80008000U ) == 0 ) {
            continue;
        }
        uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
        for ( uint32_t k = 0; k < 4U; ++k ) {
            uint32_t a = w[k] & 0xFFFFU;
            uint32_t b = w[k] >> 16;
            if ( ( a & 0x8000U ) != 0 ) {
                a = bt.perm[( a >> 9 ) & 7U][a & 0x1FFU];
            }
            if ( ( b & 0x8000U ) != 0 ) {
                b = bt.perm[( b >> 9 ) & 7U][b & 0x1FFU];
            }
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
    const size_t shared = SH_TOTAL_BYTES;
    static const int lanes = [] () {
        const char* const env = std::getenv( "IBZ2_K2B_LANES" );
        const int v = env != nullptr ? std::atoi( env ) : BITS_LANES_DEFAULT;
        return ( ( v == 1 ) || ( v == 2 ) || ( v == 4 ) || ( v == 8 ) ) ? v : BITS_LANES_DEFAULT;
    }();
    auto* const kernel = lanes == 1 ? huffmanBitsKernel<1> : lanes == 2 ? huffmanBitsKernel<2>
                         : lanes == 4 ? huffmanBitsKernel<4> : huffmanBitsKernel<8>;
    cudaFuncSetAttribute( kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>( shared ) );
    cudaMemsetAsync( counter, 0, sizeof( uint32_t ), stream );
    const uint32_t ctas = std::max( 1U, std::min( static_cast<uint32_t>( smCount() ), nblocks ) );
    const uint32_t slotsPerCta = std::min<uint32_t>( BITS_SLOTS, ( nblocks + ctas - 1U ) / ctas );
    kernel<<<ctas, ( BITS_SLOTS / lanes ) * 32, shared, stream>>>( data, nbytes, meta, nblocks, selectors, tables,
                                                                 scratch, counter, ordered ? counter + K2B_ORDER_OFFSET : nullptr, slotsPerCta );
    const dim3 grid( ( DBUF_SIZE + RESOLVE_PER_CTA - 1U ) / RESOLVE_PER_CTA, nblocks );
    resolveLongSymbolsKernel<<<grid, RESOLVE_THREADS, 0, stream>>>( meta, tables, scratch );
}
}  // namespace ibz2
