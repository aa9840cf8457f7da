/*
 * K2b -- the Huffman symbol loop of Block::readBlockData (indexed_bzip2/bzip2.hpp:709-723 and :766-768, with
 * huffman/HuffmanCodingShortBitsCached.hpp:98-150), one LANE per bzip2 block.
 *
 * The bit stream of a block is a dependent chain (window -> LUT -> code length -> shift); nothing inside one block can
 * shorten it, so the chip is filled with as many independent chains as shared memory allows:
 *   - every SM runs one persistent CTA of BITS_WARPS warps with BITS_LANES active lanes each = 32 decoders per SM;
 *   - the six 9-bit LUTs of a decoder's block (K2a's BlockTables::lut, 6 KB) sit in shared memory as
 *     u16 lut[table * 512 + index][slot]: the decoders of one warp are neighbouring slots, i.e. they read different
 *     banks whatever they look up;
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

namespace ibz2
{
namespace
{
constexpr int BITS_WARPS = 4;
constexpr int BITS_LANES = 8;
constexpr int BITS_SLOTS = BITS_WARPS * BITS_LANES;
static_assert( BITS_SLOTS == 32, "the LUT rows are 32 decoders wide" );
constexpr uint32_t LUT_ROW_BYTES = BITS_SLOTS * 2U;               /* one u16 per decoder */
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

__global__ void __launch_bounds__( BITS_WARPS * 32, 1 )
huffmanBitsKernel( const uint8_t* __restrict__ data, uint64_t nbytes, BlockMeta* __restrict__ meta, uint32_t nblocks,
                   const uint8_t* __restrict__ selectorsAll, const BlockTables* __restrict__ tablesAll,
                   uint32_t* __restrict__ scratchAll, uint32_t* __restrict__ counter )
{
    extern __shared__ __align__( 16 ) uint8_t lutShared[];   /* u16 [LUT_ENTRIES][BITS_SLOTS] */
    const uint32_t lane = threadIdx.x & 31U;
    if ( lane >= static_cast<uint32_t>( BITS_LANES ) ) {
        return;
    }
    const uint32_t slot = ( threadIdx.x >> 5 ) * BITS_LANES + lane;
    uint8_t* const myLut = lutShared + slot * 2U;
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
                const uint32_t quads = bt->groupCount * ( HUFF_LUT_SIZE / 8U );
                uint8_t* dst = myLut;
                for ( uint32_t q = 0; q < quads; ++q ) {
                    const uint4 v = __ldg( src + q );
                    const uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
                    for ( uint32_t i = 0; i < 4U; ++i ) {
                        *reinterpret_cast<uint16_t*>( dst ) = static_cast<uint16_t>( w[i] );
                        *reinterpret_cast<uint16_t*>( dst + LUT_ROW_BYTES ) = static_cast<uint16_t>( w[i] >> 16 );
                        dst += 2U * LUT_ROW_BYTES;
                    }
                }
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
            const uint8_t* const tableLut = myLut + table * ( HUFF_LUT_SIZE * LUT_ROW_BYTES );
            ++selIdx;
            uint32_t* const pairOut = reinterpret_cast<uint32_t*>( symOut + nsym );   /* nsym is a multiple of 50 */

            if ( d.wordIdx + GROUP_MAX_WORDS < fullWords ) {
                /* ---- hot loop: two symbols per iteration, no file-end checks ----
                 * Two raw words are kept in flight behind the window so that a load has two refills to land. */
                const uint32_t* wp = dataWords + d.wordIdx;
                uint32_t raw0 = __byte_perm( d.nextW, 0, 0x0123 );   /* back to memory order: swapped when used */
                uint32_t raw1 = __ldg( wp + 1 );
                uint32_t g = 0;
                uint32_t eobTail = 0;   /* bits consumed after the end-of-block symbol */
#pragma unroll 5
                for ( ; g < GROUP_SIZE / 2U; ++g ) {
                    /* The first lookup only needs the top 9 bits, which are valid before the refill (at least 15 bits
                     * are left after a pair), so it is issued first and the refill fills its latency. */
                    uint32_t e0 = *reinterpret_cast<const uint16_t*>( tableLut + ( d.hi >> ( 32 - HUFF_LUT_BITS ) ) * LUT_ROW_BYTES );
                    {
                        /* refill, written as selects so that it is predicated, not branched around: the load then
                         * lands in raw1's own register and nothing waits for it before the next refill */
                        const bool rf = d.nb <= 32;
                        const uint32_t x = __byte_perm( raw0, 0, 0x0123 );
                        const uint32_t addHi = __funnelshift_rc( x, 0U, static_cast<uint32_t>( d.nb ) );
                        const uint32_t newLo = __funnelshift_lc( 0U, x, static_cast<uint32_t>( 32 - d.nb ) );
                        d.hi |= rf ? addHi : 0U;
                        d.lo = rf ? newLo : d.lo;
                        d.nb += rf ? 32 : 0;
                        raw0 = rf ? raw1 : raw0;
                        wp += rf ? 1 : 0;
                        if ( rf ) {
                            raw1 = __ldg( wp + 1 );
                            if ( ( reinterpret_cast<uintptr_t>( wp ) & 127U ) == 0 ) {
                                prefetchL1( wp + 64 );   /* stays inside the file: see GROUP_MAX_WORDS */
                            }
                        }
                    }
                    if ( e0 == 0 ) {
                        e0 = decodeLong( bt, table, d.hi );
                        if ( e0 == 0 ) {
                            outcome = 2;
                            break;
                        }
                        consume( d, e0 & 31U );
                        if ( d.nb <= 32 ) {
                            const uint32_t x = __byte_perm( raw0, 0, 0x0123 );
                            d.hi |= __funnelshift_rc( x, 0U, static_cast<uint32_t>( d.nb ) );
                            d.lo = __funnelshift_lc( 0U, x, static_cast<uint32_t>( 32 - d.nb ) );
                            d.nb += 32;
                            raw0 = raw1;
                            raw1 = __ldg( wp + 2 );
                            ++wp;
                        }
                    } else {
                        consume( d, e0 & 31U );
                    }
                    const uint32_t sym0 = e0 >> 5;
                    uint32_t e1 = *reinterpret_cast<const uint16_t*>( tableLut + ( d.hi >> ( 32 - HUFF_LUT_BITS ) ) * LUT_ROW_BYTES );
                    if ( e1 == 0 ) {
                        if ( sym0 > symbolCount ) {
                            e1 = ( symbolCount + 2U ) << 5;   /* whatever follows the end-of-block symbol is not ours */
                        } else {
                            e1 = decodeLong( bt, table, d.hi );
                            if ( e1 == 0 ) {
                                outcome = 2;
                                break;
                            }
                        }
                        consume( d, e1 & 31U );
                        if ( d.nb <= 32 ) {
                            /* a long code can leave fewer than the 9 bits the next pair looks up before refilling */
                            const uint32_t x = __byte_perm( raw0, 0, 0x0123 );
                            d.hi |= __funnelshift_rc( x, 0U, static_cast<uint32_t>( d.nb ) );
                            d.lo = __funnelshift_lc( 0U, x, static_cast<uint32_t>( 32 - d.nb ) );
                            d.nb += 32;
                            raw0 = raw1;
                            raw1 = __ldg( wp + 2 );
                            ++wp;
                        }
                    } else {
                        consume( d, e1 & 31U );
                    }
                    const uint32_t sym1 = e1 >> 5;
                    pairOut[g] = sym0 | ( sym1 << 16 );
                    if ( max( sym0, sym1 ) > symbolCount ) {
                        /* end of block (bzip2.hpp:766-768) */
                        outcome = 1;
                        if ( sym0 > symbolCount ) {
                            nsym += 2U * g;
                            eobTail = e1 & 31U;
                        } else {
                            nsym += 2U * g + 1U;
                        }
                        break;
                    }
                }
                /* the window holds wp[-2], wp[-1]; raw0 = wp[0]: hand the position back in file terms */
                d.wordIdx = static_cast<uint64_t>( wp - dataWords );
                d.nextW = __byte_perm( raw0, 0, 0x0123 );
                if ( outcome == 0 ) {
                    nsym += GROUP_SIZE;
                } else if ( eobTail != 0 ) {
                    d.nb += static_cast<int>( eobTail );   /* only the bit position is used from here on */
                }
            } else {
                /* ---- the last groups before the end of the file: same loop with checked loads ---- */
                uint32_t g = 0;
                for ( ; g < GROUP_SIZE; ++g ) {
                    refillChecked( d, data, nbytes );
                    uint32_t e = *reinterpret_cast<const uint16_t*>( tableLut + ( d.hi >> ( 32 - HUFF_LUT_BITS ) ) * LUT_ROW_BYTES );
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
}  // namespace

void
launchHuffmanBits( const uint8_t* data, uint64_t nbytes, BlockMeta* meta, uint32_t nblocks, const uint8_t* selectors,
                   const BlockTables* tables, uint32_t* scratch, uint32_t* counter, cudaStream_t stream )
{
    const size_t shared = static_cast<size_t>( LUT_ENTRIES ) * LUT_ROW_BYTES;
    cudaFuncSetAttribute( huffmanBitsKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>( shared ) );
    cudaMemsetAsync( counter, 0, sizeof( uint32_t ), stream );
    const uint32_t ctas = static_cast<uint32_t>( std::min<uint64_t>( static_cast<uint64_t>( smCount() ),
                                                                   ( nblocks + BITS_SLOTS - 1U ) / BITS_SLOTS ) );
    huffmanBitsKernel<<<ctas, BITS_WARPS * 32, shared, stream>>>( data, nbytes, meta, nblocks, selectors, tables,
                                                                 scratch, counter );
}
}  // namespace ibz2
