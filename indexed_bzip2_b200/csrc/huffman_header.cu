/*
 * K2a -- block header, symbol map, selectors and Huffman tables, one warp per bzip2 block.
 * Replaces Block::readBlockHeader / readSymbolMaps / readSelectors / readTrees (indexed_bzip2/bzip2.hpp:479-685) and the
 * canonical Huffman coding construction (huffman/HuffmanCodingBase.hpp:46-187, HuffmanCodingSymbolsPerLength.hpp:43-124,
 * HuffmanCodingShortBitsCached.hpp:39-91).
 *
 * The header fields are bit-serial, so all 32 lanes of the warp run the SAME parse (warp-uniform control flow) and
 * the lanes serve as a distributed register file for the input window (each lane holds one 32-bit word of the current
 * and of the next 128-byte line, fetched with one shuffle).  The per-length symbol counting, the (length, symbol) sort
 * and the LUT fill are lane-parallel.  The result goes to the block's BlockTables in HBM: K2b (huffman_bits.cu) copies
 * the 9-bit LUTs into shared memory and decodes the symbol stream with one LANE per block.
 */
#include "common.cuh"

namespace ibz2
{
namespace
{
constexpr int LUT_BITS = HUFF_LUT_BITS;
constexpr int LUT_SIZE = 1 << LUT_BITS;
constexpr int WARPS_PER_CTA = 4;
constexpr uint32_t FULL = 0xFFFFFFFFU;

struct WarpShared
{
    uint16_t lut[MAX_GROUPS][LUT_SIZE];
    uint16_t perm[MAX_GROUPS][260];
    uint32_t first[MAX_GROUPS][MAX_CODE_LEN + 2];
    uint16_t count[MAX_GROUPS][MAX_CODE_LEN + 2];
    uint16_t offset[MAX_GROUPS][MAX_CODE_LEN + 2];
    uint32_t maxLen[MAX_GROUPS + 2];
    uint8_t lens[264];
    uint8_t s2b[256];
    uint8_t eobLen[MAX_GROUPS + 2];   /* code length of the end-of-block symbol per table */
};

/* MSB-first bit cursor shared by the warp (semantics of BitReader<true, uint64_t>, core/BitReader.hpp:190-290). */
struct WarpBits
{
    const uint8_t* data;
    uint64_t nbytes;
    uint64_t window;  /* next bits, left aligned */
    uint64_t gbase;   /* word index of lane 0's word in cur */
    uint32_t cur;     /* lane l: big-endian word gbase + l */
    uint32_t nxt;     /* lane l: big-endian word gbase + 32 + l */
    uint32_t widx;    /* next word of cur to move into the window */
    int nb;           /* valid bits in window */
    uint32_t lane;

    __device__ __forceinline__ void
    refill()
    {
        const uint32_t w = __shfl_sync( FULL, cur, widx );
        window |= static_cast<uint64_t>( w ) << ( 32 - nb );
        nb += 32;
        if ( ++widx == 32U ) {
            cur = nxt;
            gbase += 32U;
            nxt = loadWordBE( data, nbytes, gbase + 32U + lane );
            widx = 0;
        }
    }

    __device__ __forceinline__ void
    init( const uint8_t* d, uint64_t n, uint64_t bitOffset, uint32_t l )
    {
        data = d;
        nbytes = n;
        lane = l;
        const uint64_t w0 = bitOffset >> 5;
        gbase = w0 & ~31ULL;
        widx = static_cast<uint32_t>( w0 & 31U );
        cur = loadWordBE( data, nbytes, gbase + lane );
        nxt = loadWordBE( data, nbytes, gbase + 32U + lane );
        window = 0;
        nb = 0;
        refill();
        refill();
        const int skipBits = static_cast<int>( bitOffset & 31U );
        window <<= skipBits;
        nb -= skipBits;
    }

    /* afterwards at least 33 bits are valid */
    __device__ __forceinline__ void
    ensure()
    {
        if ( nb <= 32 ) {
            refill();
        }
    }

    __device__ __forceinline__ uint32_t
    peek( int n ) const  /* 1 <= n <= 32 */
    {
        return static_cast<uint32_t>( window >> ( 64 - n ) );
    }

    __device__ __forceinline__ void
    skip( int n )
    {
        window <<= n;
        nb -= n;
    }

    __device__ __forceinline__ uint32_t
    get( int n )
    {
        ensure();
        const uint32_t v = peek( n );
        skip( n );
        return v;
    }

    __device__ __forceinline__ uint64_t
    tell() const
    {
        return ( gbase + widx ) * 32U - static_cast<uint64_t>( nb );
    }
};

/* ---- readSelectors (bzip2.hpp:574-637) by the whole warp ----
 * A selector is j one bits and a zero bit (j < groupCount <= 6), then move-to-front over six table numbers.  Read one after
 * the other that is 160 dependent instructions per selector and 94 % of this kernel's time (18 000 selectors per block).
 * But every zero bit ends a selector: lane l takes 32 bits of the stream, counts its zeros, a prefix sum over the lanes
 * numbers them, and j is the distance to the zero in front (the ones at the end of the lane in front carry over).  The
 * move-to-front over the j values is a composition of permutations of six positions: every lane works out the permutation
 * of its share of the selectors, a scan over the lanes composes them, and a second pass over the share writes the table
 * numbers.  Anything unexpected -- a selector with groupCount or more ones -- returns false, and the caller reads the
 * selectors one at a time as before, which reports what the reference would report at the bit where it would. */
__device__ __forceinline__ uint32_t
selectorMtfStep( uint32_t& list, uint32_t j )
{
    const uint32_t uc = ( list >> ( 4U * j ) ) & 15U;
    const uint32_t low = list & ( ( 1U << ( 4U * j ) ) - 1U );
    const uint32_t high = ( j >= 5U ) ? 0U : ( list >> ( 4U * j + 4U ) ) << ( 4U * j + 4U );
    list = high | ( low << 4 ) | uc;
    return uc;
}

/* ( a then b ): the list that b makes of the list that a made of the identity */
__device__ __forceinline__ uint32_t
selectorCompose( uint32_t a, uint32_t b )
{
    uint32_t r = 0;
#pragma unroll
    for ( uint32_t i = 0; i < MAX_GROUPS; ++i ) {
        r |= ( ( a >> ( 4U * ( ( b >> ( 4U * i ) ) & 15U ) ) ) & 15U ) << ( 4U * i );
    }
    return r;
}

__device__ __forceinline__ bool
readSelectorsByWarp( const uint8_t* __restrict__ data, uint64_t nbytes, uint64_t startBit, uint32_t count, uint32_t groupCount,
                     uint8_t* __restrict__ selectors, uint32_t lane, uint64_t* endBit )
{
    /* ---- the unary part: selectors[i] = j ---- */
    uint32_t found = 0;   /* zeros seen in the chunks before this one */
    uint32_t carry = 0;   /* ones at the end of the chunk before this one */
    uint64_t end = 0;
    for ( uint64_t pos = startBit; found < count; pos += 1024U ) {
        const uint64_t bit = pos + 32U * lane;
        const uint32_t bits = __funnelshift_l( loadWordBE( data, nbytes, ( bit >> 5 ) + 1U ), loadWordBE( data, nbytes, bit >> 5 ),
                                               static_cast<uint32_t>( bit ) & 31U );
        const uint32_t zeros = ~bits;   /* bit 31 = the first bit */
        const uint32_t mine = static_cast<uint32_t>( __popc( zeros ) );
        uint32_t inclusive = mine;
#pragma unroll
        for ( int d = 1; d < 32; d <<= 1 ) {
            const uint32_t up = __shfl_up_sync( FULL, inclusive, d );
            if ( lane >= static_cast<uint32_t>( d ) ) {
                inclusive += up;
            }
        }
        const uint32_t total = __shfl_sync( FULL, inclusive, 31 );
        const uint32_t firstIndex = found + inclusive - mine;   /* number of this lane's first zero */
        /* ones behind the lane's last zero */
        const uint32_t trailing = zeros != 0 ? static_cast<uint32_t>( __ffs( static_cast<int>( zeros ) ) ) - 1U : 32U;
        uint32_t carryIn = __shfl_up_sync( FULL, trailing, 1 );
        carryIn = lane == 0 ? carry : carryIn;
        /* 32 ones in a row inside the selectors: not a selector */
        bool bad = ( zeros == 0 ) && ( firstIndex < count );
        uint64_t endHere = 0;
        uint32_t rest = zeros;
        uint32_t previous = 0;
        uint32_t index = firstIndex;
        bool first = true;
        while ( ( rest != 0 ) && ( index < count ) ) {
            const uint32_t b = static_cast<uint32_t>( __clz( static_cast<int>( rest ) ) );
            rest &= ~( 0x80000000U >> b );
            const uint32_t j = first ? b + carryIn : b - previous - 1U;
            bad = bad || ( j >= groupCount );
            selectors[index] = static_cast<uint8_t>( j );
            if ( index + 1U == count ) {
                endHere = bit + b + 1U;
            }
            previous = b;
            first = false;
            ++index;
        }
        if ( __any_sync( FULL, bad ) ) {
            return false;
        }
        const uint32_t holder = __ballot_sync( FULL, endHere != 0 );
        if ( holder != 0 ) {
            end = __shfl_sync( FULL, endHere, __ffs( static_cast<int>( holder ) ) - 1 );
        }
        found += total;
        carry = __shfl_sync( FULL, trailing, 31 );
    }
    *endBit = end;
    __syncwarp();   /* the j values are read back by other lanes */

    /* ---- move-to-front: lane l has selectors [l * share, (l + 1) * share), share a multiple of 4 ---- */
    const uint32_t share = ( ( count + 31U ) / 32U + 3U ) & ~3U;
    const uint32_t lo = min( count, lane * share );
    const uint32_t hi = min( count, lo + share );
    uint32_t* const words = reinterpret_cast<uint32_t*>( selectors );   /* 16-byte aligned; bytes behind `count` are nobody's */
    uint32_t list = 0x543210U;   /* six 4-bit entries, entry i at bits [4i, 4i + 4) */
    for ( uint32_t i = lo; i < hi; i += 4U ) {
        const uint32_t four = words[i >> 2];
#pragma unroll
        for ( uint32_t k = 0; k < 4U; ++k ) {
            if ( i + k < hi ) {
                selectorMtfStep( list, ( four >> ( 8U * k ) ) & 0xFFU );
            }
        }
    }
    uint32_t scanned = list;
#pragma unroll
    for ( int d = 1; d < 32; d <<= 1 ) {
        const uint32_t up = __shfl_up_sync( FULL, scanned, d );
        if ( lane >= static_cast<uint32_t>( d ) ) {
            scanned = selectorCompose( up, scanned );
        }
    }
    list = __shfl_up_sync( FULL, scanned, 1 );
    list = lane == 0 ? 0x543210U : list;
    for ( uint32_t i = lo; i < hi; i += 4U ) {
        const uint32_t four = words[i >> 2];
        uint32_t tables = 0;
#pragma unroll
        for ( uint32_t k = 0; k < 4U; ++k ) {
            if ( i + k < hi ) {
                tables |= selectorMtfStep( list, ( four >> ( 8U * k ) ) & 0xFFU ) << ( 8U * k );
            }
        }
        words[i >> 2] = tables;
    }
    __syncwarp();
    return true;
}

/* Canonical code of one table from its lengths in ws.lens[0..nsym) (bzip2.hpp:680-683 -> HuffmanCodingBase.hpp:165-187).
 * Returns false for an over-subscribed length set (Error::INVALID_CODE_LENGTHS); incomplete sets are accepted. */
__device__ __forceinline__ bool
buildTable( WarpShared& ws, uint32_t g, uint32_t nsym, uint32_t lane )
{
    /* lane l counts the symbols of length l */
    uint32_t c = 0;
    if ( ( lane >= 1U ) && ( lane <= MAX_CODE_LEN ) ) {
        for ( uint32_t s = 0; s < nsym; ++s ) {
            c += ( ws.lens[s] == lane ) ? 1U : 0U;
        }
    }
    const uint32_t present = __ballot_sync( FULL, c > 0 );
    const uint32_t minLen = static_cast<uint32_t>( __ffs( static_cast<int>( present ) ) ) - 1U;
    const uint32_t maxLen = 31U - static_cast<uint32_t>( __clz( static_cast<int>( present ) ) );

    /* checkCodeLengthFrequencies (HuffmanCodingBase.hpp:73-112, CHECK_OPTIMALITY = false) */
    uint32_t unused = 1U << minLen;
    bool ok = true;
    for ( uint32_t len = minLen; len <= maxLen; ++len ) {
        const uint32_t cl = __shfl_sync( FULL, c, len );
        if ( cl > unused ) {
            ok = false;
            break;
        }
        unused = ( unused - cl ) << 1;
    }
    if ( !ok ) {
        return false;
    }

    /* first code and symbol offset per length (initializeMinimumCodeValues, HuffmanCodingBase.hpp:117-149;
     * initializeSymbolsPerLength, HuffmanCodingSymbolsPerLength.hpp:43-68) */
    uint32_t code = 0;
    uint32_t off = 0;
    uint32_t myOffset = 0;
    for ( uint32_t len = 1; len <= MAX_CODE_LEN; ++len ) {
        const uint32_t cl = __shfl_sync( FULL, c, len );
        if ( lane == 0 ) {
            ws.first[g][len] = code;
            ws.count[g][len] = static_cast<uint16_t>( cl );
            ws.offset[g][len] = static_cast<uint16_t>( off );
        }
        if ( lane == len ) {
            myOffset = off;
        }
        off += cl;
        code = ( code + cl ) << 1;
    }
    if ( lane == 0 ) {
        ws.maxLen[g] = maxLen;
        ws.eobLen[g] = ws.lens[nsym - 1U];
    }
    /* symbols sorted by (length, symbol): lane l places the symbols of length l */
    if ( ( lane >= 1U ) && ( lane <= MAX_CODE_LEN ) && ( c > 0 ) ) {
        uint32_t k = myOffset;
        for ( uint32_t s = 0; s < nsym; ++s ) {
            if ( ws.lens[s] == lane ) {
                ws.perm[g][k++] = static_cast<uint16_t>( s );
            }
        }
    }
    __syncwarp();

    /* direct LUT for codes of at most LUT_BITS bits (HuffmanCodingShortBitsCached.hpp:48-91), filled per entry */
    const uint32_t lastLen = maxLen < static_cast<uint32_t>( LUT_BITS ) ? maxLen : static_cast<uint32_t>( LUT_BITS );
    for ( uint32_t idx = lane; idx < static_cast<uint32_t>( LUT_SIZE ); idx += 32U ) {
        uint32_t e = 0;
        for ( uint32_t len = minLen; len <= lastLen; ++len ) {
            const uint32_t cd = idx >> ( LUT_BITS - len );
            const uint32_t f = ws.first[g][len];
            if ( ( cd >= f ) && ( cd - f < ws.count[g][len] ) ) {
                e = ( static_cast<uint32_t>( ws.perm[g][ws.offset[g][len] + ( cd - f )] ) << 5 ) | len;
                break;
            }
        }
        ws.lut[g][idx] = static_cast<uint16_t>( e );
    }
    __syncwarp();
    return true;
}

__global__ void __launch_bounds__( WARPS_PER_CTA * 32 )
huffmanHeaderKernel( const uint8_t* __restrict__ data, uint64_t nbytes, BlockMeta* __restrict__ meta,
                     uint32_t nblocks, uint8_t* selectorsAll, BlockTables* __restrict__ tablesAll )
{
    extern __shared__ __align__( 16 ) uint8_t sharedRaw[];
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31U;
    const uint32_t block = blockIdx.x * WARPS_PER_CTA + warp;
    if ( block >= nblocks ) {
        return;
    }
    WarpShared& ws = reinterpret_cast<WarpShared*>( sharedRaw )[warp];
    BlockMeta& bm = meta[block];
    BlockTables& bt = tablesAll[block];
    uint8_t* const selectors = selectorsAll + static_cast<uint64_t>( block ) * SEL_STRIDE;
    const uint64_t nbits = nbytes * 8U;
    const uint64_t startBit = bm.bitOffset;

    uint32_t status = IBZ2_ST_OK;
    uint32_t headerCrc = 0;
    uint32_t origPtr = 0;
    uint32_t isEos = 0;
    uint32_t isEof = 0;
    uint32_t reported = 0;   /* the value the reference's message quotes for a header field it rejects (bzip2.hpp:581, 595, 627, 659) */
    uint64_t endBit = startBit;

    WarpBits br;

    /* All control flow below is warp-uniform.  `fail` jumps are expressed with a do { } while ( false ) block. */
    do {
        if ( startBit > nbits ) {
            status = IBZ2_ST_EOF;
            break;
        }
        br.init( data, nbytes, startBit, lane );

        /* readBlockHeader (bzip2.hpp:479-523) */
        const uint64_t magicHi = br.get( 24 );
        const uint64_t magicLo = br.get( 24 );
        const uint64_t magic = ( magicHi << 24 ) | magicLo;
        headerCrc = br.get( 32 );
        if ( br.tell() > nbits ) {
            status = IBZ2_ST_EOF;
            break;
        }
        if ( magic == MAGIC_EOS ) {
            isEos = 1;
            const uint32_t inByte = static_cast<uint32_t>( br.tell() & 7U );
            if ( inByte > 0 ) {
                br.get( static_cast<int>( 8U - inByte ) );
                if ( br.tell() > nbits ) {
                    status = IBZ2_ST_EOF;
                    break;
                }
            }
            endBit = br.tell();
            isEof = endBit >= nbits ? 1U : 0U;
            break;
        }
        if ( magic != MAGIC_BLOCK ) {
            status = IBZ2_ST_BAD_MAGIC;
            reported = static_cast<uint32_t>( magicHi );   /* the reference's message quotes the 48 bits (bzip2.hpp:502-507) */
            origPtr = static_cast<uint32_t>( magicLo );
            break;
        }
        const uint32_t randomised = br.get( 1 );
        if ( br.tell() > nbits ) {
            status = IBZ2_ST_EOF;
            break;
        }
        if ( randomised != 0 ) {
            status = IBZ2_ST_RANDOMISED;
            break;
        }
        origPtr = br.get( 24 );
        if ( br.tell() > nbits ) {
            status = IBZ2_ST_EOF;
            break;
        }
        if ( origPtr > DBUF_SIZE ) {
            status = IBZ2_ST_ORIGPTR_RANGE;
            break;
        }

        /* readSymbolMaps (bzip2.hpp:526-571) */
        const uint32_t usedMap = br.get( 16 );
        uint32_t symbolCount = 0;
        for ( uint32_t i = 0; i < 16U; ++i ) {
            if ( ( usedMap & ( 0x8000U >> i ) ) != 0 ) {
                const uint32_t bitmap = br.get( 16 );
                for ( uint32_t j = 0; j < 16U; ++j ) {
                    if ( ( bitmap & ( 0x8000U >> j ) ) != 0 ) {
                        if ( lane == 0 ) {
                            ws.s2b[symbolCount] = static_cast<uint8_t>( 16U * i + j );
                        }
                        ++symbolCount;
                    }
                }
            }
        }
        if ( br.tell() > nbits ) {
            status = IBZ2_ST_EOF;
            break;
        }

        /* readSelectors (bzip2.hpp:574-637) */
        const uint32_t groupCount = br.get( 3 );
        if ( br.tell() > nbits ) {
            status = IBZ2_ST_EOF;
            break;
        }
        if ( ( groupCount < 2U ) || ( groupCount > MAX_GROUPS ) ) {
            status = IBZ2_ST_GROUP_COUNT;
            reported = groupCount;
            break;
        }
        const uint32_t selectorsCount = br.get( 15 );
        if ( br.tell() > nbits ) {
            status = IBZ2_ST_EOF;
            break;
        }
        if ( selectorsCount == 0 ) {
            status = IBZ2_ST_SELECTOR_COUNT;
            reported = selectorsCount;
            break;
        }
        uint64_t behindSelectors = 0;
        if ( readSelectorsByWarp( data, nbytes, br.tell(), selectorsCount, groupCount, selectors, lane, &behindSelectors ) ) {
            br.init( data, nbytes, behindSelectors, lane );
            if ( br.tell() > nbits ) {
                status = IBZ2_ST_EOF;
                break;
            }
        } else {
            /* one at a time, from the same place: reports the selector that is wrong where the reference would */
            uint32_t mtfSel = 0x543210U;  /* six 4-bit entries, entry i at bits [4i, 4i + 4) */
            uint32_t mine = 0;
            for ( uint32_t i = 0; i < selectorsCount; ++i ) {
                br.ensure();
                const uint32_t six = br.peek( 6 );
                const uint32_t j = static_cast<uint32_t>( __clz( static_cast<int>( ~( six << 26 ) ) ) );
                br.skip( static_cast<int>( j < 6U ? j + 1U : 6U ) );
                if ( j >= groupCount ) {
                    status = IBZ2_ST_SELECTOR_UNARY;
                    reported = groupCount;
                    break;
                }
                const uint32_t uc = ( mtfSel >> ( 4U * j ) ) & 15U;
                const uint32_t low = mtfSel & ( ( 1U << ( 4U * j ) ) - 1U );
                const uint32_t high = ( j >= 5U ) ? 0U : ( mtfSel >> ( 4U * j + 4U ) ) << ( 4U * j + 4U );
                mtfSel = high | ( low << 4 ) | uc;
                if ( lane == ( i & 31U ) ) {
                    mine = uc;
                }
                if ( ( ( i & 31U ) == 31U ) || ( i + 1U == selectorsCount ) ) {
                    selectors[( i & ~31U ) + lane] = static_cast<uint8_t>( mine );
                }
            }
            if ( status != IBZ2_ST_OK ) {
                break;
            }
            if ( br.tell() > nbits ) {
                status = IBZ2_ST_EOF;
                break;
            }
        }

        /* readTrees (bzip2.hpp:644-685) */
        const uint32_t nsym = symbolCount + 2U;
        for ( uint32_t g = 0; ( g < groupCount ) && ( status == IBZ2_ST_OK ); ++g ) {
            uint32_t hh = br.get( 5 );
            for ( uint32_t s = 0; s < nsym; ++s ) {
                while ( true ) {
                    if ( hh - 1U > MAX_CODE_LEN - 1U ) {
                        status = IBZ2_ST_CODE_LENGTH;
                        reported = hh;
                        break;
                    }
                    br.ensure();
                    const uint32_t two = br.peek( 2 );
                    if ( ( two & 2U ) == 0 ) {
                        br.skip( 1 );
                        break;
                    }
                    br.skip( 2 );
                    hh += 1U - ( ( two & 1U ) << 1 );
                }
                if ( status != IBZ2_ST_OK ) {
                    break;
                }
                if ( lane == 0 ) {
                    ws.lens[s] = static_cast<uint8_t>( hh );
                }
            }
            if ( status != IBZ2_ST_OK ) {
                break;
            }
            if ( br.tell() > nbits ) {
                status = IBZ2_ST_EOF;
                break;
            }
            __syncwarp();
            if ( !buildTable( ws, g, nsym, lane ) ) {
                status = IBZ2_ST_CODE_LENGTHS;
                break;
            }
        }
        if ( status != IBZ2_ST_OK ) {
            break;
        }

        /* hand the tables to K2b */
        __syncwarp();
        {
            const uint32_t* src = reinterpret_cast<const uint32_t*>( &ws.lut[0][0] );
            uint32_t* dst = reinterpret_cast<uint32_t*>( &bt.lut[0][0] );
            for ( uint32_t i = lane; i < groupCount * ( LUT_SIZE / 2U ); i += 32U ) {
                dst[i] = src[i];
            }
            for ( uint32_t g = 0; g < groupCount; ++g ) {
                for ( uint32_t i = lane; i < nsym; i += 32U ) {
                    bt.perm[g][i] = ws.perm[g][i];
                }
                if ( lane < MAX_CODE_LEN + 2U ) {
                    bt.first[g][lane] = ws.first[g][lane];
                    bt.count[g][lane] = ws.count[g][lane];
                    bt.offset[g][lane] = ws.offset[g][lane];
                    /* left-aligned end of the codes of at most `lane` bits, and perm index = code + base (K2b decodeLong) */
                    const bool real = ( lane >= 1U ) && ( lane <= MAX_CODE_LEN );
                    bt.limit[g][lane] = real ? static_cast<uint64_t>( ws.first[g][lane] + ws.count[g][lane] ) << ( 32U - lane ) : 0U;
                    bt.base[g][lane] = real ? static_cast<int32_t>( ws.offset[g][lane] ) - static_cast<int32_t>( ws.first[g][lane] ) : 0;
                }
            }
            if ( lane < groupCount ) {
                bt.maxLen[lane] = ws.maxLen[lane];
                /* perm is sorted by (length, symbol) and the end-of-block symbol is the largest one */
                bt.eobPerm[lane] = static_cast<uint16_t>( ws.offset[lane][ws.eobLen[lane]] + ws.count[lane][ws.eobLen[lane]] - 1U );
            }
            for ( uint32_t i = lane; i < 256U; i += 32U ) {
                bt.s2b[i] = i < symbolCount ? ws.s2b[i] : static_cast<uint8_t>( i );
            }
            if ( lane == 0 ) {
                bt.dataStartBit = br.tell();
                bt.symbolCount = symbolCount;
                bt.selectorsCount = selectorsCount;
                bt.groupCount = groupCount;
            }
        }
    } while ( false );

    if ( ( status != IBZ2_ST_OK ) && ( status != IBZ2_ST_EOF ) && ( startBit <= nbits ) && ( br.tell() > nbits ) ) {
        status = IBZ2_ST_EOF;  /* the reference's bit reader would have thrown before reaching the bad field */
    }
    if ( lane == 0 ) {
        bm.status = status;
        bm.headerCrc = headerCrc;
        bm.origPtr = origPtr;
        bm.n = reported;   /* 0 unless a header field was rejected */
        bm.nCoded = 0;
        bm.endBit = endBit;
        bm.isEos = isEos;
        bm.isEof = isEof;
        bm.outSize = 0;
        bm.crcAcc = 0;
        bm.calcCrc = 0;
        bm.cycleLen = 0;
        bm.travStar = -1;
        bm.travK = 0;
        bm.writeOutput = 0;
        bm.outOffset = 0;
    }
}
}  // namespace

void
launchHuffmanHeader( const uint8_t* data, uint64_t nbytes, BlockMeta* meta, uint32_t nblocks, uint8_t* selectors,
                     BlockTables* tables, cudaStream_t stream )
{
    const size_t shared = sizeof( WarpShared ) * WARPS_PER_CTA;
    cudaFuncSetAttribute( huffmanHeaderKernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                          static_cast<int>( shared ) );
    const uint32_t ctas = ( nblocks + WARPS_PER_CTA - 1 ) / WARPS_PER_CTA;
    huffmanHeaderKernel<<<ctas, WARPS_PER_CTA * 32, shared, stream>>>( data, nbytes, meta, nblocks, selectors, tables );
}
}  // namespace ibz2
