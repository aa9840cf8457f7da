/*
 * K2x -- the exact, symbol-at-a-time decoder of one bzip2 block, one warp per block: block header, Huffman tables and the
 * Huffman / RUNA-RUNB / move-to-front loop exactly in the reference's order of events
 * (Block::readBlockHeader / readSymbolMaps / readSelectors / readTrees / readBlockData, indexed_bzip2/bzip2.hpp:479-807;
 * huffman/HuffmanCodingBase.hpp:46-187, HuffmanCodingSymbolsPerLength.hpp:43-124, HuffmanCodingShortBitsCached.hpp:39-150).
 *
 * This is NOT the throughput path.  The fast pipeline (K2a huffman_header.cu, K2b huffman_bits.cu, K2c..K2e mtf.cu) only
 * finishes blocks whose symbol stream is clean: it ends with the end-of-block symbol inside the file, holds at most
 * 900000 bytes and no RUNA/RUNB number of more than 20 digits.  Every other block (corrupt data, truncated files,
 * crafted run numbers whose 32-bit accumulation wraps) is marked ST_INTERNAL_CAREFUL and decoded again here, so that
 * the status reported for it is the exception the reference would have thrown, found at the same point of the stream.
 *
 * Execution model: all 32 lanes run the SAME decode (warp-uniform control flow) and serve as a distributed register
 * file: the input window (one 32-bit word of the current and next 128-byte line per lane), the move-to-front list
 * (entry i in lane i % 32, register i / 32; a move-to-front is two shuffles and a select per touched register instead
 * of the memmove of bzip2.hpp:781-785), run expansion and the 128-byte output rows.
 */
#include "common.cuh"

namespace ibz2
{
namespace
{
constexpr int LUT_BITS = 10;
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
    uint8_t stage[256];
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

/* Move entry ii of the lane-distributed list to the front and return it (bzip2.hpp:781-785). */
__device__ __forceinline__ uint32_t
moveToFront( uint32_t ( &m )[8], uint32_t ii, uint32_t lane )
{
    const uint32_t R = ii >> 5;
    const uint32_t c = ii & 31U;
    if ( R == 0 ) {
        const uint32_t v = __shfl_sync( FULL, m[0], c );
        const uint32_t up = __shfl_up_sync( FULL, m[0], 1 );
        m[0] = ( lane == 0 ) ? v : ( ( lane <= c ) ? up : m[0] );
        return v;
    }
    uint32_t sel = m[1];
#pragma unroll
    for ( int r = 2; r < 8; ++r ) {
        sel = ( R == static_cast<uint32_t>( r ) ) ? m[r] : sel;
    }
    const uint32_t v = __shfl_sync( FULL, sel, c );
#pragma unroll
    for ( int r = 7; r >= 0; --r ) {
        if ( static_cast<uint32_t>( r ) <= R ) {
            const uint32_t up = __shfl_up_sync( FULL, m[r], 1 );
            const uint32_t carry = ( r == 0 ) ? v : __shfl_sync( FULL, m[r > 0 ? r - 1 : 0], 31 );
            const bool affected = ( static_cast<uint32_t>( r ) < R ) || ( lane <= c );
            m[r] = ( lane == 0 ) ? carry : ( affected ? up : m[r] );
        }
    }
    return v;
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
huffmanCarefulKernel( const uint8_t* __restrict__ data, uint64_t nbytes, BlockMeta* __restrict__ meta,
                     uint32_t nblocks, uint8_t* __restrict__ lbytesAll, uint8_t* selectorsAll )
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
    if ( bm.status != ST_INTERNAL_CAREFUL ) {
        return;
    }
    uint8_t* const lbytes = lbytesAll + static_cast<uint64_t>( block ) * LBYTES_STRIDE;
    uint8_t* const selectors = selectorsAll + static_cast<uint64_t>( block ) * SEL_STRIDE;
    const uint64_t nbits = nbytes * 8U;
    const uint64_t startBit = bm.bitOffset;

    uint32_t status = IBZ2_ST_OK;
    uint32_t selectorsSeen = 0;   /* the selector count, for the message of a selector overrun */
    uint32_t overflowAt = 0;      /* the count that did not fit the block any more, for the message of an overflow */
    uint32_t headerCrc = 0;
    uint32_t origPtr = 0;
    uint32_t cnt = 0;       /* dbufCount */
    uint32_t isEos = 0;
    uint32_t isEof = 0;
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
            break;
        }
        const uint32_t selectorsCount = br.get( 15 );
        selectorsSeen = selectorsCount;
        if ( br.tell() > nbits ) {
            status = IBZ2_ST_EOF;
            break;
        }
        if ( selectorsCount == 0 ) {
            status = IBZ2_ST_SELECTOR_COUNT;
            break;
        }
        {
            uint32_t mtfSel = 0x543210U;  /* six 4-bit entries, entry i at bits [4i, 4i + 4) */
            uint32_t mine = 0;
            for ( uint32_t i = 0; i < selectorsCount; ++i ) {
                br.ensure();
                const uint32_t six = br.peek( 6 );
                const uint32_t j = static_cast<uint32_t>( __clz( static_cast<int>( ~( six << 26 ) ) ) );
                br.skip( static_cast<int>( j < 6U ? j + 1U : 6U ) );
                if ( j >= groupCount ) {
                    status = IBZ2_ST_SELECTOR_UNARY;
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

        /* readBlockData (bzip2.hpp:691-807) */
        uint32_t m[8];
        __syncwarp();
#pragma unroll
        for ( int r = 0; r < 8; ++r ) {
            const uint32_t idx = 32U * r + lane;
            m[r] = idx < symbolCount ? ws.s2b[idx] : 0U;
        }
        uint32_t* const stageWords = reinterpret_cast<uint32_t*>( ws.stage );
        uint32_t* const outWords = reinterpret_cast<uint32_t*>( lbytes );

        uint32_t hh = 0;
        uint32_t runPos = 0;
        uint32_t groupLeft = 0;
        uint32_t selIdx = 0;
        uint32_t selChunk = 0;
        uint32_t table = 0;
        uint32_t tableMaxLen = 0;

        while ( true ) {
            if ( groupLeft == 0 ) {
                if ( selIdx >= selectorsCount ) {
                    status = IBZ2_ST_SELECTOR_OVERRUN;
                    break;
                }
                if ( ( selIdx & 31U ) == 0 ) {
                    selChunk = selectors[selIdx + lane];
                    if ( br.tell() > nbits ) {
                        status = IBZ2_ST_EOF;
                        break;
                    }
                }
                table = __shfl_sync( FULL, selChunk, selIdx & 31U );
                tableMaxLen = ws.maxLen[table];
                ++selIdx;
                groupLeft = GROUP_SIZE;
            }
            --groupLeft;

            br.ensure();
            uint32_t e = ws.lut[table][br.peek( LUT_BITS )];
            uint32_t len = e & 31U;
            uint32_t sym = e >> 5;
            if ( e == 0 ) {
                /* code longer than LUT_BITS or no code at all (HuffmanCodingShortBitsCached.hpp:117-150) */
                bool found = false;
                for ( len = LUT_BITS + 1; len <= tableMaxLen; ++len ) {
                    const uint32_t cd = br.peek( static_cast<int>( len ) );
                    const uint32_t f = ws.first[table][len];
                    if ( ( cd >= f ) && ( cd - f < ws.count[table][len] ) ) {
                        sym = ws.perm[table][ws.offset[table][len] + ( cd - f )];
                        found = true;
                        break;
                    }
                }
                if ( !found ) {
                    status = ( br.tell() + tableMaxLen + 1U > nbits ) ? IBZ2_ST_EOF : IBZ2_ST_INVALID_CODE;
                    break;
                }
            }
            br.skip( static_cast<int>( len ) );

            if ( sym <= 1U ) {
                /* RUNA / RUNB (bzip2.hpp:726-743) */
                if ( runPos == 0 ) {
                    runPos = 1;
                    hh = 0;
                }
                hh += runPos << sym;
                runPos <<= 1;
                continue;
            }
            if ( runPos != 0 ) {
                /* flush the pending run of the byte at the front of the list (bzip2.hpp:749-763) */
                runPos = 0;
                if ( static_cast<uint64_t>( cnt ) + hh > DBUF_SIZE ) {
                    status = IBZ2_ST_DBUF_OVERFLOW;
                    overflowAt = static_cast<uint32_t>( cnt + hh );   /* what the reference's message quotes (bzip2.hpp:751-756: the sum in 32 bits) */
                    break;
                }
                const uint32_t uc = __shfl_sync( FULL, m[0], 0 );
                while ( hh > 0 ) {
                    if ( ( ( cnt & 127U ) == 0 ) && ( hh >= 128U ) ) {
                        outWords[( cnt >> 2 ) + lane] = uc * 0x01010101U;
                        cnt += 128U;
                        hh -= 128U;
                        continue;
                    }
                    const uint32_t space = 128U - ( cnt & 127U );
                    const uint32_t n = hh < space ? hh : space;
                    for ( uint32_t i = lane; i < n; i += 32U ) {
                        ws.stage[( cnt & 255U ) + i] = static_cast<uint8_t>( uc );
                    }
                    cnt += n;
                    hh -= n;
                    if ( ( cnt & 127U ) == 0 ) {
                        __syncwarp();
                        outWords[( ( cnt - 128U ) >> 2 ) + lane] = stageWords[( ( ( cnt - 128U ) & 255U ) >> 2 ) + lane];
                        __syncwarp();
                    }
                }
            }
            if ( sym > symbolCount ) {
                break;  /* end of block (bzip2.hpp:766-768) */
            }
            if ( cnt >= DBUF_SIZE ) {
                status = IBZ2_ST_DBUF_OVERFLOW;
                overflowAt = cnt;        /* bzip2.hpp:776-780 */
                break;
            }
            const uint32_t uc = moveToFront( m, sym - 1U, lane );
            if ( lane == 0 ) {
                ws.stage[cnt & 255U] = static_cast<uint8_t>( uc );
            }
            ++cnt;
            if ( ( cnt & 127U ) == 0 ) {
                __syncwarp();
                outWords[( ( cnt - 128U ) >> 2 ) + lane] = stageWords[( ( ( cnt - 128U ) & 255U ) >> 2 ) + lane];
                __syncwarp();
            }
        }
        if ( status != IBZ2_ST_OK ) {
            break;
        }
        if ( ( cnt & 127U ) != 0 ) {
            __syncwarp();
            outWords[( ( cnt & ~127U ) >> 2 ) + lane] = stageWords[( ( ( cnt & ~127U ) & 255U ) >> 2 ) + lane];
        }
        endBit = br.tell();
        if ( endBit > nbits ) {
            status = IBZ2_ST_EOF;
            break;
        }
        if ( origPtr >= cnt ) {
            status = IBZ2_ST_ORIGPTR;  /* bzip2.hpp:794-798 */
            break;
        }
    } while ( false );

    if ( ( status != IBZ2_ST_OK ) && ( status != IBZ2_ST_EOF ) && ( startBit <= nbits ) && ( br.tell() > nbits ) ) {
        status = IBZ2_ST_EOF;  /* the reference's bit reader would have thrown before reaching the bad field */
    }
    if ( lane == 0 ) {
        bm.status = status;
        bm.headerCrc = headerCrc;
        bm.origPtr = origPtr;
        /* a selector overrun reports the selector count, which the reference puts into its message (bzip2.hpp:714-718) */
        bm.n = ( status == IBZ2_ST_OK ) ? cnt : ( status == IBZ2_ST_SELECTOR_OVERRUN ? selectorsSeen
                                               : ( status == IBZ2_ST_DBUF_OVERFLOW ? overflowAt : 0U ) );
        bm.endBit = endBit;
        bm.isEos = isEos;
        bm.isEof = isEof;
        bm.outSize = 0;
        bm.crcAcc = 0;
        bm.calcCrc = 0;
        bm.cycleLen = 0;
        bm.writeOutput = 0;
        bm.nCoded = 0;
        bm.outOffset = 0;
    }
}
}  // namespace

void
launchHuffmanCareful( const uint8_t* data, uint64_t nbytes, BlockMeta* meta, uint32_t nblocks, uint8_t* lbytes,
                      uint8_t* selectors, cudaStream_t stream )
{
    const size_t shared = sizeof( WarpShared ) * WARPS_PER_CTA;
    cudaFuncSetAttribute( huffmanCarefulKernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                          static_cast<int>( shared ) );
    const uint32_t ctas = ( nblocks + WARPS_PER_CTA - 1 ) / WARPS_PER_CTA;
    huffmanCarefulKernel<<<ctas, WARPS_PER_CTA * 32, shared, stream>>>( data, nbytes, meta, nblocks, lbytes, selectors );
}
}  // namespace ibz2
