/*
 * K6 -- CRC-32/BZIP2 (poly 0x04C11DB7, MSB first, init/xorout 0xFFFFFFFF; indexed_bzip2/bzip2.hpp:59-91) of every
 * block's decoded bytes, and the comparison with the block header's CRC (bzip2.hpp:900-907).
 *
 * The reference updates the CRC one byte at a time inside the serial traversal loop.  Here the CRC is evaluated as a
 * polynomial over GF(2): for a message of 32-bit big-endian words w_i,
 *     R0(M) = ( sum_i w_i(x) * x^(32 * wordsAfter_i) ) * x^32  mod P,      CRC = ~( R0(M) ^ 0xFFFFFFFF * x^(8|M|) ).
 * The output buffer is cut into windows of CRC_PIECE_BYTES; a CTA of 8 warps takes one window of one block, each warp an
 * eighth of it.  The window is large (128 KiB) because what follows the byte loop -- eight multiplications mod P per
 * warp, five more and a power of x for the window -- costs as much as 100 rows of the loop.
 * Lane l accumulates the words l, l + 32, ... of its span (coalesced 128-byte loads); stepping the accumulator over the
 * 1024 bits between two of its words is four table lookups (shared-memory tables of  v * x^(1024 + 8k)).  Lanes, warps
 * and windows are then combined by multiplications with fixed powers of x; the window's weight relative to the end of
 * the block may be a negative power, which exists because P is primitive (x^(2^32 - 1) = 1).  Window results are
 * xor-accumulated into the block's meta entry; crcFinalizeKernel adds the init term and the <= 3 trailing bytes.
 * Algorithmic HBM traffic: the U output bytes are read once.
 */
#include "common.cuh"

#include <vector>

namespace ibz2
{
namespace
{
constexpr uint32_t CRC_THREADS = 256;
constexpr uint32_t CRC_WARPS = CRC_THREADS / 32U;
constexpr uint32_t SPAN_BYTES = CRC_PIECE_BYTES / CRC_WARPS;   /* per warp */
constexpr uint32_t SPAN_ROWS = SPAN_BYTES / 128U;
static_assert( ( CRC_WARPS == 8 ) && ( SPAN_BYTES % 128U == 0 ), "three warp-combine steps, whole rows" );

__device__ uint32_t g_shiftTables[4][256]; /* [k][v] = v * x^(8 * (3 - k)) * x^1024 mod P */
__device__ uint32_t g_powBits[32];         /* x^(2^k) mod P */
__device__ uint32_t g_laneWeights[5];      /* x^(32 * 2^k) */
__device__ uint32_t g_warpWeights[3];      /* x^(8 * SPAN_BYTES * 2^k) */

/* x^e mod P for an exponent in bits; negative exponents through the group order 2^32 - 1 */
__device__ __forceinline__ uint32_t
reduceExponent( int64_t exponentBits )
{
    const int64_t order = 0xFFFFFFFFLL;
    int64_t e = exponentBits % order;
    if ( e < 0 ) {
        e += order;
    }
    return static_cast<uint32_t>( e );
}

__device__ uint32_t
powX( int64_t exponentBits )
{
    const uint32_t e = reduceExponent( exponentBits );
    uint32_t r = 1;
    for ( int k = 0; k < 32; ++k ) {
        if ( ( e >> k ) & 1 ) {
            r = crcMulMod( r, g_powBits[k] );
        }
    }
    return r;
}

/* the same by a whole warp: lane k holds x^(2^k) or 1, five rounds of pairwise products (every lane gets the result) */
__device__ __forceinline__ uint32_t
powXWarp( int64_t exponentBits, uint32_t lane )
{
    const uint32_t e = reduceExponent( exponentBits );
    uint32_t r = ( ( e >> lane ) & 1U ) != 0 ? g_powBits[lane] : 1U;
#pragma unroll 1
    for ( uint32_t d = 1; d < 32U; d <<= 1 ) {
        r = crcMulMod( r, __shfl_xor_sync( 0xFFFFFFFFU, r, d ) );
    }
    return r;
}

/* Range of block b that the windowed kernel covers: [start, mainEnd), mainEnd 4-byte aligned (or == start). */
__device__ __forceinline__ void
blockRange( const BlockMeta& bm, uint64_t& start, uint64_t& mainEnd, uint64_t& end )
{
    start = bm.outOffset;
    end = bm.outOffset + bm.outSize;
    const uint64_t aligned = end & ~3ULL;
    mainEnd = aligned > start ? aligned : start;
}

__global__ void __launch_bounds__( CRC_THREADS )
crcKernel( const uint8_t* __restrict__ out, BlockMeta* __restrict__ meta, uint32_t nblocks,
           const uint32_t* __restrict__ pieceStart, uint32_t pieceBytes )
{
    __shared__ uint32_t shift[4][256];
    __shared__ uint32_t warpValues[CRC_THREADS / 32];
    for ( uint32_t i = threadIdx.x; i < 1024U; i += CRC_THREADS ) {
        shift[i >> 8][i & 255U] = g_shiftTables[i >> 8][i & 255U];
    }
    __syncthreads();

    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31U;
    const uint32_t totalPieces = pieceStart[nblocks];

    for ( uint32_t piece = blockIdx.x; piece < totalPieces; piece += gridDim.x ) {
        /* block owning this piece: last b with pieceStart[b] <= piece */
        uint32_t lo = 0;
        uint32_t hi = nblocks;
        while ( hi - lo > 1U ) {
            const uint32_t mid = ( lo + hi ) >> 1;
            if ( pieceStart[mid] <= piece ) {
                lo = mid;
            } else {
                hi = mid;
            }
        }
        const uint32_t b = lo;
        const BlockMeta& bm = meta[b];
        uint64_t start, mainEnd, end;
        blockRange( bm, start, mainEnd, end );
        const uint64_t window = start / pieceBytes + ( piece - pieceStart[b] );
        const uint64_t windowBegin = window * pieceBytes;
        const uint64_t windowEnd = windowBegin + pieceBytes;

        /* span of this warp: SPAN_ROWS rows of 32 words */
        const uint64_t spanBegin = windowBegin + static_cast<uint64_t>( warp ) * SPAN_BYTES;
        uint32_t acc = 0;
        if ( ( spanBegin < mainEnd ) && ( spanBegin + SPAN_BYTES > start ) && ( bm.writeOutput != 0 ) ) {
#pragma unroll 4
            for ( uint32_t row = 0; row < SPAN_ROWS; ++row ) {
                const uint64_t offset = spanBegin + 128U * row + 4U * lane;
                uint32_t w = 0;
                if ( ( offset + 4U > start ) && ( offset < mainEnd ) ) {
                    w = __byte_perm( __ldg( reinterpret_cast<const uint32_t*>( out + offset ) ), 0, 0x0123 );
                    if ( offset < start ) {
                        w &= 0xFFFFFFFFU >> ( 8U * static_cast<uint32_t>( start - offset ) );
                    }
                }
                acc = shift[0][acc >> 24] ^ shift[1][( acc >> 16 ) & 0xFFU] ^ shift[2][( acc >> 8 ) & 0xFFU]
                      ^ shift[3][acc & 0xFFU] ^ w;
            }
        }
        /* lanes: sum_l acc_l * x^(32 * (31 - l)) ends up in lane 31 */
#pragma unroll
        for ( int k = 0; k < 5; ++k ) {
            const uint32_t other = __shfl_up_sync( 0xFFFFFFFFU, acc, 1U << k );
            const uint32_t combined = crcMulMod( other, g_laneWeights[k] ) ^ acc;
            if ( ( lane & ( ( 2U << k ) - 1U ) ) == ( 2U << k ) - 1U ) {
                acc = combined;
            }
        }
        if ( lane == 31U ) {
            warpValues[warp] = acc;
        }
        __syncthreads();
        if ( warp == 0 ) {
            uint32_t v = lane < CRC_THREADS / 32 ? warpValues[lane] : 0U;
#pragma unroll
            for ( int k = 0; k < 3; ++k ) {
                const uint32_t other = __shfl_up_sync( 0xFFFFFFFFU, v, 1U << k );
                const uint32_t combined = crcMulMod( other, g_warpWeights[k] ) ^ v;
                if ( ( lane & ( ( 2U << k ) - 1U ) ) == ( 2U << k ) - 1U ) {
                    v = combined;
                }
            }
            /* v (lane 7) is weighted relative to windowEnd; move it to mainEnd and append the x^32 factor */
            const int64_t exponent = 8 * ( static_cast<int64_t>( mainEnd ) - static_cast<int64_t>( windowEnd ) ) + 32;
            const uint32_t weight = powXWarp( exponent, lane );
            const uint32_t contribution = crcMulMod( v, weight );
            if ( ( lane == 7U ) && ( contribution != 0 ) ) {
                atomicXor( &meta[b].crcAcc, contribution );
            }
        }
        __syncthreads();
    }
}

__global__ void
crcFinalizeKernel( const uint8_t* __restrict__ out, BlockMeta* __restrict__ meta, uint32_t nblocks )
{
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if ( b >= nblocks ) {
        return;
    }
    BlockMeta& bm = meta[b];
    if ( ( bm.status != IBZ2_ST_OK ) || ( bm.n == 0 ) || ( bm.writeOutput == 0 ) ) {
        return;
    }
    uint64_t start, mainEnd, end;
    blockRange( bm, start, mainEnd, end );
    uint32_t crc = bm.crcAcc ^ crcMulMod( 0xFFFFFFFFU, powX( 8 * static_cast<int64_t>( mainEnd - start ) ) );
    for ( uint64_t i = mainEnd; i < end; ++i ) {
        crc ^= static_cast<uint32_t>( out[i] ) << 24;
        for ( int k = 0; k < 8; ++k ) {
            crc = ( crc << 1 ) ^ ( ( crc & 0x80000000U ) ? CRC_POLY : 0U );
        }
    }
    bm.calcCrc = ~crc;
    if ( bm.calcCrc != bm.headerCrc ) {
        bm.status = IBZ2_ST_CRC;
    }
}

uint32_t
hostPowX( uint64_t exponentBits )
{
    uint32_t r = 1;
    uint32_t base = 2;
    while ( exponentBits != 0 ) {
        if ( exponentBits & 1U ) {
            r = crcMulMod( r, base );
        }
        base = crcMulMod( base, base );
        exponentBits >>= 1;
    }
    return r;
}
/* ---- stream boundaries: one thread per block ---- */
__device__ __forceinline__ uint64_t
readBitsAt( const uint8_t* __restrict__ data, uint64_t nbytes, uint64_t bit, uint32_t count )   /* count <= 48 */
{
    const uint64_t word = bit >> 5;
    const uint32_t skip = static_cast<uint32_t>( bit & 31U );
    const uint64_t w0 = loadWordBE( data, nbytes, word );
    const uint64_t w1 = loadWordBE( data, nbytes, word + 1U );
    const uint64_t w2 = loadWordBE( data, nbytes, word + 2U );
    const uint64_t hi = ( w0 << 32 ) | w1;                                   /* bits word * 32 .. + 63 */
    const uint64_t window = skip == 0 ? hi : ( ( hi << skip ) | ( w2 >> ( 32U - skip ) ) );
    return window >> ( 64U - count );
}

__global__ void
streamProbeKernel( const uint8_t* __restrict__ data, uint64_t nbytes, BlockMeta* __restrict__ meta, uint32_t nblocks )
{
    const uint32_t block = blockIdx.x * blockDim.x + threadIdx.x;
    if ( block >= nblocks ) {
        return;
    }
    BlockMeta& bm = meta[block];
    uint32_t flags = 0;
    uint32_t eosCrc = 0;
    if ( ( bm.status == IBZ2_ST_OK ) && ( bm.isEos == 0 ) ) {
        const uint64_t nbits = nbytes * 8U;
        /* "BZh1" .. "BZh9" (bzip2.hpp:114-142) in the 32 bits in front of the block magic */
        if ( bm.bitOffset >= 32U ) {
            const uint32_t header = static_cast<uint32_t>( readBitsAt( data, nbytes, bm.bitOffset - 32U, 32 ) );
            if ( ( ( header >> 8 ) == 0x425A68U ) && ( ( header & 0xFFU ) >= '1' ) && ( ( header & 0xFFU ) <= '9' ) ) {
                flags |= STREAM_FIRST_BLOCK;
            }
        }
        if ( ( bm.endBit > bm.bitOffset ) && ( bm.endBit + 80U <= nbits ) && ( readBitsAt( data, nbytes, bm.endBit, 48 ) == MAGIC_EOS ) ) {
            flags |= STREAM_LAST_BLOCK;
            eosCrc = static_cast<uint32_t>( readBitsAt( data, nbytes, bm.endBit + 48U, 32 ) );
        }
    }
    bm.streamFlags = flags;
    bm.eosCrc = eosCrc;
}
}  // namespace

void
launchStreamProbe( const uint8_t* data, uint64_t nbytes, BlockMeta* meta, uint32_t nblocks, cudaStream_t stream )
{
    if ( nblocks > 0 ) {
        streamProbeKernel<<<( nblocks + 127U ) / 128U, 128, 0, stream>>>( data, nbytes, meta, nblocks );
    }
}

void
initCrcConstants()
{
    std::vector<uint32_t> tables( 4 * 256 );
    const uint32_t x1024 = hostPowX( 1024 );
    for ( uint32_t k = 0; k < 4; ++k ) {
        for ( uint32_t v = 0; v < 256; ++v ) {
            tables[k * 256 + v] = crcMulMod( v << ( 8U * ( 3U - k ) ), x1024 );
        }
    }
    uint32_t powBits[32];
    for ( uint32_t k = 0; k < 32; ++k ) {
        powBits[k] = hostPowX( 1ULL << k );
    }
    uint32_t laneWeights[5];
    for ( uint32_t k = 0; k < 5; ++k ) {
        laneWeights[k] = hostPowX( 32ULL << k );
    }
    uint32_t warpWeights[3];
    for ( uint32_t k = 0; k < 3; ++k ) {
        warpWeights[k] = hostPowX( ( 8ULL * SPAN_BYTES ) << k );
    }
    cudaMemcpyToSymbol( g_shiftTables, tables.data(), sizeof( uint32_t ) * tables.size() );
    cudaMemcpyToSymbol( g_powBits, powBits, sizeof( powBits ) );
    cudaMemcpyToSymbol( g_laneWeights, laneWeights, sizeof( laneWeights ) );
    cudaMemcpyToSymbol( g_warpWeights, warpWeights, sizeof( warpWeights ) );
}

void
launchCrc( const uint8_t* out, BlockMeta* meta, uint32_t nblocks, const uint32_t* pieceStart, uint32_t pieceBytes,
           cudaStream_t stream )
{
    const uint32_t ctas = static_cast<uint32_t>( smCount() ) * 8U;
    crcKernel<<<ctas, CRC_THREADS, 0, stream>>>( out, meta, nblocks, pieceStart, pieceBytes );
}

void
launchCrcFinalize( const uint8_t* out, BlockMeta* meta, uint32_t nblocks, cudaStream_t stream )
{
    crcFinalizeKernel<<<( nblocks + 127 ) / 128, 128, 0, stream>>>( out, meta, nblocks );
}
}  // namespace ibz2
