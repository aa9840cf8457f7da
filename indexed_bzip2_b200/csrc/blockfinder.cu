/*
 * K1 -- BlockFinder: every bit offset at which the 48-bit block magic 0x314159265359 or the end-of-stream magic
 * 0x177245385090 starts.  Replaces BitStringFinder<48>::findBitStrings (core/BitStringFinder.hpp:158-238) and the
 * thread-pool wrapper ParallelBitStringFinder (core/ParallelBitStringFinder.hpp:159-265); both magics in one pass
 * (the reference's CLI scans the file twice, tools/ibzip2.cpp:109-123).
 *
 * Method (HBM-bound by construction, ~1 ALU op + 1/4 shared-memory lookup per byte):
 *   A 48-bit pattern starting at bit p always contains the aligned 16-bit field [32m, 32m + 16) with m = ceil(p/32),
 *   i.e. the first two bytes of the 32-bit word m, at pattern offset d = 32m - p in 0..31.  So a word can only be
 *   the anchor of a match if its leading halfword is one of the 2 x 32 values magic[d : d + 16].  Each thread streams
 *   16-byte vectors (coalesced 128-bit loads), tests the four leading halfwords against a 64 Kibit bitmap in shared
 *   memory, and only on a hit (p ~ 2^-10 per word on random data) are the 32 candidate positions p in
 *   (32m - 32, 32m] verified exactly with funnel shifts.  Every p has exactly one anchor word, so no match is reported
 *   twice.  Matches are appended through one global atomic counter (they are rare) and sorted by the caller.
 *   A hit is rare per word but not per warp (one warp iteration in eight has one when both magics are wanted), and a lane
 *   that loops over 32 positions on its own holds 31 others up for ~500 instructions: that was half of the kernel's
 *   time.  So the WARP verifies an anchor: the lane that found it broadcasts where it is, and lane s tests position s.
 */
#include "common.cuh"

namespace ibz2
{
namespace
{
constexpr int FINDER_THREADS = 256;
constexpr int FINDER_UNROLL = 4;
constexpr int BITMAP_WORDS = 65536 / 32;

__device__ __forceinline__ uint32_t
halfwordKey( uint64_t magic, uint32_t d )
{
    /* bits [d, d + 16) of the 48-bit magic (bit 0 = most significant), as the low halfword of a little-endian load */
    const uint32_t v = static_cast<uint32_t>( ( magic >> ( 32U - d ) ) & 0xFFFFU );
    return ( v >> 8 ) | ( ( v & 0xFFU ) << 8 );
}

__device__ __forceinline__ uint32_t
loadWordBEChecked( const uint8_t* __restrict__ data, uint64_t nbytes, int64_t byte )
{
    uint32_t w = 0;
#pragma unroll
    for ( int k = 0; k < 4; ++k ) {
        const int64_t b = byte + k;
        if ( ( b >= 0 ) && ( static_cast<uint64_t>( b ) < nbytes ) ) {
            w |= static_cast<uint32_t>( data[b] ) << ( 24 - 8 * k );
        }
    }
    return w;
}

/* Exact test of the 32 bit positions whose anchor word starts at byte wordByte, one position per lane of the calling
 * warp (all 32 lanes call it with the same arguments). */
__device__ __forceinline__ void
verifyAnchor( const uint8_t* __restrict__ data, uint64_t nbytes, uint64_t wordByte, uint32_t which,
              uint64_t* __restrict__ found, uint32_t* __restrict__ foundCount, uint32_t capacity )
{
    const int64_t base = static_cast<int64_t>( wordByte ) - 4;
    const uint32_t a = loadWordBEChecked( data, nbytes, base );
    const uint32_t b = loadWordBEChecked( data, nbytes, base + 4 );
    const uint32_t c = loadWordBEChecked( data, nbytes, base + 8 );
    const uint64_t nbits = nbytes * 8U;
    const uint32_t s = ( threadIdx.x & 31U ) + 1U;   /* 1..32 */
    const int64_t p = base * 8 + s;
    if ( ( p < 0 ) || ( static_cast<uint64_t>( p ) + 48U > nbits ) ) {
        return;
    }
    const uint32_t hi = ( s == 32U ) ? b : __funnelshift_l( b, a, s );
    const uint32_t mid = ( s == 32U ) ? c : __funnelshift_l( c, b, s );
    const uint64_t v = ( static_cast<uint64_t>( hi ) << 16 ) | ( mid >> 16 );
    uint64_t report = 0;
    bool hit = false;
    if ( ( ( which & IBZ2_FIND_BLOCK ) != 0 ) && ( v == MAGIC_BLOCK ) ) {
        report = static_cast<uint64_t>( p );
        hit = true;
    } else if ( ( ( which & IBZ2_FIND_EOS ) != 0 ) && ( v == MAGIC_EOS ) ) {
        report = static_cast<uint64_t>( p ) | IBZ2_OFFSET_EOS_FLAG;
        hit = true;
    }
    if ( hit ) {
        const uint32_t idx = atomicAdd( foundCount, 1U );
        if ( idx < capacity ) {
            found[idx] = report;
        }
    }
}

/* The anchors that the lanes of a warp have found in one 16-byte vector each (bit k of `hits`: word k), verified by the
 * whole warp one after the other. */
__device__ __forceinline__ void
verifyWarp( const uint8_t* __restrict__ data, uint64_t nbytes, uint64_t vector, uint32_t hits, uint32_t which,
            uint64_t* __restrict__ found, uint32_t* __restrict__ foundCount, uint32_t capacity )
{
    uint32_t pending = __ballot_sync( 0xFFFFFFFFU, hits != 0 );
    while ( pending != 0 ) {
        const int source = __ffs( static_cast<int>( pending ) ) - 1;
        pending &= pending - 1U;
        const uint64_t v = __shfl_sync( 0xFFFFFFFFU, vector, source );
        uint32_t h = __shfl_sync( 0xFFFFFFFFU, hits, source );
        while ( h != 0 ) {
            const uint32_t k = static_cast<uint32_t>( __ffs( static_cast<int>( h ) ) ) - 1U;
            h &= h - 1U;
            verifyAnchor( data, nbytes, v * 16U + 4U * k, which, found, foundCount, capacity );
        }
    }
}

__global__ void __launch_bounds__( FINDER_THREADS )
findMagicsKernel( const uint8_t* __restrict__ data, uint64_t nbytes, uint32_t which,
                  uint64_t* __restrict__ found, uint32_t* __restrict__ foundCount, uint32_t capacity )
{
    __shared__ uint32_t bitmap[BITMAP_WORDS];
    for ( int i = threadIdx.x; i < BITMAP_WORDS; i += FINDER_THREADS ) {
        bitmap[i] = 0;
    }
    __syncthreads();
    if ( threadIdx.x < 64 ) {
        const uint32_t d = threadIdx.x & 31U;
        const bool eos = threadIdx.x >= 32;
        if ( ( which & ( eos ? IBZ2_FIND_EOS : IBZ2_FIND_BLOCK ) ) != 0 ) {
            const uint32_t key = halfwordKey( eos ? MAGIC_EOS : MAGIC_BLOCK, d );
            atomicOr( &bitmap[key >> 5], 1U << ( key & 31U ) );
        }
    }
    __syncthreads();

    const uint64_t nVectors = nbytes / 16U;
    const uint4* __restrict__ vectors = reinterpret_cast<const uint4*>( data );
    const uint64_t stride = static_cast<uint64_t>( gridDim.x ) * FINDER_THREADS;
    const uint64_t tid = static_cast<uint64_t>( blockIdx.x ) * FINDER_THREADS + threadIdx.x;

    /* the trip count is the warp's (its first lane's): the verification is a warp affair */
    const uint64_t warpFirst = tid - ( threadIdx.x & 31U );
    for ( uint64_t v0 = tid, w0 = warpFirst; w0 < nVectors; v0 += stride * FINDER_UNROLL, w0 += stride * FINDER_UNROLL ) {
        uint4 q[FINDER_UNROLL];
#pragma unroll
        for ( int u = 0; u < FINDER_UNROLL; ++u ) {
            const uint64_t v = v0 + u * stride;
            q[u] = ( v < nVectors ) ? __ldg( vectors + v ) : make_uint4( 0, 0, 0, 0 );
        }
#pragma unroll
        for ( int u = 0; u < FINDER_UNROLL; ++u ) {
            const uint64_t v = v0 + u * stride;
            const uint32_t words[4] = { q[u].x, q[u].y, q[u].z, q[u].w };
            uint32_t hits = 0;
#pragma unroll
            for ( int k = 0; k < 4; ++k ) {
                const uint32_t key = words[k] & 0xFFFFU;
                hits |= ( ( bitmap[key >> 5] >> ( key & 31U ) ) & 1U ) << k;
            }
            hits = v < nVectors ? hits : 0U;
            if ( __any_sync( 0xFFFFFFFFU, hits != 0 ) ) {
                verifyWarp( data, nbytes, v, hits, which, found, foundCount, capacity );
            }
        }
    }

    /* ragged tail: the up to 4 words (last one possibly partial) behind the last full 16-byte vector, by the first warp */
    if ( tid < 32U ) {
        for ( uint64_t byte = nVectors * 16U; byte < nbytes; byte += 4U ) {
            uint32_t key = data[byte];
            if ( byte + 1U < nbytes ) {
                key |= static_cast<uint32_t>( data[byte + 1U] ) << 8;
            }
            if ( ( bitmap[key >> 5] >> ( key & 31U ) ) & 1U ) {
                verifyAnchor( data, nbytes, byte, which, found, foundCount, capacity );
            }
        }
    }
}
}  // namespace

void
launchFindMagics( const uint8_t* data, uint64_t nbytes, uint32_t which, uint64_t* found, uint32_t* foundCount,
                  uint32_t capacity, cudaStream_t stream )
{
    const uint64_t nVectors = nbytes / 16U;
    const uint64_t wanted = ( nVectors + FINDER_THREADS * FINDER_UNROLL - 1 ) / ( FINDER_THREADS * FINDER_UNROLL );
    const uint64_t maxBlocks = static_cast<uint64_t>( smCount() ) * 8U;
    const uint32_t blocks = static_cast<uint32_t>( wanted < 1 ? 1 : ( wanted > maxBlocks ? maxBlocks : wanted ) );
    findMagicsKernel<<<blocks, FINDER_THREADS, 0, stream>>>( data, nbytes, which, found, foundCount, capacity );
}
}  // namespace ibz2
