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

#include <algorithm>
#include <cstdio>
#include <mutex>

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

/* ---------------------------------------------------------------------------------------------------------------
 * ibz2_find_magics_noalloc -- K1 for callers that must not allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The match list and its counter, which ibz2_find_magics keeps in cudaMalloc'ed memory of its context,
 * live in a file-scope device array of fixed size instead: every GPU that loads the library holds these 512.25 KB
 * until the process exits.  All calls on one device share the array, so calls that overlap on the GPU race; the
 * counter is cleared at the start of every call, the match slots keep the last call's contents behind the new count.
 * --------------------------------------------------------------------------------------------------------------- */
namespace ibz2
{
namespace
{
constexpr uint32_t NOALLOC_COUNTER_BYTES = 256;   /* the counter gets an aligned slot of its own */
constexpr uint64_t NOALLOC_SCRATCH_BYTES = NOALLOC_COUNTER_BYTES + static_cast<uint64_t>( IBZ2_NOALLOC_MAX_MATCHES ) * sizeof( uint64_t );
constexpr int NOALLOC_MAX_DEVICES = 64;

static __device__ __align__( 256 ) uint8_t g_noallocScratch[NOALLOC_SCRATCH_BYTES];

/* In host code g_noallocScratch is a stand-in, not a device address: what a kernel gets is what cudaGetSymbolAddress
 * returned on that device, resolved once per device with that device current. */
struct NoallocDevice
{
    uint8_t* scratch{ nullptr };
    int sms{ 0 };
    bool ready{ false };
};
NoallocDevice g_noallocDevices[NOALLOC_MAX_DEVICES];
std::mutex g_noallocMutex;
thread_local char g_noallocMessage[320];

int
noallocFail( int rc, const char* what, const char* detail )
{
    std::snprintf( g_noallocMessage, sizeof( g_noallocMessage ), "ibz2_find_magics_noalloc: %s%s%s", what,
                   detail != nullptr ? ": " : "", detail != nullptr ? detail : "" );
    setContextFreeError( g_noallocMessage );
    return rc;
}

/* the caller's current device is put back when the call returns */
struct NoallocDeviceGuard
{
    int previous{ -1 };
    bool switched{ false };
    ~NoallocDeviceGuard()
    {
        if ( switched ) {
            cudaSetDevice( previous );
        }
    }
};

int
noallocSetUp( int device, NoallocDevice** out )
{
    const std::lock_guard<std::mutex> lock( g_noallocMutex );
    NoallocDevice& d = g_noallocDevices[device];
    if ( !d.ready ) {
        void* address = nullptr;
        cudaError_t err = cudaGetSymbolAddress( &address, g_noallocScratch );
        if ( ( err != cudaSuccess ) || ( address == nullptr ) ) {
            cudaGetLastError();
            return noallocFail( IBZ2_E_CUDA, "cudaGetSymbolAddress of the scratch array failed", cudaGetErrorString( err ) );
        }
        int sms = 0;
        err = cudaDeviceGetAttribute( &sms, cudaDevAttrMultiProcessorCount, device );
        if ( ( err != cudaSuccess ) || ( sms <= 0 ) ) {
            cudaGetLastError();
            return noallocFail( IBZ2_E_CUDA, "cannot read the SM count", cudaGetErrorString( err ) );
        }
        d.scratch = static_cast<uint8_t*>( address );
        d.sms = sms;
        d.ready = true;
    }
    *out = &d;
    return IBZ2_OK;
}
}  // namespace
}  // namespace ibz2

extern "C" int
ibz2_find_magics_noalloc( int device, void* stream, const void* device_data, uint64_t nbytes, uint64_t base_bit_offset,
                          uint32_t which, uint64_t* offsets, uint64_t capacity, uint64_t* count )
{
    using namespace ibz2;
    if ( ( count == nullptr ) || ( ( device_data == nullptr ) && ( nbytes > 0 ) )
         || ( ( which & ( IBZ2_FIND_BLOCK | IBZ2_FIND_EOS ) ) == 0 ) || ( ( offsets == nullptr ) && ( capacity > 0 ) ) ) {
        return noallocFail( IBZ2_E_INVALID_ARGUMENT, "invalid argument", nullptr );
    }
    *count = 0;
    if ( capacity > IBZ2_NOALLOC_MAX_MATCHES ) {
        char detail[160];
        std::snprintf( detail, sizeof( detail ), "%llu entries asked for, the static scratch array holds IBZ2_NOALLOC_MAX_MATCHES = %u "
                       "(use ibz2_find_magics for more)", static_cast<unsigned long long>( capacity ), IBZ2_NOALLOC_MAX_MATCHES );
        return noallocFail( IBZ2_E_CAPACITY, "capacity beyond the maximum", detail );
    }
    if ( ( device < 0 ) || ( device >= NOALLOC_MAX_DEVICES ) ) {
        return noallocFail( IBZ2_E_INVALID_ARGUMENT, "device index out of range", nullptr );
    }
    NoallocDeviceGuard guard;
    cudaError_t err = cudaGetDevice( &guard.previous );
    if ( ( err == cudaSuccess ) && ( guard.previous != device ) ) {
        err = cudaSetDevice( device );
        guard.switched = err == cudaSuccess;
    }
    if ( err != cudaSuccess ) {
        cudaGetLastError();
        return noallocFail( IBZ2_E_NO_DEVICE, "cannot make the device current (there is no CPU fallback)", cudaGetErrorString( err ) );
    }
    if ( nbytes < 6 ) {
        return IBZ2_OK;
    }
    {
        /* a host pointer would make the kernel fault (and the context unusable): look before launching */
        cudaPointerAttributes attributes{};
        err = cudaPointerGetAttributes( &attributes, device_data );
        if ( ( err != cudaSuccess ) || ( ( attributes.type != cudaMemoryTypeDevice ) && ( attributes.type != cudaMemoryTypeManaged ) ) ) {
            cudaGetLastError();
            return noallocFail( IBZ2_E_INVALID_ARGUMENT, "device_data is not device memory (this entry point stages nothing)", nullptr );
        }
    }
    NoallocDevice* d = nullptr;
    const int rc = noallocSetUp( device, &d );
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    const cudaStream_t s = static_cast<cudaStream_t>( stream );
    uint32_t* const dCount = reinterpret_cast<uint32_t*>( d->scratch );
    uint64_t* const dFound = reinterpret_cast<uint64_t*>( d->scratch + NOALLOC_COUNTER_BYTES );
    const uint32_t slots = static_cast<uint32_t>( capacity );

    /* the array keeps the last call's contents: the counter must start at zero */
    if ( ( err = cudaMemsetAsync( dCount, 0, sizeof( uint32_t ), s ) ) != cudaSuccess ) {
        return noallocFail( IBZ2_E_CUDA, "cudaMemsetAsync", cudaGetErrorString( err ) );
    }
    const uint64_t nVectors = nbytes / 16U;
    const uint64_t wanted = ( nVectors + FINDER_THREADS * FINDER_UNROLL - 1 ) / ( FINDER_THREADS * FINDER_UNROLL );
    const uint64_t maxBlocks = static_cast<uint64_t>( d->sms ) * 8U;
    const uint32_t blocks = static_cast<uint32_t>( wanted < 1 ? 1 : ( wanted > maxBlocks ? maxBlocks : wanted ) );
    findMagicsKernel<<<blocks, FINDER_THREADS, 0, s>>>( static_cast<const uint8_t*>( device_data ), nbytes, which, dFound,
                                                        dCount, slots );
    if ( ( err = cudaGetLastError() ) != cudaSuccess ) {
        return noallocFail( IBZ2_E_CUDA, "kernel launch", cudaGetErrorString( err ) );
    }
    uint32_t found = 0;
    if ( ( ( err = cudaMemcpyAsync( &found, dCount, sizeof( uint32_t ), cudaMemcpyDeviceToHost, s ) ) != cudaSuccess )
         || ( ( err = cudaStreamSynchronize( s ) ) != cudaSuccess ) ) {
        return noallocFail( IBZ2_E_CUDA, "reading the match count", cudaGetErrorString( err ) );
    }
    *count = found;
    const uint64_t n = std::min<uint64_t>( found, capacity );
    if ( n > 0 ) {
        if ( ( ( err = cudaMemcpyAsync( offsets, dFound, n * sizeof( uint64_t ), cudaMemcpyDeviceToHost, s ) ) != cudaSuccess )
             || ( ( err = cudaStreamSynchronize( s ) ) != cudaSuccess ) ) {
            return noallocFail( IBZ2_E_CUDA, "reading the matches", cudaGetErrorString( err ) );
        }
        /* ascending by bit offset like ibz2_find_magics (in place: std::sort allocates nothing) */
        std::sort( offsets, offsets + n, [] ( uint64_t a, uint64_t b ) {
            return ( a & ~IBZ2_OFFSET_EOS_FLAG ) < ( b & ~IBZ2_OFFSET_EOS_FLAG );
        } );
        for ( uint64_t i = 0; i < n; ++i ) {
            offsets[i] = ( ( offsets[i] & ~IBZ2_OFFSET_EOS_FLAG ) + base_bit_offset ) | ( offsets[i] & IBZ2_OFFSET_EOS_FLAG );
        }
    }
    if ( found > capacity ) {
        return noallocFail( IBZ2_E_CAPACITY, "more matches than capacity (offsets holds `capacity` of them, not necessarily the first)", nullptr );
    }
    return IBZ2_OK;
}
