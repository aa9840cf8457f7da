/*
 * K4a -- inverse-BWT preparation: the stable counting-sort scatter  T[cum[L[i]]++] = i, packed as
 *   dbuf[k] = ( T[k] << 8 ) | L[T[k]]
 * (the reference packs ( T[k] << 8 ) | L[k]; carrying the byte of the position the entry points TO saves the
 * traversal's first hop and needs no second random access here, since L[T[k]] = L[i] is the byte in hand),
 * i.e. BurrowsWheelerTransformData::prepare (indexed_bzip2/bzip2.hpp:810-847) including the 256-bin byte histogram
 * that the reference accumulates during readBlockData (bzip2.hpp:698,759,788).  One CTA per block.
 *
 * A direct scatter writes 4 bytes at a time into 256 slowly advancing streams per block; with hundreds of blocks in
 * flight the half-written 32-byte sectors fall out of L2 and every store turns into a DRAM read-modify-write (measured:
 * 64 B of DRAM traffic per 4-byte entry).  So the scatter is staged through shared memory, tile by tile:
 *   pass 1  histogram of the whole block (warp-aggregated per row, so runs of equal bytes cost one add per row)
 *           -> exclusive scan -> running[b] = where the next entry of byte b goes;
 *   pass 2  for every tile of TILE symbols, in order: stable local counting sort into shared memory (rank inside the
 *           warp row from ballots, per-warp counters, scan over warps and bins), then the sorted tile is written out
 *           run by run: entries of one byte value are contiguous both in shared memory and in dbuf, so the stores of
 *           a warp cover whole sectors; running[b] advances by the tile's count.
 * Algorithmic HBM traffic per block: N bytes read twice (second pass hits L2) + 4N bytes written.
 */
#include "common.cuh"

namespace ibz2
{
namespace
{
constexpr int PREP_THREADS = 1024;
constexpr int PREP_WARPS = PREP_THREADS / 32;
constexpr uint32_t ROWS = 8;                            /* rows of 32 symbols per warp and tile */
constexpr uint32_t TILE = PREP_THREADS * ROWS;          /* 8192 symbols */
constexpr uint32_t FULL = 0xFFFFFFFFU;
constexpr uint32_t INVALID_KEY = 256;

struct PrepShared
{
    uint32_t sorted[TILE];                /* the tile in (byte, position) order */
    uint16_t warpCount[PREP_WARPS][256];  /* per tile: symbols of byte b seen by warp w, then exclusive over the warps */
    uint32_t running[256];                /* next dbuf index of byte b */
    uint32_t localOff[256];               /* first index of byte b inside sorted[] */
    uint32_t blockHist[256];
    uint32_t warpSums[8];
};

/* Lanes holding the same 9-bit key (what match.any returns), from one ballot per key bit: match.any itself costs a trip
 * per DISTINCT value in the warp, which is 32 trips per row on high-entropy data. */
template<uint32_t BITS>
__device__ __forceinline__ uint32_t
matchKey( uint32_t key )
{
    uint32_t same = FULL;
#pragma unroll
    for ( uint32_t bit = 0; bit < BITS; ++bit ) {
        const bool set = ( ( key >> bit ) & 1U ) != 0;
        const uint32_t votes = __ballot_sync( FULL, set );
        same &= set ? votes : ~votes;
    }
    return same;
}

/* exclusive scan over 256 values held by threads 0..255 (one per thread); needs two __syncthreads by all threads */
__device__ __forceinline__ uint32_t
scan256( uint32_t value, uint32_t tid, uint32_t* warpSums )
{
    uint32_t incl = value;
#pragma unroll
    for ( int d = 1; d < 32; d <<= 1 ) {
        const uint32_t up = __shfl_up_sync( FULL, incl, d );
        if ( ( tid & 31U ) >= static_cast<uint32_t>( d ) ) {
            incl += up;
        }
    }
    if ( ( tid < 256U ) && ( ( tid & 31U ) == 31U ) ) {
        warpSums[tid >> 5] = incl;
    }
    __syncthreads();
    uint32_t base = incl - value;
    if ( tid < 256U ) {
        for ( uint32_t w = 0; w < ( tid >> 5 ); ++w ) {
            base += warpSums[w];
        }
    }
    __syncthreads();
    return base;
}

__global__ void __launch_bounds__( PREP_THREADS, 2 )
bwtPrepareKernel( const uint8_t* __restrict__ lbytesAll, uint32_t* __restrict__ dbufAll,
                  const BlockMeta* __restrict__ meta )
{
    extern __shared__ __align__( 16 ) uint8_t sharedRaw[];
    PrepShared& sh = *reinterpret_cast<PrepShared*>( sharedRaw );

    const BlockMeta& bm = meta[blockIdx.x];
    if ( ( bm.status != IBZ2_ST_OK ) || ( bm.n == 0 ) ) {
        return;
    }
    const uint32_t n = bm.n;
    const uint8_t* __restrict__ lbytes = lbytesAll + static_cast<uint64_t>( blockIdx.x ) * LBYTES_STRIDE;
    uint32_t* __restrict__ dbuf = dbufAll + static_cast<uint64_t>( blockIdx.x ) * DBUF_STRIDE;
    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31U;
    const uint32_t warp = tid >> 5;
    const uint32_t ltMask = ( 1U << lane ) - 1U;

    if ( tid < 256U ) {
        sh.blockHist[tid] = 0;
    }
    __syncthreads();

    /* ---- pass 1: byte histogram of the block (bzip2.hpp:759,788) ----
     * L is the last column of a BWT: equal bytes come in runs.  A lane adds the lengths of the runs inside its own four
     * bytes (one to four shared-memory adds), a warp whose 128 bytes are one run adds once. */
    for ( uint32_t base = 0; base < n; base += PREP_THREADS * 4U ) {
        const uint32_t i = base + tid * 4U;
        /* LBYTES_STRIDE pads the last word of a block */
        const uint32_t word = i < n ? *reinterpret_cast<const uint32_t*>( lbytes + i ) : 0U;
        const uint32_t key0 = word & 0xFFU;
        /* run heads and lengths from the three neighbour comparisons, no loop */
        const bool whole = i + 4U <= n;
        const uint32_t x = word ^ ( word >> 8 );
        const bool e1 = ( x & 0xFFU ) == 0;
        const bool e2 = ( x & 0xFF00U ) == 0;
        const bool e3 = ( x & 0xFF0000U ) == 0;
        /* a run that covers the warp's 128 bytes is one add instead of 32 adds to the same address */
        const uint32_t first = __shfl_sync( FULL, word, 0 );
        if ( __all_sync( FULL, whole && e1 && e2 && e3 && ( word == first ) ) ) {
            if ( lane == 0 ) {
                atomicAdd( &sh.blockHist[key0], 128U );
            }
            continue;
        }
        if ( whole ) {
            const uint32_t len3 = 1U;
            const uint32_t len2 = 1U + ( e3 ? len3 : 0U );
            const uint32_t len1 = 1U + ( e2 ? len2 : 0U );
            const uint32_t len0 = 1U + ( e1 ? len1 : 0U );
            atomicAdd( &sh.blockHist[key0], len0 );
            if ( !e1 ) {
                atomicAdd( &sh.blockHist[( word >> 8 ) & 0xFFU], len1 );
            }
            if ( !e2 ) {
                atomicAdd( &sh.blockHist[( word >> 16 ) & 0xFFU], len2 );
            }
            if ( !e3 ) {
                atomicAdd( &sh.blockHist[word >> 24], len3 );
            }
        } else {
            for ( uint32_t k = 0; i + k < n; ++k ) {
                atomicAdd( &sh.blockHist[( word >> ( 8U * k ) ) & 0xFFU], 1U );
            }
        }
    }
    __syncthreads();
    {
        /* cumulative byte counts (bzip2.hpp:814-823) */
        const uint32_t mine = tid < 256U ? sh.blockHist[tid] : 0U;
        const uint32_t base = scan256( mine, tid, sh.warpSums );
        if ( tid < 256U ) {
            sh.running[tid] = base;
        }
    }
    __syncthreads();

    /* ---- pass 2: tile-wise stable scatter ---- */
    for ( uint32_t tileBase = 0; tileBase < n; tileBase += TILE ) {
        {
            uint32_t* words = reinterpret_cast<uint32_t*>( &sh.warpCount[0][0] );
            for ( uint32_t i = tid; i < PREP_WARPS * 256U / 2U; i += PREP_THREADS ) {
                words[i] = 0;
            }
        }
        __syncthreads();

        /* a) warp w owns symbols tileBase + 256 w .. + 255 as ROWS rows of 32; rank of each symbol among the equal
         *    bytes its warp has seen so far */
        const uint32_t first = tileBase + warp * ( 32U * ROWS ) + lane;
        uint32_t keyRank[ROWS];   /* key | rank << 16 */
        uint16_t* const myCount = sh.warpCount[warp];
#pragma unroll
        for ( uint32_t r = 0; r < ROWS; ++r ) {
            const uint32_t i = first + 32U * r;
            const uint32_t key = i < n ? lbytes[i] : INVALID_KEY;
            /* L comes in runs: a row of one byte value needs no ballots at all; a row inside the block has 8-bit keys */
            const uint32_t firstKey = __shfl_sync( FULL, key, 0 );
            const uint32_t same = __all_sync( FULL, key == firstKey ) ? FULL
                                  : ( ( first - lane + 32U * r + 32U <= n ) ? matchKey<8>( key ) : matchKey<9>( key ) );
            const uint32_t leader = static_cast<uint32_t>( __ffs( static_cast<int>( same ) ) ) - 1U;
            uint32_t before = 0;
            if ( ( lane == leader ) && ( key != INVALID_KEY ) ) {
                before = myCount[key];
                myCount[key] = static_cast<uint16_t>( before + static_cast<uint32_t>( __popc( same ) ) );
            }
            before = __shfl_sync( FULL, before, leader );
            keyRank[r] = key | ( ( before + static_cast<uint32_t>( __popc( same & ltMask ) ) ) << 16 );
            __syncwarp();
        }
        __syncthreads();

        /* b) per byte value: exclusive scan over the warps, tile count, exclusive scan over the byte values */
        uint32_t tileCount = 0;
        if ( tid < 256U ) {
#pragma unroll 8
            for ( uint32_t w = 0; w < static_cast<uint32_t>( PREP_WARPS ); ++w ) {
                const uint32_t c = sh.warpCount[w][tid];
                sh.warpCount[w][tid] = static_cast<uint16_t>( tileCount );
                tileCount += c;
            }
        }
        const uint32_t off = scan256( tileCount, tid, sh.warpSums );
        if ( tid < 256U ) {
            sh.localOff[tid] = off;
        }
        __syncthreads();

        /* c) stable placement inside the tile */
#pragma unroll
        for ( uint32_t r = 0; r < ROWS; ++r ) {
            const uint32_t key = keyRank[r] & 0xFFFFU;
            if ( key != INVALID_KEY ) {
                const uint32_t pos = sh.localOff[key] + sh.warpCount[warp][key] + ( keyRank[r] >> 16 );
                sh.sorted[pos] = ( ( first + 32U * r ) << 8 ) | key;
            }
        }
        __syncthreads();

        /* d) write the runs out: neighbouring threads write neighbouring entries of one byte value */
        const uint32_t tileN = min( TILE, n - tileBase );
        for ( uint32_t pos = tid; pos < tileN; pos += PREP_THREADS ) {
            const uint32_t e = sh.sorted[pos];
            const uint32_t key = e & 0xFFU;
            dbuf[sh.running[key] + ( pos - sh.localOff[key] )] = e;
        }
        __syncthreads();
        if ( tid < 256U ) {
            sh.running[tid] += tileCount;
        }
        /* the next iteration's first __syncthreads orders this update before any use */
    }
}
}  // namespace

void
launchBwtPrepare( const uint8_t* lbytes, uint32_t* dbuf, BlockMeta* meta, uint32_t nblocks, cudaStream_t stream )
{
    cudaFuncSetAttribute( bwtPrepareKernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                          static_cast<int>( sizeof( PrepShared ) ) );
    bwtPrepareKernel<<<nblocks, PREP_THREADS, sizeof( PrepShared ), stream>>>( lbytes, dbuf, meta );
}
}  // namespace ibz2
