/*
 * K4a -- inverse-BWT preparation: the stable counting-sort scatter  T[cum[L[i]]++] = i, packed as
 *   dbuf[k] = ( T[k] << 8 ) | L[T[k]]
 * (the reference packs ( T[k] << 8 ) | L[k]; carrying the byte of the position the entry points TO saves the
 * traversal's first hop and needs no second random access here, since L[T[k]] = L[i] is the byte in hand),
 * i.e. BurrowsWheelerTransformData::prepare (indexed_bzip2/bzip2.hpp:810-847) including the 256-bin byte histogram
 * that the reference accumulates during readBlockData (bzip2.hpp:698,759,788).  One CTA per block.
 *
 * Stability is what makes the scatter order-dependent; it is obtained without atomics by giving each of the 256
 * threads a contiguous slice of L and a private 256-bin counter row in shared memory:
 *   phase 1  every thread histograms its slice (u16 counters, thread-major rows -> bank = bin/2, conflict-light);
 *   phase 2  thread b turns column b into exclusive prefixes: u16 within groups of 16 threads plus a u32 base per
 *            group and bin (a slice holds <= 3520 symbols, 16 slices < 65536);
 *   phase 3  every thread replays its slice and scatters ( i << 8 ) | byte to base + running count.
 * Algorithmic HBM traffic per block: N bytes read twice (second pass hits L2) + 4N bytes written.
 */
#include "common.cuh"

namespace ibz2
{
namespace
{
constexpr int PREP_THREADS = 256;
constexpr int GROUP_THREADS = 16;
constexpr int GROUPS = PREP_THREADS / GROUP_THREADS;

struct PrepShared
{
    uint16_t hist[PREP_THREADS][256]; /* [thread][bin] */
    uint32_t groupBase[GROUPS][256];  /* [group][bin] */
    uint32_t warpSums[PREP_THREADS / 32];
};

__global__ void __launch_bounds__( PREP_THREADS )
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

    /* slice of this thread: [begin, end), begin a multiple of 16 */
    const uint32_t chunk = ( ( ( n + PREP_THREADS - 1 ) / PREP_THREADS ) + 15U ) & ~15U;
    const uint32_t begin = min( n, tid * chunk );
    const uint32_t end = min( n, begin + chunk );

    {
        uint32_t* words = reinterpret_cast<uint32_t*>( &sh.hist[0][0] );
        for ( uint32_t i = tid; i < PREP_THREADS * 256U / 2U; i += PREP_THREADS ) {
            words[i] = 0;
        }
    }
    __syncthreads();

    /* phase 1: private histogram */
    uint16_t* const myHist = sh.hist[tid];
    for ( uint32_t i = begin; i < end; i += 16U ) {
        const uint4 v = *reinterpret_cast<const uint4*>( lbytes + i );
        const uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
        for ( uint32_t k = 0; k < 16U; ++k ) {
            if ( i + k < end ) {
                const uint32_t b = ( w[k >> 2] >> ( 8U * ( k & 3U ) ) ) & 0xFFU;
                myHist[b] = static_cast<uint16_t>( myHist[b] + 1U );
            }
        }
    }
    __syncthreads();

    /* phase 2: column tid -> exclusive prefixes over the threads */
    uint32_t total = 0;
    for ( int g = 0; g < GROUPS; ++g ) {
        uint32_t inGroup = 0;
#pragma unroll
        for ( int r = 0; r < GROUP_THREADS; ++r ) {
            const int t = g * GROUP_THREADS + r;
            const uint32_t c = sh.hist[t][tid];
            sh.hist[t][tid] = static_cast<uint16_t>( inGroup );
            inGroup += c;
        }
        sh.groupBase[g][tid] = total;
        total += inGroup;
    }
    /* exclusive scan of the bin totals over the 256 bins (cumulative byte counts, bzip2.hpp:814-823) */
    uint32_t incl = total;
#pragma unroll
    for ( int d = 1; d < 32; d <<= 1 ) {
        const uint32_t up = __shfl_up_sync( 0xFFFFFFFFU, incl, d );
        if ( ( tid & 31U ) >= static_cast<uint32_t>( d ) ) {
            incl += up;
        }
    }
    if ( ( tid & 31U ) == 31U ) {
        sh.warpSums[tid >> 5] = incl;
    }
    __syncthreads();
    uint32_t binBase = incl - total;
    for ( uint32_t w = 0; w < ( tid >> 5 ); ++w ) {
        binBase += sh.warpSums[w];
    }
    for ( int g = 0; g < GROUPS; ++g ) {
        sh.groupBase[g][tid] += binBase;
    }
    __syncthreads();

    /* phase 3: replay the slice and scatter */
    const uint32_t* const myBase = &sh.groupBase[tid / GROUP_THREADS][0];
    for ( uint32_t i = begin; i < end; i += 16U ) {
        const uint4 v = *reinterpret_cast<const uint4*>( lbytes + i );
        const uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
        for ( uint32_t k = 0; k < 16U; ++k ) {
            if ( i + k < end ) {
                const uint32_t b = ( w[k >> 2] >> ( 8U * ( k & 3U ) ) ) & 0xFFU;
                const uint32_t r = myHist[b];
                myHist[b] = static_cast<uint16_t>( r + 1U );
                dbuf[myBase[b] + r] = ( ( i + k ) << 8 ) | b;
            }
        }
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
