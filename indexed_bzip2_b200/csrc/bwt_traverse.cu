/*
 * K4b -- inverse-BWT traversal: the N-step walk  pos = T[pos], emit L[pos]  that starts at T[origPtr]
 * (with K4a's packing, entry k carries L[T[k]], so the walk reads entry origPtr first: p0 = origPtr)
 * (BurrowsWheelerTransformData::decodeBlock, indexed_bzip2/bzip2.hpp:863-879, without its RLE1 / CRC half).
 *
 * The walk is a dependent-load chain, so it is broken up by list ranking.  One CTA per block:
 *   pass A  every position that is a multiple of SPACING is a "splitter".  Each thread walks from its splitters until
 *           the next splitter, recording the segment's length and successor.  Segment lengths are geometric, so each
 *           lane keeps walking its own list of splitters (no waiting for the slowest lane).  The segment that passes
 *           through the start position p0 is remembered (sstar, kstar).
 *   rank    the successor list of the <= 14063 segments is ranked in shared memory by pointer jumping (Wyllie), with the
 *           list cut at sstar; that gives every segment on p0's cycle its output offset.
 *   copy    the bytes a thread met in pass A sit, in walking order, in the thread's own staging stream (STAGE_CAP bytes
 *           per thread in HBM, written four at a time); once the ranking has given every segment its output offset the
 *           warps copy the segments from the streams to their final positions, 32 consecutive bytes per step.  This
 *           replaces a second walk: a walk costs one 32-byte HBM sector per step (T vectors of the blocks in flight are
 *           far larger than L2), the copy moves each byte once, coalesced.
 *   pass B  only for what did not fit a staging stream (sums of ~14 geometric segment lengths rarely exceed it; runs of
 *           pathological T vectors can): those segments are walked again and written directly.
 *   fill    if p0's cycle is shorter than N (periodic data such as b"A" * n, where T splits into many short cycles) the
 *           walk laps its cycle; the remaining output is the periodic continuation of the first lap.
 * A block whose start cycle contains no splitter at all (only tiny cycles can) is walked serially by one thread.
 *
 * The output is the block's byte string BEFORE RLE1 undo ("prerle"), N bytes, written over the block's L buffer.
 * Algorithmic HBM traffic per block: 4N bytes of random 32-bit reads + N bytes staged, read back and written.
 */
#include "common.cuh"

namespace ibz2
{
namespace
{
constexpr int TRAV_THREADS = 1024;
constexpr uint32_t SPACING = 64;
constexpr uint32_t SPACING_SHIFT = 6;
constexpr uint32_t MAX_SEGMENTS = ( DBUF_SIZE + SPACING - 1 ) / SPACING; /* 14063 */
constexpr uint32_t PER_THREAD = ( MAX_SEGMENTS + TRAV_THREADS - 1 ) / TRAV_THREADS;
constexpr uint32_t NIL = 0xFFFFU;
constexpr uint32_t FULL = 0xFFFFFFFFU;
constexpr uint32_t STAGE_CAP = TRAV_STAGE_CAP;   /* bytes of staging stream per thread */

struct TraverseShared
{
    uint32_t dist[MAX_SEGMENTS + 1];
    uint16_t jump[MAX_SEGMENTS + 1];
    int sstar;
    uint32_t kstar;
};

__global__ void __launch_bounds__( TRAV_THREADS, 2 )
bwtTraverseKernel( const uint32_t* __restrict__ dbufAll, uint8_t* __restrict__ prerleAll,
                   BlockMeta* __restrict__ meta, uint8_t* __restrict__ stageAll )
{
    extern __shared__ __align__( 16 ) uint8_t sharedRaw[];
    TraverseShared& sh = *reinterpret_cast<TraverseShared*>( sharedRaw );

    BlockMeta& bm = meta[blockIdx.x];
    if ( ( bm.status != IBZ2_ST_OK ) || ( bm.n == 0 ) ) {
        return;
    }
    const uint32_t n = bm.n;
    const uint32_t* __restrict__ dbuf = dbufAll + static_cast<uint64_t>( blockIdx.x ) * DBUF_STRIDE;
    uint8_t* __restrict__ prerle = prerleAll + static_cast<uint64_t>( blockIdx.x ) * LBYTES_STRIDE;
    const uint32_t tid = threadIdx.x;
    const uint32_t segments = ( n + SPACING - 1 ) >> SPACING_SHIFT;
    uint8_t* const stageBlock = stageAll + static_cast<uint64_t>( blockIdx.x ) * TRAV_STAGE_STRIDE;
    uint32_t* const myStage = reinterpret_cast<uint32_t*>( stageBlock + static_cast<uint64_t>( tid ) * STAGE_CAP );
    uint16_t* const segLen = reinterpret_cast<uint16_t*>( stageBlock + static_cast<uint64_t>( TRAV_THREADS ) * STAGE_CAP );
    /* K4a stores ( T[k] << 8 ) | L[T[k]] (the byte of the NEXT position) instead of the reference's ( T[k] << 8 ) | L[k],
     * so the walk starts at origPtr itself and every load yields the output byte and the next position at once
     * (the reference first hops to T[origPtr], bzip2.hpp:835-838, and reads L at the position it arrives at). */
    const uint32_t p0 = bm.origPtr;

    if ( tid == 0 ) {
        sh.sstar = -1;
        sh.kstar = 0;
    }
    __syncthreads();

    /* ---- pass A: segment lengths and successors; the bytes go to the thread's staging stream ---- */
    {
        uint32_t s = tid;
        uint32_t pos = s << SPACING_SHIFT;
        uint32_t steps = 0;
        uint32_t cursor = 0;   /* bytes this thread has met so far */
        uint32_t acc = 0;
        while ( s < segments ) {
            if ( pos == p0 ) {
                sh.sstar = static_cast<int>( s );
                sh.kstar = steps;
            }
            const uint32_t e = __ldcg( dbuf + pos );
            acc |= ( e & 0xFFU ) << ( 8U * ( cursor & 3U ) );
            ++cursor;
            if ( ( cursor & 3U ) == 0 ) {
                if ( cursor <= STAGE_CAP ) {
                    myStage[( cursor >> 2 ) - 1U] = acc;
                }
                acc = 0;
            }
            pos = e >> 8;
            ++steps;
            if ( ( pos & ( SPACING - 1 ) ) == 0 ) {
                sh.jump[s] = static_cast<uint16_t>( pos >> SPACING_SHIFT );
                sh.dist[s] = steps;
                segLen[s] = static_cast<uint16_t>( min( steps, 0xFFFFU ) );
                s += TRAV_THREADS;
                pos = s << SPACING_SHIFT;
                steps = 0;
            }
        }
        if ( ( ( cursor & 3U ) != 0 ) && ( cursor < STAGE_CAP ) ) {
            myStage[cursor >> 2] = acc;
        }
    }
    __syncthreads();

    const int sstar = sh.sstar;
    uint32_t cycleLen = 0;

    if ( sstar < 0 ) {
        /* no splitter on the start cycle: serial walk of one lap by one thread */
        if ( tid == 0 ) {
            uint32_t pos = p0;
            uint32_t k = 0;
            do {
                const uint32_t e = __ldcg( dbuf + pos );
                prerle[k++] = static_cast<uint8_t>( e );
                pos = e >> 8;
            } while ( ( pos != p0 ) && ( k < n ) );
            sh.kstar = k;
        }
        __syncthreads();
        cycleLen = sh.kstar;
    } else {
        const uint32_t kstar = sh.kstar;
        const uint32_t lenStar = sh.dist[sstar];
        const uint32_t first = sh.jump[sstar];  /* the splitter reached from p0 */
        __syncthreads();
        if ( tid == 0 ) {
            sh.jump[sstar] = NIL;
            sh.dist[sstar] = 0;
        }
        __syncthreads();

        /* ---- pointer jumping: dist[s] = sum of the segment lengths from s up to (excluding) sstar ---- */
        for ( uint32_t span = 1; span < segments; span <<= 1 ) {
            uint32_t newDist[PER_THREAD];
            uint16_t newJump[PER_THREAD];
#pragma unroll
            for ( uint32_t i = 0; i < PER_THREAD; ++i ) {
                const uint32_t s = tid + i * TRAV_THREADS;
                if ( s < segments ) {
                    const uint32_t j = sh.jump[s];
                    uint32_t d = sh.dist[s];
                    uint32_t nj = j;
                    if ( j != NIL ) {
                        d += sh.dist[j];
                        nj = sh.jump[j];
                    }
                    newDist[i] = d;
                    newJump[i] = static_cast<uint16_t>( nj );
                }
            }
            __syncthreads();
#pragma unroll
            for ( uint32_t i = 0; i < PER_THREAD; ++i ) {
                const uint32_t s = tid + i * TRAV_THREADS;
                if ( s < segments ) {
                    sh.dist[s] = newDist[i];
                    sh.jump[s] = newJump[i];
                }
            }
            __syncthreads();
        }

        const uint32_t tailLen = lenStar - kstar;          /* p0 .. end of sstar's segment, output offset 0 */
        const uint32_t ranked = sh.dist[first];            /* 0 when first == sstar */
        cycleLen = ranked + lenStar;

        /* ---- copy: warp w moves the segments of its 32 threads from their staging streams to the output ---- */
        {
            const uint32_t lane = tid & 31U;
            const uint32_t warpFirst = tid & ~31U;
            uint32_t done = 0;   /* bytes of MY stream that belong to the segments handled so far */
            for ( uint32_t sBase = warpFirst; sBase < segments; sBase += TRAV_THREADS ) {
                /* the segments sBase .. sBase + 31 belong to lanes 0 .. 31 */
                const uint32_t s = sBase + lane;
                uint32_t len = 0;
                uint32_t base = 0;
                bool onCycle = false;   /* a segment that reaches sstar lies on p0's cycle */
                if ( s < segments ) {
                    len = segLen[s];
                    onCycle = sh.jump[s] == NIL;
                    base = tailLen + ranked - sh.dist[s];
                }
                const bool fits = ( len < 0xFFFFU ) && ( done + len <= STAGE_CAP );
                const uint32_t srcOffset = ( tid << TRAV_STAGE_CAP_SHIFT ) + done;
                done += len;
                uint32_t todo = __ballot_sync( FULL, onCycle && fits && ( len > 0 ) );
                while ( todo != 0 ) {
                    const uint32_t src = static_cast<uint32_t>( __ffs( static_cast<int>( todo ) ) ) - 1U;
                    todo &= todo - 1U;
                    const uint32_t l = __shfl_sync( FULL, len, src );
                    const uint32_t b = __shfl_sync( FULL, base, src );
                    const uint32_t o = __shfl_sync( FULL, srcOffset, src );
                    const bool star = static_cast<int>( sBase + src ) == sstar;
                    for ( uint32_t k = lane; k < l; k += 32U ) {
                        const uint32_t idx = ( star && ( k >= kstar ) ) ? ( k - kstar ) : ( b + k );
                        if ( idx < n ) {
                            prerle[idx] = stageBlock[o + k];
                        }
                    }
                }
                /* ---- pass B for the segments whose bytes did not fit the staging stream ---- */
                if ( onCycle && !fits ) {
                    const bool star = static_cast<int>( s ) == sstar;
                    uint32_t pos = s << SPACING_SHIFT;
                    uint32_t k = 0;
                    do {
                        const uint32_t e = __ldcg( dbuf + pos );
                        const uint32_t idx = ( star && ( k >= kstar ) ) ? ( k - kstar ) : ( base + k );
                        if ( idx < n ) {
                            prerle[idx] = static_cast<uint8_t>( e );
                        }
                        pos = e >> 8;
                        ++k;
                    } while ( ( pos & ( SPACING - 1 ) ) != 0 );
                }
            }
        }
        __syncthreads();
    }

    /* ---- periodic continuation when the start cycle is shorter than the block ---- */
    if ( cycleLen < n ) {
        for ( uint32_t j = cycleLen + tid; j < n; j += TRAV_THREADS ) {
            prerle[j] = prerle[j % cycleLen];
        }
    }
    if ( tid == 0 ) {
        bm.cycleLen = cycleLen;
    }
}
}  // namespace

void
launchBwtTraverse( const uint32_t* dbuf, uint8_t* prerle, BlockMeta* meta, uint32_t nblocks, uint8_t* stage,
                   cudaStream_t stream )
{
    cudaFuncSetAttribute( bwtTraverseKernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                          static_cast<int>( sizeof( TraverseShared ) ) );
    bwtTraverseKernel<<<nblocks, TRAV_THREADS, sizeof( TraverseShared ), stream>>>( dbuf, prerle, meta, stage );
}
}  // namespace ibz2
