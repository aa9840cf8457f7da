/*
 * K4b -- inverse-BWT traversal: the N-step walk  pos = T[pos], emit L[pos]  that starts at T[origPtr]
 * (with K4a's packing, entry k carries L[T[k]], so the walk reads entry origPtr first: p0 = origPtr)
 * (BurrowsWheelerTransformData::decodeBlock, indexed_bzip2/bzip2.hpp:863-879, without its RLE1 / CRC half).
 *
 * The walk is a dependent-load chain, so it is broken up by list ranking: every position that is a multiple of
 * TRAV_SPACING (16) is a "splitter", the stretch from one splitter to the next one met is a segment (geometric lengths,
 * mean TRAV_SPACING), segments are walked independently and put in order afterwards.
 *
 *   bwtWalkKernel   ONE persistent grid for all blocks of the (sub-)wave.  A random 4-byte read costs HBM a whole
 *                   128-byte line (measured: 114 B of DRAM traffic per entry when the T vectors of ~300 blocks are
 *                   walked at once), but all 32 entries of a line are wanted eventually -- so the walk keeps the T
 *                   vectors it is working on INSIDE L2: segments are handed out in block order from one global counter
 *                   (a warp takes TRAV_CHUNK at a time, its lanes take the next one whenever they finish, nobody waits
 *                   for a slow lane), so the blocks being walked are exponentially distributed behind the head of the
 *                   queue with a mean of threads in flight / segments per block = 151 552 / 56 250 = 2.7 blocks =
 *                   10 MB of T.  (With a splitter every 64 entries it was 5.4 blocks at half the threads, and the walk
 *                   took twice as long.)  What bounds it now is the L2 -> SM path: a 32-byte sector per 4-byte step.
 *                   The bytes a segment meets go to the segment's own slot (64 bytes, 16 at a time: two slots per
 *                   line), those of the 1.8 % of longer segments go on in a second, 128-byte slot; of a segment longer
 *                   than both only the position after 192 steps is kept.
 *   bwtRankKernel   one CTA per block.  56 250 segments do not fit a shared-memory ranking, so only every fourth
 *                   splitter is ranked: a thread follows the chain of segments from such a "coarse" splitter to the
 *                   next one (one uint2 load per hop), the <= 14 063 chains are ranked in shared memory by pointer
 *                   jumping (Wyllie), cut at the chain that passes through p0; that gives every chain on p0's cycle
 *                   its output offset.
 *   bwtPlaceKernel  one thread per chain: the slots of its segments go to one contiguous stretch of the output
 *                   (aligned 4-byte stores); what did not fit a slot is walked again.
 *   bwtFillKernel   if p0's cycle is shorter than N (periodic data such as b"A" * n, where T splits into many short
 *                   cycles) the walk laps its cycle; the remaining output is the periodic continuation of the first lap.
 * A block whose start cycle contains no coarse splitter at all (only short cycles can) is walked serially by one thread.
 *
 * The output is the block's byte string BEFORE RLE1 undo ("prerle"), N bytes, written over the block's L buffer.
 * Algorithmic HBM traffic per block: 4N bytes of T read once + N bytes staged, read back and written.
 */
#include "common.cuh"

#include <algorithm>
#include <cstdlib>

namespace ibz2
{
namespace
{
constexpr uint32_t SPACING = TRAV_SPACING;
constexpr uint32_t SPACING_SHIFT = TRAV_SPACING_SHIFT;
constexpr uint32_t MAX_SEGMENTS = TRAV_MAX_SEGMENTS;
/* ranking works on every RATIO-th splitter ("coarse" segments: few enough for shared memory), the walk on all of them */
constexpr uint32_t COARSE_SHIFT = TRAV_COARSE_SHIFT;
constexpr uint32_t RATIO_SHIFT = COARSE_SHIFT - SPACING_SHIFT;
constexpr uint32_t RATIO_MASK = ( 1U << RATIO_SHIFT ) - 1U;
constexpr uint32_t MAX_COARSE = ( DBUF_SIZE + ( 1U << COARSE_SHIFT ) - 1U ) >> COARSE_SHIFT;   /* 14063 */
constexpr uint32_t STAR_COARSE_SLOT = MAX_COARSE + 1U;
constexpr uint32_t STAR_STEP_SLOT = MAX_COARSE + 2U;
static_assert( STAR_STEP_SLOT < TRAV_SEG_CAP, "kept in the segBase array" );
constexpr uint32_t SLOT = TRAV_SLOT_BYTES;
constexpr uint32_t SLOT2 = TRAV_SLOT2_BYTES;
constexpr uint32_t STAGED_MAX = SLOT + SLOT2;   /* bytes of a segment that are staged; the rest is walked again */
constexpr uint32_t NIL = 0xFFFFU;
constexpr uint32_t NOT_ON_CYCLE = 0xFFFFFFFFU;
constexpr uint32_t FULL = 0xFFFFFFFFU;

constexpr uint32_t TRAV_CHUNK = 64;                                              /* segments a warp takes at a time */
constexpr uint32_t CHUNKS_PER_BLOCK = ( MAX_SEGMENTS + TRAV_CHUNK - 1U ) / TRAV_CHUNK;   /* 220 */
constexpr int WALK_THREADS_DEFAULT = 1024;                                       /* per SM */
constexpr uint32_t WALK_STEPS = 8;                                               /* steps per look at the pool */

struct BlockStage
{
    uint8_t* slots;      /* [segment][SLOT] */
    uint8_t* slots2;     /* [segment][SLOT2]: bytes SLOT .. SLOT + SLOT2 - 1 of the segments that have them */
    uint2* segLink;      /* x: steps from the segment's splitter to the next splitter, y: that splitter (segment index) */
    uint32_t* ovfPos;    /* position reached after STAGED_MAX steps (only meaningful when the segment is longer) */
    uint32_t* segBase;   /* bwtRankKernel: output offset of the first byte of every COARSE segment on p0's cycle,
                          * NOT_ON_CYCLE for the others; behind them the coarse segment whose chain passes through p0
                          * and the step at which it does */
};

__device__ __forceinline__ BlockStage
blockStage( uint8_t* stageAll, uint32_t block )
{
    uint8_t* const base = stageAll + TRAV_STAGE_HEADER + static_cast<uint64_t>( block ) * TRAV_STAGE_STRIDE;
    BlockStage st;
    st.slots = base;
    st.slots2 = base + TRAV_SLOT2_OFFSET;
    st.segLink = reinterpret_cast<uint2*>( base + TRAV_LINK_OFFSET );
    st.ovfPos = reinterpret_cast<uint32_t*>( base + TRAV_OVF_OFFSET );
    st.segBase = reinterpret_cast<uint32_t*>( base + TRAV_BASE_OFFSET );
    return st;
}

/* ------------------------------------------------------------------ walk ------------------------------------ */
__global__ void __launch_bounds__( 1024, 1 )
bwtWalkKernel( const uint32_t* __restrict__ dbufAll, BlockMeta* __restrict__ meta, uint8_t* __restrict__ stageAll,
               uint32_t nblocks )
{
    uint32_t* const chunkCounter = reinterpret_cast<uint32_t*>( stageAll );
    const uint32_t lane = threadIdx.x & 31U;
    const uint32_t lowerLanes = ( 1U << lane ) - 1U;
    const uint32_t totalChunks = nblocks * CHUNKS_PER_BLOCK;

    /* the warp's pool (warp-uniform): segments poolNext .. poolEnd - 1 of block poolBlock */
    uint32_t poolBlock = 0;
    uint32_t poolNext = 0;
    uint32_t poolEnd = 0;
    uint32_t poolP0 = 0;
    bool exhausted = false;

    /* the lane's segment */
    bool active = false;
    const uint32_t* dbuf = nullptr;
    uint4* slotQuads = nullptr;
    uint4* slot2Quads = nullptr;
    uint4 quad = make_uint4( 0, 0, 0, 0 );
    uint32_t myBlock = 0;
    uint32_t s = 0;
    uint32_t p0 = 0;
    uint32_t pos = 0;
    uint32_t steps = 0;
    uint32_t acc = 0;

    while ( true ) {
        uint32_t need = __ballot_sync( FULL, !active );
        while ( ( need != 0 ) && !exhausted ) {
            if ( poolNext == poolEnd ) {
                uint32_t c = 0;
                if ( lane == 0 ) {
                    c = atomicAdd( chunkCounter, 1U );
                }
                c = __shfl_sync( FULL, c, 0 );
                if ( c >= totalChunks ) {
                    exhausted = true;
                    break;
                }
                const uint32_t b = c / CHUNKS_PER_BLOCK;
                const uint32_t first = ( c - b * CHUNKS_PER_BLOCK ) * TRAV_CHUNK;
                const BlockMeta& bm = meta[b];
                if ( ( bm.status != IBZ2_ST_OK ) || ( bm.n == 0 ) ) {
                    continue;
                }
                const uint32_t segments = ( bm.n + SPACING - 1U ) >> SPACING_SHIFT;
                if ( first >= segments ) {
                    continue;
                }
                poolBlock = b;
                poolNext = first;
                poolEnd = min( first + TRAV_CHUNK, segments );
                poolP0 = bm.origPtr;
            }
            const uint32_t avail = poolEnd - poolNext;
            const uint32_t rank = __popc( need & lowerLanes );
            if ( ( ( ( need >> lane ) & 1U ) != 0 ) && ( rank < avail ) ) {
                s = poolNext + rank;
                myBlock = poolBlock;
                p0 = poolP0;
                dbuf = dbufAll + static_cast<uint64_t>( poolBlock ) * DBUF_STRIDE;
                {
                    uint8_t* const blockBase = stageAll + TRAV_STAGE_HEADER + static_cast<uint64_t>( poolBlock ) * TRAV_STAGE_STRIDE;
                    slotQuads = reinterpret_cast<uint4*>( blockBase + ( static_cast<uint64_t>( s ) << TRAV_SLOT_SHIFT ) );
                    slot2Quads = reinterpret_cast<uint4*>( blockBase + TRAV_SLOT2_OFFSET
                                                           + ( static_cast<uint64_t>( s ) << TRAV_SLOT2_SHIFT ) );
                }
                pos = s << SPACING_SHIFT;
                steps = 0;
                acc = 0;
                active = true;
            }
            poolNext += min( avail, static_cast<uint32_t>( __popc( need ) ) );
            need = __ballot_sync( FULL, !active );
        }
        if ( need == FULL ) {
            break;   /* nothing left to hand out and nobody walking */
        }
        /* K4a stores ( T[k] << 8 ) | L[T[k]] (the byte of the NEXT position) instead of the reference's
         * ( T[k] << 8 ) | L[k], so the walk starts at origPtr itself and every load yields the output byte and the
         * next position at once (the reference first hops to T[origPtr], bzip2.hpp:835-838).
         * WALK_STEPS steps between two looks at the pool: a lane that finishes early idles for at most seven steps (2 / 4 / 8 steps: 48.1 / 45.6 / 44.7 ms). */
#pragma unroll
        for ( uint32_t u = 0; u < WALK_STEPS; ++u ) {
            if ( active ) {
                if ( pos == p0 ) {
                    meta[myBlock].travStar = static_cast<int32_t>( s );
                    meta[myBlock].travK = steps;
                }
                const uint32_t e = __ldcg( dbuf + pos );
                acc |= ( e & 0xFFU ) << ( 8U * ( steps & 3U ) );
                ++steps;
                if ( ( steps & 3U ) == 0 ) {
                    /* four words make one 16-byte streaming store: the slots are written once and read once, they
                     * should not push T vectors out of L2 */
                    const uint32_t w = ( ( steps >> 2 ) - 1U ) & 3U;
                    quad.x = w == 0 ? acc : quad.x;
                    quad.y = w == 1 ? acc : quad.y;
                    quad.z = w == 2 ? acc : quad.z;
                    quad.w = w == 3 ? acc : quad.w;
                    if ( ( w == 3 ) && ( steps <= STAGED_MAX ) ) {
                        const uint32_t q = ( steps >> 4 ) - 1U;
                        __stcs( q < SLOT / 16U ? slotQuads + q : slot2Quads + ( q - SLOT / 16U ), quad );
                    }
                    acc = 0;
                }
                pos = e >> 8;
                if ( steps == STAGED_MAX ) {
                    blockStage( stageAll, myBlock ).ovfPos[s] = pos;
                }
                if ( ( pos & ( SPACING - 1U ) ) == 0 ) {
                    const BlockStage st = blockStage( stageAll, myBlock );
                    st.segLink[s] = make_uint2( steps, pos >> SPACING_SHIFT );
                    if ( ( ( steps & 15U ) != 0 ) && ( steps < STAGED_MAX ) ) {
                        /* the last, partly filled quad (what lies behind the segment's bytes is never read) */
                        const uint32_t w = ( steps >> 2 ) & 3U;
                        quad.x = w == 0 ? acc : quad.x;
                        quad.y = w == 1 ? acc : quad.y;
                        quad.z = w == 2 ? acc : quad.z;
                        quad.w = w == 3 ? acc : quad.w;
                        const uint32_t q = steps >> 4;
                        __stcs( q < SLOT / 16U ? slotQuads + q : slot2Quads + ( q - SLOT / 16U ), quad );
                    }
                    active = false;
                }
            }
        }
    }
}

/* ------------------------------------------------------------------ rank ------------------------------------ */
constexpr int RANK_THREADS = 1024;
constexpr uint32_t PER_THREAD = ( MAX_COARSE + RANK_THREADS - 1 ) / RANK_THREADS;

struct RankShared
{
    uint32_t dist[MAX_COARSE + 1];
    uint16_t jump[MAX_COARSE + 1];
    int starCoarse;       /* coarse segment whose chain of segments passes through origPtr, -1: none */
    uint32_t starStep;    /* ... after this many steps */
};

__global__ void __launch_bounds__( RANK_THREADS, 2 )
bwtRankKernel( const uint32_t* __restrict__ dbufAll, uint8_t* __restrict__ prerleAll, BlockMeta* __restrict__ meta,
               uint8_t* __restrict__ stageAll )
{
    extern __shared__ __align__( 16 ) uint8_t sharedRaw[];
    RankShared& sh = *reinterpret_cast<RankShared*>( sharedRaw );

    BlockMeta& bm = meta[blockIdx.x];
    if ( ( bm.status != IBZ2_ST_OK ) || ( bm.n == 0 ) ) {
        return;
    }
    const uint32_t n = bm.n;
    const BlockStage st = blockStage( stageAll, blockIdx.x );
    const uint32_t tid = threadIdx.x;
    const uint32_t segments = ( n + ( 1U << COARSE_SHIFT ) - 1U ) >> COARSE_SHIFT;   /* coarse */
    const int starFine = bm.travStar;

    /* ---- coarse segments: from every RATIO-th splitter along the chain of (fine) segments to the next such splitter.
     * The chain returns to where it started at the latest, so the loop ends; four hops on average, each one a pair of
     * loads that hit L2 (the walk has just written them). ---- */
    if ( tid == 0 ) {
        sh.starCoarse = -1;
        sh.starStep = 0;
    }
    __syncthreads();
    if ( starFine >= 0 ) {
        for ( uint32_t c = tid; c < segments; c += RANK_THREADS ) {
            uint32_t f = c << RATIO_SHIFT;
            uint32_t d = 0;
            do {
                if ( static_cast<int>( f ) == starFine ) {
                    sh.starCoarse = static_cast<int>( c );
                    sh.starStep = d + bm.travK;
                }
                const uint2 link = st.segLink[f];
                d += link.x;
                f = link.y;
            } while ( ( f & RATIO_MASK ) != 0 );
            sh.dist[c] = d;
            sh.jump[c] = static_cast<uint16_t>( f >> RATIO_SHIFT );
        }
    }
    __syncthreads();
    const int sstar = sh.starCoarse;

    if ( sstar < 0 ) {
        /* no coarse splitter on the start cycle (it is a short one then): serial walk of one lap by one thread */
        if ( tid == 0 ) {
            const uint32_t* __restrict__ dbuf = dbufAll + static_cast<uint64_t>( blockIdx.x ) * DBUF_STRIDE;
            uint8_t* __restrict__ prerle = prerleAll + static_cast<uint64_t>( blockIdx.x ) * LBYTES_STRIDE;
            const uint32_t p0 = bm.origPtr;
            uint32_t pos = p0;
            uint32_t k = 0;
            do {
                const uint32_t e = __ldcg( dbuf + pos );
                prerle[k++] = static_cast<uint8_t>( e );
                pos = e >> 8;
            } while ( ( pos != p0 ) && ( k < n ) );
            bm.cycleLen = k;
            bm.travStar = -1;   /* tells the place kernel that there is nothing to place */
        }
        return;
    }

    const uint32_t kstar = sh.starStep;
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
            const uint32_t s = tid + i * RANK_THREADS;
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
            const uint32_t s = tid + i * RANK_THREADS;
            if ( s < segments ) {
                sh.dist[s] = newDist[i];
                sh.jump[s] = newJump[i];
            }
        }
        __syncthreads();
    }

    const uint32_t tailLen = lenStar - kstar;          /* p0 .. end of sstar's segment, output offset 0 */
    const uint32_t ranked = sh.dist[first];            /* 0 when first == sstar */
    __syncthreads();
    for ( uint32_t c = tid; c < segments; c += RANK_THREADS ) {
        /* a coarse segment that reaches sstar lies on p0's cycle; its first byte goes to ... */
        st.segBase[c] = sh.jump[c] == NIL ? tailLen + ranked - sh.dist[c] : NOT_ON_CYCLE;
    }
    if ( tid == 0 ) {
        st.segBase[STAR_COARSE_SLOT] = static_cast<uint32_t>( sstar );
        st.segBase[STAR_STEP_SLOT] = kstar;
    }
    if ( tid == 0 ) {
        bm.cycleLen = ranked + lenStar;
    }
}

/* ------------------------------------------------------------------ place ----------------------------------- */
constexpr int PLACE_THREADS = 64;
constexpr uint32_t PLACE_CTAS_PER_BLOCK = ( MAX_COARSE + PLACE_THREADS - 1 ) / PLACE_THREADS;

/* len staged bytes of a segment (slot is 64-byte aligned) -> dst[0 .. len): bytes up to the first 4-byte boundary of
 * dst, then whole words put together from two neighbouring slot words, 16 slot bytes per round of loads. */
__device__ __forceinline__ void
placeStaged( const uint8_t* __restrict__ slot, uint8_t* __restrict__ dst, uint32_t len )
{
    const uint32_t* const slotWords = reinterpret_cast<const uint32_t*>( slot );
    const uint4* const slotQuads = reinterpret_cast<const uint4*>( slot );
    const uint32_t head = min( len, ( 4U - static_cast<uint32_t>( reinterpret_cast<uintptr_t>( dst ) & 3U ) ) & 3U );
    const uint32_t w0 = __ldg( slotWords );
    for ( uint32_t k = 0; k < head; ++k ) {
        dst[k] = static_cast<uint8_t>( w0 >> ( 8U * k ) );
    }
    const uint32_t shift = 8U * head;
    const uint32_t outWords = ( len - head ) >> 2;
    uint32_t* const out = reinterpret_cast<uint32_t*>( dst + head );
    /* 16 slot bytes per round: most segments are done after one or two */
    for ( uint32_t done = 0; done < outWords; done += 4U ) {
        const uint4 v = __ldg( slotQuads + ( done >> 2 ) );
        /* what lies behind the bytes wanted is never used (and still inside the block's staging area) */
        const uint32_t next = __ldg( slotWords + done + 4U );
        const uint32_t w[5] = { v.x, v.y, v.z, v.w, next };
#pragma unroll
        for ( uint32_t i = 0; i < 4U; ++i ) {
            if ( done + i < outWords ) {
                out[done + i] = __funnelshift_r( w[i], w[i + 1U], shift );
            }
        }
    }
    for ( uint32_t k = head + 4U * outWords; k < len; ++k ) {
        dst[k] = slot[k];
    }
}

/* One thread per coarse segment on p0's cycle: its bytes are one contiguous stretch of the output, put together from the
 * slots of the (four, on average) segments of its chain.  In the chain that passes through p0 the output begins at p0:
 * what comes after it goes to offset 0, what comes before it closes the lap. */
__global__ void __launch_bounds__( PLACE_THREADS )
bwtPlaceKernel( const uint32_t* __restrict__ dbufAll, uint8_t* __restrict__ prerleAll,
                const BlockMeta* __restrict__ meta, uint8_t* __restrict__ stageAll )
{
    const uint32_t block = blockIdx.y;
    const BlockMeta& bm = meta[block];
    if ( ( bm.status != IBZ2_ST_OK ) || ( bm.n == 0 ) || ( bm.travStar < 0 ) ) {
        return;
    }
    const uint32_t n = bm.n;
    const uint32_t coarseSegments = ( n + ( 1U << COARSE_SHIFT ) - 1U ) >> COARSE_SHIFT;
    const uint32_t c = blockIdx.x * PLACE_THREADS + threadIdx.x;
    if ( c >= coarseSegments ) {
        return;
    }
    const BlockStage st = blockStage( stageAll, block );
    const uint32_t coarseBase = st.segBase[c];
    if ( coarseBase == NOT_ON_CYCLE ) {
        return;
    }
    const uint32_t* __restrict__ dbuf = dbufAll + static_cast<uint64_t>( block ) * DBUF_STRIDE;
    uint8_t* __restrict__ prerle = prerleAll + static_cast<uint64_t>( block ) * LBYTES_STRIDE;
    const uint32_t starFine = static_cast<uint32_t>( bm.travStar );
    const uint32_t kstarFine = bm.travK;
    const bool starChain = c == st.segBase[STAR_COARSE_SLOT];
    const uint32_t kstarCoarse = st.segBase[STAR_STEP_SLOT];

    uint32_t s = c << RATIO_SHIFT;
    uint32_t d = 0;
    do {
        const uint2 link = st.segLink[s];
        const uint32_t len = link.x;
        const bool star = s == starFine;
        const bool behindP0 = starChain && ( d >= kstarCoarse ) && !star;
        const uint32_t base = behindP0 ? d - kstarCoarse : coarseBase + d;
        const uint8_t* const slot = st.slots + ( static_cast<uint64_t>( s ) << TRAV_SLOT_SHIFT );
        const uint8_t* const slot2 = st.slots2 + ( static_cast<uint64_t>( s ) << TRAV_SLOT2_SHIFT );
        const uint32_t staged = min( len, STAGED_MAX );
        if ( star ) {
            /* steps kstarFine .. len - 1 open the output, steps 0 .. kstarFine - 1 close the lap */
            for ( uint32_t k = 0; k < staged; ++k ) {
                const uint32_t idx = k >= kstarFine ? k - kstarFine : base + k;
                if ( idx < n ) {
                    prerle[idx] = k < SLOT ? slot[k] : slot2[k - SLOT];
                }
            }
        } else if ( base < n ) {
            const uint32_t fits = min( staged, n - base );
            placeStaged( slot, prerle + base, min( fits, SLOT ) );
            if ( fits > SLOT ) {
                placeStaged( slot2, prerle + base + SLOT, fits - SLOT );
            }
        }
        if ( len > STAGED_MAX ) {
            /* the rest of a very long segment: walked again from where the second slot ended */
            uint32_t pos = st.ovfPos[s];
            for ( uint32_t k = STAGED_MAX; k < len; ++k ) {
                const uint32_t e = __ldcg( dbuf + pos );
                const uint32_t idx = ( star && ( k >= kstarFine ) ) ? ( k - kstarFine ) : ( base + k );
                if ( idx < n ) {
                    prerle[idx] = static_cast<uint8_t>( e );
                }
                pos = e >> 8;
            }
        }
        d += len;
        s = link.y;
    } while ( ( s & RATIO_MASK ) != 0 );
}

/* ---- periodic continuation when the start cycle is shorter than the block ---- */
__global__ void __launch_bounds__( PLACE_THREADS )
bwtFillKernel( uint8_t* __restrict__ prerleAll, const BlockMeta* __restrict__ meta )
{
    const BlockMeta& bm = meta[blockIdx.x];
    if ( ( bm.status != IBZ2_ST_OK ) || ( bm.n == 0 ) || ( bm.cycleLen >= bm.n ) || ( bm.cycleLen == 0 ) ) {
        return;
    }
    uint8_t* __restrict__ prerle = prerleAll + static_cast<uint64_t>( blockIdx.x ) * LBYTES_STRIDE;
    const uint32_t cycleLen = bm.cycleLen;
    for ( uint32_t j = cycleLen + threadIdx.x; j < bm.n; j += PLACE_THREADS ) {
        prerle[j] = prerle[j % cycleLen];
    }
}
}  // namespace

void
launchBwtTraverse( const uint32_t* dbuf, uint8_t* prerle, BlockMeta* meta, uint32_t nblocks, uint8_t* stage,
                   cudaStream_t stream, cudaEvent_t walkStop )
{
    static const int walkThreads = [] () {
        const char* const env = std::getenv( "IBZ2_WALK_THREADS" );
        const int v = env != nullptr ? std::atoi( env ) : WALK_THREADS_DEFAULT;
        return ( ( v >= 32 ) && ( v <= 1024 ) ) ? ( v / 32 ) * 32 : WALK_THREADS_DEFAULT;
    }();
    cudaMemsetAsync( stage, 0, sizeof( uint32_t ), stream );
    /* small waves: no more warps than there are chunks of work */
    const uint64_t chunks = static_cast<uint64_t>( nblocks ) * CHUNKS_PER_BLOCK;
    const uint32_t ctas = static_cast<uint32_t>( std::min<uint64_t>( static_cast<uint64_t>( smCount() ),
                                                                   ( chunks * 32U + walkThreads - 1U ) / walkThreads ) );
    bwtWalkKernel<<<std::max( 1U, ctas ), walkThreads, 0, stream>>>( dbuf, meta, stage, nblocks );
    if ( walkStop != nullptr ) {
        cudaEventRecord( walkStop, stream );
    }
    cudaFuncSetAttribute( bwtRankKernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                          static_cast<int>( sizeof( RankShared ) ) );
    bwtRankKernel<<<nblocks, RANK_THREADS, sizeof( RankShared ), stream>>>( dbuf, prerle, meta, stage );
    bwtPlaceKernel<<<dim3( PLACE_CTAS_PER_BLOCK, nblocks ), PLACE_THREADS, 0, stream>>>( dbuf, prerle, meta, stage );
    bwtFillKernel<<<nblocks, PLACE_THREADS, 0, stream>>>( prerle, meta );
}
}  // namespace ibz2
