/*
 * K2c..K2e -- move-to-front and RUNA/RUNB expansion of Block::readBlockData (indexed_bzip2/bzip2.hpp:726-790), i.e. the
 * half of the symbol loop that K2b (huffman_bits.cu) leaves out, with chunk-level parallelism inside every block.
 *
 * Move-to-front looks serial (the list depends on every earlier symbol), but a chunk of symbols can be decoded against
 * an UNKNOWN start list symbolically: start from the identity list, so that every output is "the entry that was at
 * position p of the chunk's start list", and the list left behind is a permutation of start positions.  Permutations
 * compose, so a cheap serial pass over the chunk permutations recovers every chunk's real start list.
 *   K2c  mtfChunkKernel   one LANE per chunk of MTF_CHUNK symbols: symbolic move-to-front on a private list held in
 *                         the lane's own shared-memory column (16-byte quad q of lane t at [q][t]; the first 16 entries
 *                         stay in registers); writes one symbolic byte per symbol and the chunk's final permutation.
 *   K2d  mtfListsKernel   one WARP per block: start list of chunk j+1 = start list of chunk j o permutation j, seeded
 *                         with symbolToByte; overwrites every permutation with the chunk's start list in place.
 *   K2e  mtfExpandKernel  one WARP per chunk: symbolic byte -> real byte through the start list, RUNA/RUNB symbol at
 *                         position k of its number -> ( symbol + 1 ) << k copies (the 32-bit sum of bzip2.hpp:742 is the
 *                         sum of these shares), exclusive scan from the chunk checkpoint, bytes staged in shared memory
 *                         and written to L in 128-byte rows.
 * Algorithmic HBM traffic per block: 2 B/symbol read twice, 1 B/symbol written and read, 256 B per chunk twice,
 * N bytes of L written.
 */
#include "common.cuh"

namespace ibz2
{
namespace
{
constexpr uint32_t FULL = 0xFFFFFFFFU;
constexpr int MTFC_THREADS = 128;
constexpr uint32_t MAX_RUN_DIGITS = 20;   /* 2^20 - 1 > 900000: a longer RUNA/RUNB number overflows the block or wraps */
constexpr uint32_t LONG_RUN_FLAG = 0x80000000U;

__device__ __forceinline__ const uint8_t*
k2Scratch( const uint32_t* scratchAll, uint32_t block )
{
    return reinterpret_cast<const uint8_t*>( scratchAll + static_cast<uint64_t>( block ) * DBUF_STRIDE );
}

__device__ __forceinline__ uint8_t*
k2Scratch( uint32_t* scratchAll, uint32_t block )
{
    return reinterpret_cast<uint8_t*>( scratchAll + static_cast<uint64_t>( block ) * DBUF_STRIDE );
}

/* ---------------------------------------------------------------- K2c ---------------------------------------- */
/* A move-to-front step costs what it moves: with the list as one array that is the index itself, 128 bytes on average
 * for incompressible data, and a warp waits for the lane with the largest index (the first version of this kernel
 * shifted 16-byte quads through shared memory: 47 G warp instructions per 8 GB, 15 of 32 lanes active, LSU-bound).
 * Here a step costs the same whatever the index is:
 *   F   the entries moved to the front since the last merge, newest first, in registers (at most 15 of them);
 *   M   the other entries in shared memory, 16 rows of 16, each row's live entries packed at its start.  Taking an entry
 *       out of M closes the gap inside its row only and leaves the row one entry short (a "hole").
 * The list is F followed by the live entries of the rows in order.  An index below |F| is served inside the registers;
 * any other index j - |F| falls into row r0 = j' / 16 or the row after it, because there are at most 15 holes in
 * total -- `before[k]`, the number of holes in front of row k, is kept as 16 nibbles, so that one funnel shift yields
 * before[r0] and before[r0 + 1].  When some lane of the warp has collected 14 entries in F, every lane merges: F and
 * the rows are streamed into full rows again (byte permutes with a carry word, word stores at a running address).
 * Rows are stored as words, [word][thread]: a lane only ever touches its own bank, whatever row it is working on. */
constexpr uint32_t MTF_MERGE_AT = 14;   /* F may grow by two entries between two looks at it: never more than 15 */

__device__ __forceinline__ uint32_t
quadWord( const uint4& q, uint32_t w )
{
    return w == 0 ? q.x : ( w == 1 ? q.y : ( w == 2 ? q.z : q.w ) );
}

/* entries move up by one, `in` becomes entry 0 */
__device__ __forceinline__ void
shiftQuad( uint4& q, uint32_t in )
{
    q.w = __funnelshift_l( q.z, q.w, 8 );
    q.z = __funnelshift_l( q.y, q.z, 8 );
    q.y = __funnelshift_l( q.x, q.y, 8 );
    q.x = ( q.x << 8 ) | in;
}

/* entries 0 .. e - 1 of the quad move up by one (entry e is overwritten), `in` becomes entry 0; e = 4 w + sh / 8 */
__device__ __forceinline__ void
shiftQuadBelow( uint4& q, uint32_t w, uint32_t sh, uint32_t in )
{
    uint4 y = q;
    shiftQuad( y, in );
    const uint32_t keep = ( sh == 24U ) ? 0U : ( 0xFFFFFFFFU << ( sh + 8U ) );   /* entries above e inside its word */
    const uint32_t k0 = w == 0 ? keep : 0U;
    const uint32_t k1 = w == 1 ? keep : ( w > 1 ? 0U : 0xFFFFFFFFU );
    const uint32_t k2 = w == 2 ? keep : ( w > 2 ? 0U : 0xFFFFFFFFU );
    const uint32_t k3 = w == 3 ? keep : 0xFFFFFFFFU;
    q.x = ( q.x & k0 ) | ( y.x & ~k0 );
    q.y = ( q.y & k1 ) | ( y.y & ~k1 );
    q.z = ( q.z & k2 ) | ( y.z & ~k2 );
    q.w = ( q.w & k3 ) | ( y.w & ~k3 );
}

__global__ void __launch_bounds__( MTFC_THREADS )
mtfChunkKernel( const BlockMeta* __restrict__ meta, uint32_t* __restrict__ scratchAll )
{
    /* word 4 r + k of a lane's list = entries 16 r + 4 k .. + 3 of M; eight more words because a step reads the four words
     * from the one that holds its entry on, and a merge writes whole groups of four at its output position */
    __shared__ uint32_t rows[72][MTFC_THREADS];

    const uint32_t block = blockIdx.y;
    const BlockMeta& bm = meta[block];
    if ( ( bm.status != IBZ2_ST_OK ) || ( bm.isEos != 0 ) ) {
        return;
    }
    const uint32_t nCoded = bm.nCoded;
    const uint32_t chunk = blockIdx.x * MTFC_THREADS + threadIdx.x;
    const uint32_t begin = chunk << MTF_CHUNK_SHIFT;
    if ( ( blockIdx.x * MTFC_THREADS ) << MTF_CHUNK_SHIFT >= nCoded ) {
        return;   /* the whole CTA */
    }
    /* every lane of a warp stays until the warp's longest chunk is done: the merges are decided by votes */
    const uint32_t count = begin < nCoded ? min( MTF_CHUNK, nCoded - begin ) : 0U;
    const uint32_t warpCount = __reduce_max_sync( FULL, count );
    uint8_t* const scratch = k2Scratch( scratchAll, block );
    const uint16_t* const sym16 = reinterpret_cast<const uint16_t*>( scratch + K2_SYM16_OFFSET ) + begin;
    uint32_t* const sym8 = reinterpret_cast<uint32_t*>( scratch + K2_SYM8_OFFSET + begin );
    uint4* const listOut = reinterpret_cast<uint4*>( scratch + K2_LISTS_OFFSET + static_cast<uint64_t>( chunk ) * 256U );
    ChunkCheckpoint* const ckpt = reinterpret_cast<ChunkCheckpoint*>( scratch + K2_CKPT_OFFSET ) + chunk;

    /* digits of a RUNA/RUNB number directly in front of the chunk (bzip2.hpp:726-743: runPos = 1 << digits) */
    uint32_t runDigits = 0;
    if ( count > 0 ) {
        while ( ( runDigits <= MAX_RUN_DIGITS ) && ( runDigits < begin ) && ( sym16[-1 - static_cast<int>( runDigits )] <= 1U ) ) {
            ++runDigits;
        }
    }
    const uint32_t runShift = runDigits;
    uint32_t mostDigits = runDigits;
    uint32_t produced = 0;   /* bytes of L the chunk's symbols stand for */

    uint32_t* const my = &rows[0][threadIdx.x];
    constexpr uint32_t STRIDE = MTFC_THREADS;
#pragma unroll
    for ( uint32_t w = 0; w < 64U; ++w ) {
        my[w * STRIDE] = 0x03020100U + 0x04040404U * w;   /* the identity list */
    }
    uint4 head = make_uint4( 0, 0, 0, 0 );   /* F */
    uint32_t f = 0;
    uint32_t beforeLo = 0;                   /* before[k] = holes in front of row k, one nibble each */
    uint32_t beforeHi = 0;
    uint32_t front = 0;                      /* entry 0 of the list */

    /* F and the rows -> full rows.  Reads run one row ahead of the writes, which can reach that far. */
    auto merge = [&] () {
        uint32_t carry = 0;
        uint32_t carryBytes = 0;
        uint32_t* out = my;
        const uint32_t* in = my;
        uint32_t s0 = head.x;
        uint32_t s1 = head.y;
        uint32_t s2 = head.z;
        uint32_t s3 = head.w;
        uint32_t live = f;
        uint32_t holesLo = beforeLo;
        uint32_t holesHi = beforeHi;
        uint32_t holesBefore = 0;
#pragma unroll 1
        for ( uint32_t r = 0; r <= 16U; ++r ) {
            uint32_t n0 = 0;
            uint32_t n1 = 0;
            uint32_t n2 = 0;
            uint32_t n3 = 0;
            uint32_t nextLive = 0;
            if ( r < 16U ) {
                n0 = in[0];
                n1 = in[STRIDE];
                n2 = in[2U * STRIDE];
                n3 = in[3U * STRIDE];
                in += 4U * STRIDE;
                holesLo = __funnelshift_r( holesLo, holesHi, 4 );
                holesHi >>= 4;
                const uint32_t holesBehind = r == 15U ? f : ( holesLo & 15U );   /* before[r + 1]; all of them behind the last row */
                nextLive = 16U - ( holesBehind - holesBefore );
                holesBefore = holesBehind;
            }
            /* the source's bytes go behind the carry's: word i of the result = bytes 4 i - carryBytes .. of the source */
            const uint32_t selB = 0x7654U - 0x1111U * carryBytes;
            const uint32_t lowNibbles = ( 1U << ( 4U * carryBytes ) ) - 1U;
            const uint32_t selA = ( selB & ~lowNibbles ) | ( 0x3210U & lowNibbles );
            const uint32_t y0 = __byte_perm( carry, s0, selA );
            const uint32_t y1 = __byte_perm( s0, s1, selB );
            const uint32_t y2 = __byte_perm( s1, s2, selB );
            const uint32_t y3 = __byte_perm( s2, s3, selB );
            const uint32_t y4 = __byte_perm( s3, 0U, selB );
            const uint32_t total = carryBytes + live;   /* <= 3 + 16 */
            const uint32_t full = total >> 2;
            /* all four words go out; the ones behind the `full` complete ones are overwritten by the next source, and the
             * first of them is this source's carry (the reads are a row ahead, the writes stay inside that row) */
            out[0] = y0;
            out[STRIDE] = y1;
            out[2U * STRIDE] = y2;
            out[3U * STRIDE] = y3;
            out += full * STRIDE;
            carry = full == 4U ? y4 : out[0];
            carryBytes = total & 3U;
            s0 = n0;
            s1 = n1;
            s2 = n2;
            s3 = n3;
            live = nextLive;
        }
        f = 0;
        beforeLo = 0;
        beforeHi = 0;
    };

    for ( uint32_t base = 0; base < warpCount; base += 8U ) {
        /* 8 symbols per 16-byte load; `begin` is a multiple of 1024 symbols, so the loads are aligned */
        const uint4 v = base < count ? *reinterpret_cast<const uint4*>( sym16 + base ) : make_uint4( 0, 0, 0, 0 );
        const uint32_t packed[4] = { v.x, v.y, v.z, v.w };
        uint32_t out[2] = { 0, 0 };
#pragma unroll
        for ( uint32_t k = 0; k < 8U; ++k ) {
            const uint32_t sym = ( packed[k >> 1] >> ( 16U * ( k & 1U ) ) ) & 0xFFFFU;
            const bool valid = base + k < count;
            if ( valid ) {
                /* a number of more than 20 digits sends the block to the exact decoder whatever `produced` says, so the
                 * shift may simply run out (clamped funnel shift: 0 from 32 digits on) */
                const bool isRun = sym <= 1U;
                produced += isRun ? __funnelshift_lc( 0U, sym + 1U, runDigits ) : 1U;
                runDigits = isRun ? runDigits + 1U : 0U;
                mostDigits = max( mostDigits, runDigits );
            }
            if ( ( sym >= 2U ) && valid ) {
                const uint32_t idx = sym - 1U;   /* 1..255 */
                if ( idx < f ) {
                    /* inside F */
                    const uint32_t w = ( idx >> 2 ) & 3U;
                    const uint32_t sh = 8U * ( idx & 3U );
                    if ( idx < 4U ) {
                        front = ( head.x >> sh ) & 0xFFU;
                        const uint32_t keep = ( idx == 3U ) ? 0U : ( 0xFFFFFFFFU << ( sh + 8U ) );
                        head.x = ( head.x & keep ) | ( ( ( head.x << 8 ) | front ) & ~keep );
                    } else {
                        front = ( quadWord( head, w ) >> sh ) & 0xFFU;
                        shiftQuadBelow( head, w, sh, front );
                    }
                } else {
                    /* entry j of M: in row j / 16 or, when the holes in front of it add up, the row behind */
                    const uint32_t j = idx - f;
                    const uint32_t r0 = j >> 4;
                    const bool low = r0 < 8U;
                    const uint32_t window = __funnelshift_r( low ? beforeLo : beforeHi, low ? beforeHi : 0U, 4U * r0 );
                    const uint32_t liveBeforeNext = 16U * ( r0 + 1U ) - ( ( window >> 4 ) & 15U );
                    const bool next = j >= liveBeforeNext;
                    const uint32_t row = r0 + ( next ? 1U : 0U );
                    const uint32_t p = j - ( next ? liveBeforeNext : 16U * r0 - ( window & 15U ) );   /* 0..15 */
                    /* the four words from the one that holds entry p on (they may reach into the next row: read, never
                     * written back).  Entry p leaves; what is behind it in the row moves down by one. */
                    const uint32_t w = p >> 2;
                    const uint32_t c = p & 3U;
                    uint32_t* const wordPtr = my + ( row * 4U + w ) * STRIDE;
                    const uint32_t x0 = wordPtr[0];
                    const uint32_t x1 = wordPtr[STRIDE];
                    const uint32_t x2 = wordPtr[2U * STRIDE];
                    const uint32_t x3 = wordPtr[3U * STRIDE];
                    front = __byte_perm( x0, 0U, 0x4440U | c );
                    const uint32_t selMixed = 0x4321U - ( 0x1111U & ( ( 1U << ( 4U * c ) ) - 1U ) );   /* drops byte c */
                    wordPtr[0] = __byte_perm( x0, x1, selMixed );
                    if ( w < 3U ) {
                        wordPtr[STRIDE] = __byte_perm( x1, x2, 0x4321U );
                    }
                    if ( w < 2U ) {
                        wordPtr[2U * STRIDE] = __byte_perm( x2, x3, 0x4321U );
                    }
                    if ( w < 1U ) {
                        wordPtr[3U * STRIDE] = x3 >> 8;
                    }
                    /* one more hole in front of every row behind this one */
                    const uint32_t s = 4U * ( row + 1U );   /* 4..64 */
                    beforeLo += s < 32U ? ( 0x11111111U << ( s & 31U ) ) : 0U;
                    beforeHi += s < 32U ? 0x11111111U : ( s < 64U ? ( 0x11111111U << ( s & 31U ) ) : 0U );
                    shiftQuad( head, front );
                    ++f;
                }
            }
            out[k >> 2] |= front << ( 8U * ( k & 3U ) );
            if ( ( k & 1U ) == 1U ) {
                if ( __any_sync( FULL, f >= MTF_MERGE_AT ) ) {
                    merge();
                }
            }
        }
        if ( base < count ) {
            sym8[( base >> 2 ) + 0U] = out[0];
            sym8[( base >> 2 ) + 1U] = out[1];
        }
    }
    merge();

    if ( count > 0 ) {
        *ckpt = ChunkCheckpoint{ produced, runShift | ( mostDigits > MAX_RUN_DIGITS ? LONG_RUN_FLAG : 0U ) };
#pragma unroll 4
        for ( uint32_t q = 0; q < 16U; ++q ) {
            listOut[q] = make_uint4( my[( 4U * q ) * STRIDE], my[( 4U * q + 1U ) * STRIDE], my[( 4U * q + 2U ) * STRIDE],
                                     my[( 4U * q + 3U ) * STRIDE] );
        }
    }
}

/* ---------------------------------------------------------------- K2d ---------------------------------------- */
constexpr int MTFL_WARPS = 4;

__global__ void __launch_bounds__( MTFL_WARPS * 32 )
mtfListsKernel( BlockMeta* __restrict__ meta, uint32_t nblocks, const BlockTables* __restrict__ tablesAll,
                uint32_t* __restrict__ scratchAll )
{
    __shared__ __align__( 16 ) uint8_t startList[MTFL_WARPS][256];

    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31U;
    const uint32_t block = blockIdx.x * MTFL_WARPS + warp;
    if ( block >= nblocks ) {
        return;
    }
    BlockMeta& bm = meta[block];
    if ( ( bm.status != IBZ2_ST_OK ) || ( bm.isEos != 0 ) ) {
        return;
    }
    const uint32_t chunks = ( bm.nCoded + MTF_CHUNK - 1U ) >> MTF_CHUNK_SHIFT;
    ChunkCheckpoint* const ckpt = reinterpret_cast<ChunkCheckpoint*>( k2Scratch( scratchAll, block ) + K2_CKPT_OFFSET );
    uint64_t produced = 0;
    uint32_t flags = 0;
    uint2* const lists = reinterpret_cast<uint2*>( k2Scratch( scratchAll, block ) + K2_LISTS_OFFSET );
    uint8_t* const cur = startList[warp];
    uint2* const curWords = reinterpret_cast<uint2*>( cur );

    /* list position p holds byte symbolToByte[p] at block start (bzip2.hpp:699 + :786) */
    curWords[lane] = reinterpret_cast<const uint2*>( tablesAll[block].s2b )[lane];
    __syncwarp();

    /* what the loop needs of chunk j + 1 is loaded while chunk j is composed: the chain from one chunk to the next is the
     * shared-memory gather alone (the checkpoint's load used to sit on it: 1030 cycles per chunk) */
    uint2 perm = chunks > 0 ? lists[lane] : make_uint2( 0, 0 );
    ChunkCheckpoint ckNext = chunks > 0 ? ckpt[0] : ChunkCheckpoint{ 0, 0 };
    for ( uint32_t j = 0; j < chunks; ++j ) {
        const uint2 nextPerm = ( j + 1U < chunks ) ? lists[static_cast<uint64_t>( j + 1U ) * 32U + lane] : make_uint2( 0, 0 );
        const ChunkCheckpoint ck = ckNext;
        if ( j + 1U < chunks ) {
            ckNext = ckpt[j + 1U];
        }
        const uint2 mine = curWords[lane];
        uint2 composed;
        composed.x = static_cast<uint32_t>( cur[perm.x & 0xFFU] ) | ( static_cast<uint32_t>( cur[( perm.x >> 8 ) & 0xFFU] ) << 8 )
                     | ( static_cast<uint32_t>( cur[( perm.x >> 16 ) & 0xFFU] ) << 16 )
                     | ( static_cast<uint32_t>( cur[perm.x >> 24] ) << 24 );
        composed.y = static_cast<uint32_t>( cur[perm.y & 0xFFU] ) | ( static_cast<uint32_t>( cur[( perm.y >> 8 ) & 0xFFU] ) << 8 )
                     | ( static_cast<uint32_t>( cur[( perm.y >> 16 ) & 0xFFU] ) << 16 )
                     | ( static_cast<uint32_t>( cur[perm.y >> 24] ) << 24 );
        lists[static_cast<uint64_t>( j ) * 32U + lane] = mine;   /* the chunk's start list replaces its permutation */
        {
            /* chunk byte counts -> where each chunk's bytes start in L (every lane keeps the same running sum) */
            flags |= ck.runShift;
            if ( lane == 0 ) {
                ckpt[j] = ChunkCheckpoint{ static_cast<uint32_t>( min( produced, static_cast<uint64_t>( 0xFFFFFFFFU ) ) ),
                                           ck.runShift & ~LONG_RUN_FLAG };
            }
            produced += ck.outOffset;
        }
        __syncwarp();
        curWords[lane] = composed;
        __syncwarp();
        perm = nextPerm;
    }
    if ( lane == 0 ) {
        if ( ( ( flags & LONG_RUN_FLAG ) != 0 ) || ( produced > DBUF_SIZE ) ) {
            bm.status = ST_INTERNAL_CAREFUL;   /* bzip2.hpp:751-756 / :776-780 or a wrapped run number: the exact decoder decides */
        } else if ( bm.origPtr >= produced ) {
            bm.status = IBZ2_ST_ORIGPTR;       /* bzip2.hpp:794-798 */
            bm.endBit = bm.bitOffset;
        } else {
            bm.n = static_cast<uint32_t>( produced );
        }
    }
}

/* ---------------------------------------------------------------- K2e ---------------------------------------- */
constexpr int MTFE_WARPS = 4;
constexpr uint32_t STAGE_BYTES = 4096;
constexpr uint32_t PER_LANE = MTF_CHUNK / 32U;   /* 32 symbols per lane */
/* A lane owns 32 CONSECUTIVE symbols, so with a dense layout the 32 lanes of one load sit 64 (or 32) bytes apart: two
 * (four) banks for the whole warp.  Each lane's row gets one word of padding, which puts the rows of a warp in 32
 * different banks, and the lane reads its row into registers once instead of once per pass. */
constexpr uint32_t SYM16_ROW = PER_LANE / 2U + 1U;   /* words */
constexpr uint32_t SYM8_ROW = PER_LANE / 4U + 1U;
/* the assembled bytes: lanes write about 32 bytes apart, so one pad word goes in after every 8 */
constexpr uint32_t STAGE_PHYS_BYTES = STAGE_BYTES + ( STAGE_BYTES >> 5 ) * 4U + 32U;

struct ExpandShared
{
    uint32_t sym8[32U * SYM8_ROW];
    uint8_t list[256];
    /* the symbols are condensed into two bit masks per lane before the first byte is staged, and the long-run path,
     * which reads them again, does not stage */
    union
    {
        uint32_t sym16[32U * SYM16_ROW];
        uint8_t stage[STAGE_PHYS_BYTES];
    };
};

__device__ __forceinline__ uint32_t
stagePhys( uint32_t p )
{
    return p + ( ( p >> 5 ) << 2 );
}

__device__ __forceinline__ uint32_t
runShare( uint32_t sym, uint32_t k )
{
    return k < 32U ? ( ( sym + 1U ) << k ) : 0U;
}

__global__ void __launch_bounds__( MTFE_WARPS * 32 )
mtfExpandKernel( const BlockMeta* __restrict__ meta, const uint32_t* __restrict__ scratchAll,
                 uint8_t* __restrict__ lbytesAll )
{
    __shared__ __align__( 16 ) ExpandShared shared[MTFE_WARPS];

    const uint32_t block = blockIdx.y;
    const BlockMeta& bm = meta[block];
    if ( ( bm.status != IBZ2_ST_OK ) || ( bm.isEos != 0 ) ) {
        return;
    }
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31U;
    const uint32_t nCoded = bm.nCoded;
    const uint32_t n = bm.n;
    const uint32_t chunk = blockIdx.x * MTFE_WARPS + warp;
    const uint32_t begin = chunk << MTF_CHUNK_SHIFT;
    if ( begin >= nCoded ) {
        return;
    }
    const uint32_t count = min( MTF_CHUNK, nCoded - begin );
    ExpandShared& sh = shared[warp];
    const uint8_t* const scratch = k2Scratch( scratchAll, block );
    uint8_t* const lbytes = lbytesAll + static_cast<uint64_t>( block ) * LBYTES_STRIDE;

    /* stage the chunk: 2 KB of symbols, 1 KB of symbolic bytes, the start list */
    {
        const uint4* s16 = reinterpret_cast<const uint4*>( scratch + K2_SYM16_OFFSET + static_cast<uint64_t>( begin ) * 2U );
        const uint4* s8 = reinterpret_cast<const uint4*>( scratch + K2_SYM8_OFFSET + begin );
        const uint2* ls = reinterpret_cast<const uint2*>( scratch + K2_LISTS_OFFSET + static_cast<uint64_t>( chunk ) * 256U );
#pragma unroll
        for ( uint32_t i = 0; i < 4U; ++i ) {
            const uint32_t q = i * 32U + lane;   /* 16 bytes = 8 symbols; four of them per row */
            const uint4 v = s16[q];
            uint32_t* const d = sh.sym16 + ( q >> 2 ) * SYM16_ROW + ( q & 3U ) * 4U;
            d[0] = v.x;
            d[1] = v.y;
            d[2] = v.z;
            d[3] = v.w;
        }
#pragma unroll
        for ( uint32_t i = 0; i < 2U; ++i ) {
            const uint32_t q = i * 32U + lane;   /* 16 symbolic bytes; two of them per row */
            const uint4 v = s8[q];
            uint32_t* const d = sh.sym8 + ( q >> 1 ) * SYM8_ROW + ( q & 1U ) * 4U;
            d[0] = v.x;
            d[1] = v.y;
            d[2] = v.z;
            d[3] = v.w;
        }
        reinterpret_cast<uint2*>( sh.list )[lane] = ls[lane];
    }
    const ChunkCheckpoint ck = reinterpret_cast<const ChunkCheckpoint*>( scratch + K2_CKPT_OFFSET )[chunk];
    __syncwarp();

    /* lane l owns symbols [32 l, 32 l + 32) of the chunk */
    const uint32_t first = lane * PER_LANE;
    const uint32_t mine = first >= count ? 0U : min( PER_LANE, count - first );
    const uint32_t valid = mine >= 32U ? FULL : ( ( 1U << mine ) - 1U );

    /* bit i of runMask: symbol i is RUNA/RUNB; bit i of runB: it is RUNB */
    uint32_t runMask = 0;
    uint32_t runB = 0;
    {
        const uint32_t* const row = sh.sym16 + lane * SYM16_ROW;
#pragma unroll 4
        for ( uint32_t j = 0; j < PER_LANE / 2U; ++j ) {
            const uint32_t pair = row[j];
            const uint32_t isRun = ( ( ( pair & 0xFFFFU ) <= 1U ) ? 1U : 0U ) | ( ( ( pair >> 16 ) <= 1U ) ? 2U : 0U );
            const uint32_t isB = ( pair & 1U ) | ( ( pair >> 15 ) & 2U );
            runMask |= isRun << ( 2U * j );
            runB |= isB << ( 2U * j );
        }
    }
    runMask &= valid;
    __syncwarp();   /* sym16 may be overwritten by stage from here on */

    /* position inside a RUNA/RUNB number at the lane's first symbol */
    const uint32_t notRun = ~runMask & valid;
    const bool allRun = notRun == 0;
    /* run symbols behind the lane's last non-run symbol (only read from lanes that have one) */
    const uint32_t trailing = allRun ? mine : mine + static_cast<uint32_t>( __clz( static_cast<int>( notRun ) ) ) - 32U;
    const uint32_t allRunMask = __ballot_sync( FULL, allRun );
    uint32_t k0;
    {
        /* nearest lane below that contains a non-run symbol */
        const uint32_t below = ~allRunMask & ( ( 1U << lane ) - 1U );
        const uint32_t src = below == 0 ? 0U : 31U - static_cast<uint32_t>( __clz( static_cast<int>( below ) ) );
        const uint32_t t = __shfl_sync( FULL, trailing, src );
        k0 = below == 0 ? ck.runShift + PER_LANE * lane : t + PER_LANE * ( lane - 1U - src );
    }

    /* bytes produced by the lane's symbols: one per ordinary symbol, ( digit + 1 ) << position per RUNA/RUNB */
    uint32_t total = static_cast<uint32_t>( __popc( notRun ) );
    {
        uint32_t pending = runMask;
        while ( pending != 0 ) {
            const uint32_t i = static_cast<uint32_t>( __ffs( static_cast<int>( pending ) ) ) - 1U;
            pending &= pending - 1U;
            /* run symbols directly below i inside the lane; the number continues from k0 when they reach bit 0 */
            const uint32_t lower = ~runMask & ( ( 1U << i ) - 1U );
            const uint32_t k = lower == 0 ? k0 + i : i + static_cast<uint32_t>( __clz( static_cast<int>( lower ) ) ) - 32U;
            total += runShare( ( runB >> i ) & 1U, k );
        }
    }
    uint32_t incl = total;
#pragma unroll
    for ( int d = 1; d < 32; d <<= 1 ) {
        const uint32_t up = __shfl_up_sync( FULL, incl, d );
        if ( lane >= static_cast<uint32_t>( d ) ) {
            incl += up;
        }
    }
    const uint32_t chunkTotal = __shfl_sync( FULL, incl, 31 );
    const uint32_t chunkBase = ck.outOffset;
    uint32_t pos = incl - total;   /* relative to chunkBase */

    if ( chunkTotal <= STAGE_BYTES ) {
        /* common case: assemble the chunk's bytes in shared memory, then write 128-byte rows */
        uint32_t k = k0;
        const uint32_t* const row8 = sh.sym8 + lane * SYM8_ROW;
#pragma unroll 1
        for ( uint32_t j = 0; j < PER_LANE / 4U; ++j ) {
            const uint32_t four = row8[j];
#pragma unroll
            for ( uint32_t t = 0; t < 4U; ++t ) {
                const uint32_t i = 4U * j + t;
                const uint8_t byte = sh.list[( four >> ( 8U * t ) ) & 0xFFU];
                const bool isRun = ( ( runMask >> i ) & 1U ) != 0;
                const uint32_t c = isRun ? runShare( ( runB >> i ) & 1U, k ) : ( ( valid >> i ) & 1U );
                k = isRun ? k + 1U : 0U;
                for ( uint32_t r = 0; r < c; ++r ) {
                    sh.stage[stagePhys( pos + r )] = byte;
                }
                pos += c;
            }
        }
        __syncwarp();
        const uint32_t limit = chunkBase < n ? min( chunkTotal, n - chunkBase ) : 0U;
        uint8_t* const dst = lbytes + chunkBase;
        const uint32_t head = min( limit, ( 4U - ( chunkBase & 3U ) ) & 3U );
        if ( lane < head ) {
            dst[lane] = sh.stage[stagePhys( lane )];
        }
        const uint32_t words = ( limit - head ) >> 2;
        const uint32_t* const stageWords = reinterpret_cast<const uint32_t*>( sh.stage );
        uint32_t* const dstWords = reinterpret_cast<uint32_t*>( dst + head );
        for ( uint32_t w = lane; w < words; w += 32U ) {
            /* stage bytes head + 4 w .. + 3: unaligned in shared memory whenever head != 0 */
            const uint32_t lo = stageWords[w + ( w >> 3 )];
            const uint32_t hi = stageWords[( w + 1U ) + ( ( w + 1U ) >> 3 )];
            dstWords[w] = __funnelshift_r( lo, hi, 8U * head );
        }
        const uint32_t tail = head + 4U * words;
        if ( tail + lane < limit ) {
            dst[tail + lane] = sh.stage[stagePhys( tail + lane )];
        }
    } else {
        /* long runs: short shares are stored by their lane, long ones by the whole warp */
        uint32_t k = k0;
        for ( uint32_t i = 0; i < PER_LANE; ++i ) {
            uint32_t c = 0;
            uint32_t byte = 0;
            if ( i < mine ) {
                const uint32_t sym = ( sh.sym16[lane * SYM16_ROW + ( i >> 1 )] >> ( 16U * ( i & 1U ) ) ) & 0xFFFFU;
                byte = sh.list[( sh.sym8[lane * SYM8_ROW + ( i >> 2 )] >> ( 8U * ( i & 3U ) ) ) & 0xFFU];
                c = 1;
                if ( sym <= 1U ) {
                    c = runShare( sym, k );
                    ++k;
                } else {
                    k = 0;
                }
            }
            const uint32_t at = chunkBase + pos;
            pos += c;
            const bool isLong = c >= 64U;
            if ( !isLong ) {
                for ( uint32_t r = 0; r < c; ++r ) {
                    if ( at + r < n ) {
                        lbytes[at + r] = static_cast<uint8_t>( byte );
                    }
                }
            }
            uint32_t pending = __ballot_sync( FULL, isLong );
            while ( pending != 0 ) {
                const uint32_t src = static_cast<uint32_t>( __ffs( static_cast<int>( pending ) ) ) - 1U;
                pending &= pending - 1U;
                const uint32_t a = __shfl_sync( FULL, at, src );
                const uint32_t cc = __shfl_sync( FULL, c, src );
                const uint32_t bb = __shfl_sync( FULL, byte, src );
                const uint32_t end = ( a < n ) ? a + min( cc, n - a ) : a;
                /* byte stores up to a 4-byte boundary, then words, then the tail */
                const uint32_t alignedBegin = min( end, ( a + 3U ) & ~3U );
                const uint32_t alignedEnd = alignedBegin + ( ( end - alignedBegin ) & ~3U );
                if ( a + lane < alignedBegin ) {
                    lbytes[a + lane] = static_cast<uint8_t>( bb );
                }
                uint32_t* const words = reinterpret_cast<uint32_t*>( lbytes );
                for ( uint32_t w = ( alignedBegin >> 2 ) + lane; w < ( alignedEnd >> 2 ); w += 32U ) {
                    words[w] = bb * 0x01010101U;
                }
                if ( alignedEnd + lane < end ) {
                    lbytes[alignedEnd + lane] = static_cast<uint8_t>( bb );
                }
            }
        }
    }
}
}  // namespace

void
launchMtf( BlockMeta* meta, uint32_t nblocks, const BlockTables* tables, uint32_t* scratch, uint8_t* lbytes,
           cudaStream_t stream, cudaEvent_t chunkStop )
{
    if ( nblocks == 0 ) {
        return;
    }
    {
        const dim3 grid( ( MAX_CHUNKS + MTFC_THREADS - 1U ) / MTFC_THREADS, nblocks );
        mtfChunkKernel<<<grid, MTFC_THREADS, 0, stream>>>( meta, scratch );
    }
    if ( chunkStop != nullptr ) {
        cudaEventRecord( chunkStop, stream );
    }
    mtfListsKernel<<<( nblocks + MTFL_WARPS - 1U ) / MTFL_WARPS, MTFL_WARPS * 32, 0, stream>>>( meta, nblocks, tables,
                                                                                             scratch );
    {
        const dim3 grid( ( MAX_CHUNKS + MTFE_WARPS - 1U ) / MTFE_WARPS, nblocks );
        cudaFuncSetAttribute( mtfExpandKernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared );
        mtfExpandKernel<<<grid, MTFE_WARPS * 32, 0, stream>>>( meta, scratch, lbytes );
    }
}

void
launchHuffmanDecode( const uint8_t* data, uint64_t nbytes, BlockMeta* meta, uint32_t nblocks, uint8_t* lbytes,
                     uint8_t* selectors, BlockTables* tables, uint32_t* scratch, uint32_t* counter, bool ordered, cudaStream_t stream,
                     cudaEvent_t bitsStart, cudaEvent_t bitsStop, cudaEvent_t chunkStop )
{
    launchHuffmanHeader( data, nbytes, meta, nblocks, selectors, tables, stream );
    if ( bitsStart != nullptr ) {
        cudaEventRecord( bitsStart, stream );
    }
    launchHuffmanBits( data, nbytes, meta, nblocks, selectors, tables, scratch, counter, ordered, stream );
    if ( bitsStop != nullptr ) {
        cudaEventRecord( bitsStop, stream );
    }
    launchMtf( meta, nblocks, tables, scratch, lbytes, stream, chunkStop );
    launchHuffmanCareful( data, nbytes, meta, nblocks, lbytes, selectors, stream );
    launchStreamProbe( data, nbytes, meta, nblocks, stream );
}
}  // namespace ibz2
