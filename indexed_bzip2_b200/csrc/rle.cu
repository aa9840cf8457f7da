/*
 * K5 -- undo of bzip2's first run-length stage (the RLE1 half of BurrowsWheelerTransformData::decodeBlock,
 * indexed_bzip2/bzip2.hpp:881-896): after four equal bytes the next byte is a repeat count 0..255.
 *
 * The reference runs this as a byte-serial state machine.  Here it is a parallel scan:
 *   state r in {0..4} before each element: r = number of equal data bytes seen (r = writeRun + 1 in the reference's
 *   variables), r == 4 means "this element is a count byte", r == 0 "fresh start" (block start or after a count byte).
 *   One element maps r -> r' depending only on eq = ( byte == previous byte ), so a run of elements is a function
 *   {0..4} -> {0..4}; functions compose associatively and are packed in 15 bits.
 * K5a (rleScanKernel): per block, 32 warps each own a contiguous element range.  Pass 1 reduces the range functions,
 *   one thread chains the 32 start states, pass 2 re-scans each row of 128 elements with known start state and writes
 *   per-element output offsets (relative to the warp's range; bit 31 flags count bytes) plus the block's output size.
 * layoutKernel: exclusive scan of the block sizes -> output offsets (blocks are laid out back to back), capacity check,
 *   and the piece table for the CRC kernel.
 * K5b (rleWriteKernel): lane-per-element scatter of the data bytes (neighbouring lanes hit neighbouring bytes) and
 *   warp-cooperative expansion of the count bytes (32 bytes per store instruction).
 * Algorithmic HBM traffic per block: N (bytes) + 4N (offsets) written and read once, U output bytes written.
 */
#include "common.cuh"

namespace ibz2
{
namespace
{
constexpr int RLE_THREADS = 1024;
constexpr uint32_t FULL = 0xFFFFFFFFU;
/* a transfer function {0..4} -> {0..4} as five nibbles: nibble s = f( s ) */
constexpr uint32_t IDENT = 0x43210U;
constexpr uint32_t STEP_NE = 0x01111U;
constexpr uint32_t STEP_EQ = 0x04321U;
constexpr uint32_t COUNT_FLAG = 0x80000000U;
/* word 0 of a row's 128 offsets: the row is "plain" (plainRowEndState) and the next word holds all there is to know
 * about it; offsets are relative to a range of one block's output and stay far below 2^30 */
constexpr uint32_t PLAIN_FLAG = 0x40000000U;

__device__ __forceinline__ uint32_t
applyFn( uint32_t f, uint32_t state )
{
    return ( f >> ( 4U * state ) ) & 0xFU;
}

/* first f, then g.  g becomes a byte table ( g(0) .. g(3) | g(4) ), f's nibbles are exactly the selector of a byte
 * permute that looks them up in it: ~15 instructions instead of five nested shift-and-mask lookups. */
__device__ __forceinline__ uint32_t
composeFn( uint32_t f, uint32_t g )
{
    uint32_t table = g & 0xFFFFU;
    table = ( table | ( table << 8 ) ) & 0x00FF00FFU;
    table = ( table | ( table << 4 ) ) & 0x0F0F0F0FU;
    const uint32_t g4 = g >> 16;
    uint32_t h = __byte_perm( table, g4, f );                        /* bytes g(f(0)) .. g(f(3)) */
    const uint32_t h4 = __byte_perm( table, g4, f >> 16 ) & 0xFU;    /* g(f(4)) */
    h = ( h | ( h >> 4 ) ) & 0x00FF00FFU;
    h = ( h | ( h >> 8 ) ) & 0xFFFFU;
    return h | ( h4 << 16 );
}

__device__ __forceinline__ uint32_t
stepState( uint32_t r, bool eq )
{
    return ( r == 4U ) ? 0U : ( ( eq && ( r != 0U ) ) ? r + 1U : 1U );
}

struct RowInput
{
    uint32_t bytes;   /* 4 elements, little endian */
    uint32_t eqMask;  /* bit k: element k equals its predecessor */
    uint32_t valid;   /* number of valid elements 0..4 */
    uint32_t fn;      /* packed transfer function of the lane's elements */
};

/* Loads the lane's 4 elements of the row starting at element `row` (multiple of 128) of a range ending at `end`.
 * carry = last byte of the previous row (or the byte before the range), carryValid = such a byte exists. */
__device__ __forceinline__ RowInput
loadRow( const uint8_t* __restrict__ prerle, uint32_t row, uint32_t end, uint32_t lane, uint32_t carry,
         const uint32_t* __restrict__ fnTable )
{
    RowInput in;
    const uint32_t j = row + 4U * lane;
    in.valid = j >= end ? 0U : ( end - j >= 4U ? 4U : end - j );
    in.bytes = in.valid > 0 ? *reinterpret_cast<const uint32_t*>( prerle + j ) : 0U;
    uint32_t prev = __shfl_up_sync( FULL, in.bytes >> 24, 1 );
    if ( lane == 0 ) {
        prev = carry;
    }
    const uint32_t b0 = in.bytes & 0xFFU;
    const uint32_t b1 = ( in.bytes >> 8 ) & 0xFFU;
    const uint32_t b2 = ( in.bytes >> 16 ) & 0xFFU;
    const uint32_t b3 = in.bytes >> 24;
    in.eqMask = ( b0 == prev ? 1U : 0U ) | ( b1 == b0 ? 2U : 0U ) | ( b2 == b1 ? 4U : 0U ) | ( b3 == b2 ? 8U : 0U );
    in.fn = IDENT;
    return in;
}

/* transfer function of the lane's elements */
__device__ __forceinline__ uint32_t
laneFn( const RowInput& in, const uint32_t* __restrict__ fnTable )
{
    if ( in.valid == 4U ) {
        return fnTable[in.eqMask];
    }
    uint32_t f = IDENT;
    for ( uint32_t k = 0; k < in.valid; ++k ) {
        f = composeFn( f, ( ( in.eqMask >> k ) & 1U ) ? STEP_EQ : STEP_NE );
    }
    return f;
}

/* A full row whose first element differs from its predecessor and that holds no three "equal" marks in a row cannot
 * reach state 4 from the inside: the only possible count byte is its first element (when the state before the row is
 * 4), every other element is one output byte, and the state behind the row is 1 + the number of marks at its end
 * whatever the state before it was.  That is nearly every row of data without long runs, and it needs neither the
 * scan of transfer functions nor the scan of output sizes.  Returns 1 + trailing marks, or 0 for "take the scan". */
__device__ __forceinline__ uint32_t
plainRowEndState( const RowInput& in, bool fullRow, uint32_t lane )
{
    const uint32_t up = __shfl_up_sync( FULL, in.eqMask, 1 );
    const uint32_t window = ( in.eqMask << 2 ) | ( lane == 0 ? 0U : ( up >> 2 ) );   /* marks -2 .. 3 of the lane */
    const bool triple = ( window & ( window >> 1 ) & ( window >> 2 ) ) != 0;
    const bool firstMarked = ( lane == 0 ) && ( ( in.eqMask & 1U ) != 0 );
    const uint32_t lastMask = __shfl_sync( FULL, in.eqMask, 31 );
    if ( !fullRow || __any_sync( FULL, triple || firstMarked ) ) {
        return 0;
    }
    return ( lastMask & 8U ) == 0 ? 1U : ( ( lastMask & 4U ) == 0 ? 2U : 3U );
}

__device__ __forceinline__ uint32_t
warpInclusiveCompose( uint32_t f, uint32_t lane )
{
#pragma unroll
    for ( int d = 1; d < 32; d <<= 1 ) {
        const uint32_t g = __shfl_up_sync( FULL, f, d );
        if ( lane >= static_cast<uint32_t>( d ) ) {
            f = composeFn( g, f );
        }
    }
    return f;
}

__global__ void __launch_bounds__( RLE_THREADS, 2 )
rleScanKernel( const uint8_t* __restrict__ prerleAll, uint32_t* __restrict__ offsetsAll,
               uint32_t* __restrict__ warpBaseAll, BlockMeta* __restrict__ meta )
{
    __shared__ uint32_t fnTable[16];
    __shared__ uint32_t rangeFn[RLE_RANGES];
    __shared__ uint32_t rangeStart[RLE_RANGES];
    __shared__ uint32_t rangeTotal[RLE_RANGES];

    BlockMeta& bm = meta[blockIdx.x];
    if ( ( bm.status != IBZ2_ST_OK ) || ( bm.n == 0 ) ) {
        return;
    }
    const uint32_t n = bm.n;
    const uint8_t* __restrict__ prerle = prerleAll + static_cast<uint64_t>( blockIdx.x ) * LBYTES_STRIDE;
    uint32_t* __restrict__ offsets = offsetsAll + static_cast<uint64_t>( blockIdx.x ) * DBUF_STRIDE;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31U;

    if ( threadIdx.x < 16 ) {
        uint32_t f = IDENT;
        for ( uint32_t k = 0; k < 4U; ++k ) {
            f = composeFn( f, ( ( threadIdx.x >> k ) & 1U ) ? STEP_EQ : STEP_NE );
        }
        fnTable[threadIdx.x] = f;
    }
    __syncthreads();

    const uint32_t rangeLen = ( ( ( n + RLE_RANGES - 1 ) / RLE_RANGES ) + 127U ) & ~127U;
    const uint32_t begin = min( n, warp * rangeLen );
    const uint32_t end = min( n, begin + rangeLen );
    const uint32_t firstCarry = begin > 0 ? prerle[begin - 1] : 0x100U;  /* 0x100 never equals a byte */

    /* pass 1: transfer function of the whole range */
    {
        uint32_t f = IDENT;
        uint32_t carry = firstCarry;
        for ( uint32_t row = begin; row < end; row += 128U ) {
            const RowInput in = loadRow( prerle, row, end, lane, carry, fnTable );
            carry = __shfl_sync( FULL, in.bytes >> 24, 31 );
            const uint32_t plainEnd = plainRowEndState( in, row + 128U <= end, lane );
            if ( plainEnd != 0 ) {
                f = plainEnd * 0x11111U;   /* a constant function swallows what came before it */
                continue;
            }
            const uint32_t incl = warpInclusiveCompose( laneFn( in, fnTable ), lane );
            f = composeFn( f, __shfl_sync( FULL, incl, 31 ) );
        }
        if ( lane == 0 ) {
            rangeFn[warp] = f;
        }
    }
    __syncthreads();
    if ( threadIdx.x == 0 ) {
        uint32_t state = 0;  /* writeRun = -1 at the start of a block (bzip2.hpp:841-844) */
        for ( uint32_t w = 0; w < RLE_RANGES; ++w ) {
            rangeStart[w] = state;
            state = applyFn( rangeFn[w], state );
        }
    }
    __syncthreads();

    /* pass 2: per-element output offsets */
    {
        uint32_t rowState = rangeStart[warp];
        uint32_t running = 0;
        uint32_t carry = firstCarry;
        for ( uint32_t row = begin; row < end; row += 128U ) {
            const RowInput in = loadRow( prerle, row, end, lane, carry, fnTable );
            carry = __shfl_sync( FULL, in.bytes >> 24, 31 );
            const uint32_t plainEnd = plainRowEndState( in, row + 128U <= end, lane );
            if ( plainEnd != 0 ) {
                /* element 0 is a count byte when the row starts in state 4, everything else is one byte */
                const bool countFirst = rowState == 4U;
                const uint32_t firstByte = __shfl_sync( FULL, in.bytes & 0xFFU, 0 );
                const uint32_t firstEmit = countFirst ? firstByte : 1U;
                if ( lane == 0 ) {
                    /* K5b works the offsets out itself: element 0 at `running` (firstEmit copies when it is a count
                     * byte), element e > 0 at running + firstEmit - 1 + e */
                    *reinterpret_cast<uint4*>( offsets + row ) =
                        make_uint4( PLAIN_FLAG | running, firstEmit | ( countFirst ? COUNT_FLAG : 0U ), 0U, 0U );
                }
                running += 127U + firstEmit;
                rowState = plainEnd;
                continue;
            }
            const uint32_t incl = warpInclusiveCompose( laneFn( in, fnTable ), lane );
            uint32_t excl = __shfl_up_sync( FULL, incl, 1 );
            if ( lane == 0 ) {
                excl = IDENT;
            }
            uint32_t r = applyFn( excl, rowState );

            uint32_t emit[4];
            uint32_t flags = 0;
            uint32_t laneTotal = 0;
#pragma unroll
            for ( uint32_t k = 0; k < 4U; ++k ) {
                const uint32_t b = ( in.bytes >> ( 8U * k ) ) & 0xFFU;
                uint32_t e = 0;
                if ( k < in.valid ) {
                    if ( r == 4U ) {
                        e = b;
                        flags |= 1U << k;
                    } else {
                        e = 1U;
                    }
                    r = stepState( r, ( ( in.eqMask >> k ) & 1U ) != 0 );
                }
                emit[k] = e;
                laneTotal += e;
            }
            uint32_t scan = laneTotal;
#pragma unroll
            for ( int d = 1; d < 32; d <<= 1 ) {
                const uint32_t up = __shfl_up_sync( FULL, scan, d );
                if ( lane >= static_cast<uint32_t>( d ) ) {
                    scan += up;
                }
            }
            uint32_t o = running + scan - laneTotal;
            uint4 packed;
            packed.x = o | ( ( flags & 1U ) ? COUNT_FLAG : 0U );
            o += emit[0];
            packed.y = o | ( ( flags & 2U ) ? COUNT_FLAG : 0U );
            o += emit[1];
            packed.z = o | ( ( flags & 4U ) ? COUNT_FLAG : 0U );
            o += emit[2];
            packed.w = o | ( ( flags & 8U ) ? COUNT_FLAG : 0U );
            if ( in.valid > 0 ) {
                *reinterpret_cast<uint4*>( offsets + row + 4U * lane ) = packed;
            }
            running += __shfl_sync( FULL, scan, 31 );
            rowState = applyFn( __shfl_sync( FULL, incl, 31 ), rowState );
        }
        if ( lane == 0 ) {
            rangeTotal[warp] = running;
        }
    }
    __syncthreads();
    if ( threadIdx.x == 0 ) {
        uint32_t total = 0;
        uint32_t* const warpBase = warpBaseAll + static_cast<uint64_t>( blockIdx.x ) * RLE_RANGES;
        for ( uint32_t w = 0; w < RLE_RANGES; ++w ) {
            warpBase[w] = total;
            total += rangeTotal[w];
        }
        bm.outSize = total;
    }
}

/* One CTA: output offsets of the wave's blocks, capacity check, CRC window ("piece") table. */
constexpr int LAYOUT_THREADS = 1024;

__device__ __forceinline__ uint32_t
piecesOfBlock( uint64_t start, uint32_t size, uint32_t pieceBytes )
{
    /* must match blockRange() in crc.cu: windows of pieceBytes that intersect [start, mainEnd) */
    const uint64_t end = start + size;
    const uint64_t aligned = end & ~3ULL;
    if ( aligned <= start ) {
        return 0;
    }
    return static_cast<uint32_t>( ( aligned - 1U ) / pieceBytes - start / pieceBytes + 1U );
}

__global__ void __launch_bounds__( LAYOUT_THREADS )
layoutKernel( BlockMeta* __restrict__ meta, uint32_t nblocks, uint64_t* __restrict__ running, uint64_t outCapacity,
              int haveOutput, uint32_t* __restrict__ pieceStart, uint32_t pieceBytes )
{
    __shared__ uint64_t sums[LAYOUT_THREADS];
    const uint32_t tid = threadIdx.x;
    const uint32_t per = ( nblocks + LAYOUT_THREADS - 1 ) / LAYOUT_THREADS;
    const uint32_t first = min( nblocks, tid * per );
    const uint32_t last = min( nblocks, first + per );

    /* scan 1: output sizes -> offsets */
    uint64_t mine = 0;
    for ( uint32_t b = first; b < last; ++b ) {
        const bool ok = ( meta[b].status == IBZ2_ST_OK ) && ( meta[b].n > 0 );
        mine += ok ? meta[b].outSize : 0U;
    }
    sums[tid] = mine;
    __syncthreads();
    for ( uint32_t d = 1; d < LAYOUT_THREADS; d <<= 1 ) {
        const uint64_t s = tid >= d ? sums[tid - d] : 0U;
        __syncthreads();
        sums[tid] += s;
        __syncthreads();
    }
    const uint64_t base = running[0];
    const uint64_t grandTotal = sums[LAYOUT_THREADS - 1];
    uint64_t offset = base + sums[tid] - mine;
    uint64_t myPieces = 0;
    for ( uint32_t b = first; b < last; ++b ) {
        const bool ok = ( meta[b].status == IBZ2_ST_OK ) && ( meta[b].n > 0 );
        const uint32_t size = ok ? meta[b].outSize : 0U;
        meta[b].outOffset = offset;
        uint32_t write = 0;
        if ( ok && ( haveOutput != 0 ) ) {
            if ( offset + size <= outCapacity ) {
                write = 1;
            } else {
                meta[b].status = IBZ2_ST_OUTPUT_CAPACITY;
            }
        }
        meta[b].writeOutput = write;
        if ( write != 0 ) {
            myPieces += piecesOfBlock( offset, size, pieceBytes );
        }
        offset += size;
    }
    __syncthreads();

    /* scan 2: CRC windows per block */
    sums[tid] = myPieces;
    __syncthreads();
    for ( uint32_t d = 1; d < LAYOUT_THREADS; d <<= 1 ) {
        const uint64_t s = tid >= d ? sums[tid - d] : 0U;
        __syncthreads();
        sums[tid] += s;
        __syncthreads();
    }
    uint32_t piece = static_cast<uint32_t>( sums[tid] - myPieces );
    for ( uint32_t b = first; b < last; ++b ) {
        pieceStart[b] = piece;
        if ( meta[b].writeOutput != 0 ) {
            piece += piecesOfBlock( meta[b].outOffset, meta[b].outSize, pieceBytes );
        }
    }
    if ( tid == LAYOUT_THREADS - 1 ) {
        running[0] = base + grandTotal;
        pieceStart[nblocks] = static_cast<uint32_t>( sums[tid] );
    }
}

__global__ void __launch_bounds__( RLE_THREADS )
rleWriteKernel( const uint8_t* __restrict__ prerleAll, const uint32_t* __restrict__ offsetsAll,
                const uint32_t* __restrict__ warpBaseAll, const BlockMeta* __restrict__ meta,
                uint8_t* __restrict__ out )
{
    const BlockMeta& bm = meta[blockIdx.x];
    if ( ( bm.status != IBZ2_ST_OK ) || ( bm.n == 0 ) || ( bm.writeOutput == 0 ) ) {
        return;
    }
    const uint32_t n = bm.n;
    const uint8_t* __restrict__ prerle = prerleAll + static_cast<uint64_t>( blockIdx.x ) * LBYTES_STRIDE;
    const uint32_t* __restrict__ offsets = offsetsAll + static_cast<uint64_t>( blockIdx.x ) * DBUF_STRIDE;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31U;

    const uint32_t rangeLen = ( ( ( n + RLE_RANGES - 1 ) / RLE_RANGES ) + 127U ) & ~127U;
    const uint32_t begin = min( n, warp * rangeLen );
    const uint32_t end = min( n, begin + rangeLen );
    uint8_t* __restrict__ dst = out + bm.outOffset + warpBaseAll[static_cast<uint64_t>( blockIdx.x ) * RLE_RANGES + warp];

    uint32_t carry = begin > 0 ? prerle[begin - 1] : 0U;
    for ( uint32_t row = begin; row < end; row += 128U ) {
        const uint32_t j = row + 4U * lane;
        const uint32_t valid = j >= end ? 0U : ( end - j >= 4U ? 4U : end - j );
        uint32_t bytes = 0;
        uint4 offs = make_uint4( 0, 0, 0, 0 );
        const uint4 head = *reinterpret_cast<const uint4*>( offsets + row );   /* the same 16 bytes for every lane */
        if ( ( head.x & PLAIN_FLAG ) != 0 ) {
            /* plain row (always a full one): see rleScanKernel */
            bytes = *reinterpret_cast<const uint32_t*>( prerle + j );
            const uint32_t base = head.x & ~PLAIN_FLAG;
            const uint32_t o = base + ( head.y & ~COUNT_FLAG ) - 1U + 4U * lane;
            offs = make_uint4( lane == 0 ? ( base | ( head.y & COUNT_FLAG ) ) : o, o + 1U, o + 2U, o + 3U );
        } else if ( valid > 0 ) {
            bytes = *reinterpret_cast<const uint32_t*>( prerle + j );
            offs = *reinterpret_cast<const uint4*>( offsets + j );
        }
        uint32_t prev = __shfl_up_sync( FULL, bytes >> 24, 1 );
        if ( lane == 0 ) {
            prev = carry;
        }
        const uint32_t off[4] = { offs.x, offs.y, offs.z, offs.w };
#pragma unroll
        for ( uint32_t k = 0; k < 4U; ++k ) {
            const uint32_t b = ( bytes >> ( 8U * k ) ) & 0xFFU;
            const bool isCount = ( off[k] & COUNT_FLAG ) != 0;
            const uint32_t o = off[k] & ~COUNT_FLAG;
            const bool live = k < valid;
            if ( live && !isCount ) {
                dst[o] = static_cast<uint8_t>( b );
            }
            /* count bytes: the warp expands them together, 32 bytes per store */
            uint32_t pending = __ballot_sync( FULL, live && isCount && ( b > 0 ) );
            while ( pending != 0 ) {
                const int src = __ffs( static_cast<int>( pending ) ) - 1;
                pending &= pending - 1U;
                const uint32_t runOffset = __shfl_sync( FULL, o, src );
                const uint32_t runLength = __shfl_sync( FULL, b, src );
                const uint32_t runByte = __shfl_sync( FULL, prev, src );
                for ( uint32_t i = lane; i < runLength; i += 32U ) {
                    dst[runOffset + i] = static_cast<uint8_t>( runByte );
                }
            }
            prev = b;
        }
        carry = __shfl_sync( FULL, bytes >> 24, 31 );
    }
}
}  // namespace

void
launchRleScan( const uint8_t* prerle, uint32_t* offsets, uint32_t* warpBase, BlockMeta* meta, uint32_t nblocks,
               cudaStream_t stream )
{
    rleScanKernel<<<nblocks, RLE_THREADS, 0, stream>>>( prerle, offsets, warpBase, meta );
}

void
launchOutputLayout( BlockMeta* meta, uint32_t nblocks, uint64_t* running, uint64_t outCapacity, int haveOutput,
                    uint32_t* pieceStart, uint32_t pieceBytes, cudaStream_t stream )
{
    layoutKernel<<<1, LAYOUT_THREADS, 0, stream>>>( meta, nblocks, running, outCapacity, haveOutput, pieceStart,
                                                    pieceBytes );
}

void
launchRleWrite( const uint8_t* prerle, const uint32_t* offsets, const uint32_t* warpBase, const BlockMeta* meta,
                uint32_t nblocks, uint8_t* out, cudaStream_t stream )
{
    rleWriteKernel<<<nblocks, RLE_THREADS, 0, stream>>>( prerle, offsets, warpBase, meta, out );
}
}  // namespace ibz2
