/*
 * K2b for launches with few blocks, second version: one CTA per bzip2 block as in huffmanGroupsKernel, but the lookups
 * for the NEXT group of 50 symbols run while the chain of the CURRENT one is followed.  Included by huffman_bits.cu.
 *
 * What a ring entry says -- "a symbol that starts at this bit ends there" -- depends on the bit position and on the table,
 * not on where the groups start.  The table of every group is known in advance (the selectors), and the next group starts
 * between 50 and 450 bits behind the current one.  So while warp 16 follows the chain of group g through ring set g % 3,
 * the 16 builder warps look up all 1024 bit offsets from ( start of g + 50 ) & ~31 on with the table of group g + 1 and
 * double them to hops of 2, 4, 8 symbols into set ( g + 1 ) % 3, and warps 17-18 store the entries of group g - 1 out
 * of set ( g - 1 ) % 3.  One CTA barrier per group; the builders use a barrier of their own between the doubling rounds.
 * A group that is not 50 plain symbols of at most 9 bits (long code, end of block) is walked one symbol at a time by the
 * chain warp; if it ends outside the range that was looked up for its successor, one round goes to looking up again.
 * Same entries, same exits to the careful decoder as the other kernels.
 */
constexpr int PIPE_BUILD_THREADS = 512;
constexpr uint32_t PIPE_CHAIN_WARP = PIPE_BUILD_THREADS / 32;
constexpr uint32_t PIPE_OUT_WARP = PIPE_CHAIN_WARP + 1U;   /* and the one behind it */
constexpr int PIPE_THREADS = PIPE_BUILD_THREADS + 3 * 32;
constexpr uint32_t PIPE_RING = 1024;             /* bit offsets per set */
#ifndef IBZ2_PIPE_WINDOW
#define IBZ2_PIPE_WINDOW 0
#endif
#ifndef IBZ2_PIPE_LEVELS
#define IBZ2_PIPE_LEVELS 4
#endif
constexpr uint32_t PIPE_LEVELS = IBZ2_PIPE_LEVELS;            /* 4: hops of 1, 2, 4, 8 symbols; 3: up to 4 */
constexpr uint32_t PIPE_HOP = 1U << ( PIPE_LEVELS - 1U );     /* symbols of the longest hop */
constexpr uint32_t PIPE_HOPS = ( GROUP_SIZE - 2U ) / PIPE_HOP; /* long hops per group; one hop of 2 behind them */
static_assert( PIPE_HOPS * PIPE_HOP + 2U == GROUP_SIZE, "48 symbols in long hops, 2 in a hop of two" );
constexpr uint32_t PIPE_SETS = 3;
constexpr uint32_t PIPE_FLAG = 0x8000U;          /* the hop is whole: 2^level plain symbols */
constexpr uint32_t PIPE_TARGET = 0x07FFU;        /* offset behind the hop; >= PIPE_RING: outside the set */
constexpr uint32_t PIPE_WORDS = 256;             /* staged big-endian words of the file (a ring) */
constexpr uint32_t PIPE_REFILL = 64;
constexpr uint32_t PIPE_AHEAD = 128;             /* words staged in front of the set being built, at least */
/* a group may start this far into a set: 48 symbols of 9 bits behind it must have entries */
constexpr uint32_t PIPE_MAX_START = PIPE_RING - ( GROUP_SIZE - 2U ) * HUFF_LUT_BITS - 32U;
constexpr uint32_t PIPE_SKIP = GROUP_SIZE;       /* the next group starts at least this many bits behind the current one */
static_assert( ( GROUP_SIZE * HUFF_LUT_BITS - PIPE_SKIP ) + 31U <= PIPE_MAX_START, "the successor of a plain group is covered" );

/* what the chain warp tells the others at the end of a round: one 16-byte store, one 16-byte load */
constexpr uint32_t PIPE_OUTCOME = 3U;            /* 0 = go on, 1 = end-of-block symbol seen, 2 = careful decoder */
constexpr uint32_t PIPE_REBUILD = 4U;            /* group g has no usable set: look it up from its first bit, no chain this round */
constexpr uint32_t PIPE_STORE = 8U;              /* group g - 1 was 50 plain symbols: their entries are still to be stored */
/* The bit offsets of a set that are worth looking up.  A group of 50 plain symbols takes between 50 x the shortest and
 * 50 x the longest code that the block's 9-bit tables hold, so the first bit of the next group and its last one lie in a
 * range that is the same for every group of the block (about 700 of the 1024 offsets for data whose codes are all of
 * similar length); it is worked out once per block.  A group with an exceptional symbol may end anywhere: the chain warp
 * checks that the next group's first bit is inside the range and has it looked up again from that bit if it is not. */
struct PipeWindow
{
    uint32_t lo;   /* multiples of 32 */
    uint32_t hi;
};

struct PipeLengths
{
    uint32_t shortest;   /* over all tables of the block: codes of at most 9 bits, without the end-of-block symbol */
    uint32_t longest;
};

__device__ __forceinline__ PipeWindow
pipeWindow( const PipeLengths lengths, bool fromFirstBit )
{
    PipeWindow w;
    w.lo = 0;
    w.hi = PIPE_RING;
    /* Measured on config 1 (windows of ~740 offsets): 13.5 ms with the window, 11.9 ms with all 1024 offsets -- a builder warp
     * with two rows of offsets takes as long as before and the warps with one row wait for it, so a window only pays once
     * it is at most 512 offsets wide.  Kept for that case; off by default. */
    if ( ( IBZ2_PIPE_WINDOW != 0 ) && ( lengths.shortest <= lengths.longest ) ) {
        if ( fromFirstBit ) {
            /* a group from the word its first bit is in */
            w.hi = min( PIPE_RING, ( 31U + GROUP_SIZE * lengths.longest + 63U ) & ~31U );
        } else {
            /* a group from 50 bits behind the first bit of the group in front of it */
            w.lo = ( GROUP_SIZE * lengths.shortest - PIPE_SKIP ) & ~31U;
            w.hi = min( PIPE_RING, ( 31U + GROUP_SIZE * lengths.longest - PIPE_SKIP + GROUP_SIZE * lengths.longest + 63U ) & ~31U );
        }
    }
    return w;
}

struct PipeState
{
    unsigned long long X;   /* first bit of group g */
    uint32_t g;             /* the group to chain next */
    uint32_t flags;
};

struct PipeShared
{
    alignas( 16 ) uint8_t dec[( DEC_STRIDE + 15U ) & ~15U];
    alignas( 16 ) PipeState state[2];
    uint32_t block;
    uint32_t skip;
    uint32_t nsym;          /* symbols in front of state.X, written when the block ends */
    PipeLengths lengths;
    uint32_t words[PIPE_WORDS];
    uint16_t ring[PIPE_SETS][PIPE_LEVELS][PIPE_RING];
    uint16_t idx[PIPE_SETS][PIPE_RING];
    uint16_t coarse[PIPE_SETS][16];
    alignas( 16 ) uint8_t selectors[MAX_SELECTORS + 64U];
};

__device__ __forceinline__ void
pipeBuilderBarrier()
{
    asm volatile( "bar.sync 1, %0;" ::"n"( PIPE_BUILD_THREADS ) : "memory" );
}

__global__ void __launch_bounds__( PIPE_THREADS, 2 )
huffmanPipelineKernel( const uint8_t* __restrict__ data, uint64_t nbytes, BlockMeta* __restrict__ meta, uint32_t nblocks,
                       const uint8_t* __restrict__ selectorsAll, const BlockTables* __restrict__ tablesAll,
                       uint32_t* __restrict__ scratchAll, uint32_t* __restrict__ counter, const uint32_t* __restrict__ order )
{
    extern __shared__ __align__( 16 ) uint8_t pipeSharedRaw[];
    PipeShared& sh = *reinterpret_cast<PipeShared*>( pipeSharedRaw );
    const uint32_t t = threadIdx.x;
    const uint32_t warp = t >> 5;
    const uint32_t lane = t & 31U;
    const uint32_t decAddr = static_cast<uint32_t>( __cvta_generic_to_shared( sh.dec ) );
    const uint64_t nbits = nbytes * 8U;

    while ( true ) {
        __syncthreads();   /* everybody is done with the previous block's shared state */
        if ( t == 0 ) {
            sh.block = atomicAdd( counter, 1U );
        }
        __syncthreads();
        uint32_t block = sh.block;
        if ( block >= nblocks ) {
            break;
        }
        if ( order != nullptr ) {
            block = order[block];
        }
        BlockMeta& bm = meta[block];
        const BlockTables* const bt = tablesAll + block;
        if ( t == 0 ) {
            /* one thread looks and tells: what it writes must not split the CTA */
            uint32_t skip = ( ( bm.status != IBZ2_ST_OK ) || ( bm.isEos != 0 ) ) ? 1U : 0U;
            if ( ( skip == 0 ) && ( bt->symbolCount == 0 ) ) {
                bm.status = ST_INTERNAL_CAREFUL;   /* see huffmanBitsKernel */
                skip = 1;
            }
            sh.skip = skip;
        }
        __syncthreads();
        if ( sh.skip != 0 ) {
            continue;
        }
        const uint32_t symbolCount = bt->symbolCount;
        const uint32_t groupCount = bt->groupCount;
        const uint32_t selectorsCount = bt->selectorsCount;
        const uint8_t* const selectors = selectorsAll + static_cast<uint64_t>( block ) * SEL_STRIDE;
        uint16_t* const symOut = reinterpret_cast<uint16_t*>( reinterpret_cast<uint8_t*>( scratchAll + static_cast<uint64_t>( block ) * DBUF_STRIDE )
                                                              + K2_SYM16_OFFSET );
        if ( t == 0 ) {
            sh.lengths.shortest = 0xFFU;
            sh.lengths.longest = 0;
        }
        __syncthreads();
        {
            /* the same tables as one decoder of huffmanBitsKernel keeps, built by all threads */
            const uint4* const src = reinterpret_cast<const uint4*>( &bt->lut[0][0] );
            uint32_t* const dst = reinterpret_cast<uint32_t*>( sh.dec );
            for ( uint32_t q = t; q < groupCount * ( HUFF_LUT_SIZE / 8U ); q += PIPE_THREADS ) {
                const uint4 v = __ldg( src + q );
                const uint32_t w[4] = { v.x, v.y, v.z, v.w };
                uint32_t lens[2] = { 0, 0 };
#pragma unroll
                for ( uint32_t i = 0; i < 4U; ++i ) {
                    const uint32_t a = ( ( w[i] & 0xFFFFU ) >> 5 ) > symbolCount ? 0U : ( w[i] & 31U );
                    const uint32_t b = ( w[i] >> 21 ) > symbolCount ? 0U : ( ( w[i] >> 16 ) & 31U );
                    lens[i >> 1] |= ( a | ( b << 8 ) ) << ( 16U * ( i & 1U ) );
                }
                dst[2U * q] = lens[0];
                dst[2U * q + 1U] = lens[1];
                uint32_t shortest = 0xFFU;
                uint32_t longest = 0;
#pragma unroll
                for ( uint32_t i = 0; i < 8U; ++i ) {
                    const uint32_t len = ( lens[i >> 2] >> ( 8U * ( i & 3U ) ) ) & 0xFFU;
                    shortest = len != 0 ? min( shortest, len ) : shortest;
                    longest = max( longest, len );
                }
                atomicMin( &sh.lengths.shortest, shortest );
                atomicMax( &sh.lengths.longest, longest );
            }
            for ( uint32_t i = t; i < groupCount * LONG_LIMITS; i += PIPE_THREADS ) {
                const uint64_t limit = bt->limit[i / LONG_LIMITS][LONG_FIRST_LEN + i % LONG_LIMITS];
                *reinterpret_cast<uint32_t*>( sh.dec + DEC_LIM_OFFSET + i * 4U ) = static_cast<uint32_t>( limit >> 12 );
            }
            for ( uint32_t i = t; i < groupCount * LONG_OFFSETS; i += PIPE_THREADS ) {
                *reinterpret_cast<uint16_t*>( sh.dec + DEC_OFF_OFFSET + i * 2U )
                    = bt->offset[i / LONG_OFFSETS][LONG_FIRST_LEN + 1U + i % LONG_OFFSETS];
            }
            if ( t < groupCount ) {
                *reinterpret_cast<uint16_t*>( sh.dec + DEC_EOB_OFFSET + t * 2U ) = bt->eobPerm[t];
            }
        }
        /* the table of every group (bzip2.hpp:710-721), so that nobody waits for global memory to learn it */
        for ( uint32_t i = t; i < ( selectorsCount + 15U ) / 16U; i += PIPE_THREADS ) {
            reinterpret_cast<uint4*>( sh.selectors )[i] = __ldg( reinterpret_cast<const uint4*>( selectors ) + i );
        }
        const uint64_t startBit = bt->dataStartBit;
        for ( uint32_t i = t; i < PIPE_WORDS; i += PIPE_THREADS ) {
            const uint64_t w = ( startBit >> 5 ) + i;
            sh.words[w & ( PIPE_WORDS - 1U )] = bswap( loadWordRaw( data, nbytes, w ) );
        }
        if ( t == 0 ) {
            /* the first round only looks up: group 0, from the word its first bit is in */
            PipeState first;
            first.X = startBit;
            first.g = 0;
            first.flags = PIPE_REBUILD;
            sh.state[0] = first;
        }
        /* ---- what the roles keep between the rounds ---- */
        /* builders: file words in front of loadedUpTo are staged; a refill is loaded in one round and stored in the next */
        uint64_t loadedUpTo = ( startBit >> 5 ) + PIPE_WORDS;
        bool pending = false;
        uint32_t pendingRaw = 0;
        /* chain warp: offset 0 of the set of the group to chain, and the symbols in front of that group */
        uint64_t baseOfG = startBit & ~static_cast<uint64_t>( 31U );
        PipeWindow windowOfG{ 0, PIPE_RING };   /* set when the lengths are known, below */
        uint32_t nsym = 0;
        uint32_t round = 0;
        __syncthreads();
        const PipeWindow windowFirst = pipeWindow( sh.lengths, true );
        const PipeWindow windowBehind = pipeWindow( sh.lengths, false );
        windowOfG = windowFirst;

        while ( true ) {
            const uint4 stateRaw = *reinterpret_cast<const uint4*>( &sh.state[round & 1U] );
            const uint64_t X = ( static_cast<uint64_t>( stateRaw.y ) << 32 ) | stateRaw.x;
            const uint32_t g = stateRaw.z;
            const uint32_t flags = stateRaw.w;
            if ( ( flags & PIPE_OUTCOME ) != 0 ) {
                break;
            }
            const bool rebuild = ( flags & PIPE_REBUILD ) != 0;
            if ( warp < PIPE_CHAIN_WARP ) {
                /* ================= builders: group g from its first bit, or group g + 1 from 50 bits behind it ================= */
                const uint32_t group = rebuild ? g : g + 1U;
                const uint32_t set = group % PIPE_SETS;
                const uint64_t baseWord = ( rebuild ? X : X + PIPE_SKIP ) >> 5;
                const uint32_t lut = ( group < selectorsCount ? sh.selectors[group] : 0U ) * HUFF_LUT_SIZE;
                if ( pending ) {
                    if ( t < PIPE_REFILL ) {
                        /* these slots hold words more than 190 in front of the newest: nobody reads them any more */
                        sh.words[( loadedUpTo + t ) & ( PIPE_WORDS - 1U )] = bswap( pendingRaw );
                    }
                    loadedUpTo += PIPE_REFILL;
                    pending = false;
                }
                if ( loadedUpTo < baseWord + PIPE_AHEAD ) {
                    if ( t < PIPE_REFILL ) {
                        pendingRaw = loadWordRaw( data, nbytes, loadedUpTo + t );
                    }
                    pending = true;
                }
                const uint32_t slot = static_cast<uint32_t>( baseWord );
                const PipeWindow window = rebuild ? windowFirst : windowBehind;
                /* a thread looks at offsets first and first + 512 of the window, as far as they are inside; both limits are
                 * multiples of 32, so a warp takes one of three ways as a whole, and within a way the two offsets are
                 * independent instruction streams (a warp outside the window only keeps the others company at the barriers) */
                const uint32_t first = window.lo + t;
                const uint32_t count = first + PIPE_BUILD_THREADS < window.hi ? 2U : ( first < window.hi ? 1U : 0U );
                uint32_t e[2] = { 0, 0 };
                const auto lookUp = [&] ( auto n ) {
                    constexpr uint32_t N = decltype( n )::value;
                    uint32_t index[N];
#pragma unroll
                    for ( uint32_t j = 0; j < N; ++j ) {
                        const uint32_t r = first + j * PIPE_BUILD_THREADS;
                        const uint32_t w = slot + ( r >> 5 );
                        const uint32_t wa = sh.words[w & ( PIPE_WORDS - 1U )];
                        const uint32_t wb = sh.words[( w + 1U ) & ( PIPE_WORDS - 1U )];
                        index[j] = __funnelshift_l( wb, wa, r & 31U ) >> ( 32 - HUFF_LUT_BITS );
                    }
#pragma unroll
                    for ( uint32_t j = 0; j < N; ++j ) {
                        const uint32_t r = first + j * PIPE_BUILD_THREADS;
                        const uint32_t len = sh.dec[lut + index[j]];
                        e[j] = len != 0 ? ( ( r + len ) | PIPE_FLAG ) : r;
                        sh.ring[set][0][r] = static_cast<uint16_t>( e[j] );
                        sh.idx[set][r] = static_cast<uint16_t>( index[j] );
                    }
                };
                const auto doubleUp = [&] ( auto n, uint32_t level ) {
                    constexpr uint32_t N = decltype( n )::value;
                    uint32_t f[N];
#pragma unroll
                    for ( uint32_t j = 0; j < N; ++j ) {
                        f[j] = sh.ring[set][level - 1U][e[j] & ( PIPE_RING - 1U )];
                    }
#pragma unroll
                    for ( uint32_t j = 0; j < N; ++j ) {
                        /* what lies behind the window was looked up for an earlier group */
                        const bool whole = ( ( e[j] & PIPE_FLAG ) != 0 ) && ( ( e[j] & PIPE_TARGET ) < window.hi );
                        e[j] = whole ? f[j] : ( e[j] & ~PIPE_FLAG );
                        sh.ring[set][level][first + j * PIPE_BUILD_THREADS] = static_cast<uint16_t>( e[j] );
                    }
                };
                if ( count == 2U ) {
                    lookUp( std::integral_constant<uint32_t, 2>{} );
                } else if ( count == 1U ) {
                    lookUp( std::integral_constant<uint32_t, 1>{} );
                }
#pragma unroll
                for ( uint32_t level = 1; level < PIPE_LEVELS; ++level ) {
                    pipeBuilderBarrier();
                    if ( count == 2U ) {
                        doubleUp( std::integral_constant<uint32_t, 2>{}, level );
                    } else if ( count == 1U ) {
                        doubleUp( std::integral_constant<uint32_t, 1>{}, level );
                    }
                }
            } else if ( warp > PIPE_CHAIN_WARP ) {
                /* ================= output: the entries of group g - 1, chained in the round before ================= */
                const uint32_t s = lane + 32U * ( warp - PIPE_OUT_WARP );
                if ( ( ( flags & PIPE_STORE ) != 0 ) && ( s < GROUP_SIZE ) ) {
                    const uint32_t set = ( g - 1U ) % PIPE_SETS;
                    uint32_t q = sh.coarse[set][s / PIPE_HOP];
#pragma unroll
                    for ( uint32_t level = PIPE_LEVELS - 1U; level-- > 0; ) {
                        if ( ( s & ( 1U << level ) ) != 0 ) {
                            q = sh.ring[set][level][q] & PIPE_TARGET;
                        }
                    }
                    /* every group in front of this one had 50 symbols */
                    symOut[( g - 1U ) * GROUP_SIZE + s]
                        = static_cast<uint16_t>( sh.selectors[g - 1U] * HUFF_LUT_SIZE | sh.idx[set][q] );
                }
            } else {
                /* ================= the chain ================= */
                uint64_t nextX = X;
                uint32_t nextG = g;
                uint32_t nextFlags = 0;
                if ( !rebuild ) {
                    const uint32_t set = g % PIPE_SETS;
                    /* next group of 50 symbols: the exits the lane kernel has at this place (bzip2.hpp:710-721) */
                    if ( ( g >= selectorsCount ) || ( nsym >= SYM_LIMIT ) || ( X > nbits ) ) {
                        nextFlags = 2;
                    } else {
                        const uint32_t p0 = static_cast<uint32_t>( X - baseOfG );
                        const uint16_t* const hop8 = sh.ring[set][PIPE_LEVELS - 1U];
                        uint32_t p = p0;
                        uint32_t mine = 0;   /* lane k: where hop k starts */
                        uint32_t all = PIPE_FLAG;
                        uint32_t reach = 0;
#pragma unroll
                        for ( uint32_t hop = 0; hop < PIPE_HOPS; ++hop ) {
                            mine = lane == hop ? p : mine;
                            const uint32_t f = hop8[p & ( PIPE_RING - 1U )];
                            all &= f;
                            reach = max( reach, p );
                            p = f & PIPE_TARGET;
                        }
                        mine = lane == PIPE_HOPS ? p : mine;
                        reach = max( reach, p );
                        const uint32_t last = sh.ring[set][1][p & ( PIPE_RING - 1U )];
                        all &= last;
                        /* (the builders end a hop at the window's end, so a whole hop has only read entries of this group;
                         * where a hop STARTS is the chain's to check) */
                        if ( ( all != 0 ) && ( reach < windowOfG.hi ) ) {
                            if ( lane <= PIPE_HOPS ) {
                                sh.coarse[set][lane] = static_cast<uint16_t>( mine );
                            }
                            nextX = baseOfG + ( last & PIPE_TARGET );
                            nextG = g + 1U;
                            nextFlags = PIPE_STORE;
                            nsym += GROUP_SIZE;
                        } else {
                            /* one symbol at a time: plain ones through the single hops, the others through decodeOneSymbol */
                            const uint32_t table = sh.selectors[g];
                            uint32_t q = p0;
                            uint32_t done = 0;
                            uint32_t what = SYM_PLAIN;
                            uint16_t* const out = symOut + nsym;
                            while ( done < GROUP_SIZE ) {
                                if ( ( q >= windowOfG.lo ) && ( q < windowOfG.hi ) ) {
                                    const uint32_t f = sh.ring[set][0][q];
                                    if ( ( f & PIPE_FLAG ) != 0 ) {
                                        if ( lane == 0 ) {
                                            out[done] = static_cast<uint16_t>( table * HUFF_LUT_SIZE | sh.idx[set][q] );
                                        }
                                        q = f & PIPE_TARGET;
                                        ++done;
                                        continue;
                                    }
                                }
                                const uint32_t w = static_cast<uint32_t>( baseOfG >> 5 ) + ( q >> 5 );
                                uint32_t pos = q & 31U;
                                uint32_t entry = 0;
                                what = decodeOneSymbol( sh.words[w & ( PIPE_WORDS - 1U )], sh.words[( w + 1U ) & ( PIPE_WORDS - 1U )], pos,
                                                        decAddr, table, bt, &entry );
                                q = ( q & ~31U ) + pos;
                                if ( what != SYM_PLAIN ) {
                                    break;
                                }
                                if ( lane == 0 ) {
                                    out[done] = static_cast<uint16_t>( entry );
                                }
                                ++done;
                            }
                            nextX = baseOfG + q;
                            nsym += done;
                            nextFlags = what == SYM_PLAIN ? 0U : ( what == SYM_EOB ? 1U : 2U );
                            nextG = what == SYM_PLAIN ? g + 1U : g;
                        }
                    }
                    /* the set of the next group is being looked up from 50 bits behind this group's first bit; a group with
                     * long codes may end behind what that covers */
                    const uint64_t baseOfNext = ( X + PIPE_SKIP ) & ~static_cast<uint64_t>( 31U );
                    const PipeWindow windowOfNext = windowBehind;
                    if ( ( nextX < baseOfNext + windowOfNext.lo ) || ( nextX - baseOfNext > PIPE_MAX_START )
                         || ( nextX - baseOfNext >= windowOfNext.hi ) ) {
                        nextFlags |= PIPE_REBUILD;
                        baseOfG = nextX & ~static_cast<uint64_t>( 31U );
                        windowOfG = windowFirst;
                    } else {
                        baseOfG = baseOfNext;
                        windowOfG = windowOfNext;
                    }
                }
                if ( lane == 0 ) {
                    uint4 raw;
                    raw.x = static_cast<uint32_t>( nextX );
                    raw.y = static_cast<uint32_t>( nextX >> 32 );
                    raw.z = nextG;
                    raw.w = nextFlags;
                    *reinterpret_cast<uint4*>( &sh.state[( round + 1U ) & 1U] ) = raw;
                    if ( ( nextFlags & PIPE_OUTCOME ) != 0 ) {
                        sh.nsym = nsym;
                    }
                }
            }
            __syncthreads();
            ++round;
        }

        if ( t == 0 ) {
            const PipeState last = sh.state[round & 1U];
            if ( ( ( last.flags & PIPE_OUTCOME ) == 1 ) && ( last.X <= nbits ) ) {
                bm.nCoded = sh.nsym;
                bm.endBit = last.X;
            } else {
                bm.status = ST_INTERNAL_CAREFUL;
            }
        }
    }
}
