/*
 * K2b for launches with few blocks, second version: one CTA per bzip2 block as in huffmanGroupsKernel, but the lookups
 * for the NEXT group of 50 symbols run while the chain of the CURRENT one is followed.  Included by huffman_bits.cu.
 *
 * What a ring entry says -- "a symbol that starts at this bit ends there" -- depends on the bit position and on the table,
 * not on where the groups start.  The table of every group is known in advance (the selectors), and the next group starts
 * between 50 and 450 bits behind the current one.  So while warp 16 follows the chain of group g through ring set g % 3,
 * the 16 builder warps look up all 1024 bit offsets from ( start of g + 50 ) & ~31 on with the table of group g + 1 and
 * double them to hops of 2, 4, 8, 16 symbols into set ( g + 1 ) % 3, and warps 17-18 store the entries of group g - 1 out
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
constexpr uint32_t PIPE_LEVELS = 5;              /* hops of 1, 2, 4, 8, 16 symbols */
constexpr uint32_t PIPE_SETS = 3;
constexpr uint32_t PIPE_NO_SET = PIPE_SETS;
constexpr uint32_t PIPE_FLAG = 0x8000U;          /* the hop is whole: 2^level plain symbols */
constexpr uint32_t PIPE_TARGET = 0x07FFU;        /* offset behind the hop; >= PIPE_RING: outside the set */
constexpr uint32_t PIPE_WORDS = 256;             /* staged big-endian words of the file (a ring) */
constexpr uint32_t PIPE_REFILL = 64;
constexpr uint32_t PIPE_AHEAD = 128;             /* words staged in front of the set being built, at least */
/* a group may start this far into a set: 48 symbols of 9 bits behind it must have entries */
constexpr uint32_t PIPE_MAX_START = PIPE_RING - ( GROUP_SIZE - 2U ) * HUFF_LUT_BITS - 32U;
constexpr uint32_t PIPE_SKIP = GROUP_SIZE;       /* the next group starts at least this many bits behind the current one */
static_assert( ( GROUP_SIZE * HUFF_LUT_BITS - PIPE_SKIP ) + 31U <= PIPE_MAX_START, "the successor of a plain group is covered" );

struct PipeControl
{
    unsigned long long buildBase;   /* bit position (a multiple of 32) of offset 0 of the set to build */
    unsigned long long endBit;      /* first bit behind the groups chained so far */
    uint32_t outcome;               /* 0 = go on, 1 = end-of-block symbol seen, 2 = careful decoder */
    uint32_t buildSet;              /* PIPE_NO_SET: nothing to build */
    uint32_t buildLut;              /* table * HUFF_LUT_SIZE of the set to build */
    uint32_t outFast;               /* the group chained last is 50 plain symbols whose entries are still to be stored */
    uint32_t outSet;
    uint32_t outLut;
    uint32_t outNsym;               /* symbols in front of that group */
    uint32_t nsym;                  /* symbols of the groups chained so far */
};

struct PipeShared
{
    alignas( 16 ) uint8_t dec[( DEC_STRIDE + 15U ) & ~15U];
    alignas( 16 ) PipeControl ctrl[2];
    uint32_t block;
    uint32_t skip;
    uint32_t words[PIPE_WORDS];
    uint16_t ring[PIPE_SETS][PIPE_LEVELS][PIPE_RING];
    uint16_t idx[PIPE_SETS][PIPE_RING];
    uint16_t coarse[PIPE_SETS][4];
};
static_assert( sizeof( PipeShared ) <= 48U * 1024U, "static shared memory" );

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
    __shared__ PipeShared sh;
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
        const uint64_t startBit = bt->dataStartBit;
        for ( uint32_t i = t; i < PIPE_WORDS; i += PIPE_THREADS ) {
            const uint64_t w = ( startBit >> 5 ) + i;
            sh.words[w & ( PIPE_WORDS - 1U )] = bswap( loadWordRaw( data, nbytes, w ) );
        }
        if ( t == 0 ) {
            /* the first round only looks up: group 0 from the word its first bit is in */
            PipeControl c{};
            c.buildBase = startBit & ~static_cast<uint64_t>( 31U );
            c.endBit = startBit;
            c.buildSet = 0;
            c.buildLut = ( selectorsCount > 0 ? selectors[0] : 0U ) * HUFF_LUT_SIZE;
            sh.ctrl[0] = c;
        }
        /* ---- what the roles keep between the rounds ---- */
        /* builders: file words in front of loadedUpTo are staged; a refill is loaded in one round and stored in the next */
        uint64_t loadedUpTo = ( startBit >> 5 ) + PIPE_WORDS;
        bool pending = false;
        uint32_t pendingRaw = 0;
        /* chain warp: the next group to chain, where it starts, and which group each set holds, looked up from where */
        uint64_t X = startBit;
        uint32_t g = 0;
        uint32_t nsym = 0;
        constexpr uint32_t NO_GROUP = 0xFFFFFFFFU;
        uint32_t setGroup0 = 0, setGroup1 = NO_GROUP, setGroup2 = NO_GROUP;
        uint64_t setBase0 = startBit & ~static_cast<uint64_t>( 31U ), setBase1 = 0, setBase2 = 0;
        uint32_t round = 0;
        __syncthreads();

        while ( true ) {
            const PipeControl ctrl = sh.ctrl[round & 1U];
            if ( ctrl.outcome != 0 ) {
                break;
            }
            if ( warp < PIPE_CHAIN_WARP ) {
                /* ================= builders ================= */
                if ( pending ) {
                    if ( t < PIPE_REFILL ) {
                        /* these slots hold words more than 190 in front of the newest: nobody reads them any more */
                        sh.words[( loadedUpTo + t ) & ( PIPE_WORDS - 1U )] = bswap( pendingRaw );
                    }
                    loadedUpTo += PIPE_REFILL;
                    pending = false;
                }
                if ( ctrl.buildSet != PIPE_NO_SET ) {
                    const uint64_t baseWord = ctrl.buildBase >> 5;
                    if ( loadedUpTo < baseWord + PIPE_AHEAD ) {
                        if ( t < PIPE_REFILL ) {
                            pendingRaw = loadWordRaw( data, nbytes, loadedUpTo + t );
                        }
                        pending = true;
                    }
                    const uint32_t set = ctrl.buildSet;
                    const uint32_t slot = static_cast<uint32_t>( baseWord );
                    uint32_t e[2];
#pragma unroll
                    for ( uint32_t j = 0; j < 2U; ++j ) {
                        const uint32_t r = t + j * PIPE_BUILD_THREADS;
                        const uint32_t w = slot + ( r >> 5 );
                        const uint32_t wa = sh.words[w & ( PIPE_WORDS - 1U )];
                        const uint32_t wb = sh.words[( w + 1U ) & ( PIPE_WORDS - 1U )];
                        const uint32_t index = __funnelshift_l( wb, wa, r & 31U ) >> ( 32 - HUFF_LUT_BITS );
                        const uint32_t len = sh.dec[ctrl.buildLut + index];
                        e[j] = len != 0 ? ( ( r + len ) | PIPE_FLAG ) : r;
                        sh.ring[set][0][r] = static_cast<uint16_t>( e[j] );
                        sh.idx[set][r] = static_cast<uint16_t>( index );
                    }
#pragma unroll
                    for ( uint32_t level = 1; level < PIPE_LEVELS; ++level ) {
                        pipeBuilderBarrier();
#pragma unroll
                        for ( uint32_t j = 0; j < 2U; ++j ) {
                            const uint32_t target = e[j] & PIPE_TARGET;
                            const uint32_t f = sh.ring[set][level - 1U][target & ( PIPE_RING - 1U )];
                            const bool whole = ( ( e[j] & PIPE_FLAG ) != 0 ) && ( target < PIPE_RING );
                            e[j] = whole ? f : ( e[j] & ~PIPE_FLAG );
                            sh.ring[set][level][t + j * PIPE_BUILD_THREADS] = static_cast<uint16_t>( e[j] );
                        }
                    }
                }
            } else if ( warp > PIPE_CHAIN_WARP ) {
                /* ================= output: the entries of the group chained in the round before ================= */
                const uint32_t s = lane + 32U * ( warp - PIPE_OUT_WARP );
                if ( ( ctrl.outFast != 0 ) && ( s < GROUP_SIZE ) ) {
                    const uint32_t set = ctrl.outSet;
                    uint32_t q = sh.coarse[set][s >> 4];
                    if ( ( s & 8U ) != 0 ) {
                        q = sh.ring[set][3][q] & PIPE_TARGET;
                    }
                    if ( ( s & 4U ) != 0 ) {
                        q = sh.ring[set][2][q] & PIPE_TARGET;
                    }
                    if ( ( s & 2U ) != 0 ) {
                        q = sh.ring[set][1][q] & PIPE_TARGET;
                    }
                    if ( ( s & 1U ) != 0 ) {
                        q = sh.ring[set][0][q] & PIPE_TARGET;
                    }
                    symOut[ctrl.outNsym + s] = static_cast<uint16_t>( ctrl.outLut | sh.idx[set][q] );
                }
            } else {
                /* ================= the chain ================= */
                const uint32_t set = g % PIPE_SETS;
                const uint32_t heldGroup = set == 0 ? setGroup0 : ( set == 1 ? setGroup1 : setGroup2 );
                const uint64_t base = set == 0 ? setBase0 : ( set == 1 ? setBase1 : setBase2 );
                /* the set of group g is usable if it has been looked up (not in this very round) from a place that covers it */
                const bool ready = ( heldGroup == g ) && ( set != ctrl.buildSet ) && ( X >= base ) && ( X - base <= PIPE_MAX_START );
                const uint32_t tableNow = g < selectorsCount ? selectors[g] : 0U;
                const uint32_t tableNext = g + 1U < selectorsCount ? selectors[g + 1U] : 0U;
                const uint32_t tableAfter = g + 2U < selectorsCount ? selectors[g + 2U] : 0U;
                PipeControl next{};
                next.outcome = 0;
                next.outFast = 0;
                if ( ready ) {
                    /* next group of 50 symbols: the exits the lane kernel has at this place (bzip2.hpp:710-721) */
                    if ( ( g >= selectorsCount ) || ( nsym >= SYM_LIMIT ) || ( X > nbits ) ) {
                        next.outcome = 2;
                    } else {
                        const uint32_t p0 = static_cast<uint32_t>( X - base );
                        const uint16_t* const hop16 = sh.ring[set][4];
                        const uint32_t f1 = hop16[p0];
                        const uint32_t t1 = f1 & PIPE_TARGET;
                        const uint32_t f2 = hop16[t1 & ( PIPE_RING - 1U )];
                        const uint32_t t2 = f2 & PIPE_TARGET;
                        const uint32_t f3 = hop16[t2 & ( PIPE_RING - 1U )];
                        const uint32_t t3 = f3 & PIPE_TARGET;
                        const uint32_t f4 = sh.ring[set][1][t3 & ( PIPE_RING - 1U )];
                        const bool whole = ( ( f1 & f2 & f3 & f4 & PIPE_FLAG ) != 0 ) && ( ( t1 | t2 | t3 ) < PIPE_RING );
                        if ( whole ) {
                            if ( lane < 4U ) {
                                sh.coarse[set][lane] = static_cast<uint16_t>( lane == 0 ? p0 : ( lane == 1 ? t1 : ( lane == 2 ? t2 : t3 ) ) );
                            }
                            next.outFast = 1;
                            next.outSet = set;
                            next.outLut = tableNow * HUFF_LUT_SIZE;
                            next.outNsym = nsym;
                            X = base + ( f4 & PIPE_TARGET );
                            nsym += GROUP_SIZE;
                            ++g;
                        } else {
                            /* one symbol at a time: plain ones through the single hops, the others through decodeOneSymbol */
                            uint32_t p = p0;
                            uint32_t done = 0;
                            uint32_t what = SYM_PLAIN;
                            uint16_t* const out = symOut + nsym;
                            while ( done < GROUP_SIZE ) {
                                if ( p < PIPE_RING ) {
                                    const uint32_t f = sh.ring[set][0][p];
                                    if ( ( f & PIPE_FLAG ) != 0 ) {
                                        if ( lane == 0 ) {
                                            out[done] = static_cast<uint16_t>( tableNow * HUFF_LUT_SIZE | sh.idx[set][p] );
                                        }
                                        p = f & PIPE_TARGET;
                                        ++done;
                                        continue;
                                    }
                                }
                                const uint32_t w = static_cast<uint32_t>( base >> 5 ) + ( p >> 5 );
                                uint32_t pos = p & 31U;
                                uint32_t entry = 0;
                                what = decodeOneSymbol( sh.words[w & ( PIPE_WORDS - 1U )], sh.words[( w + 1U ) & ( PIPE_WORDS - 1U )], pos,
                                                        decAddr, tableNow, bt, &entry );
                                p = ( p & ~31U ) + pos;
                                if ( what != SYM_PLAIN ) {
                                    break;
                                }
                                if ( lane == 0 ) {
                                    out[done] = static_cast<uint16_t>( entry );
                                }
                                ++done;
                            }
                            X = base + p;
                            nsym += done;
                            next.outcome = what == SYM_PLAIN ? 0U : ( what == SYM_EOB ? 1U : 2U );
                            if ( what == SYM_PLAIN ) {
                                ++g;
                            }
                        }
                    }
                }
                /* ---- what the others do in the next round ---- */
                const bool advanced = ready && ( next.outcome == 0 );
                const uint32_t tableComing = advanced ? tableNext : tableNow;     /* of group g as it is now */
                const uint32_t tableBehind = advanced ? tableAfter : tableNext;   /* of group g + 1 */
                const uint32_t comingSet = g % PIPE_SETS;
                const uint32_t comingHeld = comingSet == 0 ? setGroup0 : ( comingSet == 1 ? setGroup1 : setGroup2 );
                const uint64_t comingBase = comingSet == 0 ? setBase0 : ( comingSet == 1 ? setBase1 : setBase2 );
                /* (the set being looked up in this round is complete when the next one starts) */
                const bool comingReady = ( comingHeld == g ) && ( X >= comingBase ) && ( X - comingBase <= PIPE_MAX_START );
                uint32_t buildSet = PIPE_NO_SET;
                uint32_t buildGroup = 0;
                uint64_t buildBase = 0;
                uint32_t buildTable = 0;
                if ( !comingReady ) {
                    buildSet = comingSet;
                    buildGroup = g;
                    buildBase = X & ~static_cast<uint64_t>( 31U );
                    buildTable = tableComing;
                } else {
                    const uint32_t behindSet = ( g + 1U ) % PIPE_SETS;
                    const uint32_t behindHeld = behindSet == 0 ? setGroup0 : ( behindSet == 1 ? setGroup1 : setGroup2 );
                    if ( behindHeld != g + 1U ) {
                        buildSet = behindSet;
                        buildGroup = g + 1U;
                        buildBase = ( X + PIPE_SKIP ) & ~static_cast<uint64_t>( 31U );
                        buildTable = tableBehind;
                    }
                }
                if ( buildSet != PIPE_NO_SET ) {
                    setGroup0 = buildSet == 0 ? buildGroup : setGroup0;
                    setGroup1 = buildSet == 1 ? buildGroup : setGroup1;
                    setGroup2 = buildSet == 2 ? buildGroup : setGroup2;
                    setBase0 = buildSet == 0 ? buildBase : setBase0;
                    setBase1 = buildSet == 1 ? buildBase : setBase1;
                    setBase2 = buildSet == 2 ? buildBase : setBase2;
                }
                next.buildSet = buildSet;
                next.buildBase = buildBase;
                next.buildLut = buildTable * HUFF_LUT_SIZE;
                next.endBit = X;
                next.nsym = nsym;
                if ( lane == 0 ) {
                    sh.ctrl[( round + 1U ) & 1U] = next;
                }
            }
            __syncthreads();
            ++round;
        }

        if ( t == 0 ) {
            const PipeControl last = sh.ctrl[round & 1U];
            if ( ( last.outcome == 1 ) && ( last.endBit <= nbits ) ) {
                bm.nCoded = last.nsym;
                bm.endBit = last.endBit;
            } else {
                bm.status = ST_INTERNAL_CAREFUL;
            }
        }
    }
}
