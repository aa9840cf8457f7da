/*
 * multi.cu -- one bzip2 file on several GPUs of one machine, in ONE process: the multi-device entry points of
 * include/ibz2_cuda.h.  Host code only; it drives one ibz2_ctx per device through the single-device C ABI from one
 * thread per device.  No NCCL, no peer copies: bzip2 blocks are independent, so the only things that cross devices
 * are a handful of byte counts (the exclusive scan that places every device's bytes in the caller's buffer) and the
 * block CRCs (folded into the stream CRCs on the host).
 *
 * Stands in for the way the reference scales inside one reader object: a pool of decode threads over the blocks of
 * one file (indexed_bzip2/ParallelBZ2Reader.hpp:50-69, core/BlockFetcher.hpp:176-190), with devices in place of
 * threads.  The compressed stream is cut into equal byte ranges (= equal compressed bits up to one block); a device owns
 * the blocks whose magic starts in its range, like the chunks of core/ParallelBitStringFinder.hpp:232-241.
 */
#include "../../include/ibz2_cuda.h"

#include <algorithm>
#include <chrono>
#include <cstring>
#include <memory>
#include <new>
#include <string>
#include <thread>
#include <vector>

namespace
{
/* a block never takes more than 900 000 symbols x 20 bits plus its header: what a device reads past its range */
constexpr uint64_t WINDOW_SLACK_BYTES = 4ULL << 20;

struct Share
{
    int device = 0;
    ibz2_ctx* ctx = nullptr;
    void* dOut = nullptr;
    uint64_t dOutCapacity = 0;
    std::vector<ibz2_block_result> results;
    uint64_t nblocks = 0;
    uint64_t total = 0;
    uint64_t windowStart = 0;
    int rc = IBZ2_OK;
    std::string error;
    ibz2_timings timings{};
};

double
millisecondsSince( std::chrono::steady_clock::time_point t0 )
{
    return std::chrono::duration<double, std::milli>( std::chrono::steady_clock::now() - t0 ).count();
}
}  // namespace

struct ibz2_multi
{
    std::vector<Share> shares;
    std::string lastError;
    ibz2_multi_timings timings{};
};

extern "C" {

int
ibz2_multi_create( const int* devices, int ndevices, ibz2_multi** out )
{
    if ( ( out == nullptr ) || ( devices == nullptr ) || ( ndevices <= 0 ) ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    *out = nullptr;
    auto multi = std::unique_ptr<ibz2_multi>( new ( std::nothrow ) ibz2_multi );
    if ( !multi ) {
        return IBZ2_E_OUT_OF_MEMORY;
    }
    multi->shares.resize( static_cast<size_t>( ndevices ) );
    for ( int i = 0; i < ndevices; ++i ) {
        Share& share = multi->shares[static_cast<size_t>( i )];
        share.device = devices[i];
        const int rc = ibz2_create( devices[i], &share.ctx );
        if ( rc != IBZ2_OK ) {
            for ( auto& s : multi->shares ) {
                ibz2_destroy( s.ctx );
            }
            return rc;   /* ibz2_last_error( NULL ) has the reason */
        }
        /* streams are folded over the shares laid end to end, not per device */
        ibz2_set_verify_stream_crc( share.ctx, 0 );
    }
    *out = multi.release();
    return IBZ2_OK;
}

void
ibz2_multi_destroy( ibz2_multi* multi )
{
    if ( multi == nullptr ) {
        return;
    }
    for ( auto& share : multi->shares ) {
        if ( share.dOut != nullptr ) {
            ibz2_device_free( share.ctx, share.dOut );
        }
        ibz2_destroy( share.ctx );
    }
    delete multi;
}

const char*
ibz2_multi_last_error( const ibz2_multi* multi )
{
    return multi != nullptr ? multi->lastError.c_str() : ibz2_last_error( nullptr );
}

int
ibz2_multi_device_count( const ibz2_multi* multi )
{
    return multi != nullptr ? static_cast<int>( multi->shares.size() ) : 0;
}

int
ibz2_multi_host_alloc_pinned( ibz2_multi* multi, uint64_t nbytes, void** hptr )
{
    if ( ( multi == nullptr ) || multi->shares.empty() ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    return ibz2_host_alloc_pinned( multi->shares.front().ctx, nbytes, hptr );   /* portable: every device may copy into it */
}

int
ibz2_multi_host_free_pinned( ibz2_multi* multi, void* hptr )
{
    if ( ( multi == nullptr ) || multi->shares.empty() ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    return ibz2_host_free_pinned( multi->shares.front().ctx, hptr );
}

int
ibz2_multi_get_timings( const ibz2_multi* multi, ibz2_multi_timings* timings, ibz2_timings* per_device, int capacity )
{
    if ( ( multi == nullptr ) || ( timings == nullptr ) ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    *timings = multi->timings;
    for ( int i = 0; ( per_device != nullptr ) && ( i < capacity ) && ( i < static_cast<int>( multi->shares.size() ) ); ++i ) {
        per_device[i] = multi->shares[static_cast<size_t>( i )].timings;
    }
    return IBZ2_OK;
}

int
ibz2_multi_decode_buffer( ibz2_multi* multi, const void* data, uint64_t nbytes, void* out, uint64_t out_capacity,
                          ibz2_block_result* results, uint64_t results_capacity, uint64_t* nblocks, uint64_t* total_out )
{
    if ( ( multi == nullptr ) || ( nblocks == nullptr ) || ( ( data == nullptr ) && ( nbytes > 0 ) )
         || ( ( results == nullptr ) && ( results_capacity > 0 ) ) ) {
        if ( multi != nullptr ) {
            multi->lastError = "ibz2_multi_decode_buffer: invalid argument";
        }
        return IBZ2_E_INVALID_ARGUMENT;
    }
    *nblocks = 0;
    if ( total_out != nullptr ) {
        *total_out = 0;
    }
    multi->timings = ibz2_multi_timings{};
    const auto t0 = std::chrono::steady_clock::now();
    const size_t n = multi->shares.size();
    const auto* const bytes = static_cast<const uint8_t*>( data );

    /* ---- phase 1: every device finds and decodes the blocks of its byte range into its own memory ---- */
    const auto decodeShare = [&] ( size_t i, bool wholeTail ) {
        Share& share = multi->shares[i];
        const uint64_t rangeStart = nbytes / n * i;
        const uint64_t rangeEnd = ( i + 1 == n ) ? nbytes : nbytes / n * ( i + 1 );
        const uint64_t windowEnd = wholeTail ? nbytes : std::min( nbytes, rangeEnd + WINDOW_SLACK_BYTES );
        share.windowStart = rangeStart;
        const uint64_t firstBit = 0;                                                          /* relative to the window */
        const uint64_t lastBit = ( i + 1 == n ) ? UINT64_MAX : ( rangeEnd - rangeStart ) * 8U;
        /* room for this share's bytes: its part of the caller's buffer and a margin; grown when the sizes say so */
        uint64_t wanted = std::max<uint64_t>( share.dOutCapacity, out_capacity / n + out_capacity / ( 4U * n ) + ( 64ULL << 20 ) );
        share.results.resize( std::max<size_t>( share.results.size(), static_cast<size_t>( ( windowEnd - rangeStart ) / 4096U + 1024U ) ) );
        for ( int attempt = 0; attempt < 4; ++attempt ) {
            if ( ( share.dOut == nullptr ) || ( share.dOutCapacity < wanted ) ) {
                if ( share.dOut != nullptr ) {
                    ibz2_device_free( share.ctx, share.dOut );
                    share.dOut = nullptr;
                    share.dOutCapacity = 0;
                }
                share.rc = ibz2_device_alloc( share.ctx, wanted, &share.dOut );
                if ( share.rc != IBZ2_OK ) {
                    share.error = ibz2_last_error( share.ctx );
                    return;
                }
                share.dOutCapacity = wanted;
            }
            share.rc = ibz2_decode_buffer( share.ctx, bytes + rangeStart, windowEnd - rangeStart, IBZ2_MEM_HOST, firstBit, lastBit,
                                           ( out != nullptr ) ? share.dOut : nullptr, share.dOutCapacity, IBZ2_MEM_DEVICE,
                                           share.results.data(), share.results.size(), &share.nblocks, &share.total );
            if ( share.rc == IBZ2_E_CAPACITY ) {
                share.results.resize( static_cast<size_t>( share.nblocks ) + 16U );
                continue;
            }
            if ( share.rc != IBZ2_OK ) {
                share.error = ibz2_last_error( share.ctx );
                return;
            }
            if ( ( out != nullptr ) && ( share.total > share.dOutCapacity ) ) {
                wanted = share.total + ( 1ULL << 20 );   /* e.g. long runs: RLE1 expands a block up to 51-fold */
                continue;
            }
            break;
        }
        ibz2_get_timings( share.ctx, &share.timings );
    };
    {
        std::vector<std::thread> threads;
        for ( size_t i = 0; i < n; ++i ) {
            threads.emplace_back( decodeShare, i, false );
        }
        for ( auto& t : threads ) {
            t.join();
        }
    }
    /* a last block that the window cut short: once more with everything behind the range */
    for ( size_t i = 0; i + 1 < n; ++i ) {
        Share& share = multi->shares[i];
        if ( ( share.rc == IBZ2_OK ) && ( share.nblocks > 0 )
             && ( share.results[static_cast<size_t>( share.nblocks ) - 1U].status != IBZ2_ST_OK ) ) {
            decodeShare( i, true );
        }
    }
    for ( const auto& share : multi->shares ) {
        if ( share.rc != IBZ2_OK ) {
            multi->lastError = "device " + std::to_string( share.device ) + ": " + share.error;
            return share.rc;
        }
    }
    multi->timings.decode_ms = static_cast<float>( millisecondsSince( t0 ) );

    /* ---- the exclusive scan over the devices' byte counts ---- */
    uint64_t blocks = 0;
    uint64_t total = 0;
    std::vector<uint64_t> base( n, 0 );
    for ( size_t i = 0; i < n; ++i ) {
        base[i] = total;
        total += multi->shares[i].total;
        blocks += multi->shares[i].nblocks;
    }
    *nblocks = blocks;
    if ( total_out != nullptr ) {
        *total_out = total;
    }
    if ( blocks > results_capacity ) {
        multi->lastError = "ibz2_multi_decode_buffer: results array too small";
        return IBZ2_E_CAPACITY;
    }

    /* ---- phase 2: every device copies its bytes to where they belong in the caller's buffer ---- */
    const auto t1 = std::chrono::steady_clock::now();
    if ( out != nullptr ) {
        std::vector<std::thread> threads;
        for ( size_t i = 0; i < n; ++i ) {
            threads.emplace_back( [&, i] () {
                Share& share = multi->shares[i];
                if ( base[i] >= out_capacity ) {
                    return;
                }
                const uint64_t count = std::min<uint64_t>( share.total, out_capacity - base[i] );
                if ( count > 0 ) {
                    share.rc = ibz2_memcpy_d2h( share.ctx, static_cast<uint8_t*>( out ) + base[i], share.dOut, count );
                    if ( share.rc != IBZ2_OK ) {
                        share.error = ibz2_last_error( share.ctx );
                    }
                }
            } );
        }
        for ( auto& t : threads ) {
            t.join();
        }
        for ( const auto& share : multi->shares ) {
            if ( share.rc != IBZ2_OK ) {
                multi->lastError = "device " + std::to_string( share.device ) + ": " + share.error;
                return share.rc;
            }
        }
    }
    multi->timings.copy_ms = static_cast<float>( millisecondsSince( t1 ) );

    /* ---- one result array in file order ---- */
    uint64_t k = 0;
    for ( size_t i = 0; i < n; ++i ) {
        const Share& share = multi->shares[i];
        for ( uint64_t b = 0; b < share.nblocks; ++b, ++k ) {
            ibz2_block_result r = share.results[static_cast<size_t>( b )];
            r.enc_offset_bits += share.windowStart * 8U;
            r.out_offset += base[i];
            if ( ( out != nullptr ) && ( r.status == IBZ2_ST_OK ) && ( r.out_offset + r.out_size > out_capacity ) ) {
                r.status = IBZ2_ST_OUTPUT_CAPACITY;
            }
            results[k] = r;
        }
    }
    multi->timings.stream_crc_mismatches = ibz2_fold_stream_crcs( results, blocks, 1 );
    multi->timings.total_ms = static_cast<float>( millisecondsSince( t0 ) );
    return IBZ2_OK;
}

}  // extern "C"
