/*
 * GpuDecoder -- C++ face of the C ABI in include/ibz2_cuda.h for ONE compressed file held in HBM.
 * It stands where the reference has its BlockFinder thread and BlockFetcher thread pool
 * (core/BlockFinder.hpp:36-219, core/BlockFetcher.hpp:244-317, indexed_bzip2/BZ2BlockFetcher.hpp:64-138):
 *   blockOffsets()   <- BlockFinder::get( i ) for all i          (K1 over the whole file, once)
 *   decode( ... )    <- BlockFetcher::get( offset ) for a batch  (K2..K6)
 *   probeHeader()    <- BZ2BlockFetcher::readBlockHeader          (host, a few dozen bits)
 * and turns per-block status codes back into the exceptions the reference throws at the same places.
 * There is no CPU decode path: without a usable B200 the constructor throws.
 */
#pragma once

#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <optional>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/ibz2_cuda.h"

namespace ibz2host
{
/** Mirrors rapidgzip::BitReader::EndOfFileReached (core/BitReader.hpp:82): the CLI prints a dedicated message for it. */
class EndOfFileReached : public std::exception
{
public:
    [[nodiscard]] const char*
    what() const noexcept override
    {
        return "End of file reached";
    }
};

/** Rethrows what bzip2::Block would have thrown for this status (bzip2.hpp:479-907). */
[[noreturn]] inline void
throwForStatus( const ibz2_block_result& r )
{
    std::stringstream msg;
    switch ( r.status ) {
    case IBZ2_ST_BAD_MAGIC:
        /* n_symbols and orig_ptr carry the upper and lower 24 of the 48 bits that were there (bzip2.hpp:502-507) */
        msg << "[BZip2 block header] invalid compressed magic 0x" << std::hex
            << ( ( static_cast<uint64_t>( r.n_symbols ) << 24 ) | r.orig_ptr ) << std::dec
            << " at offset " << r.enc_offset_bits / 8U << " B " << r.enc_offset_bits % 8U << " b";
        throw std::domain_error( msg.str() );
    case IBZ2_ST_RANDOMISED:
        throw std::domain_error( "[BZip2 block header] deprecated isRandomized bit is not supported" );
    case IBZ2_ST_ORIGPTR_RANGE:
        msg << "[BZip2 block header] origPtr " << r.orig_ptr << " is larger than buffer size: " << IBZ2_MAX_BLOCK_SYMBOLS;
        throw std::logic_error( msg.str() );
    /* for the header fields below n_symbols carries the value that was rejected: the reference's messages quote it
     * (bzip2.hpp:581, 595, 627, 659) */
    case IBZ2_ST_GROUP_COUNT:
        msg << "[BZip2 block header] Invalid Huffman coding group count " << r.n_symbols;
        throw std::logic_error( msg.str() );
    case IBZ2_ST_SELECTOR_COUNT:
        msg << "[BZip2 block header] The number of selectors " << r.n_symbols << " is invalid";
        throw std::logic_error( msg.str() );
    case IBZ2_ST_SELECTOR_UNARY:
        msg << "[BZip2 block header] Could not find zero termination after " << r.n_symbols << " bits";
        throw std::domain_error( msg.str() );
    case IBZ2_ST_CODE_LENGTH:
        msg << "[BZip2 block header] start_huffman_length " << r.n_symbols << " is larger than 20 or zero\n";
        throw std::logic_error( msg.str() );
    case IBZ2_ST_CODE_LENGTHS:
        throw std::domain_error( "Constructing a Huffman coding from the given code length sequence failed!" );
    case IBZ2_ST_SELECTOR_OVERRUN:
        /* n_symbols carries the block's selector count for this status: the loop stops at selector == count */
        msg << "[BZip2 block data] selector " << r.n_symbols << " out of maximum range " << r.n_symbols;
        throw std::domain_error( msg.str() );
    case IBZ2_ST_DBUF_OVERFLOW:
        /* n_symbols carries the count that did not fit: behind a run the sum the reference prints (bzip2.hpp:751-756),
         * in front of a literal the count itself, which is the buffer size then (bzip2.hpp:776-780) */
        msg << "[BZip2 block data] dbufCount " << ( r.n_symbols > IBZ2_MAX_BLOCK_SYMBOLS ? "+ hh " : "" ) << r.n_symbols
            << " > " << IBZ2_MAX_BLOCK_SYMBOLS << " dbufSize";
        throw std::domain_error( msg.str() );
    case IBZ2_ST_INVALID_CODE:
        throw std::bad_optional_access();
    case IBZ2_ST_ORIGPTR:
        msg << "[BZip2 block data] origPtr error " << r.orig_ptr;
        throw std::domain_error( msg.str() );
    case IBZ2_ST_CRC:
        msg << "Calculated CRC " << std::hex << r.calculated_crc << " for block mismatches " << r.expected_crc;
        throw std::runtime_error( msg.str() );
    case IBZ2_ST_EOF:
        throw EndOfFileReached();
    default:
        msg << "Block decode failed with status " << r.status;
        throw std::runtime_error( msg.str() );
    }
}

/** IBZ2_DEBUG_TIMING=1: wall-clock time of every call the readers make into the C ABI, on stderr */
class DebugTimer
{
public:
    explicit
    DebugTimer( const char* what, size_t amount = 0 ) :
        m_what( what ),
        m_amount( amount ),
        m_start( std::chrono::steady_clock::now() )
    {}

    ~DebugTimer()
    {
        static const bool enabled = std::getenv( "IBZ2_DEBUG_TIMING" ) != nullptr;
        if ( enabled ) {
            const double ms = std::chrono::duration<double, std::milli>( std::chrono::steady_clock::now() - m_start ).count();
            std::fprintf( stderr, "[ibz2 host] %-22s %10.3f ms  (%zu)\n", m_what, ms, m_amount );
        }
    }

private:
    const char* m_what;
    size_t m_amount;
    std::chrono::steady_clock::time_point m_start;
};

class GpuDecoder
{
public:
    explicit
    GpuDecoder( int device = 0 )
    {
        const DebugTimer timer( "create context" );
        const int rc = ibz2_create( device, &m_ctx );
        if ( rc != IBZ2_OK ) {
            throw std::runtime_error( std::string( "Cannot open the CUDA bzip2 decoder (there is no CPU fallback): " )
                                      + ibz2_last_error( nullptr ) );
        }
        /* the readers fold the block CRCs themselves, in file order across all their batches */
        ibz2_set_verify_stream_crc( m_ctx, 0 );
    }

    ~GpuDecoder()
    {
        if ( m_ctx != nullptr ) {
            if ( m_deviceData != nullptr ) {
                ibz2_device_free( m_ctx, m_deviceData );
            }
            for ( auto& buffer : m_staging ) {
                if ( buffer != nullptr ) {
                    ibz2_host_free_pinned( m_ctx, buffer );
                }
            }
            ibz2_destroy( m_ctx );
        }
    }

    /** Device memory that a batch of decoded blocks lands in; owned by whoever asked for it. */
    struct DeviceBuffer
    {
        void* data{ nullptr };
        size_t capacity{ 0 };
    };

    [[nodiscard]] DeviceBuffer
    allocDevice( size_t capacity )
    {
        const DebugTimer timer( "alloc device bytes", capacity );
        DeviceBuffer buffer;
        check( ibz2_device_alloc( m_ctx, capacity, &buffer.data ), "ibz2_device_alloc" );
        buffer.capacity = capacity;
        return buffer;
    }

    void
    freeDevice( DeviceBuffer& buffer )
    {
        if ( buffer.data != nullptr ) {
            ibz2_device_free( m_ctx, buffer.data );
        }
        buffer = DeviceBuffer{};
    }

    /**
     * Hands @p size decoded bytes that live in HBM at @p deviceData to @p sink, through two pinned staging buffers:
     * one piece travels over PCIe while the sink (a write() to a descriptor, a memcpy into the caller's buffer) works on
     * the piece before it.  Decoded blocks stay in HBM; only what a read asks for crosses the link.
     */
    template<typename Sink>
    void
    deliver( const uint8_t* deviceData, size_t size, const Sink& sink )
    {
        if ( size == 0 ) {
            return;
        }
        if ( m_staging[0] == nullptr ) {
            const DebugTimer timer( "alloc pinned bytes", 2 * STAGING_BYTES );
            for ( auto& buffer : m_staging ) {
                check( ibz2_host_alloc_pinned( m_ctx, STAGING_BYTES, &buffer ), "ibz2_host_alloc_pinned" );
            }
        }
        const size_t pieces = ( size + STAGING_BYTES - 1 ) / STAGING_BYTES;
        const auto pieceSize = [&] ( size_t k ) { return std::min( STAGING_BYTES, size - k * STAGING_BYTES ); };
        check( ibz2_memcpy_d2h_async( m_ctx, m_staging[0], deviceData, pieceSize( 0 ) ), "ibz2_memcpy_d2h_async" );
        for ( size_t k = 0; k < pieces; ++k ) {
            check( ibz2_copy_synchronize( m_ctx ), "ibz2_copy_synchronize" );   /* piece k has arrived */
            if ( k + 1 < pieces ) {
                check( ibz2_memcpy_d2h_async( m_ctx, m_staging[( k + 1 ) & 1U], deviceData + ( k + 1 ) * STAGING_BYTES,
                                              pieceSize( k + 1 ) ), "ibz2_memcpy_d2h_async" );
            }
            sink( static_cast<const uint8_t*>( m_staging[k & 1U] ), pieceSize( k ) );
        }
    }

    GpuDecoder( const GpuDecoder& ) = delete;
    GpuDecoder& operator=( const GpuDecoder& ) = delete;

    /** Copies the compressed file (pageable memory: a mapping or a vector) to the device through pinned staging. */
    void
    upload( const uint8_t* data, size_t size )
    {
        const DebugTimer timer( "upload bytes", size );
        m_size = size;
        check( ibz2_device_alloc( m_ctx, size + 16, &m_deviceData ), "ibz2_device_alloc" );
        if ( size > 0 ) {
            check( ibz2_upload_pageable( m_ctx, m_deviceData, data, size ), "ibz2_upload_pageable" );
        }
    }

    [[nodiscard]] size_t
    size() const
    {
        return m_size;
    }

    /** All data-block magics, ascending (what rapidgzip::BlockFinder collects: block magic only). */
    [[nodiscard]] std::vector<uint64_t>
    findBlocks() const
    {
        const DebugTimer timer( "find blocks in bytes", m_size );
        std::vector<uint64_t> offsets( 1024 );
        while ( true ) {
            uint64_t count = 0;
            const int rc = ibz2_find_magics( m_ctx, m_deviceData, m_size, IBZ2_MEM_DEVICE, 0, IBZ2_FIND_BLOCK,
                                             offsets.data(), offsets.size(), &count );
            if ( rc == IBZ2_E_CAPACITY ) {
                offsets.resize( count );
                continue;
            }
            check( rc, "ibz2_find_magics" );
            offsets.resize( count );
            return offsets;
        }
    }

    /**
     * Decodes the blocks at @p offsets (bit offsets of their magics) into the DEVICE buffer @p buffer, back to back; the
     * buffer is replaced by a larger one when the blocks turn out to need it.
     * @p results gets one entry per block; failing blocks keep their status and do not fail the call.
     */
    void
    decode( const uint64_t* offsets, size_t count, std::vector<ibz2_block_result>& results, DeviceBuffer& buffer )
    {
        const DebugTimer timer( "decode blocks", count );
        results.assign( count, ibz2_block_result{} );
        if ( count == 0 ) {
            return;
        }
        for ( int attempt = 0; attempt < 2; ++attempt ) {
            uint64_t total = 0;
            check( ibz2_decode_blocks( m_ctx, m_deviceData, m_size, IBZ2_MEM_DEVICE, offsets, count, buffer.data,
                                       buffer.capacity, IBZ2_MEM_DEVICE, results.data(), &total ),
                   "ibz2_decode_blocks" );
            const bool overflow = std::any_of( results.begin(), results.end(), [] ( const ibz2_block_result& r ) {
                return r.status == IBZ2_ST_OUTPUT_CAPACITY;
            } );
            if ( !overflow ) {
                return;
            }
            /* RLE1 can expand a block 51-fold: now the sizes are known */
            freeDevice( buffer );
            buffer = allocDevice( total + 65536ULL );
        }
        throw std::runtime_error( "Decoded block sizes changed between two decodes of the same blocks" );
    }

    /** BZ2BlockFetcher::readBlockHeader (BZ2BlockFetcher.hpp:64-82) on the host copy of the file. */
    [[nodiscard]] static ibz2_block_result
    probeHeader( const uint8_t* data, size_t size, uint64_t bitOffset )
    {
        ibz2_block_result result{};
        ibz2_read_block_header( data, size, bitOffset, &result );
        return result;
    }

    [[nodiscard]] ibz2_timings
    timings() const
    {
        ibz2_timings t{};
        ibz2_get_timings( m_ctx, &t );
        return t;
    }

private:
    void
    check( int rc, const char* what ) const
    {
        if ( rc != IBZ2_OK ) {
            throw std::runtime_error( std::string( what ) + " failed: " + ibz2_last_error( m_ctx ) );
        }
    }

    static constexpr size_t STAGING_BYTES = 16ULL << 20;

    ibz2_ctx* m_ctx{ nullptr };
    void* m_deviceData{ nullptr };
    size_t m_size{ 0 };
    void* m_staging[2] = { nullptr, nullptr };
};
}  // namespace ibz2host
