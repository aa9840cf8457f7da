/*
 * TEST INFRASTRUCTURE ONLY -- thin C-ABI shim around the UNMODIFIED reference headers.
 *
 * This file is ours; it merely #includes the reference's header-only classes from where they lie
 * (/root/reference/src, passed with -I by oracle/Makefile) and exposes them with plain-C signatures so
 * that tests, golden-vector generation and bench.py's cpu_baseline / --impl reference leg can call the
 * real reference in-process.  Output goes to oracle/_ref/ only (git-ignored, travels with gpurun).
 *
 * Reference entry points used:
 *   indexed_bzip2::ParallelBZ2Reader   src/indexed_bzip2/ParallelBZ2Reader.hpp:39-498
 *   indexed_bzip2::BZ2Reader           src/indexed_bzip2/BZ2Reader.hpp:29-491
 *   rapidgzip::BitStringFinder<48>     src/core/BitStringFinder.hpp:158-320
 *   rapidgzip::ParallelBitStringFinder src/core/ParallelBitStringFinder.hpp:159-265
 */
#include <chrono>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include <BitStringFinder.hpp>
#include <ParallelBitStringFinder.hpp>
#include <BZ2Reader.hpp>
#include <ParallelBZ2Reader.hpp>
#include <filereader/Standard.hpp>
#include <filereader/BufferView.hpp>

namespace
{
thread_local std::string g_lastError;

std::unique_ptr<indexed_bzip2::BZ2ReaderInterface>
makeReader( const char* path, int parallelization, int serialClass )
{
    auto file = std::make_unique<rapidgzip::StandardFileReader>( std::string( path ) );
    if ( serialClass != 0 ) {
        return std::make_unique<indexed_bzip2::BZ2Reader>( std::move( file ) );
    }
    return std::make_unique<indexed_bzip2::ParallelBZ2Reader>( std::move( file ),
                                                                static_cast<size_t>( parallelization ) );
}
}  // namespace

extern "C" {

const char*
ref_last_error()
{
    return g_lastError.c_str();
}

/**
 * Decode a whole file. @p out may be NULL (decode without sink, like read(-1, nullptr, SIZE_MAX)).
 * offsets receives (bits, bytes) pairs of blockOffsets(). Returns 0 on success, 1 on exception.
 */
int
ref_decode_file( const char* path, int parallelization, int serialClass,
                 uint8_t* out, uint64_t outCapacity, uint64_t* outSize,
                 uint64_t* offsets, uint64_t offsetsCapacityPairs, uint64_t* offsetsCount,
                 double* seconds )
{
    try {
        auto reader = makeReader( path, parallelization, serialClass );
        const auto t0 = std::chrono::steady_clock::now();
        uint64_t total = 0;
        if ( out == nullptr ) {
            total = reader->read( -1, nullptr, std::numeric_limits<size_t>::max() );
        } else {
            total = reader->read( -1, reinterpret_cast<char*>( out ), outCapacity );
            /* probe for overflow: anything left means the caller's buffer was too small */
            char probe;
            if ( ( total == outCapacity ) && ( reader->read( -1, &probe, 1 ) != 0 ) ) {
                g_lastError = "output buffer too small";
                return 2;
            }
        }
        const auto t1 = std::chrono::steady_clock::now();
        if ( seconds != nullptr ) {
            *seconds = std::chrono::duration<double>( t1 - t0 ).count();
        }
        if ( outSize != nullptr ) {
            *outSize = total;
        }
        if ( offsetsCount != nullptr ) {
            const auto map = reader->blockOffsets();
            uint64_t i = 0;
            for ( const auto& [bits, bytes] : map ) {
                if ( ( offsets != nullptr ) && ( i < offsetsCapacityPairs ) ) {
                    offsets[2 * i] = bits;
                    offsets[2 * i + 1] = bytes;
                }
                ++i;
            }
            *offsetsCount = i;
        }
        return 0;
    } catch ( const std::exception& e ) {
        g_lastError = e.what();
        return 1;
    }
}

/**
 * importIndex-style random access: setBlockOffsets(offsets) on a fresh ParallelBZ2Reader, then
 * nReads x { seek(positions[i]); read(readSize) } into out (readSize bytes, overwritten each time).
 * checksum receives the sum of all bytes read (cheap cross-check). Returns 0 / 1.
 */
int
ref_random_reads( const char* path, int parallelization,
                  const uint64_t* offsets, uint64_t offsetsCount,
                  const uint64_t* positions, uint64_t nReads, uint64_t readSize,
                  uint8_t* out, uint64_t* checksum, double* seconds )
{
    try {
        indexed_bzip2::ParallelBZ2Reader reader( std::make_unique<rapidgzip::StandardFileReader>( std::string( path ) ),
                                                 static_cast<size_t>( parallelization ) );
        if ( offsetsCount > 0 ) {
            std::map<size_t, size_t> map;
            for ( uint64_t i = 0; i < offsetsCount; ++i ) {
                map.emplace( offsets[2 * i], offsets[2 * i + 1] );
            }
            reader.setBlockOffsets( std::move( map ) );
        }
        std::vector<char> buffer( readSize );
        uint64_t sum = 0;
        const auto t0 = std::chrono::steady_clock::now();
        for ( uint64_t i = 0; i < nReads; ++i ) {
            reader.seek( static_cast<long long>( positions[i] ) );
            const auto n = reader.read( -1, buffer.data(), readSize );
            for ( size_t j = 0; j < n; ++j ) {
                sum += static_cast<uint8_t>( buffer[j] );
            }
        }
        const auto t1 = std::chrono::steady_clock::now();
        if ( out != nullptr ) {
            std::memcpy( out, buffer.data(), readSize );
        }
        if ( checksum != nullptr ) {
            *checksum = sum;
        }
        if ( seconds != nullptr ) {
            *seconds = std::chrono::duration<double>( t1 - t0 ).count();
        }
        return 0;
    } catch ( const std::exception& e ) {
        g_lastError = e.what();
        return 1;
    }
}

/**
 * Raw magic finder over a memory buffer: all bit offsets of the 48-bit @p magic, ascending,
 * as `ibzip2 -l` would list them for one magic (src/tools/ibzip2.cpp:96-143).
 */
int
ref_find_magics( const uint8_t* data, uint64_t size, uint64_t magic, int parallelization, uint64_t fileBufferBytes,
                 uint64_t* out, uint64_t capacity, uint64_t* count, double* seconds )
{
    try {
        using Finder = rapidgzip::BitStringFinder<48>;
        using ParallelFinder = rapidgzip::ParallelBitStringFinder<48>;
        auto file = std::make_unique<rapidgzip::BufferViewFileReader>( data, size );
        std::unique_ptr<Finder> finder;
        if ( parallelization <= 1 ) {
            finder = std::make_unique<Finder>( std::move( file ), magic, fileBufferBytes );
        } else {
            finder = std::make_unique<ParallelFinder>( std::move( file ), magic, parallelization, 0, fileBufferBytes );
        }
        const auto t0 = std::chrono::steady_clock::now();
        uint64_t n = 0;
        for ( auto offset = finder->find(); offset != std::numeric_limits<size_t>::max(); offset = finder->find() ) {
            if ( n < capacity ) {
                out[n] = offset;
            }
            ++n;
        }
        const auto t1 = std::chrono::steady_clock::now();
        *count = n;
        if ( seconds != nullptr ) {
            *seconds = std::chrono::duration<double>( t1 - t0 ).count();
        }
        return 0;
    } catch ( const std::exception& e ) {
        g_lastError = e.what();
        return 1;
    }
}

}  // extern "C"
