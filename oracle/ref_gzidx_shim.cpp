/*
 * TEST INFRASTRUCTURE ONLY.  The reference's gzip-architecture reader (rapidgzip::ParallelGzipReader, which also
 * decodes bzip2 through rapidgzip/chunkdecoding/Bzip2Chunk.hpp) behind a C interface, compiled UNMODIFIED from
 * /root/reference/src by oracle/Makefile (needs zlib).  It exists to check that the GZIDX files our readers write
 * (indexed_bzip2_b200/host/IndexFileFormat.hpp) are read by the reference: ParallelGzipReader::importIndex
 * (rapidgzip/ParallelGzipReader.hpp:993-1029) followed by reads at given positions.
 */
#include <chrono>
#include <cstring>
#include <fstream>
#include <limits>
#include <memory>
#include <string>

#include <filereader/Standard.hpp>
#include <rapidgzip/ParallelGzipReader.hpp>

namespace
{
thread_local std::string g_lastError;
}

extern "C" {

const char*
refgz_last_error()
{
    return g_lastError.c_str();
}

/* Opens `path` (a bzip2 file), imports the index file `indexPath` if given, reads `readSize` bytes at every position and
 * returns the byte sum of everything read plus the first read's bytes in `out`. */
int
refgz_reads_with_index( const char* path, const char* indexPath, int parallelization, const uint64_t* positions,
                        uint64_t nReads, uint64_t readSize, uint8_t* out, uint64_t* checksum, uint64_t* decodedSize )
{
    try {
        using namespace rapidgzip;
        ParallelGzipReader<ChunkData> reader( std::make_unique<StandardFileReader>( std::string( path ) ),
                                              static_cast<size_t>( parallelization ) );
        if ( ( indexPath != nullptr ) && ( indexPath[0] != '\0' ) ) {
            reader.importIndex( std::make_unique<StandardFileReader>( std::string( indexPath ) ) );
        }
        std::vector<char> buffer( readSize );
        uint64_t sum = 0;
        for ( uint64_t i = 0; i < nReads; ++i ) {
            reader.seek( static_cast<long long>( positions[i] ) );
            const auto n = reader.read( -1, buffer.data(), readSize );
            for ( size_t j = 0; j < n; ++j ) {
                sum += static_cast<uint8_t>( buffer[j] );
            }
            if ( ( i == 0 ) && ( out != nullptr ) ) {
                std::memcpy( out, buffer.data(), n );
            }
        }
        if ( checksum != nullptr ) {
            *checksum = sum;
        }
        if ( decodedSize != nullptr ) {
            reader.seek( 0, SEEK_END );
            *decodedSize = reader.tell();
        }
        return 0;
    } catch ( const std::exception& e ) {
        g_lastError = e.what();
        return 1;
    }
}

/* The chunk map of the reference's gzip-architecture reader for a bzip2 file read with chunks of `chunkSizeInBytes`
 * compressed bytes: ( encoded bit offset, decoded byte offset ) pairs incl. the end of the file.  Every entry is where
 * Bzip2Chunk::decodeChunk (rapidgzip/chunkdecoding/Bzip2Chunk.hpp:200-268) started a chunk. */
int
refgz_chunk_offsets( const char* path, int parallelization, uint64_t chunkSizeInBytes, uint64_t* pairs, uint64_t capacity,
                     uint64_t* count )
{
    try {
        using namespace rapidgzip;
        ParallelGzipReader<ChunkData> reader( std::make_unique<StandardFileReader>( std::string( path ) ),
                                              static_cast<size_t>( parallelization ), chunkSizeInBytes );
        reader.read( -1, nullptr, std::numeric_limits<size_t>::max() );
        const auto offsets = reader.blockOffsets();
        uint64_t i = 0;
        for ( const auto& [encoded, decoded] : offsets ) {
            if ( i < capacity ) {
                pairs[2 * i] = encoded;
                pairs[2 * i + 1] = decoded;
            }
            ++i;
        }
        *count = i;
        return 0;
    } catch ( const std::exception& e ) {
        g_lastError = e.what();
        return 1;
    }
}

/* The reference's own GZIDX export for a bzip2 file (chunk-level checkpoints). */
int
refgz_export_index( const char* path, const char* indexPath, int parallelization )
{
    try {
        using namespace rapidgzip;
        ParallelGzipReader<ChunkData> reader( std::make_unique<StandardFileReader>( std::string( path ) ),
                                              static_cast<size_t>( parallelization ) );
        reader.read( -1, nullptr, std::numeric_limits<size_t>::max() );
        std::ofstream out( indexPath, std::ios::binary );
        reader.exportIndex( [&out] ( const void* p, size_t n ) { out.write( static_cast<const char*>( p ), static_cast<std::streamsize>( n ) ); } );
        return out ? 0 : 1;
    } catch ( const std::exception& e ) {
        g_lastError = e.what();
        return 1;
    }
}

}  // extern "C"
