/*
 * reader_capi.cpp -- flat C entry points over ParallelBZ2Reader / BZ2Reader for FFI callers (the Python module
 * indexed_bzip2_b200/__init__.py binds them with ctypes, the way the reference's Cython wrapper binds the C++ classes,
 * python/indexed_bzip2/indexed_bzip2.pyx:26-67).  Exceptions are caught at this boundary; the kind of the last one
 * is kept so that the caller can raise the exception type Cython's "except +" would have produced.
 */
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <fstream>
#include <memory>
#include <vector>
#include <new>
#include <optional>
#include <string>

#include "BZ2Reader.hpp"
#include "Bzip2Chunk.hpp"
#include "ParallelBZ2Reader.hpp"

using namespace ibz2host;

int ibzip2CLI( int argc, char** argv );   /* tools/ibzip2.cpp */

#include <csignal>
#include <execinfo.h>
namespace
{
/* IBZ2_DEBUG_SEGV=1: print a native backtrace when the process dies inside this library */
void
segvHandler( int sig )
{
    void* frames[64];
    const int n = backtrace( frames, 64 );
    backtrace_symbols_fd( frames, n, 2 );
    signal( sig, SIG_DFL );
    raise( sig );
}

struct SegvInstaller
{
    SegvInstaller()
    {
        if ( getenv( "IBZ2_DEBUG_SEGV" ) != nullptr ) {
            signal( SIGSEGV, segvHandler );
        }
    }
} g_segvInstaller;
}  // namespace

extern "C" {
enum ibz2r_error_kind {
    IBZ2R_NONE = 0,
    IBZ2R_INVALID_ARGUMENT = 1,   /* ValueError */
    IBZ2R_DOMAIN_ERROR = 2,       /* ValueError */
    IBZ2R_OUT_OF_RANGE = 3,       /* IndexError */
    IBZ2R_RUNTIME_ERROR = 4,      /* RuntimeError */
    IBZ2R_LOGIC_ERROR = 5,        /* RuntimeError */
    IBZ2R_END_OF_FILE = 6,        /* RuntimeError (EndOfFileReached derives from std::exception) */
    IBZ2R_BAD_ALLOC = 7,          /* MemoryError */
    IBZ2R_OTHER = 8               /* RuntimeError */
};
}

struct ibz2_reader
{
    std::unique_ptr<BZ2ReaderInterface> reader;
    ParallelBZ2Reader* parallel{ nullptr };
    std::string lastError;
    int lastKind{ IBZ2R_NONE };
};

namespace
{
thread_local std::string g_openError;
thread_local int g_openKind = IBZ2R_NONE;
thread_local int g_openDevice = -1;   /* ibz2r_set_open_device; -1 = $IBZ2_DEVICE, else device 0 */
thread_local std::vector<int> g_openDevices;   /* ibz2r_set_open_devices; empty = the single device above */

int
openDevice()
{
    if ( g_openDevice >= 0 ) {
        return g_openDevice;
    }
    const char* const env = std::getenv( "IBZ2_DEVICE" );
    return env != nullptr ? std::max( 0, std::atoi( env ) ) : 0;
}

template<typename Function>
long long
guarded( std::string& message, int& kind, Function&& function )
{
    kind = IBZ2R_NONE;
    message.clear();
    try {
        return function();
    } catch ( const EndOfFileReached& e ) {
        kind = IBZ2R_END_OF_FILE;
        message = "Unexpected end of file. Truncated or invalid bzip2?";
    } catch ( const std::invalid_argument& e ) {
        kind = IBZ2R_INVALID_ARGUMENT;
        message = e.what();
    } catch ( const std::domain_error& e ) {
        kind = IBZ2R_DOMAIN_ERROR;
        message = e.what();
    } catch ( const std::out_of_range& e ) {
        kind = IBZ2R_OUT_OF_RANGE;
        message = e.what();
    } catch ( const std::logic_error& e ) {
        kind = IBZ2R_LOGIC_ERROR;
        message = e.what();
    } catch ( const std::bad_alloc& e ) {
        kind = IBZ2R_BAD_ALLOC;
        message = e.what();
    } catch ( const std::runtime_error& e ) {
        kind = IBZ2R_RUNTIME_ERROR;
        message = e.what();
    } catch ( const std::exception& e ) {
        kind = IBZ2R_OTHER;
        message = e.what();
    } catch ( ... ) {
        kind = IBZ2R_OTHER;
        message = "unknown exception";
    }
    return -1;
}

ibz2_reader*
wrap( UniqueFileReader fileReader, size_t parallelization )
{
    auto handle = std::make_unique<ibz2_reader>();
    /* parallelization == 1 selects the serial class, anything else the parallel one (indexed_bzip2.pyx:295-296) */
    if ( parallelization == 1 ) {
        handle->reader = std::make_unique<BZ2Reader>( std::move( fileReader ), openDevice() );
    } else {
        auto parallel = g_openDevices.empty()
                        ? std::make_unique<ParallelBZ2Reader>( std::move( fileReader ), parallelization, openDevice() )
                        : std::make_unique<ParallelBZ2Reader>( std::move( fileReader ), parallelization, g_openDevices );
        handle->parallel = parallel.get();
        handle->reader = std::move( parallel );
    }
    return handle.release();
}

template<typename Function>
ibz2_reader*
openWith( Function&& function )
{
    ibz2_reader* result = nullptr;
    guarded( g_openError, g_openKind, [&] () { result = function(); return 0LL; } );
    return result;
}
}  // namespace

extern "C" {
/* the CUDA device the readers opened next by this thread decode on (the C++ constructors take it as an argument) */
void
ibz2r_set_open_device( int device )
{
    g_openDevice = device;
}

/* the set of devices a parallel reader opened next by this thread spreads its batches over (count 0: one device) */
void
ibz2r_set_open_devices( const int* devices, int count )
{
    g_openDevices.assign( devices, devices + ( ( devices != nullptr ) && ( count > 0 ) ? count : 0 ) );
}

ibz2_reader*
ibz2r_open_path( const char* path, size_t parallelization )
{
    return openWith( [&] () { return wrap( std::make_unique<StandardFileReader>( std::string( path ) ), parallelization ); } );
}

ibz2_reader*
ibz2r_open_fd( int fd, size_t parallelization )
{
    return openWith( [&] () { return wrap( std::make_unique<StandardFileReader>( fd ), parallelization ); } );
}

ibz2_reader*
ibz2r_open_memory( const void* data, size_t size, size_t parallelization )
{
    return openWith( [&] () { return wrap( std::make_unique<MemoryFileReader>( data, size ), parallelization ); } );
}

void
ibz2r_destroy( ibz2_reader* r )
{
    delete r;
}

const char*
ibz2r_last_error( const ibz2_reader* r )
{
    return r != nullptr ? r->lastError.c_str() : g_openError.c_str();
}

int
ibz2r_last_error_kind( const ibz2_reader* r )
{
    return r != nullptr ? r->lastKind : g_openKind;
}

long long
ibz2r_read( ibz2_reader* r, void* buffer, size_t size )
{
    return guarded( r->lastError, r->lastKind, [&] () {
        return static_cast<long long>( r->reader->read( -1, static_cast<char*>( buffer ), size ) );
    } );
}

long long
ibz2r_read_to_fd( ibz2_reader* r, int fd, size_t size )
{
    return guarded( r->lastError, r->lastKind, [&] () {
        return static_cast<long long>( r->reader->read( fd, nullptr, size ) );
    } );
}

long long
ibz2r_seek( ibz2_reader* r, long long offset, int whence )
{
    return guarded( r->lastError, r->lastKind, [&] () { return static_cast<long long>( r->reader->seek( offset, whence ) ); } );
}

long long
ibz2r_tell( ibz2_reader* r )
{
    return guarded( r->lastError, r->lastKind, [&] () { return static_cast<long long>( r->reader->tell() ); } );
}

/* the decoded size; -1 while the block map is incomplete unless `force` asks to decode to the end first */
long long
ibz2r_size( ibz2_reader* r, int force )
{
    return guarded( r->lastError, r->lastKind, [&] () {
        if ( !r->reader->size() ) {
            if ( force == 0 ) {
                return -1LL;
            }
            const auto here = r->reader->tell();
            r->reader->seek( 0, SEEK_END );
            r->reader->seek( static_cast<long long>( here ), SEEK_SET );
        }
        return static_cast<long long>( r->reader->size().value_or( 0 ) );
    } );
}

long long
ibz2r_tell_compressed( ibz2_reader* r )
{
    return guarded( r->lastError, r->lastKind, [&] () { return static_cast<long long>( r->reader->tellCompressed() ); } );
}

int
ibz2r_eof( ibz2_reader* r )
{
    return r->reader->eof() ? 1 : 0;
}

int
ibz2r_closed( ibz2_reader* r )
{
    return r->reader->closed() ? 1 : 0;
}

int
ibz2r_close( ibz2_reader* r )
{
    return static_cast<int>( guarded( r->lastError, r->lastKind, [&] () { r->reader->close(); return 0LL; } ) );
}

int
ibz2r_block_offsets_complete( ibz2_reader* r )
{
    return r->reader->blockOffsetsComplete() ? 1 : 0;
}

/* pairs: ( encoded bits, decoded bytes ); returns the number of entries (which may exceed capacity), -1 on error */
long long
ibz2r_block_offsets( ibz2_reader* r, uint64_t* pairs, size_t capacity, int availableOnly )
{
    return guarded( r->lastError, r->lastKind, [&] () {
        const auto offsets = availableOnly != 0 ? r->reader->availableBlockOffsets() : r->reader->blockOffsets();
        size_t i = 0;
        for ( const auto& [bits, bytes] : offsets ) {
            if ( i < capacity ) {
                pairs[2 * i] = bits;
                pairs[2 * i + 1] = bytes;
            }
            ++i;
        }
        return static_cast<long long>( offsets.size() );
    } );
}

int
ibz2r_set_block_offsets( ibz2_reader* r, const uint64_t* pairs, size_t count )
{
    return static_cast<int>( guarded( r->lastError, r->lastKind, [&] () {
        std::map<size_t, size_t> offsets;
        for ( size_t i = 0; i < count; ++i ) {
            offsets.emplace( pairs[2 * i], pairs[2 * i + 1] );
        }
        r->reader->setBlockOffsets( std::move( offsets ) );
        return 0LL;
    } ) );
}

/* ---- Bzip2Chunk::decodeChunk for FFI callers and the tests: one chunk of a file on disk ----
 * info[0..3] = encodedOffsetInBits, encodedSizeInBits, decodedSizeInBytes, stoppedPreemptively;
 * boundaries / footers: ( encoded offset, decoded offset ) pairs; counts[0..1] = how many there are of each.
 * Returns 0, or -1 with ibz2r_last_error( NULL ) / ibz2r_last_error_kind( NULL ) set. */
int
ibz2r_decode_chunk( const char* path, uint64_t chunkOffset, uint64_t untilOffset, uint64_t maxDecompressedChunkSize,
                    uint64_t* info, uint64_t* boundaries, size_t boundariesCapacity, uint64_t* footers, size_t footersCapacity,
                    uint64_t* counts, uint8_t* out, uint64_t outCapacity )
{
    return static_cast<int>( guarded( g_openError, g_openKind, [&] () {
        std::ifstream in( path, std::ios::binary );
        if ( !in ) {
            throw std::invalid_argument( std::string( "Cannot open " ) + path );
        }
        const std::vector<char> file( ( std::istreambuf_iterator<char>( in ) ), std::istreambuf_iterator<char>() );
        const auto* const bytes = reinterpret_cast<const uint8_t*>( file.data() );
        GpuDecoder decoder( openDevice() );
        decoder.upload( bytes, file.size() );
        const auto blocks = decoder.findBlocks();
        auto chunk = Bzip2Chunk::decodeChunk( decoder, bytes, file.size(), blocks, chunkOffset, untilOffset,
                                              maxDecompressedChunkSize );
        info[0] = chunk.encodedOffsetInBits;
        info[1] = chunk.encodedSizeInBits;
        info[2] = chunk.decodedSizeInBytes;
        info[3] = chunk.stoppedPreemptively ? 1 : 0;
        counts[0] = chunk.blockBoundaries.size();
        counts[1] = chunk.footers.size();
        for ( size_t i = 0; i < std::min( boundariesCapacity, chunk.blockBoundaries.size() ); ++i ) {
            boundaries[2 * i] = chunk.blockBoundaries[i].encodedOffset;
            boundaries[2 * i + 1] = chunk.blockBoundaries[i].decodedOffset;
        }
        for ( size_t i = 0; i < std::min( footersCapacity, chunk.footers.size() ); ++i ) {
            footers[2 * i] = chunk.footers[i].encodedOffset;
            footers[2 * i + 1] = chunk.footers[i].decodedOffset;
        }
        if ( out != nullptr ) {
            uint64_t written = 0;
            decoder.deliver( static_cast<const uint8_t*>( chunk.data.data ), std::min<uint64_t>( chunk.decodedSizeInBytes, outCapacity ),
                             [&] ( const uint8_t* piece, size_t n ) {
                std::memcpy( out + written, piece, n );
                written += n;
            } );
        }
        decoder.freeDevice( chunk.data );
        return 0LL;
    } ) );
}

/* ---- test hooks (no GPU): the GZIDX writer and reader on plain offset pairs ---- */
int
ibz2r_gzidx_write( const uint64_t* pairs, size_t count, uint64_t compressedSizeInBytes, const char* path )
{
    try {
        std::map<size_t, size_t> offsets;
        for ( size_t i = 0; i < count; ++i ) {
            offsets.emplace( pairs[2 * i], pairs[2 * i + 1] );
        }
        std::ofstream out( path, std::ios::binary );
        writeGzidx( out, offsets, compressedSizeInBytes );
        return 0;
    } catch ( const std::exception& e ) {
        g_openError = e.what();
        return -1;
    }
}

long long
ibz2r_gzidx_read( const char* path, uint64_t* pairs, size_t capacity )
{
    try {
        std::ifstream in( path, std::ios::binary );
        const auto offsets = readGzidx( in );
        size_t i = 0;
        for ( const auto& [bits, bytes] : offsets ) {
            if ( i < capacity ) {
                pairs[2 * i] = bits;
                pairs[2 * i + 1] = bytes;
            }
            ++i;
        }
        return static_cast<long long>( offsets.size() );
    } catch ( const std::exception& e ) {
        g_openError = e.what();
        return -1;
    }
}

/* The block map to / from an index file: format 0 = the "<bits>,<bytes>" lines of `ibzip2 -L`, 1 = GZIDX (IndexFileFormat.hpp).
 * The import tells the formats apart by the magic bytes. */
int
ibz2r_export_index( ibz2_reader* r, const char* path, int format )
{
    return static_cast<int>( guarded( r->lastError, r->lastKind, [&] () {
        std::ofstream out( path, std::ios::binary );
        if ( !out ) {
            throw std::invalid_argument( std::string( "Cannot open index file for writing: " ) + path );
        }
        if ( r->parallel != nullptr ) {
            r->parallel->exportIndex( out, format == 1 ? IndexFormat::GZIDX : IndexFormat::OFFSET_LINES );
        } else if ( format == 1 ) {
            const auto offsets = r->reader->blockOffsets();
            /* the serial reader's map has no end-of-stream entries and ends at the last block's end: the same checkpoints */
            writeGzidx( out, offsets, offsets.empty() ? 0 : ( offsets.rbegin()->first + 7U ) / 8U );
        } else {
            for ( const auto& [bits, bytes] : r->reader->blockOffsets() ) {
                out << bits << "," << bytes << "\n";
            }
        }
        return 0LL;
    } ) );
}

int
ibz2r_import_index( ibz2_reader* r, const char* path )
{
    return static_cast<int>( guarded( r->lastError, r->lastKind, [&] () {
        std::ifstream in( path, std::ios::binary );
        if ( !in ) {
            throw std::invalid_argument( std::string( "Cannot open index file: " ) + path );
        }
        if ( r->parallel != nullptr ) {
            r->parallel->importIndex( in );
            return 0LL;
        }
        std::map<size_t, size_t> offsets;
        if ( in.peek() == 'G' ) {
            offsets = readGzidx( in );
        } else {
            std::string line;
            while ( std::getline( in, line ) ) {
                const auto comma = line.find( ',' );
                if ( comma != std::string::npos ) {
                    offsets.emplace( std::stoull( line.substr( 0, comma ) ), std::stoull( line.substr( comma + 1 ) ) );
                }
            }
        }
        r->reader->setBlockOffsets( std::move( offsets ) );
        return 0LL;
    } ) );
}

/* ParallelBZ2Reader::setVerifyStreamCRC; the serial class always checks (BZ2Reader.hpp:406-416) */
int
ibz2r_set_verify_stream_crc( ibz2_reader* r, int verify )
{
    if ( r->parallel != nullptr ) {
        r->parallel->setVerifyStreamCRC( verify != 0 );
    }
    return 0;
}

int
ibz2r_join_threads( ibz2_reader* r )
{
    if ( r->parallel != nullptr ) {
        r->parallel->joinThreads();
    }
    return 0;
}

/* ---- test hook (no GPU): drives a BlockIndex the way the readers do, so that the block-map rules can be checked
 * against the reference's golden maps on a machine without a GPU.
 * triples: ( encoded offset, encoded size, decoded size ) per append; returns the number of map entries, -1 on error. */
long long
ibz2r_index_simulate( const uint64_t* triples, size_t count, int finish, uint64_t* pairs, size_t capacity,
                      uint64_t probe, uint64_t* located )
{
    std::string message;
    int kind = IBZ2R_NONE;
    return guarded( message, kind, [&] () {
        BlockIndex index;
        for ( size_t i = 0; i < count; ++i ) {
            index.append( triples[3 * i], triples[3 * i + 1], triples[3 * i + 2] );
        }
        if ( finish != 0 ) {
            index.finish();
        }
        const auto offsets = index.offsets();
        size_t i = 0;
        for ( const auto& [bits, bytes] : offsets ) {
            if ( i < capacity ) {
                pairs[2 * i] = bits;
                pairs[2 * i + 1] = bytes;
            }
            ++i;
        }
        if ( located != nullptr ) {
            const auto span = index.locate( probe );
            located[0] = span.valid ? 1 : 0;
            located[1] = span.encodedOffsetInBits;
            located[2] = span.decodedOffsetInBytes;
            located[3] = span.decodedSizeInBytes;
            located[4] = index.dataBlockCount();
        }
        return static_cast<long long>( offsets.size() );
    } );
}

/* the reference's Python cli() calls ibzip2CLI with sys.argv (indexed_bzip2.pyx:343-363) */
int
ibz2r_cli( int argc, char** argv )
{
    std::string message;
    int kind = IBZ2R_NONE;
    const auto rc = guarded( message, kind, [&] () { return static_cast<long long>( ibzip2CLI( argc, argv ) ); } );
    if ( kind == IBZ2R_END_OF_FILE ) {
        std::cerr << "Unexpected end of file. Truncated or invalid bzip2?\n";
        return 1;
    }
    if ( kind != IBZ2R_NONE ) {
        std::cerr << "Caught exception: " << message << "\n";
        return 1;
    }
    return static_cast<int>( rc );
}
}  // extern "C"
