/*
 * BZ2ReaderInterface -- the abstract reader both bzip2 readers implement; same method set, argument meaning and
 * error behaviour as the reference's (indexed_bzip2/BZ2ReaderInterface.hpp:15-103), so that code written against
 * indexed_bzip2::BZ2ReaderInterface (the ibzip2 CLI, the Python wrapper) compiles against this one.
 */
#pragma once

#include <cerrno>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <functional>
#include <limits>
#include <map>
#include <sstream>
#include <stdexcept>

#include <unistd.h>

#include "FileReader.hpp"

namespace ibz2host
{
/** writeAll of core/FileUtils.hpp:804-827: to a descriptor if one is given, else to a buffer if one is given, else nowhere. */
inline int
writeAll( int outputFileDescriptor, char* outputBuffer, const void* data, uint64_t size )
{
    if ( outputFileDescriptor >= 0 ) {
        const char* p = static_cast<const char*>( data );
        uint64_t left = size;
        while ( left > 0 ) {
            const auto n = ::write( outputFileDescriptor, p, std::min<uint64_t>( left, 1ULL << 30 ) );
            if ( n < 0 ) {
                if ( errno == EINTR ) {
                    continue;
                }
                return errno;
            }
            if ( n == 0 ) {
                return EIO;
            }
            p += n;
            left -= static_cast<uint64_t>( n );
        }
    } else if ( outputBuffer != nullptr ) {
        std::memcpy( outputBuffer, data, size );
    }
    return 0;
}

class BZ2ReaderInterface : public FileReader
{
public:
    using WriteFunctor = std::function<void ( const void*, uint64_t )>;

    ~BZ2ReaderInterface() override = default;

    [[nodiscard]] size_t
    read( char* outputBuffer, size_t nBytesToRead ) final
    {
        return read( -1, outputBuffer, nBytesToRead );
    }

    /**
     * @param[out] outputBuffer should at least be large enough to hold @p nBytesToRead bytes
     * @return number of bytes written
     */
    size_t
    read( const int    outputFileDescriptor = -1,
          char* const  outputBuffer         = nullptr,
          const size_t nBytesToRead         = std::numeric_limits<size_t>::max() )
    {
        uint64_t nBytesDecoded = 0;
        const WriteFunctor writeFunctor =
            [&nBytesDecoded, outputFileDescriptor, outputBuffer] ( const void* const buffer, uint64_t const size ) {
                char* const position = outputBuffer == nullptr ? nullptr : outputBuffer + nBytesDecoded;
                const int errorCode = writeAll( outputFileDescriptor, position, buffer, size );
                if ( errorCode != 0 ) {
                    std::stringstream message;
                    message << "Failed to write all bytes because of: " << std::strerror( errorCode )
                            << " (" << errorCode << ")";
                    throw std::runtime_error( message.str() );
                }
                nBytesDecoded += size;
            };
        return read( writeFunctor, nBytesToRead );
    }

    virtual size_t
    read( const WriteFunctor& writeFunctor,
          size_t              nBytesToRead = std::numeric_limits<size_t>::max() ) = 0;

    /* BZip2 specific methods */

    [[nodiscard]] virtual bool blockOffsetsComplete() const = 0;

    /** @return compressed bit offset of every block -> offset of its first decoded byte; decodes to the end if needed. */
    virtual std::map<size_t, size_t> blockOffsets() = 0;

    /** Same as @ref blockOffsets but only what is known at call time. */
    [[nodiscard]] virtual std::map<size_t, size_t> availableBlockOffsets() const = 0;

    virtual void setBlockOffsets( std::map<size_t, size_t> offsets ) = 0;

    /** @return number of processed bits of the compressed stream (block granularity). */
    [[nodiscard]] virtual size_t tellCompressed() const = 0;

    void
    setShowProfileOnDestruction( bool showProfileOnDestruction )
    {
        m_showProfileOnDestruction = showProfileOnDestruction;
    }

protected:
    bool m_showProfileOnDestruction{ false };
};
}  // namespace ibz2host
