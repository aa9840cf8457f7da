/*
 * FileReader -- the byte source handed to the readers, with the part of the reference's virtual interface the bzip2
 * readers use (core/filereader/FileReader.hpp:24-141): read / seek / tell / size / eof / seekable / close.
 * StandardFileReader wraps a path or a file descriptor (filereader/Standard.hpp), MemoryFileReader a byte vector
 * (filereader/BufferView.hpp; also what the Python layer uses for file-like objects it has read itself).
 */
#pragma once

#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

namespace ibz2host
{
class FileReader
{
public:
    virtual ~FileReader() = default;
    [[nodiscard]] virtual bool eof() const = 0;
    [[nodiscard]] virtual bool seekable() const = 0;
    [[nodiscard]] virtual bool closed() const = 0;
    virtual void close() = 0;
    virtual size_t read( char* buffer, size_t nMaxBytesToRead ) = 0;
    virtual size_t seek( long long int offset, int origin = SEEK_SET ) = 0;
    [[nodiscard]] virtual size_t tell() const = 0;
    [[nodiscard]] virtual std::optional<size_t> size() const = 0;
};

using UniqueFileReader = std::unique_ptr<FileReader>;

class StandardFileReader final : public FileReader
{
public:
    explicit
    StandardFileReader( const std::string& path ) :
        m_fd( ::open( path.c_str(), O_RDONLY ) ),
        m_owned( true )
    {
        if ( m_fd < 0 ) {
            throw std::invalid_argument( "Opening file '" + path + "' with mode 'rb' failed!" );
        }
        init();
    }

    /** The descriptor is duplicated: the caller keeps ownership of the one passed in (filereader/Standard.hpp:49-55). */
    explicit
    StandardFileReader( int fileDescriptor ) :
        m_fd( ::dup( fileDescriptor ) ),
        m_owned( true )
    {
        if ( m_fd < 0 ) {
            throw std::invalid_argument( "Cannot duplicate the given file descriptor!" );
        }
        init();
    }

    ~StandardFileReader() override
    {
        close();
    }

    [[nodiscard]] bool eof() const override { return m_seekable ? ( m_position >= m_size ) : m_sawEnd; }
    [[nodiscard]] bool seekable() const override { return m_seekable; }
    [[nodiscard]] bool closed() const override { return m_fd < 0; }

    void
    close() override
    {
        if ( ( m_fd >= 0 ) && m_owned ) {
            ::close( m_fd );
        }
        m_fd = -1;
    }

    size_t
    read( char* buffer, size_t nMaxBytesToRead ) override
    {
        if ( m_fd < 0 ) {
            throw std::invalid_argument( "Invalid or closed file cannot be read from!" );
        }
        size_t done = 0;
        while ( done < nMaxBytesToRead ) {
            const auto n = m_seekable ? ::pread( m_fd, buffer + done, nMaxBytesToRead - done, static_cast<off_t>( m_position ) )
                                      : ::read( m_fd, buffer + done, nMaxBytesToRead - done );
            if ( n < 0 ) {
                if ( errno == EINTR ) {
                    continue;
                }
                throw std::runtime_error( std::string( "Reading from the file failed: " ) + std::strerror( errno ) );
            }
            if ( n == 0 ) {
                m_sawEnd = true;
                break;
            }
            done += static_cast<size_t>( n );
            m_position += static_cast<size_t>( n );
        }
        return done;
    }

    size_t
    seek( long long int offset, int origin = SEEK_SET ) override
    {
        if ( !m_seekable ) {
            throw std::invalid_argument( "Invalid or file can't be seeked!" );
        }
        long long int target = offset;
        if ( origin == SEEK_CUR ) {
            target += static_cast<long long int>( m_position );
        } else if ( origin == SEEK_END ) {
            target += static_cast<long long int>( m_size );
        }
        m_position = static_cast<size_t>( std::max( 0LL, std::min<long long int>( target, static_cast<long long int>( m_size ) ) ) );
        return m_position;
    }

    [[nodiscard]] size_t tell() const override { return m_position; }

    [[nodiscard]] std::optional<size_t>
    size() const override
    {
        return m_seekable ? std::optional<size_t>( m_size ) : std::nullopt;
    }

    /** The descriptor itself, for callers that map the file instead of reading it (filereader/Standard.hpp:88-95). */
    [[nodiscard]] int
    fileno() const
    {
        if ( m_fd < 0 ) {
            throw std::invalid_argument( "Trying to get fileno of an invalid file!" );
        }
        return m_fd;
    }

private:
    void
    init()
    {
        struct stat st{};
        m_seekable = ( ::fstat( m_fd, &st ) == 0 ) && S_ISREG( st.st_mode );
        if ( m_seekable ) {
            m_size = static_cast<size_t>( st.st_size );
        }
    }

    int m_fd{ -1 };
    bool m_owned{ false };
    bool m_seekable{ false };
    bool m_sawEnd{ false };
    size_t m_size{ 0 };
    size_t m_position{ 0 };
};

class MemoryFileReader final : public FileReader
{
public:
    explicit
    MemoryFileReader( std::vector<uint8_t> data ) :
        m_data( std::move( data ) )
    {}

    MemoryFileReader( const void* data, size_t size ) :
        m_data( static_cast<const uint8_t*>( data ), static_cast<const uint8_t*>( data ) + size )
    {}

    [[nodiscard]] bool eof() const override { return m_position >= m_data.size(); }
    [[nodiscard]] bool seekable() const override { return true; }
    [[nodiscard]] bool closed() const override { return m_closed; }
    void close() override { m_closed = true; }

    size_t
    read( char* buffer, size_t nMaxBytesToRead ) override
    {
        const size_t n = std::min( nMaxBytesToRead, m_data.size() - std::min( m_position, m_data.size() ) );
        if ( n > 0 ) {
            std::memcpy( buffer, m_data.data() + m_position, n );
        }
        m_position += n;
        return n;
    }

    size_t
    seek( long long int offset, int origin = SEEK_SET ) override
    {
        long long int target = offset;
        if ( origin == SEEK_CUR ) {
            target += static_cast<long long int>( m_position );
        } else if ( origin == SEEK_END ) {
            target += static_cast<long long int>( m_data.size() );
        }
        m_position = static_cast<size_t>( std::max( 0LL, std::min<long long int>( target, static_cast<long long int>( m_data.size() ) ) ) );
        return m_position;
    }

    [[nodiscard]] size_t tell() const override { return m_position; }
    [[nodiscard]] std::optional<size_t> size() const override { return m_data.size(); }

    /** Lets a reader take the bytes without another copy. */
    [[nodiscard]] std::vector<uint8_t>&
    bytes()
    {
        return m_data;
    }

private:
    std::vector<uint8_t> m_data;
    size_t m_position{ 0 };
    bool m_closed{ false };
};
}  // namespace ibz2host
