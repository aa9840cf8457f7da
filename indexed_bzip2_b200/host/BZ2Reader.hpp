/*
 * BZ2Reader -- the serial flavour of the bzip2 reader, drop-in for indexed_bzip2::BZ2Reader
 * (indexed_bzip2/BZ2Reader.hpp:29-491).  It walks the file the way the reference does -- stream header, block
 * header, block data, next header at the bit where the block ended -- and therefore produces the serial block map
 * (one entry per header it reads, data blocks AND every end-of-stream block, no end sentinel; Appendix B of
 * SURVEY.md), verifies the stream CRC of every stream (BZ2Reader.hpp:406-416, 481-484) and exposes crc().
 *
 * The walk itself is serial only on the host: the block data comes from the same BlockSource as the parallel
 * reader's, i.e. the block at the walk's position is decoded on the GPU together with the blocks the finder lists
 * behind it, and the walk then runs through the decoded batch.
 * Deviation from the reference (documented in DESIGN.md): a block whose CRC fails throws before any of its bytes are
 * handed out; the reference's serial reader emits them first (bzip2.hpp:900-907).
 */
#pragma once

#include <iostream>

#include "BlockSource.hpp"
#include "BZ2ReaderInterface.hpp"

namespace ibz2host
{
class BZ2Reader final : public BZ2ReaderInterface
{
public:
    explicit
    BZ2Reader( UniqueFileReader fileReader, int device = 0 ) :
        m_source( std::move( fileReader ), 1, std::vector<int>{ device } )
    {}

    explicit
    BZ2Reader( const std::string& filePath ) :
        BZ2Reader( std::make_unique<StandardFileReader>( filePath ) )
    {}

    explicit
    BZ2Reader( int fileDescriptor ) :
        BZ2Reader( std::make_unique<StandardFileReader>( fileDescriptor ) )
    {}

    BZ2Reader( const BZ2Reader& ) = delete;
    BZ2Reader& operator=( const BZ2Reader& ) = delete;

    /* FileReader overrides */

    [[nodiscard]] int
    fileno() const
    {
        throw std::logic_error( "This is a virtual file object, which has no corresponding file descriptor!" );
    }

    [[nodiscard]] bool seekable() const override { return true; }

    void
    close() override
    {
        m_source.close();
        m_closed = true;
    }

    [[nodiscard]] bool closed() const override { return m_closed; }
    [[nodiscard]] bool eof() const override { return m_atEndOfFile; }

    [[nodiscard]] bool
    fail() const
    {
        throw std::logic_error( "Not implemented!" );
    }

    [[nodiscard]] size_t
    tell() const override
    {
        if ( m_atEndOfFile ) {
            const auto fileSize = size();
            if ( !fileSize ) {
                throw std::logic_error( "When the file end has been reached, the block map should have been finalized "
                                        "and the file size should be available!" );
            }
            return *fileSize;
        }
        return m_currentPosition;
    }

    [[nodiscard]] std::optional<size_t>
    size() const override
    {
        if ( !m_offsetsComplete ) {
            return std::nullopt;
        }
        return m_offsets.rbegin()->second;
    }

    void
    clearerr()
    {
        m_atEndOfFile = false;
        throw std::invalid_argument( "Not fully tested!" );
    }

    /* BZip2 specific methods */

    [[nodiscard]] uint32_t crc() const { return m_calculatedStreamCRC; }
    [[nodiscard]] bool blockOffsetsComplete() const override { return m_offsetsComplete; }

    [[nodiscard]] std::map<size_t, size_t>
    blockOffsets() override
    {
        if ( !m_offsetsComplete ) {
            read();
        }
        return m_offsets;
    }

    [[nodiscard]] std::map<size_t, size_t>
    availableBlockOffsets() const override
    {
        return m_offsets;
    }

    void
    setBlockOffsets( std::map<size_t, size_t> offsets ) override
    {
        if ( offsets.size() < 2 ) {
            throw std::invalid_argument( "Block offset map must contain at least one valid block and one EOS block!" );
        }
        m_offsetsComplete = true;
        m_offsets = std::move( offsets );
        /* entries that carry bytes are data blocks; their sizes follow from the next entry */
        m_dataBlocks.clear();
        for ( auto it = m_offsets.begin(), nit = std::next( m_offsets.begin() ); nit != m_offsets.end(); ++it, ++nit ) {
            if ( nit->second > it->second ) {
                m_dataBlocks.push_back( { it->first, it->second, nit->second - it->second } );
            }
        }
    }

    [[nodiscard]] size_t
    tellCompressed() const override
    {
        return m_walkPosition;
    }

    using BZ2ReaderInterface::read;

    size_t
    read( const WriteFunctor& writeFunctor,
          const size_t        nBytesToRead = std::numeric_limits<size_t>::max() ) override
    {
        size_t nBytesDecoded = 0;
        while ( ( nBytesDecoded < nBytesToRead ) && !eof() ) {
            if ( const auto* dataBlock = findDataBlock( m_currentPosition ); dataBlock != nullptr ) {
                const auto block = m_source.fetch( dataBlock->encodedOffset );
                const auto offsetInBlock = m_currentPosition - dataBlock->decodedOffset;
                const auto n = std::min<size_t>( block.result->out_size - offsetInBlock, nBytesToRead - nBytesDecoded );
                if ( writeFunctor ) {
                    m_source.deliver( block, offsetInBlock, n, [&writeFunctor] ( const uint8_t* bytes, size_t size ) {
                        writeFunctor( bytes, size );
                    } );
                }
                nBytesDecoded += n;
                m_currentPosition += n;
                continue;
            }
            if ( m_offsetsComplete ) {
                m_atEndOfFile = true;   /* behind the last known block */
                break;
            }
            if ( !walkOn() ) {
                break;
            }
        }
        return nBytesDecoded;
    }

    size_t
    seek( long long int offset, int origin = SEEK_SET ) override
    {
        if ( ( origin == SEEK_END ) && !m_offsetsComplete ) {
            read();
        }
        long long int target = offset;
        if ( origin == SEEK_CUR ) {
            target += static_cast<long long int>( tell() );
        } else if ( origin == SEEK_END ) {
            target += static_cast<long long int>( size().value_or( 0 ) );
        }
        const auto positiveOffset = static_cast<size_t>( std::max( 0LL, target ) );
        if ( positiveOffset == tell() ) {
            return positiveOffset;
        }
        m_source.noteSeek();
        /* When block offsets are not complete yet, emulate forward seeking with a read (BZ2Reader.hpp:318-322). */
        if ( !m_offsetsComplete && ( positiveOffset > tell() ) ) {
            read( -1, nullptr, positiveOffset - tell() );
            return tell();
        }
        if ( !m_offsetsComplete ) {
            const auto here = m_currentPosition;
            m_atEndOfFile = false;
            read();   /* size() and then seeking requires the block offsets to be available */
            m_currentPosition = here;
        }
        m_currentPosition = positiveOffset;
        m_atEndOfFile = positiveOffset >= size().value_or( 0 );
        return tell();
    }

private:
    struct DataBlock
    {
        size_t encodedOffset;
        size_t decodedOffset;
        size_t decodedSize;
    };

    [[nodiscard]] const DataBlock*
    findDataBlock( size_t decodedOffset ) const
    {
        const auto after = std::upper_bound( m_dataBlocks.begin(), m_dataBlocks.end(), decodedOffset,
                                             [] ( size_t offset, const DataBlock& b ) { return offset < b.decodedOffset; } );
        if ( after == m_dataBlocks.begin() ) {
            return nullptr;
        }
        const auto& candidate = *std::prev( after );
        return decodedOffset < candidate.decodedOffset + candidate.decodedSize ? &candidate : nullptr;
    }

    /** bzip2::readBzip2Header (bzip2.hpp:114-142) on the host copy of the file, with the reference's messages. */
    void
    readStreamHeader()
    {
        const auto byteAt = [this] ( size_t bitPosition ) {
            if ( bitPosition + 8U > m_source.fileSizeInBits() ) {
                throw EndOfFileReached();
            }
            const uint8_t* const data = m_source.fileData();
            const size_t byte = bitPosition / 8U;
            const unsigned shift = static_cast<unsigned>( bitPosition % 8U );
            const unsigned value = shift == 0 ? data[byte]
                                              : ( ( data[byte] << shift ) | ( data[byte + 1] >> ( 8U - shift ) ) ) & 0xFFU;
            return static_cast<char>( static_cast<uint8_t>( value ) );
        };
        size_t position = m_walkPosition;
        for ( const char magicByte : { 'B', 'Z', 'h' } ) {
            const char readByte = byteAt( position );
            position += 8;
            if ( readByte != magicByte ) {
                std::stringstream msg;
                msg << "Input header is not BZip2 magic string 'BZh' (0x"
                    << std::hex << int( 'B' ) << int( 'Z' ) << int( 'h' ) << std::dec << "). Mismatch at bit position "
                    << position - 8 << " with " << readByte << " (0x" << std::hex
                    << static_cast<int>( static_cast<uint8_t>( readByte ) ) << ") should be " << magicByte;
                throw std::domain_error( msg.str() );
            }
        }
        const char level = byteAt( position );
        position += 8;
        if ( ( level < '1' ) || ( level > '9' ) ) {
            std::stringstream msg;
            msg << "Blocksize must be one of '0' (" << std::hex << static_cast<int>( '0' ) << ") ... '9' ("
                << static_cast<int>( '9' ) << ") but is " << level << " (" << static_cast<int>( level ) << ")";
            throw std::domain_error( msg.str() );
        }
        m_walkPosition = position;
        m_calculatedStreamCRC = 0;
    }

    /**
     * One step of the serial walk (BZ2Reader.hpp:233-264 and :395-491): a stream header if one is due, then one block
     * header; a data block is decoded and entered, an end-of-stream block closes its stream.
     * @return false when the compressed input is used up (or only garbage follows).
     */
    bool
    walkOn()
    {
        if ( m_walkPosition >= m_source.fileSizeInBits() ) {
            return false;
        }
        if ( m_walkPosition == 0 ) {
            readStreamHeader();
        } else if ( m_lastWasEndOfStream ) {
            try {
                readStreamHeader();
            } catch ( const std::domain_error& ) {
                std::cerr << "[Warning] Trailing garbage after EOF ignored!\n";
                m_atEndOfFile = true;
                m_offsetsComplete = true;
                return false;
            }
        }

        /* readBlockHeader (BZ2Reader.hpp:395-424): the map entry is made before the header is parsed */
        const auto headerOffset = m_walkPosition;
        m_offsets.insert( { headerOffset, m_decodedBytesCount } );
        const auto header = m_source.probeHeader( headerOffset );
        if ( header.status == IBZ2_ST_EOF ) {
            throw EndOfFileReached();
        }
        if ( header.status != IBZ2_ST_OK ) {
            throwForStatus( header );
        }

        if ( header.is_eos != 0 ) {
            m_streamCRC = header.expected_crc;
            if ( m_streamCRC != m_calculatedStreamCRC ) {
                std::stringstream msg;
                msg << "[BZip2 block header] Stream CRC 0x" << std::hex << m_streamCRC
                    << " does not match calculated CRC 0x" << m_calculatedStreamCRC;
                throw std::runtime_error( msg.str() );
            }
            m_atEndOfFile = header.is_eof != 0;
            if ( m_atEndOfFile ) {
                m_offsetsComplete = true;
            }
            m_walkPosition = headerOffset + header.enc_size_bits;
            m_lastWasEndOfStream = true;
            return !m_atEndOfFile;
        }

        const auto block = m_source.fetch( headerOffset );
        m_dataBlocks.push_back( { headerOffset, m_decodedBytesCount, block.result->out_size } );
        m_decodedBytesCount += block.result->out_size;
        m_calculatedStreamCRC = ibz2_stream_crc_fold( m_calculatedStreamCRC, block.result->calculated_crc );
        m_walkPosition = headerOffset + block.result->enc_size_bits;
        m_lastWasEndOfStream = false;
        return true;
    }

    BlockSource m_source;
    std::map<size_t, size_t> m_offsets;
    std::vector<DataBlock> m_dataBlocks;
    bool m_offsetsComplete{ false };

    size_t m_walkPosition{ 0 };        /* compressed bit position of the next header */
    size_t m_decodedBytesCount{ 0 };   /* decoded bytes in front of that header */
    bool m_lastWasEndOfStream{ false };
    uint32_t m_streamCRC{ 0 };
    uint32_t m_calculatedStreamCRC{ 0 };

    size_t m_currentPosition{ 0 };
    bool m_atEndOfFile{ false };
    bool m_closed{ false };
};
}  // namespace ibz2host
