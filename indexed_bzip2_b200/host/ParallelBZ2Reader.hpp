/*
 * ParallelBZ2Reader -- seekable reader over a bzip2 file whose blocks are decoded on a B200 (include/ibz2_cuda.h).
 * Drop-in for indexed_bzip2::ParallelBZ2Reader (indexed_bzip2/ParallelBZ2Reader.hpp:39-498): same constructors,
 * methods, positions, end-of-file behaviour, exceptions and -- bit for bit -- the same block map.
 *
 * Where the reference has a block-finder thread and a pool of decode threads with prefetching, this reader has
 *   - the whole compressed file in HBM (uploaded on first use),
 *   - ONE finder pass over it (K1) that yields every data-block offset,
 *   - batches: when a read needs a block that is not decoded yet, that block and the following ones are decoded in
 *     one ibz2_decode_blocks call.  Sequential reading doubles the batch up to MAX_BATCH_BLOCKS (the GPU wants
 *     thousands of blocks in flight); a seek resets it to a few blocks so that random access stays cheap.
 * The block map is filled in file order exactly as ParallelBZ2Reader.hpp:191-238 does it, including the probing for
 * end-of-stream blocks and the "trailing garbage" cut-off; it may only run AHEAD of the reference's (a decoded batch is
 * entered at once), never differ from it.
 * Like the reference, an object must not be used from several threads at once.
 */
#pragma once

#include <algorithm>
#include <iostream>
#include <memory>
#include <sstream>
#include <vector>

#include "BlockIndex.hpp"
#include "IndexFileFormat.hpp"
#include "BZ2ReaderInterface.hpp"
#include "BlockSource.hpp"

namespace ibz2host
{
class ParallelBZ2Reader final : public BZ2ReaderInterface
{
public:
    explicit
    ParallelBZ2Reader( UniqueFileReader fileReader, size_t parallelization = 0, int device = 0 ) :
        m_source( requireSeekable( std::move( fileReader ) ), parallelization, std::vector<int>{ device } )
    {}

    /**
     * New next to the reference: the GPUs of one machine that share the decoding of every large batch, one host thread
     * per device -- where the reference scales with the threads of its pool (ParallelBZ2Reader.hpp:50-69), this reader
     * scales with devices.  Positions, block map and bytes do not depend on the set.
     */
    ParallelBZ2Reader( UniqueFileReader fileReader, size_t parallelization, std::vector<int> devices ) :
        m_source( requireSeekable( std::move( fileReader ) ), parallelization, std::move( devices ) )
    {}

    explicit
    ParallelBZ2Reader( int fileDescriptor, size_t parallelization = 0 ) :
        ParallelBZ2Reader( std::make_unique<StandardFileReader>( fileDescriptor ), parallelization )
    {}

    explicit
    ParallelBZ2Reader( const std::string& filePath, size_t parallelization = 0 ) :
        ParallelBZ2Reader( std::make_unique<StandardFileReader>( filePath ), parallelization )
    {}

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
        if ( !m_index.finished() ) {
            return std::nullopt;
        }
        return m_index.back().second;
    }

    void
    clearerr()
    {
        m_atEndOfFile = false;
        throw std::invalid_argument( "Not fully tested!" );
    }

    /* BZ2ReaderInterface overrides */

    using BZ2ReaderInterface::read;

    size_t
    read( const WriteFunctor& writeFunctor,
          const size_t        nBytesToRead = std::numeric_limits<size_t>::max() ) override
    {
        if ( closed() ) {
            throw std::invalid_argument( "You may not call read on closed ParallelBZ2Reader!" );
        }
        if ( eof() || ( nBytesToRead == 0 ) ) {
            return 0;
        }

        /* with the block map complete the blocks this read touches are known: they are decoded by one call */
        if ( m_index.finished() && ( nBytesToRead != std::numeric_limits<size_t>::max() ) ) {
            m_source.expectBlocks( m_index.dataBlocksIn( m_currentPosition, nBytesToRead ) );
        }

        size_t nBytesDecoded = 0;
        while ( ( nBytesDecoded < nBytesToRead ) && !eof() ) {
            throwDeferredStreamError();
            BlockSource::Block block;

            auto span = m_index.locate( m_currentPosition );
            if ( !span.contains( m_currentPosition ) ) {
                /* a block that has not been entered into the block map yet (ParallelBZ2Reader.hpp:191-244) */
                const auto dataBlockIndex = m_index.dataBlockCount();
                if ( dataBlockIndex >= m_source.finderCount() ) {
                    m_index.finish();
                    m_atEndOfFile = true;
                    break;
                }
                /* the stream CRC is folded over the blocks in file order, exactly once each (BZ2Reader.hpp:481-484) */
                const auto encodedOffsetInBits = m_source.finderOffsets()[dataBlockIndex];
                block = m_source.fetch( encodedOffsetInBits, dataBlockIndex );
                m_index.append( block.result->enc_offset_bits, block.result->enc_size_bits, block.result->out_size );
                m_calculatedStreamCRC = ibz2_stream_crc_fold( m_calculatedStreamCRC, block.result->calculated_crc );
                probeAfter( *block.result );

                span = m_index.locate( m_currentPosition );
                if ( !span.contains( m_currentPosition ) ) {
                    continue;
                }
            } else {
                block = m_source.fetch( span.encodedOffsetInBits );
            }
            const size_t blockSize = block.result->out_size;

            const auto offsetInBlock = m_currentPosition - span.decodedOffsetInBytes;
            if ( offsetInBlock >= blockSize ) {
                throw std::logic_error( "Block does not contain the requested offset even though it "
                                        "shouldn't be according to block map!" );
            }
            const auto nBytesToDecode = std::min( blockSize - offsetInBlock, nBytesToRead - nBytesDecoded );
            if ( writeFunctor ) {
                /* the block lives in HBM: what this read wants of it comes over in pieces */
                m_source.deliver( block, offsetInBlock, nBytesToDecode, [&writeFunctor] ( const uint8_t* bytes, size_t n ) {
                    writeFunctor( bytes, n );
                } );
            }
            nBytesDecoded += nBytesToDecode;
            m_currentPosition += nBytesToDecode;
        }
        if ( eof() ) {
            throwDeferredStreamError();
        }
        return nBytesDecoded;
    }

    size_t
    seek( long long int offset, int origin = SEEK_SET ) override
    {
        if ( closed() ) {
            throw std::invalid_argument( "You may not call seek on closed ParallelBZ2Reader!" );
        }
        if ( ( origin == SEEK_END ) && !m_index.finished() ) {
            read();   /* size() needs the complete block map */
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
        m_source.noteSeek();   /* random access: start again with small batches */

        if ( positiveOffset < tell() ) {
            m_atEndOfFile = false;
            m_currentPosition = positiveOffset;
            return positiveOffset;
        }

        const auto span = m_index.locate( positiveOffset );
        if ( span.valid && ( positiveOffset < span.decodedOffsetInBytes ) ) {
            throw std::logic_error( "Block map returned unwanted block!" );
        }
        if ( span.contains( positiveOffset ) ) {
            m_atEndOfFile = false;
            m_currentPosition = positiveOffset;
            return tell();
        }
        if ( m_index.finished() ) {
            m_atEndOfFile = true;
            m_currentPosition = m_index.back().second;
            return tell();
        }

        /* beyond what is known: decode forward without a sink (ParallelBZ2Reader.hpp:318-324) */
        m_atEndOfFile = false;
        m_currentPosition = span.decodedOffsetInBytes + span.decodedSizeInBytes;
        read( -1, nullptr, positiveOffset - tell() );
        return tell();
    }

    /* BZip2 specific methods */

    [[nodiscard]] bool blockOffsetsComplete() const override { return m_index.finished(); }

    [[nodiscard]] std::map<size_t, size_t>
    blockOffsets() override
    {
        if ( !m_index.finished() ) {
            read();
            if ( !m_index.finished() ) {
                throw std::logic_error( "Reading everything should have finalized the block map!" );
            }
        }
        return m_index.offsets();
    }

    [[nodiscard]] std::map<size_t, size_t>
    availableBlockOffsets() const override
    {
        return m_index.offsets();
    }

    void
    setBlockOffsets( std::map<size_t, size_t> offsets ) override
    {
        if ( offsets.empty() ) {
            throw std::invalid_argument( "May not clear offsets. Construct a new ParallelBZ2Reader instead!" );
        }
        /* entries without data are end-of-stream blocks, the last one always is (ParallelBZ2Reader.hpp:463-472) */
        std::vector<uint64_t> dataBlocks;
        for ( auto it = offsets.begin(), nit = std::next( offsets.begin() ); nit != offsets.end(); ++it, ++nit ) {
            if ( it->second != nit->second ) {
                dataBlocks.push_back( it->first );
            }
        }
        m_source.setFinderOffsets( std::move( dataBlocks ) );
        if ( offsets.size() < 2 ) {
            throw std::invalid_argument( "Block offset map must contain at least one valid block and one EOS block!" );
        }
        m_index.assign( offsets );
    }

    [[nodiscard]] size_t
    tellCompressed() const override
    {
        const auto span = m_index.locate( m_currentPosition );
        if ( span.contains( m_currentPosition ) ) {
            return span.encodedOffsetInBits;
        }
        return m_index.back().first;
    }

    /** The reference stops its thread pools here; this reader has none, so only the decoded batch is dropped. */
    void
    joinThreads()
    {
        m_source.dropBatch();
    }

    /**
     * New next to the reference's bzip2 readers (its gzip reader has them, ParallelGzipReader.hpp:993-1029): the block
     * map as the text `ibzip2 -L` writes, one "bits,bytes" pair per line, or as a GZIDX file (IndexFileFormat.hpp).
     */
    void
    exportIndex( std::ostream& out, IndexFormat format = IndexFormat::OFFSET_LINES )
    {
        const auto offsets = blockOffsets();
        if ( format == IndexFormat::GZIDX ) {
            writeGzidx( out, offsets, m_source.fileSize() );
            return;
        }
        for ( const auto& [bits, bytes] : offsets ) {
            out << bits << "," << bytes << "\n";
        }
    }

    /** Either format, told apart by the GZIDX magic. */
    void
    importIndex( std::istream& in )
    {
        std::map<size_t, size_t> offsets;
        if ( in.peek() == 'G' ) {
            offsets = readGzidx( in );
            /* rapidgzip's checkpoints for bzip2 are chunk starts: some block boundaries, not all.  This reader decodes
             * block by block, so every data block needs its entry (as in the files exportIndex writes). */
            const auto& found = m_source.finderOffsets();
            size_t listed = 0;
            for ( const auto offset : found ) {
                listed += offsets.count( offset );
            }
            if ( listed != found.size() ) {
                throw std::invalid_argument( "The index does not list every bzip2 block of the file (a chunk-level index "
                                             "as rapidgzip writes it?); only block-level indexes can be imported." );
            }
        } else {
            std::string line;
            while ( std::getline( in, line ) ) {
                if ( line.empty() ) {
                    continue;
                }
                const auto comma = line.find( ',' );
                if ( comma == std::string::npos ) {
                    throw std::invalid_argument( "Index lines must look like <bits>,<bytes>" );
                }
                offsets.emplace( std::stoull( line.substr( 0, comma ) ), std::stoull( line.substr( comma + 1 ) ) );
            }
        }
        setBlockOffsets( std::move( offsets ) );
    }

    /**
     * New next to the reference, whose parallel reader never looks at the stream CRC (only BZ2Reader does,
     * BZ2Reader.hpp:406-416): while the block map is being built, the CRCs of the blocks are folded per stream and
     * compared with the CRC field of the end-of-stream block.  A mismatch throws what BZ2Reader throws, after the bytes
     * of the stream's last block have been delivered.  On by default; switch it off for the reference's exact behaviour.
     */
    void
    setVerifyStreamCRC( bool verify )
    {
        m_verifyStreamCRC = verify;
    }

    [[nodiscard]] ibz2_timings
    lastTimings()
    {
        return m_source.gpu().timings();
    }

private:
    static UniqueFileReader
    requireSeekable( UniqueFileReader fileReader )
    {
        if ( !fileReader ) {
            throw std::invalid_argument( "File reader must not be null!" );
        }
        if ( !fileReader->seekable() ) {
            /* ensureSharedFileReader (core/filereader/Shared.hpp:471-473) puts a SinglePassFileReader, which keeps
             * everything it has read, in front of a pipe; the GPU path wants the whole file anyway, so read it now. */
            std::vector<uint8_t> bytes;
            std::vector<char> chunk( 4U << 20 );
            while ( true ) {
                const size_t n = fileReader->read( chunk.data(), chunk.size() );
                if ( n == 0 ) {
                    break;
                }
                bytes.insert( bytes.end(), chunk.begin(), chunk.begin() + static_cast<std::ptrdiff_t>( n ) );
            }
            fileReader = std::make_unique<MemoryFileReader>( std::move( bytes ) );
        }
        if ( !fileReader->seekable() ) {
            throw std::invalid_argument( "Parallel BZ2 Reader will not work on non-seekable input like stdin (yet)!" );
        }
        return fileReader;
    }

    /** ParallelBZ2Reader.hpp:203-238: an end-of-stream block after a data block gets its own entry. */
    void
    probeAfter( const ibz2_block_result& block )
    {
        if ( block.is_eof != 0 ) {
            return;
        }
        const auto next = m_source.probeHeader( block.enc_offset_bits + block.enc_size_bits );
        if ( next.status == IBZ2_ST_EOF ) {
            throw EndOfFileReached();
        }
        if ( next.status != IBZ2_ST_OK ) {
            throwForStatus( next );
        }
        if ( next.is_eos == 0 ) {
            return;
        }
        if ( m_verifyStreamCRC && ( next.expected_crc != m_calculatedStreamCRC ) && m_deferredStreamError.empty() ) {
            std::stringstream msg;
            msg << "[BZip2 block header] Stream CRC 0x" << std::hex << next.expected_crc
                << " does not match calculated CRC 0x" << m_calculatedStreamCRC;
            m_deferredStreamError = std::move( msg ).str();
        }
        m_calculatedStreamCRC = 0;
        m_index.append( next.enc_offset_bits, next.enc_size_bits, 0 );
        if ( !m_deferredStreamError.empty() && !m_deferredStreamErrorArmed ) {
            m_deferredStreamErrorArmed = true;
            m_deferredStreamErrorAt = m_index.back().second;   /* where the stream's decoded bytes end */
        }
        const auto nextStreamOffsetInBits = next.enc_offset_bits + next.enc_size_bits;
        if ( nextStreamOffsetInBits < m_source.fileSizeInBits() ) {
            uint64_t position = nextStreamOffsetInBits;
            if ( ibz2_read_stream_header( m_source.fileData(), m_source.fileSize(), &position ) <= 0 ) {
                std::cerr << "[Warning] Trailing garbage after EOF ignored!\n";
                /* valid blocks behind a gap are not read any more: truncate the finder's list */
                m_source.truncateFinder( m_index.dataBlockCount() );
            }
        }
    }

    /** thrown where BZ2Reader would throw: once the bytes in front of the end-of-stream block have been handed out */
    void
    throwDeferredStreamError()
    {
        if ( m_deferredStreamErrorArmed && ( m_currentPosition >= m_deferredStreamErrorAt ) ) {
            const auto message = std::move( m_deferredStreamError );
            m_deferredStreamError.clear();
            m_deferredStreamErrorArmed = false;
            throw std::runtime_error( message );
        }
    }

    BlockSource m_source;
    BlockIndex m_index;
    bool m_verifyStreamCRC{ true };
    uint32_t m_calculatedStreamCRC{ 0 };
    std::string m_deferredStreamError;
    bool m_deferredStreamErrorArmed{ false };
    size_t m_deferredStreamErrorAt{ 0 };

    size_t m_currentPosition{ 0 };
    bool m_atEndOfFile{ false };
    bool m_closed{ false };
};
}  // namespace ibz2host
