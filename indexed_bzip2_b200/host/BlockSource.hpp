/*
 * BlockSource -- "give me the decoded block that starts at this bit offset" for both readers: the compressed file
 * (host copy for header probes, device copy for decoding), the finder's list of data-block offsets and the batch of
 * blocks decoded last.  It replaces the reference's BlockFetcher + cache + prefetcher
 * (core/BlockFetcher.hpp:244-317, core/Cache.hpp, core/Prefetcher.hpp) with the policy a GPU wants:
 * a miss decodes the wanted block AND the blocks behind it in one call; consecutive misses in file order double the
 * batch up to MAX_BATCH_BLOCKS, any other access pattern starts again from MIN_BATCH_BLOCKS.
 */
#pragma once

#include <algorithm>
#include <limits>
#include <memory>
#include <optional>
#include <unordered_map>
#include <vector>

#include "FileReader.hpp"
#include "GpuDecoder.hpp"

namespace ibz2host
{
class BlockSource
{
public:
    static constexpr size_t MIN_BATCH_BLOCKS = 4;
    static constexpr size_t MAX_BATCH_BLOCKS = 2048;

    BlockSource( UniqueFileReader fileReader, size_t parallelization, int device ) :
        m_parallelization( parallelization ),
        m_device( device )
    {
        if ( !fileReader ) {
            throw std::invalid_argument( "A file reader is required!" );
        }
        if ( auto* memory = dynamic_cast<MemoryFileReader*>( fileReader.get() ); memory != nullptr ) {
            m_file = std::move( memory->bytes() );
            return;
        }
        if ( fileReader->seekable() ) {
            m_file.resize( fileReader->size().value_or( 0 ) );
            fileReader->seek( 0 );
        }
        /* non-seekable input (stdin) is read to its end in growing steps */
        size_t done = 0;
        while ( true ) {
            if ( done == m_file.size() ) {
                if ( fileReader->seekable() ) {
                    break;
                }
                m_file.resize( std::max<size_t>( 1U << 20, m_file.size() * 2 ) );
            }
            const size_t n = fileReader->read( reinterpret_cast<char*>( m_file.data() ) + done, m_file.size() - done );
            if ( n == 0 ) {
                break;
            }
            done += n;
        }
        m_file.resize( done );
    }

    [[nodiscard]] const uint8_t* fileData() const { return m_file.data(); }
    [[nodiscard]] size_t fileSize() const { return m_file.size(); }
    [[nodiscard]] size_t fileSizeInBits() const { return m_file.size() * 8U; }

    /** rapidgzip::BlockFinder: every data-block magic of the file, ascending, found once by K1. */
    const std::vector<uint64_t>&
    finderOffsets()
    {
        if ( !m_finderReady ) {
            m_finderOffsets = gpu().findBlocks();
            m_finderReady = true;
        }
        return m_finderOffsets;
    }

    /** Number of finder offsets the readers may still use (BlockFinder::finalize( n ) truncates the list). */
    [[nodiscard]] size_t
    finderCount()
    {
        return std::min( finderOffsets().size(), m_finderLimit );
    }

    void
    truncateFinder( size_t count )
    {
        m_finderLimit = count;
    }

    /** BlockFinder::setBlockOffsets: the finder becomes a static list (imported index). */
    void
    setFinderOffsets( std::vector<uint64_t> offsets )
    {
        m_finderOffsets = std::move( offsets );
        m_finderReady = true;
        m_finderLimit = std::numeric_limits<size_t>::max();
        dropBatch();
    }

    void
    dropBatch()
    {
        m_results.clear();
        m_byOffset.clear();
        m_data = nullptr;
    }

    void
    noteSeek()
    {
        m_sequentialFetches = 0;
    }

    struct Block
    {
        const ibz2_block_result* result{ nullptr };
        const uint8_t* data{ nullptr };
    };

    /**
     * The decoded block at @p encodedOffset; throws what the reference's decoder would have thrown for it.
     * @p finderIndex is the block's index in the finder list when the caller knows it.  An offset the finder does
     * not list (possible for the serial reader on odd files) is decoded on its own.
     */
    Block
    fetch( uint64_t encodedOffset, std::optional<size_t> finderIndex = std::nullopt )
    {
        if ( const auto cached = m_byOffset.find( encodedOffset ); cached != m_byOffset.end() ) {
            return checked( cached->second );
        }

        const auto& offsets = finderOffsets();
        std::vector<uint64_t> single;
        const uint64_t* batchOffsets = nullptr;
        size_t count = 1;
        size_t first = finderIndex.value_or( 0 );
        bool listed = finderIndex.has_value();
        if ( !listed ) {
            const auto match = std::lower_bound( offsets.begin(), offsets.end(), encodedOffset );
            listed = ( match != offsets.end() ) && ( *match == encodedOffset );
            first = static_cast<size_t>( std::distance( offsets.begin(), match ) );
        }
        if ( listed ) {
            const bool sequential = first == m_nextSequentialBlock;
            m_sequentialFetches = sequential ? m_sequentialFetches + 1 : 0;
            const size_t floor = std::max<size_t>( MIN_BATCH_BLOCKS, std::min<size_t>( m_parallelization, 64 ) );
            const size_t wanted = std::min<size_t>( MAX_BATCH_BLOCKS, floor << std::min<size_t>( m_sequentialFetches, 12 ) );
            const size_t limit = finderCount();
            count = std::min( wanted, limit > first ? limit - first : 1 );
            m_nextSequentialBlock = first + count;
            batchOffsets = offsets.data() + first;
        } else {
            single.push_back( encodedOffset );
            batchOffsets = single.data();
        }

        dropBatch();
        uint8_t* data = nullptr;
        size_t dataSize = 0;
        gpu().decode( batchOffsets, count, m_results, data, dataSize );
        m_data = data;
        for ( size_t i = 0; i < m_results.size(); ++i ) {
            m_byOffset.emplace( m_results[i].enc_offset_bits, i );
        }
        return checked( m_byOffset.at( encodedOffset ) );
    }

    /** BZ2BlockFetcher::readBlockHeader (BZ2BlockFetcher.hpp:64-82): magic, CRC and EOS padding only. */
    [[nodiscard]] ibz2_block_result
    probeHeader( uint64_t bitOffset ) const
    {
        return GpuDecoder::probeHeader( m_file.data(), m_file.size(), bitOffset );
    }

    GpuDecoder&
    gpu()
    {
        if ( !m_gpu ) {
            m_gpu = std::make_unique<GpuDecoder>( m_device );
            m_gpu->upload( m_file.data(), m_file.size() );
        }
        return *m_gpu;
    }

    void
    close()
    {
        dropBatch();
        m_gpu.reset();
    }

private:
    /** A block that failed to decode throws where the reference's worker exception would surface: on access. */
    Block
    checked( size_t i ) const
    {
        const auto& r = m_results[i];
        if ( ( r.status == IBZ2_ST_OK ) && ( r.is_eos == 0 ) && ( r.calculated_crc != r.expected_crc ) ) {
            ibz2_block_result failed = r;
            failed.status = IBZ2_ST_CRC;
            throwForStatus( failed );
        }
        if ( r.status != IBZ2_ST_OK ) {
            throwForStatus( r );
        }
        return Block{ &r, m_data + r.out_offset };
    }

    std::vector<uint8_t> m_file;
    const size_t m_parallelization;
    const int m_device;
    std::unique_ptr<GpuDecoder> m_gpu;

    std::vector<uint64_t> m_finderOffsets;
    bool m_finderReady{ false };
    size_t m_finderLimit{ std::numeric_limits<size_t>::max() };

    std::vector<ibz2_block_result> m_results;
    std::unordered_map<uint64_t, size_t> m_byOffset;
    const uint8_t* m_data{ nullptr };
    size_t m_nextSequentialBlock{ 0 };
    size_t m_sequentialFetches{ 0 };
};
}  // namespace ibz2host
