/*
 * BlockSource -- "give me the decoded block that starts at this bit offset" for both readers: the compressed file
 * (memory-mapped on the host for the header probes, in HBM for decoding), the finder's list of data-block offsets and
 * a cache of decoded batches.  It stands where the reference has BlockFetcher + Cache + Prefetcher
 * (core/BlockFetcher.hpp:244-317 and :446-559, core/Cache.hpp:122-297, core/Prefetcher.hpp:82-223), with the policy a
 * GPU wants:
 *   - the unit of work is a BATCH of consecutive blocks decoded by one ibz2_decode_blocks call (the GPU needs thousands
 *     of blocks in flight; one block alone costs as long as its sequential Huffman chain);
 *   - prefetching = batch size.  Like FetchNextAdaptive (Prefetcher.hpp:159-176) nothing is prefetched on random
 *     access and more and more on consecutive access: every miss that continues where the last batch ended decodes
 *     eight times as many blocks as the one before, up to MAX_BATCH_BLOCKS; any other access starts small again.  A
 *     caller that knows how far its read reaches (block map complete) says so, and the blocks of ONE read are decoded
 *     by ONE call.  Once the batches are large, the next one is decoded by a second host thread while the caller
 *     consumes the current one;
 *   - decoded batches stay in a least-recently-used cache (the reference keeps max(16, P) blocks + 2 P prefetched
 *     ones, BlockFetcher.hpp:180-182; here up to MAX_CACHED_BATCHES batches inside a byte budget) IN HBM, where they
 *     were decoded; the buffers of evicted batches are reused.  Only the bytes a read asks for cross PCIe, through two
 *     small pinned staging buffers (GpuDecoder::deliver): page-locking host memory for whole batches cost more than
 *     decoding them (0.72 s per 2 GB against 0.125 s).
 * The file is not read into a private copy: paths and descriptors are memory-mapped, and the mapping goes to the GPU
 * through pinned staging buffers (ibz2_upload_pageable).  It must fit in HBM next to the decode scratch.
 */
#pragma once

#include <algorithm>
#include <cstdlib>
#include <limits>
#include <list>
#include <memory>
#include <exception>
#include <optional>
#include <thread>
#include <unordered_map>
#include <vector>

#include <sys/mman.h>

#include "FileReader.hpp"
#include "GpuDecoder.hpp"

namespace ibz2host
{
class BlockSource
{
public:
    static constexpr size_t MIN_BATCH_BLOCKS = 4;
    static constexpr size_t MAX_BATCH_BLOCKS = 2048;      /* 1.8 GB of decoded bytes at 900 kB blocks */
    static constexpr size_t BATCH_GROWTH_SHIFT = 3;       /* consecutive misses: 4, 32, 256, 2048 blocks */
    static constexpr size_t MAX_CACHED_BATCHES = 16;
    static constexpr size_t MIN_BLOCKS_PER_DEVICE = 64;   /* below this a batch stays on the first device */
    static constexpr size_t LARGE_BATCH_BYTES = 256ULL << 20;
    static constexpr size_t MAX_LARGE_BATCHES = 4;
    static constexpr size_t DEFAULT_CACHE_BYTES = 12ULL << 30;   /* of HBM, per device; $IBZ2_CACHE_BYTES */

    /** @p devices: the CUDA devices that share the decoding of every large batch (one entry: the usual case; an entry
     *  may repeat, which gives one GPU two contexts -- how the tests exercise the multi-device path on one GPU). */
    BlockSource( UniqueFileReader fileReader, size_t parallelization, std::vector<int> devices ) :
        m_parallelization( parallelization ),
        m_devices( devices.empty() ? std::vector<int>{ 0 } : std::move( devices ) ),
        m_gpus( m_devices.size() )
    {
        if ( !fileReader ) {
            throw std::invalid_argument( "A file reader is required!" );
        }
        if ( const char* const env = std::getenv( "IBZ2_CACHE_BYTES" ); env != nullptr ) {
            m_cacheBudget = std::max<size_t>( 1, std::strtoull( env, nullptr, 10 ) );
        }
        if ( const char* const env = std::getenv( "IBZ2_PREFETCH_MIN_BLOCKS" ); env != nullptr ) {
            m_prefetchMinBlocks = std::strtoull( env, nullptr, 10 );
        }
        if ( const char* const env = std::getenv( "IBZ2_MIN_BLOCKS_PER_DEVICE" ); env != nullptr ) {
            m_minBlocksPerDevice = std::max<size_t>( 1, std::strtoull( env, nullptr, 10 ) );
        }
        if ( auto* memory = dynamic_cast<MemoryFileReader*>( fileReader.get() ); memory != nullptr ) {
            m_owned = std::move( memory->bytes() );
            m_data = m_owned.data();
            m_size = m_owned.size();
            return;
        }
        if ( auto* standard = dynamic_cast<StandardFileReader*>( fileReader.get() );
             ( standard != nullptr ) && standard->seekable() ) {
            /* a regular file: mapped, not copied (the reference shares one descriptor between its threads instead,
             * core/filereader/Shared.hpp:267-300) */
            m_size = standard->size().value_or( 0 );
            if ( m_size == 0 ) {
                return;
            }
            void* const mapping = ::mmap( nullptr, m_size, PROT_READ, MAP_SHARED, standard->fileno(), 0 );
            if ( mapping != MAP_FAILED ) {
                m_mapping = mapping;
                m_data = static_cast<const uint8_t*>( mapping );
                ::madvise( mapping, m_size, MADV_SEQUENTIAL );
                return;
            }
            m_size = 0;   /* fall through: read it */
        }
        if ( fileReader->seekable() ) {
            m_owned.resize( fileReader->size().value_or( 0 ) );
            fileReader->seek( 0 );
        }
        /* anything else (a pipe) is read to its end in growing steps */
        size_t done = 0;
        while ( true ) {
            if ( done == m_owned.size() ) {
                if ( fileReader->seekable() ) {
                    break;
                }
                m_owned.resize( std::max<size_t>( 1U << 20, m_owned.size() * 2 ) );
            }
            const size_t n = fileReader->read( reinterpret_cast<char*>( m_owned.data() ) + done, m_owned.size() - done );
            if ( n == 0 ) {
                break;
            }
            done += n;
        }
        m_owned.resize( done );
        m_data = m_owned.data();
        m_size = m_owned.size();
    }

    ~BlockSource()
    {
        close();
        if ( m_mapping != nullptr ) {
            ::munmap( m_mapping, m_size );
        }
    }

    BlockSource( const BlockSource& ) = delete;
    BlockSource& operator=( const BlockSource& ) = delete;

    [[nodiscard]] const uint8_t* fileData() const { return m_data; }
    [[nodiscard]] size_t fileSize() const { return m_size; }
    [[nodiscard]] size_t fileSizeInBits() const { return m_size * 8U; }

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
        joinPrefetch( false );   /* it reads the list that is about to be replaced */
        m_finderOffsets = std::move( offsets );
        m_finderReady = true;
        m_finderLimit = std::numeric_limits<size_t>::max();
        dropBatch();
    }

    /** Forgets every decoded block (their pinned buffers stay for reuse). */
    void
    dropBatch()
    {
        joinPrefetch( false );
        while ( !m_batches.empty() ) {
            evictOldest();
        }
    }

    void
    noteSeek()
    {
        m_sequentialFetches = 0;
    }

    /** The caller's next fetches will ask for this many consecutive data blocks (0: unknown). */
    void
    expectBlocks( size_t count )
    {
        m_expected = count;
    }

    struct Block
    {
        const ibz2_block_result* result{ nullptr };
        const uint8_t* deviceData{ nullptr };   /* the block's decoded bytes, in the memory of device `decoder` */
        size_t decoder{ 0 };
    };

    /** Bytes [offset, offset + size) of a fetched block to @p sink( const uint8_t*, size_t ), in order, in pieces. */
    template<typename Sink>
    void
    deliver( const Block& block, size_t offset, size_t size, const Sink& sink )
    {
        m_gpus[block.decoder]->deliver( block.deviceData + offset, size, sink );
    }

    /**
     * The decoded block at @p encodedOffset; throws what the reference's decoder would have thrown for it.
     * @p finderIndex is the block's index in the finder list when the caller knows it.  An offset the finder does
     * not list (possible for the serial reader on odd files) is decoded on its own.
     * The returned pointers stay valid until the next fetch.
     */
    Block
    fetch( uint64_t encodedOffset, std::optional<size_t> finderIndex = std::nullopt )
    {
        if ( const auto cached = m_byOffset.find( encodedOffset ); cached != m_byOffset.end() ) {
            /* Cache::get moves a hit to the young end (core/Cache.hpp:160-178) */
            m_batches.splice( m_batches.begin(), m_batches, cached->second.first );
            ++m_hits;
            return checked( *cached->second.first, cached->second.second );
        }
        /* the batch a second thread has been decoding while the caller worked on the last one (Prefetcher.hpp: the
         * prefetch cache is looked at before anything is decoded on demand, BlockFetcher.hpp:262-280) */
        if ( joinPrefetch( true ) ) {
            if ( const auto cached = m_byOffset.find( encodedOffset ); cached != m_byOffset.end() ) {
                m_batches.splice( m_batches.begin(), m_batches, cached->second.first );
                ++m_prefetchHits;
                startPrefetch();
                return checked( *cached->second.first, cached->second.second );
            }
        }
        ++m_misses;

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
            const size_t grown = floor << std::min<size_t>( BATCH_GROWTH_SHIFT * m_sequentialFetches, 20 );
            /* a read whose extent is known decodes what it needs in one call, and no more than that on random access */
            const size_t wanted = ( m_expected > 0 ) && !sequential ? m_expected : std::max( grown, m_expected );
            const size_t limit = finderCount();
            /* a file of a few thousand blocks is not worth the largest buffers: an eighth of the file per batch */
            const size_t largest = std::min<size_t>( MAX_BATCH_BLOCKS, std::max<size_t>( 256, roundUpToPowerOfTwo( limit / 8U ) ) );
            count = std::min( std::min<size_t>( wanted, largest ), limit > first ? limit - first : 1 );
            /* stop in front of a block that is cached already */
            for ( size_t i = 1; i < count; ++i ) {
                if ( m_byOffset.count( offsets[first + i] ) != 0 ) {
                    count = i;
                    break;
                }
            }
            m_nextSequentialBlock = first + count;
            batchOffsets = offsets.data() + first;
        } else {
            single.push_back( encodedOffset );
            batchOffsets = single.data();
        }
        m_expected = 0;

        /* Several devices: each takes a contiguous part of the batch (blocks are independent: no collective, the parts
         * land in separate pinned buffers and enter the cache as batches of their own), decoded by one host thread per
         * device like the worker threads of the reference's pool (core/BlockFetcher.hpp:176-190).  Small batches are
         * not worth it: a decode call costs one block's Huffman chain however few blocks it holds. */
        const size_t parts = ( listed && ( count >= m_minBlocksPerDevice * m_gpus.size() ) ) ? m_gpus.size() : 1;
        /* (a prefetch, which only exists with one device, has been joined above: the contexts are this thread's) */
        std::vector<Batch> decoded( parts );
        if ( parts == 1 ) {
            decoded[0].buffer = takeBuffer( 0, count * 1000000ULL + 65536ULL );
            gpu().decode( batchOffsets, count, decoded[0].results, decoded[0].buffer );
        } else {
            std::vector<std::thread> workers;
            std::vector<std::exception_ptr> failures( parts );
            for ( size_t d = 0; d < parts; ++d ) {
                const size_t begin = count * d / parts;
                const size_t end = count * ( d + 1 ) / parts;
                gpu( d );   /* contexts and uploads are created here, one after the other */
                decoded[d].decoder = d;
                decoded[d].buffer = takeBuffer( d, ( end - begin ) * 1000000ULL + 65536ULL );
                workers.emplace_back( [this, d, begin, end, batchOffsets, &decoded, &failures] () {
                    try {
                        m_gpus[d]->decode( batchOffsets + begin, end - begin, decoded[d].results, decoded[d].buffer );
                    } catch ( ... ) {
                        failures[d] = std::current_exception();
                    }
                } );
            }
            for ( auto& worker : workers ) {
                worker.join();
            }
            for ( const auto& failure : failures ) {
                if ( failure ) {
                    std::rethrow_exception( failure );
                }
            }
        }
        /* youngest last: the part that holds the wanted block ends up at the front */
        for ( size_t d = parts; d-- > 0; ) {
            decoded[d].bytes = decoded[d].buffer.capacity;
            m_batches.push_front( std::move( decoded[d] ) );
            const auto inserted = m_batches.begin();
            for ( size_t i = 0; i < inserted->results.size(); ++i ) {
                m_byOffset[inserted->results[i].enc_offset_bits] = { inserted, i };
            }
            m_cachedBytes += inserted->bytes;
        }
        const auto wantedEntry = m_byOffset.at( encodedOffset );
        m_batches.splice( m_batches.begin(), m_batches, wantedEntry.first );
        const auto inserted = wantedEntry.first;
        while ( ( m_batches.size() > parts )
                && ( ( m_batches.size() > MAX_CACHED_BATCHES ) || ( m_cachedBytes > m_cacheBudget )
                     || ( largeBatches() > MAX_LARGE_BATCHES * parts ) ) ) {
            evictOldest();
        }
        if ( listed && ( parts == 1 ) && ( m_sequentialFetches > 0 ) ) {
            startPrefetch();
        }
        return checked( *inserted, wantedEntry.second );
    }

    /** BZ2BlockFetcher::readBlockHeader (BZ2BlockFetcher.hpp:64-82): magic, CRC and EOS padding only. */
    [[nodiscard]] ibz2_block_result
    probeHeader( uint64_t bitOffset ) const
    {
        return GpuDecoder::probeHeader( m_data, m_size, bitOffset );
    }

    /** The decoder of device @p index of the set (created, and given the file, on first use). */
    GpuDecoder&
    gpu( size_t index = 0 )
    {
        if ( !m_gpus[index] ) {
            m_gpus[index] = std::make_unique<GpuDecoder>( m_devices[index] );
            m_gpus[index]->upload( m_data, m_size );
        }
        return *m_gpus[index];
    }

    [[nodiscard]] size_t
    deviceCount() const
    {
        return m_gpus.size();
    }

    void
    close()
    {
        dropBatch();
        for ( auto& spare : m_freeBuffers ) {
            if ( m_gpus[spare.first] ) {
                m_gpus[spare.first]->freeDevice( spare.second );
            }
        }
        m_freeBuffers.clear();
        for ( auto& decoder : m_gpus ) {
            decoder.reset();
        }
    }

    /** cache statistics like BlockFetcher::statistics() (core/BlockFetcher.hpp:62-126): hits, misses */
    [[nodiscard]] std::pair<size_t, size_t>
    cacheStatistics() const
    {
        return { m_hits, m_misses };
    }

private:
    struct Batch
    {
        std::vector<ibz2_block_result> results;
        GpuDecoder::DeviceBuffer buffer;   /* in the memory of device `decoder` */
        size_t decoder{ 0 };
        size_t bytes{ 0 };
    };
    using BatchList = std::list<Batch>;

    /** A block that failed to decode throws where the reference's worker exception would surface: on access. */
    static Block
    checked( const Batch& batch, size_t i )
    {
        const auto& r = batch.results[i];
        if ( ( r.status == IBZ2_ST_OK ) && ( r.is_eos == 0 ) && ( r.calculated_crc != r.expected_crc ) ) {
            ibz2_block_result failed = r;
            failed.status = IBZ2_ST_CRC;
            throwForStatus( failed );
        }
        if ( r.status != IBZ2_ST_OK ) {
            throwForStatus( r );
        }
        return Block{ &r, static_cast<const uint8_t*>( batch.buffer.data ) + r.out_offset, batch.decoder };
    }

    /**
     * Consecutive access with large batches: the batch behind the last one is decoded by a second host thread while the
     * caller consumes the current one (its bytes leave through the copy stream and the staging buffers, which the decode
     * does not touch).  One device only; like the reference's prefetches it is dropped when the access pattern changes.
     */
    void
    startPrefetch()
    {
        if ( ( m_gpus.size() != 1 ) || m_prefetchPending || ( m_prefetchMinBlocks == 0 ) ) {
            return;
        }
        const auto& offsets = finderOffsets();
        const size_t limit = finderCount();
        const size_t first = m_nextSequentialBlock;
        if ( first >= limit ) {
            return;
        }
        const size_t largest = std::min<size_t>( MAX_BATCH_BLOCKS, std::max<size_t>( 256, roundUpToPowerOfTwo( limit / 8U ) ) );
        size_t count = std::min( largest, limit - first );
        for ( size_t i = 0; i < count; ++i ) {
            if ( m_byOffset.count( offsets[first + i] ) != 0 ) {
                count = i;
                break;
            }
        }
        if ( count < std::min( m_prefetchMinBlocks, limit - first ) ) {
            return;
        }
        m_prefetched = Batch{};
        m_prefetched.buffer = takeBuffer( 0, count * 1000000ULL + 65536ULL );
        m_prefetchFailure = nullptr;
        m_prefetchFirst = first;
        m_prefetchCount = count;
        m_prefetchPending = true;
        GpuDecoder* const decoder = &gpu();
        const uint64_t* const batchOffsets = offsets.data() + first;
        m_prefetchThread = std::thread( [this, decoder, batchOffsets, count] () {
            try {
                decoder->decode( batchOffsets, count, m_prefetched.results, m_prefetched.buffer );
            } catch ( ... ) {
                m_prefetchFailure = std::current_exception();
            }
        } );
    }

    /** Waits for a running prefetch; @p keep: its batch enters the cache.  @return whether one did. */
    bool
    joinPrefetch( bool keep )
    {
        if ( !m_prefetchPending ) {
            return false;
        }
        m_prefetchThread.join();
        m_prefetchPending = false;
        if ( !keep || m_prefetchFailure ) {
            /* a failed prefetch is simply forgotten: the blocks are decoded again when they are asked for, and then the
             * failure surfaces where the reference's would */
            if ( m_prefetched.buffer.data != nullptr ) {
                m_freeBuffers.emplace_back( 0, m_prefetched.buffer );
            }
            m_prefetched = Batch{};
            return false;
        }
        m_prefetched.bytes = m_prefetched.buffer.capacity;
        m_batches.push_front( std::move( m_prefetched ) );
        m_prefetched = Batch{};
        const auto inserted = m_batches.begin();
        for ( size_t i = 0; i < inserted->results.size(); ++i ) {
            m_byOffset[inserted->results[i].enc_offset_bits] = { inserted, i };
        }
        m_cachedBytes += inserted->bytes;
        m_nextSequentialBlock = m_prefetchFirst + m_prefetchCount;
        while ( ( m_batches.size() > 1 )
                && ( ( m_batches.size() > MAX_CACHED_BATCHES ) || ( m_cachedBytes > m_cacheBudget )
                     || ( largeBatches() > MAX_LARGE_BATCHES ) ) ) {
            evictOldest();
        }
        return true;
    }

    [[nodiscard]] static size_t
    roundUpToPowerOfTwo( size_t value )
    {
        size_t result = 1;
        while ( result < value ) {
            result <<= 1;
        }
        return result;
    }

    [[nodiscard]] size_t
    largeBatches() const
    {
        return static_cast<size_t>( std::count_if( m_batches.begin(), m_batches.end(),
                                                   [] ( const Batch& b ) { return b.bytes >= LARGE_BATCH_BYTES; } ) );
    }

    void
    evictOldest()
    {
        auto& oldest = m_batches.back();
        for ( const auto& r : oldest.results ) {
            const auto entry = m_byOffset.find( r.enc_offset_bits );
            if ( ( entry != m_byOffset.end() ) && ( &*entry->second.first == &oldest ) ) {
                m_byOffset.erase( entry );
            }
        }
        m_cachedBytes -= std::min( m_cachedBytes, oldest.bytes );
        if ( oldest.buffer.data != nullptr ) {
            m_freeBuffers.emplace_back( oldest.decoder, oldest.buffer );
        }
        m_batches.pop_back();
        /* a few spare buffers are worth keeping, the largest ones first */
        if ( m_freeBuffers.size() > 4 * m_gpus.size() ) {
            const auto smallest = std::min_element( m_freeBuffers.begin(), m_freeBuffers.end(), [] ( const auto& a, const auto& b ) {
                return a.second.capacity < b.second.capacity;
            } );
            m_gpus[smallest->first]->freeDevice( smallest->second );
            m_freeBuffers.erase( smallest );
        }
    }

    /** The smallest spare buffer of device @p decoder that is large enough, else a new one. */
    GpuDecoder::DeviceBuffer
    takeBuffer( size_t decoder, size_t capacity )
    {
        auto best = m_freeBuffers.end();
        for ( auto it = m_freeBuffers.begin(); it != m_freeBuffers.end(); ++it ) {
            if ( ( it->first == decoder ) && ( it->second.capacity >= capacity )
                 && ( ( best == m_freeBuffers.end() ) || ( it->second.capacity < best->second.capacity ) ) ) {
                best = it;
            }
        }
        if ( best != m_freeBuffers.end() ) {
            const auto buffer = best->second;
            m_freeBuffers.erase( best );
            return buffer;
        }
        return gpu( decoder ).allocDevice( capacity );
    }

    /* the compressed file on the host: a mapping, or bytes this object owns */
    const uint8_t* m_data{ nullptr };
    size_t m_size{ 0 };
    void* m_mapping{ nullptr };
    std::vector<uint8_t> m_owned;

    const size_t m_parallelization;
    const std::vector<int> m_devices;
    std::vector<std::unique_ptr<GpuDecoder>> m_gpus;

    std::vector<uint64_t> m_finderOffsets;
    bool m_finderReady{ false };
    size_t m_finderLimit{ std::numeric_limits<size_t>::max() };

    BatchList m_batches;   /* youngest first */
    std::unordered_map<uint64_t, std::pair<BatchList::iterator, size_t>> m_byOffset;
    std::vector<std::pair<size_t, GpuDecoder::DeviceBuffer>> m_freeBuffers;   /* ( device of the set, buffer ) */
    size_t m_cachedBytes{ 0 };
    size_t m_cacheBudget{ DEFAULT_CACHE_BYTES };
    size_t m_minBlocksPerDevice{ MIN_BLOCKS_PER_DEVICE };
    size_t m_nextSequentialBlock{ 0 };
    size_t m_sequentialFetches{ 0 };
    size_t m_expected{ 0 };
    size_t m_hits{ 0 };
    size_t m_misses{ 0 };
    size_t m_prefetchHits{ 0 };

    std::thread m_prefetchThread;
    Batch m_prefetched;
    std::exception_ptr m_prefetchFailure;
    size_t m_prefetchFirst{ 0 };
    size_t m_prefetchCount{ 0 };
    size_t m_prefetchMinBlocks{ 256 };   /* batches smaller than this are not prefetched; $IBZ2_PREFETCH_MIN_BLOCKS, 0 = never */
    bool m_prefetchPending{ false };
};
}  // namespace ibz2host
