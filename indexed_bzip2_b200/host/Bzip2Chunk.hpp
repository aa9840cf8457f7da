/*
 * Bzip2Chunk -- the bzip2 chunk decoder of the reference's gzip-architecture reader
 * (rapidgzip/chunkdecoding/Bzip2Chunk.hpp:34-268: decodeChunk / decodeUnknownBzip2Chunk), on the GPU.
 *
 * The chunk fetcher of rapidgzip::ParallelGzipReader asks a worker to "decode what starts at bit chunkOffset and stop at
 * the first block that starts at or behind untilOffset".  For bzip2 that is: all blocks whose magic starts in
 * [ first block at or behind chunkOffset, untilOffset ), the end-of-stream blocks and stream headers between and behind
 * them included, until the next data block.  Here ONE ibz2_decode_buffer call does the finding and the decoding; the
 * host then lays out what the reference's ChunkData carries for bzip2:
 *   encodedOffsetInBits / encodedSizeInBits   where the chunk really starts and how far it reaches
 *   decodedSizeInBytes + the bytes            (the bytes stay in HBM; deliver() hands them out)
 *   blockBoundaries                           ( encoded offset, decoded offset ) of every block, end-of-stream blocks
 *                                             included, in front of which the chunk already holds data
 *                                             (ChunkData::appendDeflateBlockBoundary, Bzip2Chunk.hpp:119-121)
 *   footers                                   ( bit position behind an end-of-stream block, decoded offset )
 *                                             (Bzip2Chunk.hpp:125-131)
 *   stoppedPreemptively                       the decoded size reached maxDecompressedChunkSize (Bzip2Chunk.hpp:88-91)
 * Subchunks (the on-the-fly splitting for the window map) need no port: a bzip2 block needs no window, every block
 * boundary is a possible split point, and the caller has them all.
 */
#pragma once

#include <algorithm>
#include <cstdint>
#include <limits>
#include <sstream>
#include <stdexcept>
#include <vector>

#include "GpuDecoder.hpp"

namespace ibz2host
{
/** rapidgzip::NoBlockInRange (core/BlockFetcher.hpp): nothing decodable starts inside the requested range */
class NoBlockInRange : public std::domain_error
{
public:
    using std::domain_error::domain_error;
};

struct Bzip2ChunkData
{
    struct Boundary
    {
        uint64_t encodedOffset{ 0 };
        uint64_t decodedOffset{ 0 };
    };

    uint64_t encodedOffsetInBits{ 0 };
    uint64_t encodedSizeInBits{ 0 };
    uint64_t decodedSizeInBytes{ 0 };
    std::vector<Boundary> blockBoundaries;
    std::vector<Boundary> footers;
    std::vector<uint32_t> blockCrcs;     /* calculated CRC of every data block, in order (ChunkData keeps stream CRCs) */
    bool stoppedPreemptively{ false };
    GpuDecoder::DeviceBuffer data;       /* decodedSizeInBytes bytes in HBM; owned by the caller (GpuDecoder::freeDevice) */
};

class Bzip2Chunk
{
public:
    /**
     * @param decoder      holds the compressed file in HBM (GpuDecoder::upload)
     * @param fileData     the same bytes on the host, for the header probes between blocks
     * @param allBlocks    GpuDecoder::findBlocks() of the file: every data-block magic, ascending
     * Throws NoBlockInRange like the reference when no block starts in [chunkOffset, untilOffset), and what the
     * reference's Block would throw for a damaged block.
     */
    [[nodiscard]] static Bzip2ChunkData
    decodeChunk( GpuDecoder& decoder, const uint8_t* fileData, size_t fileSize, const std::vector<uint64_t>& allBlocks,
                 uint64_t chunkOffset, uint64_t untilOffset,
                 uint64_t maxDecompressedChunkSize = std::numeric_limits<uint64_t>::max() )
    {
        /* the data blocks that start inside the range; the reference tries chunkOffset itself and then looks for the
         * next magic (Bzip2Chunk.hpp:225-258), which comes to the same list */
        Bzip2ChunkData result;
        /* end-of-stream blocks and stream headers from @p position on belong to the chunk (Bzip2Chunk.hpp:123-138),
         * but nothing that starts exactly at untilOffset does (the second half of the stop condition,
         * Bzip2Chunk.hpp:112-114): an end-of-stream block there opens the next chunk */
        const auto skipStreamEnds = [&] ( uint64_t position, uint64_t decodedSoFar ) {
            while ( ( position + 48U + 32U <= fileSize * 8ULL ) && ( ( position != untilOffset ) || ( position == chunkOffset ) ) ) {
                const auto header = GpuDecoder::probeHeader( fileData, fileSize, position );
                if ( ( header.status != IBZ2_ST_OK ) || ( header.is_eos == 0 ) ) {
                    break;
                }
                if ( decodedSoFar > 0 ) {
                    result.blockBoundaries.push_back( { position, decodedSoFar } );
                }
                position = header.enc_offset_bits + header.enc_size_bits;
                result.footers.push_back( { position, decodedSoFar } );
                if ( header.is_eof != 0 ) {
                    break;
                }
                uint64_t behindHeader = position;
                if ( ibz2_read_stream_header( fileData, fileSize, &behindHeader ) <= 0 ) {
                    break;   /* trailing garbage: the chunk ends behind the end-of-stream block */
                }
                position = behindHeader;
            }
            return position;
        };

        /* a chunk may start at an end-of-stream block (the reference's chunk map has such entries: the tail of a file
         * whose last data block ended the chunk in front of it); it then begins with footers and may hold nothing else */
        const uint64_t firstPossibleBlock = skipStreamEnds( chunkOffset, 0 );
        const bool startsAtStreamEnd = firstPossibleBlock != chunkOffset;
        const std::vector<uint64_t> blocks( std::lower_bound( allBlocks.begin(), allBlocks.end(), firstPossibleBlock ),
                                            std::lower_bound( allBlocks.begin(), allBlocks.end(), untilOffset ) );
        if ( blocks.empty() ) {
            if ( startsAtStreamEnd ) {
                result.encodedOffsetInBits = chunkOffset;
                result.encodedSizeInBits = firstPossibleBlock - chunkOffset;
                return result;
            }
            std::stringstream message;
            message << "Failed to find any valid bzip2 block in [" << chunkOffset << " b, " << untilOffset << " b)";
            throw NoBlockInRange( message.str() );
        }

        result.encodedOffsetInBits = startsAtStreamEnd ? chunkOffset : blocks.front();
        std::vector<ibz2_block_result> results;
        result.data = decoder.allocDevice( blocks.size() * 1000000ULL + 65536ULL );
        try {
            decoder.decode( blocks.data(), blocks.size(), results, result.data );
        } catch ( ... ) {
            decoder.freeDevice( result.data );
            throw;
        }

        uint64_t decoded = 0;
        uint64_t nextBlockOffset = blocks.front();
        for ( size_t i = 0; i < results.size(); ++i ) {
            const auto& r = results[i];
            if ( decoded >= maxDecompressedChunkSize ) {
                result.stoppedPreemptively = true;
                break;
            }
            if ( ( r.status == IBZ2_ST_OK ) && ( r.calculated_crc != r.expected_crc ) ) {
                ibz2_block_result failed = r;
                failed.status = IBZ2_ST_CRC;
                decoder.freeDevice( result.data );
                throwForStatus( failed );
            }
            if ( r.status != IBZ2_ST_OK ) {
                decoder.freeDevice( result.data );
                throwForStatus( r );
            }
            if ( decoded > 0 ) {
                result.blockBoundaries.push_back( { r.enc_offset_bits, decoded } );
            }
            decoded += r.out_size;
            result.blockCrcs.push_back( r.calculated_crc );
            nextBlockOffset = skipStreamEnds( r.enc_offset_bits + r.enc_size_bits, decoded );
        }
        result.decodedSizeInBytes = decoded;
        result.encodedSizeInBits = nextBlockOffset - result.encodedOffsetInBits;
        return result;
    }
};
}  // namespace ibz2host
