/*
 * IndexFileFormat -- the block map of a bzip2 file in the index formats the reference's tools exchange:
 *   - the "<encoded bits>,<decoded bytes>" lines that `ibzip2 -L` writes (tools/ibzip2.cpp:439-452);
 *   - GZIDX version 1, the indexed_gzip format that rapidgzip writes and reads for every compression it decodes,
 *     bzip2 included (rapidgzip/IndexFileFormat.hpp:37-64 layout, :596-660 writer, :274-412 reader; a bzip2 index has
 *     no window data: a bzip2 block needs nothing in front of it).
 * A checkpoint of a GZIDX file is ( compressed byte offset rounded UP, decoded offset, bits = 8 - offset % 8 ): the
 * bit offset is byte * 8 - bits.  This writer emits one checkpoint per data block plus the end of the file, the shape
 * rapidgzip's own bzip2 indexes have (its checkpoints are chunk starts, every one of them a block boundary).
 */
#pragma once

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <istream>
#include <limits>
#include <map>
#include <ostream>
#include <stdexcept>
#include <string>
#include <vector>

namespace ibz2host
{
enum class IndexFormat
{
    OFFSET_LINES,
    GZIDX,
};

constexpr char GZIDX_MAGIC[5] = { 'G', 'Z', 'I', 'D', 'X' };
constexpr uint32_t GZIDX_WINDOW_SIZE = 32768;

namespace detail
{
template<typename T>
void
putLittleEndian( std::ostream& out, T value )
{
    char bytes[sizeof( T )];
    for ( size_t i = 0; i < sizeof( T ); ++i ) {
        bytes[i] = static_cast<char>( ( static_cast<uint64_t>( value ) >> ( 8U * i ) ) & 0xFFU );
    }
    out.write( bytes, sizeof( T ) );
}

template<typename T>
T
getLittleEndian( std::istream& in )
{
    char bytes[sizeof( T )];
    in.read( bytes, sizeof( T ) );
    if ( static_cast<size_t>( in.gcount() ) != sizeof( T ) ) {
        throw std::invalid_argument( "Premature end of index file!" );
    }
    uint64_t value = 0;
    for ( size_t i = 0; i < sizeof( T ); ++i ) {
        value |= static_cast<uint64_t>( static_cast<uint8_t>( bytes[i] ) ) << ( 8U * i );
    }
    return static_cast<T>( value );
}
}  // namespace detail

/**
 * @param offsets the reader's blockOffsets(): every block (data and end-of-stream) and the end sentinel.
 * Entries without decoded bytes of their own (end-of-stream blocks) carry no information a decoder needs and are
 * left out, like in the reference's own GZIDX files for bzip2.
 */
inline void
writeGzidx( std::ostream& out, const std::map<size_t, size_t>& offsets, uint64_t compressedSizeInBytes )
{
    if ( offsets.empty() ) {
        throw std::invalid_argument( "Cannot write an index without the end-of-file entry!" );
    }
    std::vector<std::pair<uint64_t, uint64_t>> checkpoints;
    for ( auto it = offsets.begin(); it != offsets.end(); ++it ) {
        const auto next = std::next( it );
        if ( ( next == offsets.end() ) || ( next->second != it->second ) ) {
            checkpoints.emplace_back( it->first, it->second );   /* a data block, or the end sentinel */
        }
    }
    const uint64_t decodedSize = offsets.rbegin()->second;
    uint64_t spacing = std::numeric_limits<uint64_t>::max();
    for ( size_t i = 1; i < checkpoints.size(); ++i ) {
        spacing = std::min( spacing, checkpoints[i].second - checkpoints[i - 1].second );
    }
    out.write( GZIDX_MAGIC, sizeof( GZIDX_MAGIC ) );
    out.put( '\x01' );   /* format version */
    out.put( '\x00' );   /* flags, unused */
    detail::putLittleEndian<uint64_t>( out, compressedSizeInBytes );
    detail::putLittleEndian<uint64_t>( out, decodedSize );
    detail::putLittleEndian<uint32_t>( out, static_cast<uint32_t>( std::max<uint64_t>( GZIDX_WINDOW_SIZE,
                                            std::min<uint64_t>( spacing, std::numeric_limits<uint32_t>::max() ) ) ) );
    detail::putLittleEndian<uint32_t>( out, GZIDX_WINDOW_SIZE );
    detail::putLittleEndian<uint32_t>( out, static_cast<uint32_t>( checkpoints.size() ) );
    for ( const auto& [bits, bytes] : checkpoints ) {
        const auto remainder = bits % 8U;
        detail::putLittleEndian<uint64_t>( out, bits / 8U + ( remainder == 0 ? 0U : 1U ) );
        detail::putLittleEndian<uint64_t>( out, bytes );
        detail::putLittleEndian<uint8_t>( out, static_cast<uint8_t>( remainder == 0 ? 0U : 8U - remainder ) );
        detail::putLittleEndian<uint8_t>( out, 0 );   /* no window data */
    }
    if ( !out ) {
        throw std::runtime_error( "Writing the index failed!" );
    }
}

/** @return checkpoint bit offset -> decoded byte offset, the end-of-file entry included; window data is skipped. */
[[nodiscard]] inline std::map<size_t, size_t>
readGzidx( std::istream& in )
{
    char magic[sizeof( GZIDX_MAGIC )] = {};
    in.read( magic, sizeof( magic ) );
    if ( ( static_cast<size_t>( in.gcount() ) != sizeof( magic ) ) || ( std::memcmp( magic, GZIDX_MAGIC, sizeof( magic ) ) != 0 ) ) {
        throw std::invalid_argument( "Magic bytes do not match! Expected 'GZIDX'." );
    }
    const auto version = detail::getLittleEndian<uint8_t>( in );
    if ( version > 1 ) {
        throw std::invalid_argument( "Index was written with a newer indexed_gzip version than supported!" );
    }
    detail::getLittleEndian<uint8_t>( in );   /* flags */
    const auto compressedSize = detail::getLittleEndian<uint64_t>( in );
    const auto decodedSize = detail::getLittleEndian<uint64_t>( in );
    detail::getLittleEndian<uint32_t>( in );   /* spacing: only a hint for writers */
    const auto windowSize = detail::getLittleEndian<uint32_t>( in );
    const auto count = detail::getLittleEndian<uint32_t>( in );
    std::map<size_t, size_t> offsets;
    for ( uint32_t i = 0; i < count; ++i ) {
        const auto byteOffset = detail::getLittleEndian<uint64_t>( in );
        const auto decodedOffset = detail::getLittleEndian<uint64_t>( in );
        const auto bits = detail::getLittleEndian<uint8_t>( in );
        if ( bits >= 8 ) {
            throw std::invalid_argument( "Denormal compressed offset for checkpoint. Bits must be in [0, 7]." );
        }
        if ( version >= 1 ) {
            detail::getLittleEndian<uint8_t>( in );   /* window flag: bzip2 blocks need none; window bytes trail the table */
        }
        const uint64_t bitOffset = byteOffset * 8U - bits;
        if ( ( bitOffset > compressedSize * 8U ) || ( decodedOffset > decodedSize ) ) {
            throw std::invalid_argument( "Checkpoint lies outside of the file the index describes!" );
        }
        offsets[bitOffset] = decodedOffset;
    }
    (void)windowSize;
    if ( offsets.empty() || ( offsets.rbegin()->second != decodedSize ) ) {
        offsets[compressedSize * 8U] = decodedSize;   /* indexed_gzip files end with the last window, not the file */
    }
    return offsets;
}
}  // namespace ibz2host
