/*
 * BlockIndex -- the seek index of a bzip2 file: compressed bit offset of every block <-> offset of its first decoded byte.
 * Host-side mirror of the reference's block map semantics (core/BlockMap.hpp:69-230), which decide what
 * ParallelBZ2Reader::blockOffsets() returns (Appendix B of SURVEY.md):
 *   - append( encodedOffset, encodedSize, decodedSize ) in file order; an entry with decodedSize 0 is an end-of-stream
 *     block (BlockMap.hpp:69-93);
 *   - re-appending a known offset is accepted when its size agrees (BlockMap.hpp:95-119);
 *   - locate( decodedOffset ) = the LAST entry that starts at or before the offset (BlockMap.hpp:121-137), so that a
 *     position on a block boundary belongs to the following data block, not to an EOS entry in between;
 *   - finish() appends the sentinel ( end of the last block, total bytes ), or ( 0, 0 ) for a file without any block
 *     (BlockMap.hpp:174-198);
 *   - assign( map ) imports an index: entries whose decoded offset equals the next one's are end-of-stream blocks, the
 *     last entry always is (BlockMap.hpp:207-230).
 * Not thread-safe: the readers that own it are documented as single-caller objects (ParallelBZ2Reader.hpp:36-38).
 */
#pragma once

#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <limits>
#include <map>
#include <stdexcept>
#include <utility>
#include <vector>

namespace ibz2host
{
struct BlockSpan
{
    size_t encodedOffsetInBits{ 0 };
    size_t encodedSizeInBits{ 0 };
    size_t decodedOffsetInBytes{ 0 };
    size_t decodedSizeInBytes{ 0 };
    size_t blockIndex{ 0 };   /* position inside the index, EOS entries included */
    bool valid{ false };

    [[nodiscard]] bool
    contains( size_t decodedOffset ) const
    {
        return valid && ( decodedOffsetInBytes <= decodedOffset )
               && ( decodedOffset < decodedOffsetInBytes + decodedSizeInBytes );
    }
};

class BlockIndex
{
public:
    using Entry = std::pair<size_t /* encoded bits */, size_t /* decoded bytes */>;

    size_t
    append( size_t encodedOffset, size_t encodedSize, size_t decodedSize )
    {
        if ( m_finished ) {
            throw std::invalid_argument( "May not insert into finalized block map!" );
        }
        if ( m_entries.empty() || ( encodedOffset > m_entries.back().first ) ) {
            const size_t decodedOffset = m_entries.empty() ? 0 : m_entries.back().second + m_lastDecodedSize;
            m_entries.emplace_back( encodedOffset, decodedOffset );
            if ( decodedSize == 0 ) {
                ++m_eosCount;
            }
            m_lastDecodedSize = decodedSize;
            m_lastEncodedSize = encodedSize;
            return decodedOffset;
        }

        /* a block seen before: it must sit where it sat and be as large as it was */
        const auto match = std::lower_bound( m_entries.begin(), m_entries.end(), encodedOffset,
                                             [] ( const Entry& e, size_t offset ) { return e.first < offset; } );
        if ( ( match == m_entries.end() ) || ( match->first != encodedOffset ) ) {
            throw std::invalid_argument( "Inserted block offsets should be strictly increasing!" );
        }
        if ( std::next( match ) == m_entries.end() ) {
            throw std::logic_error( "In this case, the new block should already have been appended above!" );
        }
        if ( std::next( match )->second - match->second != decodedSize ) {
            throw std::invalid_argument( "Got duplicate block offset with inconsistent size!" );
        }
        return match->second;
    }

    [[nodiscard]] BlockSpan
    locate( size_t decodedOffset ) const
    {
        /* entries are sorted by decoded offset; several may share one (EOS blocks): take the last of them */
        const auto after = std::upper_bound( m_entries.begin(), m_entries.end(), decodedOffset,
                                             [] ( size_t offset, const Entry& e ) { return offset < e.second; } );
        if ( after == m_entries.begin() ) {
            return {};
        }
        return span( static_cast<size_t>( std::distance( m_entries.begin(), after ) ) - 1 );
    }

    [[nodiscard]] BlockSpan
    locateEncoded( size_t encodedOffset ) const
    {
        const auto match = std::lower_bound( m_entries.begin(), m_entries.end(), encodedOffset,
                                             [] ( const Entry& e, size_t offset ) { return e.first < offset; } );
        if ( ( match == m_entries.end() ) || ( match->first != encodedOffset ) ) {
            return {};
        }
        return span( static_cast<size_t>( std::distance( m_entries.begin(), match ) ) );
    }

    [[nodiscard]] size_t
    dataBlockCount() const
    {
        return m_entries.size() - m_eosCount;
    }

    /** Data blocks that hold decoded bytes [offset, offset + size); gives up (and says so with 0) beyond 4096 entries. */
    [[nodiscard]] size_t
    dataBlocksIn( size_t decodedOffset, size_t size ) const
    {
        if ( ( size == 0 ) || m_entries.empty() ) {
            return 0;
        }
        const auto end = ( size > std::numeric_limits<size_t>::max() - decodedOffset ) ? std::numeric_limits<size_t>::max()
                                                                                         : decodedOffset + size;
        auto it = std::upper_bound( m_entries.begin(), m_entries.end(), decodedOffset,
                                    [] ( size_t offset, const Entry& e ) { return offset < e.second; } );
        if ( it != m_entries.begin() ) {
            --it;
        }
        size_t count = 0;
        for ( size_t seen = 0; ( it != m_entries.end() ) && ( it->second < end ); ++it, ++seen ) {
            if ( seen > 4096 ) {
                return 0;
            }
            const auto next = std::next( it );
            if ( ( next != m_entries.end() ) && ( next->second > it->second ) ) {
                ++count;   /* an entry with bytes of its own: a data block */
            }
        }
        return count;
    }

    void
    finish()
    {
        if ( m_finished ) {
            return;
        }
        if ( m_entries.empty() ) {
            m_entries.emplace_back( 0, 0 );
        } else if ( ( m_lastEncodedSize != 0 ) || ( m_lastDecodedSize != 0 ) ) {
            m_entries.emplace_back( m_entries.back().first + m_lastEncodedSize,
                                    m_entries.back().second + m_lastDecodedSize );
        }
        m_lastEncodedSize = 0;
        m_lastDecodedSize = 0;
        m_finished = true;
    }

    [[nodiscard]] bool
    finished() const
    {
        return m_finished;
    }

    void
    assign( const std::map<size_t, size_t>& offsets )
    {
        m_entries.assign( offsets.begin(), offsets.end() );
        m_lastEncodedSize = 0;
        m_lastDecodedSize = 0;
        m_eosCount = 0;
        for ( size_t i = 0; i + 1 < m_entries.size(); ++i ) {
            if ( m_entries[i].second == m_entries[i + 1].second ) {
                ++m_eosCount;
            }
        }
        ++m_eosCount;   /* the last entry is taken to be an end-of-stream block */
        m_finished = true;
    }

    [[nodiscard]] std::map<size_t, size_t>
    offsets() const
    {
        return { m_entries.begin(), m_entries.end() };
    }

    [[nodiscard]] Entry
    back() const
    {
        if ( m_entries.empty() ) {
            throw std::out_of_range( "Can not return last element of empty block map!" );
        }
        return m_entries.back();
    }

    [[nodiscard]] bool
    empty() const
    {
        return m_entries.empty();
    }

    [[nodiscard]] size_t
    size() const
    {
        return m_entries.size();
    }

private:
    [[nodiscard]] BlockSpan
    span( size_t i ) const
    {
        BlockSpan result;
        result.valid = true;
        result.blockIndex = i;
        result.encodedOffsetInBits = m_entries[i].first;
        result.decodedOffsetInBytes = m_entries[i].second;
        if ( i + 1 == m_entries.size() ) {
            result.decodedSizeInBytes = m_lastDecodedSize;
            result.encodedSizeInBits = m_lastEncodedSize;
        } else {
            if ( m_entries[i + 1].second < m_entries[i].second ) {
                throw std::logic_error( "Data offsets are not monotonically increasing!" );
            }
            result.decodedSizeInBytes = m_entries[i + 1].second - m_entries[i].second;
            result.encodedSizeInBits = m_entries[i + 1].first - m_entries[i].first;
        }
        return result;
    }

    std::vector<Entry> m_entries;
    size_t m_eosCount{ 0 };
    size_t m_lastEncodedSize{ 0 };
    size_t m_lastDecodedSize{ 0 };
    bool m_finished{ false };
};
}  // namespace ibz2host
