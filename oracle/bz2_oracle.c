/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement (plain C11) of the reference's bzip2 decode path.
 * See bz2_oracle.h for the contract and the parity-pinning statement.  The product path under
 * indexed_bzip2_b200/ never includes, links or calls this file.
 *
 * Citations are relative to /root/reference/src.
 */
#include "bz2_oracle.h"

#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------------
 * CRC-32/BZIP2: poly 0x04C11DB7, MSB-first, no reflection (indexed_bzip2/bzip2.hpp:59-91)
 * ---------------------------------------------------------------------------------------------- */
static uint32_t g_crcTable[256];
static int g_crcTableReady = 0;

static void
crc_init( void )
{
    if ( g_crcTableReady ) {
        return;
    }
    for ( uint32_t i = 0; i < 256; ++i ) {
        uint32_t c = i << 24;
        for ( int j = 0; j < 8; ++j ) {
            c = ( c & 0x80000000U ) ? ( c << 1 ) ^ 0x04C11DB7U : ( c << 1 );
        }
        g_crcTable[i] = c;
    }
    g_crcTableReady = 1;
}

uint32_t
orc_crc32_update( uint32_t crc, const uint8_t* data, size_t n )
{
    crc_init();
    for ( size_t i = 0; i < n; ++i ) {
        crc = ( crc << 8 ) ^ g_crcTable[( ( crc >> 24 ) ^ data[i] ) & 0xFFU]; /* bzip2.hpp:90 */
    }
    return crc;
}

uint32_t
orc_crc32( const uint8_t* data, size_t n )
{
    return ~orc_crc32_update( 0xFFFFFFFFU, data, n ); /* bzip2.hpp:833, 901 */
}

uint32_t
orc_stream_crc_fold( uint32_t streamCrc, uint32_t blockCrc )
{
    return ( ( streamCrc << 1 ) | ( streamCrc >> 31 ) ) ^ blockCrc; /* indexed_bzip2/BZ2Reader.hpp:481-484 */
}

/* ------------------------------------------------------------------------------------------------
 * MSB-first bit access (semantics of core/BitReader.hpp:190-278 in bzip2 mode; reading past the end
 * is EndOfFileReached, core/BitReader.hpp:82)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    const uint8_t* buf;
    uint64_t nbits;
    uint64_t pos;
    int eof; /* sticky: a read ran past the end */
} bitrd;

static uint32_t
rd_bits( bitrd* r, unsigned n )
{
    if ( r->pos + n > r->nbits ) {
        r->eof = 1;
        r->pos = r->nbits;
        return 0;
    }
    uint32_t v = 0;
    for ( unsigned i = 0; i < n; ++i ) {
        const uint64_t p = r->pos + i;
        v = ( v << 1 ) | ( ( r->buf[p >> 3] >> ( 7 - ( p & 7 ) ) ) & 1U );
    }
    r->pos += n;
    return v;
}

/* ------------------------------------------------------------------------------------------------
 * Magic finder (core/BitStringFinder.hpp:158-238): a match needs all 48 bits inside the buffer; the
 * reference finds them with 8 shifted 5-byte searches plus head/tail bit checks, which is equivalent to
 * testing every bit offset.  Offsets are returned ascending (BitStringFinder.hpp:274-281 after the
 * caller's sort, tools/ibzip2.cpp:125).
 * ---------------------------------------------------------------------------------------------- */
size_t
orc_find_magic( const uint8_t* buf, size_t nbytes, uint64_t magic48, uint64_t* out, size_t cap )
{
    size_t count = 0;
    if ( nbytes < 6 ) {
        return 0;
    }
    const uint64_t mask = 0xFFFFFFFFFFFFULL;
    magic48 &= mask;
    for ( size_t byte = 0; byte + 6 <= nbytes; ++byte ) {
        /* 56-bit window: bytes [byte, byte + 7), zero-padded past the end */
        uint64_t w = 0;
        for ( size_t k = 0; k < 7; ++k ) {
            w = ( w << 8 ) | ( byte + k < nbytes ? buf[byte + k] : 0 );
        }
        for ( unsigned s = 0; s < 8; ++s ) {
            if ( ( s > 0 ) && ( byte + 7 > nbytes ) ) {
                break; /* tail bits would lie outside the buffer (tailMatches, BitStringFinder.hpp:210-222) */
            }
            if ( ( ( w >> ( 8 - s ) ) & mask ) == magic48 ) {
                if ( count < cap ) {
                    out[count] = (uint64_t)byte * 8 + s;
                }
                ++count;
            }
        }
    }
    return count;
}

/* ------------------------------------------------------------------------------------------------
 * RUNA/RUNB (indexed_bzip2/bzip2.hpp:726-742)
 * ---------------------------------------------------------------------------------------------- */
uint32_t
orc_run_length( const uint8_t* runSymbols, size_t n )
{
    uint32_t hh = 0;
    uint32_t runPos = 0;
    for ( size_t i = 0; i < n; ++i ) {
        if ( runPos == 0 ) {
            runPos = 1;
            hh = 0;
        }
        hh += runPos << runSymbols[i];
        runPos <<= 1;
    }
    return hh;
}

/* ------------------------------------------------------------------------------------------------
 * Stream header (indexed_bzip2/bzip2.hpp:114-142)
 * ---------------------------------------------------------------------------------------------- */
static int
read_stream_header( bitrd* r )
{
    static const char magic[3] = { 'B', 'Z', 'h' };
    for ( int i = 0; i < 3; ++i ) {
        const uint32_t b = rd_bits( r, 8 );
        if ( r->eof ) {
            return -1; /* EndOfFileReached */
        }
        if ( b != (uint8_t)magic[i] ) {
            return 0; /* domain_error */
        }
    }
    const uint32_t lvl = rd_bits( r, 8 );
    if ( r->eof ) {
        return -1;
    }
    if ( ( lvl < '1' ) || ( lvl > '9' ) ) {
        return 0;
    }
    return (int)( lvl - '0' );
}

int
orc_read_stream_header( const uint8_t* buf, size_t nbytes, uint64_t* bitpos )
{
    bitrd r = { buf, (uint64_t)nbytes * 8, *bitpos, 0 };
    const int lvl = read_stream_header( &r );
    *bitpos = r.pos;
    return lvl > 0 ? lvl : 0;
}

/* ------------------------------------------------------------------------------------------------
 * Canonical Huffman coding from code lengths
 *   huffman/HuffmanCodingBase.hpp:46-187, HuffmanCodingSymbolsPerLength.hpp:43-124,
 *   HuffmanCodingShortBitsCached.hpp:39-150 (the 12-bit LUT is an acceleration of the same mapping).
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    uint8_t minLen, maxLen;
    uint32_t minCode[ORC_MAX_CODE_LEN + 2]; /* indexed by length - minLen (HuffmanCodingBase.hpp:143-148) */
    uint16_t offsets[ORC_MAX_CODE_LEN + 3]; /* indexed by length - minLen (SymbolsPerLength.hpp:47-53) */
    uint16_t symbolsPerLength[ORC_MAX_SYMBOLS];
} huffcoding;

static int
huff_init( huffcoding* h, const uint8_t* lengths, unsigned count )
{
    /* initializeMinMaxCodeLengths (HuffmanCodingBase.hpp:46-71): getMax / getMinPositive */
    unsigned maxLen = 0;
    unsigned minLen = 255;
    uint32_t freq[256];
    memset( freq, 0, sizeof( freq ) );
    for ( unsigned i = 0; i < count; ++i ) {
        if ( lengths[i] > maxLen ) {
            maxLen = lengths[i];
        }
        if ( ( lengths[i] > 0 ) && ( lengths[i] < minLen ) ) {
            minLen = lengths[i];
        }
        ++freq[lengths[i]];
    }
    h->minLen = (uint8_t)minLen;
    h->maxLen = (uint8_t)maxLen;

    /* checkCodeLengthFrequencies (HuffmanCodingBase.hpp:73-112) with CHECK_OPTIMALITY = false */
    uint32_t unused = 1U << minLen;
    for ( unsigned len = minLen; len <= maxLen; ++len ) {
        if ( freq[len] > unused ) {
            return ORC_ERR_CODE_LENGTHS;
        }
        unused -= freq[len];
        unused *= 2;
    }

    /* initializeMinimumCodeValues (HuffmanCodingBase.hpp:117-149) */
    freq[0] = 0;
    uint32_t minCode = 0;
    for ( unsigned bits = ( minLen > 1 ? minLen : 1 ); bits <= maxLen; ++bits ) {
        minCode = ( minCode + freq[bits - 1] ) << 1;
        h->minCode[bits - minLen] = minCode;
    }

    /* initializeSymbolsPerLength (HuffmanCodingSymbolsPerLength.hpp:43-68) */
    unsigned sum = 0;
    for ( unsigned len = minLen; len <= maxLen; ++len ) {
        h->offsets[len - minLen] = (uint16_t)sum;
        sum += freq[len];
    }
    h->offsets[maxLen - minLen + 1] = (uint16_t)sum;
    uint16_t sizes[ORC_MAX_CODE_LEN + 3];
    memcpy( sizes, h->offsets, sizeof( sizes ) );
    for ( unsigned s = 0; s < count; ++s ) {
        if ( lengths[s] != 0 ) {
            h->symbolsPerLength[sizes[lengths[s] - minLen]++] = (uint16_t)s;
        }
    }
    return ORC_OK;
}

/* HuffmanCodingSymbolsPerLength::decode (:97-124); returns symbol, -1 for nullopt, -2 for EOF */
static int
huff_decode( const huffcoding* h, bitrd* r )
{
    uint32_t code = 0;
    for ( unsigned i = 0; i < h->minLen; ++i ) {
        code = ( code << 1 ) | rd_bits( r, 1 );
    }
    if ( r->eof ) {
        return -2;
    }
    for ( unsigned k = 0; k <= (unsigned)( h->maxLen - h->minLen ); ++k ) {
        const uint32_t minCode = h->minCode[k];
        if ( minCode <= code ) {
            const uint32_t sub = h->offsets[k] + ( code - minCode );
            if ( sub < h->offsets[k + 1] ) {
                return h->symbolsPerLength[sub];
            }
        }
        code = ( code << 1 ) | rd_bits( r, 1 );
        if ( r->eof ) {
            return -2;
        }
    }
    return -1;
}

/* ------------------------------------------------------------------------------------------------
 * Block header + trees (indexed_bzip2/bzip2.hpp:479-685)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    huffcoding codings[ORC_MAX_GROUPS];
    uint8_t selectors[ORC_MAX_SELECTORS];
} blocktables;

static int
parse_header( bitrd* r, orc_block_info* info, blocktables* t )
{
    info->enc_offset_bits = r->pos;
    info->enc_size_bits = 0;
    info->is_eos = 0;
    info->at_eof = 0;

    /* readBlockHeader (bzip2.hpp:479-523) */
    const uint64_t magic = ( (uint64_t)rd_bits( r, 24 ) << 24 ) | rd_bits( r, 24 );
    info->header_crc = rd_bits( r, 32 );
    if ( r->eof ) {
        return ORC_ERR_EOF;
    }
    if ( magic == ORC_MAGIC_EOS ) {
        info->is_eos = 1;
        const unsigned inByte = (unsigned)( r->pos & 7U );
        if ( inByte > 0 ) {
            rd_bits( r, 8 - inByte );
            if ( r->eof ) {
                return ORC_ERR_EOF;
            }
        }
        info->enc_size_bits = r->pos - info->enc_offset_bits;
        info->at_eof = r->pos >= r->nbits;
        return ORC_OK;
    }
    if ( magic != ORC_MAGIC_BLOCK ) {
        return ORC_ERR_BAD_MAGIC;
    }
    if ( rd_bits( r, 1 ) != 0 ) {
        return r->eof ? ORC_ERR_EOF : ORC_ERR_RANDOMISED;
    }
    info->orig_ptr = rd_bits( r, 24 );
    if ( r->eof ) {
        return ORC_ERR_EOF;
    }
    if ( info->orig_ptr > ORC_DBUF_SIZE ) {
        return ORC_ERR_ORIGPTR_RANGE;
    }

    /* readSymbolMaps (bzip2.hpp:526-571) */
    const uint32_t usedMap = rd_bits( r, 16 );
    info->symbol_count = 0;
    for ( int i = 0; i < 16; ++i ) {
        if ( usedMap & ( 1U << ( 15 - i ) ) ) {
            const uint32_t bitmap = rd_bits( r, 16 );
            for ( int j = 0; j < 16; ++j ) {
                if ( bitmap & ( 1U << ( 15 - j ) ) ) {
                    info->symbol_to_byte[info->symbol_count++] = (uint8_t)( 16 * i + j );
                }
            }
        }
    }
    if ( r->eof ) {
        return ORC_ERR_EOF;
    }

    /* readSelectors (bzip2.hpp:574-637) */
    info->group_count = rd_bits( r, 3 );
    if ( r->eof ) {
        return ORC_ERR_EOF;
    }
    if ( ( info->group_count < 2 ) || ( info->group_count > ORC_MAX_GROUPS ) ) {
        return ORC_ERR_GROUP_COUNT;
    }
    info->selectors_count = rd_bits( r, 15 );
    if ( r->eof ) {
        return ORC_ERR_EOF;
    }
    if ( info->selectors_count == 0 ) {
        return ORC_ERR_SELECTOR_COUNT;
    }
    uint8_t mtf[ORC_MAX_GROUPS];
    for ( unsigned i = 0; i < info->group_count; ++i ) {
        mtf[i] = (uint8_t)i;
    }
    for ( unsigned i = 0; i < info->selectors_count; ++i ) {
        /* unary: j ones then a zero; the reference peeks 6 bits (bzip2.hpp:602-624), which near the end of
         * the file needs those bits to exist -- valid files always have >= 80 bits of footer behind. */
        unsigned j = 0;
        while ( 1 ) {
            const uint32_t bit = rd_bits( r, 1 );
            if ( r->eof ) {
                return ORC_ERR_EOF;
            }
            if ( bit == 0 ) {
                break;
            }
            ++j;
            if ( j >= ORC_MAX_GROUPS ) {
                break; /* 111111 -> selector 6 */
            }
        }
        if ( j >= info->group_count ) {
            return ORC_ERR_SELECTOR_UNARY;
        }
        const uint8_t uc = mtf[j];
        memmove( mtf + 1, mtf, j );
        mtf[0] = uc;
        t->selectors[i] = uc;
    }

    /* readTrees (bzip2.hpp:644-685) */
    const unsigned symCount = info->symbol_count + 2;
    for ( unsigned g = 0; g < info->group_count; ++g ) {
        uint8_t* lengths = info->code_lengths[g];
        memset( lengths, 0, ORC_MAX_SYMBOLS );
        unsigned hh = rd_bits( r, 5 );
        for ( unsigned s = 0; s < symCount; ++s ) {
            while ( 1 ) {
                if ( r->eof ) {
                    return ORC_ERR_EOF;
                }
                if ( ORC_MAX_CODE_LEN - 1 < hh - 1 ) { /* !hh || hh > 20, bzip2.hpp:657 */
                    return ORC_ERR_CODE_LENGTH;
                }
                if ( rd_bits( r, 1 ) != 0 ) {
                    hh += 1 - ( rd_bits( r, 1 ) << 1 );
                } else {
                    break;
                }
            }
            if ( r->eof ) {
                return ORC_ERR_EOF;
            }
            lengths[s] = (uint8_t)hh;
        }
        const int err = huff_init( &t->codings[g], lengths, symCount );
        if ( err != ORC_OK ) {
            return err;
        }
    }
    return ORC_OK;
}

int
orc_decode_block( const uint8_t* buf, size_t nbytes, uint64_t bit_offset, orc_block_info* info,
                  uint8_t* selectorsOut, uint16_t* symbols, uint8_t* lbytes, uint32_t* histOut,
                  uint32_t* tvecOut, uint8_t* prerle, uint8_t* out, uint64_t out_cap )
{
    memset( info, 0, sizeof( *info ) );
    bitrd r = { buf, (uint64_t)nbytes * 8, bit_offset, 0 };
    if ( bit_offset > r.nbits ) {
        info->status = ORC_ERR_EOF;
        return info->status;
    }

    blocktables* t = (blocktables*)malloc( sizeof( blocktables ) );
    uint32_t* dbuf = (uint32_t*)malloc( sizeof( uint32_t ) * ORC_DBUF_SIZE );
    int status = parse_header( &r, info, t );
    if ( ( status != ORC_OK ) || info->is_eos ) {
        goto done;
    }
    if ( selectorsOut != NULL ) {
        memcpy( selectorsOut, t->selectors, info->selectors_count );
    }

    /* readBlockData (bzip2.hpp:691-807) */
    {
        uint32_t byteCount[256];
        memset( byteCount, 0, sizeof( byteCount ) );
        uint8_t mtf[256];
        for ( int i = 0; i < 256; ++i ) {
            mtf[i] = (uint8_t)i;
        }
        uint32_t dbufCount = 0;
        uint32_t nsym = 0;
        const huffcoding* coding = &t->codings[0];
        uint32_t hh = 0, runPos = 0, symCount = 0, selector = 0;
        while ( 1 ) {
            if ( symCount-- == 0 ) {
                symCount = ORC_GROUP_SIZE - 1;
                if ( selector >= info->selectors_count ) {
                    status = ORC_ERR_SELECTOR_OVERRUN;
                    goto done;
                }
                coding = &t->codings[t->selectors[selector]];
                selector++;
            }
            const int nextSym = huff_decode( coding, &r );
            if ( nextSym == -2 ) {
                status = ORC_ERR_EOF;
                goto done;
            }
            if ( nextSym < 0 ) {
                status = ORC_ERR_INVALID_CODE;
                goto done;
            }
            if ( ( symbols != NULL ) && ( nsym < ORC_SYMBOL_CAPACITY ) ) {
                symbols[nsym] = (uint16_t)nextSym;
            }
            ++nsym;

            if ( nextSym <= 1 ) {
                if ( runPos == 0 ) {
                    runPos = 1;
                    hh = 0;
                }
                hh += runPos << nextSym;
                runPos <<= 1;
                continue;
            }
            if ( runPos != 0 ) {
                runPos = 0;
                /* The reference adds in uint32 (bzip2.hpp:751); a wrapped sum would run off dbuf there (UB).
                 * We treat any run that does not fit as the overflow error. */
                if ( (uint64_t)dbufCount + hh > ORC_DBUF_SIZE ) {
                    status = ORC_ERR_DBUF_OVERFLOW;
                    goto done;
                }
                const uint8_t uc = info->symbol_to_byte[mtf[0]];
                byteCount[uc] += hh;
                while ( hh-- != 0 ) {
                    dbuf[dbufCount++] = uc;
                }
            }
            if ( (uint32_t)nextSym > info->symbol_count ) {
                break;
            }
            if ( dbufCount >= ORC_DBUF_SIZE ) {
                status = ORC_ERR_DBUF_OVERFLOW;
                goto done;
            }
            const int ii = nextSym - 1;
            uint8_t uc = mtf[ii];
            memmove( mtf + 1, mtf, (size_t)ii );
            mtf[0] = uc;
            uc = info->symbol_to_byte[uc];
            byteCount[uc]++;
            dbuf[dbufCount++] = uc;
        }
        info->nsymbols = nsym;
        info->n = dbufCount;
        if ( info->orig_ptr >= dbufCount ) {
            status = ORC_ERR_ORIGPTR;
            goto done;
        }
        info->enc_size_bits = r.pos - info->enc_offset_bits; /* bzip2.hpp:806 */

        if ( lbytes != NULL ) {
            for ( uint32_t i = 0; i < dbufCount; ++i ) {
                lbytes[i] = (uint8_t)dbuf[i];
            }
        }
        if ( histOut != NULL ) {
            memcpy( histOut, byteCount, sizeof( byteCount ) );
        }

        /* prepare (bzip2.hpp:810-847) */
        uint32_t cumulative = 0;
        for ( int i = 0; i < 256; ++i ) {
            const uint32_t next = cumulative + byteCount[i];
            byteCount[i] = cumulative;
            cumulative = next;
        }
        for ( uint32_t i = 0; i < dbufCount; ++i ) {
            const uint8_t uc = (uint8_t)dbuf[i];
            dbuf[byteCount[uc]] |= i << 8;
            byteCount[uc]++;
        }
        if ( tvecOut != NULL ) {
            memcpy( tvecOut, dbuf, sizeof( uint32_t ) * dbufCount );
        }
    }

    /* decodeBlock (bzip2.hpp:850-910) */
    {
        uint32_t crc = 0xFFFFFFFFU;
        uint32_t writeCount = info->n;
        uint32_t writePos = dbuf[info->orig_ptr];
        int writeCurrent = (int)( writePos & 0xFFU );
        writePos >>= 8;
        int writeRun = -1;
        uint64_t nOut = 0;
        uint32_t nPre = 0;
        while ( writeCount > 0 ) {
            writeCount--;
            const int previous = writeCurrent;
            writePos = dbuf[writePos];
            writeCurrent = (int)( writePos & 0xFFU );
            writePos >>= 8;
            if ( prerle != NULL ) {
                prerle[nPre] = (uint8_t)writeCurrent;
            }
            ++nPre;
            if ( writeRun < 3 ) {
                const uint8_t b = (uint8_t)writeCurrent;
                if ( ( out != NULL ) && ( nOut < out_cap ) ) {
                    out[nOut] = b;
                }
                ++nOut;
                crc = orc_crc32_update( crc, &b, 1 );
                if ( writeCurrent != previous ) {
                    writeRun = 0;
                } else {
                    ++writeRun;
                }
            } else {
                const uint8_t b = (uint8_t)previous;
                for ( int k = 0; k < writeCurrent; ++k ) {
                    if ( ( out != NULL ) && ( nOut < out_cap ) ) {
                        out[nOut] = b;
                    }
                    ++nOut;
                    crc = orc_crc32_update( crc, &b, 1 );
                }
                writeCurrent = -1;
                writeRun = 0;
            }
        }
        info->out_size = nOut;
        info->calc_crc = ~crc;
        if ( ( out != NULL ) && ( nOut > out_cap ) ) {
            status = ORC_ERR_OUTPUT_CAPACITY;
            goto done;
        }
        if ( info->calc_crc != info->header_crc ) {
            status = ORC_ERR_CRC;
            goto done;
        }
    }

done:
    free( dbuf );
    free( t );
    info->status = status;
    return status;
}

/* ------------------------------------------------------------------------------------------------
 * core/BlockMap.hpp:69-119,174-198 (push / finalize), only what the readers need
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    uint64_t* pairs; /* caller's buffer, may be NULL */
    size_t cap;
    size_t count;
    uint64_t lastEnc, lastDec; /* back() */
    uint64_t lastEncSize, lastDecSize;
    int error;
} blockmap;

static void
map_append( blockmap* m, uint64_t enc, uint64_t dec )
{
    if ( ( m->pairs != NULL ) && ( m->count < m->cap ) ) {
        m->pairs[2 * m->count] = enc;
        m->pairs[2 * m->count + 1] = dec;
    }
    m->count++;
    m->lastEnc = enc;
    m->lastDec = dec;
}

static void
map_push( blockmap* m, uint64_t enc, uint64_t encSize, uint64_t decSize )
{
    if ( m->count == 0 ) {
        map_append( m, enc, 0 );
    } else if ( enc > m->lastEnc ) {
        map_append( m, enc, m->lastDec + m->lastDecSize );
    } else {
        /* BlockMap.hpp:97-118: duplicates are tolerated, anything else throws; the readers below never
         * re-push, so every non-increasing offset is the "strictly increasing" invalid_argument. */
        m->error = 1;
        return;
    }
    m->lastDecSize = decSize;
    m->lastEncSize = encSize;
}

static void
map_finalize( blockmap* m )
{
    if ( m->count == 0 ) {
        map_append( m, 0, 0 ); /* BlockMap.hpp:184-187 */
    } else if ( ( m->lastEncSize != 0 ) || ( m->lastDecSize != 0 ) ) {
        map_append( m, m->lastEnc + m->lastEncSize, m->lastDec + m->lastDecSize ); /* :188-192 */
    }
    m->lastEncSize = 0;
    m->lastDecSize = 0;
}

int
orc_parallel_reader( const uint8_t* buf, size_t nbytes, uint8_t* out, uint64_t out_cap, uint64_t* out_size,
                     uint64_t* offsets, size_t offsets_cap_pairs, size_t* offsets_count, int* warnTrailing )
{
    blockmap map = { offsets, offsets_cap_pairs, 0, 0, 0, 0, 0, 0 };
    if ( warnTrailing != NULL ) {
        *warnTrailing = 0;
    }
    int status = ORC_OK;
    uint64_t total = 0;

    /* block finder: data-block magic only (ParallelBZ2Reader.hpp:55-64) */
    size_t nFound = orc_find_magic( buf, nbytes, ORC_MAGIC_BLOCK, NULL, 0 );
    uint64_t* found = (uint64_t*)malloc( sizeof( uint64_t ) * ( nFound + 1 ) );
    orc_find_magic( buf, nbytes, ORC_MAGIC_BLOCK, found, nFound );

    orc_block_info* info = (orc_block_info*)malloc( sizeof( orc_block_info ) );
    orc_block_info* next = (orc_block_info*)malloc( sizeof( orc_block_info ) );
    blocktables* scratch = (blocktables*)malloc( sizeof( blocktables ) );

    for ( size_t dataBlockIndex = 0; ; ++dataBlockIndex ) {
        if ( dataBlockIndex >= nFound ) {
            break; /* ParallelBZ2Reader.hpp:193-197 */
        }
        if ( dataBlockIndex == 0 ) {
            /* BZ2BlockFetcher ctor reads the stream header at bit 0 (BZ2BlockFetcher.hpp:55-56) */
            bitrd r = { buf, (uint64_t)nbytes * 8, 0, 0 };
            const int lvl = read_stream_header( &r );
            if ( lvl <= 0 ) {
                status = lvl < 0 ? ORC_ERR_EOF : ORC_ERR_STREAM_HEADER;
                break;
            }
        }
        const uint64_t off = found[dataBlockIndex];
        uint8_t* dst = ( out != NULL ) && ( total < out_cap ) ? out + total : NULL;
        const uint64_t dstCap = ( out != NULL ) && ( total < out_cap ) ? out_cap - total : 0;
        status = orc_decode_block( buf, nbytes, off, info, NULL, NULL, NULL, NULL, NULL, NULL,
                                   dst, dst != NULL ? dstCap : 0 );
        if ( ( status == ORC_ERR_OUTPUT_CAPACITY ) && ( out == NULL ) ) {
            status = ORC_OK;
        }
        if ( status != ORC_OK ) {
            break;
        }
        map_push( &map, info->enc_offset_bits, info->enc_size_bits, info->out_size );
        if ( map.error ) {
            status = ORC_ERR_BLOCK_ORDER;
            break;
        }
        total += info->out_size;

        /* probe the following header (ParallelBZ2Reader.hpp:207-238; BZ2BlockFetcher.hpp:64-82 builds a full
         * Block, i.e. parses symbol maps, selectors and trees of a following data block, too) */
        {
            bitrd r = { buf, (uint64_t)nbytes * 8, info->enc_offset_bits + info->enc_size_bits, 0 };
            memset( next, 0, sizeof( *next ) );
            const int st = parse_header( &r, next, scratch );
            if ( st != ORC_OK ) {
                status = st;
                break;
            }
            if ( next->is_eos ) {
                map_push( &map, next->enc_offset_bits, next->enc_size_bits, 0 );
                if ( map.error ) {
                    status = ORC_ERR_BLOCK_ORDER;
                    break;
                }
                const uint64_t nextStream = next->enc_offset_bits + next->enc_size_bits;
                if ( nextStream < (uint64_t)nbytes * 8 ) {
                    bitrd r2 = { buf, (uint64_t)nbytes * 8, nextStream, 0 };
                    if ( read_stream_header( &r2 ) <= 0 ) {
                        if ( warnTrailing != NULL ) {
                            *warnTrailing = 1;
                        }
                        nFound = dataBlockIndex + 1; /* m_blockFinder->finalize( dataBlockCount ), :234 */
                    }
                }
            }
        }
    }
    if ( status == ORC_OK ) {
        map_finalize( &map );
    }
    if ( out_size != NULL ) {
        *out_size = total;
    }
    if ( offsets_count != NULL ) {
        *offsets_count = map.count;
    }
    free( scratch );
    free( next );
    free( info );
    free( found );
    return status;
}

int
orc_serial_reader( const uint8_t* buf, size_t nbytes, uint8_t* out, uint64_t out_cap, uint64_t* out_size,
                   uint64_t* offsets, size_t offsets_cap_pairs, size_t* offsets_count, int* warnTrailing,
                   uint32_t* stream_crc )
{
    blockmap map = { offsets, offsets_cap_pairs, 0, 0, 0, 0, 0, 0 };
    if ( warnTrailing != NULL ) {
        *warnTrailing = 0;
    }
    int status = ORC_OK;
    uint64_t total = 0;
    const uint64_t nbits = (uint64_t)nbytes * 8;
    uint64_t pos = 0;
    int lastWasEos = 0;
    int atEof = 0;
    uint32_t calcStreamCrc = 0;
    orc_block_info* info = (orc_block_info*)malloc( sizeof( orc_block_info ) );

    /* BZ2Reader::read (BZ2Reader.hpp:233-264) */
    while ( ( pos < nbits ) && !atEof && ( status == ORC_OK ) ) {
        if ( ( pos == 0 ) || lastWasEos ) {
            bitrd r = { buf, nbits, pos, 0 };
            const int lvl = read_stream_header( &r );
            if ( lvl < 0 ) {
                status = ORC_ERR_EOF; /* EndOfFileReached is not a domain_error: propagates */
                break;
            }
            if ( lvl == 0 ) {
                if ( pos == 0 ) {
                    status = ORC_ERR_STREAM_HEADER;
                    break;
                }
                if ( warnTrailing != NULL ) {
                    *warnTrailing = 1;
                }
                break;
            }
            pos = r.pos;
            calcStreamCrc = 0; /* BZ2Reader.hpp:295-299 */
            lastWasEos = 0;
        }
        /* decodeStream (BZ2Reader.hpp:447-491) until the stream's EOS */
        while ( status == ORC_OK ) {
            /* readBlockHeader (BZ2Reader.hpp:395-424): the map entry is inserted before parsing */
            if ( ( map.pairs != NULL ) && ( map.count < map.cap ) ) {
                map.pairs[2 * map.count] = pos;
                map.pairs[2 * map.count + 1] = total;
            }
            map.count++;

            uint8_t* dst = ( out != NULL ) && ( total < out_cap ) ? out + total : NULL;
            const uint64_t dstCap = dst != NULL ? out_cap - total : 0;
            status = orc_decode_block( buf, nbytes, pos, info, NULL, NULL, NULL, NULL, NULL, NULL, dst, dstCap );
            if ( ( status == ORC_ERR_OUTPUT_CAPACITY ) && ( out == NULL ) ) {
                status = ORC_OK;
            }
            if ( status != ORC_OK ) {
                break;
            }
            pos = info->enc_offset_bits + info->enc_size_bits;
            if ( info->is_eos ) {
                if ( info->header_crc != calcStreamCrc ) {
                    status = ORC_ERR_STREAM_CRC;
                    break;
                }
                atEof = info->at_eof;
                lastWasEos = 1;
                break;
            }
            total += info->out_size;
            calcStreamCrc = orc_stream_crc_fold( calcStreamCrc, info->calc_crc );
        }
    }
    if ( out_size != NULL ) {
        *out_size = total;
    }
    if ( offsets_count != NULL ) {
        *offsets_count = map.count;
    }
    if ( stream_crc != NULL ) {
        *stream_crc = calcStreamCrc;
    }
    free( info );
    return status;
}
