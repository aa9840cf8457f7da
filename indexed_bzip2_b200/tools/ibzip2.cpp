/*
 * ibzip2 -- command-line front end over the GPU readers; same flags, outputs and exit codes as the reference's tool
 * (tools/ibzip2.cpp:172-504): -d -c -f -i -o -k -t -p -P -h -q -v -V -l -L --buffer-size.
 *   -l FILE   raw finder output: every block AND end-of-stream magic, one decimal bit offset per line (ibzip2.cpp:96-143)
 *   -L FILE   the block map as "<encoded bits>,<decoded bytes>" lines (ibzip2.cpp:82-93)
 *   -P 1      serial reader (BZ2Reader), any other value the parallel one (ibzip2.cpp:374-379)
 * The decoding happens on the GPU either way; -P only picks the reader flavour (and with it the block-map flavour).
 */
#include <algorithm>
#include <climits>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <string>
#include <string_view>
#include <thread>
#include <typeinfo>
#include <vector>

#include <fcntl.h>
#include <poll.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../host/BZ2Reader.hpp"
#include <cstdlib>

#include "../host/ParallelBZ2Reader.hpp"

using namespace ibz2host;

namespace
{
struct Arguments
{
    std::map<std::string, std::string> values;
    std::map<std::string, int> counts;
    std::vector<std::string> positional;

    [[nodiscard]] int count( const std::string& name ) const
    {
        const auto match = counts.find( name );
        return match == counts.end() ? 0 : match->second;
    }

    [[nodiscard]] std::string value( const std::string& name, const std::string& fallback = "" ) const
    {
        const auto match = values.find( name );
        return match == values.end() ? fallback : match->second;
    }
};

const std::map<char, std::string> SHORT_NAMES = {
    { 'c', "stdout" }, { 'd', "decompress" }, { 'f', "force" }, { 'i', "input" }, { 'o', "output" }, { 'k', "keep" },
    { 't', "test" }, { 'p', "block-finder-parallelism" }, { 'P', "decoder-parallelism" }, { 'h', "help" },
    { 'q', "quiet" }, { 'v', "verbose" }, { 'V', "version" }, { 'l', "list-compressed-offsets" }, { 'L', "list-offsets" },
};
const std::vector<std::string> WITH_VALUE = { "input", "output", "block-finder-parallelism", "decoder-parallelism",
                                              "list-compressed-offsets", "list-offsets", "buffer-size", "gpus" };

bool
takesValue( const std::string& name )
{
    return std::find( WITH_VALUE.begin(), WITH_VALUE.end(), name ) != WITH_VALUE.end();
}

Arguments
parseArguments( int argc, char** argv )
{
    Arguments result;
    for ( int i = 1; i < argc; ++i ) {
        const std::string argument = argv[i];
        const auto take = [&] ( const std::string& name, std::string inlineValue, bool hasInline ) {
            result.counts[name] += 1;
            if ( !takesValue( name ) ) {
                return;
            }
            if ( !hasInline ) {
                if ( i + 1 >= argc ) {
                    throw std::invalid_argument( "Option '" + name + "' is missing an argument" );
                }
                inlineValue = argv[++i];
            }
            result.values[name] = inlineValue;
        };
        if ( ( argument.size() > 2 ) && ( argument.rfind( "--", 0 ) == 0 ) ) {
            const auto equals = argument.find( '=' );
            const auto name = argument.substr( 2, equals == std::string::npos ? std::string::npos : equals - 2 );
            const bool known = takesValue( name ) || ( name == "oss-attributions" ) || ( name == "oss-attributions-yaml" )
                               || std::any_of( SHORT_NAMES.begin(), SHORT_NAMES.end(),
                                               [&] ( const auto& kv ) { return kv.second == name; } );
            if ( !known ) {
                throw std::invalid_argument( "Option '" + name + "' does not exist" );
            }
            take( name, equals == std::string::npos ? "" : argument.substr( equals + 1 ), equals != std::string::npos );
        } else if ( ( argument.size() > 1 ) && ( argument[0] == '-' ) && ( argument != "-" ) ) {
            for ( size_t k = 1; k < argument.size(); ++k ) {
                const auto match = SHORT_NAMES.find( argument[k] );
                if ( match == SHORT_NAMES.end() ) {
                    throw std::invalid_argument( std::string( "Option '" ) + argument[k] + "' does not exist" );
                }
                if ( takesValue( match->second ) ) {
                    const auto rest = argument.substr( k + 1 );
                    take( match->second, rest, !rest.empty() );
                    break;
                }
                take( match->second, "", false );
            }
        } else {
            result.positional.push_back( argument );
        }
    }
    if ( ( result.count( "input" ) == 0 ) && !result.positional.empty() ) {
        result.values["input"] = result.positional.front();
        result.counts["input"] = static_cast<int>( result.positional.size() );
    }
    return result;
}

/** '-' means standard output, which is spelled as the empty path (ibzip2.cpp getFilePath). */
std::string
getFilePath( const Arguments& arguments, const std::string& name )
{
    if ( arguments.count( name ) == 0 ) {
        return {};
    }
    const auto path = arguments.value( name );
    return path == "-" ? std::string() : path;
}

bool
fileExists( const std::string& path )
{
    struct stat st{};
    return !path.empty() && ( ::stat( path.c_str(), &st ) == 0 );
}

bool
stdinHasInput()
{
    pollfd fds{ STDIN_FILENO, POLLIN, 0 };
    return ::poll( &fds, 1, 100 ) == 1;
}

bool
endsWithNoCase( const std::string& text, const std::string& suffix )
{
    return ( text.size() >= suffix.size() )
           && std::equal( suffix.rbegin(), suffix.rend(), text.rbegin(),
                          [] ( char a, char b ) { return std::tolower( a ) == std::tolower( b ); } );
}

UniqueFileReader
openFileOrStdin( const std::string& path )
{
    if ( path.empty() ) {
        return std::make_unique<StandardFileReader>( STDIN_FILENO );
    }
    return std::make_unique<StandardFileReader>( path );
}

void
printHelp()
{
    std::cout <<
        "A bzip2 decompressor tool with the indexed_bzip2 interface, decoding on a B200 GPU\n"
        "Usage:\n  ibzip2 [OPTION...] positional parameters\n\n"
        " Decompression options:\n"
        "  -c, --stdout                  Output to standard output. This is the default, when reading from standard input.\n"
        "  -d, --decompress              Force decompression. Only for compatibility. No compression supported anyways.\n"
        "  -f, --force                   Force overwriting existing output files.\n"
        "  -i, --input arg               Input file. If none is given, data is read from standard input.\n"
        "  -o, --output arg              Output file. If none is given, use the input file name with '.bz2' stripped or\n"
        "                                '<input file>.out'.\n"
        "  -k, --keep                    Keep (do not delete) input file. Only for compatibility.\n"
        "  -t, --test                    Test compressed file integrity.\n"
        "  -p, --block-finder-parallelism arg  Only for compatibility: the block finder is one GPU kernel.\n"
        "  -P, --decoder-parallelism arg  1 selects the serial reader (BZ2Reader), anything else the parallel reader\n"
        "                                (ParallelBZ2Reader). The blocks are decoded on the GPU either way.\n"
        "      --gpus arg                Number of GPUs (devices 0 .. arg - 1) that share the decoding of the parallel\n"
        "                                reader. Not in the reference. Default: $IBZ2_GPUS, else 1.\n\n"
        " Output options:\n"
        "  -h, --help                    Print this help message.\n"
        "  -q, --quiet                   Suppress noncritical error messages.\n"
        "  -v, --verbose                 Be verbose.\n"
        "  -V, --version                 Display software version.\n"
        "  -l, --list-compressed-offsets arg  List only the bzip2 block offsets given in bits one per line to the\n"
        "                                specified output file. '-' writes to stdout.\n"
        "  -L, --list-offsets arg        List bzip2 block offsets in bits and the corresponding offsets in the decoded\n"
        "                                data in bytes as '<encoded bits>,<decoded bytes>' lines. '-' writes to stdout.\n\n"
        " Advanced options:\n"
        "      --buffer-size arg         Output buffer size between calls to the decoder and writing to the output.\n\n"
        "If no file names are given, ibzip2 decompresses from standard input to standard output.\n"
        "Examples:\n\n"
        "Decompress a file:\n  ibzip2 -d file.bz2\n\n"
        "Decompress a file in parallel:\n  ibzip2 -d -P 0 file.bz2\n\n"
        "List block offsets in bits:\n  ibzip2 -l blockoffsets.dat -- file.bz2\n\n"
        "List block offsets in bits and decoded bytes:\n  ibzip2 -L blockoffsets.dat -- file.bz2\n";
}

/** checkOffsets of ibzip2.cpp:30-65: every listed offset must carry one of the two magics. */
void
checkOffsets( const std::vector<uint8_t>& file, const std::vector<size_t>& offsets )
{
    for ( const auto offset : offsets ) {
        if ( offset + 48U > file.size() * CHAR_BIT ) {
            continue;   /* the end sentinel of the parallel block map */
        }
        uint64_t magic = 0;
        for ( unsigned bit = 0; bit < 48U; ++bit ) {
            const auto position = offset + bit;
            magic = ( magic << 1U ) | ( ( file[position / CHAR_BIT] >> ( 7U - position % CHAR_BIT ) ) & 1U );
        }
        if ( ( magic != IBZ2_MAGIC_BLOCK ) && ( magic != IBZ2_MAGIC_EOS ) ) {
            std::stringstream msg;
            msg << "Magic bytes at offset " << ( offset / CHAR_BIT ) << " B " << ( offset % CHAR_BIT ) << "b "
                << "do not match bzip2 magic bytes!";
            throw std::logic_error( msg.str() );
        }
    }
}

std::vector<uint8_t>
readWholeFile( const std::string& path )
{
    auto reader = openFileOrStdin( path );
    std::vector<uint8_t> data;
    std::vector<char> chunk( 1U << 20 );
    while ( true ) {
        const auto n = reader->read( chunk.data(), chunk.size() );
        if ( n == 0 ) {
            break;
        }
        data.insert( data.end(), chunk.begin(), chunk.begin() + static_cast<std::ptrdiff_t>( n ) );
    }
    return data;
}

template<typename Offsets>
void
dumpOffsets( std::ostream& out, const Offsets& offsets )
{
    if ( !out.good() ) {
        return;
    }
    for ( const auto& entry : offsets ) {
        if constexpr ( std::is_same_v<std::decay_t<decltype( entry )>, size_t>
                       || std::is_same_v<std::decay_t<decltype( entry )>, uint64_t> ) {
            out << entry << "\n";
        } else {
            out << entry.first << "," << entry.second << "\n";
        }
    }
}

/** findCompressedBlocks of ibzip2.cpp:96-143: both magics, sorted, straight from the finder kernel. */
void
findCompressedBlocks( const std::string& inputFilePath, const std::string& outputFilePath, bool test, bool verbose )
{
    const auto file = readWholeFile( inputFilePath );
    std::vector<uint64_t> found( 4096 );
    uint64_t count = 0;
    ibz2_ctx* ctx = nullptr;
    if ( ibz2_create( 0, &ctx ) != IBZ2_OK ) {
        throw std::runtime_error( std::string( "Cannot open the CUDA bzip2 decoder: " ) + ibz2_last_error( nullptr ) );
    }
    while ( true ) {
        const int rc = ibz2_find_magics( ctx, file.data(), file.size(), IBZ2_MEM_HOST, 0, IBZ2_FIND_BLOCK | IBZ2_FIND_EOS,
                                         found.data(), found.size(), &count );
        if ( rc == IBZ2_E_CAPACITY ) {
            found.resize( count );
            continue;
        }
        if ( rc != IBZ2_OK ) {
            const std::string message = ibz2_last_error( ctx );
            ibz2_destroy( ctx );
            throw std::runtime_error( "ibz2_find_magics failed: " + message );
        }
        break;
    }
    ibz2_destroy( ctx );
    std::vector<size_t> offsets;
    offsets.reserve( count );
    for ( uint64_t i = 0; i < count; ++i ) {
        offsets.push_back( found[i] & ~IBZ2_OFFSET_EOS_FLAG );
    }
    std::sort( offsets.begin(), offsets.end() );
    if ( test ) {
        checkOffsets( file, offsets );
    }
    if ( outputFilePath.empty() ) {
        dumpOffsets( std::cout, offsets );
    } else {
        std::ofstream out( outputFilePath );
        dumpOffsets( out, offsets );
    }
    if ( verbose ) {
        std::cout << "Found " << offsets.size() << " blocks\n";
    }
}
}  // namespace

int
ibzip2CLI( int argc, char** argv )
{
    const auto arguments = parseArguments( argc, argv );
    const bool force = arguments.count( "force" ) > 0;
    const bool quiet = arguments.count( "quiet" ) > 0;
    const bool test = arguments.count( "test" ) > 0;
    const bool verbose = arguments.count( "verbose" ) > 0;
    const auto cores = std::max( 1U, std::thread::hardware_concurrency() );
    const auto parallelismOf = [&] ( const std::string& name, unsigned fallback ) {
        const auto p = static_cast<unsigned>( std::stoul( arguments.value( name, std::to_string( fallback ) ) ) );
        return p > 0 ? p : cores;
    };
    parallelismOf( "block-finder-parallelism", 1 );
    const auto decoderParallelism = parallelismOf( "decoder-parallelism", 0 );

    if ( arguments.count( "help" ) > 0 ) {
        printHelp();
        return 0;
    }
    if ( arguments.count( "version" ) > 0 ) {
        std::cout << "ibzip2, CLI to the indexed and seekable bzip2 decoding library indexed-bzip2 version 1.7.0.\n";
        return 0;
    }
    if ( arguments.count( "input" ) > 1 ) {
        std::cerr << "One or none bzip2 filename to decompress must be specified!\n";
        return 1;
    }
    if ( ( arguments.count( "input" ) != 1 ) && !stdinHasInput() ) {
        std::cerr << "Either stdin must have input, e.g., by piping to it, or an input file must be specified!\n";
        return 1;
    }
    const std::string inputFilePath = arguments.count( "input" ) == 1 ? arguments.value( "input" ) : std::string();

    auto outputFilePath = getFilePath( arguments, "output" );
    if ( ( arguments.count( "stdout" ) == 0 ) && outputFilePath.empty() && !inputFilePath.empty() ) {
        if ( endsWithNoCase( inputFilePath, ".bz2" ) ) {
            outputFilePath = inputFilePath.substr( 0, inputFilePath.size() - 4 );
        } else {
            outputFilePath = inputFilePath + ".out";
            if ( !quiet ) {
                std::cerr << "Could not deduce output file name. Will write to '" << outputFilePath << "'\n";
            }
        }
    }

    const bool listCompressedOffsets = arguments.count( "list-compressed-offsets" ) > 0;
    const bool listOffsets = arguments.count( "list-offsets" ) > 0;
    const bool decompress = arguments.count( "decompress" ) > 0;

    if ( decompress && ( outputFilePath != "/dev/null" ) && fileExists( outputFilePath ) && !force ) {
        std::cerr << "Output file '" << outputFilePath << "' already exists! Use --force to overwrite.\n";
        return 1;
    }
    const auto bufferSize = static_cast<unsigned>( std::stoul( arguments.value( "buffer-size", "0" ) ) );
    const auto offsetsFilePath = getFilePath( arguments, "list-offsets" );
    if ( !offsetsFilePath.empty() && fileExists( offsetsFilePath ) && !force ) {
        std::cerr << "Output file for offsets'" << offsetsFilePath
                  << "' for offsets already exists! Use --force to overwrite.\n";
        return 1;
    }
    const auto compressedOffsetsFilePath = getFilePath( arguments, "list-compressed-offsets" );
    if ( !compressedOffsetsFilePath.empty() && fileExists( compressedOffsetsFilePath ) && !force ) {
        std::cerr << "Output file compressed offsets '" << compressedOffsetsFilePath
                  << "' for offsets already exists! Use --force to overwrite.\n";
        return 1;
    }

    if ( decompress || listOffsets ) {
        if ( verbose ) {
            std::cerr << "Decompress " << inputFilePath << " -> " << outputFilePath << " with " << decoderParallelism
                      << " threads\n";
        }
        auto fileReader = openFileOrStdin( inputFilePath );

        int outputFileDescriptor = -1;
        bool ownsOutput = false;
        if ( decompress ) {
            if ( outputFilePath.empty() ) {
                outputFileDescriptor = STDOUT_FILENO;
            } else {
                outputFileDescriptor = ::open( outputFilePath.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644 );
                if ( outputFileDescriptor < 0 ) {
                    std::cerr << "Could not open output file: " << outputFilePath << " for writing!\n";
                    return 1;
                }
                ownsOutput = true;
            }
        }
        const bool writingToStdout = decompress && outputFilePath.empty();

        std::unique_ptr<BZ2ReaderInterface> reader;
        if ( decoderParallelism == 1 ) {
            reader = std::make_unique<BZ2Reader>( std::move( fileReader ) );
        } else {
            /* --gpus N (or $IBZ2_GPUS): devices 0 .. N - 1 share every large batch; new next to the reference */
            std::vector<int> devices;
            {
                const char* const env = std::getenv( "IBZ2_GPUS" );
                const auto gpus = std::stoul( arguments.value( "gpus", env != nullptr ? env : "1" ) );
                for ( unsigned long d = 0; d < std::max( 1UL, gpus ); ++d ) {
                    devices.push_back( static_cast<int>( d ) );
                }
            }
            auto parallelReader = std::make_unique<ParallelBZ2Reader>( std::move( fileReader ), decoderParallelism, devices );
            /* the reference's parallel reader never looks at the stream CRC; this one does unless told not to */
            if ( const char* const env = std::getenv( "IBZ2_VERIFY_STREAM_CRC" ); ( env != nullptr ) && ( env[0] == '0' ) ) {
                parallelReader->setVerifyStreamCRC( false );
            }
            reader = std::move( parallelReader );
        }
        reader->setShowProfileOnDestruction( verbose );

        size_t nBytesWrittenTotal = 0;
        if ( bufferSize > 0 ) {
            do {
                std::vector<char> buffer( bufferSize, 0 );
                const auto nBytesRead = reader->read( -1, buffer.data(), buffer.size() );
                if ( ( outputFileDescriptor >= 0 ) && ( writeAll( outputFileDescriptor, nullptr, buffer.data(), nBytesRead ) != 0 ) ) {
                    std::cerr << "Could not write all the decoded data to the specified output!\n";
                }
                nBytesWrittenTotal += nBytesRead;
            } while ( !reader->eof() );
        } else {
            nBytesWrittenTotal = reader->read( outputFileDescriptor );
        }
        auto& out = writingToStdout ? std::cerr : std::cout;
        if ( ownsOutput ) {
            ::close( outputFileDescriptor );
        }

        const auto offsets = reader->blockOffsets();
        std::vector<size_t> compressedOffsets;
        compressedOffsets.reserve( offsets.size() );
        for ( const auto& [offset, _] : offsets ) {
            compressedOffsets.push_back( offset );
        }
        if ( verbose ) {
            out << "Found " << offsets.size() << " blocks\n";
        }
        if ( test ) {
            checkOffsets( readWholeFile( inputFilePath ), compressedOffsets );
            const auto readerSize = reader->size();
            if ( !readerSize ) {
                throw std::logic_error( "Bzip2 reader size should be available at this point!" );
            }
            if ( nBytesWrittenTotal != *readerSize ) {
                std::stringstream msg;
                msg << "Wrote less bytes (" << nBytesWrittenTotal << " B) than decoded stream is large("
                    << *readerSize << " B)!";
                throw std::logic_error( msg.str() );
            }
        }
        if ( listOffsets ) {
            if ( !offsetsFilePath.empty() ) {
                std::ofstream file( offsetsFilePath );
                dumpOffsets( file, offsets );
            } else if ( outputFilePath.empty() ) {
                dumpOffsets( std::cerr, offsets );
            } else {
                dumpOffsets( std::cout, offsets );
            }
        }
        if ( listCompressedOffsets ) {
            if ( !compressedOffsetsFilePath.empty() ) {
                std::ofstream file( compressedOffsetsFilePath );
                dumpOffsets( file, compressedOffsets );
            } else if ( outputFilePath.empty() ) {
                dumpOffsets( std::cerr, compressedOffsets );
            } else {
                dumpOffsets( std::cout, compressedOffsets );
            }
        }
        return 0;
    }

    if ( listCompressedOffsets ) {
        if ( verbose ) {
            std::cerr << "Find block offsets\n";
        }
        findCompressedBlocks( inputFilePath, compressedOffsetsFilePath, test, verbose );
        return 0;
    }

    std::cerr << "No suitable arguments were given. Please refer to the help!\n\n";
    printHelp();
    return 1;
}

#ifndef IBZ2_NO_MAIN
int
main( int argc, char** argv )
{
    try {
        return ibzip2CLI( argc, argv );
    } catch ( const EndOfFileReached& ) {
        std::cerr << "Unexpected end of file. Truncated or invalid bzip2?\n";
        return 1;
    } catch ( const std::exception& exception ) {
        const std::string_view message{ exception.what() };
        if ( message.empty() ) {
            std::cerr << "Caught exception with typeid: " << typeid( exception ).name() << "\n";
        } else {
            std::cerr << "Caught exception: " << message << "\n";
        }
        return 1;
    }
    return 1;
}
#endif
