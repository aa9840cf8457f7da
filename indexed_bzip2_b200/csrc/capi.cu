/*
 * capi.cu -- the C ABI declared in include/ibz2_cuda.h: context, scratch memory, stream orchestration.
 * All bzip2 arithmetic lives in the kernels (blockfinder.cu, huffman_decode.cu, bwt_prepare.cu, bwt_traverse.cu,
 * rle.cu, crc.cu); this file only moves buffers, launches, and converts BlockMeta into ibz2_block_result.
 */
#include "common.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <new>
#include <string>
#include <thread>
#include <utility>
#include <vector>

namespace
{
thread_local std::string g_createError;
}

struct ibz2_ctx
{
    int device = 0;
    int sms = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t copyStream = nullptr;   /* device -> host copies of finished waves, overlapped with the next wave */
    cudaStream_t inStream = nullptr;     /* host -> device copy of the second half of the input, overlapped with wave 1 */
    cudaEvent_t inFirstHalf = nullptr;
    cudaEvent_t inSecondHalf = nullptr;
    uint64_t pipelineCopied = 0;         /* output bytes already on their way to the host (split-input calls) */
    uint64_t* hRunning = nullptr;        /* pinned: output bytes produced up to the end of the last finished wave */
    void* hUpload[2] = { nullptr, nullptr };          /* pinned staging of ibz2_upload_pageable */
    cudaEvent_t uploadDone[2] = { nullptr, nullptr };
    std::string lastError;
    uint32_t batchBlocks = 0;
    bool verifyStreamCrc = true;

    /* per-wave scratch */
    uint32_t scratchBlocks = 0;
    ibz2::BlockMeta* dMeta = nullptr;
    uint8_t* dLbytes = nullptr;
    uint32_t* dDbuf = nullptr;
    uint8_t* dSelectors = nullptr;
    ibz2::BlockTables* dTables = nullptr;
    uint8_t* dStage = nullptr;           /* K4b staging of ONE sub-wave (subWaveBlocks()), reused by the next */
    uint32_t stageBlocks = 0;
    uint32_t* dCounter = nullptr;
    uint32_t* dWarpBase = nullptr;
    uint32_t* dPieceStart = nullptr;
    uint64_t* dRunning = nullptr;
    ibz2::BlockMeta* hMeta = nullptr; /* pinned, scratchBlocks entries */
    uint32_t* hOrder = nullptr;       /* pinned, scratchBlocks entries: K2b's block order of the wave */

    /* staging for host-side inputs / outputs */
    uint8_t* dIn = nullptr;
    uint64_t dInCapacity = 0;
    uint8_t* dOut = nullptr;
    uint64_t dOutCapacity = 0;

    /* finder */
    uint64_t* dFound = nullptr;
    uint32_t* dFoundCount = nullptr;
    uint32_t foundCapacity = 0;

    std::vector<cudaEvent_t> events;
    cudaEvent_t timerStart = nullptr;
    cudaEvent_t timerStop = nullptr;
    size_t eventsUsed = 0;
    ibz2_timings timings{};
};

namespace
{
/* what the launch wrappers size their grids with: the SM count of the context whose call runs on this thread (contexts on
 * devices with different SM counts may live in one process; every entry point sets it before it launches anything) */
thread_local int g_smCount = 148;

#define IBZ2_CUDA_CHECK( ctx, call )                                                                          \
    do {                                                                                                      \
        const cudaError_t err__ = ( call );                                                                   \
        if ( err__ != cudaSuccess ) {                                                                         \
            ( ctx )->lastError = std::string( #call ) + ": " + cudaGetErrorString( err__ );                   \
            return err__ == cudaErrorMemoryAllocation ? IBZ2_E_OUT_OF_MEMORY : IBZ2_E_CUDA;                   \
        }                                                                                                     \
    } while ( false )

/* an event of the pool that the callee records.  Events only carry timings: when one cannot be created the first event
 * of the pool stands in (the timings of that call are then meaningless, its results are not), and the error stays
 * readable through ibz2_last_error. */
cudaEvent_t
spareEvent( ibz2_ctx* ctx )
{
    if ( ctx->eventsUsed == ctx->events.size() ) {
        cudaEvent_t e = nullptr;
        const cudaError_t err = cudaEventCreate( &e );
        if ( err != cudaSuccess ) {
            ctx->lastError = std::string( "cudaEventCreate: " ) + cudaGetErrorString( err );
            cudaGetLastError();
            return ctx->events.empty() ? nullptr : ctx->events.front();
        }
        ctx->events.push_back( e );
    }
    return ctx->events[ctx->eventsUsed++];
}

cudaEvent_t
nextEvent( ibz2_ctx* ctx )
{
    cudaEvent_t e = spareEvent( ctx );
    if ( e != nullptr ) {
        cudaEventRecord( e, ctx->stream );
    }
    return e;
}

float
elapsed( cudaEvent_t a, cudaEvent_t b )
{
    float ms = 0;
    if ( ( a == nullptr ) || ( b == nullptr ) || ( cudaEventElapsedTime( &ms, a, b ) != cudaSuccess ) ) {
        cudaGetLastError();
        return 0;
    }
    return ms;
}

void
freeScratch( ibz2_ctx* ctx )
{
    cudaFree( ctx->dMeta );
    cudaFree( ctx->dLbytes );
    cudaFree( ctx->dDbuf );
    cudaFree( ctx->dSelectors );
    cudaFree( ctx->dTables );
    cudaFree( ctx->dStage );
    cudaFree( ctx->dCounter );
    cudaFree( ctx->dWarpBase );
    cudaFree( ctx->dPieceStart );
    cudaFreeHost( ctx->hMeta );
    cudaFreeHost( ctx->hOrder );
    ctx->dMeta = nullptr;
    ctx->dLbytes = nullptr;
    ctx->dDbuf = nullptr;
    ctx->dSelectors = nullptr;
    ctx->dTables = nullptr;
    ctx->dStage = nullptr;
    ctx->stageBlocks = 0;
    ctx->dCounter = nullptr;
    ctx->dWarpBase = nullptr;
    ctx->dPieceStart = nullptr;
    ctx->hMeta = nullptr;
    ctx->hOrder = nullptr;
    ctx->scratchBlocks = 0;
}

/* K4..K6 run over at most this many blocks at a time: an eighth of a wave when finished output travels to the host
 * meanwhile (the copy nothing hides is then one sub-wave), half a wave otherwise (every sub-wave ends with a round of
 * partly idle SMs: 6 ms per call on the 8 GB workload with eighths).  The K4b staging area holds one sub-wave. */
uint32_t
subWaveBlocks( const ibz2_ctx* ctx, bool pipelined = false )
{
    return std::max<uint32_t>( 1U, ctx->batchBlocks / ( pipelined ? 8U : 2U ) );
}

int
ensureStage( ibz2_ctx* ctx, uint32_t blocks )
{
    if ( blocks <= ctx->stageBlocks ) {
        return IBZ2_OK;
    }
    cudaFree( ctx->dStage );
    ctx->dStage = nullptr;
    ctx->stageBlocks = 0;
    IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dStage, ibz2::TRAV_STAGE_HEADER + static_cast<uint64_t>( blocks ) * ibz2::TRAV_STAGE_STRIDE ) );
    ctx->stageBlocks = blocks;
    return IBZ2_OK;
}

int
ensureScratch( ibz2_ctx* ctx, uint32_t blocks )
{
    if ( blocks <= ctx->scratchBlocks ) {
        return ensureStage( ctx, std::min( blocks, subWaveBlocks( ctx ) ) );
    }
    freeScratch( ctx );
    const uint64_t b = blocks;
    IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dMeta, b * sizeof( ibz2::BlockMeta ) ) );
    IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dLbytes, b * ibz2::LBYTES_STRIDE ) );
    IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dDbuf, b * ibz2::DBUF_STRIDE * sizeof( uint32_t ) ) );
    IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dSelectors, b * ibz2::SEL_STRIDE ) );
    IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dTables, b * sizeof( ibz2::BlockTables ) ) );
    IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dCounter, ( b + 16 ) * sizeof( uint32_t ) ) );
    IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dWarpBase, b * ibz2::RLE_RANGES * sizeof( uint32_t ) ) );
    IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dPieceStart, ( b + 1 ) * sizeof( uint32_t ) ) );
    IBZ2_CUDA_CHECK( ctx, cudaMallocHost( &ctx->hMeta, b * sizeof( ibz2::BlockMeta ) ) );
    IBZ2_CUDA_CHECK( ctx, cudaMallocHost( &ctx->hOrder, b * sizeof( uint32_t ) ) );
    ctx->scratchBlocks = blocks;
    return ensureStage( ctx, std::min( blocks, subWaveBlocks( ctx ) ) );
}

int
ensureInput( ibz2_ctx* ctx, uint64_t nbytes )
{
    if ( nbytes + 256U <= ctx->dInCapacity ) {
        return IBZ2_OK;
    }
    cudaFree( ctx->dIn );
    ctx->dIn = nullptr;
    ctx->dInCapacity = 0;
    IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dIn, nbytes + 256U ) );
    ctx->dInCapacity = nbytes + 256U;
    return IBZ2_OK;
}

int
ensureOutput( ibz2_ctx* ctx, uint64_t nbytes )
{
    if ( nbytes + 256U <= ctx->dOutCapacity ) {
        return IBZ2_OK;
    }
    cudaFree( ctx->dOut );
    ctx->dOut = nullptr;
    ctx->dOutCapacity = 0;
    IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dOut, nbytes + 256U ) );
    ctx->dOutCapacity = nbytes + 256U;
    return IBZ2_OK;
}

void
fillResult( const ibz2::BlockMeta& m, ibz2_block_result* r )
{
    r->enc_offset_bits = m.bitOffset;
    r->enc_size_bits = m.endBit >= m.bitOffset ? m.endBit - m.bitOffset : 0;
    r->out_offset = m.outOffset;
    r->out_size = ( ( m.status == IBZ2_ST_OK ) || ( m.status == IBZ2_ST_OUTPUT_CAPACITY ) || ( m.status == IBZ2_ST_CRC ) )
                  ? m.outSize : 0;
    r->stream_flags = m.streamFlags & ( IBZ2_STREAM_FIRST_BLOCK | IBZ2_STREAM_LAST_BLOCK );
    r->stream_crc_expected = m.eosCrc;
    r->stream_crc_calculated = 0;
    r->n_coded = m.nCoded;
    r->reserved = 0;
    r->expected_crc = m.headerCrc;
    r->calculated_crc = m.calcCrc;
    r->status = m.status;
    r->is_eos = m.isEos;
    r->is_eof = m.isEof;
    r->n_symbols = m.n;
    r->orig_ptr = m.origPtr;
}

/* MSB-first bit extraction on the host, for the two header probes only. */
struct HostBits
{
    const uint8_t* data;
    uint64_t nbits;
    uint64_t pos;
    bool eof;

    uint64_t
    read( unsigned n )
    {
        if ( pos + n > nbits ) {
            eof = true;
            pos = nbits;
            return 0;
        }
        uint64_t v = 0;
        for ( unsigned i = 0; i < n; ++i, ++pos ) {
            v = ( v << 1 ) | ( ( data[pos >> 3] >> ( 7U - ( pos & 7U ) ) ) & 1U );
        }
        return v;
    }
};
}  // namespace

namespace ibz2
{
int
smCount()
{
    return g_smCount;
}

/* what ibz2_last_error( NULL ) returns on this thread, for entry points that have no context (blockfinder.cu) */
void
setContextFreeError( const char* message )
{
    g_createError = message;
}
}  // namespace ibz2

extern "C" {

int
ibz2_abi_version( void )
{
    return IBZ2_ABI_VERSION;
}

int
ibz2_device_count( void )
{
    int n = 0;
    if ( cudaGetDeviceCount( &n ) != cudaSuccess ) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

const char*
ibz2_last_error( const ibz2_ctx* ctx )
{
    return ctx != nullptr ? ctx->lastError.c_str() : g_createError.c_str();
}

int
ibz2_create( int device, ibz2_ctx** out )
{
    if ( out == nullptr ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    *out = nullptr;
    int n = 0;
    cudaError_t err = cudaGetDeviceCount( &n );
    if ( ( err != cudaSuccess ) || ( n == 0 ) ) {
        cudaGetLastError();
        g_createError = std::string( "no usable CUDA device (there is no CPU fallback): " )
                        + ( err != cudaSuccess ? cudaGetErrorString( err ) : "device count is 0" );
        return IBZ2_E_NO_DEVICE;
    }
    if ( ( device < 0 ) || ( device >= n ) ) {
        g_createError = "device index out of range";
        return IBZ2_E_INVALID_ARGUMENT;
    }
    ibz2_ctx* ctx = new ( std::nothrow ) ibz2_ctx;
    if ( ctx == nullptr ) {
        return IBZ2_E_OUT_OF_MEMORY;
    }
    ctx->device = device;
    cudaDeviceProp prop{};
    /* whatever has been created when a step fails is released by ibz2_destroy, which looks at every handle */
    if ( ( ( err = cudaSetDevice( device ) ) != cudaSuccess )
         || ( ( err = cudaGetDeviceProperties( &prop, device ) ) != cudaSuccess ) ) {
        g_createError = std::string( "CUDA initialisation failed: " ) + cudaGetErrorString( err );
        cudaGetLastError();
        ibz2_destroy( ctx );
        return IBZ2_E_CUDA;
    }
    if ( prop.major < 10 ) {
        g_createError = "this library is built for sm_100a (Blackwell B200) only; device is sm_"
                        + std::to_string( prop.major ) + std::to_string( prop.minor );
        ibz2_destroy( ctx );
        return IBZ2_E_NO_DEVICE;
    }
    if ( ( ( err = cudaStreamCreateWithFlags( &ctx->stream, cudaStreamNonBlocking ) ) != cudaSuccess )
         || ( ( err = cudaStreamCreateWithFlags( &ctx->copyStream, cudaStreamNonBlocking ) ) != cudaSuccess )
         || ( ( err = cudaStreamCreateWithFlags( &ctx->inStream, cudaStreamNonBlocking ) ) != cudaSuccess )
         || ( ( err = cudaEventCreateWithFlags( &ctx->inFirstHalf, cudaEventDefault ) ) != cudaSuccess )
         || ( ( err = cudaEventCreateWithFlags( &ctx->inSecondHalf, cudaEventDefault ) ) != cudaSuccess )
         || ( ( err = cudaMallocHost( &ctx->hRunning, 16 * sizeof( uint64_t ) ) ) != cudaSuccess )
         || ( ( err = cudaMalloc( &ctx->dRunning, 2 * sizeof( uint64_t ) ) ) != cudaSuccess ) ) {
        g_createError = std::string( "CUDA initialisation failed: " ) + cudaGetErrorString( err );
        cudaGetLastError();
        ibz2_destroy( ctx );
        return IBZ2_E_CUDA;
    }
    ctx->sms = prop.multiProcessorCount;
    g_smCount = ctx->sms;
    ctx->batchBlocks = static_cast<uint32_t>( ctx->sms ) * 64U;   /* one round of K2b's 64 decoders per SM */
    ibz2::initCrcConstants();
    if ( ( err = cudaGetLastError() ) != cudaSuccess ) {
        g_createError = std::string( "constant upload failed: " ) + cudaGetErrorString( err );
        ibz2_destroy( ctx );
        return IBZ2_E_CUDA;
    }
    *out = ctx;
    return IBZ2_OK;
}

void
ibz2_destroy( ibz2_ctx* ctx )
{
    if ( ctx == nullptr ) {
        return;
    }
    cudaSetDevice( ctx->device );
    g_smCount = ctx->sms;
    if ( ctx->stream != nullptr ) {
        cudaStreamSynchronize( ctx->stream );
    }
    freeScratch( ctx );
    cudaFree( ctx->dIn );
    cudaFree( ctx->dOut );
    cudaFree( ctx->dFound );
    cudaFree( ctx->dFoundCount );
    cudaFree( ctx->dRunning );
    cudaFreeHost( ctx->hRunning );
    for ( int i = 0; i < 2; ++i ) {
        cudaFreeHost( ctx->hUpload[i] );
        if ( ctx->uploadDone[i] != nullptr ) {
            cudaEventDestroy( ctx->uploadDone[i] );
        }
    }
    if ( ctx->copyStream != nullptr ) {
        cudaStreamDestroy( ctx->copyStream );
    }
    if ( ctx->inStream != nullptr ) {
        cudaStreamDestroy( ctx->inStream );
    }
    if ( ctx->inFirstHalf != nullptr ) {
        cudaEventDestroy( ctx->inFirstHalf );
    }
    if ( ctx->inSecondHalf != nullptr ) {
        cudaEventDestroy( ctx->inSecondHalf );
    }
    for ( cudaEvent_t e : ctx->events ) {
        cudaEventDestroy( e );
    }
    if ( ctx->timerStart != nullptr ) {
        cudaEventDestroy( ctx->timerStart );
    }
    if ( ctx->timerStop != nullptr ) {
        cudaEventDestroy( ctx->timerStop );
    }
    if ( ctx->stream != nullptr ) {
        cudaStreamDestroy( ctx->stream );
    }
    cudaGetLastError();   /* a half-built context may have handed null handles to the calls above */
    delete ctx;
}

int
ibz2_get_timings( const ibz2_ctx* ctx, ibz2_timings* timings )
{
    if ( ( ctx == nullptr ) || ( timings == nullptr ) ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    *timings = ctx->timings;
    return IBZ2_OK;
}

int
ibz2_set_batch_blocks( ibz2_ctx* ctx, uint32_t blocks )
{
    if ( ( ctx == nullptr ) || ( blocks == 0 ) ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    /* several kernels launch with dim3( x, blocks of a sub-wave ), and a sub-wave is up to half a wave: gridDim.y <= 65535 */
    ctx->batchBlocks = std::min<uint32_t>( blocks, 2U * 65535U );
    return IBZ2_OK;
}

int
ibz2_set_verify_stream_crc( ibz2_ctx* ctx, int verify )
{
    if ( ctx == nullptr ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    ctx->verifyStreamCrc = verify != 0;
    return IBZ2_OK;
}

uint64_t
ibz2_fold_stream_crcs( ibz2_block_result* results, uint64_t nblocks, int verify )
{
    /* BZ2Reader.hpp:481-484 per data block, :406-416 at the end-of-stream block.  A fold is only trusted from a block that
     * directly follows a stream header on, over blocks that follow one another without a gap and decoded cleanly. */
    uint64_t mismatches = 0;
    bool tracking = false;
    uint32_t crc = 0;
    uint64_t expectedNext = 0;
    for ( uint64_t i = 0; i < nblocks; ++i ) {
        ibz2_block_result& r = results[i];
        r.stream_flags &= ~IBZ2_STREAM_CRC_CHECKED;
        if ( r.is_eos != 0 ) {
            tracking = false;
            continue;
        }
        if ( ( r.stream_flags & IBZ2_STREAM_FIRST_BLOCK ) != 0 ) {
            tracking = true;
            crc = 0;
        } else if ( tracking && ( r.enc_offset_bits != expectedNext ) ) {
            tracking = false;
        }
        const bool clean = ( ( r.status == IBZ2_ST_OK ) || ( r.status == IBZ2_ST_OUTPUT_CAPACITY ) || ( r.status == IBZ2_ST_STREAM_CRC ) )
                           && ( r.calculated_crc == r.expected_crc );
        tracking = tracking && clean;
        if ( tracking ) {
            crc = ibz2_stream_crc_fold( crc, r.calculated_crc );
            expectedNext = r.enc_offset_bits + r.enc_size_bits;
            if ( ( r.stream_flags & IBZ2_STREAM_LAST_BLOCK ) != 0 ) {
                r.stream_flags |= IBZ2_STREAM_CRC_CHECKED;
                r.stream_crc_calculated = crc;
                if ( crc != r.stream_crc_expected ) {
                    ++mismatches;
                    if ( ( verify != 0 ) && ( r.status == IBZ2_ST_OK ) ) {
                        r.status = IBZ2_ST_STREAM_CRC;
                    }
                }
                tracking = false;
            }
        }
    }
    return mismatches;
}

int
ibz2_device_alloc( ibz2_ctx* ctx, uint64_t nbytes, void** dptr )
{
    if ( ( ctx == nullptr ) || ( dptr == nullptr ) ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    cudaSetDevice( ctx->device );
    g_smCount = ctx->sms;
    IBZ2_CUDA_CHECK( ctx, cudaMalloc( dptr, nbytes + 256U ) );
    return IBZ2_OK;
}

int
ibz2_device_free( ibz2_ctx* ctx, void* dptr )
{
    if ( ctx == nullptr ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    cudaSetDevice( ctx->device );
    g_smCount = ctx->sms;
    IBZ2_CUDA_CHECK( ctx, cudaFree( dptr ) );
    return IBZ2_OK;
}

int
ibz2_host_alloc_pinned( ibz2_ctx* ctx, uint64_t nbytes, void** hptr )
{
    if ( ( ctx == nullptr ) || ( hptr == nullptr ) ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    cudaSetDevice( ctx->device );
    g_smCount = ctx->sms;
    IBZ2_CUDA_CHECK( ctx, cudaHostAlloc( hptr, nbytes > 0 ? nbytes : 1, cudaHostAllocPortable ) );
    return IBZ2_OK;
}

int
ibz2_host_free_pinned( ibz2_ctx* ctx, void* hptr )
{
    if ( ctx == nullptr ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    IBZ2_CUDA_CHECK( ctx, cudaFreeHost( hptr ) );
    return IBZ2_OK;
}

int
ibz2_memcpy_h2d( ibz2_ctx* ctx, void* dptr, const void* hptr, uint64_t nbytes )
{
    if ( ctx == nullptr ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    cudaSetDevice( ctx->device );
    g_smCount = ctx->sms;
    IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( dptr, hptr, nbytes, cudaMemcpyHostToDevice, ctx->stream ) );
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->stream ) );
    return IBZ2_OK;
}

/* Pageable host memory (a memory-mapped file, a vector) -> device through two pinned staging buffers: host threads fill
 * one while the copy engine empties the other.  A plain cudaMemcpy from pageable memory does the same thing with one
 * thread and small buffers (about 6 GB/s here); this reaches what the page cache and PCIe give. */
int
ibz2_upload_pageable( ibz2_ctx* ctx, void* dptr, const void* hptr, uint64_t nbytes )
{
    if ( ( ctx == nullptr ) || ( ( nbytes > 0 ) && ( ( dptr == nullptr ) || ( hptr == nullptr ) ) ) ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    cudaSetDevice( ctx->device );
    g_smCount = ctx->sms;
    constexpr uint64_t CHUNK = 32ULL << 20;
    constexpr unsigned THREADS = 4;
    if ( nbytes <= ( 4ULL << 20 ) ) {
        return ibz2_memcpy_h2d( ctx, dptr, hptr, nbytes );
    }
    if ( ctx->hUpload[0] == nullptr ) {
        for ( int i = 0; i < 2; ++i ) {
            IBZ2_CUDA_CHECK( ctx, cudaMallocHost( &ctx->hUpload[i], CHUNK ) );
            IBZ2_CUDA_CHECK( ctx, cudaEventCreateWithFlags( &ctx->uploadDone[i], cudaEventDisableTiming ) );
        }
    }
    const auto* const src = static_cast<const uint8_t*>( hptr );
    auto* const dst = static_cast<uint8_t*>( dptr );
    uint64_t k = 0;
    for ( uint64_t offset = 0; offset < nbytes; offset += CHUNK, ++k ) {
        const uint64_t n = std::min<uint64_t>( CHUNK, nbytes - offset );
        const int slot = static_cast<int>( k & 1U );
        if ( k >= 2 ) {
            IBZ2_CUDA_CHECK( ctx, cudaEventSynchronize( ctx->uploadDone[slot] ) );   /* the copy that last read this buffer */
        }
        auto* const staging = static_cast<uint8_t*>( ctx->hUpload[slot] );
        std::vector<std::thread> workers;
        const uint64_t piece = ( n + THREADS - 1 ) / THREADS;
        for ( unsigned t = 1; t < THREADS; ++t ) {
            const uint64_t begin = std::min<uint64_t>( n, piece * t );
            const uint64_t end = std::min<uint64_t>( n, piece * ( t + 1 ) );
            if ( end > begin ) {
                workers.emplace_back( [=] () { std::memcpy( staging + begin, src + offset + begin, end - begin ); } );
            }
        }
        std::memcpy( staging, src + offset, std::min<uint64_t>( n, piece ) );
        for ( auto& w : workers ) {
            w.join();
        }
        IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( dst + offset, staging, n, cudaMemcpyHostToDevice, ctx->inStream ) );
        IBZ2_CUDA_CHECK( ctx, cudaEventRecord( ctx->uploadDone[slot], ctx->inStream ) );
    }
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->inStream ) );
    return IBZ2_OK;
}

int
ibz2_memcpy_d2h( ibz2_ctx* ctx, void* hptr, const void* dptr, uint64_t nbytes )
{
    if ( ctx == nullptr ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    cudaSetDevice( ctx->device );
    g_smCount = ctx->sms;
    IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( hptr, dptr, nbytes, cudaMemcpyDeviceToHost, ctx->stream ) );
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->stream ) );
    return IBZ2_OK;
}

/* device -> pinned host on the context's copy stream, without waiting: the building block of a consumer that hands a
 * decoded batch on piece by piece (one piece travels while the previous one is written out) */
int
ibz2_memcpy_d2h_async( ibz2_ctx* ctx, void* hptr, const void* dptr, uint64_t nbytes )
{
    if ( ctx == nullptr ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    cudaSetDevice( ctx->device );
    IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( hptr, dptr, nbytes, cudaMemcpyDeviceToHost, ctx->copyStream ) );
    return IBZ2_OK;
}

int
ibz2_copy_synchronize( ibz2_ctx* ctx )
{
    if ( ctx == nullptr ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    cudaSetDevice( ctx->device );
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->copyStream ) );
    return IBZ2_OK;
}

int
ibz2_synchronize( ibz2_ctx* ctx )
{
    if ( ctx == nullptr ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    cudaSetDevice( ctx->device );
    g_smCount = ctx->sms;
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->stream ) );
    return IBZ2_OK;
}

/* ---------------- internal building blocks of the three decode entry points ---------------- */

/* Makes the compressed bytes available on the device: copies host input into the context's staging buffer. */
static int
stageInput( ibz2_ctx* ctx, const void* data, uint64_t nbytes, int data_mem, const uint8_t** dData, const char* who )
{
    *dData = static_cast<const uint8_t*>( data );
    if ( data_mem == IBZ2_MEM_HOST ) {
        const int rc = ensureInput( ctx, nbytes );
        if ( rc != IBZ2_OK ) {
            return rc;
        }
        if ( nbytes > 0 ) {
            IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( ctx->dIn, data, nbytes, cudaMemcpyHostToDevice, ctx->stream ) );
        }
        *dData = ctx->dIn;
    } else if ( ( reinterpret_cast<uintptr_t>( data ) & 15U ) != 0 ) {
        ctx->lastError = std::string( who ) + ": device input must be 16-byte aligned";
        return IBZ2_E_INVALID_ARGUMENT;
    }
    return IBZ2_OK;
}

/* K1 on device-resident bytes; `host` receives the matches sorted by bit offset (EOS flag in bit 63). */
static int
findOnDevice( ibz2_ctx* ctx, const uint8_t* dData, uint64_t nbytes, uint32_t which, std::vector<uint64_t>& host )
{
    host.clear();
    if ( nbytes < 6 ) {
        return IBZ2_OK;
    }
    /* candidates are rare (one per block plus ~2^-48 per bit of noise); grow and rescan if a file is adversarial */
    uint32_t wanted = std::max<uint32_t>( ctx->foundCapacity, 1U << 16 );
    uint32_t found = 0;
    while ( true ) {
        if ( wanted > ctx->foundCapacity ) {
            cudaFree( ctx->dFound );
            cudaFree( ctx->dFoundCount );
            ctx->dFound = nullptr;
            ctx->dFoundCount = nullptr;
            ctx->foundCapacity = 0;
            IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dFound, static_cast<uint64_t>( wanted ) * sizeof( uint64_t ) ) );
            IBZ2_CUDA_CHECK( ctx, cudaMalloc( &ctx->dFoundCount, sizeof( uint32_t ) ) );
            ctx->foundCapacity = wanted;
        }
        IBZ2_CUDA_CHECK( ctx, cudaMemsetAsync( ctx->dFoundCount, 0, sizeof( uint32_t ), ctx->stream ) );
        const cudaEvent_t e0 = nextEvent( ctx );
        ibz2::launchFindMagics( dData, nbytes, which, ctx->dFound, ctx->dFoundCount, ctx->foundCapacity, ctx->stream );
        ctx->timings.launches += 1;
        const cudaEvent_t e1 = nextEvent( ctx );
        IBZ2_CUDA_CHECK( ctx, cudaGetLastError() );
        IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( &found, ctx->dFoundCount, sizeof( uint32_t ), cudaMemcpyDeviceToHost,
                                               ctx->stream ) );
        IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->stream ) );
        ctx->timings.find_ms += elapsed( e0, e1 );
        if ( found <= ctx->foundCapacity ) {
            break;
        }
        wanted = found;
    }
    host.resize( found );
    if ( found > 0 ) {
        IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( host.data(), ctx->dFound, static_cast<uint64_t>( found ) * sizeof( uint64_t ),
                                               cudaMemcpyDeviceToHost, ctx->stream ) );
        IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->stream ) );
    }
    /* ascending by bit offset (the EOS flag is bit 63, so mask it for the comparison) */
    std::sort( host.begin(), host.end(), [] ( uint64_t a, uint64_t b ) {
        return ( a & ~IBZ2_OFFSET_EOS_FLAG ) < ( b & ~IBZ2_OFFSET_EOS_FLAG );
    } );
    return IBZ2_OK;
}

/* K2..K6 over device-resident input, in waves of at most ctx->batchBlocks blocks; dOut may be null (sizes only). */
static int
decodeOnDevice( ibz2_ctx* ctx, const uint8_t* dData, uint64_t nbytes, const uint64_t* offsets, uint64_t nblocks,
                uint8_t* dOut, uint64_t outCapacity, ibz2_block_result* results, uint64_t* totalOut,
                void* hostOut = nullptr, bool continueOutput = false, bool finishCopy = true,
                const std::function<int()>* afterFirstUpload = nullptr )
{
    const bool haveOutput = dOut != nullptr;
    uint32_t wave = static_cast<uint32_t>( std::min<uint64_t>( nblocks, ctx->batchBlocks ) );
    const bool pipelined = haveOutput && ( hostOut != nullptr );
    /* output bytes already on their way to the host; a split-input call decodes its halves with two calls of this
     * function that share one output: the second goes on where the first stopped */
    uint64_t copied = continueOutput ? ctx->pipelineCopied : 0;
    if ( wave > ctx->scratchBlocks ) {
        /* scratch is about 4.7 MB per block in flight plus the staging of one sub-wave: stay inside what the device has left */
        size_t freeBytes = 0;
        size_t totalBytes = 0;
        if ( cudaMemGetInfo( &freeBytes, &totalBytes ) == cudaSuccess ) {
            const uint64_t perBlock = ibz2::LBYTES_STRIDE + ibz2::DBUF_STRIDE * sizeof( uint32_t ) + ibz2::SEL_STRIDE
                                      + sizeof( ibz2::BlockTables ) + 4096U;
            const uint64_t reusable = static_cast<uint64_t>( ctx->scratchBlocks ) * perBlock;
            const uint64_t stageNeed = ctx->stageBlocks >= subWaveBlocks( ctx ) ? 0U
                                       : static_cast<uint64_t>( subWaveBlocks( ctx ) ) * ibz2::TRAV_STAGE_STRIDE;
            const uint64_t usable = static_cast<uint64_t>( freeBytes ) + reusable;
            const uint64_t fit = ( usable > stageNeed ? usable - stageNeed : 0U ) / 10U * 9U / perBlock;
            wave = static_cast<uint32_t>( std::max<uint64_t>( 1, std::min<uint64_t>( wave, fit ) ) );
        }
    }
    {
        const int rc = ensureScratch( ctx, wave );
        if ( rc != IBZ2_OK ) {
            return rc;
        }
    }
    cudaStream_t s = ctx->stream;
    if ( !continueOutput ) {
        IBZ2_CUDA_CHECK( ctx, cudaMemsetAsync( ctx->dRunning, 0, 2 * sizeof( uint64_t ), s ) );
    }

    struct WaveEvents { cudaEvent_t e[7]; cudaEvent_t bits[3]; cudaEvent_t walk; };
    std::vector<WaveEvents> waveEvents;   /* K2 per wave: e[0], e[1], bits */
    std::vector<WaveEvents> subEvents;    /* K4..K6 per sub-wave: e[1]..e[6], walk */

    for ( uint64_t first = 0; first < nblocks; first += wave ) {
        const uint32_t count = static_cast<uint32_t>( std::min<uint64_t>( wave, nblocks - first ) );
        /* the previous wave's results must have left the pinned mirror before it is refilled */
        IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( s ) );
        if ( pipelined && ( ( first > 0 ) || continueOutput ) ) {
            const uint64_t produced = std::min<uint64_t>( ctx->hRunning[0], outCapacity );
            if ( produced > copied ) {
                IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( static_cast<uint8_t*>( hostOut ) + copied, dOut + copied, produced - copied,
                                                       cudaMemcpyDeviceToHost, ctx->copyStream ) );
                copied = produced;
            }
        }
        if ( first > 0 ) {
            for ( uint32_t i = 0; i < wave; ++i ) {
                fillResult( ctx->hMeta[i], &results[first - wave + i] );
            }
        }
        std::memset( ctx->hMeta, 0, sizeof( ibz2::BlockMeta ) * count );
        for ( uint32_t i = 0; i < count; ++i ) {
            ctx->hMeta[i].bitOffset = offsets[first + i];
        }
        IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( ctx->dMeta, ctx->hMeta, sizeof( ibz2::BlockMeta ) * count,
                                               cudaMemcpyHostToDevice, s ) );
        {
            /* K2b hands its decoders the blocks of the wave longest first (compressed size stands in for the symbol
             * count, which nobody knows yet): one block is one sequential chain, so the wave ends when the decoder
             * with the most work ends */
            std::vector<std::pair<uint64_t, uint32_t>> bySize( count );
            for ( uint32_t i = 0; i < count; ++i ) {
                const uint64_t begin = offsets[first + i] & ~IBZ2_OFFSET_EOS_FLAG;
                const uint64_t end = ( first + i + 1 < nblocks ) ? ( offsets[first + i + 1] & ~IBZ2_OFFSET_EOS_FLAG ) : nbytes * 8U;
                bySize[i] = { end > begin ? end - begin : 0U, i };
            }
            std::sort( bySize.begin(), bySize.end(), [] ( const auto& a, const auto& b ) {
                return ( a.first != b.first ) ? ( a.first > b.first ) : ( a.second < b.second );
            } );
            for ( uint32_t i = 0; i < count; ++i ) {
                ctx->hOrder[i] = bySize[i].second;
            }
            IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( ctx->dCounter + ibz2::K2B_ORDER_OFFSET, ctx->hOrder, sizeof( uint32_t ) * count,
                                                   cudaMemcpyHostToDevice, s ) );
        }
        if ( ( first == 0 ) && ( afterFirstUpload != nullptr ) ) {
            /* bulk host -> device traffic of the caller goes behind this wave's own small uploads in the copy engine's
             * queue, not in front of them */
            const int rc = ( *afterFirstUpload )();
            if ( rc != IBZ2_OK ) {
                return rc;
            }
        }
        WaveEvents we{};
        we.e[0] = nextEvent( ctx );
        we.bits[0] = spareEvent( ctx );
        we.bits[1] = spareEvent( ctx );
        we.bits[2] = spareEvent( ctx );
        ibz2::launchHuffmanDecode( dData, nbytes, ctx->dMeta, count, ctx->dLbytes, ctx->dSelectors, ctx->dTables,
                                   ctx->dDbuf, ctx->dCounter, true, s, we.bits[0], we.bits[1], we.bits[2] );
        we.e[1] = nextEvent( ctx );
        waveEvents.push_back( we );
        ctx->timings.launches += ibz2::HUFFMAN_LAUNCHES;

        /* K4..K6 in sub-waves: with host output a finished sub-wave travels while the next one is computed, and the part
         * of the copy that nothing hides shrinks to one sub-wave */
        /* with host output at least eight sub-waves whatever the wave holds (the copy of a sub-wave hides behind the
         * next one's kernels: a rank of an 8-GPU job has 1133 blocks, fewer than one fixed sub-wave), but not so small
         * that the SMs idle between launches */
        const uint32_t sub = pipelined ? std::min<uint32_t>( subWaveBlocks( ctx, true ),
                                                             std::max<uint32_t>( ( count + 7U ) / 8U, static_cast<uint32_t>( ctx->sms ) ) )
                                       : std::min<uint32_t>( count, subWaveBlocks( ctx, false ) );
        cudaEvent_t previousDone = nullptr;
        uint32_t previousSlot = 0;
        uint32_t k = 0;
        for ( uint32_t subFirst = 0; subFirst < count; subFirst += sub, ++k ) {
            const uint32_t subCount = std::min<uint32_t>( sub, count - subFirst );
            ibz2::BlockMeta* const meta = ctx->dMeta + subFirst;
            uint8_t* const lbytes = ctx->dLbytes + static_cast<uint64_t>( subFirst ) * ibz2::LBYTES_STRIDE;
            uint32_t* const dbuf = ctx->dDbuf + static_cast<uint64_t>( subFirst ) * ibz2::DBUF_STRIDE;
            uint8_t* const stage = ctx->dStage;   /* every sub-wave stages in the same place */
            uint32_t* const warpBase = ctx->dWarpBase + static_cast<uint64_t>( subFirst ) * ibz2::RLE_RANGES;
            WaveEvents se{};
            se.e[1] = nextEvent( ctx );
            se.walk = spareEvent( ctx );
            ibz2::launchBwtPrepare( lbytes, dbuf, meta, subCount, s );
            se.e[2] = nextEvent( ctx );
            ibz2::launchBwtTraverse( dbuf, lbytes, meta, subCount, stage, s, se.walk );
            se.e[3] = nextEvent( ctx );
            ibz2::launchRleScan( lbytes, dbuf, warpBase, meta, subCount, s );
            ibz2::launchOutputLayout( meta, subCount, ctx->dRunning, outCapacity, haveOutput ? 1 : 0,
                                      ctx->dPieceStart, ibz2::CRC_PIECE_BYTES, s );
            se.e[4] = nextEvent( ctx );
            ctx->timings.launches += 7;
            if ( haveOutput ) {
                ibz2::launchRleWrite( lbytes, dbuf, warpBase, meta, subCount, dOut, s );
                se.e[5] = nextEvent( ctx );
                ibz2::launchCrc( dOut, meta, subCount, ctx->dPieceStart, ibz2::CRC_PIECE_BYTES, s );
                ibz2::launchCrcFinalize( dOut, meta, subCount, s );
                se.e[6] = nextEvent( ctx );
                ctx->timings.launches += 3;
            } else {
                se.e[5] = se.e[4];
                se.e[6] = se.e[4];
            }
            subEvents.push_back( se );
            if ( pipelined ) {
                /* the host learns how far the output has grown; it looks only after the NEXT sub-wave is queued */
                const uint32_t slot = 1U + ( k % 8U );
                IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( ctx->hRunning + slot, ctx->dRunning, sizeof( uint64_t ), cudaMemcpyDeviceToHost, s ) );
                const cudaEvent_t done = spareEvent( ctx );
                IBZ2_CUDA_CHECK( ctx, cudaEventRecord( done, s ) );
                if ( previousDone != nullptr ) {
                    IBZ2_CUDA_CHECK( ctx, cudaEventSynchronize( previousDone ) );
                    const uint64_t produced = std::min<uint64_t>( ctx->hRunning[previousSlot], outCapacity );
                    if ( produced > copied ) {
                        IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( static_cast<uint8_t*>( hostOut ) + copied, dOut + copied, produced - copied,
                                                               cudaMemcpyDeviceToHost, ctx->copyStream ) );
                        copied = produced;
                    }
                }
                previousDone = done;
                previousSlot = slot;
            }
        }
        IBZ2_CUDA_CHECK( ctx, cudaGetLastError() );
        IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( ctx->hMeta, ctx->dMeta, sizeof( ibz2::BlockMeta ) * count,
                                               cudaMemcpyDeviceToHost, s ) );
        if ( pipelined ) {
            IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( ctx->hRunning, ctx->dRunning, sizeof( uint64_t ), cudaMemcpyDeviceToHost, s ) );
        }
    }
    uint64_t total = 0;
    IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( &total, ctx->dRunning, sizeof( uint64_t ), cudaMemcpyDeviceToHost, s ) );
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( s ) );
    {
        const uint64_t lastFirst = ( ( nblocks - 1 ) / wave ) * wave;
        const uint32_t count = static_cast<uint32_t>( nblocks - lastFirst );
        for ( uint32_t i = 0; i < count; ++i ) {
            fillResult( ctx->hMeta[i], &results[lastFirst + i] );
        }
    }
    *totalOut = total;
    ctx->pipelineCopied = copied;
    if ( pipelined && finishCopy ) {
        const cudaEvent_t e0 = nextEvent( ctx );   /* on the decode stream: what is left of the copy after the last kernel */
        const uint64_t produced = std::min<uint64_t>( total, outCapacity );
        if ( produced > copied ) {
            IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( static_cast<uint8_t*>( hostOut ) + copied, dOut + copied, produced - copied,
                                                   cudaMemcpyDeviceToHost, ctx->copyStream ) );
        }
        IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->copyStream ) );
        const cudaEvent_t e1 = nextEvent( ctx );
        IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( s ) );
        ctx->timings.d2h_ms = elapsed( e0, e1 );
    }
    for ( const auto& we : waveEvents ) {
        ctx->timings.huffman_ms += elapsed( we.e[0], we.e[1] );
        ctx->timings.huffman_bits_ms += elapsed( we.bits[0], we.bits[1] );
        ctx->timings.mtf_chunk_ms += elapsed( we.bits[1], we.bits[2] );
    }
    for ( const auto& se : subEvents ) {
        ctx->timings.bwt_prepare_ms += elapsed( se.e[1], se.e[2] );
        ctx->timings.bwt_traverse_ms += elapsed( se.e[2], se.e[3] );
        ctx->timings.bwt_walk_ms += elapsed( se.e[2], se.walk );
        ctx->timings.rle_scan_ms += elapsed( se.e[3], se.e[4] );
        if ( haveOutput ) {
            ctx->timings.rle_write_ms += elapsed( se.e[4], se.e[5] );
            ctx->timings.crc_ms += elapsed( se.e[5], se.e[6] );
        }
    }
    return IBZ2_OK;
}

/* Resolves the output buffer: device pointer as is, host pointer -> the context's device staging buffer. */
static int
stageOutput( ibz2_ctx* ctx, void* out, uint64_t out_capacity, int out_mem, uint8_t** dOut, const char* who )
{
    *dOut = static_cast<uint8_t*>( out );
    if ( out == nullptr ) {
        return IBZ2_OK;
    }
    if ( out_mem == IBZ2_MEM_HOST ) {
        const int rc = ensureOutput( ctx, out_capacity );
        if ( rc != IBZ2_OK ) {
            return rc;
        }
        *dOut = ctx->dOut;
    } else if ( ( reinterpret_cast<uintptr_t>( out ) & 15U ) != 0 ) {
        ctx->lastError = std::string( who ) + ": device output must be 16-byte aligned";
        return IBZ2_E_INVALID_ARGUMENT;
    }
    return IBZ2_OK;
}

static int
copyOutputToHost( ibz2_ctx* ctx, void* out, uint64_t out_capacity, int out_mem, uint64_t total )
{
    const cudaEvent_t e0 = nextEvent( ctx );
    if ( ( out != nullptr ) && ( out_mem == IBZ2_MEM_HOST ) ) {
        const uint64_t n = std::min<uint64_t>( total, out_capacity );
        if ( n > 0 ) {
            IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( out, ctx->dOut, n, cudaMemcpyDeviceToHost, ctx->stream ) );
        }
    }
    const cudaEvent_t e1 = nextEvent( ctx );
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->stream ) );
    ctx->timings.d2h_ms = elapsed( e0, e1 );
    return IBZ2_OK;
}

static void
beginCall( ibz2_ctx* ctx )
{
    cudaSetDevice( ctx->device );
    g_smCount = ctx->sms;
    g_smCount = ctx->sms;
    ctx->timings = ibz2_timings{};
    ctx->eventsUsed = 0;
}

int
ibz2_find_magics( ibz2_ctx* ctx, const void* data, uint64_t nbytes, int data_mem, uint64_t base_bit_offset,
                  uint32_t which, uint64_t* offsets, uint64_t capacity, uint64_t* count )
{
    if ( ( ctx == nullptr ) || ( count == nullptr ) || ( ( data == nullptr ) && ( nbytes > 0 ) )
         || ( ( which & ( IBZ2_FIND_BLOCK | IBZ2_FIND_EOS ) ) == 0 ) || ( ( offsets == nullptr ) && ( capacity > 0 ) ) ) {
        if ( ctx != nullptr ) {
            ctx->lastError = "ibz2_find_magics: invalid argument";
        }
        return IBZ2_E_INVALID_ARGUMENT;
    }
    *count = 0;
    beginCall( ctx );
    if ( nbytes < 6 ) {
        return IBZ2_OK;
    }
    const cudaEvent_t e0 = nextEvent( ctx );
    const uint8_t* dData = nullptr;
    int rc = stageInput( ctx, data, nbytes, data_mem, &dData, "ibz2_find_magics" );
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    const cudaEvent_t e1 = nextEvent( ctx );
    std::vector<uint64_t> host;
    rc = findOnDevice( ctx, dData, nbytes, which, host );
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    const cudaEvent_t e2 = nextEvent( ctx );
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->stream ) );
    ctx->timings.h2d_ms = elapsed( e0, e1 );
    ctx->timings.total_ms = elapsed( e0, e2 );

    *count = host.size();
    const uint64_t n = std::min<uint64_t>( host.size(), capacity );
    for ( uint64_t i = 0; i < n; ++i ) {
        offsets[i] = ( ( host[i] & ~IBZ2_OFFSET_EOS_FLAG ) + base_bit_offset ) | ( host[i] & IBZ2_OFFSET_EOS_FLAG );
    }
    return host.size() > capacity ? IBZ2_E_CAPACITY : IBZ2_OK;
}

int
ibz2_decode_blocks( ibz2_ctx* ctx, const void* data, uint64_t nbytes, int data_mem,
                    const uint64_t* block_bit_offsets, uint64_t nblocks,
                    void* out, uint64_t out_capacity, int out_mem,
                    ibz2_block_result* results, uint64_t* total_out )
{
    if ( ( ctx == nullptr ) || ( ( data == nullptr ) && ( nbytes > 0 ) ) || ( ( block_bit_offsets == nullptr ) && ( nblocks > 0 ) )
         || ( ( results == nullptr ) && ( nblocks > 0 ) ) ) {
        if ( ctx != nullptr ) {
            ctx->lastError = "ibz2_decode_blocks: invalid argument";
        }
        return IBZ2_E_INVALID_ARGUMENT;
    }
    if ( total_out != nullptr ) {
        *total_out = 0;
    }
    beginCall( ctx );
    if ( nblocks == 0 ) {
        return IBZ2_OK;
    }
    const cudaEvent_t eStart = nextEvent( ctx );
    const uint8_t* dData = nullptr;
    int rc = stageInput( ctx, data, nbytes, data_mem, &dData, "ibz2_decode_blocks" );
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    const cudaEvent_t eH2D = nextEvent( ctx );
    uint8_t* dOut = nullptr;
    rc = stageOutput( ctx, out, out_capacity, out_mem, &dOut, "ibz2_decode_blocks" );
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    uint64_t total = 0;
    rc = decodeOnDevice( ctx, dData, nbytes, block_bit_offsets, nblocks, dOut, out_capacity, results, &total );
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    ibz2_fold_stream_crcs( results, nblocks, ctx->verifyStreamCrc ? 1 : 0 );
    if ( total_out != nullptr ) {
        *total_out = total;
    }
    rc = copyOutputToHost( ctx, out, out_capacity, out_mem, total );
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    ctx->timings.h2d_ms = elapsed( eStart, eH2D );
    ctx->timings.total_ms = elapsed( eStart, ctx->events[ctx->eventsUsed - 1] );
    return IBZ2_OK;
}

/* Host-to-host calls of at least IBZ2_SPLIT_INPUT_BYTES input bytes (default 4 GiB) upload the input in two parts, the
 * first one IBZ2_SPLIT_PERCENT of it (default 25), and decode the blocks of the first part while the second one travels.
 * It hides most of the upload, but K2b then runs once per part, and a K2b launch never takes less than its longest
 * block (~50 ms), so it only pays when the second part still fills every decoder slot more than once.  Measured
 * through bench.py's e2e (ms per call, no split / 25 %): 8 GB raw 539 / 520 (20 %: 529, 30 %: 541, 50 %: 566),
 * 4 GB raw 302 / 320, 2 GB raw 187 / 209 -- hence the threshold; with several GPUs per file each rank's window decides. */
static uint64_t
splitInputBytes()
{
    const char* const env = std::getenv( "IBZ2_SPLIT_INPUT_BYTES" );
    return env != nullptr ? std::strtoull( env, nullptr, 10 ) : ( 4ULL << 30 );
}

/* ibz2_decode_buffer for host input and host output: input in two halves.  *done = false (with IBZ2_OK) asks the caller
 * to run the plain path instead. */
static int
decodeBufferSplit( ibz2_ctx* ctx, const uint8_t* data, uint64_t nbytes, uint64_t first_bit, uint64_t last_bit,
                   void* out, uint64_t out_capacity, ibz2_block_result* results, uint64_t results_capacity,
                   uint64_t* nblocks, uint64_t* total_out, bool* done )
{
    *done = false;
    cudaStream_t s = ctx->stream;
    int rc = ensureInput( ctx, nbytes );
    if ( rc == IBZ2_OK ) {
        rc = ensureOutput( ctx, out_capacity );
    }
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    const char* const percentEnv = std::getenv( "IBZ2_SPLIT_PERCENT" );
    const uint64_t percent = percentEnv != nullptr
                             ? std::min<uint64_t>( 95U, std::max<uint64_t>( 5U, std::strtoull( percentEnv, nullptr, 10 ) ) )
                             : 25U;
    const uint64_t half = ( nbytes / 100U * percent ) & ~127ULL;
    const cudaEvent_t e0 = nextEvent( ctx );
    IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( ctx->dIn, data, half, cudaMemcpyHostToDevice, s ) );
    IBZ2_CUDA_CHECK( ctx, cudaEventRecord( ctx->inFirstHalf, s ) );
    const cudaEvent_t e1 = nextEvent( ctx );
    bool secondHalfQueued = false;
    /* whichever way this function is left, nothing reads the caller's input or writes the caller's output afterwards */
    struct DrainCopies
    {
        ibz2_ctx* ctx;
        const bool* queued;
        ~DrainCopies()
        {
            if ( *queued ) {
                cudaStreamSynchronize( ctx->inStream );
            }
            cudaStreamSynchronize( ctx->copyStream );
        }
    } drainCopies{ ctx, &secondHalfQueued };
    const std::function<int()> queueSecondHalf = [&] () -> int {
        if ( !secondHalfQueued ) {
            secondHalfQueued = true;
            IBZ2_CUDA_CHECK( ctx, cudaStreamWaitEvent( ctx->inStream, ctx->inFirstHalf, 0 ) );
            IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( ctx->dIn + half, data + half, nbytes - half, cudaMemcpyHostToDevice,
                                                   ctx->inStream ) );
            IBZ2_CUDA_CHECK( ctx, cudaEventRecord( ctx->inSecondHalf, ctx->inStream ) );
        }
        return IBZ2_OK;
    };

    /* ---- first half: every block whose successor also starts inside it is complete ---- */
    std::vector<uint64_t> foundA;
    rc = findOnDevice( ctx, ctx->dIn, half, IBZ2_FIND_BLOCK, foundA );
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    const bool haveCarry = !foundA.empty();
    const uint64_t carry = haveCarry ? foundA.back() : 0;   /* may reach into the second half: decoded with it */
    std::vector<uint64_t> mineA;
    for ( size_t i = 0; i + 1 < foundA.size(); ++i ) {
        if ( ( foundA[i] >= first_bit ) && ( foundA[i] < last_bit ) ) {
            mineA.push_back( foundA[i] );
        }
    }
    if ( mineA.size() > results_capacity ) {
        ctx->lastError = "ibz2_decode_buffer: results array too small";
        *nblocks = mineA.size();
        return IBZ2_E_CAPACITY;
    }
    uint64_t total = 0;
    ctx->pipelineCopied = 0;
    IBZ2_CUDA_CHECK( ctx, cudaMemsetAsync( ctx->dRunning, 0, 2 * sizeof( uint64_t ), s ) );
    ctx->hRunning[0] = 0;
    if ( !mineA.empty() ) {
        rc = decodeOnDevice( ctx, ctx->dIn, half, mineA.data(), mineA.size(), ctx->dOut, out_capacity, results, &total, out,
                             true, false, &queueSecondHalf );
        if ( rc != IBZ2_OK ) {
            cudaStreamSynchronize( ctx->inStream );
            return rc;
        }
        for ( size_t i = 0; i < mineA.size(); ++i ) {
            if ( results[i].status != IBZ2_ST_OK ) {
                IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->copyStream ) );
                IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->inStream ) );
                return IBZ2_OK;   /* *done stays false */
            }
        }
    }
    rc = queueSecondHalf();   /* no block in the first half: nothing was queued yet */
    if ( rc != IBZ2_OK ) {
        return rc;
    }

    /* ---- second half ---- */
    const cudaEvent_t eAfterA = nextEvent( ctx );
    IBZ2_CUDA_CHECK( ctx, cudaStreamWaitEvent( s, ctx->inSecondHalf, 0 ) );
    const uint64_t lead = std::min<uint64_t>( half, 128U );   /* a magic may straddle the cut */
    std::vector<uint64_t> foundB;
    rc = findOnDevice( ctx, ctx->dIn + ( half - lead ), nbytes - ( half - lead ), IBZ2_FIND_BLOCK, foundB );
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    std::vector<uint64_t> mineB;
    if ( haveCarry && ( carry >= first_bit ) && ( carry < last_bit ) ) {
        mineB.push_back( carry );
    }
    for ( const uint64_t f : foundB ) {
        const uint64_t bit = f + ( half - lead ) * 8U;
        if ( ( !haveCarry || ( bit > carry ) ) && ( bit >= first_bit ) && ( bit < last_bit ) ) {
            mineB.push_back( bit );
        }
    }
    *nblocks = mineA.size() + mineB.size();
    if ( *nblocks > results_capacity ) {
        cudaStreamSynchronize( ctx->copyStream );
        ctx->lastError = "ibz2_decode_buffer: results array too small";
        return IBZ2_E_CAPACITY;
    }
    const cudaEvent_t eBeforeB = nextEvent( ctx );
    if ( !mineB.empty() ) {
        rc = decodeOnDevice( ctx, ctx->dIn, nbytes, mineB.data(), mineB.size(), ctx->dOut, out_capacity,
                             results + mineA.size(), &total, out, true, true );
        if ( rc != IBZ2_OK ) {
            return rc;
        }
    } else if ( !mineA.empty() ) {
        /* nothing in the second half: the first half's last wave still has to reach the host */
        const uint64_t produced = std::min<uint64_t>( total, out_capacity );
        if ( produced > ctx->pipelineCopied ) {
            IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( static_cast<uint8_t*>( out ) + ctx->pipelineCopied, ctx->dOut + ctx->pipelineCopied,
                                                   produced - ctx->pipelineCopied, cudaMemcpyDeviceToHost, ctx->copyStream ) );
        }
        IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->copyStream ) );
    }
    if ( total_out != nullptr ) {
        *total_out = total;
    }
    ctx->timings.h2d_ms = elapsed( e0, e1 );
    if ( std::getenv( "IBZ2_DEBUG_TIMING" ) != nullptr ) {
        const cudaEvent_t eAfterB = nextEvent( ctx );
        cudaEventSynchronize( eAfterB );
        std::fprintf( stderr, "[ibz2] split input: first half %.1f ms, second half %.1f ms, second half done %.1f ms after the call began; "
                              "first half decoded at %.1f ms, second half started at %.1f ms, all done at %.1f ms (%zu + %zu blocks)\n",
                      elapsed( e0, e1 ), elapsed( ctx->inFirstHalf, ctx->inSecondHalf ), elapsed( e0, ctx->inSecondHalf ),
                      elapsed( e0, eAfterA ), elapsed( e0, eBeforeB ), elapsed( e0, eAfterB ), mineA.size(), mineB.size() );
    }
    ibz2_fold_stream_crcs( results, *nblocks, ctx->verifyStreamCrc ? 1 : 0 );
    *done = true;
    return IBZ2_OK;
}

int
ibz2_decode_buffer( ibz2_ctx* ctx, const void* data, uint64_t nbytes, int data_mem,
                    uint64_t first_bit, uint64_t last_bit,
                    void* out, uint64_t out_capacity, int out_mem,
                    ibz2_block_result* results, uint64_t results_capacity, uint64_t* nblocks, uint64_t* total_out )
{
    if ( ( ctx == nullptr ) || ( nblocks == nullptr ) || ( ( data == nullptr ) && ( nbytes > 0 ) )
         || ( ( results == nullptr ) && ( results_capacity > 0 ) ) ) {
        if ( ctx != nullptr ) {
            ctx->lastError = "ibz2_decode_buffer: invalid argument";
        }
        return IBZ2_E_INVALID_ARGUMENT;
    }
    *nblocks = 0;
    if ( total_out != nullptr ) {
        *total_out = 0;
    }
    beginCall( ctx );
    const cudaEvent_t eStart = nextEvent( ctx );
    if ( ( data_mem == IBZ2_MEM_HOST ) && ( out != nullptr ) && ( out_mem == IBZ2_MEM_HOST ) && ( nbytes >= splitInputBytes() ) ) {
        /* host to host: the second half of the input travels while the blocks of the first half decode */
        bool done = false;
        const int rcSplit = decodeBufferSplit( ctx, static_cast<const uint8_t*>( data ), nbytes, first_bit, last_bit, out,
                                               out_capacity, results, results_capacity, nblocks, total_out, &done );
        if ( ( rcSplit != IBZ2_OK ) || done ) {
            if ( rcSplit == IBZ2_OK ) {
                const cudaEvent_t eEnd = nextEvent( ctx );
                IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->stream ) );
                ctx->timings.total_ms = elapsed( eStart, eEnd );
            }
            return rcSplit;
        }
        /* a block of the first half did not decode cleanly: what the reference reports for it may depend on bytes behind
         * the half, so the whole call is done again the plain way (the input is on the device by now) */
        *nblocks = 0;
        ctx->timings = ibz2_timings{};
    }
    const uint8_t* dData = nullptr;
    int rc = stageInput( ctx, data, nbytes, data_mem, &dData, "ibz2_decode_buffer" );
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    const cudaEvent_t eH2D = nextEvent( ctx );
    std::vector<uint64_t> found;
    rc = findOnDevice( ctx, dData, nbytes, IBZ2_FIND_BLOCK, found );
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    std::vector<uint64_t> mine;
    mine.reserve( found.size() );
    for ( const uint64_t f : found ) {
        if ( ( f >= first_bit ) && ( f < last_bit ) ) {
            mine.push_back( f );
        }
    }
    *nblocks = mine.size();
    if ( mine.size() > results_capacity ) {
        ctx->lastError = "ibz2_decode_buffer: results array too small";
        return IBZ2_E_CAPACITY;
    }
    uint64_t total = 0;
    if ( !mine.empty() ) {
        uint8_t* dOut = nullptr;
        rc = stageOutput( ctx, out, out_capacity, out_mem, &dOut, "ibz2_decode_buffer" );
        if ( rc != IBZ2_OK ) {
            return rc;
        }
        const bool toHost = ( out != nullptr ) && ( out_mem == IBZ2_MEM_HOST );
        rc = decodeOnDevice( ctx, dData, nbytes, mine.data(), mine.size(), dOut, out_capacity, results, &total,
                             toHost ? out : nullptr );
        if ( rc != IBZ2_OK ) {
            return rc;
        }
        ibz2_fold_stream_crcs( results, mine.size(), ctx->verifyStreamCrc ? 1 : 0 );
    }
    if ( total_out != nullptr ) {
        *total_out = total;
    }
    const cudaEvent_t eEnd = nextEvent( ctx );
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->stream ) );
    ctx->timings.h2d_ms = elapsed( eStart, eH2D );
    ctx->timings.total_ms = elapsed( eStart, eEnd );
    return IBZ2_OK;
}

int
ibz2_timer_start( ibz2_ctx* ctx )
{
    if ( ctx == nullptr ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    cudaSetDevice( ctx->device );
    g_smCount = ctx->sms;
    if ( ctx->timerStart == nullptr ) {
        IBZ2_CUDA_CHECK( ctx, cudaEventCreate( &ctx->timerStart ) );
        IBZ2_CUDA_CHECK( ctx, cudaEventCreate( &ctx->timerStop ) );
    }
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( ctx->stream ) );
    IBZ2_CUDA_CHECK( ctx, cudaEventRecord( ctx->timerStart, ctx->stream ) );
    return IBZ2_OK;
}

int
ibz2_timer_stop( ibz2_ctx* ctx, float* milliseconds )
{
    if ( ( ctx == nullptr ) || ( milliseconds == nullptr ) || ( ctx->timerStart == nullptr ) ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    cudaSetDevice( ctx->device );
    g_smCount = ctx->sms;
    IBZ2_CUDA_CHECK( ctx, cudaEventRecord( ctx->timerStop, ctx->stream ) );
    IBZ2_CUDA_CHECK( ctx, cudaEventSynchronize( ctx->timerStop ) );
    IBZ2_CUDA_CHECK( ctx, cudaEventElapsedTime( milliseconds, ctx->timerStart, ctx->timerStop ) );
    return IBZ2_OK;
}

int
ibz2_debug_decode_block( ibz2_ctx* ctx, const void* data, uint64_t nbytes, uint64_t bit_offset,
                         ibz2_block_result* result, uint8_t* lbytes, uint32_t* tvec, uint8_t* prerle,
                         uint8_t* out, uint64_t out_capacity )
{
    if ( ( ctx == nullptr ) || ( data == nullptr ) || ( result == nullptr ) ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    cudaSetDevice( ctx->device );
    g_smCount = ctx->sms;
    cudaStream_t s = ctx->stream;
    int rc = ensureInput( ctx, nbytes );
    if ( rc == IBZ2_OK ) {
        rc = ensureScratch( ctx, 1 );
    }
    if ( rc == IBZ2_OK ) {
        rc = ensureOutput( ctx, out_capacity );
    }
    if ( rc != IBZ2_OK ) {
        return rc;
    }
    IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( ctx->dIn, data, nbytes, cudaMemcpyHostToDevice, s ) );
    IBZ2_CUDA_CHECK( ctx, cudaMemsetAsync( ctx->dRunning, 0, 2 * sizeof( uint64_t ), s ) );
    std::memset( ctx->hMeta, 0, sizeof( ibz2::BlockMeta ) );
    ctx->hMeta[0].bitOffset = bit_offset;
    IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( ctx->dMeta, ctx->hMeta, sizeof( ibz2::BlockMeta ), cudaMemcpyHostToDevice, s ) );

    ibz2::launchHuffmanDecode( ctx->dIn, nbytes, ctx->dMeta, 1, ctx->dLbytes, ctx->dSelectors, ctx->dTables, ctx->dDbuf,
                               ctx->dCounter, false, s );
    IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( ctx->hMeta, ctx->dMeta, sizeof( ibz2::BlockMeta ), cudaMemcpyDeviceToHost, s ) );
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( s ) );
    IBZ2_CUDA_CHECK( ctx, cudaGetLastError() );
    /* (for a block that failed, n carries what the reference's message quotes, not a symbol count) */
    const uint32_t n = ctx->hMeta[0].status == IBZ2_ST_OK ? ctx->hMeta[0].n : 0U;
    if ( ( lbytes != nullptr ) && ( n > 0 ) ) {
        IBZ2_CUDA_CHECK( ctx, cudaMemcpy( lbytes, ctx->dLbytes, n, cudaMemcpyDeviceToHost ) );
    }
    ibz2::launchBwtPrepare( ctx->dLbytes, ctx->dDbuf, ctx->dMeta, 1, s );
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( s ) );
    IBZ2_CUDA_CHECK( ctx, cudaGetLastError() );
    if ( ( tvec != nullptr ) && ( n > 0 ) ) {
        IBZ2_CUDA_CHECK( ctx, cudaMemcpy( tvec, ctx->dDbuf, static_cast<uint64_t>( n ) * sizeof( uint32_t ),
                                          cudaMemcpyDeviceToHost ) );
    }
    ibz2::launchBwtTraverse( ctx->dDbuf, ctx->dLbytes, ctx->dMeta, 1, ctx->dStage, s );
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( s ) );
    IBZ2_CUDA_CHECK( ctx, cudaGetLastError() );
    if ( ( prerle != nullptr ) && ( n > 0 ) ) {
        IBZ2_CUDA_CHECK( ctx, cudaMemcpy( prerle, ctx->dLbytes, n, cudaMemcpyDeviceToHost ) );
    }
    ibz2::launchRleScan( ctx->dLbytes, ctx->dDbuf, ctx->dWarpBase, ctx->dMeta, 1, s );
    ibz2::launchOutputLayout( ctx->dMeta, 1, ctx->dRunning, out_capacity, out != nullptr ? 1 : 0, ctx->dPieceStart,
                              ibz2::CRC_PIECE_BYTES, s );
    if ( out != nullptr ) {
        ibz2::launchRleWrite( ctx->dLbytes, ctx->dDbuf, ctx->dWarpBase, ctx->dMeta, 1, ctx->dOut, s );
        ibz2::launchCrc( ctx->dOut, ctx->dMeta, 1, ctx->dPieceStart, ibz2::CRC_PIECE_BYTES, s );
        ibz2::launchCrcFinalize( ctx->dOut, ctx->dMeta, 1, s );
    }
    IBZ2_CUDA_CHECK( ctx, cudaMemcpyAsync( ctx->hMeta, ctx->dMeta, sizeof( ibz2::BlockMeta ), cudaMemcpyDeviceToHost, s ) );
    IBZ2_CUDA_CHECK( ctx, cudaStreamSynchronize( s ) );
    IBZ2_CUDA_CHECK( ctx, cudaGetLastError() );
    fillResult( ctx->hMeta[0], result );
    if ( ( out != nullptr ) && ( ctx->hMeta[0].writeOutput != 0 ) && ( ctx->hMeta[0].outSize > 0 ) ) {
        IBZ2_CUDA_CHECK( ctx, cudaMemcpy( out, ctx->dOut, ctx->hMeta[0].outSize, cudaMemcpyDeviceToHost ) );
    }
    return IBZ2_OK;
}

int
ibz2_read_stream_header( const uint8_t* data, uint64_t nbytes, uint64_t* bit_pos )
{
    if ( ( data == nullptr ) || ( bit_pos == nullptr ) ) {
        return 0;
    }
    HostBits r{ data, nbytes * 8U, *bit_pos, false };
    static const char magic[3] = { 'B', 'Z', 'h' };
    int level = 0;
    for ( int i = 0; i < 4; ++i ) {
        const uint64_t b = r.read( 8 );
        if ( r.eof ) {
            *bit_pos = r.pos;
            return -1;
        }
        if ( i < 3 ) {
            if ( b != static_cast<uint8_t>( magic[i] ) ) {
                *bit_pos = r.pos;
                return 0;
            }
        } else {
            level = ( ( b >= '1' ) && ( b <= '9' ) ) ? static_cast<int>( b - '0' ) : 0;
        }
    }
    *bit_pos = r.pos;
    return level;
}

int
ibz2_read_block_header( const uint8_t* data, uint64_t nbytes, uint64_t bit_offset, ibz2_block_result* result )
{
    if ( ( data == nullptr ) || ( result == nullptr ) ) {
        return IBZ2_E_INVALID_ARGUMENT;
    }
    std::memset( result, 0, sizeof( *result ) );
    result->enc_offset_bits = bit_offset;
    HostBits r{ data, nbytes * 8U, bit_offset, false };
    if ( bit_offset > r.nbits ) {
        result->status = IBZ2_ST_EOF;
        return IBZ2_OK;
    }
    const uint64_t magic = r.read( 48 );
    result->expected_crc = static_cast<uint32_t>( r.read( 32 ) );
    if ( r.eof ) {
        result->status = IBZ2_ST_EOF;
        return IBZ2_OK;
    }
    if ( magic == IBZ2_MAGIC_EOS ) {
        result->is_eos = 1;
        const unsigned inByte = static_cast<unsigned>( r.pos & 7U );
        if ( inByte > 0 ) {
            r.read( 8U - inByte );
            if ( r.eof ) {
                result->status = IBZ2_ST_EOF;
                return IBZ2_OK;
            }
        }
        result->enc_size_bits = r.pos - bit_offset;
        result->is_eof = r.pos >= r.nbits ? 1U : 0U;
        return IBZ2_OK;
    }
    if ( magic != IBZ2_MAGIC_BLOCK ) {
        result->status = IBZ2_ST_BAD_MAGIC;
        result->n_symbols = static_cast<uint32_t>( magic >> 24 );   /* what was there instead, for the message */
        result->orig_ptr = static_cast<uint32_t>( magic & 0xFFFFFFU );
    }
    return IBZ2_OK;
}

uint32_t
ibz2_stream_crc_fold( uint32_t stream_crc, uint32_t block_crc )
{
    return ( ( stream_crc << 1 ) | ( stream_crc >> 31 ) ) ^ block_crc;
}

}  // extern "C"
