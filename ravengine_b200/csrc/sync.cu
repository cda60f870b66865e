// sync.cu — dirty-range upload of host-shared per-entity arrays: the step BEFORE the render-prep
// path. Mirrors BufferedVRAMVector / BufferedVRAMStructureBase::EncodeSync
// (include/RavEngine/BufferedVRAMVector.hpp:33-209, src/BufferedVRAMVector.cpp:27-144).
//
// The reference walks one byte flag per element, merges dirty elements whose gaps are <= maxGapSize (5) into
// ranges and records one CopyBufferToBuffer per range. That is reproduced exactly by RVS_SYNC_RANGES (the range
// list is the same, each range is one async copy). With many small ranges (scattered moving objects) the copy
// count explodes, so RVS_SYNC_PACKED - the B200-first path - sends long ranges as direct copies and packs all
// short ranges' modified rows (not the clean gap rows a range drags along) into ONE pinned staging block
// [rows][u32 row indices], uploads it with one copy and scatters it with one kernel (16-B lanes, fully coalesced
// on the staging side). The flag scan skips clean spans eight
// flags at a time. Both modes leave the private buffer byte-identical to the host buffer.
#include <cstring>

#include <fcntl.h>
#include <sys/mman.h>
#include <unistd.h>

#include "frame_internal.h"
#include <algorithm>

#include "internal.h"

namespace rve {

// one thread per 16-B chunk of a row (rows are multiples of 4 B; the 16-B path needs 16-B aligned rows)
__global__ void scatter_rows16_kernel(uint4* __restrict__ dst, const uint32_t* __restrict__ idx, const uint4* __restrict__ rows,
                                      uint32_t count, uint32_t chunksPerRow) {
    const uint64_t t = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint64_t total = uint64_t(count) * chunksPerRow;
    for (uint64_t i = t; i < total; i += uint64_t(gridDim.x) * blockDim.x) {
        const uint32_t r = uint32_t(i / chunksPerRow), c = uint32_t(i % chunksPerRow);
        dst[uint64_t(__ldg(idx + r)) * chunksPerRow + c] = __ldg(rows + i);
    }
}
__global__ void scatter_rows4_kernel(uint32_t* __restrict__ dst, const uint32_t* __restrict__ idx, const uint32_t* __restrict__ rows,
                                     uint32_t count, uint32_t wordsPerRow) {
    const uint64_t t = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint64_t total = uint64_t(count) * wordsPerRow;
    for (uint64_t i = t; i < total; i += uint64_t(gridDim.x) * blockDim.x) {
        const uint32_t r = uint32_t(i / wordsPerRow), c = uint32_t(i % wordsPerRow);
        dst[uint64_t(__ldg(idx + r)) * wordsPerRow + c] = __ldg(rows + i);
    }
}
// 2-byte rows (perObjectAttributes): one thread per row
__global__ void scatter_rows2_kernel(uint16_t* __restrict__ dst, const uint32_t* __restrict__ idx, const uint16_t* __restrict__ rows,
                                     uint32_t count) {
    const uint64_t t = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    for (uint64_t i = t; i < count; i += uint64_t(gridDim.x) * blockDim.x) dst[__ldg(idx + i)] = rows[i];
}

// RVS_SYNC_PULL: the GPU walks the modification flags itself (pinned host memory, read over PCIe), copies every
// flagged row straight from the pinned host array into the private buffer and clears the flag - no host scan, no
// host packing, one launch. A warp owns 32 consecutive rows; flagged rows are copied by groups of lanes.
template <typename Chunk>
__global__ void pull_rows_kernel(uint8_t* __restrict__ flags, const Chunk* __restrict__ src, Chunk* __restrict__ dst, uint32_t n,
                                 uint32_t chunksPerRow, uint32_t lanesPerRow, uint32_t* __restrict__ rowsCopied) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warpsTotal = gridDim.x * (blockDim.x >> 5);
    const uint32_t groups = (n + 31) / 32;
    const uint32_t rowsPerIter = 32 / lanesPerRow;
    uint32_t copied = 0;
    for (uint32_t grp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); grp < groups; grp += warpsTotal) {
        const uint32_t row0 = grp * 32;
        uint8_t f = 0;
        if (row0 + lane < n) f = flags[row0 + lane];
        const uint32_t mask = __ballot_sync(0xffffffffu, f != 0);
        if (mask == 0) continue;
        if (f) flags[row0 + lane] = 0;
        const uint32_t cnt = __popc(mask);
        copied += cnt;
        for (uint32_t i = 0; i < cnt; i += rowsPerIter) {
            const uint32_t j = i + lane / lanesPerRow;
            if (j < cnt) {
                const uint64_t row = row0 + (__fns(mask, 0, int(j) + 1));
                for (uint32_t c = lane % lanesPerRow; c < chunksPerRow; c += lanesPerRow) dst[row * chunksPerRow + c] = src[row * chunksPerRow + c];
            }
        }
    }
    if (lane == 0 && copied) atomicAdd(rowsCopied, copied);
}

static uint32_t pow2_at_most_32(uint32_t v) {
    uint32_t p = 1;
    while (p * 2 <= v && p < 32) p *= 2;
    return p;
}

int launch_pull(rve_ctx* ctx, uint8_t* flags, const unsigned char* host, unsigned char* dev, uint32_t n, uint32_t elemBytes,
                uint32_t* rowsCopied) {
    const int threads = 256;
    const uint32_t groups = (n + 31) / 32;
    uint32_t blocks = (groups + 7) / 8;
    const uint32_t maxBlocks = uint32_t(ctx->sm_count) * 8u;
    if (blocks > maxBlocks) blocks = maxBlocks;
    const bool a16 = elemBytes % 16 == 0 && ((reinterpret_cast<uintptr_t>(host) | reinterpret_cast<uintptr_t>(dev)) & 15) == 0;
    if (a16) {
        const uint32_t cpr = elemBytes / 16;
        pull_rows_kernel<uint4><<<blocks, threads, 0, ctx->stream>>>(flags, reinterpret_cast<const uint4*>(host), reinterpret_cast<uint4*>(dev), n,
                                                                    cpr, pow2_at_most_32(cpr), rowsCopied);
    } else if (elemBytes % 4 == 0) {
        const uint32_t cpr = elemBytes / 4;
        pull_rows_kernel<uint32_t><<<blocks, threads, 0, ctx->stream>>>(flags, reinterpret_cast<const uint32_t*>(host),
                                                                       reinterpret_cast<uint32_t*>(dev), n, cpr, pow2_at_most_32(cpr), rowsCopied);
    } else {
        const uint32_t cpr = elemBytes / 2;
        pull_rows_kernel<uint16_t><<<blocks, threads, 0, ctx->stream>>>(flags, reinterpret_cast<const uint16_t*>(host),
                                                                       reinterpret_cast<uint16_t*>(dev), n, cpr, pow2_at_most_32(cpr), rowsCopied);
    }
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}


// ---- affine rows (RVS_VECTOR_AFFINE_ROWS): host rows are glm::mat4x3 (48 B: columns c0..c3, xyz each), the private
// buffer keeps the reference's mat4 rows (64 B, column-major, w = 0,0,0,1) - the layout defaultcull.csh and the vertex
// shaders read. A world matrix is affine, so the fourth row carries no information: 25 % fewer bytes over PCIe. ----

// column `col` (0..3) of a 64-B row from the 12 floats of a 48-B row
__device__ __forceinline__ float4 affine_column(const float* r12, uint32_t col) {
    return make_float4(r12[col * 3], r12[col * 3 + 1], r12[col * 3 + 2], col == 3 ? 1.0f : 0.0f);
}

// Whole-array upload, one warp per 32 rows: three fully coalesced 512-B loads of the 48-B rows (device staging memory,
// or pinned HOST memory read over PCIe - no staging copy at all), re-formed through shared memory into four fully
// coalesced 512-B stores of 64-B rows. EXPAND = false: plain coalesced copy of 16-B chunks (any row size).
template <bool EXPAND>
__global__ void __launch_bounds__(256) full_rows_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, uint64_t rows, uint64_t chunks) {
    if constexpr (EXPAND) {
        __shared__ uint4 s_buf[8][96];
        const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        const uint64_t warpsTotal = uint64_t(gridDim.x) * 8u;
        const uint64_t groups = (rows + 31) / 32;
        const float* sf = reinterpret_cast<const float*>(s_buf[warp]);
        for (uint64_t g = uint64_t(blockIdx.x) * 8u + warp; g < groups; g += warpsTotal) {
            const uint64_t c0 = g * 96;  // first 16-B chunk of the group
            const uint64_t cEnd = rows * 3;
            uint4 v[3];
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const uint64_t c = c0 + uint64_t(j * 32) + lane;
                v[j] = c < cEnd ? __ldg(src + c) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int j = 0; j < 3; ++j) s_buf[warp][j * 32 + lane] = v[j];
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t idx = uint32_t(j * 32) + lane;
                const uint64_t row = g * 32 + (idx >> 2);
                if (row < rows) {
                    const float4 o = affine_column(sf + (idx >> 2) * 12, idx & 3);
                    reinterpret_cast<float4*>(dst)[row * 4 + (idx & 3)] = o;
                }
            }
            __syncwarp();
        }
    } else {
        const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
        for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < chunks; i += stride) dst[i] = __ldg(src + i);
    }
}

// packed staging block [48-B rows][u32 indices] -> 64-B rows at dst[idx[r]]: one thread per (row, column)
__global__ void scatter_rows_affine_kernel(float4* __restrict__ dst, const uint32_t* __restrict__ idx, const float* __restrict__ rows,
                                           uint32_t count) {
    const uint64_t total = uint64_t(count) * 4;
    for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += uint64_t(gridDim.x) * blockDim.x) {
        const uint32_t r = uint32_t(i >> 2), c = uint32_t(i & 3);
        dst[uint64_t(__ldg(idx + r)) * 4 + c] = affine_column(rows + uint64_t(r) * 12, c);
    }
}

// contiguous run of 48-B rows in device staging memory -> 64-B rows dst[first ..) (the direct-copy ranges of an affine vector)
__global__ void expand_range_affine_kernel(float4* __restrict__ dst, const float* __restrict__ rows, uint64_t first, uint32_t count) {
    const uint64_t total = uint64_t(count) * 4;
    for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += uint64_t(gridDim.x) * blockDim.x)
        dst[first * 4 + i] = affine_column(rows + (i >> 2) * 12, uint32_t(i & 3));
}

// RVS_SYNC_PULL for affine rows: like pull_rows_kernel, four lanes per flagged row; lanes 0..2 fetch the row's three
// 16-B chunks from the pinned host array, the group shares them by shuffle and every lane writes one 64-B-row column.
__global__ void pull_rows_affine_kernel(uint8_t* __restrict__ flags, const uint4* __restrict__ src, float4* __restrict__ dst, uint32_t n,
                                        uint32_t* __restrict__ rowsCopied) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warpsTotal = gridDim.x * (blockDim.x >> 5);
    const uint32_t groups = (n + 31) / 32;
    uint32_t copied = 0;
    for (uint32_t grp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); grp < groups; grp += warpsTotal) {
        const uint32_t row0 = grp * 32;
        uint8_t f = 0;
        if (row0 + lane < n) f = flags[row0 + lane];
        const uint32_t mask = __ballot_sync(0xffffffffu, f != 0);
        if (mask == 0) continue;
        if (f) flags[row0 + lane] = 0;
        const uint32_t cnt = __popc(mask);
        copied += cnt;
        for (uint32_t i = 0; i < cnt; i += 8) {  // eight rows per iteration, four lanes each
            const uint32_t j = i + (lane >> 2), c = lane & 3;
            const bool on = j < cnt;
            const uint64_t row = on ? row0 + __fns(mask, 0, int(j) + 1) : 0;
            uint4 ch = make_uint4(0, 0, 0, 0);
            if (on && c < 3) ch = src[row * 3 + c];
            float r12[12];
            const uint32_t base = lane & ~3u;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                r12[4 * k + 0] = __uint_as_float(__shfl_sync(0xffffffffu, ch.x, int(base) + k));
                r12[4 * k + 1] = __uint_as_float(__shfl_sync(0xffffffffu, ch.y, int(base) + k));
                r12[4 * k + 2] = __uint_as_float(__shfl_sync(0xffffffffu, ch.z, int(base) + k));
                r12[4 * k + 3] = __uint_as_float(__shfl_sync(0xffffffffu, ch.w, int(base) + k));
            }
            if (on) {
                // column c without dynamic register indexing
                float4 o;
                if (c == 0) o = make_float4(r12[0], r12[1], r12[2], 0.0f);
                else if (c == 1) o = make_float4(r12[3], r12[4], r12[5], 0.0f);
                else if (c == 2) o = make_float4(r12[6], r12[7], r12[8], 0.0f);
                else o = make_float4(r12[9], r12[10], r12[11], 1.0f);
                dst[row * 4 + c] = o;
            }
        }
    }
    if (lane == 0 && copied) atomicAdd(rowsCopied, copied);
}

int launch_full(rve_ctx* ctx, cudaStream_t stream, const unsigned char* src, unsigned char* dst, uint64_t rows, uint32_t srcRowBytes, bool expand) {
    if (!rows) return 0;
    const uint32_t maxBlocks = uint32_t(ctx->sm_count) * 8u;
    if (expand) {
        const uint64_t groups = (rows + 31) / 32;
        const uint32_t blocks = uint32_t(std::min<uint64_t>((groups + 7) / 8, maxBlocks));
        full_rows_kernel<true><<<blocks, 256, 0, stream>>>(reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(dst), rows, 0);
    } else {
        const uint64_t chunks = rows * srcRowBytes / 16;
        const uint32_t blocks = uint32_t(std::min<uint64_t>((chunks + 255) / 256, maxBlocks));
        full_rows_kernel<false><<<blocks, 256, 0, stream>>>(reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(dst), rows, chunks);
    }
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace rve

namespace rve {
// rvs_vectors_encode_sync_batch: the modified rows of up to kSyncBatchSegs vectors, staged as one block of 4-byte words
// [rows of segment 0 | rows of segment 1 | ..] + one u32 row index per staged row, scattered by ONE launch
constexpr uint32_t kSyncBatchSegs = 16;
struct SyncBatchSeg {
    uint32_t* dst;       // private buffer of the vector
    uint32_t rowWords;   // 4-byte words per row
    uint32_t rowStart;   // first staged row of the segment (index array)
    uint32_t wordStart;  // first staged word of the segment
    uint32_t words;      // rows * rowWords
    uint32_t pad;
};
struct SyncBatch {
    SyncBatchSeg s[kSyncBatchSegs];
    uint32_t n, totalWords;
    const uint32_t* data;
    const uint32_t* idx;
};
__global__ void scatter_rows_batch_kernel(const SyncBatch b) {
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < b.totalWords; t += gridDim.x * blockDim.x) {
        uint32_t k = 0;
        while (k + 1 < b.n && t >= b.s[k + 1].wordStart) ++k;
        const SyncBatchSeg& g = b.s[k];
        const uint32_t local = t - g.wordStart;
        const uint32_t row = local / g.rowWords, w = local - row * g.rowWords;
        g.dst[size_t(__ldg(b.idx + g.rowStart + row)) * g.rowWords + w] = __ldg(b.data + t);
    }
}
}  // namespace rve

using namespace rve;

extern "C" int rve_scatter_rows(rve_ctx* ctx, rve_buffer dst, uint32_t elem_bytes, rve_buffer staging, uint64_t indices_offset,
                                uint64_t rows_offset, uint32_t count) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "null context");
    if (count == 0) return 0;
    DeviceGuard g(ctx->device);
    BufferRec* bd = get_buffer(ctx, dst, "scatter destination");
    BufferRec* bs = get_buffer(ctx, staging, "scatter staging");
    if (!bd || !bs) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE(elem_bytes == 2 || (elem_bytes % 4) == 0, "row size must be 2 or a multiple of 4 bytes");
    RVE_REQUIRE(indices_offset % 4 == 0 && indices_offset + uint64_t(count) * 4 <= bs->bytes, "index list outside the staging buffer");
    RVE_REQUIRE(rows_offset + uint64_t(count) * elem_bytes <= bs->bytes, "row payload outside the staging buffer");
    const auto* base = static_cast<const unsigned char*>(bs->ptr);
    const auto* idx = reinterpret_cast<const uint32_t*>(base + indices_offset);
    const unsigned char* rows = base + rows_offset;
    const int threads = 256;
    const uint32_t maxBlocks = uint32_t(ctx->sm_count) * 8u;
    if (elem_bytes == 2) {
        RVE_REQUIRE(rows_offset % 2 == 0, "row payload must be 2-byte aligned");
        const uint32_t blocks = std::min<uint64_t>((uint64_t(count) + threads - 1) / threads, maxBlocks);
        scatter_rows2_kernel<<<blocks, threads, 0, ctx->stream>>>(static_cast<uint16_t*>(bd->ptr), idx,
                                                                 reinterpret_cast<const uint16_t*>(rows), count);
    } else if (elem_bytes % 16 == 0 && rows_offset % 16 == 0 && (reinterpret_cast<uintptr_t>(bd->ptr) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(bs->ptr) & 15) == 0) {
        const uint32_t cpr = elem_bytes / 16;
        const uint32_t blocks = std::min<uint64_t>((uint64_t(count) * cpr + threads - 1) / threads, maxBlocks);
        scatter_rows16_kernel<<<blocks, threads, 0, ctx->stream>>>(static_cast<uint4*>(bd->ptr), idx, reinterpret_cast<const uint4*>(rows),
                                                                  count, cpr);
    } else {
        RVE_REQUIRE(rows_offset % 4 == 0, "row payload must be 4-byte aligned");
        const uint32_t wpr = elem_bytes / 4;
        const uint32_t blocks = std::min<uint64_t>((uint64_t(count) * wpr + threads - 1) / threads, maxBlocks);
        scatter_rows4_kernel<<<blocks, threads, 0, ctx->stream>>>(static_cast<uint32_t*>(bd->ptr), idx,
                                                                 reinterpret_cast<const uint32_t*>(rows), count, wpr);
    }
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// BufferedVRAMVector stand-in (host-shared buffer + modification flags + device-private buffer)
// ---------------------------------------------------------------------------------------------------------
struct rvs_vector {
    rve_ctx* ctx = nullptr;
    uint32_t elemBytes = 0;     // host row
    uint32_t devElemBytes = 0;  // private-buffer row: elemBytes, or 64 for affine rows (48-B host rows)
    bool affine = false;
    bool sharedHost = false;  // RVS_VECTOR_SHARED_HOST: the host block is a memfd mapping another rank can map and pin too
    int hostFd = -1;
    size_t hostMapBytes = 0;
    bool anyDirty = false;  // some flag may be set (SetValueAt / mark / a caller holding the tracking pointer): a clean vector syncs for free
    uint32_t count = 0, capacity = 0;
    unsigned char* host = nullptr;  // pinned: VRAMVector's host-mapped Shared buffer
    // RVS_SYNC_FULL of an affine vector: copy-engine chunks into device staging, expanded on a second stream while
    // the next chunk is in flight
    rve_buffer fullStage = RVE_BUFFER_INVALID;
    size_t fullStageBytes = 0;
    cudaStream_t copyStream = nullptr, expandStream = nullptr;
    cudaEvent_t evBegin = nullptr, evDone = nullptr;
    std::vector<cudaEvent_t> evChunk;
    // syncTrackingBuffer: pinned, capacity rounded up to 32 flags, double-buffered so that RVS_SYNC_PULL can hand
    // the flags of this frame to the GPU while the host already records the next frame's modifications
    uint8_t* trackBuf[2] = {nullptr, nullptr};
    uint8_t* track = nullptr;  // == trackBuf[cur]
    int cur = 0;
    cudaEvent_t pullDone[2] = {nullptr, nullptr};  // the pull kernel that read (and cleared) trackBuf[i] has finished
    bool pullBusy[2] = {false, false};
    uint32_t* pullRows = nullptr;  // device counter: rows copied by pull kernels since creation
    rve_buffer priv = RVE_BUFFER_INVALID;
    size_t privBytes = 0;
    unsigned char* stageHost = nullptr;  // packed mode: pinned staging block [rows][indices]
    size_t stageHostBytes = 0;
    rve_buffer stageDev = RVE_BUFFER_INVALID;
    size_t stageDevBytes = 0;
    cudaEvent_t stageFree = nullptr;  // upload of the previous staging block has completed
    bool stageBusy = false;
    std::vector<uint32_t> ranges;  // (begin, end) pairs of the last sync
    // Indices flagged since the last sync, while every flag was set through SetValueAt / mark / Resize (touchedExact): the next
    // sync derives the reference's ranges from them instead of scanning all flags - a 2.5 M-entity list with a few hundred
    // changes costs microseconds of host time instead of ~0.4 ms. A caller holding the raw tracking pointer, or too many
    // touches, turn it off until the next sync (which then scans).
    std::vector<uint32_t> touched;
    bool touchedExact = true;
    bool rangesReady = false;  // `ranges` already holds this sync's ranges (computed by rvs_vectors_encode_sync_batch before it handed the vector on)
};
constexpr size_t kTouchedMax = size_t(1) << 16;
static inline void note_touched(rvs_vector* v, uint32_t begin, uint32_t end) {
    if (!v->touchedExact) return;
    if (v->touched.size() + (end - begin) > kTouchedMax) {
        v->touchedExact = false;
        v->touched.clear();
        return;
    }
    for (uint32_t i = begin; i < end; ++i) v->touched.push_back(i);
}

namespace {

constexpr int kMaxGapSize = 5;  // src/BufferedVRAMVector.cpp:106

uint32_t next_pow2(uint32_t v) {
    uint32_t p = 1;
    while (p < v) p <<= 1;
    return p;
}

int vec_reserve(rvs_vector* v, uint32_t newCap) {
    if (newCap <= v->capacity) return 0;
    DeviceGuard g(v->ctx->device);
    unsigned char* nh = nullptr;
    if (v->sharedHost) {
        // one allocation for the vector's life: peers hold a mapping of this very block
        if (v->host) return set_error(RVE_ERR_UNSUPPORTED, "a shared-host vector cannot grow past its first size");
        const size_t bytes = (size_t(newCap) * v->elemBytes + 4095) & ~size_t(4095);
        const int fd = memfd_create("rvs_vector", 0);
        if (fd < 0 || ftruncate(fd, off_t(bytes)) != 0) {
            if (fd >= 0) close(fd);
            return set_error(RVE_ERR_OUT_OF_MEMORY, "memfd_create / ftruncate failed for a shared-host vector");
        }
        void* m = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
        if (m == MAP_FAILED) {
            close(fd);
            return set_error(RVE_ERR_OUT_OF_MEMORY, "mmap failed for a shared-host vector");
        }
        const int rc = check_cuda(cudaHostRegister(m, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped), "cudaHostRegister");
        if (rc) {
            munmap(m, bytes);
            close(fd);
            return rc;
        }
        nh = static_cast<unsigned char*>(m);
        v->hostFd = fd;
        v->hostMapBytes = bytes;
    } else {
        RVE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&nh), size_t(newCap) * v->elemBytes, cudaHostAllocDefault));
    }
    if (v->host) {
        // copies enqueued from the old block must have been consumed before it is released
        RVE_CUDA(cudaStreamSynchronize(v->ctx->stream));
        std::memcpy(nh, v->host, size_t(v->count) * v->elemBytes);
        cudaFreeHost(v->host);
    }
    v->host = nh;
    for (int i = 0; i < 2; ++i) {
        uint8_t* nt = nullptr;
        RVE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&nt), size_t(newCap) + 32, cudaHostAllocDefault));
        std::memset(nt, 0, size_t(newCap) + 32);
        if (v->trackBuf[i]) {
            if (v->pullBusy[i]) cudaEventSynchronize(v->pullDone[i]);
            v->pullBusy[i] = false;
            std::memcpy(nt, v->trackBuf[i], v->count);
            cudaFreeHost(v->trackBuf[i]);
        }
        v->trackBuf[i] = nt;
    }
    v->track = v->trackBuf[v->cur];
    v->capacity = newCap;
    return 0;
}

// The reference's range computation (src/BufferedVRAMVector.cpp:108-139); clean spans are skipped eight flags at a
// time while no range is open (same result: they neither open nor close a range). The flags are cleared by the
// caller, range by range (the reference clears them inside this loop; every set flag lies inside some range).
// The same ranges from the list of flagged indices (ascending, unique, all < n, all with their flag still set): a range runs
// from a flagged index to kMaxGapSize past the last flagged index that lies within kMaxGapSize of its predecessor, clipped at n -
// exactly what the loop below produces (tests/test_vram_vector.py compares both with the restated reference loop).
void ranges_from_indices(const std::vector<uint32_t>& idx, uint32_t n, std::vector<uint32_t>& out) {
    out.clear();
    size_t i = 0;
    while (i < idx.size()) {
        const uint32_t begin = idx[i];
        uint32_t last = begin;
        ++i;
        while (i < idx.size() && idx[i] <= last + uint32_t(kMaxGapSize)) last = idx[i++];
        const uint64_t end = uint64_t(last) + uint32_t(kMaxGapSize);
        out.push_back(begin);
        out.push_back(end < n ? uint32_t(end) : n);
    }
}

// ranges of the vector's flags for a sync: from the touched list when it is exact, by the scan otherwise; restarts the list
void vector_ranges(rvs_vector* v);

void compute_ranges(const uint8_t* track, uint32_t n, std::vector<uint32_t>& out) {
    out.clear();
    int gap = 0;
    uint32_t begin = 0, i = 0;
    while (i < n) {
        if (gap == 0 && (i & 7u) == 0) {
            while (i + 8 <= n) {
                uint64_t w;
                std::memcpy(&w, track + i, 8);
                if (w) break;
                i += 8;
            }
            if (i >= n) break;
        }
        if (track[i]) {
            if (gap == 0) begin = i;
            gap = kMaxGapSize;
        } else {
            const int prev = gap;
            gap = gap > 0 ? gap - 1 : 0;
            if (gap == 0 && prev != 0) {
                out.push_back(begin);
                out.push_back(i);
            }
        }
        ++i;
    }
    if (gap != 0) {
        out.push_back(begin);
        out.push_back(n);
    }
}

void vector_ranges(rvs_vector* v) {
    if (v->touchedExact) {
        std::vector<uint32_t>& t = v->touched;
        std::sort(t.begin(), t.end());
        t.erase(std::unique(t.begin(), t.end()), t.end());
        // a Resize below a flagged index cleared its flag; indices past the count are gone with it
        t.erase(std::remove_if(t.begin(), t.end(), [v](uint32_t i) { return i >= v->count || !v->track[i]; }), t.end());
        ranges_from_indices(t, v->count, v->ranges);
    } else {
        compute_ranges(v->track, v->count, v->ranges);
    }
    v->touched.clear();
    v->touchedExact = true;
}

}  // namespace

extern "C" {

int rvs_vector_create(rve_ctx* ctx, uint32_t elem_bytes, uint32_t count, rvs_vector** out) {
    return rvs_vector_create_ex(ctx, elem_bytes, count, 0u, out);
}

int rvs_vector_create_ex(rve_ctx* ctx, uint32_t elem_bytes, uint32_t count, uint32_t flags, rvs_vector** out) {
    if (!ctx || !out) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_REQUIRE(elem_bytes == 2 || (elem_bytes % 4 == 0 && elem_bytes > 0), "element size must be 2 or a multiple of 4 bytes");
    RVE_REQUIRE(!(flags & ~(RVS_VECTOR_AFFINE_ROWS | RVS_VECTOR_SHARED_HOST)), "unknown vector flags");
    RVE_REQUIRE(!(flags & RVS_VECTOR_AFFINE_ROWS) || elem_bytes == 48, "affine rows are 48-byte mat4x3 host rows");
    auto* v = new rvs_vector();
    v->ctx = ctx;
    v->elemBytes = elem_bytes;
    v->affine = (flags & RVS_VECTOR_AFFINE_ROWS) != 0;
    v->devElemBytes = v->affine ? 64u : elem_bytes;
    v->sharedHost = (flags & RVS_VECTOR_SHARED_HOST) != 0;
    {
        DeviceGuard g(ctx->device);
        if (cudaEventCreateWithFlags(&v->stageFree, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&v->pullDone[0], cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&v->pullDone[1], cudaEventDisableTiming) != cudaSuccess ||
            cudaMalloc(reinterpret_cast<void**>(&v->pullRows), sizeof(uint32_t)) != cudaSuccess ||
            cudaMemsetAsync(v->pullRows, 0, sizeof(uint32_t), ctx->stream) != cudaSuccess) {
            delete v;
            return set_error(RVE_ERR_CUDA, "rvs_vector_create: CUDA resource creation failed");
        }
    }
    *out = v;
    return count ? rvs_vector_resize(v, count) : 0;
}

int rvs_vector_destroy(rvs_vector* v) {
    if (!v) return 0;
    DeviceGuard g(v->ctx->device);
    cudaStreamSynchronize(v->ctx->stream);
    if (v->priv != RVE_BUFFER_INVALID) rve_buffer_destroy(v->ctx, v->priv);
    if (v->stageDev != RVE_BUFFER_INVALID) rve_buffer_destroy(v->ctx, v->stageDev);
    if (v->copyStream) cudaStreamSynchronize(v->copyStream);
    if (v->expandStream) cudaStreamSynchronize(v->expandStream);
    if (v->fullStage != RVE_BUFFER_INVALID) rve_buffer_destroy(v->ctx, v->fullStage);
    if (v->copyStream) cudaStreamDestroy(v->copyStream);
    if (v->expandStream) cudaStreamDestroy(v->expandStream);
    if (v->evBegin) cudaEventDestroy(v->evBegin);
    if (v->evDone) cudaEventDestroy(v->evDone);
    for (cudaEvent_t e : v->evChunk) cudaEventDestroy(e);
    if (v->host && v->sharedHost) {
        cudaHostUnregister(v->host);
        munmap(v->host, v->hostMapBytes);
        close(v->hostFd);
    } else if (v->host) {
        cudaFreeHost(v->host);
    }
    if (v->stageHost) cudaFreeHost(v->stageHost);
    if (v->stageFree) cudaEventDestroy(v->stageFree);
    for (int i = 0; i < 2; ++i) {
        if (v->trackBuf[i]) cudaFreeHost(v->trackBuf[i]);
        if (v->pullDone[i]) cudaEventDestroy(v->pullDone[i]);
    }
    if (v->pullRows) cudaFree(v->pullRows);
    delete v;
    return 0;
}

// BufferedVRAMVector::Resize (BufferedVRAMVector.hpp:64-69): new elements start out modified
int rvs_vector_resize(rvs_vector* v, uint32_t count) {
    if (!v) return set_error(RVE_ERR_INVALID_ARGUMENT, "null vector");
    if (count == v->count) return 0;
    if (count > v->capacity) {
        // growth by powers of two; a vector sized once at creation (the bench's 10 M-entity arrays) takes exactly what it needs
        const int rc = vec_reserve(v, v->capacity == 0 ? count : next_pow2(count));
        if (rc) return rc;
    }
    const uint32_t old = v->count;
    if (count > old) {
        std::memset(v->host + size_t(old) * v->elemBytes, 0, size_t(count - old) * v->elemBytes);
        std::memset(v->track + old, 1, count - old);
        v->anyDirty = true;
        note_touched(v, old, count);
    } else {
        std::memset(v->track + count, 0, old - count);
    }
    v->count = count;
    return 0;
}

uint32_t rvs_vector_size(rvs_vector* v) { return v ? v->count : 0; }
void* rvs_vector_host(rvs_vector* v) { return v ? v->host : nullptr; }
uint8_t* rvs_vector_tracking(rvs_vector* v) {
    if (!v) return nullptr;
    v->anyDirty = true;  // the caller may set flags through this pointer: the next sync looks at them ...
    v->touchedExact = false;  // ... all of them: what it flags is not recorded
    v->touched.clear();
    return v->track;
}
rve_buffer rvs_vector_private_buffer(rvs_vector* v) { return v ? v->priv : RVE_BUFFER_INVALID; }

// SetValueAt (BufferedVRAMVector.hpp:113-116)
int rvs_vector_set(rvs_vector* v, uint32_t index, const void* value) {
    if (!v || !value) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_REQUIRE(index < v->count, "index out of range");
    std::memcpy(v->host + size_t(index) * v->elemBytes, value, v->elemBytes);
    if (!v->track[index]) note_touched(v, index, index + 1);
    v->track[index] = 1;
    v->anyDirty = true;
    return 0;
}

// GetValueAtForWriting over [begin, end): the caller writes through rvs_vector_host()
int rvs_vector_mark(rvs_vector* v, uint32_t begin, uint32_t end) {
    if (!v) return set_error(RVE_ERR_INVALID_ARGUMENT, "null vector");
    RVE_REQUIRE(begin <= end && end <= v->count, "range out of bounds");
    std::memset(v->track + begin, 1, end - begin);
    if (end > begin) {
        v->anyDirty = true;
        note_touched(v, begin, end);
    }
    return 0;
}

// rows copied by RVS_SYNC_PULL syncs since creation (reads a device counter: synchronises the stream)
uint64_t rvs_vector_pulled_rows(rvs_vector* v) {
    if (!v || !v->pullRows) return 0;
    DeviceGuard g(v->ctx->device);
    uint32_t r = 0;
    if (cudaMemcpyAsync(&r, v->pullRows, sizeof(r), cudaMemcpyDeviceToHost, v->ctx->stream) != cudaSuccess) return 0;
    cudaStreamSynchronize(v->ctx->stream);
    return r;
}

uint32_t rvs_vector_last_ranges(rvs_vector* v, uint32_t* out_pairs, uint32_t max_pairs) {
    if (!v) return 0;
    const uint32_t n = uint32_t(v->ranges.size() / 2);
    for (uint32_t i = 0; i < n && i < max_pairs; ++i) {
        out_pairs[2 * i] = v->ranges[2 * i];
        out_pairs[2 * i + 1] = v->ranges[2 * i + 1];
    }
    return n;
}

// The two range derivations as plain host functions (no device, no vector): what a sync computes from a flag array by the
// reference's scan, and from the ascending list of flagged indices. Returns the number of (begin, end) pairs; writes at most
// max_pairs of them. For tests that compare the two with the restated reference loop without a GPU.
uint32_t rvs_sync_ranges_from_flags(const uint8_t* flags, uint32_t count, uint32_t* out_pairs, uint32_t max_pairs) {
    std::vector<uint32_t> r;
    if (flags) compute_ranges(flags, count, r);
    for (size_t i = 0; i < r.size() && i < size_t(max_pairs) * 2; ++i) out_pairs[i] = r[i];
    return uint32_t(r.size() / 2);
}
uint32_t rvs_sync_ranges_from_indices(const uint32_t* indices, uint32_t n, uint32_t count, uint32_t* out_pairs, uint32_t max_pairs) {
    std::vector<uint32_t> idx(indices, indices + (indices ? n : 0)), r;
    ranges_from_indices(idx, count, r);
    for (size_t i = 0; i < r.size() && i < size_t(max_pairs) * 2; ++i) out_pairs[i] = r[i];
    return uint32_t(r.size() / 2);
}

// RVS_SYNC_FULL / RVS_SYNC_FULL_PULL: every row goes up, flags are not consulted (and cleared when any may be set)
static int sync_full(rvs_vector* v, uint32_t mode, rvs_sync_stats* st, uint32_t rowBegin = 0, uint32_t rowEnd = 0xFFFFFFFFu) {
    rve_ctx* ctx = v->ctx;
    auto* dev0 = static_cast<unsigned char*>(rve_buffer_device_ptr(ctx, v->priv));
    if (!dev0) return RVE_ERR_BAD_HANDLE;
    if (rowEnd > v->count) rowEnd = v->count;
    if (rowBegin >= rowEnd) return 0;
    const size_t eb = v->elemBytes;
    // the rows [rowBegin, rowEnd) as a vector of their own
    struct Sub {
        unsigned char* host;
        unsigned char* dev;
        uint32_t count;
    } sub{v->host + size_t(rowBegin) * eb, dev0 + size_t(rowBegin) * v->devElemBytes, rowEnd - rowBegin};
    unsigned char* dev = sub.dev;
    const size_t bytes = size_t(sub.count) * eb;
    const bool whole = rowBegin == 0 && rowEnd == v->count;
    const bool chunked16 = (eb % 16 == 0) && ((reinterpret_cast<uintptr_t>(sub.host) | reinterpret_cast<uintptr_t>(dev)) & 15) == 0;
    if (mode == RVS_SYNC_FULL_PULL && (v->affine || chunked16)) {
        // one kernel reads the pinned host rows over PCIe and writes the private rows: no copy-engine operation, no staging
        const int rc = launch_full(ctx, ctx->stream, sub.host, dev, sub.count, v->elemBytes, v->affine);
        if (rc) return rc;
    } else if (!v->affine) {
        RVE_CUDA(cudaMemcpyAsync(dev, sub.host, bytes, cudaMemcpyHostToDevice, ctx->stream));
        st->direct_copies = 1;
    } else {
        // copy engine -> device staging (48-B rows) in chunks on a copy stream; each chunk is expanded into the 64-B
        // private rows on a second stream while the next chunk is still crossing PCIe
        if (!v->copyStream) {
            RVE_CUDA(cudaStreamCreateWithFlags(&v->copyStream, cudaStreamNonBlocking));
            RVE_CUDA(cudaStreamCreateWithFlags(&v->expandStream, cudaStreamNonBlocking));
            RVE_CUDA(cudaEventCreateWithFlags(&v->evBegin, cudaEventDisableTiming));
            RVE_CUDA(cudaEventCreateWithFlags(&v->evDone, cudaEventDisableTiming));
        }
        const size_t want = size_t(v->capacity) * eb;
        if (v->fullStage == RVE_BUFFER_INVALID || v->fullStageBytes != want) {
            if (v->fullStage != RVE_BUFFER_INVALID) {
                RVE_CUDA(cudaStreamSynchronize(v->expandStream));
                rve_buffer_destroy(ctx, v->fullStage);
            }
            const int rc = rve_buffer_create(ctx, want, &v->fullStage);
            if (rc) return rc;
            v->fullStageBytes = want;
        }
        auto* stage = static_cast<unsigned char*>(rve_buffer_device_ptr(ctx, v->fullStage)) + size_t(rowBegin) * eb;
        constexpr uint32_t kChunkRows = 1u << 20;  // 48 MB per copy: ~0.9 ms on the link, ~25 us to expand
        const uint32_t nChunks = (sub.count + kChunkRows - 1) / kChunkRows;
        while (v->evChunk.size() < nChunks) {
            cudaEvent_t e;
            RVE_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            v->evChunk.push_back(e);
        }
        // everything enqueued so far (last frame's readers of the private buffer, this buffer's creation) comes first
        RVE_CUDA(cudaEventRecord(v->evBegin, ctx->stream));
        RVE_CUDA(cudaStreamWaitEvent(v->copyStream, v->evBegin, 0));
        RVE_CUDA(cudaStreamWaitEvent(v->expandStream, v->evBegin, 0));
        for (uint32_t c = 0; c < nChunks; ++c) {
            const uint32_t r0 = c * kChunkRows;
            const uint32_t rows = sub.count - r0 < kChunkRows ? sub.count - r0 : kChunkRows;
            RVE_CUDA(cudaMemcpyAsync(stage + size_t(r0) * eb, sub.host + size_t(r0) * eb, size_t(rows) * eb, cudaMemcpyHostToDevice, v->copyStream));
            RVE_CUDA(cudaEventRecord(v->evChunk[c], v->copyStream));
            RVE_CUDA(cudaStreamWaitEvent(v->expandStream, v->evChunk[c], 0));
            const int rc = launch_full(ctx, v->expandStream, stage + size_t(r0) * eb, dev + size_t(r0) * 64, rows, v->elemBytes, true);
            if (rc) return rc;
        }
        RVE_CUDA(cudaEventRecord(v->evDone, v->expandStream));
        RVE_CUDA(cudaStreamWaitEvent(ctx->stream, v->evDone, 0));
        st->direct_copies = nChunks;
    }
    st->direct_bytes = bytes;
    st->rows_in_ranges = sub.count;
    st->ranges = 1;
    v->ranges.assign({rowBegin, rowEnd});
    if (v->anyDirty) std::memset(v->track + rowBegin, 0, sub.count);
    if (whole) {
        v->anyDirty = false;
        v->touched.clear();  // every flag was cleared
        v->touchedExact = true;
    }
    pyramid_note_buffer_write(ctx, v->priv);
    return 0;
}

// BufferedVRAMStructureBase::EncodeSync (src/BufferedVRAMVector.cpp:27-144)
int rvs_vector_encode_sync(rvs_vector* v, uint32_t mode, uint32_t direct_min_rows, rvs_sync_stats* stats) {
    if (!v) return set_error(RVE_ERR_INVALID_ARGUMENT, "null vector");
    RVE_PROFILE_FN_N("EncodeSync");
    RVE_REQUIRE(mode <= RVS_SYNC_FULL_PULL, "unknown sync mode");
    rve_ctx* ctx = v->ctx;
    DeviceGuard g(ctx->device);
    rvs_sync_stats st{};
    const size_t eb = v->elemBytes, deb = v->devElemBytes;
    const bool rangesReady = v->rangesReady && mode == RVS_SYNC_PACKED;
    v->rangesReady = false;

    // private buffer follows the host buffer's size; old contents are carried over (:31-81)
    const size_t wantBytes = size_t(v->capacity) * deb;
    if (wantBytes && (v->priv == RVE_BUFFER_INVALID || v->privBytes != wantBytes)) {
        rve_buffer nb;
        int rc = rve_buffer_create(ctx, wantBytes, &nb);
        if (rc) return rc;
        if (v->priv != RVE_BUFFER_INVALID) {
            rc = rve_buffer_copy(ctx, v->priv, 0, nb, 0, v->privBytes < wantBytes ? v->privBytes : wantBytes);
            if (rc) return rc;
            rc = rve_buffer_destroy(ctx, v->priv);  // deferred to stream order, like gcBuffers
            if (rc) return rc;
        }
        v->priv = nb;
        v->privBytes = wantBytes;
    }
    if (v->count == 0) {
        v->ranges.clear();
        if (stats) *stats = st;
        return 0;
    }
    if (mode == RVS_SYNC_FULL || mode == RVS_SYNC_FULL_PULL) {
        const int rc = sync_full(v, mode, &st);
        if (stats) *stats = st;
        return rc;
    }
    if (!v->anyDirty) {  // nothing was written since the last sync (the common state of ids / layers / attributes)
        v->ranges.clear();
        if (stats) *stats = st;
        return 0;
    }

    if (mode == RVS_SYNC_PULL) {
        v->ranges.clear();
        auto* devp = static_cast<unsigned char*>(rve_buffer_device_ptr(ctx, v->priv));
        if (!devp) return RVE_ERR_BAD_HANDLE;
        int rc = 0;
        if (v->affine) {
            const uint32_t groups = (v->count + 31) / 32;
            uint32_t blocks = (groups + 7) / 8;
            const uint32_t maxBlocks = uint32_t(ctx->sm_count) * 8u;
            if (blocks > maxBlocks) blocks = maxBlocks;
            pull_rows_affine_kernel<<<blocks, 256, 0, ctx->stream>>>(v->track, reinterpret_cast<const uint4*>(v->host), reinterpret_cast<float4*>(devp),
                                                                    v->count, v->pullRows);
            ctx->launches++;
            RVE_CUDA(cudaGetLastError());
        } else {
            rc = launch_pull(ctx, v->track, v->host, devp, v->count, v->elemBytes, v->pullRows);
        }
        if (rc) return rc;
        RVE_CUDA(cudaEventRecord(v->pullDone[v->cur], ctx->stream));
        v->pullBusy[v->cur] = true;
        // the host records the next frame's modifications in the other flag block; the kernel that last read it
        // (two syncs ago) must have finished clearing it
        v->cur ^= 1;
        if (v->pullBusy[v->cur]) RVE_CUDA(cudaEventSynchronize(v->pullDone[v->cur]));
        v->pullBusy[v->cur] = false;
        v->track = v->trackBuf[v->cur];
        v->anyDirty = false;
        v->touched.clear();  // the fresh flag block is clean
        v->touchedExact = true;
        st.packed_bytes = uint64_t(v->count);  // flag bytes the kernel reads; row bytes depend on the flags (rvs_vector_pulled_rows)
        pyramid_note_buffer_write(ctx, v->priv);
        if (stats) *stats = st;
        return 0;
    }

    if (!rangesReady) vector_ranges(v);
    v->anyDirty = false;
    const uint32_t nRanges = uint32_t(v->ranges.size() / 2);
    st.ranges = nRanges;
    if (nRanges == 0) {
        if (stats) *stats = st;
        return 0;
    }
    auto* dev = static_cast<unsigned char*>(rve_buffer_device_ptr(ctx, v->priv));
    if (!dev) return RVE_ERR_BAD_HANDLE;

    // Affine rows cannot be copied range by range into the 64-B private rows: every mode stages. Short ranges (PACKED)
    // stage only their modified rows + indices; direct ranges stage all their rows, contiguously, behind them.
    const bool packShort = mode == RVS_SYNC_PACKED;
    uint64_t packedRows = 0, directRows = 0;
    for (uint32_t r = 0; r < nRanges; ++r) {
        const uint32_t len = v->ranges[2 * r + 1] - v->ranges[2 * r];
        st.rows_in_ranges += len;
        if (packShort && len < direct_min_rows) {
            for (uint32_t k = v->ranges[2 * r]; k < v->ranges[2 * r + 1]; ++k) packedRows += v->track[k] ? 1u : 0u;
        } else {
            directRows += len;
        }
    }
    const uint64_t stagedDirect = v->affine ? directRows : 0;  // rows of direct ranges that travel through the staging block
    unsigned char* rowsDst = nullptr;
    uint32_t* idxDst = nullptr;
    size_t idxOffset = 0, directOffset = 0;
    if (packedRows || stagedDirect) {
        RVE_REQUIRE(packedRows <= 0xFFFFFFFFull, "too many packed rows");
        idxOffset = (size_t(packedRows) * eb + 15) & ~size_t(15);
        directOffset = (idxOffset + size_t(packedRows) * 4 + 15) & ~size_t(15);
        const size_t need = directOffset + size_t(stagedDirect) * eb;
        if (need > v->stageHostBytes) {
            if (v->stageBusy) RVE_CUDA(cudaEventSynchronize(v->stageFree));
            v->stageBusy = false;
            if (v->stageHost) cudaFreeHost(v->stageHost);
            const size_t cap = need + need / 2;
            RVE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&v->stageHost), cap, cudaHostAllocDefault));
            v->stageHostBytes = cap;
        }
        if (need > v->stageDevBytes) {
            if (v->stageDev != RVE_BUFFER_INVALID) rve_buffer_destroy(ctx, v->stageDev);
            const size_t cap = need + need / 2;
            const int rc = rve_buffer_create(ctx, cap, &v->stageDev);
            if (rc) return rc;
            v->stageDevBytes = cap;
        }
        if (v->stageBusy) RVE_CUDA(cudaEventSynchronize(v->stageFree));  // previous upload consumed the block
        v->stageBusy = false;
        rowsDst = v->stageHost;
        idxDst = reinterpret_cast<uint32_t*>(v->stageHost + idxOffset);
    }
    uint64_t pr = 0, dr = 0;
    struct DirectRun {
        uint32_t first, rows;
        uint64_t stagedAt;
    };
    std::vector<DirectRun> runs;
    for (uint32_t r = 0; r < nRanges; ++r) {
        const uint32_t b = v->ranges[2 * r], e = v->ranges[2 * r + 1], len = e - b;
        if (packShort && len < direct_min_rows) {
            for (uint32_t k = b; k < e; ++k)
                if (v->track[k]) {
                    std::memcpy(rowsDst + pr * eb, v->host + size_t(k) * eb, eb);
                    idxDst[pr++] = k;
                }
        } else if (v->affine) {
            std::memcpy(v->stageHost + directOffset + dr * eb, v->host + size_t(b) * eb, size_t(len) * eb);
            runs.push_back({b, len, dr});
            dr += len;
            st.direct_copies++;
            st.direct_bytes += uint64_t(len) * eb;
        } else {
            RVE_CUDA(cudaMemcpyAsync(dev + size_t(b) * eb, v->host + size_t(b) * eb, size_t(len) * eb, cudaMemcpyHostToDevice, ctx->stream));
            st.direct_copies++;
            st.direct_bytes += uint64_t(len) * eb;
        }
        std::memset(v->track + b, 0, len);
    }
    if (packedRows || stagedDirect) {
        auto* sdev = static_cast<unsigned char*>(rve_buffer_device_ptr(ctx, v->stageDev));
        const size_t bytes = directOffset + size_t(stagedDirect) * eb;
        RVE_CUDA(cudaMemcpyAsync(sdev, v->stageHost, bytes, cudaMemcpyHostToDevice, ctx->stream));
        RVE_CUDA(cudaEventRecord(v->stageFree, ctx->stream));
        v->stageBusy = true;
        const uint32_t maxBlocks = uint32_t(ctx->sm_count) * 8u;
        if (packedRows) {
            if (v->affine) {
                const uint32_t blocks = uint32_t(std::min<uint64_t>((packedRows * 4 + 255) / 256, maxBlocks));
                scatter_rows_affine_kernel<<<blocks, 256, 0, ctx->stream>>>(reinterpret_cast<float4*>(dev), reinterpret_cast<const uint32_t*>(sdev + idxOffset),
                                                                           reinterpret_cast<const float*>(sdev), uint32_t(packedRows));
                ctx->launches++;
                RVE_CUDA(cudaGetLastError());
            } else {
                const int rc = rve_scatter_rows(ctx, v->priv, v->elemBytes, v->stageDev, idxOffset, 0, uint32_t(packedRows));
                if (rc) return rc;
            }
        }
        for (const DirectRun& d : runs) {
            const uint32_t blocks = uint32_t(std::min<uint64_t>((uint64_t(d.rows) * 4 + 255) / 256, maxBlocks));
            expand_range_affine_kernel<<<blocks, 256, 0, ctx->stream>>>(reinterpret_cast<float4*>(dev),
                                                                       reinterpret_cast<const float*>(sdev + directOffset + d.stagedAt * eb), d.first, d.rows);
            ctx->launches++;
            RVE_CUDA(cudaGetLastError());
        }
        st.packed_rows = uint32_t(packedRows);
        st.packed_bytes = bytes;
    }
    pyramid_note_buffer_write(ctx, v->priv);
    if (stats) *stats = st;
    return 0;
}

// EncodeSync of SEVERAL vectors at once. Frames in which many structures change a little (entity lists, layers, attributes: a
// few hundred rows each) pay per vector one staging copy and one scatter launch - ~10 us of latency each. Here the modified
// rows of all vectors whose changes are small (<= 64 Ki modified rows, 4-byte-multiple rows, no reallocation due)
// travel in ONE pinned staging block and ONE copy, and are scattered by ONE launch per 16 vectors; every other vector takes
// its own rvs_vector_encode_sync(RVS_SYNC_PACKED). stats: one entry per vector (may be null).
int rvs_vectors_encode_sync_batch(rvs_vector* const* vs, uint32_t n, uint32_t direct_min_rows, rvs_sync_stats* stats) {
    if (n == 0) return 0;
    if (!vs) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_PROFILE_FN_N("EncodeSync (batch)");
    rve_ctx* ctx = nullptr;
    struct Item {
        uint32_t i;
        uint64_t rows;
    };
    std::vector<Item> batched;
    uint64_t totalRows = 0, totalWords = 0;
    for (uint32_t i = 0; i < n; ++i) {
        rvs_vector* v = vs[i];
        if (!v) return set_error(RVE_ERR_INVALID_ARGUMENT, "null vector");
        if (stats) stats[i] = rvs_sync_stats{};
        if (!ctx) ctx = v->ctx;
        bool small = v->ctx == ctx && !v->affine && (v->elemBytes % 4) == 0 && v->count > 0 && v->priv != RVE_BUFFER_INVALID &&
                     v->privBytes == size_t(v->capacity) * v->devElemBytes;
        uint64_t rows = 0;
        if (small && v->anyDirty) {
            vector_ranges(v);  // consumes the touched list: should this vector fall back to its own sync below, that sync takes these ranges
            v->rangesReady = true;
            // small = few modified rows in total, whatever the shape of their ranges (a run of appended rows is packed like the rest:
            // direct_min_rows only steers the single-vector path of the vectors that are handed on)
            for (size_t r = 0; r + 1 < v->ranges.size() && rows <= 65536; r += 2)
                for (uint32_t k = v->ranges[r]; k < v->ranges[r + 1]; ++k) rows += v->track[k] ? 1u : 0u;
            if (rows > 65536) small = false;
        }
        if (!small) {
            const int rc = rvs_vector_encode_sync(v, RVS_SYNC_PACKED, direct_min_rows, stats ? &stats[i] : nullptr);
            if (rc) return rc;
            continue;
        }
        if (!v->anyDirty || v->ranges.empty()) {  // clean: free, like the single-vector call
            v->rangesReady = false;
            v->anyDirty = false;
            v->ranges.clear();
            continue;
        }
        batched.push_back({i, rows});
        totalRows += rows;
        totalWords += rows * (v->elemBytes / 4);
    }
    if (batched.empty()) return 0;
    DeviceGuard g(ctx->device);
    const size_t idxOffset = size_t(totalWords) * 4;
    const size_t need = idxOffset + size_t(totalRows) * 4;
    if (need > ctx->sync_stage_bytes) {
        if (ctx->sync_stage_busy) RVE_CUDA(cudaEventSynchronize(ctx->sync_stage_free));
        ctx->sync_stage_busy = false;
        if (ctx->sync_stage_host) cudaFreeHost(ctx->sync_stage_host);
        if (ctx->sync_stage_dev) RVE_CUDA(cudaFreeAsync(ctx->sync_stage_dev, ctx->stream));
        const size_t cap = need + need / 2 + 4096;
        RVE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&ctx->sync_stage_host), cap, cudaHostAllocDefault));
        RVE_CUDA(cudaMallocAsync(&ctx->sync_stage_dev, cap, ctx->stream));
        ctx->sync_stage_bytes = cap;
        if (!ctx->sync_stage_free) RVE_CUDA(cudaEventCreateWithFlags(&ctx->sync_stage_free, cudaEventDisableTiming));
    }
    if (ctx->sync_stage_busy) RVE_CUDA(cudaEventSynchronize(ctx->sync_stage_free));  // the previous batch's copy consumed the block
    ctx->sync_stage_busy = false;
    auto* words = reinterpret_cast<uint32_t*>(ctx->sync_stage_host);
    auto* idx = reinterpret_cast<uint32_t*>(ctx->sync_stage_host + idxOffset);
    std::vector<SyncBatch> launches;
    uint64_t rowAt = 0, wordAt = 0;
    for (const Item& it : batched) {
        rvs_vector* v = vs[it.i];
        const uint32_t rowWords = v->elemBytes / 4;
        if (launches.empty() || launches.back().n == kSyncBatchSegs) launches.emplace_back();
        SyncBatch& b = launches.back();
        SyncBatchSeg& sg = b.s[b.n++];
        sg.dst = static_cast<uint32_t*>(rve_buffer_device_ptr(ctx, v->priv));
        if (!sg.dst) return RVE_ERR_BAD_HANDLE;
        sg.rowWords = rowWords;
        sg.rowStart = uint32_t(rowAt);
        sg.wordStart = uint32_t(wordAt);
        sg.words = uint32_t(it.rows * rowWords);
        rvs_sync_stats st{};
        st.ranges = uint32_t(v->ranges.size() / 2);
        for (size_t r = 0; r + 1 < v->ranges.size(); r += 2) {
            const uint32_t rb = v->ranges[r], re = v->ranges[r + 1];
            st.rows_in_ranges += re - rb;
            for (uint32_t k = rb; k < re; ++k)
                if (v->track[k]) {
                    std::memcpy(words + wordAt, v->host + size_t(k) * v->elemBytes, v->elemBytes);
                    wordAt += rowWords;
                    idx[rowAt++] = k;
                }
            std::memset(v->track + rb, 0, re - rb);
        }
        st.packed_rows = uint32_t(it.rows);
        st.packed_bytes = it.rows * (v->elemBytes + 4);
        if (stats) stats[it.i] = st;
        v->anyDirty = false;
        v->rangesReady = false;
        pyramid_note_buffer_write(ctx, v->priv);
    }
    RVE_CUDA(cudaMemcpyAsync(ctx->sync_stage_dev, ctx->sync_stage_host, need, cudaMemcpyHostToDevice, ctx->stream));
    RVE_CUDA(cudaEventRecord(ctx->sync_stage_free, ctx->stream));
    ctx->sync_stage_busy = true;
    const auto* dwords = static_cast<const uint32_t*>(ctx->sync_stage_dev);
    const auto* didx = reinterpret_cast<const uint32_t*>(static_cast<const unsigned char*>(ctx->sync_stage_dev) + idxOffset);
    for (SyncBatch& b : launches) {
        const uint32_t first = b.s[0].wordStart, last = b.s[b.n - 1].wordStart + b.s[b.n - 1].words;
        // the kernel indexes the staged words of ITS segments from zero
        for (uint32_t k = 0; k < b.n; ++k) b.s[k].wordStart -= first;
        b.totalWords = last - first;
        b.data = dwords + first;
        b.idx = didx;
        const uint32_t blocks = std::max<uint32_t>(1u, std::min<uint32_t>((b.totalWords + 255) / 256, uint32_t(ctx->sm_count) * 8u));
        scatter_rows_batch_kernel<<<blocks, 256, 0, ctx->stream>>>(b);
        ctx->launches++;
        RVE_CUDA(cudaGetLastError());
    }
    return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------
// BufferedVRAMSparseSet<entity_id_t, entity_id_t> stand-in: the per-command entity lists the cull walks
// (include/RavEngine/World.hpp:342,355; include/RavEngine/SparseSet.hpp:6-104; include/RavEngine/BufferedVRAMVector.hpp:129-170).
// Dense array = an rvs_vector of u32 (pinned host block + flags + private buffer), sparse -> dense map and reverse map
// on the host. Erase is a swap-remove: the last dense element moves into the hole, which is the only flag it sets.
// ---------------------------------------------------------------------------------------------------------
struct rvs_sparse_set {
    rvs_vector* dense = nullptr;
    std::vector<uint32_t> sparse{0xFFFFFFFFu};  // sparse_set{ default_index } (SparseSet.hpp:15)
    std::vector<uint32_t> reverse;              // reverse_map: dense index -> sparse index
    // change log for a consumer that maintains something per VALUE (the shim's sorted device copy of the list): (value << 1) |
    // entered, in call order. Off until sparse_set_track_changes; gives up (overflow) past kChangeLogMax entries.
    uint64_t version = 0;  // bumped by every Emplace / Erase that changes the set: consumers compare it instead of relying on a sync's statistics
    bool track = false, overflow = false;
    uint64_t logOwner = 0;  // the one consumer the log is kept for (taking it empties it: a second reader would miss changes)
    std::vector<uint64_t> log;
};
constexpr size_t kChangeLogMax = size_t(1) << 22;

namespace rve {
uint64_t sparse_set_version(rvs_sparse_set* s) { return s->version; }

bool sparse_set_track_changes(rvs_sparse_set* s, uint64_t owner) {
    if (s->track && s->logOwner != owner) return false;  // somebody else's log: this consumer keeps rebuilding from the list
    s->track = true;
    s->logOwner = owner;
    s->overflow = false;
    s->log.clear();
    return true;
}

bool sparse_set_take_changes(rvs_sparse_set* s, uint64_t owner, std::vector<uint32_t>* entered, std::vector<uint32_t>* left) {
    entered->clear();
    left->clear();
    if (!s->track || s->logOwner != owner) return false;
    const bool ok = !s->overflow;
    if (ok) {
        // values are unique inside the set, so the calls on one value alternate: the net effect is the sum of +1 / -1 per value
        std::sort(s->log.begin(), s->log.end());
        for (size_t i = 0; i < s->log.size();) {
            const uint32_t value = uint32_t(s->log[i] >> 1);
            int net = 0;
            for (; i < s->log.size() && uint32_t(s->log[i] >> 1) == value; ++i) net += (s->log[i] & 1u) ? 1 : -1;
            if (net > 0) entered->push_back(value);
            else if (net < 0) left->push_back(value);
        }
    }
    s->overflow = false;
    s->log.clear();
    return ok;
}
}  // namespace rve

static void log_change(rvs_sparse_set* s, uint32_t value, bool entered) {
    ++s->version;
    if (!s->track || s->overflow) return;
    if (s->log.size() >= kChangeLogMax) {
        s->overflow = true;
        s->log.clear();
        s->log.shrink_to_fit();
        return;
    }
    s->log.push_back((uint64_t(value) << 1) | (entered ? 1u : 0u));
}

extern "C" {

int rvs_sparse_set_create(rve_ctx* ctx, rvs_sparse_set** out) {
    if (!ctx || !out) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    auto* s = new rvs_sparse_set();
    const int rc = rvs_vector_create(ctx, 4, 0, &s->dense);
    if (rc) {
        delete s;
        return rc;
    }
    *out = s;
    return 0;
}

int rvs_sparse_set_destroy(rvs_sparse_set* s) {
    if (!s) return 0;
    rvs_vector_destroy(s->dense);
    delete s;
    return 0;
}

uint32_t rvs_sparse_set_dense_size(rvs_sparse_set* s) { return s ? rvs_vector_size(s->dense) : 0; }

int rvs_sparse_set_has(rvs_sparse_set* s, uint32_t sparse_index) {
    return s && sparse_index < s->sparse.size() && s->sparse[sparse_index] != 0xFFFFFFFFu;  // HasForSparseIndex (SparseSet.hpp:58-60)
}

// Emplace (SparseSet.hpp:22-32 + BufferedVRAMVector.hpp:153-158): no-op when the index already has a record
int rvs_sparse_set_emplace(rvs_sparse_set* s, uint32_t sparse_index, uint32_t value) {
    if (!s) return set_error(RVE_ERR_INVALID_ARGUMENT, "null set");
    RVE_REQUIRE(sparse_index != 0xFFFFFFFFu, "0xFFFFFFFF is the invalid index");
    if (rvs_sparse_set_has(s, sparse_index)) return 0;
    const uint32_t n = rvs_vector_size(s->dense);
    int rc = rvs_vector_resize(s->dense, n + 1);  // the new element starts out modified
    if (rc) return rc;
    rc = rvs_vector_set(s->dense, n, &value);
    if (rc) return rc;
    s->reverse.push_back(sparse_index);
    if (sparse_index >= s->sparse.size()) {
        const size_t want = (size_t(sparse_index) + 1 + 1) / 2 * 2;  // closest_multiple_of<int>(sparse_index + 1, 2)
        s->sparse.resize(want, 0xFFFFFFFFu);
    }
    s->sparse[sparse_index] = n;
    log_change(s, value, true);
    return 0;
}

// EraseAtSparseIndex (BufferedVRAMVector.hpp:160-163 + SparseSet.hpp:34-47): swap-remove
int rvs_sparse_set_erase(rvs_sparse_set* s, uint32_t sparse_index) {
    if (!s) return set_error(RVE_ERR_INVALID_ARGUMENT, "null set");
    RVE_REQUIRE(rvs_sparse_set_has(s, sparse_index), "no record for this sparse index");  // the reference asserts
    const uint32_t denseidx = s->sparse[sparse_index];
    const uint32_t last = rvs_vector_size(s->dense) - 1;
    auto* host = static_cast<uint32_t*>(rvs_vector_host(s->dense));
    log_change(s, host[denseidx], false);
    if (denseidx < last) {  // a move happens: the last element lands in the hole
        const uint32_t moved = host[last];
        const int rc = rvs_vector_set(s->dense, denseidx, &moved);  // marks the dense index modified
        if (rc) return rc;
        const uint32_t ownerOfMoved = s->reverse.back();
        s->sparse[ownerOfMoved] = denseidx;
        s->reverse[denseidx] = ownerOfMoved;
    }
    s->reverse.pop_back();
    s->sparse[sparse_index] = 0xFFFFFFFFu;
    return rvs_vector_resize(s->dense, last);
}

int rvs_sparse_set_emplace_many(rvs_sparse_set* s, const uint32_t* sparse_indices, const uint32_t* values, uint32_t n) {
    if (!s || (n && (!sparse_indices || !values))) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    for (uint32_t i = 0; i < n; ++i) {
        const int rc = rvs_sparse_set_emplace(s, sparse_indices[i], values[i]);
        if (rc) return rc;
    }
    return 0;
}

int rvs_sparse_set_erase_many(rvs_sparse_set* s, const uint32_t* sparse_indices, uint32_t n) {
    if (!s || (n && !sparse_indices)) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    for (uint32_t i = 0; i < n; ++i) {
        const int rc = rvs_sparse_set_erase(s, sparse_indices[i]);
        if (rc) return rc;
    }
    return 0;
}

const uint32_t* rvs_sparse_set_dense(rvs_sparse_set* s) { return s ? static_cast<const uint32_t*>(rvs_vector_host(s->dense)) : nullptr; }
const uint32_t* rvs_sparse_set_reverse_map(rvs_sparse_set* s) { return s && !s->reverse.empty() ? s->reverse.data() : nullptr; }
rvs_vector* rvs_sparse_set_vector(rvs_sparse_set* s) { return s ? s->dense : nullptr; }

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------
// Upload forwarding between ranks of one box (rve_shim.h): shared host blocks, flags, the forwarding copy
// ---------------------------------------------------------------------------------------------------------
namespace rve {

__global__ void flag_signal_kernel(uint32_t* word, uint32_t value) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(word), "r"(value) : "memory");
}

__global__ void flag_wait_kernel(const uint32_t* word, uint32_t value, uint32_t* status, unsigned long long timeoutCycles) {
    const long long t0 = clock64();
    for (;;) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(word) : "memory");
        if (int32_t(v - value) >= 0) break;
        if (uint64_t(clock64() - t0) > timeoutCycles) {
            atomicOr(status, 1u);
            break;
        }
        __nanosleep(200);
    }
}

// loaded when the context is created: a first (lazy-loading) launch of one of these while a peer's wait kernel already spins
// on this device would stall until that wait times out (see p2p_preload)
int sync_preload() {
    cudaFuncAttributes a{};
    RVE_CUDA(cudaFuncGetAttributes(&a, flag_signal_kernel));
    RVE_CUDA(cudaFuncGetAttributes(&a, flag_wait_kernel));
    RVE_CUDA(cudaFuncGetAttributes(&a, full_rows_kernel<true>));
    RVE_CUDA(cudaFuncGetAttributes(&a, full_rows_kernel<false>));
    return 0;
}

}  // namespace rve

extern "C" {

int rvs_vector_encode_sync_rows(rvs_vector* v, uint32_t mode, uint32_t row_begin, uint32_t row_end, rvs_sync_stats* stats) {
    if (!v) return set_error(RVE_ERR_INVALID_ARGUMENT, "null vector");
    RVE_PROFILE_FN_N("EncodeSync");
    RVE_REQUIRE(mode == RVS_SYNC_FULL || mode == RVS_SYNC_FULL_PULL, "row-range syncs are whole-row uploads (RVS_SYNC_FULL / _FULL_PULL)");
    RVE_REQUIRE(row_begin <= row_end && row_end <= v->count, "row range out of bounds");
    RVE_REQUIRE(v->priv != RVE_BUFFER_INVALID && v->privBytes == size_t(v->capacity) * v->devElemBytes,
                "sync the whole vector once before row-range syncs (the private buffer does not exist yet)");
    DeviceGuard g(v->ctx->device);
    rvs_sync_stats st{};
    const int rc = sync_full(v, mode, &st, row_begin, row_end);
    if (stats) *stats = st;
    return rc;
}

int rvs_vector_share(rvs_vector* v, int32_t* pid, int32_t* fd, uint64_t* bytes) {
    if (!v || !pid || !fd || !bytes) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_REQUIRE(v->sharedHost && v->host, "not a RVS_VECTOR_SHARED_HOST vector (or still empty)");
    *pid = int32_t(getpid());
    *fd = v->hostFd;
    *bytes = v->hostMapBytes;
    return 0;
}

int rve_host_import(rve_ctx* ctx, int32_t pid, int32_t fd, uint64_t bytes, void** out_ptr) {
    if (!ctx || !out_ptr) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    char path[64];
    snprintf(path, sizeof(path), "/proc/%d/fd/%d", int(pid), int(fd));
    const int myfd = open(path, O_RDWR);
    if (myfd < 0) return set_error(RVE_ERR_INVALID_ARGUMENT, "cannot open %s (the owner's host block)", path);
    void* m = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, myfd, 0);
    close(myfd);  // the mapping keeps the memory alive
    if (m == MAP_FAILED) return set_error(RVE_ERR_OUT_OF_MEMORY, "mmap of the owner's host block failed");
    const int rc = check_cuda(cudaHostRegister(m, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped), "cudaHostRegister");
    if (rc) {
        munmap(m, bytes);
        return rc;
    }
    *out_ptr = m;
    return 0;
}

int rve_host_release(rve_ctx* ctx, void* ptr, uint64_t bytes) {
    if (!ctx || !ptr) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    RVE_CUDA(cudaStreamSynchronize(ctx->stream));
    cudaHostUnregister(ptr);
    munmap(ptr, bytes);
    return 0;
}

int rve_flag_signal(rve_ctx* ctx, rve_buffer flags, uint32_t word, uint32_t value) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    DeviceGuard g(ctx->device);
    BufferRec* b = get_buffer(ctx, flags, "flags");
    if (!b) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE((uint64_t(word) + 1) * 4 <= b->bytes, "flag word outside the buffer");
    flag_signal_kernel<<<1, 1, 0, ctx->stream>>>(static_cast<uint32_t*>(b->ptr) + word, value);
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

int rve_flag_wait(rve_ctx* ctx, rve_buffer flags, uint32_t word, uint32_t value, rve_buffer status) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    DeviceGuard g(ctx->device);
    BufferRec* b = get_buffer(ctx, flags, "flags");
    BufferRec* s = get_buffer(ctx, status, "status");
    if (!b || !s) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE((uint64_t(word) + 1) * 4 <= b->bytes && s->bytes >= 4, "flag word outside the buffer");
    flag_wait_kernel<<<1, 1, 0, ctx->stream>>>(static_cast<const uint32_t*>(b->ptr) + word, value, static_cast<uint32_t*>(s->ptr), 4000000000ull);
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

int rvs_forward_rows(rve_ctx* ctx, const void* host_rows, uint32_t row_bytes, uint32_t affine, uint64_t first_row, uint64_t rows, rve_buffer dst) {
    if (!ctx || !host_rows) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_PROFILE_FN_N("EncodeSync (forwarded rows)");
    DeviceGuard g(ctx->device);
    if (!rows) return 0;
    BufferRec* d = get_buffer(ctx, dst, "forward destination");
    if (!d) return RVE_ERR_BAD_HANDLE;
    const uint32_t devRow = affine ? 64u : row_bytes;
    RVE_REQUIRE(!affine || row_bytes == 48, "affine rows are 48-byte host rows");
    RVE_REQUIRE(affine || row_bytes % 16 == 0, "forwarded rows are copied in 16-byte chunks");
    RVE_REQUIRE((first_row + rows) * devRow <= d->bytes, "forwarded rows past the end of the destination");
    const auto* src = static_cast<const unsigned char*>(host_rows) + first_row * row_bytes;
    auto* out = static_cast<unsigned char*>(d->ptr) + first_row * devRow;
    RVE_REQUIRE(((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(out)) & 15) == 0, "forwarded rows must start 16-byte aligned");
    return launch_full(ctx, ctx->stream, src, out, rows, row_bytes, affine != 0);
}

}  // extern "C"
