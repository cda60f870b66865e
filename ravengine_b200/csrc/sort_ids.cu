// sort_ids.cu — ascending copy of an entity-id list (list maintenance for the cull).
//
// The cull's result does not depend on the ORDER of a command's entity list (defaultcull.csh:150-170 maps invocation ->
// list position -> entity id; the appends are unordered anyway) - except in single-instance mode, where the position is
// the output slot. The engine's lists are BufferedVRAMSparseSet dense arrays: every erase swap-removes, so a live list
// is a shuffle of short runs (10 % churn of a 2.5 M-entity list leaves runs of ~3.7 ids) and the cull falls from its
// TMA tile path (0.81 of the HBM peak) to the gather path (0.54; 0.26 for a full shuffle: random 64-B DRAM accesses).
// Keeping an ascending copy of the list on the device restores runs wherever the ids are consecutive and sector /
// DRAM-page locality everywhere else. It is rebuilt only when the list changed.
//
// Ids of a list are unique (sparse-set semantics), so sorting is a bitmap pass: set one bit per id over the entity id
// space, count bits per 1024-word block, scan the block counts, expand the bits in order. Five small launches,
// O(id space / 32 + list) words of traffic; no comparison sort, no library call.
#include "internal.h"

namespace rve {

constexpr uint32_t kSortBlock = 1024;  // bitmap words per block (32 768 ids)

__global__ void ids_set_bits_kernel(const uint32_t* __restrict__ ids, uint32_t count, uint32_t idLimit, uint32_t* __restrict__ bitmap,
                                    uint32_t* __restrict__ outOfRange) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const uint32_t id = __ldg(ids + i);
        if (id < idLimit) atomicOr(bitmap + (id >> 5), 1u << (id & 31u));
        else atomicAdd(outOfRange, 1u);
    }
}

// inclusive scan of one value per thread over a 1024-thread block
__device__ __forceinline__ uint32_t block_scan_1024(uint32_t x, uint32_t* s_warp) {
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= uint32_t(o)) x += y;
    }
    __syncthreads();
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = s_warp[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= uint32_t(o)) w += y;
        }
        s_warp[lane] = w;
    }
    __syncthreads();
    return x + (warp ? s_warp[warp - 1] : 0u);
}

__global__ void __launch_bounds__(kSortBlock) ids_block_counts_kernel(const uint32_t* __restrict__ bitmap, uint32_t words, uint32_t* __restrict__ blockSums) {
    __shared__ uint32_t s_warp[32];
    const uint32_t w = blockIdx.x * kSortBlock + threadIdx.x;
    const uint32_t incl = block_scan_1024(w < words ? uint32_t(__popc(bitmap[w])) : 0u, s_warp);
    if (threadIdx.x == kSortBlock - 1) blockSums[blockIdx.x] = incl;
}

// exclusive scan of the block counts in place, one block; total -> blockSums[blocks]
__global__ void __launch_bounds__(kSortBlock) ids_scan_blocks_kernel(uint32_t* __restrict__ blockSums, uint32_t blocks) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < blocks; base += kSortBlock) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < blocks ? blockSums[i] : 0u;
        const uint32_t incl = block_scan_1024(v, s_warp) + s_carry;
        if (i < blocks) blockSums[i] = incl - v;
        __syncthreads();
        if (threadIdx.x == kSortBlock - 1) s_carry = incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) blockSums[blocks] = s_carry;
}

__global__ void __launch_bounds__(kSortBlock) ids_expand_kernel(const uint32_t* __restrict__ bitmap, uint32_t words, const uint32_t* __restrict__ blockSums,
                                                                 uint32_t* __restrict__ out, uint32_t count) {
    __shared__ uint32_t s_warp[32];
    const uint32_t w = blockIdx.x * kSortBlock + threadIdx.x;
    uint32_t bits = w < words ? bitmap[w] : 0u;
    const uint32_t n = uint32_t(__popc(bits));
    uint32_t pos = blockSums[blockIdx.x] + block_scan_1024(n, s_warp) - n;
    while (bits) {
        const uint32_t b = uint32_t(__ffs(int(bits))) - 1u;
        if (pos < count) out[pos] = (w << 5) + b;
        ++pos;
        bits &= bits - 1u;
    }
}

}  // namespace rve

using namespace rve;

extern "C" int rve_sort_entity_ids(rve_ctx* ctx, rve_buffer entity_ids, uint32_t count, uint32_t id_limit, rve_buffer sorted_out) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    RVE_PROFILE_FN_N("Sort entity ids");
    DeviceGuard g(ctx->device);
    if (count == 0) return 0;
    BufferRec* bi = get_buffer(ctx, entity_ids, "entity ids");
    BufferRec* bo = get_buffer(ctx, sorted_out, "sorted entity ids");
    if (!bi || !bo) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE(uint64_t(count) * 4 <= bi->bytes && uint64_t(count) * 4 <= bo->bytes, "id buffers smaller than count entries");
    RVE_REQUIRE(id_limit >= 1, "id_limit is the size of the entity id space (ids are < id_limit)");
    const uint32_t words = (id_limit + 31u) / 32u;
    const uint32_t blocks = (words + kSortBlock - 1) / kSortBlock;
    // scratch: bitmap words, block sums (+ total), an out-of-range counter
    const size_t need = (size_t(blocks) * kSortBlock + blocks + 2) * sizeof(uint32_t);
    int rc = ensure_scratch(ctx, &ctx->d_sort, &ctx->d_sort_cap, need);
    if (rc) return rc;
    uint32_t* bitmap = static_cast<uint32_t*>(ctx->d_sort);
    uint32_t* sums = bitmap + size_t(blocks) * kSortBlock;
    uint32_t* bad = sums + blocks + 1;
    RVE_CUDA(cudaMemsetAsync(ctx->d_sort, 0, need, ctx->stream));
    const uint32_t setBlocks = std::min<uint32_t>((count + 255) / 256, uint32_t(ctx->sm_count) * 8u);
    ids_set_bits_kernel<<<setBlocks, 256, 0, ctx->stream>>>(static_cast<const uint32_t*>(bi->ptr), count, id_limit, bitmap, bad);
    ids_block_counts_kernel<<<blocks, kSortBlock, 0, ctx->stream>>>(bitmap, words, sums);
    ids_scan_blocks_kernel<<<1, kSortBlock, 0, ctx->stream>>>(sums, blocks);
    ids_expand_kernel<<<blocks, kSortBlock, 0, ctx->stream>>>(bitmap, words, sums, static_cast<uint32_t*>(bo->ptr), count);
    ctx->launches += 4;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

// ---- rve_sort_entity_ids_static: the same bitmap pass with its scratch in a file-scope device array ----
//
// For callers that must not reach cudaMalloc (stream capture, serving loops, harnesses that forbid it). The scratch of
// the largest id space the variant accepts (RVE_SORT_STATIC_MAX_IDS ids: bitmap words, block sums + total, the
// out-of-range counter) lives in g_sort_static; every GPU that loads the library holds it until the process exits.
// In host code `g_sort_static` is only a stand-in for the symbol: the kernels are given the device address that
// cudaGetSymbolAddress reports, looked up once per device with that device current.
namespace rve {

constexpr uint32_t kStaticMaxIds = RVE_SORT_STATIC_MAX_IDS;
constexpr uint32_t kStaticMaxWords = kStaticMaxIds / 32u;
constexpr uint32_t kStaticMaxBlocks = kStaticMaxWords / kSortBlock;
static_assert(kStaticMaxIds % (32u * kSortBlock) == 0, "RVE_SORT_STATIC_MAX_IDS must be a whole number of bitmap blocks");
constexpr size_t kStaticWords = size_t(kStaticMaxBlocks) * kSortBlock + kStaticMaxBlocks + 2;

static __device__ __align__(256) uint32_t g_sort_static[kStaticWords];

constexpr int kStaticMaxDevices = 64;
static std::mutex g_sort_static_mutex;
static uint32_t* g_sort_static_addr[kStaticMaxDevices] = {};

// device address of g_sort_static on `device` (which must be current); nullptr + error code on failure
static int sort_static_address(int device, uint32_t** out) {
    if (device < 0 || device >= kStaticMaxDevices) return set_error(RVE_ERR_UNSUPPORTED, "device ordinal %d is past the %d this entry point tracks", device, kStaticMaxDevices);
    std::lock_guard<std::mutex> lock(g_sort_static_mutex);
    if (!g_sort_static_addr[device]) {
        void* p = nullptr;
        RVE_CUDA(cudaGetSymbolAddress(&p, g_sort_static));
        if (!p) return set_error(RVE_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static sort scratch");
        g_sort_static_addr[device] = static_cast<uint32_t*>(p);
    }
    *out = g_sort_static_addr[device];
    return 0;
}

}  // namespace rve

extern "C" int rve_sort_entity_ids_static(rve_ctx* ctx, rve_buffer entity_ids, uint32_t count, uint32_t id_limit, rve_buffer sorted_out) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    DeviceGuard g(ctx->device);
    if (count == 0) return 0;
    BufferRec* bi = get_buffer(ctx, entity_ids, "entity ids");
    BufferRec* bo = get_buffer(ctx, sorted_out, "sorted entity ids");
    if (!bi || !bo) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE(uint64_t(count) * 4 <= bi->bytes && uint64_t(count) * 4 <= bo->bytes, "id buffers smaller than count entries");
    RVE_REQUIRE(id_limit >= 1, "id_limit is the size of the entity id space (ids are < id_limit)");
    if (id_limit > kStaticMaxIds)
        return set_error(RVE_ERR_UNSUPPORTED, "rve_sort_entity_ids_static: id_limit %u is past RVE_SORT_STATIC_MAX_IDS (%u); use rve_sort_entity_ids", id_limit,
                         kStaticMaxIds);
    const uint32_t words = (id_limit + 31u) / 32u;
    const uint32_t blocks = (words + kSortBlock - 1) / kSortBlock;
    uint32_t* scratch = nullptr;
    int rc = sort_static_address(ctx->device, &scratch);
    if (rc) return rc;
    // same layout as the allocating entry point, inside the fixed array
    uint32_t* bitmap = scratch;
    uint32_t* sums = bitmap + size_t(blocks) * kSortBlock;
    uint32_t* bad = sums + blocks + 1;
    // the array keeps the last call's contents: the bitmap and the counter must start at zero (the sums are overwritten)
    RVE_CUDA(cudaMemsetAsync(scratch, 0, (size_t(blocks) * kSortBlock + blocks + 2) * sizeof(uint32_t), ctx->stream));
    const uint32_t setBlocks = std::min<uint32_t>((count + 255) / 256, uint32_t(ctx->sm_count) * 8u);
    ids_set_bits_kernel<<<setBlocks, 256, 0, ctx->stream>>>(static_cast<const uint32_t*>(bi->ptr), count, id_limit, bitmap, bad);
    ids_block_counts_kernel<<<blocks, kSortBlock, 0, ctx->stream>>>(bitmap, words, sums);
    ids_scan_blocks_kernel<<<1, kSortBlock, 0, ctx->stream>>>(sums, blocks);
    ids_expand_kernel<<<blocks, kSortBlock, 0, ctx->stream>>>(bitmap, words, sums, static_cast<uint32_t*>(bo->ptr), count);
    ctx->launches += 4;
    RVE_CUDA(cudaGetLastError());
    return 0;
}
