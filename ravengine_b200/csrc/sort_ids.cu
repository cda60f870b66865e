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
// space, count bits per 1024-word block, scan the block counts, expand the bits in order. O(id space / 32 + list) words
// of traffic; no comparison sort, no library call. rve_sort_entity_ids / _batch do it in two launches for up to 16 lists
// (further down); the four-kernel sequence right below is what rve_sort_entity_ids_static still runs over its fixed scratch.
#include <cstring>

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

// ---- the same sort in TWO launches, for up to kSortBatch lists at once (r02m) ----
// One list cost five stream operations (memset, set bits, block counts, scan, expand: 48 us for 2.5 M ids, mostly launch
// latency), and the shim issued them list by list. Now: (1) one launch sets the bits of every list of the batch and resets
// the look-back state; (2) one launch does counts, scan and expansion in a single pass - every block publishes the popcount
// of its 1024 bitmap words, obtains the sum of its predecessors by a warp-wide decoupled look-back (32 predecessors per
// step; blocks take their index from a ticket, so a predecessor has always started) and expands its bits. The bitmap words
// are cleared as they are read: the scratch is left ready for the next sort, no memset.
constexpr uint32_t kSortBatch = 16;
struct SortList {
    const uint32_t* ids;    // full build: the list itself
    uint32_t* out;
    uint32_t* region;       // bitmap | look-back state | ticket, out-of-range count (sort_region_words)
    const uint32_t* delta;  // persistent bitmap, small change: nSet ids that entered the list, then nClear ids that left it
    uint32_t count;         // entries of the list (= ids the expansion writes)
    uint32_t nSet, nClear;
    uint32_t keep;          // 1: the bitmap belongs to the list and survives the expansion (next change: a delta); 0: scratch, cleared as it is read
};
struct SortBatch {
    SortList l[kSortBatch];
};
// per-list scratch region (words): bitmap [blocks * 1024] | look-back state, u64 per block [2 * blocks] | ticket, out-of-range count
__host__ __device__ inline uint32_t sort_region_words(uint32_t blocks) { return blocks * kSortBlock + 2u * blocks + 2u; }

__global__ void ids_set_bits_batch_kernel(const SortBatch b, uint32_t idLimit, uint32_t blocks) {
    const SortList& L = b.l[blockIdx.y];
    uint32_t* region = L.region;
    uint32_t* ctl = region + size_t(blocks) * kSortBlock + 2u * blocks;
    // look-back state and ticket of the expand launch that follows (nobody reads them during this one)
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < 2u * blocks; i += gridDim.x * blockDim.x) region[size_t(blocks) * kSortBlock + i] = 0u;
    if (blockIdx.x == 0 && threadIdx.x == 0) ctl[0] = 0u, ctl[1] = 0u;
    // Engine lists are runs of nearly consecutive ids: neighbouring lanes of a warp mostly hit the same bitmap word, and
    // same-address atomics serialise. Runs of ADJACENT lanes with the same word are merged - heads by a neighbour compare, a
    // segmented OR over five shuffle steps - and the head issues ONE atomicOr per run (match_any would also merge lanes that
    // are not adjacent, but costs one pass per distinct word: 73 us on random ids, where this costs nothing extra).
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * blockDim.x;
    if (L.delta) {
        // the list's own bitmap, a small change: ids that entered are set, ids that left are cleared (disjoint by construction)
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < L.nSet + L.nClear; i += stride) {
            const uint32_t id = __ldg(L.delta + i);
            if (id >= idLimit) {
                if (i < L.nSet) atomicAdd(ctl + 1, 1u);
            } else if (i < L.nSet) {
                atomicOr(region + (id >> 5), 1u << (id & 31u));
            } else {
                atomicAnd(region + (id >> 5), ~(1u << (id & 31u)));
            }
        }
        return;
    }
    for (uint32_t base = blockIdx.x * blockDim.x + (threadIdx.x & ~31u); base < L.count; base += stride) {  // warp-uniform trip count
        const uint32_t i = base + lane;
        uint32_t id = 0xFFFFFFFFu;
        if (i < L.count) id = __ldg(L.ids + i);
        const bool valid = i < L.count && id < idLimit;
        if (i < L.count && !valid) atomicAdd(ctl + 1, 1u);
        const uint32_t key = valid ? (id >> 5) : (0x80000000u | lane);  // invalid lanes: a run of their own
        const uint32_t prev = __shfl_up_sync(0xffffffffu, key, 1);
        const uint32_t heads = __ballot_sync(0xffffffffu, lane == 0u || key != prev);
        const uint32_t after = heads & ~((2u << lane) - 1u);                        // heads behind this lane
        const uint32_t last = after ? uint32_t(__ffs(int(after))) - 2u : 31u;       // last lane of this lane's run
        uint32_t bits = valid ? (1u << (id & 31u)) : 0u;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t other = __shfl_down_sync(0xffffffffu, bits, o);
            if (lane + uint32_t(o) <= last) bits |= other;
        }
        if (valid && ((heads >> lane) & 1u)) atomicOr(region + (id >> 5), bits);
    }
}

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// One CTA covers kLbSub sub-blocks of 1024 bitmap words: shorter look-back chains against fewer, heavier CTAs. Measured on the
// four churned lists of C2 (ncu, one launch): 47.5 us at 1, 38.4 us at 2, 41.8 us at 4 - the dense id range of a list keeps
// only a quarter of its CTAs busy (busiest SM active 86 k cycles, mean 36 k), so fewer CTAs soon cost more than the chain saves.
constexpr uint32_t kLbSub = 2;
__global__ void __launch_bounds__(kSortBlock) ids_expand_lookback_kernel(const SortBatch b, uint32_t words, uint32_t blocks) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_blk, s_prefix, s_tot[kLbSub];
    const SortList& L = b.l[blockIdx.y];
    uint32_t* region = L.region;
    unsigned long long* state = reinterpret_cast<unsigned long long*>(region + size_t(blocks) * kSortBlock);  // (status << 32) | value; status 1 aggregate, 2 prefix
    uint32_t* ctl = region + size_t(blocks) * kSortBlock + 2u * blocks;
    if (threadIdx.x == 0) s_blk = atomicAdd(ctl, 1u);
    __syncthreads();
    const uint32_t blk = s_blk;  // CTA index in start order: its predecessors have started
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t bits[kLbSub], n[kLbSub], incl[kLbSub];
#pragma unroll
    for (uint32_t k = 0; k < kLbSub; ++k) {
        const uint32_t w = (blk * kLbSub + k) * kSortBlock + threadIdx.x;
        bits[k] = 0;
        if (w < words) {
            bits[k] = region[w];
            if (bits[k] && !L.keep) region[w] = 0u;  // scratch bitmaps are clean again when this launch ends
        }
        n[k] = uint32_t(__popc(bits[k]));
    }
#pragma unroll
    for (uint32_t k = 0; k < kLbSub; ++k) {
        incl[k] = block_scan_1024(n[k], s_warp);
        if (threadIdx.x == kSortBlock - 1) s_tot[k] = incl[k];
    }
    __syncthreads();
    if ((threadIdx.x >> 5) == 31u) {  // the last warp does the look-back
        uint32_t total = 0;
#pragma unroll
        for (uint32_t k = 0; k < kLbSub; ++k) total += s_tot[k];
        uint32_t prefix = 0;
        if (blk == 0) {
            if (lane == 0) st_release_u64(state, (2ull << 32) | total);
        } else {
            if (lane == 0) st_release_u64(state + blk, (1ull << 32) | total);
            int base = int(blk) - 1;
            for (;;) {
                const int j = base - int(lane);
                const unsigned long long v = j >= 0 ? ld_acquire_u64(state + j) : (2ull << 32);
                const uint32_t status = uint32_t(v >> 32);
                const uint32_t ready = __ballot_sync(0xffffffffu, status != 0u), closed = __ballot_sync(0xffffffffu, status == 2u);
                const uint32_t upto = closed ? uint32_t(__ffs(int(closed))) - 1u : 31u;  // last lane that contributes
                const uint32_t need = upto == 31u ? 0xffffffffu : ((2u << upto) - 1u);
                if ((ready & need) != need) continue;  // a predecessor in the window has not published yet: read again
                uint32_t val = lane <= upto ? uint32_t(v) : 0u;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
                prefix += val;
                if (closed) break;
                base -= 32;
            }
            if (lane == 0) st_release_u64(state + blk, (2ull << 32) | (prefix + total));
        }
        if (lane == 0) s_prefix = prefix;
    }
    __syncthreads();
    // expansion, a warp at a time over its 32 words of every sub-block: the lane whose bit is set writes that id at its rank inside
    // the word, so the ids of one word leave as ONE coalesced store (a thread expanding its own word wrote 32 scattered sectors per store)
    uint32_t* __restrict__ out = L.out;
    uint32_t before = s_prefix;
#pragma unroll
    for (uint32_t k = 0; k < kLbSub; ++k) {
        const uint32_t pos = before + incl[k] - n[k];
        const uint32_t wordBase = (blk * kLbSub + k) * kSortBlock + threadIdx.x - lane;
        const uint32_t nonEmpty = __ballot_sync(0xffffffffu, bits[k] != 0u);
        for (uint32_t todo = nonEmpty; todo; todo &= todo - 1u) {
            const uint32_t wi = uint32_t(__ffs(int(todo))) - 1u;
            const uint32_t bw = __shfl_sync(0xffffffffu, bits[k], wi);
            const uint32_t pw = __shfl_sync(0xffffffffu, pos, wi);
            if ((bw >> lane) & 1u) {
                const uint32_t at = pw + uint32_t(__popc(bw & ((1u << lane) - 1u)));
                if (at < L.count) out[at] = ((wordBase + wi) << 5) + lane;
            }
        }
        before += s_tot[k];
    }
}

// sorts lists[0..n) (n <= kSortBatch) of one id space with two launches; the context's sort scratch holds one region per list
static int sort_batch(rve_ctx* ctx, SortBatch& batch, uint32_t n, uint32_t maxCount, uint32_t id_limit) {
    const uint32_t words = (id_limit + 31u) / 32u;
    const uint32_t blocks = (words + kSortBlock - 1) / kSortBlock;
    const size_t need = size_t(n) * sort_region_words(blocks) * sizeof(uint32_t);
    void* before = ctx->d_sort;
    int rc = ensure_scratch(ctx, &ctx->d_sort, &ctx->d_sort_cap, need);
    if (rc) return rc;
    // the bitmaps must be zero when the first launch starts: they are after every completed sort of the same geometry
    const uint64_t geometry = (uint64_t(blocks) << 8) | n;
    if (ctx->d_sort != before || ctx->sort_clean_geometry != geometry || ctx->sort_dirty) {
        RVE_CUDA(cudaMemsetAsync(ctx->d_sort, 0, ctx->d_sort_cap, ctx->stream));
        ctx->sort_clean_geometry = geometry;
    }
    ctx->sort_dirty = true;
    uint32_t* scratch = static_cast<uint32_t*>(ctx->d_sort);
    for (uint32_t i = 0; i < n; ++i) batch.l[i].region = scratch + size_t(i) * sort_region_words(blocks);
    const uint32_t setBlocks = std::max<uint32_t>(1u, std::min<uint32_t>((maxCount + 255) / 256, uint32_t(ctx->sm_count) * 8u / n + 1u));
    ids_set_bits_batch_kernel<<<dim3(setBlocks, n), 256, 0, ctx->stream>>>(batch, id_limit, blocks);
    ids_expand_lookback_kernel<<<dim3((blocks + kLbSub - 1) / kLbSub, n), kSortBlock, 0, ctx->stream>>>(batch, words, blocks);
    ctx->launches += 2;
    RVE_CUDA(cudaGetLastError());
    ctx->sort_dirty = false;
    return 0;
}

// Lists that own their bitmap (internal.h: SortedListUpdate): a full build zeroes the bitmap and sets the bits of the whole
// list, a delta only touches the ids that entered or left; the expansion leaves the bitmap in place. Two launches per 16 lists.
size_t sort_region_bytes(uint32_t id_limit) {
    const uint32_t words = (id_limit + 31u) / 32u;
    return size_t(sort_region_words((words + kSortBlock - 1) / kSortBlock)) * sizeof(uint32_t);
}

int sort_lists_persistent(rve_ctx* ctx, const SortedListUpdate* u, uint32_t n_lists, uint32_t id_limit) {
    if (n_lists == 0) return 0;
    DeviceGuard g(ctx->device);
    const uint32_t words = (id_limit + 31u) / 32u;
    const uint32_t blocks = (words + kSortBlock - 1) / kSortBlock;
    for (uint32_t first = 0; first < n_lists; first += kSortBatch) {
        SortBatch batch{};
        uint32_t n = 0, maxWork = 0;
        size_t deltaWords = 0;
        for (uint32_t i = first; i < n_lists && i < first + kSortBatch; ++i)
            if (!u[i].full) deltaWords += size_t(u[i].nSet) + u[i].nClear;
        // the deltas of the batch travel in one staged copy
        std::vector<uint32_t> host(deltaWords);
        int rc = ensure_scratch(ctx, &ctx->d_sort_delta, &ctx->d_sort_delta_cap, std::max<size_t>(deltaWords, 1) * sizeof(uint32_t));
        if (rc) return rc;
        size_t at = 0;
        for (uint32_t i = first; i < n_lists && i < first + kSortBatch; ++i) {
            const SortedListUpdate& x = u[i];
            if (x.count == 0) continue;
            BufferRec* bo = get_buffer(ctx, x.out, "sorted entity ids");
            BufferRec* br = get_buffer(ctx, x.bitmap, "entity id bitmap");
            if (!bo || !br) return RVE_ERR_BAD_HANDLE;
            RVE_REQUIRE(uint64_t(x.count) * 4 <= bo->bytes && sort_region_bytes(id_limit) <= br->bytes, "sorted id / bitmap buffers too small");
            SortList& L = batch.l[n];
            L = SortList{};
            L.out = static_cast<uint32_t*>(bo->ptr);
            L.region = static_cast<uint32_t*>(br->ptr);
            L.count = x.count;
            L.keep = 1u;
            if (x.full) {
                BufferRec* bi = get_buffer(ctx, x.ids, "entity ids");
                if (!bi) return RVE_ERR_BAD_HANDLE;
                RVE_REQUIRE(uint64_t(x.count) * 4 <= bi->bytes, "id buffer smaller than count entries");
                L.ids = static_cast<const uint32_t*>(bi->ptr);
                RVE_CUDA(cudaMemsetAsync(br->ptr, 0, size_t(blocks) * kSortBlock * sizeof(uint32_t), ctx->stream));
                maxWork = std::max(maxWork, x.count);
            } else {
                if (x.nSet) std::memcpy(host.data() + at, x.setIds, size_t(x.nSet) * 4);
                if (x.nClear) std::memcpy(host.data() + at + x.nSet, x.clearIds, size_t(x.nClear) * 4);
                L.delta = static_cast<const uint32_t*>(ctx->d_sort_delta) + at;
                L.nSet = x.nSet;
                L.nClear = x.nClear;
                at += size_t(x.nSet) + x.nClear;
                maxWork = std::max(maxWork, x.nSet + x.nClear);
            }
            ++n;
        }
        if (!n) continue;
        if (at) {
            rc = staged_h2d(ctx, ctx->d_sort_delta, host.data(), at * sizeof(uint32_t));
            if (rc) return rc;
        }
        const uint32_t setBlocks = std::max<uint32_t>(1u, std::min<uint32_t>((maxWork + 255) / 256, uint32_t(ctx->sm_count) * 8u / n + 1u));
        ids_set_bits_batch_kernel<<<dim3(setBlocks, n), 256, 0, ctx->stream>>>(batch, id_limit, blocks);
        ids_expand_lookback_kernel<<<dim3((blocks + kLbSub - 1) / kLbSub, n), kSortBlock, 0, ctx->stream>>>(batch, words, blocks);
        ctx->launches += 2;
        RVE_CUDA(cudaGetLastError());
    }
    return 0;
}

}  // namespace rve

using namespace rve;

extern "C" int rve_sort_entity_ids_batch(rve_ctx* ctx, uint32_t n_lists, const rve_buffer* entity_ids, const uint32_t* counts, uint32_t id_limit,
                                         const rve_buffer* sorted_out) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    if (n_lists == 0) return 0;
    if (!entity_ids || !counts || !sorted_out) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_PROFILE_FN_N("Sort entity ids");
    DeviceGuard g(ctx->device);
    RVE_REQUIRE(id_limit >= 1, "id_limit is the size of the entity id space (ids are < id_limit)");
    for (uint32_t first = 0; first < n_lists; first += kSortBatch) {
        SortBatch batch{};
        uint32_t n = 0, maxCount = 0;
        for (uint32_t i = first; i < n_lists && i < first + kSortBatch; ++i) {
            if (counts[i] == 0) continue;
            BufferRec* bi = get_buffer(ctx, entity_ids[i], "entity ids");
            BufferRec* bo = get_buffer(ctx, sorted_out[i], "sorted entity ids");
            if (!bi || !bo) return RVE_ERR_BAD_HANDLE;
            RVE_REQUIRE(uint64_t(counts[i]) * 4 <= bi->bytes && uint64_t(counts[i]) * 4 <= bo->bytes, "id buffers smaller than count entries");
            batch.l[n] = SortList{};
            batch.l[n].ids = static_cast<const uint32_t*>(bi->ptr);
            batch.l[n].out = static_cast<uint32_t*>(bo->ptr);
            batch.l[n].count = counts[i];
            maxCount = std::max(maxCount, counts[i]);
            ++n;
        }
        if (!n) continue;
        const int rc = sort_batch(ctx, batch, n, maxCount, id_limit);
        if (rc) return rc;
    }
    return 0;
}

extern "C" int rve_sort_entity_ids(rve_ctx* ctx, rve_buffer entity_ids, uint32_t count, uint32_t id_limit, rve_buffer sorted_out) {
    return rve_sort_entity_ids_batch(ctx, 1, &entity_ids, &count, id_limit, &sorted_out);
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
