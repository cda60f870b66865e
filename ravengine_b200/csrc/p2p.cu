// p2p.cu — the exchange steps of instance-range sharding as kernels over NVLink peer memory, no NCCL call
// and no host round trip:
//   count_exchange_kernel: gather this rank's per-slot visible counts out of the indirect buffers, store them into every
//     rank's mailbox (peer stores), signal, wait for every peer's signal, exclusive scan over ranks -> this rank's output
//     offsets and the global instanceCount. One launch per frame on a side stream.
//   gather_ids_kernel: (only when a single consumer needs the ids) every rank writes its compacted slices straight into
//     the consumer's GLOBAL culling buffers at cullingBufferOffset + numObjects * lod + offset[rank][slot]
//     (shaders/defaultcull.csh:305-312) over the mapped peer memory, then signals; consumers wait for every rank.
//
// Mailbox of a rank (u32 words; geometry fixed at creation, independent of the per-call slot count): 64 header words =
// count flags[set][sender] (0..15), slot counts announced by the sender [set][sender] (16..31), gather-done
// flags[set][sender] (32..47), then counts[set][sender][slot]. Epochs start at 1 and grow by one per exchange;
// set = epoch & 1. Two sets are enough: a rank can only pass the wait of epoch e+1 after every peer finished epoch e
// (its e+1 signal is issued by a kernel ordered after its epoch-e kernel), so nobody overwrites set (e & 1) with epoch
// e+2 while a peer still reads epoch e from it.
#include <cstring>

#include "internal.h"

namespace rve {

constexpr uint32_t kMaxWorld = 8;      // one NVSwitch domain
constexpr uint32_t kMaxGather = 64;    // indirect buffers per call (one per material)
constexpr uint32_t kMbHeader = 64;     // header words
constexpr uint32_t kMbCountFlag = 0, kMbSlots = 16, kMbGatherFlag = 32;
constexpr uint32_t kSentinel = 0xFFFFFFFFu;

struct ExchangeParams {
    const RveIndirectIndexedCommand* cmds[kMaxGather];
    uint32_t first[kMaxGather], n[kMaxGather], outStart[kMaxGather];
    uint32_t nBuffers, slots;
    const uint32_t* dense;  // when nBuffers == 0: the counts already gathered into a dense u32[slots]
    uint32_t* mailbox[kMaxWorld];  // [rank] = this rank's own
    uint32_t rank, world, set, epoch;
    uint32_t* offsets;  // u32[world][slots]
    uint32_t* totals;   // u32[slots]
    uint32_t* status;   // bit 0: a peer never signalled (timeout); bit 1: a peer announced a different slot count
    unsigned long long timeoutCycles;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_volatile(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// waits until *f reaches `epoch` (wrap-safe); false on timeout
__device__ __forceinline__ bool wait_flag(const uint32_t* f, uint32_t epoch, unsigned long long timeoutCycles) {
    const long long t0 = clock64();
    while (int32_t(ld_acquire_sys(f) - epoch) < 0) {
        if (uint64_t(clock64() - t0) > timeoutCycles) return false;
        __nanosleep(200);
    }
    return true;
}

__global__ void __launch_bounds__(256) count_exchange_kernel(const ExchangeParams p) {
    __shared__ uint32_t s_bad;
    const uint32_t slots = p.slots, world = p.world;
    if (threadIdx.x == 0) s_bad = 0;
    // phase 1: gather + peer stores
    for (uint32_t t = threadIdx.x; t < slots; t += blockDim.x) {
        uint32_t c;
        if (p.nBuffers == 0) {
            c = p.dense[t];
        } else {
            uint32_t b = 0;
            while (b + 1 < p.nBuffers && t >= p.outStart[b + 1]) ++b;
            c = p.cmds[b][p.first[b] + (t - p.outStart[b])].instanceCount;
        }
        for (uint32_t q = 0; q < world; ++q) p.mailbox[q][kMbHeader + (p.set * world + p.rank) * slots + t] = c;
    }
    if (threadIdx.x < world) p.mailbox[threadIdx.x][kMbSlots + p.set * kMaxWorld + p.rank] = slots;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < world) st_release_sys(p.mailbox[threadIdx.x] + kMbCountFlag + p.set * kMaxWorld + p.rank, p.epoch);
    // phase 2: wait for every sender's signal of this epoch
    uint32_t* mine = p.mailbox[p.rank];
    if (threadIdx.x < world) {
        if (!wait_flag(mine + kMbCountFlag + p.set * kMaxWorld + threadIdx.x, p.epoch, p.timeoutCycles)) atomicOr(&s_bad, 1u);
        else if (ld_volatile(mine + kMbSlots + p.set * kMaxWorld + threadIdx.x) != slots) atomicOr(&s_bad, 2u);
    }
    __syncthreads();
    const uint32_t bad = s_bad;
    if (bad && threadIdx.x == 0) atomicOr(p.status, bad);
    // exclusive scan over ranks (sentinels instead of offsets derived from stale counts when the exchange failed)
    for (uint32_t t = threadIdx.x; t < slots; t += blockDim.x) {
        uint32_t run = 0;
        for (uint32_t q = 0; q < world; ++q) {
            p.offsets[q * slots + t] = bad ? kSentinel : run;
            run += ld_volatile(mine + kMbHeader + (p.set * world + q) * slots + t);
        }
        p.totals[t] = bad ? kSentinel : run;
    }
}

constexpr uint32_t kGatherMaxSlots = 64, kGatherMaxBuffers = 16;
struct GatherParams {
    const uint32_t* src[kGatherMaxBuffers];            // this rank's culling buffers
    uint32_t* dst[kMaxWorld][kGatherMaxBuffers];       // consumer q's global culling buffers (nullptr: q is no consumer)
    uint32_t srcStart[kGatherMaxSlots], dstStart[kGatherMaxSlots];  // first id of the slot's local / global slice
    uint8_t bufOf[kGatherMaxSlots];
    uint32_t slots, nBuffers, rank, world, set, epoch, idBase, consumers;  // consumers: bit q set = rank q receives
    const uint32_t* offsets;  // u32[world][slots] of this epoch's count exchange
    const uint32_t* totals;   // u32[slots]
    uint32_t* mailbox[kMaxWorld];
    uint32_t* ticket;   // CTAs done (self-resetting)
    uint32_t* status;
    unsigned long long timeoutCycles;
};

// Every CTA walks all slots and copies a grid-strided share of each local slice into every consumer's global slice.
// The last CTA to finish publishes this rank's "done" flag in the consumers' mailboxes; on a consumer it then waits for
// every rank's flag, so that work enqueued behind this kernel on the consumer's stream sees the complete buffers.
__global__ void __launch_bounds__(256) gather_ids_kernel(const GatherParams p) {
    __shared__ uint32_t s_last;
    const uint32_t slots = p.slots, world = p.world;
    const bool failed = ld_volatile(p.status) != 0;  // the count exchange of this epoch did not complete: nothing to place
    if (!failed) {
        for (uint32_t s = 0; s < slots; ++s) {
            const uint32_t off = p.offsets[p.rank * slots + s];
            const uint32_t next = p.rank + 1 < world ? p.offsets[(p.rank + 1) * slots + s] : p.totals[s];
            const uint32_t n = next - off;
            const uint32_t b = p.bufOf[s];
            const uint32_t* src = p.src[b] + p.srcStart[s];
            for (uint32_t q = 0; q < world; ++q) {
                if (!((p.consumers >> q) & 1u)) continue;
                uint32_t* dst = p.dst[q][b] + p.dstStart[s] + off;
                // four independent loads in flight per thread (the loop is latency-bound: NVLink stores are fire-and-forget)
                const uint32_t stride = gridDim.x * blockDim.x;
                uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
                for (; i + 3 * stride < n; i += 4 * stride) {
                    const uint32_t a = __ldg(src + i), b = __ldg(src + i + stride), c = __ldg(src + i + 2 * stride), d = __ldg(src + i + 3 * stride);
                    dst[i] = a + p.idBase;
                    dst[i + stride] = b + p.idBase;
                    dst[i + 2 * stride] = c + p.idBase;
                    dst[i + 3 * stride] = d + p.idBase;
                }
                for (; i < n; i += stride) dst[i] = __ldg(src + i) + p.idBase;
            }
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(p.ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (!s_last) return;
    if (threadIdx.x == 0) *p.ticket = 0;
    __threadfence_system();
    if (threadIdx.x < world && ((p.consumers >> threadIdx.x) & 1u))
        st_release_sys(p.mailbox[threadIdx.x] + kMbGatherFlag + p.set * kMaxWorld + p.rank, p.epoch);
    if ((p.consumers >> p.rank) & 1u) {
        if (threadIdx.x < world && !wait_flag(p.mailbox[p.rank] + kMbGatherFlag + p.set * kMaxWorld + threadIdx.x, p.epoch, p.timeoutCycles))
            atomicOr(p.status, 1u);
    }
}

// CUDA loads kernels lazily, and loading one may synchronise the device. A first launch of these kernels while a peer
// rank's kernel already spins on this rank's signal (two ranks on one device, or a consumer waiting for this one) would
// stall until the peer's timeout: load them when the context is created.
int p2p_preload() {
    cudaFuncAttributes a{};
    RVE_CUDA(cudaFuncGetAttributes(&a, count_exchange_kernel));
    RVE_CUDA(cudaFuncGetAttributes(&a, gather_ids_kernel));
    return 0;
}

}  // namespace rve

using namespace rve;

extern "C" {

int rve_ipc_export(rve_ctx* ctx, rve_buffer buf, unsigned char out_handle[64]) {
    if (!ctx || !out_handle) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t size");
    DeviceGuard g(ctx->device);
    BufferRec* b = get_buffer(ctx, buf, "ipc export");
    if (!b) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE(b->owned, "only buffers created by rve_buffer_create (whole cudaMalloc allocations) can be exported");
    cudaIpcMemHandle_t h;
    RVE_CUDA(cudaIpcGetMemHandle(&h, b->ptr));
    memcpy(out_handle, &h, 64);
    return 0;
}

int rve_ipc_import(rve_ctx* ctx, const unsigned char handle[64], size_t bytes, rve_buffer* out) {
    if (!ctx || !handle || !out) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    void* p = nullptr;
    RVE_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    const int rc = rve_buffer_wrap(ctx, p, bytes, out);  // not owned; unmapped by rve_buffer_destroy
    if (rc) {
        cudaIpcCloseMemHandle(p);
        return rc;
    }
    ctx->buffers[*out].ipc_imported = true;
    return 0;
}

size_t rve_count_mailbox_bytes(uint32_t world, uint32_t total_slots) { return (size_t(kMbHeader) + size_t(2) * world * total_slots) * sizeof(uint32_t); }

int rve_exchange_counts_p2p(rve_ctx* ctx, const rve_buffer* indirect, const uint32_t* first_slot, const uint32_t* num_slots,
                            uint32_t n_buffers, rve_buffer dense_counts, uint32_t dense_slots, const rve_buffer* mailboxes, uint32_t rank,
                            uint32_t world, uint32_t epoch, rve_buffer offsets_out, rve_buffer totals_out, rve_buffer status) {
    if (!ctx || !mailboxes) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    RVE_REQUIRE(world >= 1 && world <= kMaxWorld && rank < world, "world size must be 1..8 (one NVSwitch domain)");
    RVE_REQUIRE(n_buffers <= kMaxGather, "at most 64 indirect buffers per call");
    RVE_REQUIRE(n_buffers == 0 || (indirect && first_slot && num_slots), "null indirect buffer list");
    RVE_REQUIRE(epoch != 0, "epochs start at 1");
    ExchangeParams p{};
    uint32_t total = 0;
    if (n_buffers == 0) {
        BufferRec* d = get_buffer(ctx, dense_counts, "dense counts");
        if (!d) return RVE_ERR_BAD_HANDLE;
        RVE_REQUIRE(size_t(dense_slots) * 4 <= d->bytes, "dense count buffer too small");
        p.dense = static_cast<const uint32_t*>(d->ptr);
        total = dense_slots;
    }
    for (uint32_t i = 0; i < n_buffers; ++i) {
        BufferRec* b = get_buffer(ctx, indirect[i], "indirect");
        if (!b) return RVE_ERR_BAD_HANDLE;
        RVE_REQUIRE((uint64_t(first_slot[i]) + num_slots[i]) * 20 <= b->bytes, "indirect buffer too small");
        p.cmds[i] = static_cast<const RveIndirectIndexedCommand*>(b->ptr);
        p.first[i] = first_slot[i];
        p.n[i] = num_slots[i];
        p.outStart[i] = total;
        total += num_slots[i];
    }
    if (!total) return 0;
    p.nBuffers = n_buffers;
    p.slots = total;
    for (uint32_t q = 0; q < world; ++q) {
        BufferRec* m = get_buffer(ctx, mailboxes[q], "mailbox");
        if (!m) return RVE_ERR_BAD_HANDLE;
        RVE_REQUIRE(m->bytes >= rve_count_mailbox_bytes(world, total), "mailbox smaller than rve_count_mailbox_bytes(world, slots)");
        p.mailbox[q] = static_cast<uint32_t*>(m->ptr);
    }
    BufferRec* bo = get_buffer(ctx, offsets_out, "offsets");
    BufferRec* bt = get_buffer(ctx, totals_out, "totals");
    BufferRec* bs = get_buffer(ctx, status, "status");
    if (!bo || !bt || !bs) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE(bo->bytes >= size_t(world) * total * 4 && bt->bytes >= size_t(total) * 4 && bs->bytes >= 4, "output buffers too small");
    p.rank = rank;
    p.world = world;
    p.epoch = epoch;
    p.set = epoch & 1u;
    p.offsets = static_cast<uint32_t*>(bo->ptr);
    p.totals = static_cast<uint32_t*>(bt->ptr);
    p.status = static_cast<uint32_t*>(bs->ptr);
    p.timeoutCycles = 4000000000ull;  // ~2 s at 2 GHz: a peer that never arrives is reported, the GPU is not left spinning
    count_exchange_kernel<<<1, 256, 0, ctx->stream>>>(p);
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}


int rve_gather_ids_p2p(rve_ctx* ctx, const rve_buffer* src_culling, uint32_t n_buffers, const uint32_t* slot_buffer, const uint32_t* src_start,
                       const uint32_t* dst_start, uint32_t slots, const rve_buffer* dst_culling, uint32_t consumers, uint32_t id_base,
                       rve_buffer offsets, rve_buffer totals, const rve_buffer* mailboxes, uint32_t rank, uint32_t world, uint32_t epoch,
                       rve_buffer status) {
    if (!ctx || !src_culling || !slot_buffer || !src_start || !dst_start || !dst_culling || !mailboxes)
        return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    RVE_REQUIRE(world >= 1 && world <= kMaxWorld && rank < world, "world size must be 1..8 (one NVSwitch domain)");
    RVE_REQUIRE(n_buffers >= 1 && n_buffers <= kGatherMaxBuffers, "1..16 culling buffers per call");
    RVE_REQUIRE(slots >= 1 && slots <= kGatherMaxSlots, "1..64 slots per call");
    RVE_REQUIRE(epoch != 0, "epochs start at 1");
    RVE_REQUIRE(consumers != 0 && (consumers >> world) == 0, "consumer mask names ranks outside the world");
    GatherParams p{};
    for (uint32_t b = 0; b < n_buffers; ++b) {
        BufferRec* sb = get_buffer(ctx, src_culling[b], "local culling buffer");
        if (!sb) return RVE_ERR_BAD_HANDLE;
        p.src[b] = static_cast<const uint32_t*>(sb->ptr);
        for (uint32_t q = 0; q < world; ++q) {
            if (!((consumers >> q) & 1u)) continue;
            BufferRec* db = get_buffer(ctx, dst_culling[q * n_buffers + b], "consumer culling buffer");
            if (!db) return RVE_ERR_BAD_HANDLE;
            p.dst[q][b] = static_cast<uint32_t*>(db->ptr);
        }
    }
    for (uint32_t s = 0; s < slots; ++s) {
        RVE_REQUIRE(slot_buffer[s] < n_buffers, "slot names a culling buffer outside the list");
        p.bufOf[s] = uint8_t(slot_buffer[s]);
        p.srcStart[s] = src_start[s];
        p.dstStart[s] = dst_start[s];
    }
    BufferRec* bo = get_buffer(ctx, offsets, "offsets");
    BufferRec* bt = get_buffer(ctx, totals, "totals");
    BufferRec* bs = get_buffer(ctx, status, "status");
    if (!bo || !bt || !bs) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE(bo->bytes >= size_t(world) * slots * 4 && bt->bytes >= size_t(slots) * 4 && bs->bytes >= 4, "offset / total / status buffers too small");
    for (uint32_t q = 0; q < world; ++q) {
        BufferRec* m = get_buffer(ctx, mailboxes[q], "mailbox");
        if (!m) return RVE_ERR_BAD_HANDLE;
        RVE_REQUIRE(m->bytes >= kMbHeader * sizeof(uint32_t), "mailbox smaller than its header");
        p.mailbox[q] = static_cast<uint32_t*>(m->ptr);
    }
    RVE_REQUIRE(ctx->d_gather_ticket != nullptr, "context has no gather ticket");  // allocated by rve_ctx_create: a cudaMalloc here
                                                                                     // would synchronise the device under a peer's spin
    p.slots = slots;
    p.nBuffers = n_buffers;
    p.rank = rank;
    p.world = world;
    p.set = epoch & 1u;
    p.epoch = epoch;
    p.idBase = id_base;
    p.consumers = consumers;
    p.offsets = static_cast<const uint32_t*>(bo->ptr);
    p.totals = static_cast<const uint32_t*>(bt->ptr);
    p.ticket = ctx->d_gather_ticket;
    p.status = static_cast<uint32_t*>(bs->ptr);
    p.timeoutCycles = 4000000000ull;
    // enough CTAs to keep NVLink stores in flight; every rank's kernel must be able to run while a consumer waits
    const uint32_t blocks = uint32_t(ctx->sm_count) * 2u;
    gather_ids_kernel<<<blocks, 256, 0, ctx->stream>>>(p);
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
