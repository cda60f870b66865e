// p2p.cu — the one exchange step of instance-range sharding as ONE kernel over NVLink peer memory:
// gather this rank's per-slot visible counts out of the indirect buffers, store them into every rank's mailbox
// (peer stores), signal, wait for every peer's signal, exclusive scan over ranks -> this rank's output offsets and
// the global instanceCount. No NCCL call, no host round trip: one launch per frame on a side stream.
//
// Mailbox of a rank (u32 words): counts[set][sender][slot] for set in {0,1}, then flags[set][sender]. Epochs start
// at 1 and grow by one per exchange; set = epoch & 1. Two sets are enough: a rank can only pass the wait of epoch
// e+1 after every peer finished epoch e (its e+1 signal is issued by a kernel ordered after its epoch-e kernel), so
// nobody overwrites set (e & 1) with epoch e+2 while a peer still reads epoch e from it.
#include <cstring>

#include "internal.h"

namespace rve {

constexpr uint32_t kMaxWorld = 8;      // one NVSwitch domain
constexpr uint32_t kMaxGather = 64;    // indirect buffers per call (one per material)

struct ExchangeParams {
    const RveIndirectIndexedCommand* cmds[kMaxGather];
    uint32_t first[kMaxGather], n[kMaxGather], outStart[kMaxGather];
    uint32_t nBuffers, slots;
    const uint32_t* dense;  // when nBuffers == 0: the counts already gathered into a dense u32[slots]
    uint32_t* mailbox[kMaxWorld];  // [rank] = this rank's own
    uint32_t rank, world, set, epoch;
    uint32_t* offsets;  // u32[world][slots]
    uint32_t* totals;   // u32[slots]
    uint32_t* status;   // set to 1 when a peer never signalled (timeout)
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

__global__ void __launch_bounds__(256) count_exchange_kernel(const ExchangeParams p) {
    const uint32_t slots = p.slots, world = p.world;
    const uint32_t flagBase = 2u * world * slots;
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
        for (uint32_t q = 0; q < world; ++q) p.mailbox[q][(p.set * world + p.rank) * slots + t] = c;
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < world) st_release_sys(p.mailbox[threadIdx.x] + flagBase + p.set * world + p.rank, p.epoch);
    // phase 2: wait for every sender's signal of this epoch
    uint32_t* mine = p.mailbox[p.rank];
    if (threadIdx.x < world) {
        const uint32_t* f = mine + flagBase + p.set * world + threadIdx.x;
        const long long t0 = clock64();
        while (int32_t(ld_acquire_sys(f) - p.epoch) < 0) {
            if (uint64_t(clock64() - t0) > p.timeoutCycles) {
                *p.status = 1u;
                break;
            }
            __nanosleep(200);
        }
    }
    __syncthreads();
    // exclusive scan over ranks
    for (uint32_t t = threadIdx.x; t < slots; t += blockDim.x) {
        uint32_t run = 0;
        for (uint32_t q = 0; q < world; ++q) {
            p.offsets[q * slots + t] = run;
            run += ld_volatile(mine + (p.set * world + q) * slots + t);
        }
        p.totals[t] = run;
    }
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
    const int rc = rve_buffer_wrap(ctx, p, bytes, out);  // not owned: stays mapped for the life of the process
    if (rc) cudaIpcCloseMemHandle(p);
    return rc;
}

size_t rve_count_mailbox_bytes(uint32_t world, uint32_t total_slots) { return (size_t(2) * world * total_slots + size_t(2) * world) * sizeof(uint32_t); }

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

}  // extern "C"
