// ctx.cu — context, buffer table ("bindless" handles) and error plumbing of the C-ABI.
// Replaces the RGL objects the render-prep path touches: IDevice/ICommandQueue
// (src/RenderEngine.cpp:304-307) and IBuffer (deps/RGL/include/RGL/Buffer.hpp:50-81).
#include <cctype>
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include <sched.h>
#include <sys/syscall.h>
#include <unistd.h>

#include "internal.h"

namespace rve {

static thread_local std::string g_last_error;

int set_error(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    const int code = (e == cudaErrorMemoryAllocation) ? RVE_ERR_OUT_OF_MEMORY : RVE_ERR_CUDA;
    return set_error(code, "CUDA error %d (%s) in %s", int(e), cudaGetErrorString(e), what);
}

BufferRec* get_buffer(rve_ctx* ctx, rve_buffer h, const char* role) {
    if (h >= ctx->buffers.size() || !ctx->buffers[h].live) {
        set_error(RVE_ERR_BAD_HANDLE, "invalid buffer handle %u for '%s'", h, role);
        return nullptr;
    }
    return &ctx->buffers[h];
}

PyramidRec* get_pyramid(rve_ctx* ctx, rve_pyramid h) {
    if (h >= ctx->pyramids.size() || !ctx->pyramids[h].live) {
        set_error(RVE_ERR_BAD_HANDLE, "invalid pyramid handle %u", h);
        return nullptr;
    }
    return &ctx->pyramids[h];
}

int ensure_scratch(rve_ctx* ctx, void** ptr, size_t* cap, size_t bytes) {
    if (bytes <= *cap) return 0;
    size_t ncap = *cap ? *cap : 4096;
    while (ncap < bytes) ncap *= 2;
    void* np = nullptr;
    RVE_CUDA(cudaMallocAsync(&np, ncap, ctx->stream));
    if (*ptr) RVE_CUDA(cudaFreeAsync(*ptr, ctx->stream));
    *ptr = np;
    *cap = ncap;
    return 0;
}

int staged_h2d(rve_ctx* ctx, void* dst, const void* src, size_t bytes) {
    if (!bytes) return 0;
    if (ctx->capture) {
        void* pin = nullptr;
        RVE_CUDA(cudaHostAlloc(&pin, bytes, cudaHostAllocDefault));
        memcpy(pin, src, bytes);
        ctx->capture->pinned.push_back(pin);
        src = pin;
    }
    RVE_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return 0;
}

int sync_table(rve_ctx* ctx) {
    if (!ctx->table_dirty) return 0;
    const size_t n = ctx->buffers.size();
    if (n > ctx->d_table_cap) {
        size_t ncap = ctx->d_table_cap ? ctx->d_table_cap : 256;
        while (ncap < n) ncap *= 2;
        void** np = nullptr;
        RVE_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&np), ncap * sizeof(void*), ctx->stream));
        if (ctx->d_table) RVE_CUDA(cudaFreeAsync(ctx->d_table, ctx->stream));
        ctx->d_table = np;
        ctx->d_table_cap = ncap;
    }
    if (n) {
        std::vector<void*> host(n);
        for (size_t i = 0; i < n; ++i) host[i] = ctx->buffers[i].live ? ctx->buffers[i].ptr : nullptr;
        // pageable source: the runtime stages it before returning, so `host` may die afterwards
        const int rc = staged_h2d(ctx, ctx->d_table, host.data(), n * sizeof(void*));
        if (rc) return rc;
    }
    ctx->table_dirty = false;
    return 0;
}

static int alloc_slot(rve_ctx* ctx, void* ptr, size_t bytes, bool owned, rve_buffer* out) {
    uint32_t slot;
    if (!ctx->free_slots.empty()) {
        slot = ctx->free_slots.back();
        ctx->free_slots.pop_back();
    } else {
        slot = uint32_t(ctx->buffers.size());
        ctx->buffers.emplace_back();
    }
    ctx->buffers[slot] = BufferRec{ptr, bytes, owned, true, false};
    ctx->table_dirty = true;
    *out = slot;
    return 0;
}

}  // namespace rve

using namespace rve;

namespace rve {
// one CTA per segment; the payload is read from the parameter (constant) bank
__global__ void upload_params_kernel(const __grid_constant__ UploadBlock blk) {
    pdl_trigger();  // a programmatic dependent (cubo_prepare_kernel) may become resident; it waits for this grid's completion
    const uint32_t s = blockIdx.x;
    const uint32_t* src = blk.data + blk.first[s];
    uint32_t* dst = blk.dst[s];
    for (uint32_t i = threadIdx.x; i < blk.count[s]; i += blockDim.x) dst[i] = src[i];
}

int make_upload_block(const RawSegment* segs, uint32_t n, UploadBlock* blk) {
    blk->n = 0;
    if (n == 0) return 0;
    if (n > kUploadMaxSegs) return RVE_ERR_UNSUPPORTED;
    uint32_t words = 0;
    for (uint32_t i = 0; i < n; ++i) {
        if ((segs[i].bytes & 3) || (reinterpret_cast<uintptr_t>(segs[i].dst) & 3) || words + segs[i].bytes / 4 > kUploadMaxWords)
            return RVE_ERR_UNSUPPORTED;
        blk->dst[i] = static_cast<uint32_t*>(segs[i].dst);
        blk->first[i] = words;
        blk->count[i] = uint32_t(segs[i].bytes / 4);
        if (segs[i].bytes) memcpy(blk->data + words, segs[i].src, segs[i].bytes);
        words += uint32_t(segs[i].bytes / 4);
    }
    if (words) blk->n = n;
    return 0;
}

int upload_block(rve_ctx* ctx, const UploadBlock& blk) {
    if (!blk.n) return 0;
    upload_params_kernel<<<blk.n, 128, 0, ctx->stream>>>(blk);
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

int upload_batch_raw(rve_ctx* ctx, const RawSegment* segs, uint32_t n) {
    UploadBlock blk{};
    const int rc = make_upload_block(segs, n, &blk);
    if (rc) return rc;
    return upload_block(ctx, blk);
}
}  // namespace rve

extern "C" {

const char* rve_last_error_string(void) { return g_last_error.c_str(); }
const char* rve_version(void) { return "rve_b200 0.1 (sm_100a)"; }

int rve_ctx_create(int device_ordinal, void* cuda_stream, rve_ctx** out_ctx) {
    if (!out_ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "out_ctx is null");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return set_error(RVE_ERR_NO_DEVICE, "no CUDA device available (%s); this library has no CPU fallback",
                         cudaGetErrorString(e));
    if (device_ordinal < 0 || device_ordinal >= count)
        return set_error(RVE_ERR_INVALID_ARGUMENT, "device ordinal %d out of range [0,%d)", device_ordinal, count);
    DeviceGuard g(device_ordinal);
    auto* ctx = new rve_ctx();
    ctx->device = device_ordinal;
    if (cuda_stream) {
        ctx->stream = static_cast<cudaStream_t>(cuda_stream);
    } else {
        int rc = check_cuda(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking), "cudaStreamCreate");
        if (rc) {
            delete ctx;
            return rc;
        }
        ctx->owns_stream = true;
    }
    cudaDeviceProp prop{};
    int rc = check_cuda(cudaGetDeviceProperties(&prop, device_ordinal), "cudaGetDeviceProperties");
    if (rc) {
        delete ctx;
        return rc;
    }
    ctx->sm_count = prop.multiProcessorCount;
    ctx->l2_bytes = size_t(prop.l2CacheSize);
    rc = check_cuda(cudaMalloc(reinterpret_cast<void**>(&ctx->d_gather_ticket), 2 * sizeof(uint32_t)), "cudaMalloc");
    if (rc == 0) rc = check_cuda(cudaMemset(ctx->d_gather_ticket, 0, 2 * sizeof(uint32_t)), "cudaMemset");
    if (rc == 0) rc = p2p_preload();
    if (rc == 0) rc = sync_preload();
    if (rc) {
        delete ctx;
        return rc;
    }
    *out_ctx = ctx;
    return 0;
}

int rve_ctx_destroy(rve_ctx* ctx) {
    if (!ctx) return 0;
    DeviceGuard g(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto& b : ctx->buffers)
        if (b.live && b.owned && b.ptr) cudaFree(b.ptr);
    if (ctx->d_table) cudaFree(ctx->d_table);
    if (ctx->d_cubo_start) cudaFree(ctx->d_cubo_start);
    if (ctx->d_cubo_res) cudaFree(ctx->d_cubo_res);
    if (ctx->d_cubo_lod) cudaFree(ctx->d_cubo_lod);
    if (ctx->d_view_index) cudaFree(ctx->d_view_index);
    if (ctx->d_view_stride) cudaFree(ctx->d_view_stride);
    if (ctx->d_views) cudaFree(ctx->d_views);
    if (ctx->d_lights) cudaFree(ctx->d_lights);
    if (ctx->d_slices) cudaFree(ctx->d_slices);
    if (ctx->d_sort) cudaFree(ctx->d_sort);
    if (ctx->d_sort_delta) cudaFree(ctx->d_sort_delta);
    if (ctx->sync_stage_host) cudaFreeHost(ctx->sync_stage_host);
    if (ctx->sync_stage_dev) cudaFree(ctx->sync_stage_dev);
    if (ctx->sync_stage_free) cudaEventDestroy(ctx->sync_stage_free);
    if (ctx->d_gather_ticket) cudaFree(ctx->d_gather_ticket);
    for (auto& b : ctx->buffers)
        if (b.live && b.ipc_imported && b.ptr) cudaIpcCloseMemHandle(b.ptr);
    if (ctx->owns_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return 0;
}

int rve_ctx_synchronize(rve_ctx* ctx) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    DeviceGuard g(ctx->device);
    RVE_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int rve_ctx_set_profiling(rve_ctx* ctx, int enabled) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    ctx->profiling = enabled != 0;
    return 0;
}

int rve_ctx_profile_read(rve_ctx* ctx, int kernel, double* total_ms, uint32_t* launches) {
    if (!ctx || kernel < 0 || kernel > 3) return set_error(RVE_ERR_INVALID_ARGUMENT, "bad profile query");
    DeviceGuard g(ctx->device);
    RVE_CUDA(cudaStreamSynchronize(ctx->stream));
    double sum = 0;
    for (auto& pr : ctx->prof[kernel]) {
        float ms = 0;
        RVE_CUDA(cudaEventElapsedTime(&ms, pr.a, pr.b));
        sum += ms;
        cudaEventDestroy(pr.a);
        cudaEventDestroy(pr.b);
    }
    if (total_ms) *total_ms = sum;
    if (launches) *launches = uint32_t(ctx->prof[kernel].size());
    ctx->prof[kernel].clear();
    return 0;
}

// Host-side placement for the upload path: the pinned host arrays a rank uploads from every frame should live on the
// NUMA node its GPU hangs off, and the thread that fills them should run there. Reads the GPU's PCI address from the
// runtime and the node / CPU list from sysfs; no libnuma, no NVML.
int rve_ctx_bind_host_to_device(rve_ctx* ctx, uint32_t flags, rve_host_affinity* out) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    rve_host_affinity info{};
    info.numa_node = -1;
    char bus[32] = {0};
    RVE_CUDA(cudaDeviceGetPCIBusId(bus, sizeof(bus), ctx->device));
    for (char* c = bus; *c; ++c) *c = char(tolower(*c));
    snprintf(info.pci_bus_id, sizeof(info.pci_bus_id), "%s", bus);
    char path[128];
    snprintf(path, sizeof(path), "/sys/bus/pci/devices/%s/numa_node", bus);
    if (FILE* f = fopen(path, "r")) {
        if (fscanf(f, "%d", &info.numa_node) != 1) info.numa_node = -1;
        fclose(f);
    }
    // CPUs local to the device, intersected with what this thread may use (cgroup / launcher restrictions)
    cpu_set_t allowed, want;
    CPU_ZERO(&allowed);
    CPU_ZERO(&want);
    sched_getaffinity(0, sizeof(allowed), &allowed);
    info.cpus_allowed = uint32_t(CPU_COUNT(&allowed));
    snprintf(path, sizeof(path), "/sys/bus/pci/devices/%s/local_cpulist", bus);
    if (FILE* f = fopen(path, "r")) {
        char line[4096] = {0};
        if (fgets(line, sizeof(line), f)) {
            for (char* tok = strtok(line, ",\n"); tok; tok = strtok(nullptr, ",\n")) {
                int a = 0, b = 0;
                const int k = sscanf(tok, "%d-%d", &a, &b);
                if (k == 1) b = a;
                if (k >= 1)
                    for (int c = a; c <= b && c < CPU_SETSIZE; ++c)
                        if (CPU_ISSET(c, &allowed)) CPU_SET(c, &want);
            }
        }
        fclose(f);
    }
    info.cpus_local = uint32_t(CPU_COUNT(&want));
    if ((flags & RVE_BIND_CPUS) && info.cpus_local > 0 && info.cpus_local < info.cpus_allowed)
        info.cpus_bound = sched_setaffinity(0, sizeof(want), &want) == 0 ? 1u : 0u;
    if ((flags & RVE_BIND_MEMORY) && info.numa_node >= 0 && info.numa_node < 1024) {
        // set_mempolicy(MPOL_PREFERRED, {node}): pages this thread touches first - the pinned allocations that follow -
        // come from the GPU's node when it has room, from elsewhere otherwise (never fails an allocation)
        unsigned long mask[16] = {0};
        mask[info.numa_node / (8 * sizeof(unsigned long))] |= 1ul << (info.numa_node % (8 * sizeof(unsigned long)));
        constexpr int kMpolPreferred = 1;
        info.memory_bound = syscall(SYS_set_mempolicy, kMpolPreferred, mask, sizeof(mask) * 8) == 0 ? 1u : 0u;
    }
    if (out) *out = info;
    return 0;
}

uint32_t rve_ctx_bad_cubo_handles(rve_ctx* ctx) {
    if (!ctx) return 0;
    DeviceGuard g(ctx->device);
    uint32_t v = 0;
    if (cudaMemcpyAsync(&v, ctx->d_gather_ticket + 1, sizeof(v), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) return 0;
    cudaStreamSynchronize(ctx->stream);
    return v;
}

// ---- one frame = one submission: the reference records its whole frame into one command buffer and commits it once
// (src/App_priv.hpp:475-483). Here: capture the rve_* / rvs_* enqueues of a frame into a CUDA graph, replay it with one launch.
int rve_ctx_capture_begin(rve_ctx* ctx) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    RVE_REQUIRE(!ctx->capture, "a capture is already in progress on this context");
    DeviceGuard g(ctx->device);
    auto* gr = new rve_graph();
    gr->device = ctx->device;
    // relaxed mode: the library may allocate (pinned staging, scratch growth) while the stream records
    const int rc = check_cuda(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed), "cudaStreamBeginCapture");
    if (rc) {
        delete gr;
        return rc;
    }
    ctx->capture = gr;
    return 0;
}

int rve_ctx_capture_end(rve_ctx* ctx, rve_graph** out) {
    if (!ctx || !out) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_REQUIRE(ctx->capture, "no capture in progress");
    DeviceGuard g(ctx->device);
    rve_graph* gr = ctx->capture;
    ctx->capture = nullptr;
    int rc = check_cuda(cudaStreamEndCapture(ctx->stream, &gr->graph), "cudaStreamEndCapture");
    if (rc == 0) rc = check_cuda(cudaGraphInstantiate(&gr->exec, gr->graph, 0), "cudaGraphInstantiate");
    if (rc) {
        rve_graph_destroy(gr);
        return rc;
    }
    *out = gr;
    return 0;
}

int rve_graph_launch(rve_ctx* ctx, rve_graph* graph) {
    if (!ctx || !graph || !graph->exec) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_REQUIRE(ctx->device == graph->device, "graph captured on another device");
    DeviceGuard g(ctx->device);
    RVE_CUDA(cudaGraphLaunch(graph->exec, ctx->stream));
    return 0;
}

int rve_graph_destroy(rve_graph* graph) {
    if (!graph) return 0;
    DeviceGuard g(graph->device);
    if (graph->exec) cudaGraphExecDestroy(graph->exec);
    if (graph->graph) cudaGraphDestroy(graph->graph);
    for (void* p : graph->pinned) cudaFreeHost(p);
    delete graph;
    return 0;
}

void* rve_ctx_stream(rve_ctx* ctx) { return ctx ? ctx->stream : nullptr; }
int rve_ctx_sm_count(rve_ctx* ctx) { return ctx ? ctx->sm_count : 0; }
uint64_t rve_ctx_launch_count(rve_ctx* ctx) { return ctx ? ctx->launches : 0; }

int rve_buffer_create(rve_ctx* ctx, size_t bytes, rve_buffer* out) {
    if (!ctx || !out) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    void* p = nullptr;
    const size_t alloc = bytes ? bytes : 16;
    RVE_CUDA(cudaMalloc(&p, alloc));
    int rc = check_cuda(cudaMemsetAsync(p, 0, alloc, ctx->stream), "cudaMemsetAsync");
    if (rc) {
        cudaFree(p);
        return rc;
    }
    std::lock_guard<std::mutex> lk(ctx->mu);
    return alloc_slot(ctx, p, bytes, true, out);
}

int rve_buffer_wrap(rve_ctx* ctx, void* device_ptr, size_t bytes, rve_buffer* out) {
    if (!ctx || !out) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    if (!device_ptr && bytes) return set_error(RVE_ERR_INVALID_ARGUMENT, "null device pointer with non-zero size");
    std::lock_guard<std::mutex> lk(ctx->mu);
    return alloc_slot(ctx, device_ptr, bytes, false, out);
}

int rve_buffer_destroy(rve_ctx* ctx, rve_buffer buf) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    DeviceGuard g(ctx->device);
    std::lock_guard<std::mutex> lk(ctx->mu);
    BufferRec* b = get_buffer(ctx, buf, "destroy");
    if (!b) return RVE_ERR_BAD_HANDLE;
    // stream-ordered release: work already enqueued may still read the memory
    if (b->owned && b->ptr) RVE_CUDA(cudaFreeAsync(b->ptr, ctx->stream));
    if (b->ipc_imported && b->ptr) {  // a peer's allocation: work enqueued so far may still touch it
        RVE_CUDA(cudaStreamSynchronize(ctx->stream));
        RVE_CUDA(cudaIpcCloseMemHandle(b->ptr));
    }
    *b = BufferRec{};
    ctx->free_slots.push_back(buf);
    ctx->table_dirty = true;
    return 0;
}

int rve_buffer_upload(rve_ctx* ctx, rve_buffer buf, size_t dst_offset, const void* host_src, size_t bytes) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    DeviceGuard g(ctx->device);
    BufferRec* b = get_buffer(ctx, buf, "upload");
    if (!b) return RVE_ERR_BAD_HANDLE;
    if (dst_offset + bytes > b->bytes) return set_error(RVE_ERR_INVALID_ARGUMENT, "upload out of range");
    if (!bytes) return 0;
    const int rc = staged_h2d(ctx, static_cast<char*>(b->ptr) + dst_offset, host_src, bytes);
    if (rc) return rc;
    pyramid_note_buffer_write(ctx, buf);
    return 0;
}

int rve_buffer_upload_batch(rve_ctx* ctx, const rve_upload_segment* segs, uint32_t n) {
    if (!ctx || (!segs && n)) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    // Small tables (the per-pass CUBO and indirect-init tables) travel as KERNEL PARAMETERS: one launch writes every
    // segment, instead of one staged host->device copy (a DMA operation of ~2 us on the stream) per segment.
    std::vector<RawSegment> raw(n);
    for (uint32_t i = 0; i < n; ++i) {
        BufferRec* b = get_buffer(ctx, segs[i].dst, "upload batch");
        if (!b) return RVE_ERR_BAD_HANDLE;
        if (segs[i].dst_offset + segs[i].bytes > b->bytes) return set_error(RVE_ERR_INVALID_ARGUMENT, "upload out of range");
        raw[i] = RawSegment{static_cast<char*>(b->ptr) + segs[i].dst_offset, segs[i].src, segs[i].bytes};
    }
    const int rc = upload_batch_raw(ctx, raw.data(), n);
    if (rc == RVE_ERR_UNSUPPORTED) {  // too large or unaligned for one parameter block: plain copies
        for (uint32_t i = 0; i < n; ++i) {
            const int rc2 = rve_buffer_upload(ctx, segs[i].dst, segs[i].dst_offset, segs[i].src, segs[i].bytes);
            if (rc2) return rc2;
        }
        return 0;
    }
    if (rc) return rc;
    for (uint32_t i = 0; i < n; ++i) pyramid_note_buffer_write(ctx, segs[i].dst);
    return 0;
}

int rve_buffer_download(rve_ctx* ctx, rve_buffer buf, size_t src_offset, void* host_dst, size_t bytes) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    DeviceGuard g(ctx->device);
    BufferRec* b = get_buffer(ctx, buf, "download");
    if (!b) return RVE_ERR_BAD_HANDLE;
    if (src_offset + bytes > b->bytes) return set_error(RVE_ERR_INVALID_ARGUMENT, "download out of range");
    if (!bytes) return 0;
    const int prc = pyramid_before_buffer_read(ctx, buf);
    if (prc) return prc;
    RVE_CUDA(cudaMemcpyAsync(host_dst, static_cast<const char*>(b->ptr) + src_offset, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    RVE_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int rve_buffer_copy(rve_ctx* ctx, rve_buffer src, size_t src_offset, rve_buffer dst, size_t dst_offset, size_t bytes) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    DeviceGuard g(ctx->device);
    BufferRec* s = get_buffer(ctx, src, "copy src");
    BufferRec* d = get_buffer(ctx, dst, "copy dst");
    if (!s || !d) return RVE_ERR_BAD_HANDLE;
    if (src_offset + bytes > s->bytes || dst_offset + bytes > d->bytes)
        return set_error(RVE_ERR_INVALID_ARGUMENT, "copy out of range");
    if (!bytes) return 0;
    const int prc = pyramid_before_buffer_read(ctx, src);
    if (prc) return prc;
    RVE_CUDA(cudaMemcpyAsync(static_cast<char*>(d->ptr) + dst_offset, static_cast<const char*>(s->ptr) + src_offset, bytes,
                             cudaMemcpyDeviceToDevice, ctx->stream));
    pyramid_note_buffer_write(ctx, dst);
    return 0;
}

int rve_buffer_fill_zero(rve_ctx* ctx, rve_buffer buf, size_t offset, size_t bytes) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    DeviceGuard g(ctx->device);
    BufferRec* b = get_buffer(ctx, buf, "fill");
    if (!b) return RVE_ERR_BAD_HANDLE;
    if (offset + bytes > b->bytes) return set_error(RVE_ERR_INVALID_ARGUMENT, "fill out of range");
    if (!bytes) return 0;
    RVE_CUDA(cudaMemsetAsync(static_cast<char*>(b->ptr) + offset, 0, bytes, ctx->stream));
    pyramid_note_buffer_write(ctx, buf);
    return 0;
}

void* rve_buffer_device_ptr(rve_ctx* ctx, rve_buffer buf) {
    if (!ctx) return nullptr;
    BufferRec* b = get_buffer(ctx, buf, "device_ptr");
    return b ? b->ptr : nullptr;
}

size_t rve_buffer_size(rve_ctx* ctx, rve_buffer buf) {
    if (!ctx) return 0;
    BufferRec* b = get_buffer(ctx, buf, "size");
    return b ? b->bytes : 0;
}

uint32_t rve_buffer_bindless_index(rve_ctx* ctx, rve_buffer buf) {
    if (!ctx) return RVE_BUFFER_INVALID;
    BufferRec* b = get_buffer(ctx, buf, "bindless_index");
    return b ? buf : RVE_BUFFER_INVALID;
}

}  // extern "C"
