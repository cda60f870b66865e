// internal.h — shared by the translation units of librve_b200.so (not part of the ABI)
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>  // header-only NVTX 3: a no-op unless a profiler is attached

#include <cstdint>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/rve_b200.h"

namespace rve {

struct BufferRec {
    void* ptr = nullptr;
    size_t bytes = 0;
    bool owned = false;
    bool live = false;
    bool ipc_imported = false;  // mapped by rve_ipc_import: unmapped (cudaIpcCloseMemHandle) when the handle is destroyed
};

struct PyramidRec {
    uint32_t dim = 0, levels = 0;
    rve_buffer texels = RVE_BUFFER_INVALID;
    // device words of the persistent build's grid barrier (arrivals [0], generation [32]) and of a deferred upper part
    // (worker tickets [64], workers done [96]: zeroed by the build, counted by the cull that finishes the levels)
    rve_buffer counter = RVE_BUFFER_INVALID;
    bool upperPending = false;  // texel levels >= 6 and footprint-min levels >= 1 are still to be derived (see pyramid_build_impl)
    // Footprint-min copy of the chain (acceleration structure of rve_cull, derived from `texels`): per level of n x n
    // texels an (n+1)^2 image F[j][i] = min over the clamped 2x2 footprint {i-1,i} x {j-1,j}, i.e. the value of one
    // min-reduction sample whose footprint starts at texel (i-1, j-1). One load per sample instead of four.
    rve_buffer footmin = RVE_BUFFER_INVALID;
    bool footmin_valid = true;  // both buffers start zeroed, which is consistent
    bool live = false;
    // TMA path of rve_depth_pyramid_build: tensor map of the last depth image built from, and the tile geometry for it
    alignas(64) unsigned char depthMap[128] = {};
    const void* mapBase = nullptr;
    uint32_t mapW = 0, mapH = 0, boxW = 0, boxH = 0;
    // persistent single-launch build: geometry chosen for the depth image `persFor` (grid 0: does not qualify)
    const void* persFor = nullptr;
    uint32_t persBox = 0, persGrid = 0, persStages = 0, persStageFloats = 0;
    size_t persSmem = 0;
};

// Programmatic dependent launch: the kernel may become resident while its predecessor on the stream is still running;
// it must execute pdl_wait() before touching anything the predecessor wrote. pdl_trigger() in the predecessor lets the
// dependent launch early (without it the dependent launches when the predecessor completes: still correct).
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

}  // namespace rve

// A captured frame (rve_ctx_capture_begin / _end): the CUDA graph, its executable instance and the pinned blocks that hold
// the host data its copy nodes read at every replay.
struct rve_graph {
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    std::vector<void*> pinned;
    int device = 0;
};

struct rve_ctx {
    int device = 0;
    rve_graph* capture = nullptr;  // non-null between rve_ctx_capture_begin and rve_ctx_capture_end
    cudaStream_t stream = nullptr;
    bool owns_stream = false;
    int sm_count = 0;
    size_t l2_bytes = 0;
    uint64_t launches = 0;

    bool profiling = false;
    struct ProfPair {
        cudaEvent_t a, b;
    };
    std::vector<ProfPair> prof[4];

    std::mutex mu;  // guards the tables (destroy may come from another thread)
    std::vector<rve::BufferRec> buffers;
    std::vector<uint32_t> free_slots;
    std::vector<rve::PyramidRec> pyramids;

    // device mirror of buffers[i].ptr — the "bindless descriptor array"
    void** d_table = nullptr;
    size_t d_table_cap = 0;
    bool table_dirty = true;

    // scratch: cubo prefix sums for rve_cull (grown on demand)
    uint32_t* d_cubo_start = nullptr;
    size_t d_cubo_start_cap = 0;
    void* d_cubo_res = nullptr;  // resolved CUBO records (cull.cu CuboRes)
    size_t d_cubo_res_cap = 0;
    void* d_view_index = nullptr;  // rve_cull_views_matrices: per-view matrix indices
    size_t d_view_index_cap = 0;
    const float* view_matrices = nullptr;         // set for the duration of one rve_cull_views_matrices call
    const uint32_t* view_matrix_index = nullptr;  // (host array of that call)
    void* d_cubo_lod = nullptr;  // per-CUBO squared LOD thresholds
    size_t d_cubo_lod_cap = 0;
    void* d_views = nullptr;  // view table of a view-batched cull
    size_t d_views_cap = 0;
    void* d_view_stride = nullptr;  // per-CUBO output strides of a view-batched cull
    size_t d_view_stride_cap = 0;
    void* d_sort = nullptr;  // rve_sort_entity_ids: bitmap + block sums
    size_t d_sort_cap = 0;
    // staging of rvs_vectors_encode_sync_batch: one pinned block + its device twin, reused once the previous copy has completed
    unsigned char* sync_stage_host = nullptr;
    void* sync_stage_dev = nullptr;
    size_t sync_stage_bytes = 0;
    cudaEvent_t sync_stage_free = nullptr;
    bool sync_stage_busy = false;
    void* d_sort_delta = nullptr;  // ids that entered / left lists with a persistent bitmap (sort_lists_persistent)
    size_t d_sort_delta_cap = 0;
    uint64_t sort_clean_geometry = 0;  // (blocks, lists) for which the sort scratch's bitmaps are known to be zero
    bool sort_dirty = false;           // a sort was enqueued only in part: clear the scratch before the next one
    int cull_blocks_per_sm = 0;  // occupancy of cull_kernel, queried once
    size_t pyramid_smem_opt_in = 48 * 1024;  // dynamic shared memory pyramid_build_tma_kernel has been opted in for
    size_t pyramid_pers_smem_opt_in = 48 * 1024;  // ... and pyramid_persistent_kernel
    bool views_attr_set = false;  // cull_views_kernel's dynamic shared memory opt-in (per device, hence per context)
    uint32_t* d_gather_ticket = nullptr;  // [0] rve_gather_ids_p2p: CTAs-done counter; [1] CUBO entries with bad handles (cubo_prepare_kernel)
    // scratch: view-space light records for rve_cluster_assign
    float4* d_lights = nullptr;
    size_t d_lights_cap = 0;
    // scratch: cluster slice depth table
    float* d_slices = nullptr;
    size_t d_slices_cap = 0;
};

namespace rve {

int set_error(int code, const char* fmt, ...);
int p2p_preload();   // p2p.cu: forces the exchange / gather kernels to be loaded
int sync_preload();  // sync.cu: the same for the flag and row-forwarding kernels
int check_cuda(cudaError_t e, const char* what);
// resolves a handle; returns nullptr (and sets the error) when invalid
BufferRec* get_buffer(rve_ctx* ctx, rve_buffer h, const char* role);
PyramidRec* get_pyramid(rve_ctx* ctx, rve_pyramid h);
// makes sure the device pointer table reflects the host table (enqueued on the stream)
int sync_table(rve_ctx* ctx);
int ensure_scratch(rve_ctx* ctx, void** ptr, size_t* cap, size_t bytes);
// host -> device copy of a small table on the context's stream. Outside a capture: a plain (pageable, runtime-staged) copy.
// While a frame is being captured the bytes are first copied into a pinned block owned by the graph, so that the copy
// node has something to read at every replay.
int staged_h2d(rve_ctx* ctx, void* dst, const void* src, size_t bytes);
// Small host tables written by ONE kernel whose payload travels as kernel parameters (ctx.cu). Raw destinations; returns
// RVE_ERR_UNSUPPORTED (nothing enqueued) when the segments do not fit one launch.
struct RawSegment {
    void* dst;
    const void* src;
    size_t bytes;
};
int upload_batch_raw(rve_ctx* ctx, const RawSegment* segs, uint32_t n);
// The parameter block such an upload travels in (one CTA - or CTA-sized group of threads - per segment). A kernel that is
// launched anyway just ahead of the tables' consumer can carry it instead of upload_params_kernel (the persistent pyramid
// kernel of a fused encode does: one launch less per frame).
constexpr uint32_t kUploadMaxSegs = 16;
constexpr uint32_t kUploadMaxWords = 896;  // 3.5 KB of payload: the whole block stays under the 4 KB parameter limit
struct UploadBlock {
    uint32_t* dst[kUploadMaxSegs];
    uint32_t first[kUploadMaxSegs], count[kUploadMaxSegs];
    uint32_t n;
    uint32_t data[kUploadMaxWords];
};
int make_upload_block(const RawSegment* segs, uint32_t n, UploadBlock* blk);  // 0 (blk->n == 0: nothing to write) or RVE_ERR_UNSUPPORTED
int upload_block(rve_ctx* ctx, const UploadBlock& blk);                       // upload_params_kernel
// The tables cubo_prepare_kernel derives from a CUBO array, built on the host instead (cull.cu) for a caller that has the
// CUBOs and the LOD distances in host memory (the shim): appended as raw segments aimed at the context's scratch tables.
// `keep` owns the staging memory until the upload call returns.
struct HostCuboStaging {
    std::vector<uint32_t> start;
    std::vector<unsigned char> res;
    std::vector<float> lodSq;
};
int cull_host_tables(rve_ctx* ctx, const RveCullingUBOinstance* cubos, uint32_t n, const float* const* lodDistHost, HostCuboStaging* keep,
                     std::vector<RawSegment>* segs);
constexpr uint32_t kCullTablesPrepared = 1u;  // cull_launch flag: the scratch tables of this CUBO array are already uploaded
constexpr uint32_t kCullLateWait = 2u;        // ... and were complete before the grids directly ahead of the cull (a pyramid build) started:
                                              // the cull waits for those only before its first Hi-Z batch
int cull_launch(rve_ctx* ctx, const RveCullingUBO* ubo, uint32_t num_views, const rve_pyramid* pyramids, const RveViewStride* strides,
                rve_buffer cubos, rve_buffer model_matrices, rve_buffer render_layers, rve_buffer per_object_flags, uint32_t num_groups_x,
                uint32_t flags);
// cull.cu: CUtensorMap (128 bytes, 64-byte aligned, passed as void*) of a row-major f32 image, boxes of boxRows x boxCols
int make_image_map(const void* base, uint64_t cols, uint64_t rows, uint32_t boxCols, uint32_t boxRows, void* out);
// (re)derives the footprint-min chain of a pyramid from its texels when they changed since the last derivation
// Ascending copies of entity lists that OWN their bitmap (csrc/sort_ids.cu; the shim's sparse-set-backed commands): `full`
// rebuilds the bitmap from the list, otherwise only the ids that entered (setIds) / left (clearIds) the list since the last
// update are applied; either way `out` receives the `count` ids in ascending order. Two launches per 16 lists.
struct SortedListUpdate {
    rve_buffer ids;     // the list (read by a full build)
    uint32_t count;     // its length now
    rve_buffer out;     // >= count ids
    rve_buffer bitmap;  // >= sort_region_bytes(id_limit), owned by the list
    const uint32_t* setIds;
    uint32_t nSet;
    const uint32_t* clearIds;
    uint32_t nClear;
    bool full;
};
size_t sort_region_bytes(uint32_t id_limit);
int sort_lists_persistent(rve_ctx* ctx, const SortedListUpdate* u, uint32_t n_lists, uint32_t id_limit);
int pyramid_refresh_footmin(rve_ctx* ctx, PyramidRec* p, uint32_t firstLevel = 0);
// rve_depth_pyramid_build with an option: allowDefer lets the build stop after the tile levels (mips 0..5, footprint-min
// level 0) and leave PyramidRec::upperPending set - the caller enqueues a cull next, whose kernel finishes the upper levels
// in its start-up (cull_launch merges them, or runs pyramid_finish_upper first when it cannot)
// `carry` (deferring builds only): tables to be written ahead of that cull; the build's kernel writes them, or they go out first
int pyramid_build_impl(rve_ctx* ctx, rve_pyramid pyr, rve_buffer depth, uint32_t width, uint32_t height, bool allowDefer,
                       const UploadBlock* carry = nullptr);
int pyramid_finish_upper(rve_ctx* ctx, PyramidRec* p);  // no-op unless upperPending
// marks pyramids whose texel buffer is `buf` as changed (buffer upload / copy / fill paths)
void pyramid_note_buffer_write(rve_ctx* ctx, rve_buffer buf);
// ... and before the library reads a buffer on the caller's behalf (download, copy source): a pyramid whose upper levels a
// deferring build left open is completed first when `buf` is its texel buffer
int pyramid_before_buffer_read(rve_ctx* ctx, rve_buffer buf);

// RAII: brackets a kernel launch with events when ctx->profiling
struct ProfScope {
    rve_ctx* ctx;
    int tag;
    cudaEvent_t b = nullptr;
    ProfScope(rve_ctx* c, int t) : ctx(c), tag(t) {
        if (!ctx->profiling) return;
        cudaEvent_t a;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        cudaEventRecord(a, ctx->stream);
        ctx->prof[tag].push_back({a, b});
    }
    ~ProfScope() {
        if (b) cudaEventRecord(b, ctx->stream);
    }
};

// Host-side profiler zone around an encode block, named like the reference's Tracy zones for the same block
// (RVE_PROFILE_FN_N, include/RavEngine/Profile.hpp:5-30; zones at src/RenderEngine_Draw.cpp:100,578,911,1068,1222,2477).
struct ProfileZone {
    explicit ProfileZone(const char* name) { nvtxRangePushA(name); }
    ~ProfileZone() { nvtxRangePop(); }
    ProfileZone(const ProfileZone&) = delete;
    ProfileZone& operator=(const ProfileZone&) = delete;
};
#define RVE_ZONE_CAT2(a, b) a##b
#define RVE_ZONE_CAT(a, b) RVE_ZONE_CAT2(a, b)
#define RVE_PROFILE_FN_N(name) rve::ProfileZone RVE_ZONE_CAT(rve_zone_, __LINE__)(name)

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

#define RVE_CUDA(expr)                                        \
    do {                                                      \
        int _rc = rve::check_cuda((expr), #expr);             \
        if (_rc != 0) return _rc;                             \
    } while (0)

#define RVE_REQUIRE(cond, msg)                                              \
    do {                                                                    \
        if (!(cond)) return rve::set_error(RVE_ERR_INVALID_ARGUMENT, "%s", msg); \
    } while (0)

// ---- device helpers ----
// Arithmetic contract shared with oracle/oracle.cpp (see its header and DESIGN.md): the library is
// compiled with -fmad=false so ptxas never fuses on its own; every fused multiply-add is an
// explicit __fmaf_rn in the same place the oracle has one. min/max: the oracle states GLSL's
// (y<x?y:x); fminf/fmaxf (one FMNMX) give the same value for every non-NaN input.
#ifdef __CUDACC__
__device__ __forceinline__ float gmin(float x, float y) { return fminf(x, y); }
__device__ __forceinline__ float gmax(float x, float y) { return fmaxf(x, y); }
__device__ __forceinline__ float dot3(float ax, float ay, float az, float bx, float by, float bz) {
    return __fmaf_rn(az, bz, __fmaf_rn(ay, by, ax * bx));
}

// Packed pairs of IEEE f32 (sm_100 FFMA2 / FMUL2 / FADD2): two independent, individually rounded operations per
// issued instruction - same bits as two scalar operations, half the issue slots. The cull kernel is issue-bound.
struct f2 {
    float x, y;
};
__device__ __forceinline__ unsigned long long f2_pack(f2 a) {
    unsigned long long r;
    asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a.x), "f"(a.y));
    return r;
}
__device__ __forceinline__ f2 f2_unpack(unsigned long long r) {
    f2 d;
    asm("mov.b64 {%0,%1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(r));
    return d;
}
#ifdef RVE_F2_SCALAR  // tuning aid: the same operations as scalar instructions
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { return f2{__fmaf_rn(a.x, b.x, c.x), __fmaf_rn(a.y, b.y, c.y)}; }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) { return f2{a.x * b.x, a.y * b.y}; }
__device__ __forceinline__ f2 add2(f2 a, f2 b) { return f2{a.x + b.x, a.y + b.y}; }
#else
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(f2_pack(a)), "l"(f2_pack(b)), "l"(f2_pack(c)));
    return f2_unpack(d);
}
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
    unsigned long long d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(f2_pack(a)), "l"(f2_pack(b)));
    return f2_unpack(d);
}
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
    unsigned long long d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(f2_pack(a)), "l"(f2_pack(b)));
    return f2_unpack(d);
}
#endif
__device__ __forceinline__ f2 splat2(float s) { return f2{s, s}; }

// 128-bit read-only load that does not allocate in L1 (streamed once)
__device__ __forceinline__ float4 ld_stream_f4(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ uint64_t pyr_level_offset(uint32_t dim, uint32_t level) {
    const uint64_t d0 = dim, dl = dim >> level;
    return 4ull * (d0 * d0 - dl * dl) / 3ull;
}
#endif
// start of level `level` in the footprint-min chain: sum over l < level of ((dim>>l)+1)^2
__host__ __device__ inline uint64_t footmin_level_offset(uint32_t dim, uint32_t level) {
    const uint64_t d0 = dim, dl = dim >> level;
    return 4ull * (d0 * d0 - dl * dl) / 3ull + 4ull * (d0 - dl) + level;
}

// Level starts of a pyramid, computed once on the host (per-thread 64-bit divisions by 3 otherwise): px texel chain,
// fm footprint-min chain ([levels] = total), run = first work item of each level for footmin_items.
struct PyrOffsets {
    uint32_t px[16], fm[16], run[16];
};
// Work items of footmin_items per level of n x n texels: runs of 8 consecutive entries of one row of the (n+1)^2
// footprint-min image (n/8 + 1 runs per row; the last one is the single entry fx = n), or single entries while n < 8.
inline uint32_t footmin_items_of_level(uint32_t n) { return (n & 7u) ? (n + 1) * (n + 1) : (n + 1) * (n / 8 + 1); }
inline PyrOffsets make_pyr_offsets(uint32_t dim, uint32_t levels) {
    PyrOffsets o{};
    uint32_t items = 0;
    for (uint32_t l = 0; l <= levels && l < 16; ++l) {
        const uint64_t d0 = dim, dl = l < 32 ? (dim >> l) : 0;
        o.px[l] = uint32_t(4ull * (d0 * d0 - dl * dl) / 3ull);
        o.fm[l] = uint32_t(footmin_level_offset(dim, l));
        o.run[l] = items;
        if (l >= 1 && l < levels) items += footmin_items_of_level(dim >> l);
    }
    return o;
}
#ifdef __CUDACC__
// Texel levels topLevel .. levels-1 (what is left above the tiles: an image of at most 32 x 32) and their footprint-min
// levels, by ONE group of NT threads (tid < NT) in `scratch` (shared memory, >= n^2 * 4/3 + 2 floats for the n x n level
// topLevel-1); sync() is the group's barrier. Same operand order as depthpyramid.csh's 2x2 min (reduce_level).
template <int NT, class Sync>
__device__ __forceinline__ void pyramid_top_chain(float* __restrict__ pyr, float* __restrict__ fm, const PyrOffsets& offs, uint32_t dim, uint32_t levels,
                                                  uint32_t topLevel, float* scratch, uint32_t tid, Sync sync) {
    const uint32_t r = dim >> (topLevel - 1);
    const float* top = pyr + offs.px[topLevel - 1];
    for (uint32_t k = tid; k < r * r; k += NT) scratch[k] = __ldcg(top + k);
    sync();
    uint32_t in_off = 0, n = r, level = topLevel;
    while (n > 1 && level < levels) {
        const uint32_t n_out = n / 2, out_off = in_off + n * n;
        const float* lv = scratch + in_off;
        float* out = scratch + out_off;
        float* gdst = pyr + offs.px[level];
        for (uint32_t k = tid; k < n_out * n_out; k += NT) {
            const uint32_t x = k % n_out, y = k / n_out;
            const float v = fminf(fminf(lv[(2 * y) * n + 2 * x], lv[(2 * y) * n + 2 * x + 1]), fminf(lv[(2 * y + 1) * n + 2 * x], lv[(2 * y + 1) * n + 2 * x + 1]));
            out[k] = v;
            gdst[k] = v;
        }
        sync();
        float* f = fm + offs.fm[level];
        const uint32_t m1 = n_out + 1;
        for (uint32_t k = tid; k < m1 * m1; k += NT) {
            const uint32_t fx = k % m1, fy = k / m1;
            const uint32_t x0 = fx ? fx - 1 : 0, x1 = fx < n_out ? fx : n_out - 1;
            const uint32_t y0 = fy ? fy - 1 : 0, y1 = fy < n_out ? fy : n_out - 1;
            f[k] = fminf(fminf(out[y0 * n_out + x0], out[y0 * n_out + x1]), fminf(out[y1 * n_out + x0], out[y1 * n_out + x1]));
        }
        in_off = out_off;
        n = n_out;
        ++level;
    }
}
// Footprint-min levels 1 .. lastFm-1 from their texel levels (already complete and visible; read through L2): worker
// `worker` of `workers` takes items worker, worker + workers, ... A run of 8 entries costs one scalar and two 128-bit
// loads per texel row instead of 32 scalar loads; entry (fx, fy) = min over texels {fx-1, fx} x {fy-1, fy} clamped, in
// the same operand order as the one-entry-per-thread kernels.
__device__ __forceinline__ void footmin_items(const float* __restrict__ pyr, float* __restrict__ fm, const PyrOffsets& offs, uint32_t dim,
                                              uint32_t lastFm, uint32_t worker, uint32_t workers) {
    const uint32_t end = offs.run[lastFm];
    for (uint32_t it = worker; it < end; it += workers) {
        uint32_t level = 1;
        while (level + 1 < lastFm && it >= offs.run[level + 1]) ++level;
        const uint32_t n = dim >> level, n1 = n + 1, local = it - offs.run[level];
        const float* img = pyr + offs.px[level];
        float* out = fm + offs.fm[level];
        if (n & 7u) {
            const uint32_t fy = local / n1, fx = local - fy * n1;
            const uint32_t x0 = fx ? fx - 1 : 0, x1 = fx < n ? fx : n - 1;
            const uint32_t y0 = fy ? fy - 1 : 0, y1 = fy < n ? fy : n - 1;
            out[local] = fminf(fminf(__ldcg(img + y0 * n + x0), __ldcg(img + y0 * n + x1)), fminf(__ldcg(img + y1 * n + x0), __ldcg(img + y1 * n + x1)));
            continue;
        }
        const uint32_t perRow = n / 8 + 1;
        const uint32_t fy = local / perRow, fx0 = (local - fy * perRow) * 8;
        const uint32_t y0 = fy ? fy - 1 : 0, y1 = fy < n ? fy : n - 1;
        const float* ra = img + y0 * n;
        const float* rb = img + y1 * n;
        const uint32_t xl = fx0 ? fx0 - 1 : 0;
        const float al = __ldcg(ra + xl), bl = __ldcg(rb + xl);
        float* o = out + fy * n1 + fx0;
        if (fx0 == n) {  // the entry past the last texel column: both columns clamp to n - 1
            o[0] = fminf(fminf(al, al), fminf(bl, bl));
            continue;
        }
        const float4 a0 = __ldcg(reinterpret_cast<const float4*>(ra + fx0)), a1 = __ldcg(reinterpret_cast<const float4*>(ra + fx0 + 4));
        const float4 b0 = __ldcg(reinterpret_cast<const float4*>(rb + fx0)), b1 = __ldcg(reinterpret_cast<const float4*>(rb + fx0 + 4));
        o[0] = fminf(fminf(al, a0.x), fminf(bl, b0.x));
        o[1] = fminf(fminf(a0.x, a0.y), fminf(b0.x, b0.y));
        o[2] = fminf(fminf(a0.y, a0.z), fminf(b0.y, b0.z));
        o[3] = fminf(fminf(a0.z, a0.w), fminf(b0.z, b0.w));
        o[4] = fminf(fminf(a0.w, a1.x), fminf(b0.w, b1.x));
        o[5] = fminf(fminf(a1.x, a1.y), fminf(b1.x, b1.y));
        o[6] = fminf(fminf(a1.y, a1.z), fminf(b1.y, b1.z));
        o[7] = fminf(fminf(a1.z, a1.w), fminf(b1.z, b1.w));
    }
}
#endif

}  // namespace rve
