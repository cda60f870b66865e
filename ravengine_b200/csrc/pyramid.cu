// pyramid.cu — Hi-Z depth pyramid.
// Replaces generatePyramid (src/RenderEngine_Draw.cpp:2475-2514): one raster pass of
// depthpyramidcopy.fsh (mip 0) + one depthpyramid.csh dispatch per further mip (10 launches
// with barriers for 1024^2) by TWO launches: in the first every CTA produces a 32x32 tile of mip 0
// from the depth buffer and reduces it through shared memory down to a single texel (mips 0..5,
// plus level 0 of the cull's footprint-min chain); the second (programmatic dependent launch: its
// blocks are resident before the first one drains) derives footprint-min levels 1..5 with one thread
// per entry while one extra block reduces the remaining 32x32 image (mips 6..10) and their
// footprint-min levels out of shared memory. No ticket counter, no grid-wide fence, no serial tail
// inside the big kernel (the r01l ncu capture had 45 % of its samples there).
// r02f: pyramid_persistent_kernel (further down) builds the tile levels from a TMA ring with register / shuffle tiles; as
// part of a fused encode it stops there and the cull kernel that follows derives the upper levels in its start-up
// (pyramid_build_impl / PyramidRec::upperPending).
//
// Sampler model = oracle.cpp sample_min(): min over the clamped 2x2 bilinear footprint
// floor(u*W-0.5)+{0,1} (src/RenderEngine.cpp:323-331, reduction mode Minimum).
#include <cmath>

#include <cuda.h>  // CUtensorMap (type only)

#include "internal.h"

namespace rve {

constexpr int kPyrThreads = 256;
constexpr int kPyrTile = 32;

__device__ __forceinline__ void footprint(float x, uint32_t n, uint32_t& i0, uint32_t& i1) {
    const float fx = floorf(x);
    if (!(fx >= 0.0f)) {
        i0 = 0;
        i1 = 0;
    } else if (fx >= float(n - 1)) {
        i0 = n - 1;
        i1 = n - 1;
    } else {
        i0 = uint32_t(fx);
        i1 = i0 + 1;
    }
}

__device__ __forceinline__ float sample_min_src(const float* __restrict__ img, uint32_t w, uint32_t h, float u, float v) {
    const float x = __fmaf_rn(u, float(w), -0.5f);  // arithmetic contract: u*W - 0.5 is one fma
    const float y = __fmaf_rn(v, float(h), -0.5f);
    uint32_t x0, x1, y0, y1;
    footprint(x, w, x0, x1);
    footprint(y, h, y0, y1);
    const float a = __ldg(img + uint64_t(y0) * w + x0);
    const float b = __ldg(img + uint64_t(y0) * w + x1);
    const float c = __ldg(img + uint64_t(y1) * w + x0);
    const float d = __ldg(img + uint64_t(y1) * w + x1);
    return gmin(gmin(a, b), gmin(c, d));
}

// reduce `lv` (n x n, shared) into `out` (n/2 x n/2, shared) and global level
__device__ __forceinline__ void reduce_level(const float* lv, uint32_t n_in, float* out, uint32_t n_out, float* __restrict__ gdst,
                                             uint32_t g_dim_out, uint32_t gx0, uint32_t gy0) {
    const uint32_t sh = uint32_t(__ffs(int(n_out))) - 1u;  // n_out is a power of two
    for (uint32_t i = threadIdx.x; i < n_out * n_out; i += blockDim.x) {
        const uint32_t x = i & (n_out - 1u), y = i >> sh;
        const float a = lv[(2 * y) * n_in + 2 * x];
        const float b = lv[(2 * y) * n_in + 2 * x + 1];
        const float c = lv[(2 * y + 1) * n_in + 2 * x];
        const float d = lv[(2 * y + 1) * n_in + 2 * x + 1];
        const float m = gmin(gmin(a, b), gmin(c, d));
        out[i] = m;
        gdst[uint64_t(gy0 + y) * g_dim_out + (gx0 + x)] = m;
    }
}

// Everything a tile does after its (tile+1)^2 mip-0 samples sit in s_lv: level 0 of the footprint-min chain and the
// in-tile mips 1..log2(tile), each written to the chain and reduced further from shared memory (depthpyramid.csh).
// FULL: the tile is kPyrTile wide (index arithmetic with compile-time constants: the kernel is instruction-bound)
template <bool FULL>
__device__ __forceinline__ void tile_levels(float* s_lv, float* __restrict__ pyr, float* __restrict__ fm, uint32_t dim, uint32_t levels,
                                            uint32_t tileArg, uint32_t tx0, uint32_t ty0) {
    const uint32_t tile = FULL ? uint32_t(kPyrTile) : tileArg;
    const uint32_t S = tile + 1;
    __syncthreads();
    {   // footprint-min level 0: F[fy][fx] = min of texels {fx-1,fx} x {fy-1,fy} (clamped), (dim+1)^2 entries
        const uint32_t n1 = dim + 1;
        for (uint32_t i = threadIdx.x; i < tile * tile; i += blockDim.x) {
            const uint32_t lx = i % tile + 1, ly = i / tile + 1;
            fm[uint64_t(ty0 + ly) * n1 + (tx0 + lx)] =
                gmin(gmin(s_lv[(ly - 1) * S + lx - 1], s_lv[(ly - 1) * S + lx]), gmin(s_lv[ly * S + lx - 1], s_lv[ly * S + lx]));
        }
        if (tx0 == 0)
            for (uint32_t ly = threadIdx.x + 1; ly <= tile; ly += blockDim.x) fm[uint64_t(ty0 + ly) * n1] = gmin(s_lv[(ly - 1) * S], s_lv[ly * S]);
        if (ty0 == 0)
            for (uint32_t lx = threadIdx.x + 1; lx <= tile; lx += blockDim.x) fm[tx0 + lx] = gmin(s_lv[lx - 1], s_lv[lx]);
        if (tx0 == 0 && ty0 == 0 && threadIdx.x == 0) fm[0] = s_lv[0];
    }

    // mips 1..log2(tile) from shared memory — depthpyramid.csh
    uint32_t in_off = 0, n = tile, n_stride = S, level = 1;
    while (n > 1 && level < levels) {
        const uint32_t n_out = n / 2;
        const uint32_t out_off = in_off + (level == 1 ? S * S : n * n);
        reduce_level(s_lv + in_off, n_stride, s_lv + out_off, n_out, pyr + pyr_level_offset(dim, level), dim >> level,
                     tx0 >> level, ty0 >> level);
        __syncthreads();
        in_off = out_off;
        n = n_out;
        n_stride = n_out;
        ++level;
    }
}

__global__ void __launch_bounds__(kPyrThreads)
pyramid_build_kernel(const float* __restrict__ depth, uint32_t w, uint32_t h, float* __restrict__ pyr, float* __restrict__ fm,
                     uint32_t dim, uint32_t levels, uint32_t tile) {
    // mip 0 of the tile plus one halo column / row ((tile+1)^2, stride tile+1), then the in-tile levels 1.. back to back
    __shared__ float s_lv[(kPyrTile + 1) * (kPyrTile + 1) + (kPyrTile * kPyrTile) / 3 + 2];

    // let pyramid_upper_kernel's blocks become resident as this grid's CTAs retire (they wait in griddepcontrol.wait)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const uint32_t tiles_per_row = dim / tile;
    const uint32_t tx0 = (blockIdx.x % tiles_per_row) * tile;
    const uint32_t ty0 = (blockIdx.x / tiles_per_row) * tile;

    // mip 0 — depthpyramidcopy.fsh: texcoord = gl_FragCoord.xy / targetDim (pixel centres). The tile is sampled with
    // one extra column and row (clamped at the image edge) so that level 0 of the footprint-min chain - 3/4 of it -
    // is produced here, from shared memory, instead of by a second pass over the texels.
    const uint32_t S = tile + 1;
    for (uint32_t i = threadIdx.x; i < S * S; i += blockDim.x) {
        const uint32_t lx = i % S, ly = i / S;
        uint32_t x = tx0 + lx, y = ty0 + ly;
        x = x < dim ? x : dim - 1;
        y = y < dim ? y : dim - 1;
        const float u = (float(x) + 0.5f) / float(dim);
        const float v = (float(y) + 0.5f) / float(dim);
        const float m = sample_min_src(depth, w, h, u, v);
        s_lv[i] = m;
        if (lx < tile && ly < tile) pyr[uint64_t(y) * dim + x] = m;
    }
    tile_levels<false>(s_lv, pyr, fm, dim, levels, tile, tx0, ty0);
    // levels above log2(tile) and the footprint-min levels >= 1: pyramid_upper_kernel
}


namespace pyr {
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "PYR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra PYR_DONE;\n"
        "bra PYR_WAIT;\n"
        "PYR_DONE:\n"
        "}\n" ::"r"(smem_u32(b)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_box(void* dst, const void* map, int32_t x, int32_t y, uint64_t* b) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(dst)),
                 "l"(map), "r"(x), "r"(y), "r"(smem_u32(b))
                 : "memory");
}
}  // namespace pyr

// first texel of the clamped 2x2 footprint of mip-0 sample `t` along an axis of n source texels (sample_min_src's
// arithmetic); host and device evaluate the same IEEE operations, so the host can size the source box exactly
__host__ __device__ inline uint32_t footprint_start(uint32_t t, uint32_t dim, uint32_t n) {
    const float u = (float(t) + 0.5f) / float(dim);
#ifdef __CUDA_ARCH__
    const float fx = floorf(__fmaf_rn(u, float(n), -0.5f));
#else
    const float fx = floorf(fmaf(u, float(n), -0.5f));
#endif
    if (!(fx >= 0.0f)) return 0;
    if (fx >= float(n - 1)) return n - 1;
    return uint32_t(fx);
}

// The same tile work fed by ONE TMA tensor copy per CTA: the source rectangle that the tile's 33 x 33 samples touch
// (124 x 71 texels for 3840x2160 -> 1024^2) lands in shared memory asynchronously, in full lines, with the copy of every
// resident CTA in flight at once; CTAs then finish at different times, so the sampling / reduction of one overlaps the
// DRAM traffic of the others (the LSU-fed kernel moved through load -> five barrier-separated reductions in lock step).
// The box must start at a 16-byte aligned global address (x a multiple of 4 texels): anything else raises "illegal
// instruction" on sm_100a (tools/tma_probe.cu); boxes up to 256 x 128 texels in one copy are fine.
__global__ void __launch_bounds__(kPyrThreads)
pyramid_build_tma_kernel(const __grid_constant__ CUtensorMap depthMap, uint32_t w, uint32_t h, float* __restrict__ pyr, float* __restrict__ fm,
                         uint32_t dim, uint32_t levels, uint32_t boxW, uint32_t boxH, uint32_t* __restrict__ deferSync) {
    extern __shared__ __align__(128) unsigned char s_dyn[];
    __shared__ uint64_t s_bar;
    if (deferSync && blockIdx.x == 0 && threadIdx.x == 0) deferSync[64] = 0, deferSync[96] = 0;  // the upper part is left to the next cull
    float* s_src = reinterpret_cast<float*>(s_dyn);
    float* s_lv = s_src + boxW * boxH;
    constexpr uint32_t tile = kPyrTile, S = kPyrTile + 1;
    const uint32_t tiles_per_row = dim / tile;
    const uint32_t tx0 = (blockIdx.x % tiles_per_row) * tile;
    const uint32_t ty0 = (blockIdx.x / tiles_per_row) * tile;
    // TMA needs a 16-byte aligned global start address: the box begins at the 4-texel boundary at or left of the first footprint
    const uint32_t sx0 = footprint_start(tx0, dim, w) & ~3u, sy0 = footprint_start(ty0, dim, h);
    if (threadIdx.x == 0) {
        pyr::mbar_init(&s_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        pyr::mbar_expect_tx(&s_bar, boxW * boxH * 4u);
        pyr::tma_load_box(s_src, &depthMap, int32_t(sx0), int32_t(sy0), &s_bar);
    }
    // The footprints are separable: sample (x, y) reads source columns {fx0[x], fx1[x]} and rows {fy0[y], fy1[y]}. The 2 x 33
    // pairs of the tile are derived once (box-local), while the tensor copy is in flight - the per-sample work shrinks
    // from two IEEE divisions, two floors and two clamps to four table reads (the kernel was instruction-bound: ncu r02).
    __shared__ uint16_t s_f[4][S];
    if (threadIdx.x < 2 * S) {
        const uint32_t axis = threadIdx.x / S, l = threadIdx.x % S;  // 0: x, 1: y
        uint32_t t = (axis ? ty0 : tx0) + l;
        t = t < dim ? t : dim - 1;
        const uint32_t n = axis ? h : w;
        const float u = (float(t) + 0.5f) / float(dim);
        uint32_t i0, i1;
        footprint(__fmaf_rn(u, float(n), -0.5f), n, i0, i1);
        const uint32_t origin = axis ? sy0 : sx0;
        s_f[2 * axis][l] = uint16_t(i0 - origin);
        s_f[2 * axis + 1][l] = uint16_t(i1 - origin);
    }
    __syncthreads();  // the barrier is initialised and the tables are written for everybody
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    pyr::mbar_wait(&s_bar, 0);

    for (uint32_t i = threadIdx.x; i < S * S; i += kPyrThreads) {
        const uint32_t lx = i % S, ly = i / S;
        const uint32_t r0 = uint32_t(s_f[2][ly]) * boxW, r1 = uint32_t(s_f[3][ly]) * boxW;
        const uint32_t c0 = s_f[0][lx], c1 = s_f[1][lx];
        const float m = gmin(gmin(s_src[r0 + c0], s_src[r0 + c1]), gmin(s_src[r1 + c0], s_src[r1 + c1]));
        s_lv[i] = m;
        if (lx < tile && ly < tile) pyr[uint64_t(ty0 + ly) * dim + (tx0 + lx)] = m;
    }
    tile_levels<true>(s_lv, pyr, fm, dim, levels, tile, tx0, ty0);
}

__device__ __forceinline__ float footmin_entry(const float* img, uint32_t n, uint32_t fx, uint32_t fy) {
    const uint32_t x0 = fx ? fx - 1 : 0, x1 = fx < n ? fx : n - 1;
    const uint32_t y0 = fy ? fy - 1 : 0, y1 = fy < n ? fy : n - 1;
    return gmin(gmin(img[y0 * n + x0], img[y0 * n + x1]), gmin(img[y1 * n + x0], img[y1 * n + x1]));
}

// Second launch of a build. Block 0: texel levels topLevel.. (the (dim >> (topLevel-1))^2 <= 32x32 image left by the
// tiles, reduced in shared memory) and their footprint-min levels. Blocks 1..: footprint-min levels 1..topLevel-1,
// one thread per entry. Launched with programmatic stream serialization: everything before griddepcontrol.wait
// overlaps the tail of pyramid_build_kernel.
__global__ void __launch_bounds__(kPyrThreads)
pyramid_upper_kernel(float* __restrict__ pyr, float* __restrict__ fm, uint32_t dim, uint32_t levels, uint32_t topLevel, uint32_t earlyTrigger,
                     const __grid_constant__ PyrOffsets offs) {
    __shared__ float s_lv[kPyrTile * kPyrTile + (kPyrTile * kPyrTile) / 3 + 2];
    // earlyTrigger (RVE_UPPER_TRIGGER=1, off by default): a cull enqueued behind the build becomes resident now and waits for
    // this grid before its first Hi-Z batch. Measured on B200 (profiles/r02_fused_variants.txt): the persistent cull then
    // takes 64 us LONGER - its CTAs are placed while pyramid CTAs still hold SM resources and the late ones carry a full
    // static share - so by default the cull launches when this grid has drained.
    if (earlyTrigger) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (blockIdx.x == 0) {
        if (topLevel >= levels) return;
        const uint32_t r = dim >> (topLevel - 1);  // <= 32
        const float* src = pyr + pyr_level_offset(dim, topLevel - 1);
        asm volatile("griddepcontrol.wait;" ::: "memory");
        for (uint32_t i = threadIdx.x; i < r * r; i += blockDim.x) s_lv[i] = __ldcg(src + i);
        __syncthreads();
        uint32_t in_off = 0, n = r, level = topLevel;
        while (n > 1 && level < levels) {
            const uint32_t n_out = n / 2;
            const uint32_t out_off = in_off + n * n;
            reduce_level(s_lv + in_off, n, s_lv + out_off, n_out, pyr + pyr_level_offset(dim, level), dim >> level, 0, 0);
            __syncthreads();
            // footprint-min of the level just produced, from shared memory
            float* f = fm + footmin_level_offset(dim, level);
            for (uint32_t i = threadIdx.x; i < (n_out + 1) * (n_out + 1); i += blockDim.x)
                f[i] = footmin_entry(s_lv + out_off, n_out, i % (n_out + 1), i / (n_out + 1));
            in_off = out_off;
            n = n_out;
            ++level;
        }
        return;
    }
    // level offsets come from the host (PyrOffsets): deriving them per thread costs 64-bit divisions by 3
    const uint32_t lastFm = topLevel < levels ? topLevel : levels;
    const uint32_t first = offs.fm[1], end = offs.fm[lastFm];
    // two entries per thread (i, i + blockDim.x): the whole grid is resident in one wave at dim 1024
    const uint32_t base = (blockIdx.x - 1) * (2 * blockDim.x) + threadIdx.x + first;
    uint32_t idx[2], n[2], fx[2], fy[2];
    const float* img[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const uint32_t i = base + k * blockDim.x;
        uint32_t level = 1;
        while (level + 1 < levels && i >= offs.fm[level + 1]) ++level;
        n[k] = dim >> level;
        const uint32_t local = i - offs.fm[level];
        fy[k] = local / (n[k] + 1);
        fx[k] = local - fy[k] * (n[k] + 1);
        img[k] = pyr + offs.px[level];
        idx[k] = i;
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    float v[2];
#pragma unroll
    for (int k = 0; k < 2; ++k)
        if (idx[k] < end) v[k] = footmin_entry(img[k], n[k], fx[k], fy[k]);
#pragma unroll
    for (int k = 0; k < 2; ++k)
        if (idx[k] < end) fm[idx[k]] = v[k];
}

// ---- the whole build in ONE persistent launch -------------------------------------------------------------------------
// Two CTAs per SM, all resident. A producer warp streams the tiles' source boxes through a ring of shared-memory stages
// (TMA, full / empty mbarriers): with 3 stages per CTA nearly the whole 33 MB source is requested before the first tile is
// sampled. Eight compute warps work in registers: warp w samples rows 4w .. 4w+4 of the 33 x 33 sample grid (lane = column;
// the halo column by lanes 0..4), and from those five registers per lane come, by shuffles alone, mip 0, level 0 of the
// footprint-min chain, mip 1 (two rows per warp) and mip 2 (one row per warp); the 8 x 8 mip-2 image crosses warps through
// 256 B of shared memory and one warp finishes mips 3..5. ~6x fewer instructions per tile than the shared-memory ladder of
// tile_levels (which the ncu capture showed instruction-bound), no __syncthreads ladder, one named barrier per tile.
// Then a grid-wide barrier (sense-reversing, self re-arming - safe under graph replay) replaces the second launch: CTA 0
// reduces the remaining 32 x 32 image (mips 6..10 + their footprint-min levels), the others derive footprint-min levels 1..5.
// u = (t + 0.5) / dim is evaluated as (t + 0.5) * (1 / dim): dim is a power of two, so both are exact and equal.
constexpr int kPersWarps = 8;
constexpr int kPersComputeThreads = kPersWarps * 32;
constexpr int kPersThreads = kPersComputeThreads + 32;  // + the producer warp
constexpr int kPersMaxStages = 6;

namespace pyr {
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void compute_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kPersComputeThreads) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(uint32_t* p, uint32_t v) { asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
// clamped 2x2 footprint of sample t along an axis of n source texels (sample_min_src's arithmetic)
__device__ __forceinline__ void axis_footprint(uint32_t t, float invDim, uint32_t n, uint32_t& i0, uint32_t& i1) {
    const float u = (float(t) + 0.5f) * invDim;
    footprint(__fmaf_rn(u, float(n), -0.5f), n, i0, i1);
}
}  // namespace pyr

// all compute threads of all CTAs of the grid; sync[0] arrivals of the current generation, sync[32] the generation
__device__ __forceinline__ void pers_grid_barrier(uint32_t* sync, uint32_t G) {
    pyr::compute_sync();
    if (threadIdx.x == 0) {
        const uint32_t gen = pyr::ld_acquire(sync + 32);  // cannot advance before this CTA has arrived
        __threadfence();
        const uint32_t old = atomicInc(sync, G - 1);  // wraps to 0 with the last arrival: re-armed for the next launch
        if (old == G - 1) pyr::st_release(sync + 32, gen + 1);
        else
            while (pyr::ld_acquire(sync + 32) == gen) __nanosleep(32);
        __threadfence();
    }
    pyr::compute_sync();
}

__device__ __forceinline__ void
pyramid_persistent_body(const CUtensorMap& depthMap, uint32_t w, uint32_t h, float* __restrict__ pyr, float* __restrict__ fm, uint32_t dim, uint32_t levels,
                        uint32_t boxW, uint32_t boxH, uint32_t stages, uint32_t stageFloats, uint32_t topLevel, uint32_t* __restrict__ gridSync,
                        const PyrOffsets& offs, uint32_t debug, uint32_t deferUpper) {
    extern __shared__ __align__(128) unsigned char s_dyn[];
    __shared__ uint64_t s_full[kPersMaxStages], s_empty[kPersMaxStages];
    __shared__ float s_l2[2][kPersWarps][8];
    float* s_src = reinterpret_cast<float*>(s_dyn);
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t tilesPerRow = dim / kPyrTile, numTiles = tilesPerRow * tilesPerRow;
    const uint32_t G = gridDim.x;
    const float invDim = 1.0f / float(dim);
    constexpr uint32_t kFull = 0xffffffffu;
    if (threadIdx.x == 0) {
        if (deferUpper && blockIdx.x == 0) gridSync[64] = 0, gridSync[96] = 0;  // the upper part is left to the next cull
        for (uint32_t s = 0; s < stages; ++s) {
            pyr::mbar_init(&s_full[s], 1);
            pyr::mbar_init(&s_empty[s], kPersWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp == kPersWarps) {  // producer: one lane feeds the ring, then the warp is done
        if (lane == 0 && !(debug & 8u)) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&depthMap) : "memory");
            uint32_t i = 0;
            for (uint32_t t = blockIdx.x; t < numTiles; t += G, ++i) {
                const uint32_t s = i % stages, use = i / stages;
                if (use) pyr::mbar_wait(&s_empty[s], (use - 1) & 1);
                const uint32_t tx0 = (t % tilesPerRow) * kPyrTile, ty0 = (t / tilesPerRow) * kPyrTile;
                uint32_t a, b, sx0, sy0;
                pyr::axis_footprint(tx0, invDim, w, sx0, a);
                pyr::axis_footprint(ty0, invDim, h, sy0, b);
                sx0 &= ~3u;  // 16-byte aligned box start (see pyramid_build_tma_kernel)
                pyr::mbar_expect_tx(&s_full[s], boxW * boxH * 4u);
                pyr::tma_load_box(s_src + size_t(s) * stageFloats, &depthMap, int32_t(sx0), int32_t(sy0), &s_full[s]);
            }
        }
        return;
    }

    const uint32_t n1 = dim + 1;
    uint32_t i = 0;
    for (uint32_t t = blockIdx.x; t < numTiles; t += G, ++i) {
        const uint32_t s = i % stages, use = i / stages;
        const uint32_t tx0 = (t % tilesPerRow) * kPyrTile, ty0 = (t / tilesPerRow) * kPyrTile;
        uint32_t sx0, sy0, c0, c1, ch0, ch1, ra, rb;
        pyr::axis_footprint(tx0, invDim, w, sx0, c0);
        pyr::axis_footprint(ty0, invDim, h, sy0, c0);
        sx0 &= ~3u;
        // box-local footprints: this lane's sample column, the halo column (sample 32), and - in lanes 0..4 - one of the warp's rows
        pyr::axis_footprint(tx0 + lane, invDim, w, c0, c1);
        const uint32_t xh = tx0 + kPyrTile < dim ? tx0 + kPyrTile : dim - 1;
        pyr::axis_footprint(xh, invDim, w, ch0, ch1);
        uint32_t yr = ty0 + 4 * warp + (lane < 4 ? lane : 4);
        yr = yr < dim ? yr : dim - 1;
        pyr::axis_footprint(yr, invDim, h, ra, rb);
        c0 -= sx0, c1 -= sx0, ch0 -= sx0, ch1 -= sx0;
        ra = (ra - sy0) * boxW, rb = (rb - sy0) * boxW;

        if (debug & 8u) continue;
        pyr::mbar_wait(&s_full[s], use & 1);
        if (debug & 4u) {  // RVE_PYRAMID_DEBUG (development only): loads without the tile work
            __syncwarp();
            if (lane == 0) pyr::mbar_arrive(&s_empty[s]);
            continue;
        }
        const float* src = s_src + size_t(s) * stageFloats;
        float m[5];
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const uint32_t r0 = __shfl_sync(kFull, ra, j), r1 = __shfl_sync(kFull, rb, j);
            m[j] = gmin(gmin(src[r0 + c0], src[r0 + c1]), gmin(src[r1 + c0], src[r1 + c1]));
        }
        const float e = gmin(gmin(src[ra + ch0], src[ra + ch1]), gmin(src[rb + ch0], src[rb + ch1]));  // halo column, row 4w + lane (lanes 0..4)
        __syncwarp();
        if (lane == 0) pyr::mbar_arrive(&s_empty[s]);  // the stage may be refilled once all eight warps have read it

        const uint32_t y0 = ty0 + 4 * warp;
#pragma unroll
        for (int j = 0; j < 4; ++j) pyr[size_t(y0 + j) * dim + tx0 + lane] = m[j];
        // level 0 of the footprint-min chain: F[y][x] = min of samples {x-1, x} x {y-1, y}; lane holds x - 1
        float hd[5];
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            float nb = __shfl_down_sync(kFull, m[j], 1);
            const float m32 = __shfl_sync(kFull, e, j);
            nb = lane == 31 ? m32 : nb;
            hd[j] = gmin(m[j], nb);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) fm[size_t(y0 + 1 + j) * n1 + tx0 + lane + 1] = gmin(hd[j], hd[j + 1]);
        if (tx0 == 0 && lane == 0) {
#pragma unroll
            for (int j = 0; j < 4; ++j) fm[size_t(y0 + 1 + j) * n1] = gmin(m[j], m[j + 1]);
        }
        if (ty0 == 0 && warp == 0) {
            fm[tx0 + lane + 1] = hd[0];
            if (tx0 == 0 && lane == 0) fm[0] = m[0];
        }
        // mip 1: rows 2w, 2w+1 of the tile's 16 x 16 (even lanes); mip 2: row w of its 8 x 8 (every fourth lane) - depthpyramid.csh
        const float p0 = gmin(m[0], __shfl_xor_sync(kFull, m[0], 1)), p1 = gmin(m[1], __shfl_xor_sync(kFull, m[1], 1));
        const float p2 = gmin(m[2], __shfl_xor_sync(kFull, m[2], 1)), p3 = gmin(m[3], __shfl_xor_sync(kFull, m[3], 1));
        const float l1a = gmin(p0, p1), l1b = gmin(p2, p3);
        if ((lane & 1u) == 0) {
            float* d1 = pyr + offs.px[1] + size_t((ty0 >> 1) + 2 * warp) * (dim >> 1) + (tx0 >> 1) + (lane >> 1);
            d1[0] = l1a;
            d1[dim >> 1] = l1b;
        }
        const float l2 = gmin(gmin(l1a, __shfl_xor_sync(kFull, l1a, 2)), gmin(l1b, __shfl_xor_sync(kFull, l1b, 2)));
        if ((lane & 3u) == 0) {
            pyr[offs.px[2] + size_t((ty0 >> 2) + warp) * (dim >> 2) + (tx0 >> 2) + (lane >> 2)] = l2;
            s_l2[i & 1][warp][lane >> 2] = l2;
        }
        pyr::compute_sync();
        if (warp == 0) {  // mips 3..5 of the tile from its 8 x 8 mip 2 (double-buffered: the next tile's writes go to the other copy)
            const uint32_t x = lane & 3u, y = (lane >> 2) & 3u;
            const float* q = &s_l2[i & 1][0][0];
            const float v3 = gmin(gmin(q[(2 * y) * 8 + 2 * x], q[(2 * y) * 8 + 2 * x + 1]), gmin(q[(2 * y + 1) * 8 + 2 * x], q[(2 * y + 1) * 8 + 2 * x + 1]));
            if (lane < 16) pyr[offs.px[3] + size_t((ty0 >> 3) + y) * (dim >> 3) + (tx0 >> 3) + x] = v3;
            const float h4 = gmin(v3, __shfl_xor_sync(kFull, v3, 1));
            const float v4 = gmin(h4, __shfl_xor_sync(kFull, h4, 4));
            if (lane < 16 && (lane & 5u) == 0) pyr[offs.px[4] + size_t((ty0 >> 4) + (lane >> 3)) * (dim >> 4) + (tx0 >> 4) + ((lane >> 1) & 1u)] = v4;
            const float h5 = gmin(v4, __shfl_xor_sync(kFull, v4, 2));
            const float v5 = gmin(h5, __shfl_xor_sync(kFull, h5, 8));
            if (lane == 0) pyr[offs.px[5] + size_t(ty0 >> 5) * (dim >> 5) + (tx0 >> 5)] = v5;
        }
    }

    if ((debug & 2u) || deferUpper) return;
    pers_grid_barrier(gridSync, G);
    if (debug & 1u) return;

    const uint32_t tid = threadIdx.x;  // < kPersComputeThreads
    if (blockIdx.x == 0 && G > 1 && topLevel >= levels) return;  // (cannot happen: more than one tile means levels > topLevel)
    if (blockIdx.x == 0 && topLevel < levels) {
        // texel levels topLevel.. (the 32 x 32 or smaller image the tiles left) and their footprint-min levels; the ring is idle now
        pyramid_top_chain<kPersComputeThreads>(pyr, fm, offs, dim, levels, topLevel, s_src, tid, [] { pyr::compute_sync(); });
        if (G > 1) return;
    }
    // footprint-min levels 1 .. topLevel-1 from the texel levels every CTA wrote before the barrier (read through L2)
    const uint32_t lastFm = topLevel < levels ? topLevel : levels;
    footmin_items(pyr, fm, offs, dim, lastFm, (G > 1 ? blockIdx.x - 1 : 0) * kPersComputeThreads + tid, (G > 1 ? G - 1 : 1) * kPersComputeThreads);
}

__global__ void __launch_bounds__(kPersThreads, 2)
pyramid_persistent_kernel(const __grid_constant__ CUtensorMap depthMap, uint32_t w, uint32_t h, float* __restrict__ pyr, float* __restrict__ fm,
                          uint32_t dim, uint32_t levels, uint32_t boxW, uint32_t boxH, uint32_t stages, uint32_t stageFloats, uint32_t topLevel,
                          uint32_t* __restrict__ gridSync, const __grid_constant__ PyrOffsets offs, uint32_t debug, uint32_t deferUpper) {
    pyramid_persistent_body(depthMap, w, h, pyr, fm, dim, levels, boxW, boxH, stages, stageFloats, topLevel, gridSync, offs, debug, deferUpper);
}

// The same build carrying a table upload (UploadBlock, internal.h) for the kernel that follows it: the fused encode's
// indirect-command / CUBO / prepared cull tables are written here instead of by a launch of their own.
__global__ void __launch_bounds__(kPersThreads, 2)
pyramid_persistent_carry_kernel(const __grid_constant__ CUtensorMap depthMap, uint32_t w, uint32_t h, float* __restrict__ pyr, float* __restrict__ fm,
                                uint32_t dim, uint32_t levels, uint32_t boxW, uint32_t boxH, uint32_t stages, uint32_t stageFloats, uint32_t topLevel,
                                uint32_t* __restrict__ gridSync, const __grid_constant__ PyrOffsets offs, uint32_t debug, uint32_t deferUpper,
                                const __grid_constant__ UploadBlock carry) {
    for (uint32_t seg = blockIdx.x; seg < carry.n; seg += gridDim.x) {
        const uint32_t* src = carry.data + carry.first[seg];
        uint32_t* dst = carry.dst[seg];
        for (uint32_t i = threadIdx.x; i < carry.count[seg]; i += blockDim.x) dst[i] = src[i];
    }
    pyramid_persistent_body(depthMap, w, h, pyr, fm, dim, levels, boxW, boxH, stages, stageFloats, topLevel, gridSync, offs, debug, deferUpper);
}

// footprint-min chain (see PyramidRec::footmin): one thread per output, levels firstLevel.. in one launch
__global__ void pyramid_footmin_kernel(const float* __restrict__ pyr, float* __restrict__ fm, uint32_t dim, uint32_t levels,
                                       uint32_t firstLevel, uint32_t total) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x + uint32_t(footmin_level_offset(dim, firstLevel));
    if (i >= total) return;
    uint32_t level = firstLevel;
    while (level + 1 < levels && i >= uint32_t(footmin_level_offset(dim, level + 1))) ++level;
    const uint32_t n = dim >> level, n1 = n + 1;
    const uint32_t local = i - uint32_t(footmin_level_offset(dim, level));
    const uint32_t fx = local % n1, fy = local / n1;
    const uint32_t x0 = fx ? fx - 1 : 0, x1 = fx < n ? fx : n - 1;
    const uint32_t y0 = fy ? fy - 1 : 0, y1 = fy < n ? fy : n - 1;
    const float* img = pyr + pyr_level_offset(dim, level);
    fm[i] = gmin(gmin(img[y0 * n + x0], img[y0 * n + x1]), gmin(img[y1 * n + x0], img[y1 * n + x1]));
}

int pyramid_refresh_footmin(rve_ctx* ctx, PyramidRec* p, uint32_t firstLevel) {
    if (p->footmin_valid) return 0;
    BufferRec* t = get_buffer(ctx, p->texels, "pyramid texels");
    BufferRec* f = get_buffer(ctx, p->footmin, "pyramid footprint-min chain");
    if (!t || !f) return RVE_ERR_BAD_HANDLE;
    const uint32_t total = uint32_t(footmin_level_offset(p->dim, p->levels));
    const uint32_t todo = firstLevel < p->levels ? total - uint32_t(footmin_level_offset(p->dim, firstLevel)) : 0u;
    if (todo) {
        pyramid_footmin_kernel<<<(todo + 255) / 256, 256, 0, ctx->stream>>>(static_cast<const float*>(t->ptr), static_cast<float*>(f->ptr),
                                                                           p->dim, p->levels, firstLevel, total);
        ctx->launches++;
        RVE_CUDA(cudaGetLastError());
    }
    p->footmin_valid = true;
    return 0;
}

int pyramid_before_buffer_read(rve_ctx* ctx, rve_buffer buf) {
    for (auto& p : ctx->pyramids)
        if (p.live && p.upperPending && (p.texels == buf || p.footmin == buf)) {
            const int rc = pyramid_finish_upper(ctx, &p);
            if (rc) return rc;
        }
    return 0;
}

void pyramid_note_buffer_write(rve_ctx* ctx, rve_buffer buf) {
    for (auto& p : ctx->pyramids)
        if (p.live && p.texels == buf) p.footmin_valid = false;
}

}  // namespace rve

using namespace rve;

extern "C" {

int rve_depth_pyramid_create(rve_ctx* ctx, uint32_t width, rve_pyramid* out) {
    if (!ctx || !out) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_REQUIRE(width >= 1, "pyramid width must be >= 1");
    // src/DepthPyramid.cpp:10-17
    const uint32_t w = width < 1024 ? width : 1024;
    uint32_t dim = 1, l = 0;
    while (dim * 2 <= w) {
        dim *= 2;
        ++l;
    }
    PyramidRec rec;
    rec.dim = dim;
    rec.levels = 1 + l;
    const size_t texels = (4ull * dim * dim - 1) / 3;
    int rc = rve_buffer_create(ctx, texels * sizeof(float), &rec.texels);  // zero = "far" in reverse-Z: nothing occluded
    if (rc) return rc;
    rc = rve_buffer_create(ctx, 128 * sizeof(uint32_t), &rec.counter);  // PyramidRec::counter
    if (rc) return rc;
    rc = rve_buffer_create(ctx, footmin_level_offset(dim, rec.levels) * sizeof(float), &rec.footmin);
    if (rc) return rc;
    rec.live = true;
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->pyramids.push_back(rec);
    *out = rve_pyramid(ctx->pyramids.size() - 1);
    return 0;
}

int rve_depth_pyramid_destroy(rve_ctx* ctx, rve_pyramid pyr) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    PyramidRec* p = get_pyramid(ctx, pyr);
    if (!p) return RVE_ERR_BAD_HANDLE;
    rve_buffer_destroy(ctx, p->texels);
    rve_buffer_destroy(ctx, p->counter);
    rve_buffer_destroy(ctx, p->footmin);
    p->live = false;
    return 0;
}

int rve_depth_pyramid_texels_changed(rve_ctx* ctx, rve_pyramid pyr) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    PyramidRec* p = get_pyramid(ctx, pyr);
    if (!p) return RVE_ERR_BAD_HANDLE;
    p->footmin_valid = false;
    return 0;
}

int rve_depth_pyramid_info(rve_ctx* ctx, rve_pyramid pyr, uint32_t* dim, uint32_t* num_levels, rve_buffer* texels) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    PyramidRec* p = get_pyramid(ctx, pyr);
    if (!p) return RVE_ERR_BAD_HANDLE;
    if (dim) *dim = p->dim;
    if (num_levels) *num_levels = p->levels;
    if (texels) *texels = p->texels;
    return 0;
}

int rve_depth_pyramid_footmin(rve_ctx* ctx, rve_pyramid pyr, rve_buffer* footmin) {
    if (!ctx || !footmin) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    PyramidRec* p = get_pyramid(ctx, pyr);
    if (!p) return RVE_ERR_BAD_HANDLE;
    int rc = pyramid_finish_upper(ctx, p);  // (a deferring build is always followed by its cull; kept for safety)
    if (rc) return rc;
    rc = pyramid_refresh_footmin(ctx, p, 0);  // texels written behind the library's back since the last build
    if (rc) return rc;
    *footmin = p->footmin;
    return 0;
}

int rve_depth_pyramid_build(rve_ctx* ctx, rve_pyramid pyr, rve_buffer depth, uint32_t width, uint32_t height) {
    return pyramid_build_impl(ctx, pyr, depth, width, height, false);
}

int rve_depth_pyramid_build_deferred(rve_ctx* ctx, rve_pyramid pyr, rve_buffer depth, uint32_t width, uint32_t height) {
    return pyramid_build_impl(ctx, pyr, depth, width, height, true);
}

}  // extern "C"

namespace rve {

static int launch_upper(rve_ctx* ctx, PyramidRec* p, float* pyrPtr, float* fmPtr, uint32_t topLevel, bool programmatic) {
    if (p->levels <= 1) return 0;
    const uint32_t lastFm = topLevel < p->levels ? topLevel : p->levels;
    const uint32_t entries = uint32_t(footmin_level_offset(p->dim, lastFm) - footmin_level_offset(p->dim, 1));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(1 + (entries + 2 * kPyrThreads - 1) / (2 * kPyrThreads));
    cfg.blockDim = dim3(kPyrThreads);
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = programmatic ? 1 : 0;
    static const uint32_t early = [] { const char* e = getenv("RVE_UPPER_TRIGGER"); return (e && atoi(e) != 0) ? 1u : 0u; }();
    const PyrOffsets offs = make_pyr_offsets(p->dim, p->levels);
    RVE_CUDA(cudaLaunchKernelEx(&cfg, pyramid_upper_kernel, pyrPtr, fmPtr, p->dim, p->levels, topLevel, programmatic ? early : 0u, offs));
    ctx->launches++;
    return 0;
}

// the part a deferring build left out, as a launch of its own (the consumer is not a cull that can take it along)
int pyramid_finish_upper(rve_ctx* ctx, PyramidRec* p) {
    if (!p->upperPending) return 0;
    BufferRec* t = get_buffer(ctx, p->texels, "pyramid texels");
    BufferRec* f = get_buffer(ctx, p->footmin, "pyramid footprint-min chain");
    if (!t || !f) return RVE_ERR_BAD_HANDLE;
    uint32_t topLevel = 1;
    while ((1u << topLevel) <= uint32_t(kPyrTile)) ++topLevel;
    const int rc = launch_upper(ctx, p, static_cast<float*>(t->ptr), static_cast<float*>(f->ptr), topLevel, false);
    if (rc) return rc;
    p->upperPending = false;
    return 0;
}

int pyramid_build_impl(rve_ctx* ctx, rve_pyramid pyr, rve_buffer depth, uint32_t width, uint32_t height, bool allowDefer, const UploadBlock* carry) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    RVE_PROFILE_FN_N("generatePyramid");
    DeviceGuard g(ctx->device);
    PyramidRec* p = get_pyramid(ctx, pyr);
    if (!p) return RVE_ERR_BAD_HANDLE;
    BufferRec* d = get_buffer(ctx, depth, "depth");
    if (!d) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE(width >= 1 && height >= 1, "depth buffer must be at least 1x1");
    RVE_REQUIRE(size_t(width) * height * sizeof(float) <= d->bytes, "depth buffer smaller than width*height floats");
    BufferRec* t = get_buffer(ctx, p->texels, "pyramid texels");
    BufferRec* f = get_buffer(ctx, p->footmin, "pyramid footprint-min chain");
    if (!t || !f) return RVE_ERR_BAD_HANDLE;
    const uint32_t tile = p->dim < uint32_t(kPyrTile) ? p->dim : uint32_t(kPyrTile);
    const uint32_t blocks = (p->dim / tile) * (p->dim / tile);
    uint32_t topLevel = 1;  // first level the tiles do not produce
    while ((1u << topLevel) <= tile) ++topLevel;
    float* pyrPtr = static_cast<float*>(t->ptr);
    float* fmPtr = static_cast<float*>(f->ptr);
    // Deferral (allowDefer: a cull follows at once): the build ends with the tile levels; the cull kernel derives the rest while
    // its first rounds are in flight. RVE_PYRAMID_DEFER=0 switches it off, RVE_PYRAMID_DEFER_TILES=tma picks the one-box-per-CTA
    // tile kernel instead of the persistent one (A/B runs).
    const char* deferEnv = getenv("RVE_PYRAMID_DEFER");
    const bool wantDefer = allowDefer && !(deferEnv && atoi(deferEnv) == 0) && tile == uint32_t(kPyrTile) && p->levels > 1;
    const char* deferTiles = getenv("RVE_PYRAMID_DEFER_TILES");
    const bool deferWithTma = wantDefer && deferTiles && deferTiles[0] == 't';
    BufferRec* sync = get_buffer(ctx, p->counter, "pyramid counters");
    if (!sync) return RVE_ERR_BAD_HANDLE;
    p->upperPending = false;  // whatever an earlier build left open is rewritten by this one (arguments are valid: it will launch)
    {
        ProfScope prof(ctx, RVE_PROF_PYRAMID);
        // TMA-fed tiles when the image qualifies (16-byte aligned rows, full 32-texel tiles, a source box that fits)
        bool tma = false;
        if (tile == uint32_t(kPyrTile) && (width & 3u) == 0 && (reinterpret_cast<uintptr_t>(d->ptr) & 15) == 0 && !getenv("RVE_PYRAMID_NO_TMA")) {
            if (p->mapBase != d->ptr || p->mapW != width || p->mapH != height) {
                // exact extent of the source rectangle a tile's 33 x 33 samples touch, over all tiles (same float math as the kernel)
                uint32_t bw = 0, bh = 0;
                for (uint32_t t0 = 0; t0 < p->dim; t0 += tile) {
                    const uint32_t last = t0 + tile < p->dim ? t0 + tile : p->dim - 1;
                    const uint32_t ex = footprint_start(last, p->dim, width) + 1, ey = footprint_start(last, p->dim, height) + 1;
                    const uint32_t x1 = ex < width ? ex : width - 1, y1 = ey < height ? ey : height - 1;
                    bw = std::max(bw, x1 - (footprint_start(t0, p->dim, width) & ~3u) + 1);  // from the aligned box start
                    bh = std::max(bh, y1 - footprint_start(t0, p->dim, height) + 1);
                }
                bw = (bw + 3u) & ~3u;  // rows of whole 16-byte units
                if ((bw & 31u) == 0) bw += 4;  // a row pitch of 32 banks would put a footprint's two rows on the same banks
                p->mapBase = nullptr;
                if (bw <= 256 && bh <= 256) {
                    const int rc = make_image_map(d->ptr, width, height, bw, bh, p->depthMap);
                    if (rc) return rc;
                    p->mapBase = d->ptr;
                    p->mapW = width, p->mapH = height, p->boxW = bw, p->boxH = bh;
                }
            }
            const size_t smem = size_t(p->boxW) * p->boxH * 4 + sizeof(float) * ((kPyrTile + 1) * (kPyrTile + 1) + (kPyrTile * kPyrTile) / 3 + 2);
            // The persistent kernel (a ring of >= 2 source boxes must fit a CTA). Deferring builds use it for the tile levels.
            // As a complete build in one launch (tiles + grid barrier + upper levels) it measures 20.8 us against 19.6 us
            // for the two launches below (profiles/r02f_pyramid_variants.txt: the grid barrier and the upper phase sit on the
            // critical path either way), so a plain rve_depth_pyramid_build takes it only with RVE_PYRAMID_PERSISTENT=1.
            const char* persEnv = getenv("RVE_PYRAMID_PERSISTENT");
            const bool persistentOff = wantDefer ? deferWithTma : !(persEnv && atoi(persEnv) != 0);
            if (p->mapBase == d->ptr && !persistentOff && p->levels > uint32_t(__builtin_ctz(kPyrTile))) {
                if (p->persFor != d->ptr || p->persBox != p->boxW * p->boxH) {
                    p->persGrid = 0;
                    p->persFor = d->ptr;
                    p->persBox = p->boxW * p->boxH;
                    const size_t stageBytes = (size_t(p->boxW) * p->boxH * 4 + 127) & ~size_t(127);
                    const char* ctasEnv = getenv("RVE_PYRAMID_PERS_CTAS");      // development: CTAs per SM (1 or 2)
                    const char* stagesEnv = getenv("RVE_PYRAMID_PERS_STAGES");  // development: ring depth
                    for (uint32_t perSm = ctasEnv ? uint32_t(atoi(ctasEnv)) : 2u; perSm >= 1 && !p->persGrid; --perSm) {
                        const size_t budget = std::min<size_t>(227 * 1024, (228 * 1024) / perSm - 1024) - 1024;  // static shared memory + slack
                        const uint32_t grid = std::min<uint32_t>(blocks, perSm * uint32_t(ctx->sm_count));
                        const uint32_t tilesPerCta = (blocks + grid - 1) / grid;
                        uint32_t st = uint32_t(std::min<size_t>(std::min<size_t>(budget / stageBytes, kPersMaxStages), tilesPerCta));
                        if (stagesEnv) st = std::min<uint32_t>(st, uint32_t(atoi(stagesEnv)));
                        if (st < 2 && st < tilesPerCta) continue;
                        const size_t bytes = std::max<size_t>(st * stageBytes, sizeof(float) * (kPyrTile * kPyrTile + (kPyrTile * kPyrTile) / 3 + 2));
                        if (bytes > ctx->pyramid_pers_smem_opt_in) {
                            RVE_CUDA(cudaFuncSetAttribute(pyramid_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes)));
                            RVE_CUDA(cudaFuncSetAttribute(pyramid_persistent_carry_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes)));
                            ctx->pyramid_pers_smem_opt_in = bytes;
                        }
                        int occ = 0;  // the grid barrier needs every CTA resident
                        RVE_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pyramid_persistent_kernel, kPersThreads, bytes));
                        if (occ < int(perSm)) continue;
                        p->persGrid = grid, p->persStages = st, p->persStageFloats = uint32_t(stageBytes / 4), p->persSmem = bytes;
                    }
                }
                if (p->persGrid) {
                    const PyrOffsets offs = make_pyr_offsets(p->dim, p->levels);
                    const char* dbgEnv = getenv("RVE_PYRAMID_DEBUG");  // development only: 1 no upper phase, 2 no grid barrier, 4 no tile work, 8 no loads
                    const uint32_t dbg = dbgEnv ? uint32_t(atoi(dbgEnv)) : 0u;
                    // a carried upload is written by this launch only when the consumer waits for the launch's completion at its
                    // entry - the cull of a deferring build does
                    if (carry && carry->n && wantDefer) {
                        pyramid_persistent_carry_kernel<<<p->persGrid, kPersThreads, p->persSmem, ctx->stream>>>(
                            *reinterpret_cast<const CUtensorMap*>(p->depthMap), width, height, pyrPtr, fmPtr, p->dim, p->levels, p->boxW, p->boxH,
                            p->persStages, p->persStageFloats, topLevel, static_cast<uint32_t*>(sync->ptr), offs, dbg, 1u, *carry);
                    } else {
                        if (carry) {
                            const int urc = upload_block(ctx, *carry);
                            if (urc) return urc;
                        }
                        pyramid_persistent_kernel<<<p->persGrid, kPersThreads, p->persSmem, ctx->stream>>>(
                            *reinterpret_cast<const CUtensorMap*>(p->depthMap), width, height, pyrPtr, fmPtr, p->dim, p->levels, p->boxW, p->boxH,
                            p->persStages, p->persStageFloats, topLevel, static_cast<uint32_t*>(sync->ptr), offs, dbg, wantDefer ? 1u : 0u);
                    }
                    ctx->launches++;
                    RVE_CUDA(cudaGetLastError());
                    p->footmin_valid = true;
                    p->upperPending = wantDefer;
                    return 0;
                }
            }
            if (carry) {  // no persistent launch took it along
                const int urc = upload_block(ctx, *carry);
                if (urc) return urc;
                carry = nullptr;
            }
            if (p->mapBase == d->ptr && smem <= 96 * 1024) {
                if (smem > ctx->pyramid_smem_opt_in) {
                    RVE_CUDA(cudaFuncSetAttribute(pyramid_build_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
                    ctx->pyramid_smem_opt_in = smem;
                }
                pyramid_build_tma_kernel<<<blocks, kPyrThreads, smem, ctx->stream>>>(*reinterpret_cast<const CUtensorMap*>(p->depthMap), width, height,
                                                                                   pyrPtr, fmPtr, p->dim, p->levels, p->boxW, p->boxH,
                                                                                   wantDefer ? static_cast<uint32_t*>(sync->ptr) : nullptr);
                tma = true;
            }
        }
        if (carry) {  // (images that do not qualify for TMA at all)
            const int urc = upload_block(ctx, *carry);
            if (urc) return urc;
        }
        if (!tma)
            pyramid_build_kernel<<<blocks, kPyrThreads, 0, ctx->stream>>>(static_cast<const float*>(d->ptr), width, height, pyrPtr, fmPtr, p->dim,
                                                                          p->levels, tile);
        ctx->launches++;
        RVE_CUDA(cudaGetLastError());
        if (tma && wantDefer) {
            p->upperPending = true;
        } else {
            const int rc = launch_upper(ctx, p, pyrPtr, fmPtr, topLevel, true);
            if (rc) return rc;
        }
    }
    p->footmin_valid = true;
    return 0;
}

}  // namespace rve
