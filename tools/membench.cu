// membench.cu — read-only streaming microbenchmarks on B200 (development tool, not part of the library).
// Answers one design question for the cull kernel: which load mechanism / occupancy sustains HBM read
// bandwidth for 2-KB-per-warp tiles — LDG.128 into registers, per-lane cp.async (LDGSTS) pipelines, or
// cp.async.bulk (TMA 1D) with mbarrier completion.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/membench tools/membench.cu && /tmp/membench
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <vector>

#define CK(x)                                                                      \
    do {                                                                           \
        cudaError_t e_ = (x);                                                      \
        if (e_ != cudaSuccess) {                                                   \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            return 1;                                                              \
        }                                                                          \
    } while (0)

__device__ __forceinline__ float4 ldg_stream(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }

// ---- A: LDG.128, U independent loads per thread in flight ----
template <int U>
__global__ void k_ldg(const float4* __restrict__ src, size_t n4, float* out) {
    float acc = 0.f;
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < n4; i += U * stride) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = ldg_stream(src + i + u * stride);
#pragma unroll
        for (int u = 0; u < U; ++u) acc += v[u].x + v[u].y + v[u].z + v[u].w;
    }
    for (; i < n4; i += stride) {
        float4 v = ldg_stream(src + i);
        acc += v.x + v.y + v.z + v.w;
    }
    if (acc == 123.456f) *out = acc;
}

// ---- B: per-warp cp.async pipeline, 2 KB per stage ----
template <int S>
__global__ void k_cpasync(const float4* __restrict__ src, size_t n4, float* out) {
    extern __shared__ __align__(128) unsigned char raw[];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    float4* st = reinterpret_cast<float4*>(raw) + size_t(warp) * S * 128;
    const size_t rounds = n4 / 128;
    const size_t W = size_t(gridDim.x) * wpb;
    const size_t first = size_t(blockIdx.x) * wpb + warp;
    if (first >= rounds) return;
    const size_t iters = (rounds - first + W - 1) / W;
    auto issue = [&](size_t k, int s) {
        const float4* g = src + (first + k * W) * 128;
#pragma unroll
        for (int m = 0; m < 4; ++m)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(st + s * 128 + m * 32 + lane)), "l"(g + m * 32 + lane) : "memory");
    };
    for (int s = 0; s < S - 1; ++s) {
        if (size_t(s) < iters) issue(s, s);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    float acc = 0.f;
    int si = S - 1, su = 0;
    for (size_t k = 0; k < iters; ++k) {
        if (k + S - 1 < iters) issue(k + S - 1, si);
        asm volatile("cp.async.commit_group;" ::: "memory");
        si = (si + 1 == S) ? 0 : si + 1;
        asm volatile("cp.async.wait_group %0;" ::"n"(S - 1) : "memory");
        __syncwarp();
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            float4 v = st[su * 128 + m * 32 + lane];
            acc += v.x + v.y + v.z + v.w;
        }
        __syncwarp();
        su = (su + 1 == S) ? 0 : su + 1;
    }
    if (acc == 123.456f) *out = acc;
}

// ---- C: per-warp cp.async.bulk pipeline (TMA 1D), CHUNK bytes per stage, mbarrier completion ----
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
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(b)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(b))
                 : "memory");
}

template <int S, int CHUNK>
__global__ void k_bulk(const float4* __restrict__ src, size_t n4, float* out) {
    extern __shared__ __align__(128) unsigned char raw[];
    constexpr int F4 = CHUNK / 16;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    float4* st = reinterpret_cast<float4*>(raw) + size_t(warp) * S * F4;
    uint64_t* bars = reinterpret_cast<uint64_t*>(raw + size_t(wpb) * S * CHUNK) + warp * S;
    if (lane == 0)
        for (int s = 0; s < S; ++s) mbar_init(bars + s, 1);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    const size_t rounds = n4 / F4;
    const size_t W = size_t(gridDim.x) * wpb;
    const size_t first = size_t(blockIdx.x) * wpb + warp;
    if (first >= rounds) return;
    const size_t iters = (rounds - first + W - 1) / W;
    auto issue = [&](size_t k, int s) {
        if (lane == 0) {
            mbar_expect_tx(bars + s, CHUNK);
            bulk_g2s(st + s * F4, src + (first + k * W) * F4, CHUNK, bars + s);
        }
    };
    for (int s = 0; s < S - 1; ++s)
        if (size_t(s) < iters) issue(s, s);
    float acc = 0.f;
    int si = S - 1, su = 0;
    uint32_t phase = 0;
    for (size_t k = 0; k < iters; ++k) {
        if (k + S - 1 < iters) issue(k + S - 1, si);
        si = (si + 1 == S) ? 0 : si + 1;
        mbar_wait(bars + su, (phase >> su) & 1);
        phase ^= 1u << su;
#pragma unroll
        for (int m = 0; m < F4 / 32; ++m) {
            float4 v = st[su * F4 + m * 32 + lane];
            acc += v.x + v.y + v.z + v.w;
        }
        __syncwarp();
        su = (su + 1 == S) ? 0 : su + 1;
    }
    if (acc == 123.456f) *out = acc;
}

template <typename F>
static float time_ms(F&& launch, int reps = 10) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    for (int i = 0; i < 3; ++i) launch();
    cudaEventRecord(a);
    for (int i = 0; i < reps; ++i) launch();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    return ms / reps;
}

int main() {
    const size_t bytes = size_t(768) << 20;
    const size_t n4 = bytes / 16;
    float4* src;
    float* out;
    CK(cudaMalloc(&src, bytes));
    CK(cudaMalloc(&out, 4));
    CK(cudaMemset(src, 0, bytes));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sm = prop.multiProcessorCount;
    printf("device %s, %d SMs, buffer %zu MiB\n", prop.name, sm, bytes >> 20);
    auto report = [&](const char* name, float ms) { printf("%-44s %8.3f ms  %8.1f GB/s\n", name, ms, bytes / ms / 1e6); };

    char name[128];
    for (int bps : {2, 4, 8}) {
        snprintf(name, sizeof name, "ldg U=4  256thr x %d blocks/SM", bps);
        report(name, time_ms([&] { k_ldg<4><<<sm * bps, 256>>>(src, n4, out); }));
        snprintf(name, sizeof name, "ldg U=8  256thr x %d blocks/SM", bps);
        report(name, time_ms([&] { k_ldg<8><<<sm * bps, 256>>>(src, n4, out); }));
    }
    report("ldg U=4  1024thr x 2 blocks/SM", time_ms([&] { k_ldg<4><<<sm * 2, 1024>>>(src, n4, out); }));
    CK(cudaGetLastError());

    auto run_cp = [&](auto kern, int S, int wpb, int bps) {
        const size_t smem = size_t(wpb) * S * 2048;
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        snprintf(name, sizeof name, "cp.async S=%d %d warps/blk x %d blk/SM (%d w/SM)", S, wpb, bps, wpb * bps);
        report(name, time_ms([&] { kern<<<sm * bps, wpb * 32, smem>>>(src, n4, out); }));
    };
    for (int bps : {3, 5, 8}) run_cp(k_cpasync<3>, 3, 4, bps);
    for (int bps : {2, 4}) run_cp(k_cpasync<3>, 3, 8, bps);
    for (int bps : {3, 5, 8}) run_cp(k_cpasync<4>, 4, 4, bps);
    for (int bps : {3, 6}) run_cp(k_cpasync<6>, 6, 4, bps);
    run_cp(k_cpasync<2>, 2, 4, 8);
    run_cp(k_cpasync<2>, 2, 4, 12);
    CK(cudaGetLastError());

    auto run_bulk = [&](auto kern, int S, int chunk, int wpb, int bps) {
        const size_t smem = size_t(wpb) * S * chunk + size_t(wpb) * S * 8;
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        snprintf(name, sizeof name, "bulk S=%d chunk=%d %d w/blk x %d blk/SM (%d w/SM)", S, chunk, wpb, bps, wpb * bps);
        report(name, time_ms([&] { kern<<<sm * bps, wpb * 32, smem>>>(src, n4, out); }));
    };
    for (int bps : {2, 4, 8}) run_bulk(k_bulk<3, 2048>, 3, 2048, 4, bps);
    for (int bps : {2, 4, 6}) run_bulk(k_bulk<4, 2048>, 4, 2048, 4, bps);
    for (int bps : {1, 2, 4}) run_bulk(k_bulk<3, 4096>, 3, 4096, 4, bps);
    for (int bps : {1, 2}) run_bulk(k_bulk<3, 8192>, 3, 8192, 4, bps);
    for (int bps : {1, 2}) run_bulk(k_bulk<6, 2048>, 6, 2048, 4, bps);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    return 0;
}
