// tma_gather4_probe.cu — does sm_100a take cp.async.bulk.tensor.2d ... tile::gather4 (four arbitrary rows of a 2-D tensor in
// one TMA instruction) over the cull's transform array (rows of 16 f32 = 64 B), with which box shape, and how do the rows
// land in shared memory? (development probe for the gather path of cull.cu; no documentation on this box)
//   nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/g4probe tools/tma_gather4_probe.cu && /tmp/g4probe
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }

__global__ void probe(const __grid_constant__ CUtensorMap map, int4 rows, uint32_t expectBytes, float* out, uint32_t* status) {
    __shared__ __align__(1024) float tile[16 * 16];
    __shared__ uint64_t bar;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) tile[i] = -1.0f;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(expectBytes) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(
                         smem_u32(tile)),
                     "l"(&map), "r"(0), "r"(rows.x), "r"(rows.y), "r"(rows.z), "r"(rows.w), "r"(smem_u32(&bar))
                     : "memory");
        uint32_t done = 0;
        for (int it = 0; it < 2000000 && !done; ++it)
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
        *status = done;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 256; i += blockDim.x) out[i] = tile[i];
}

typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                        CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q);
    Enc enc = (Enc)f;
    const uint64_t N = 200000;
    std::vector<float> h(N * 16);
    for (uint64_t r = 0; r < N; ++r)
        for (int c = 0; c < 16; ++c) h[r * 16 + c] = float(r) + c * 0.01f;
    struct Case { uint32_t boxRows; CUtensorMapSwizzle sw; const char* name; uint32_t bytes; };
    std::vector<Case> cases = {{1, CU_TENSOR_MAP_SWIZZLE_NONE, "box 16x1, no swizzle", 256}, {1, CU_TENSOR_MAP_SWIZZLE_64B, "box 16x1, 64-B swizzle", 256},
                               {4, CU_TENSOR_MAP_SWIZZLE_NONE, "box 16x4, no swizzle", 256}};
    for (auto c : cases) {
        float *mat, *out;
        uint32_t* st;
        cudaMalloc(&mat, N * 64);
        cudaMemcpy(mat, h.data(), N * 64, cudaMemcpyHostToDevice);
        cudaMalloc(&out, 1024);
        cudaMalloc(&st, 4);
        cudaMemset(st, 0, 4);
        CUtensorMap map;
        cuuint64_t gdim[2] = {16, N}, gstride[1] = {64};
        cuuint32_t box[2] = {16, c.boxRows}, estr[2] = {1, 1};
        CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, mat, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, c.sw,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        probe<<<1, 128>>>(map, make_int4(5, 100000, 7, 199999), c.bytes, out, st);
        cudaError_t e = cudaDeviceSynchronize();
        float o[256];
        uint32_t done = 0;
        if (e == cudaSuccess) {
            cudaMemcpy(o, out, 1024, cudaMemcpyDeviceToHost);
            cudaMemcpy(&done, st, 4, cudaMemcpyDeviceToHost);
        }
        printf("%-26s encode %d -> %s, barrier %s\n", c.name, int(r), cudaGetErrorString(e), done ? "completed" : "TIMED OUT");
        if (e == cudaSuccess)
            for (int row = 0; row < 5; ++row) {
                printf("   smem floats %3d..%3d:", row * 16, row * 16 + 15);
                for (int k = 0; k < 16; ++k) printf(" %.2f", o[row * 16 + k]);
                printf("\n");
            }
        if (e != cudaSuccess) cudaDeviceReset();
    }
    return 0;
}
