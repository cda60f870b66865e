// scenegen.cu — procedural synthetic scenes on the device (C2 / C5 inputs are
// generated per rank from (seed, global index), so a 1 B-instance scene needs no 64 GB
// host->device copy) and the small multi-GPU helper kernels.
// The generator is bit-identical to the oracle's orc_gen_instances (tests check this).
#include "internal.h"

namespace rve {

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ float u01(uint64_t seed, uint64_t idx, uint32_t k) {
    const uint64_t h = splitmix64(seed ^ splitmix64(idx * 16 + k));
    return float(uint32_t(h >> 40)) * (1.0f / 16777216.0f);
}

__global__ void gen_instances_kernel(const RveSceneGenParams p, uint64_t first_index, uint32_t count,
                                     float4* __restrict__ model, uint32_t* __restrict__ layers,
                                     uint16_t* __restrict__ flags, uint32_t* __restrict__ ids) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint64_t g = first_index + i;
    const uint32_t id = p.id_base + i;
    const float px = (2.0f * u01(p.seed, g, 0) - 1.0f) * p.extent_xy;
    const float pz = (2.0f * u01(p.seed, g, 1) - 1.0f) * p.extent_xy;
    const float py = (2.0f * u01(p.seed, g, 2) - 1.0f) * p.extent_z;
    float qx = 2.0f * u01(p.seed, g, 3) - 1.0f;
    float qy = 2.0f * u01(p.seed, g, 4) - 1.0f;
    float qz = 2.0f * u01(p.seed, g, 5) - 1.0f;
    float qw = 2.0f * u01(p.seed, g, 6) - 1.0f;
    const float n2 = ((qx * qx + qy * qy) + qz * qz) + qw * qw;
    if (n2 < 1e-6f) {
        qx = qy = qz = 0.0f;
        qw = 1.0f;
    } else {
        const float n = sqrtf(n2);
        qx = qx / n;
        qy = qy / n;
        qz = qz / n;
        qw = qw / n;
    }
    const float s = p.scale_min + (p.scale_max - p.scale_min) * u01(p.seed, g, 7);
    const float xx = qx * qx, yy = qy * qy, zz = qz * qz;
    const float xy = qx * qy, xz = qx * qz, yz = qy * qz;
    const float wx = qw * qx, wy = qw * qy, wz = qw * qz;
    float4* m = model + uint64_t(id) * 4;
    m[0] = make_float4((1.0f - 2.0f * (yy + zz)) * s, (2.0f * (xy + wz)) * s, (2.0f * (xz - wy)) * s, 0.0f);
    m[1] = make_float4((2.0f * (xy - wz)) * s, (1.0f - 2.0f * (xx + zz)) * s, (2.0f * (yz + wx)) * s, 0.0f);
    m[2] = make_float4((2.0f * (xz + wy)) * s, (2.0f * (yz - wx)) * s, (1.0f - 2.0f * (xx + yy)) * s, 0.0f);
    m[3] = make_float4(px, py, pz, 1.0f);
    layers[id] = (u01(p.seed, g, 8) < p.frac_layer_zero) ? 0u : 0xFFFFFFFFu;
    flags[id] = (u01(p.seed, g, 9) < p.frac_no_cull_bits) ? uint16_t(0xFFFC) : uint16_t(0xFFFF);
    ids[i] = id;
}

__global__ void shard_offsets_kernel(const uint32_t* __restrict__ counts, uint32_t world, uint32_t slots,
                                     uint32_t* __restrict__ offsets, uint32_t* __restrict__ totals) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= slots) return;
    uint32_t run = 0;
    for (uint32_t r = 0; r < world; ++r) {
        offsets[uint64_t(r) * slots + s] = run;
        run += counts[uint64_t(r) * slots + s];
    }
    totals[s] = run;
}

__global__ void gather_counts_kernel(const RveIndirectIndexedCommand* __restrict__ cmds, uint32_t first, uint32_t n,
                                     uint32_t* __restrict__ out) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n) out[s] = cmds[first + s].instanceCount;
}

struct GatherMulti {
    const RveIndirectIndexedCommand* cmds[64];
    uint32_t first[64], n[64], outStart[64];
    uint32_t count, total;
};
__global__ void gather_counts_multi_kernel(const GatherMulti g, uint32_t* __restrict__ out) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= g.total) return;
    uint32_t b = 0;
    while (b + 1 < g.count && s >= g.outStart[b + 1]) ++b;
    out[s] = g.cmds[b][g.first[b] + (s - g.outStart[b])].instanceCount;
}

}  // namespace rve

using namespace rve;

extern "C" {

int rve_gen_instances(rve_ctx* ctx, const RveSceneGenParams* p, uint64_t first_index, uint32_t count,
                      rve_buffer model_matrices, rve_buffer render_layers, rve_buffer per_object_flags,
                      rve_buffer entity_ids) {
    if (!ctx || !p) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    if (!count) return 0;
    BufferRec* bm = get_buffer(ctx, model_matrices, "model matrices");
    BufferRec* bl = get_buffer(ctx, render_layers, "render layers");
    BufferRec* bf = get_buffer(ctx, per_object_flags, "per-object flags");
    BufferRec* bi = get_buffer(ctx, entity_ids, "entity ids");
    if (!bm || !bl || !bf || !bi) return RVE_ERR_BAD_HANDLE;
    const uint64_t maxId = uint64_t(p->id_base) + count;
    RVE_REQUIRE(maxId <= 0xFFFFFFFFull, "entity ids exceed 32 bits");
    RVE_REQUIRE(maxId * 64 <= bm->bytes && maxId * 4 <= bl->bytes && maxId * 2 <= bf->bytes && uint64_t(count) * 4 <= bi->bytes,
                "scene buffers too small for id_base + count");
    const uint32_t threads = 256;
    gen_instances_kernel<<<(count + threads - 1) / threads, threads, 0, ctx->stream>>>(
        *p, first_index, count, static_cast<float4*>(bm->ptr), static_cast<uint32_t*>(bl->ptr),
        static_cast<uint16_t*>(bf->ptr), static_cast<uint32_t*>(bi->ptr));
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

int rve_shard_offsets(rve_ctx* ctx, rve_buffer counts_all, uint32_t world, uint32_t num_slots, rve_buffer offsets_out,
                      rve_buffer totals_out) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    DeviceGuard g(ctx->device);
    if (!world || !num_slots) return 0;
    BufferRec* c = get_buffer(ctx, counts_all, "counts");
    BufferRec* o = get_buffer(ctx, offsets_out, "offsets");
    BufferRec* t = get_buffer(ctx, totals_out, "totals");
    if (!c || !o || !t) return RVE_ERR_BAD_HANDLE;
    const uint64_t n = uint64_t(world) * num_slots * 4;
    RVE_REQUIRE(n <= c->bytes && n <= o->bytes && uint64_t(num_slots) * 4 <= t->bytes, "shard offset buffers too small");
    const uint32_t threads = 128;
    shard_offsets_kernel<<<(num_slots + threads - 1) / threads, threads, 0, ctx->stream>>>(
        static_cast<const uint32_t*>(c->ptr), world, num_slots, static_cast<uint32_t*>(o->ptr), static_cast<uint32_t*>(t->ptr));
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

int rve_gather_instance_counts(rve_ctx* ctx, rve_buffer indirect, uint32_t first_slot, uint32_t num_slots,
                               rve_buffer counts_out, size_t counts_offset_bytes) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "ctx is null");
    DeviceGuard g(ctx->device);
    if (!num_slots) return 0;
    BufferRec* i = get_buffer(ctx, indirect, "indirect");
    BufferRec* o = get_buffer(ctx, counts_out, "counts");
    if (!i || !o) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE((uint64_t(first_slot) + num_slots) * 20 <= i->bytes, "indirect buffer too small");
    RVE_REQUIRE(counts_offset_bytes % 4 == 0 && counts_offset_bytes + uint64_t(num_slots) * 4 <= o->bytes, "counts buffer too small");
    const uint32_t threads = 128;
    gather_counts_kernel<<<(num_slots + threads - 1) / threads, threads, 0, ctx->stream>>>(
        static_cast<const RveIndirectIndexedCommand*>(i->ptr), first_slot, num_slots,
        reinterpret_cast<uint32_t*>(static_cast<char*>(o->ptr) + counts_offset_bytes));
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

int rve_gather_instance_counts_multi(rve_ctx* ctx, const rve_buffer* indirect, const uint32_t* first_slot, const uint32_t* num_slots,
                                     uint32_t n_buffers, rve_buffer counts_out, size_t counts_offset_bytes) {
    if (!ctx || !indirect || !first_slot || !num_slots) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    if (!n_buffers) return 0;
    RVE_REQUIRE(n_buffers <= 64, "at most 64 indirect buffers per call");
    BufferRec* o = get_buffer(ctx, counts_out, "counts");
    if (!o) return RVE_ERR_BAD_HANDLE;
    GatherMulti gm{};
    gm.count = n_buffers;
    uint32_t total = 0;
    for (uint32_t i = 0; i < n_buffers; ++i) {
        BufferRec* b = get_buffer(ctx, indirect[i], "indirect");
        if (!b) return RVE_ERR_BAD_HANDLE;
        RVE_REQUIRE((uint64_t(first_slot[i]) + num_slots[i]) * 20 <= b->bytes, "indirect buffer too small");
        gm.cmds[i] = static_cast<const RveIndirectIndexedCommand*>(b->ptr);
        gm.first[i] = first_slot[i];
        gm.n[i] = num_slots[i];
        gm.outStart[i] = total;
        total += num_slots[i];
    }
    gm.total = total;
    if (!total) return 0;
    RVE_REQUIRE(counts_offset_bytes % 4 == 0 && counts_offset_bytes + uint64_t(total) * 4 <= o->bytes, "counts buffer too small");
    gather_counts_multi_kernel<<<(total + 127) / 128, 128, 0, ctx->stream>>>(
        gm, reinterpret_cast<uint32_t*>(static_cast<char*>(o->ptr) + counts_offset_bytes));
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
