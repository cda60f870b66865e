// frame_cull.cu — cullTheRenderData (src/RenderEngine_Draw.cpp:1221-1407) and the clustered-light
// prologue of renderFromPerspective (:928-974), restated over the rve_* C-ABI. Host C++ only.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "frame_internal.h"
#include "internal.h"

using namespace rve;
using namespace rvs;

extern "C" {

}  // extern "C"

// cullTheRenderData for `nviews` views at once (nviews == 0: the reference's single pass through rve_cull). With
// views, every per-material output buffer holds nviews consecutive copies of the single-view layout.
struct PyramidSource {  // generatePyramid to run between the table upload and the cull (single-view pass only)
    rve_buffer depth = RVE_BUFFER_INVALID;
    uint32_t width = 0, height = 0;
};

static int cull_static_impl(rvs_world* w, uint32_t nviews, const rvs_view* views, const float view_proj[16], const float cam_pos[3],
                            uint32_t camera_layers, uint32_t shadow_pass, uint32_t pass_mask, const rve_pyramid* pyramids,
                            const PyramidSource* build = nullptr) {
    rve_ctx* ctx = w->ctx;
    const uint32_t copies = nviews ? nviews : 1;
    // :1223-1229
    RveCullingUBO globalCubo{};
    std::memcpy(globalCubo.viewProj, nviews ? views[0].view_proj : view_proj, 64);
    std::memcpy(globalCubo.camPos, nviews ? views[0].cam_pos : cam_pos, 12);
    globalCubo.numCubos = 0;
    globalCubo.cameraRenderLayers = nviews ? views[0].camera_layers : camera_layers;
    globalCubo.singleInstanceModeAndShadowMode = (nviews ? views[0].shadow_pass : shadow_pass) ? RVE_CULL_SHADOW_MODE : 0u;

    // entity lists held in sparse sets: EncodeSync first (the reference syncs its BufferedVRAM structures ahead of the encode),
    // then the command reads the set's private buffer and DenseSize()
    {
        // (all sets in one rvs_vectors_encode_sync_batch: lists with small changes share one staging copy and one scatter launch)
        std::vector<rvs_vector*> vecs;
        std::vector<decltype(&w->materials[0].commands[0])> owners;
        for (auto& m : w->materials)
            for (auto& c : m.commands)
                if (c.set) {
                    vecs.push_back(rvs_sparse_set_vector(c.set));
                    owners.push_back(&c);
                }
        if (!vecs.empty()) {
            std::vector<rvs_sync_stats> sts(vecs.size());
            const int src = rvs_vectors_encode_sync_batch(vecs.data(), uint32_t(vecs.size()), 256, sts.data());
            if (src) return src;
            for (size_t i = 0; i < vecs.size(); ++i) {
                auto& c = *owners[i];
                // the list changed: by the set's version, not by this sync's statistics - the application may have synced the set itself
                // (rvs_vectors_encode_sync_batch over all of a world's structures) before this call
                if (sts[i].ranges || sparse_set_version(c.set) != c.sortedVersion) c.sortedValid = false;
                c.entityIds = rvs_vector_private_buffer(vecs[i]);
                c.count = rvs_sparse_set_dense_size(c.set);
            }
        }
    }
    // RVS_SORT_ENTITY_LISTS: the cull reads an ascending copy of every list (its result does not depend on list order;
    // a list shuffled by swap-remove erases costs the cull its tile path), rebuilt only when the list changed
    if (w->flags & RVS_SORT_ENTITY_LISTS) {
        const uint64_t idLimit = rve_buffer_size(w->ctx, w->model) / 64;
        const uint32_t idLimit32 = uint32_t(std::min<uint64_t>(idLimit, 0xFFFFFFFFull));
        // every list that changed goes into ONE batch (two launches per 16 lists): caller-owned lists through the context's
        // scratch bitmaps (rve_sort_entity_ids_batch), sparse-set-backed lists through a bitmap of their own, to which a small
        // change is applied as a delta (the ids the set reports as entered / left) instead of rebuilding it from the list
        std::vector<rve_buffer> srcs, dsts;
        std::vector<uint32_t> counts;
        std::vector<decltype(&w->materials[0].commands[0])> stale;
        std::vector<SortedListUpdate> owned;
        std::vector<std::vector<uint32_t>> deltas;  // entered / left ids of the owned lists, alive until the batch is enqueued
        for (auto& m : w->materials)
            for (auto& c : m.commands) {
                if (!c.count) continue;
                if (c.sortedValid && c.sortedOf == c.entityIds && c.sortedCount == c.count) continue;
                int src = realloc_buffer(w, &c.sortedIds, &c.sortedBytes, c.count, sizeof(uint32_t));
                if (src) return src;
                stale.push_back(&c);
                if (!c.set) {
                    srcs.push_back(c.entityIds);
                    dsts.push_back(c.sortedIds);
                    counts.push_back(c.count);
                    continue;
                }
                const size_t before = c.sortBitmapBytes;
                src = realloc_buffer(w, &c.sortBitmap, &c.sortBitmapBytes, sort_region_bytes(idLimit32), 1);
                if (src) return src;
                SortedListUpdate u{c.entityIds, c.count, c.sortedIds, c.sortBitmap, nullptr, 0, nullptr, 0, true};
                deltas.emplace_back();
                deltas.emplace_back();
                std::vector<uint32_t>&entered = deltas[deltas.size() - 2], &left = deltas.back();
                static std::atomic<uint64_t> nextOwner{1};
                if (!c.sortOwnerId) c.sortOwnerId = nextOwner.fetch_add(1);
                const bool tracked = sparse_set_take_changes(c.set, c.sortOwnerId, &entered, &left);
                if (tracked && c.bitmapValid && c.bitmapIdLimit == idLimit32 && c.sortBitmapBytes == before &&
                    entered.size() + left.size() <= size_t(c.count) / 4) {
                    u.full = false;
                    u.setIds = entered.data(), u.nSet = uint32_t(entered.size());
                    u.clearIds = left.data(), u.nClear = uint32_t(left.size());
                } else if (!tracked) {
                    sparse_set_track_changes(c.set, c.sortOwnerId);  // from this full build on the set reports its changes to this command
                }
                c.bitmapValid = false;  // until the batch is enqueued
                owned.push_back(u);
            }
        if (!srcs.empty()) {
            const int src = rve_sort_entity_ids_batch(w->ctx, uint32_t(srcs.size()), srcs.data(), counts.data(), idLimit32, dsts.data());
            if (src) return src;
        }
        if (!owned.empty()) {
            const int src = sort_lists_persistent(w->ctx, owned.data(), uint32_t(owned.size()), idLimit32);
            if (src) return src;
        }
        w->sortFullRebuilds += uint32_t(srcs.size());
        for (const SortedListUpdate& u : owned) (u.full ? w->sortFullRebuilds : w->sortDeltaUpdates) += 1;
        for (auto* c : stale) {
            c->sortedOf = c->entityIds;
            c->sortedCount = c->count;
            c->sortedValid = true;
            if (c->set) c->bitmapValid = true, c->bitmapIdLimit = idLimit32, c->sortedVersion = sparse_set_version(c->set);
        }
    }
    // :1231-1244 — numCubos counts LODs (not commands); totalEntities sizes the dispatch
    uint32_t totalEntities = 0;
    for (auto& m : w->materials) {
        if (!(m.kind & pass_mask)) continue;
        for (auto& c : m.commands) {
            globalCubo.numCubos += uint32_t(w->meshes[c.mesh].lods.size());
            totalEntities += c.count;
        }
    }
    w->lastNumCubos = globalCubo.numCubos;
    w->lastTotalEntities = totalEntities;
    if (globalCubo.numCubos == 0)  // :1245-1247 (the pyramid of the fused encode is still built)
        return build ? rve_depth_pyramid_build(ctx, pyramids[0], build->depth, build->width, build->height) : 0;

    // :1249-1251 cuboBuffer + staging (entries past the written ones are zero here; stale in the reference)
    int rc = realloc_buffer(w, &w->cuboBuffer, &w->cuboBytes, globalCubo.numCubos, sizeof(RveCullingUBOinstance));
    if (rc) return rc;
    std::vector<RveCullingUBOinstance> cuboStaging(globalCubo.numCubos);
    std::memset(cuboStaging.data(), 0, cuboStaging.size() * sizeof(RveCullingUBOinstance));

    uint32_t cuboIdx = 0;
    // every table of the pass goes out in one stream operation (rve_buffer_upload_batch) at the end
    std::vector<std::vector<RveIndirectIndexedCommand>> indirectTables;
    std::vector<rve_upload_segment> uploads;
    std::vector<RveViewStride> strides(globalCubo.numCubos);
    std::vector<const float*> lodHost(globalCubo.numCubos, nullptr);  // per written CUBO: its mesh's LOD distances
    for (auto& m : w->materials) {
        if (!(m.kind & pass_mask)) continue;
        // :1265-1286 prepass
        uint32_t numLODs = 0, numEntities = 0;
        for (auto& c : m.commands) {
            numLODs += uint32_t(w->meshes[c.mesh].lods.size());
            numEntities += c.count;
        }
        m.numLODs = numLODs;
        m.numEntities = numEntities;
        // :1288-1291
        const size_t cullingbufferTotalSlots = size_t(numEntities) * numLODs;
        RVE_REQUIRE(cullingbufferTotalSlots * copies <= 0xFFFFFFFFull, "culling buffer exceeds 2^32 slots (u32 indices in the CUBO layout)");
        rc = realloc_buffer(w, &m.cullingBuffer, &m.cullingBytes, cullingbufferTotalSlots * copies, sizeof(uint32_t));
        if (rc) return rc;
        rc = realloc_buffer(w, &m.indirectBuffer, &m.indirectBytes, size_t(numLODs) * copies, sizeof(RveIndirectIndexedCommand));
        if (rc) return rc;
        if (numLODs == 0) continue;

        // :1295-1320 initial populate of the drawcall buffer: one command per mesh per LOD
        indirectTables.emplace_back(numLODs, RveIndirectIndexedCommand{0, 0, 0, 0, 0});
        std::vector<RveIndirectIndexedCommand>& indirectStaging = indirectTables.back();
        {
            uint32_t meshID = 0, baseInstance = 0, lodBase = 0;
            for (auto& c : m.commands) {
                const MeshCollection& mesh = w->meshes[c.mesh];
                for (uint32_t lodID = 0; lodID < mesh.lods.size(); lodID++) {
                    const LODMesh& lm = mesh.lods[lodID];
                    const RveIndirectIndexedCommand initData{lm.indexCount, 0, lm.indexStart, lm.baseVertex, baseInstance};
                    baseInstance += c.count;
                    // reference writes slot (meshID + lodID) (:1314); the shader reads cubo.indirectBufferOffset + lod
                    const uint32_t slot = (w->flags & RVS_QUIRK_INDIRECT_SLOT) ? (meshID + lodID) : (lodBase + lodID);
                    if (slot < numLODs) indirectStaging[slot] = initData;
                }
                lodBase += uint32_t(mesh.lods.size());
                meshID++;
            }
        }
        // :1321-1329 CopyBufferToBuffer(indirectStagingBuffer -> indirectBuffer)
        for (uint32_t v = 1; v < copies; ++v)  // the same initial commands for every view
            for (uint32_t k = 0; k < numLODs; ++k) indirectStaging.push_back(indirectStaging[k]);
        uploads.push_back({m.indirectBuffer, 0, indirectStaging.data(), indirectStaging.size() * sizeof(RveIndirectIndexedCommand)});

        // :1334-1362 one CUBO per command; offsets restart per material
        RveCullingUBOinstance cubo{};
        cubo.indirectBufferOffset = 0;
        for (auto& c : m.commands) {
            const MeshCollection& mesh = w->meshes[c.mesh];
            const uint32_t lodsForThisMesh = uint32_t(mesh.lods.size());
            cubo.numObjects = c.count;
            cubo.radius = mesh.radius;
            cubo.numLODs = lodsForThisMesh;
            cubo.idOutputBufferBindlessHandle = rve_buffer_bindless_index(ctx, m.cullingBuffer);
            cubo.indirectOutputBufferBindlessHandle = rve_buffer_bindless_index(ctx, m.indirectBuffer);
            cubo.lodDistanceBufferBindlessHandle = rve_buffer_bindless_index(ctx, mesh.lodDistances);
            cubo.entityIDInputBufferBindlessHandle =
                rve_buffer_bindless_index(ctx, ((w->flags & RVS_SORT_ENTITY_LISTS) && c.sortedValid && c.count) ? c.sortedIds : c.entityIds);
            strides[cuboIdx] = RveViewStride{numLODs, uint32_t(cullingbufferTotalSlots)};
            lodHost[cuboIdx] = mesh.lodDistancesHost.data();
            cuboStaging[cuboIdx++] = cubo;
            cubo.indirectBufferOffset += lodsForThisMesh;
            cubo.cullingBufferOffset += lodsForThisMesh * c.count;
        }
    }
    // Below the measured crossover the view-batched kernel loses to one launch per view (profiles/r01r_views.jsonl: 4 views
    // 0.65x, 16 views 1.21x; its fixed cost is ~0.3-0.5 ms): small batches run pass by pass - the reference's own schedule -
    // into the same per-view layout, each view with its CUBO offsets displaced by its copy of the output buffers.
    const char* batchEnv = getenv("RVS_VIEW_BATCH_MIN");  // development / tests: force either path
    const uint32_t kBatchMin = batchEnv ? uint32_t(atoi(batchEnv)) : 8u;
    const bool passByPass = nviews != 0 && nviews < kBatchMin;
    const uint32_t groups = uint32_t(std::ceil(totalEntities / 64.f));
    const uint32_t passes = passByPass ? nviews : 1u;
    uint32_t cullFlags = 0;
    UploadBlock carryBlk;  // tables of a fused encode, written by its pyramid launch
    carryBlk.n = 0;
    bool haveCarry = false;
    for (uint32_t pass = 0; pass < passes; ++pass) {
        std::vector<RveCullingUBOinstance> cuboPass = cuboStaging;
        if (passByPass)
            for (uint32_t c = 0; c < cuboIdx; ++c) {
                cuboPass[c].indirectBufferOffset += pass * strides[c].indirect_stride;
                cuboPass[c].cullingBufferOffset += pass * strides[c].culling_stride;
            }
        std::vector<rve_upload_segment> ups;
        if (pass == 0) ups = uploads;  // the indirect-init tables (all views' copies) go up once
        // :1367-1375 copy to the cubo buffer
        ups.push_back({w->cuboBuffer, 0, cuboPass.data(), cuboPass.size() * sizeof(RveCullingUBOinstance)});
        // The CUBO table was built right here, so the tables the cull derives from it (prefix sums, resolved handles, squared
        // LOD thresholds: cubo_prepare_kernel) are built on the host too and ride in the same parameter-block upload - one
        // launch and one dependent kernel less per pass. Entries past the written ones (numCubos counts LODs) are zero:
        // numObjects 0, never selected. Falls back to the plain path when the block would not fit.
        cullFlags = 0;
        {
            std::vector<RawSegment> raw;
            HostCuboStaging hostTables;
            bool ok = true;
            for (const rve_upload_segment& u : ups) {
                char* base = static_cast<char*>(rve_buffer_device_ptr(ctx, u.dst));
                if (!base || u.dst_offset + u.bytes > rve_buffer_size(ctx, u.dst)) ok = false;
                raw.push_back(RawSegment{base + u.dst_offset, u.src, u.bytes});
            }
            static const float kNoLods[4] = {0.f, 0.f, 0.f, 0.f};
            for (auto& lp : lodHost)
                if (!lp) lp = kNoLods;
            if (ok && cull_host_tables(ctx, cuboPass.data(), globalCubo.numCubos, lodHost.data(), &hostTables, &raw) == 0) {
                if (build && !nviews && !passByPass && groups > 0) {
                    // fused encode: the pyramid launch below carries the block (pyramid_build_impl) - no launch of its own
                    rc = make_upload_block(raw.data(), uint32_t(raw.size()), &carryBlk);
                    haveCarry = rc == 0;
                } else {
                    rc = upload_batch_raw(ctx, raw.data(), uint32_t(raw.size()));
                }
                if (rc == 0) cullFlags = kCullTablesPrepared;
                else if (rc != RVE_ERR_UNSUPPORTED) return rc;
            }
        }
        if (!cullFlags) {
            rc = rve_buffer_upload_batch(ctx, ups.data(), uint32_t(ups.size()));
            if (rc) return rc;
        }
        if (passByPass) {
            RveCullingUBO u = globalCubo;
            std::memcpy(u.viewProj, views[pass].view_proj, 64);
            std::memcpy(u.camPos, views[pass].cam_pos, 12);
            u.cameraRenderLayers = views[pass].camera_layers;
            u.singleInstanceModeAndShadowMode = views[pass].shadow_pass ? RVE_CULL_SHADOW_MODE : 0u;
            rc = cull_launch(ctx, &u, 0, &pyramids[pass], nullptr, w->cuboBuffer, w->model, w->layers, w->flagsBuf, groups, cullFlags);
            if (rc) return rc;
        }
    }
    if (passByPass) return 0;
    // :1377-1403 bind + DispatchCompute(ceil(totalEntities / 64), 1, 1, 64, 1, 1)
    if (build) {
        // upload -> pyramid build -> pyramid upper -> cull: the build is a plain launch (it starts once the tables are
        // complete), the cull a programmatic dependent of the pyramid kernels that needs them only for its Hi-Z batches
        // single view with work to do: the cull kernel below finishes the pyramid's upper levels in its start-up
        rc = pyramid_build_impl(ctx, pyramids[0], build->depth, build->width, build->height, !nviews && groups > 0, haveCarry ? &carryBlk : nullptr);
        if (rc) return rc;
        static const bool late = [] { const char* e = getenv("RVE_CULL_LATE_WAIT"); return !e || atoi(e) != 0; }();
        if ((cullFlags & kCullTablesPrepared) && late) cullFlags |= kCullLateWait;
    }
    if (!nviews) return cull_launch(ctx, &globalCubo, 0, &pyramids[0], nullptr, w->cuboBuffer, w->model, w->layers, w->flagsBuf, groups, cullFlags);
    std::vector<RveCullingUBO> ubos(nviews, globalCubo);
    for (uint32_t v = 0; v < nviews; ++v) {
        std::memcpy(ubos[v].viewProj, views[v].view_proj, 64);
        std::memcpy(ubos[v].camPos, views[v].cam_pos, 12);
        ubos[v].cameraRenderLayers = views[v].camera_layers;
        ubos[v].singleInstanceModeAndShadowMode = views[v].shadow_pass ? RVE_CULL_SHADOW_MODE : 0u;
    }
    RVE_REQUIRE(nviews <= 128, "1..128 views per batch");
    return cull_launch(ctx, ubos.data(), nviews, pyramids, strides.data(), w->cuboBuffer, w->model, w->layers, w->flagsBuf, groups, cullFlags);
}

extern "C" {

int rvs_cull_static(rvs_world* w, const float view_proj[16], const float cam_pos[3], uint32_t camera_layers,
                    uint32_t shadow_pass, uint32_t pass_mask, rve_pyramid pyramid) {
    if (!w || !view_proj || !cam_pos) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    return cull_static_impl(w, 0, nullptr, view_proj, cam_pos, camera_layers, shadow_pass, pass_mask, &pyramid);
}

int rvs_cull_static_with_pyramid(rvs_world* w, const float view_proj[16], const float cam_pos[3], uint32_t camera_layers, uint32_t shadow_pass,
                                 uint32_t pass_mask, rve_pyramid pyramid, rve_buffer depth, uint32_t depth_width, uint32_t depth_height) {
    if (!w || !view_proj || !cam_pos) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    const PyramidSource src{depth, depth_width, depth_height};
    return cull_static_impl(w, 0, nullptr, view_proj, cam_pos, camera_layers, shadow_pass, pass_mask, &pyramid, &src);
}

int rvs_cull_static_views(rvs_world* w, uint32_t num_views, const rvs_view* views, const rve_pyramid* pyramids, uint32_t pass_mask) {
    if (!w || !views || !pyramids) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_REQUIRE(num_views >= 1 && num_views <= 128, "1..128 views per batch");
    return cull_static_impl(w, num_views, views, nullptr, nullptr, 0, 0, pass_mask, pyramids);
}

// the clustered-light prologue into a given cluster buffer (the world's single lightClusterBuffer, or one per camera
// of a frame)
static int bin_lights_into(rvs_world* w, rve_buffer* buf, size_t* bytes, const float inv_proj[16], const float view_mat[16],
                           const uint32_t grid_size[3], const uint32_t screen_dim[2], float z_near, float z_far, rve_buffer point_lights,
                           uint32_t point_light_count, rve_buffer spot_lights, uint32_t spot_light_count) {
    RVE_PROFILE_FN_N("RenderFromPerspective: cluster lights");
    const uint64_t n = uint64_t(grid_size[0]) * grid_size[1] * grid_size[2];
    RVE_REQUIRE(n > 0 && n < 0x7FFFFFFFull, "bad cluster grid");
    // lightClusterBuffer: numClusters * sizeof(Cluster) (src/RenderEngine.cpp:839-845)
    if (*bytes != n * sizeof(RveCluster)) {
        int rc = realloc_buffer(w, buf, bytes, n, sizeof(RveCluster));
        if (rc) return rc;
    }
    // :930 — only when there is at least one point or spot light
    if (point_light_count == 0 && spot_light_count == 0) return 0;
    // :935-950 build the grid
    RveGridBuildUBO g{};
    std::memcpy(g.invProj, inv_proj, 64);
    std::memcpy(g.gridSize, grid_size, 12);
    g.zNear = z_near;
    g.screenDim[0] = screen_dim[0];
    g.screenDim[1] = screen_dim[1];
    g.zFar = z_far;
    int rc = rve_cluster_build(w->ctx, &g, *buf);
    if (rc) return rc;
    // :953-970 assign lights; the reference dispatches numClusters/128 groups of 128
    RveGridAssignUBO a{};
    std::memcpy(a.viewMat, view_mat, 64);
    a.pointLightCount = point_light_count;
    a.spotLightCount = spot_light_count;
    const uint32_t invocations = uint32_t(n / 128) * 128;
    return rve_cluster_assign(w->ctx, &a, *buf, point_lights, spot_lights, invocations);
}

int rvs_bin_lights(rvs_world* w, const float inv_proj[16], const float view_mat[16], const uint32_t grid_size[3],
                   const uint32_t screen_dim[2], float z_near, float z_far, rve_buffer point_lights,
                   uint32_t point_light_count, rve_buffer spot_lights, uint32_t spot_light_count,
                   rve_buffer* out_cluster_buffer) {
    if (!w || !inv_proj || !view_mat || !grid_size || !screen_dim) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    const int rc = bin_lights_into(w, &w->clusterBuffer, &w->clusterBytes, inv_proj, view_mat, grid_size, screen_dim, z_near, z_far,
                                   point_lights, point_light_count, spot_lights, spot_light_count);
    if (out_cluster_buffer) *out_cluster_buffer = w->clusterBuffer;
    return rc;
}

// One frame's render-prep in one submission. The reference's Draw walks, per camera, six
// renderFromPerspective passes (lit depth prepass + lit opaque, unlit depth prepass + unlit opaque, lit and unlit
// transparent: src/RenderEngine_Draw.cpp:2177-2209, 2567-2655), each with its own cullTheRenderData over the
// materials of the pass, re-bins the lights in every lit pass (:928-974, three times per camera) and runs one more
// cull per shadow view (:1769-1866). The prepass and the colour pass of a material class see the same view, the same
// pyramid and the same materials, and a material belongs to exactly one class, so: every pyramid is built once, ALL
// views (cameras, then shadow views) are culled in ONE rve_cull_views launch over all requested materials, and the
// lights are binned once per camera into that camera's own cluster buffer.
int rvs_frame_encode(rvs_world* w, const rvs_frame_desc* f) {
    if (!w || !f) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_PROFILE_FN_N("RenderEngine::Draw");
    RVE_REQUIRE(f->num_cameras == 0 || f->cameras, "cameras is null");
    RVE_REQUIRE(f->num_shadow_views == 0 || (f->shadow_views && f->shadow_pyramids), "shadow views / pyramids are null");
    const uint32_t nviews = f->num_cameras + f->num_shadow_views;
    RVE_REQUIRE(nviews >= 1 && nviews <= 128, "a frame holds 1..128 views");
    int rc;
    // generatePyramid for every camera that brought a depth buffer (last frame's depth: :2475-2514)
    for (uint32_t c = 0; c < f->num_cameras; ++c) {
        const rvs_frame_camera& cam = f->cameras[c];
        if (cam.depth == RVE_BUFFER_INVALID) continue;
        rc = rve_depth_pyramid_build(w->ctx, cam.pyramid, cam.depth, cam.depth_width, cam.depth_height);
        if (rc) return rc;
    }
    // view-independent skinned work, once per frame (:530-668)
    if (!w->skinnedMaterials.empty() && w->arena[0] != RVE_BUFFER_INVALID) {
        uint32_t exist = 0;
        rc = rvs_prepare_skeletal_mesh_buffers(w, &exist);
        if (rc) return rc;
        if (exist) {
            rc = rvs_pose_skeletal_meshes(w);
            if (rc) return rc;
        }
    }
    std::vector<rvs_view> views(nviews);
    std::vector<rve_pyramid> pyramids(nviews);
    for (uint32_t c = 0; c < f->num_cameras; ++c) {
        const rvs_frame_camera& cam = f->cameras[c];
        std::memcpy(views[c].view_proj, cam.view_proj, 64);
        std::memcpy(views[c].cam_pos, cam.cam_pos, 12);
        views[c].camera_layers = cam.layers;
        views[c].shadow_pass = 0;
        pyramids[c] = cam.pyramid;
    }
    for (uint32_t s = 0; s < f->num_shadow_views; ++s) {
        views[f->num_cameras + s] = f->shadow_views[s];
        pyramids[f->num_cameras + s] = f->shadow_pyramids[s];
    }
    rc = cull_static_impl(w, nviews, views.data(), nullptr, nullptr, 0, 0, f->pass_mask ? f->pass_mask : RVS_MAT_ALL, pyramids.data());
    if (rc) return rc;
    // lights: once per camera
    if (w->frameClusters.size() < f->num_cameras) w->frameClusters.resize(f->num_cameras, std::make_pair(rve_buffer(RVE_BUFFER_INVALID), size_t(0)));
    for (uint32_t c = 0; c < f->num_cameras; ++c) {
        const rvs_frame_camera& cam = f->cameras[c];
        rc = bin_lights_into(w, &w->frameClusters[c].first, &w->frameClusters[c].second, cam.inv_proj, cam.view, f->grid_size, cam.screen_dim,
                             cam.z_near, cam.z_far, f->point_lights, f->point_light_count, f->spot_lights, f->spot_light_count);
        if (rc) return rc;
    }
    w->frameViews = nviews;
    return 0;
}

int rvs_frame_cluster_buffer(rvs_world* w, uint32_t camera, rve_buffer* out) {
    if (!w || !out) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_REQUIRE(camera < w->frameClusters.size(), "no such camera in the last frame");
    *out = w->frameClusters[camera].first;
    return 0;
}

}  // extern "C"
