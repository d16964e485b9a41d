// crt_api.cu — the C ABI of include/crt.h: context, scene upload, dispatch.
#include "crt_internal.h"
#include <algorithm>
#include <stdlib.h>

#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <vector>

namespace crt {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line)
{
    set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    (void)cudaGetLastError();
    return e == cudaErrorMemoryAllocation ? CRT_ERR_NOMEM : CRT_ERR_CUDA;
}

static int upload(DevBuf& buf, const void* host, size_t bytes, cudaStream_t st)
{
    if (bytes == 0) return CRT_OK;
    CRT_CUDA(buf.reserve(bytes));
    CRT_CUDA(cudaMemcpyAsync(buf.p, host, bytes, cudaMemcpyHostToDevice, st));
    return CRT_OK;
}

static int upload_ids(DevBuf& buf, const uint32_t* ids, size_t n, uint32_t first, cudaStream_t st)
{
    if (n == 0) return CRT_OK;
    if (ids) return upload(buf, ids, n * 4, st);
    std::vector<uint32_t> seq(n);
    for (size_t i = 0; i < n; i++) seq[i] = first + (uint32_t)i;
    int r = upload(buf, seq.data(), n * 4, st);
    if (r == CRT_OK) CRT_CUDA(cudaStreamSynchronize(st));   // seq dies with this scope
    return r;
}

// Triangle::SetVertices (shapes/triangle.h:40-45) for a range of triangles: the
// reference's update_pose runs this loop on one device thread
// (createScene.h:99-109); here one thread per vertex component.
__global__ void __launch_bounds__(256)
k_update_vertices(const float* __restrict__ verts, size_t n_values, size_t first, float* __restrict__ tri_in)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_values) return;
    const size_t t = i / 9, k = i - t * 9;
    tri_in[(first + t) * 12 + k] = verts[i];
}

// A rigid part (bone) moved by a translation: vertex = rest vertex + T[part]. One
// thread per vertex component; plain rounded adds so the host can predict every bit.
// The reference keeps a BoneBVHNode per bone whose boxes live in the bone's frame
// (hitable/BoneBVH.h:83-87,130-135) and still visits every node per frame, one
// thread per bone (createScene.h:69-76); here the pose is one streaming kernel and
// the ordinary refit follows.
__global__ void __launch_bounds__(256)
k_move_parts(const float* __restrict__ rest, const uint32_t* __restrict__ part, const float* __restrict__ T,
             size_t n_values, float* __restrict__ tri_in)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_values) return;
    const size_t t = i / 9, k = i - t * 9;
    tri_in[t * 12 + k] = __fadd_rn(rest[i], T[(size_t)part[t] * 3 + k % 3]);
}

// update_pose (createScene.h:99-109) for an indexed mesh: vertex k of triangle
// first + t = points[idx[3 t + k]]; pure copies, one thread per vertex component.
__global__ void __launch_bounds__(256)
k_gather_points(const float* __restrict__ points, const uint32_t* __restrict__ idx, size_t n_values, size_t first,
                float* __restrict__ tri_in)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_values) return;
    const size_t t = i / 9, k = i - t * 9;
    tri_in[(first + t) * 12 + k] = points[(size_t)idx[t * 3 + k / 3] * 3 + k % 3];
}

__global__ void __launch_bounds__(256)
k_capture_rest(const float* __restrict__ tri_in, size_t n_values, float* __restrict__ rest)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_values) return;
    const size_t t = i / 9, k = i - t * 9;
    rest[i] = tri_in[t * 12 + k];
}

}  // namespace crt

using namespace crt;

#define CRT_CHECK_CTX(c)                                                \
    do {                                                                \
        if (!(c)) { set_error("null context"); return CRT_ERR_INVALID; } \
        cudaError_t e__ = cudaSetDevice((c)->device);                   \
        if (e__ != cudaSuccess) return cuda_fail(e__, "cudaSetDevice", __FILE__, __LINE__); \
    } while (0)

// A pose queued by crt_scene_pose_async is still in flight on the upload stream: wait for it before anything other
// than a render touches the scene, the tree or the staging buffers.
static int settle_pose(crt_ctx* c)
{
    if (!c->pose_pending) return CRT_OK;
    c->pose_pending = false;
    CRT_CUDA(cudaStreamSynchronize(c->up_stream));
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->ev_up0, c->ev_up1) == cudaSuccess) c->host_counters.refit_ms = ms;
    return CRT_OK;
}

#define CRT_CHECK_SETTLED(c)                       \
    do {                                           \
        CRT_CHECK_CTX(c);                          \
        int s__ = settle_pose(c);                  \
        if (s__ != CRT_OK) return s__;             \
    } while (0)

extern "C" {

int crt_abi_version(void) { return CRT_ABI_VERSION; }

const char* crt_last_error(void) { return g_err; }

int crt_create(int device, crt_ctx** out)
{
    if (!out) { set_error("crt_create: out is null"); return CRT_ERR_INVALID; }
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        (void)cudaGetLastError();
        set_error("crt_create: no CUDA device (%s); this library has no CPU path",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return CRT_ERR_CUDA;
    }
    if (device < 0 || device >= count) { set_error("crt_create: device %d out of range [0, %d)", device, count); return CRT_ERR_INVALID; }
    CRT_CUDA(cudaSetDevice(device));
    crt_ctx* c = new crt_ctx();
    c->device = device;
    cudaDeviceProp prop;
    CRT_CUDA(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    if (const char* e = getenv("CRT_BUILDER")) c->builder = atoi(e) == CRT_BUILDER_PLOC ? CRT_BUILDER_PLOC : CRT_BUILDER_LBVH;   // development aids
    if (const char* e = getenv("CRT_PLOC_RADIUS")) c->ploc_radius = std::max(1, std::min(32, atoi(e)));
    // Stream priorities, measured at v9 (profiles/r01_v9_stream_priorities.txt). The upload + refit stream gets the
    // highest priority: the closest-hit launches of every slice wait for it while ray generation fills the GPU, and
    // with equal priorities its kernels queue behind the ray generation blocks (points frame 9.70 -> 9.56 ms).
    // Giving the slices of a frame (lanes) descending priorities - so that they finish one after the other and each
    // one's read-back overlaps the next one's rendering - serialises them and costs 0.7 ms per frame (8.95 -> 9.65):
    // the slices stay equal. CRT_LANE_PRIO (development aid): 0 = no priorities, 1 = descending lanes too, 2 = default.
    int prio_lo = 0, prio_hi = 0;
    CRT_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));   // numerically lower = higher priority
    const int prio_mode = getenv("CRT_LANE_PRIO") ? atoi(getenv("CRT_LANE_PRIO")) : 2;
    const bool use_prio = prio_mode == 1;
    auto lane_prio = [&](int l) { return use_prio ? std::max(prio_hi, std::min(prio_lo, prio_lo - (CRT_MAX_LANES - 1 - l))) : prio_lo; };
    CRT_CUDA(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, lane_prio(0)));
    CRT_CUDA(cudaEventCreate(&c->ev0));
    CRT_CUDA(cudaEventCreate(&c->ev1));
    CRT_CUDA(cudaEventCreate(&c->ev_r0));
    CRT_CUDA(cudaEventCreate(&c->ev_r1));
    CRT_CUDA(cudaEventCreate(&c->ev_fork));
    CRT_CUDA(cudaStreamCreateWithPriority(&c->up_stream, cudaStreamNonBlocking, prio_mode != 0 ? prio_hi : prio_lo));
    CRT_CUDA(cudaEventCreate(&c->ev_up0));
    CRT_CUDA(cudaEventCreate(&c->ev_up1));
    CRT_CUDA(cudaEventCreate(&c->ev_frame));
    for (int l = 0; l < CRT_MAX_LANES; l++) {
        crt_lane& L = c->lane[l];
        if (l == 0) L.stream = c->stream;
        else CRT_CUDA(cudaStreamCreateWithPriority(&L.stream, cudaStreamNonBlocking, lane_prio(l)));
        CRT_CUDA(cudaEventCreate(&L.ev_begin));
        CRT_CUDA(cudaEventCreate(&L.ev_end));
        CRT_CUDA(cudaEventCreateWithFlags(&L.ev_join, cudaEventDisableTiming));
        for (auto& e : L.ev_trace) CRT_CUDA(cudaEventCreate(&e));
        CRT_CUDA(cudaMallocHost((void**)&L.h_counts, sizeof(uint32_t) * (CRT_MAX_DEPTH + 3)));
        CRT_CUDA(cudaHostAlloc((void**)&L.tail_hint, sizeof(uint32_t) * CRT_TAIL_HINTS, cudaHostAllocMapped));
        for (int i = 0; i < CRT_TAIL_HINTS; i++) L.tail_hint[i] = 0u;
        CRT_CUDA(cudaHostGetDevicePointer((void**)&L.tail_hint_dev, L.tail_hint, 0));
    }
    CRT_CUDA(c->counters.reserve(sizeof(unsigned long long) * CNT_COUNT));
    CRT_CUDA(cudaMemsetAsync(c->counters.p, 0, sizeof(unsigned long long) * CNT_COUNT, c->stream));
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    *out = c;
    return CRT_OK;
}

void crt_destroy(crt_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->comm) crt_comm_destroy(c);
    DevBuf* bufs[] = {&c->tri_in, &c->tri_mat, &c->tri_id, &c->sph_in, &c->sph_mat, &c->sph_id, &c->mats,
                      &c->keys[0], &c->keys[1], &c->vals[0], &c->vals[1], &c->hist, &c->nodes, &c->recs,
                      &c->child, &c->parent, &c->leaf_parent, &c->range, &c->node_box, &c->leaf_box, &c->flags,
                      &c->bounds, &c->depth_out, &c->ploc_box[0], &c->ploc_box[1], &c->ploc_nn, &c->ploc_tmp, &c->ploc_blk, &c->work_ctr, &c->accum, &c->rgb,
                      &c->rays_tmp, &c->hits_tmp, &c->counters, &c->rest, &c->part_id, &c->part_T, &c->tri_idx, &c->node_cone, &c->frames, &c->tri_frame, &c->sph_frame, &c->rect_mat, &c->rect_id, &c->rect_frame, &c->tile_stage};
    for (DevBuf* b : bufs) b->release();
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->ev_r0) cudaEventDestroy(c->ev_r0);
    if (c->ev_r1) cudaEventDestroy(c->ev_r1);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->up_stream) { cudaStreamSynchronize(c->up_stream); cudaStreamDestroy(c->up_stream); }
    if (c->ev_up0) cudaEventDestroy(c->ev_up0);
    if (c->ev_up1) cudaEventDestroy(c->ev_up1);
    if (c->ev_frame) cudaEventDestroy(c->ev_frame);
    for (int l = 0; l < CRT_MAX_LANES; l++) {
        crt_lane& L = c->lane[l];
        if (L.stream) cudaStreamSynchronize(L.stream);
        DevBuf* lb[] = {&L.queue[0], &L.queue[1], &L.qcount, &L.work_ctr, &L.hitbuf, &L.rad, &L.deferred};
        for (DevBuf* b : lb) b->release();
        if (L.ev_begin) cudaEventDestroy(L.ev_begin);
        if (L.ev_end) cudaEventDestroy(L.ev_end);
        if (L.ev_join) cudaEventDestroy(L.ev_join);
        for (auto& e : L.ev_trace) if (e) cudaEventDestroy(e);
        if (L.h_counts) cudaFreeHost(L.h_counts);
        if (L.tail_hint) cudaFreeHost(L.tail_hint);
        if (l > 0 && L.stream) cudaStreamDestroy(L.stream);
    }
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int crt_set_option(crt_ctx* c, int option, int64_t value)
{
    CRT_CHECK_SETTLED(c);
    switch (option) {
    case CRT_OPT_LEAF_MAX:
        if (value < 1 || value > 8) { set_error("CRT_OPT_LEAF_MAX must be in [1, 8]"); return CRT_ERR_INVALID; }
        c->leaf_max = (int)value;
        return CRT_OK;
    case CRT_OPT_COUNTERS:
        c->count_work = value != 0;
        return CRT_OK;
    case CRT_OPT_REFILL:
        if (value < 1 || value > 32) { set_error("CRT_OPT_REFILL must be in [1, 32]"); return CRT_ERR_INVALID; }
        c->refill = (int)value;
        return CRT_OK;
    case CRT_OPT_LEAF_THRESH:
        if (value < 1 || value > 32) { set_error("CRT_OPT_LEAF_THRESH must be in [1, 32]"); return CRT_ERR_INVALID; }
        c->leaf_thresh = (int)value;
        return CRT_OK;
    case CRT_OPT_PACKET_PRIMARY:
        c->packet_primary = value == 2 ? 2 : (value != 0);   // 2: packets without the interval test (development aid)
        return CRT_OK;
    case CRT_OPT_BUILDER:
        if (value != CRT_BUILDER_LBVH && value != CRT_BUILDER_PLOC) { set_error("CRT_OPT_BUILDER must be CRT_BUILDER_LBVH or CRT_BUILDER_PLOC"); return CRT_ERR_INVALID; }
        c->builder = (int)value;
        return CRT_OK;
    case CRT_OPT_WIDE_BELOW:
        if (value < 0 || value > (1ll << 31) - 1) { set_error("CRT_OPT_WIDE_BELOW must be in [0, 2^31)"); return CRT_ERR_INVALID; }
        c->wide_below = (int64_t)value;
        return CRT_OK;
    case CRT_OPT_LANES:
        if (value < 1 || value > CRT_MAX_LANES) { set_error("CRT_OPT_LANES must be in [1, %d]", CRT_MAX_LANES); return CRT_ERR_INVALID; }
        c->n_lanes = (int)value;
        return CRT_OK;
    case CRT_OPT_TIMING:
        c->timing = value != 0;
        return CRT_OK;
    case CRT_OPT_TAIL_HINT:
        c->tail_hint = value != 0;
        return CRT_OK;
    case CRT_OPT_PACKET_DEFER_BELOW:
        if (value < 0) { set_error("CRT_OPT_PACKET_DEFER_BELOW must be >= 0"); return CRT_ERR_INVALID; }
        c->packet_defer_below = value;
        return CRT_OK;
    case CRT_OPT_PACKET_DEFER:
        if (value < 0 || value > (1 << 30)) { set_error("CRT_OPT_PACKET_DEFER must be in [0, 2^30]"); return CRT_ERR_INVALID; }
        c->packet_defer = (int)value;
        return CRT_OK;
    case CRT_OPT_SLICE_MIN:
        if (value < 1) { set_error("CRT_OPT_SLICE_MIN must be >= 1"); return CRT_ERR_INVALID; }
        c->slice_min_samples = value;
        return CRT_OK;
    case CRT_OPT_BATCH_PATHS:
        if (value < 1) { set_error("CRT_OPT_BATCH_PATHS must be >= 1"); return CRT_ERR_INVALID; }
        c->batch_paths = value;
        return CRT_OK;
    case CRT_OPT_BATCH_SAMPLES:
        if (value < 0 || value > (1 << 20)) { set_error("CRT_OPT_BATCH_SAMPLES must be in [0, 2^20]"); return CRT_ERR_INVALID; }
        c->batch_samples = (int)value;
        return CRT_OK;
    }
    set_error("crt_set_option: unknown option %d", option);
    return CRT_ERR_INVALID;
}

int crt_sync(crt_ctx* c)
{
    CRT_CHECK_SETTLED(c);
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    return CRT_OK;
}

void* crt_stream(crt_ctx* c) { return c ? (void*)c->stream : nullptr; }

int crt_scene_set_materials(crt_ctx* c, const crt_material* mats, size_t n)
{
    CRT_CHECK_SETTLED(c);
    if (n > 0 && !mats) { set_error("crt_scene_set_materials: null table"); return CRT_ERR_INVALID; }
    if (n > 32768) { set_error("crt_scene_set_materials: at most 32768 materials (got %zu)", n); return CRT_ERR_INVALID; }
    std::vector<DMat> d(n);
    for (size_t i = 0; i < n; i++) {
        if (mats[i].type > CRT_MAT_DIFFUSE_LIGHT) { set_error("crt_scene_set_materials: material %zu has unknown type %u", i, mats[i].type); return CRT_ERR_INVALID; }
        d[i].r = mats[i].albedo[0]; d[i].g = mats[i].albedo[1]; d[i].b = mats[i].albedo[2];
        d[i].param = mats[i].param;
        d[i].type = mats[i].type;
        d[i].r2 = d[i].g2 = d[i].b2 = 0.0f;
        if (mats[i].texture == CRT_TEX_CHECKER) {
            if (mats[i].type != CRT_MAT_LAMBERTIAN && mats[i].type != CRT_MAT_DIFFUSE_LIGHT) { set_error("crt_scene_set_materials: material %zu: only Lambertian and DiffuseLight take a texture (material.h:59,148)", i); return CRT_ERR_INVALID; }
            d[i].type = CRT_DMAT_CHECKER + mats[i].type;
            d[i].r2 = mats[i].albedo2[0]; d[i].g2 = mats[i].albedo2[1]; d[i].b2 = mats[i].albedo2[2];
        } else if (mats[i].texture != CRT_TEX_CONSTANT) {
            set_error("crt_scene_set_materials: material %zu has unknown texture %u", i, mats[i].texture);
            return CRT_ERR_INVALID;
        }
    }
    int r = upload(c->mats, d.data(), n * sizeof(DMat), c->stream);
    if (r != CRT_OK) return r;
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    c->n_mats = n;
    return CRT_OK;
}

static int check_mat_ids(const crt_ctx* c, const uint32_t* mat_id, size_t n, const char* who)
{
    if (!mat_id) return CRT_OK;
    for (size_t i = 0; i < n; i++)
        if (mat_id[i] >= c->n_mats) { set_error("%s: primitive %zu uses material %u but only %zu are set", who, i, mat_id[i], c->n_mats); return CRT_ERR_INVALID; }
    return CRT_OK;
}

int crt_scene_set_triangles(crt_ctx* c, const float* tri, const uint32_t* mat_id, const uint32_t* prim_id, size_t n)
{
    CRT_CHECK_SETTLED(c);
    if (n > 0 && !tri) { set_error("crt_scene_set_triangles: null vertex array"); return CRT_ERR_INVALID; }
    if (n >= (1ull << 28)) { set_error("crt_scene_set_triangles: too many triangles"); return CRT_ERR_INVALID; }
    int r = check_mat_ids(c, mat_id, n, "crt_scene_set_triangles");
    if (r != CRT_OK) return r;
    if ((r = upload(c->tri_in, tri, n * 48, c->stream)) != CRT_OK) return r;
    if (mat_id) { if ((r = upload(c->tri_mat, mat_id, n * 4, c->stream)) != CRT_OK) return r; }
    else if (n) { CRT_CUDA(c->tri_mat.reserve(n * 4)); CRT_CUDA(cudaMemsetAsync(c->tri_mat.p, 0, n * 4, c->stream)); }
    if ((r = upload_ids(c->tri_id, prim_id, n, 0u, c->stream)) != CRT_OK) return r;
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    c->n_tri = n;
    c->tri_has_chain = false;
    c->n_parts = 0;
    c->n_idx_tri = 0;
    c->n_points = 0;
    c->accel_valid = false;
    // default sphere ids follow the triangles: renumber them if they were defaulted against another triangle count
    if (c->sph_id_default && c->n_sph > 0 && c->sph_id_first != (uint32_t)n) {
        if ((r = upload_ids(c->sph_id, nullptr, c->n_sph, (uint32_t)n, c->stream)) != CRT_OK) return r;
        CRT_CUDA(cudaStreamSynchronize(c->stream));
        c->sph_id_first = (uint32_t)n;
    }
    return CRT_OK;
}

int crt_scene_set_spheres(crt_ctx* c, const float* sph, const uint32_t* mat_id, const uint32_t* prim_id, size_t n)
{
    CRT_CHECK_SETTLED(c);
    if (n > 0 && !sph) { set_error("crt_scene_set_spheres: null array"); return CRT_ERR_INVALID; }
    if (n >= (1ull << 28)) { set_error("crt_scene_set_spheres: too many spheres"); return CRT_ERR_INVALID; }
    int r = check_mat_ids(c, mat_id, n, "crt_scene_set_spheres");
    if (r != CRT_OK) return r;
    if ((r = upload(c->sph_in, sph, n * 16, c->stream)) != CRT_OK) return r;
    if (mat_id) { if ((r = upload(c->sph_mat, mat_id, n * 4, c->stream)) != CRT_OK) return r; }
    else if (n) { CRT_CUDA(c->sph_mat.reserve(n * 4)); CRT_CUDA(cudaMemsetAsync(c->sph_mat.p, 0, n * 4, c->stream)); }
    if ((r = upload_ids(c->sph_id, prim_id, n, (uint32_t)c->n_tri, c->stream)) != CRT_OK) return r;
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    c->n_sph = n;
    c->sph_has_chain = false;
    c->sph_id_default = prim_id == nullptr;
    c->sph_id_first = (uint32_t)c->n_tri;
    c->accel_valid = false;
    return CRT_OK;
}

// ---- chains of Transforms and rectangles ----
static int check_chains(const crt_ctx* c, const uint32_t* chain, size_t n, bool need, const char* who)
{
    for (size_t i = 0; i < n; i++) {
        const uint32_t first = chain[i] >> 8, count = chain[i] & 255u;
        if (chain[i] == 0u && !need) continue;
        if (count == 0u || (size_t)first + count > c->n_frames) {
            set_error("%s: primitive %zu names frames [%u, %u) of %zu", who, i, first, first + count, c->n_frames);
            return CRT_ERR_INVALID;
        }
    }
    return CRT_OK;
}

int crt_scene_set_frames(crt_ctx* c, const crt_frame* frames, size_t n)
{
    CRT_CHECK_SETTLED(c);
    if (n > 0 && !frames) { set_error("crt_scene_set_frames: null table"); return CRT_ERR_INVALID; }
    if (n >= (1u << 24)) { set_error("crt_scene_set_frames: at most 2^24 frames"); return CRT_ERR_INVALID; }
    std::vector<DFrame> d(n);
    for (size_t i = 0; i < n; i++) {
        memcpy(d[i].m, frames[i].rows, sizeof d[i].m);
        d[i].px = frames[i].position[0]; d[i].py = frames[i].position[1]; d[i].pz = frames[i].position[2];
        d[i].has_rot = frames[i].has_rotation ? 1u : 0u;
        d[i].pad[0] = d[i].pad[1] = d[i].pad[2] = 0u;
    }
    int r = upload(c->frames, d.data(), n * sizeof(DFrame), c->stream);
    if (r != CRT_OK) return r;
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    c->n_frames = n;
    // chains name frames by number: a new table invalidates them
    c->tri_has_chain = c->sph_has_chain = false;
    c->n_rect = 0;
    c->accel_valid = false;
    return CRT_OK;
}

int crt_scene_set_triangle_chains(crt_ctx* c, const uint32_t* chain, size_t n)
{
    CRT_CHECK_SETTLED(c);
    if (!chain || n == 0) { c->tri_has_chain = false; c->accel_valid = false; return CRT_OK; }
    if (n != c->n_tri) { set_error("crt_scene_set_triangle_chains: %zu chains for %zu triangles", n, c->n_tri); return CRT_ERR_INVALID; }
    int r = check_chains(c, chain, n, false, "crt_scene_set_triangle_chains");
    if (r != CRT_OK) return r;
    if ((r = upload(c->tri_frame, chain, n * 4, c->stream)) != CRT_OK) return r;
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    bool any = false;
    for (size_t i = 0; i < n && !any; i++) any = chain[i] != 0u;
    c->tri_has_chain = any;
    c->accel_valid = false;
    return CRT_OK;
}

int crt_scene_set_sphere_chains(crt_ctx* c, const uint32_t* chain, size_t n)
{
    CRT_CHECK_SETTLED(c);
    if (!chain || n == 0) { c->sph_has_chain = false; c->accel_valid = false; return CRT_OK; }
    if (n != c->n_sph) { set_error("crt_scene_set_sphere_chains: %zu chains for %zu spheres", n, c->n_sph); return CRT_ERR_INVALID; }
    int r = check_chains(c, chain, n, false, "crt_scene_set_sphere_chains");
    if (r != CRT_OK) return r;
    if ((r = upload(c->sph_frame, chain, n * 4, c->stream)) != CRT_OK) return r;
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    bool any = false;
    for (size_t i = 0; i < n && !any; i++) any = chain[i] != 0u;
    c->sph_has_chain = any;
    c->accel_valid = false;
    return CRT_OK;
}

int crt_scene_set_rectangles(crt_ctx* c, const uint32_t* chain, const uint8_t* flip_normal, const uint32_t* mat_id,
                             const uint32_t* prim_id, size_t n)
{
    CRT_CHECK_SETTLED(c);
    if (n == 0) { c->n_rect = 0; c->accel_valid = false; return CRT_OK; }
    if (!chain || !prim_id) { set_error("crt_scene_set_rectangles: a rectangle needs its chain (it is the unit square of a frame) and its id"); return CRT_ERR_INVALID; }
    if (n >= (1ull << 28)) { set_error("crt_scene_set_rectangles: too many rectangles"); return CRT_ERR_INVALID; }
    int r = check_mat_ids(c, mat_id, n, "crt_scene_set_rectangles");
    if (r != CRT_OK) return r;
    if ((r = check_chains(c, chain, n, true, "crt_scene_set_rectangles")) != CRT_OK) return r;
    std::vector<uint32_t> m(n);
    for (size_t i = 0; i < n; i++) m[i] = (mat_id ? mat_id[i] : 0u) | ((flip_normal && flip_normal[i]) ? CRT_RECT_FLIP : 0u);
    if ((r = upload(c->rect_mat, m.data(), n * 4, c->stream)) != CRT_OK) return r;
    if ((r = upload(c->rect_id, prim_id, n * 4, c->stream)) != CRT_OK) return r;
    if ((r = upload(c->rect_frame, chain, n * 4, c->stream)) != CRT_OK) return r;
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    c->n_rect = n;
    c->accel_valid = false;
    return CRT_OK;
}

// Enqueue one pose update on `st` (no synchronisation): the H2D copy of the caller's array into the staging buffer and
// the kernel that rewrites tri_in.
static int enqueue_pose(crt_ctx* c, int kind, const float* data, size_t first, size_t n, cudaStream_t st, const char* who)
{
    if (n == 0) return CRT_OK;
    if (!data) { set_error("%s: null array", who); return CRT_ERR_INVALID; }
    size_t nv = 0;
    switch (kind) {
    case CRT_POSE_VERTICES:
        if (first + n > c->n_tri) { set_error("%s: range [%zu, %zu) exceeds %zu triangles", who, first, first + n, c->n_tri); return CRT_ERR_INVALID; }
        CRT_CUDA(c->rays_tmp.reserve(n * 36));
        CRT_CUDA(cudaMemcpyAsync(c->rays_tmp.p, data, n * 36, cudaMemcpyHostToDevice, st));
        nv = n * 9;
        k_update_vertices<<<(unsigned int)((nv + 255) / 256), 256, 0, st>>>(c->rays_tmp.as<float>(), nv, first, c->tri_in.as<float>());
        break;
    case CRT_POSE_POINTS:
        if (c->n_idx_tri == 0) { set_error("%s: no index buffer (crt_scene_set_indices)", who); return CRT_ERR_STATE; }
        if (first != 0 || n != c->n_points) { set_error("%s: %zu points from %zu given, the index buffer addresses %zu from 0", who, n, first, c->n_points); return CRT_ERR_INVALID; }
        CRT_CUDA(c->rays_tmp.reserve(n * 12));
        CRT_CUDA(cudaMemcpyAsync(c->rays_tmp.p, data, n * 12, cudaMemcpyHostToDevice, st));
        nv = c->n_idx_tri * 9;
        k_gather_points<<<(unsigned int)((nv + 255) / 256), 256, 0, st>>>(c->rays_tmp.as<float>(), c->tri_idx.as<uint32_t>(), nv,
                                                                          c->idx_first, c->tri_in.as<float>());
        break;
    case CRT_POSE_PARTS:
        if (c->n_parts == 0) { set_error("%s: no parts set (crt_scene_set_parts)", who); return CRT_ERR_STATE; }
        if (first != 0 || n != c->n_parts) { set_error("%s: %zu translations from %zu given, the scene has %zu parts", who, n, first, c->n_parts); return CRT_ERR_INVALID; }
        CRT_CUDA(cudaMemcpyAsync(c->part_T.p, data, n * 12, cudaMemcpyHostToDevice, st));
        nv = c->n_tri * 9;
        k_move_parts<<<(unsigned int)((nv + 255) / 256), 256, 0, st>>>(c->rest.as<float>(), c->part_id.as<uint32_t>(),
                                                                       c->part_T.as<float>(), nv, c->tri_in.as<float>());
        break;
    default:
        set_error("%s: unknown pose kind %d", who, kind);
        return CRT_ERR_INVALID;
    }
    CRT_CUDA(cudaGetLastError());
    c->host_counters.launches++;
    return CRT_OK;
}

int crt_scene_update_vertices(crt_ctx* c, const float* verts, size_t first, size_t n)
{
    CRT_CHECK_SETTLED(c);
    // staged through the (otherwise idle) ray upload buffer
    int r = enqueue_pose(c, CRT_POSE_VERTICES, verts, first, n, c->stream, "crt_scene_update_vertices");
    if (r != CRT_OK) return r;
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    return CRT_OK;
}

int crt_scene_set_indices(crt_ctx* c, const uint32_t* idx, size_t first, size_t n_tri, size_t n_points)
{
    CRT_CHECK_SETTLED(c);
    if (n_tri == 0 || n_points == 0) { c->n_idx_tri = 0; c->n_points = 0; return CRT_OK; }
    if (!idx) { set_error("crt_scene_set_indices: null array"); return CRT_ERR_INVALID; }
    if (first + n_tri > c->n_tri) { set_error("crt_scene_set_indices: range [%zu, %zu) exceeds %zu triangles", first, first + n_tri, c->n_tri); return CRT_ERR_INVALID; }
    if (n_points >= (1ull << 32)) { set_error("crt_scene_set_indices: too many points"); return CRT_ERR_INVALID; }
    for (size_t i = 0; i < n_tri * 3; i++)
        if (idx[i] >= n_points) { set_error("crt_scene_set_indices: triangle %zu names point %u of %zu", i / 3, idx[i], n_points); return CRT_ERR_INVALID; }
    int r = upload(c->tri_idx, idx, n_tri * 12, c->stream);
    if (r != CRT_OK) return r;
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    c->idx_first = first;
    c->n_idx_tri = n_tri;
    c->n_points = n_points;
    return CRT_OK;
}

int crt_scene_update_points(crt_ctx* c, const float* points, size_t n_points)
{
    CRT_CHECK_SETTLED(c);
    int r = enqueue_pose(c, CRT_POSE_POINTS, points, 0, n_points, c->stream, "crt_scene_update_points");
    if (r != CRT_OK) return r;
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    return CRT_OK;
}

int crt_scene_pose_async(crt_ctx* c, int pose_kind, const float* pose, size_t first, size_t n)
{
    CRT_CHECK_SETTLED(c);
    if (!c->accel_valid || c->n_prims != (uint32_t)(c->n_tri + c->n_sph + c->n_rect)) {
        set_error("crt_scene_pose_async: call crt_accel_build first (a pose only refits)");
        return CRT_ERR_STATE;
    }
    cudaStream_t up = c->up_stream;
    // behind whatever the context's stream still does with the scene (a render queued by a _device call has
    // finished when that call returned; a crt_trace_device may not have)
    CRT_CUDA(cudaEventRecord(c->ev_fork, c->stream));
    CRT_CUDA(cudaStreamWaitEvent(up, c->ev_fork, 0));
    int r = enqueue_pose(c, pose_kind, pose, first, n, up, "crt_scene_pose_async");
    if (r != CRT_OK) return r;
    r = lbvh_build(c, true, up);
    if (r != CRT_OK) { cudaStreamSynchronize(up); return r; }
    CRT_CUDA(cudaEventRecord(c->ev_frame, up));
    c->pose_pending = true;
    return CRT_OK;
}

int crt_scene_set_parts(crt_ctx* c, const uint32_t* part_of_tri, size_t n_tri, size_t n_parts)
{
    CRT_CHECK_SETTLED(c);
    if (n_tri != c->n_tri) { set_error("crt_scene_set_parts: %zu part ids for %zu triangles", n_tri, c->n_tri); return CRT_ERR_INVALID; }
    if (n_tri == 0 || n_parts == 0) { c->n_parts = 0; return CRT_OK; }
    if (!part_of_tri) { set_error("crt_scene_set_parts: null array"); return CRT_ERR_INVALID; }
    for (size_t i = 0; i < n_tri; i++)
        if (part_of_tri[i] >= n_parts) { set_error("crt_scene_set_parts: triangle %zu is in part %u of %zu", i, part_of_tri[i], n_parts); return CRT_ERR_INVALID; }
    int r = upload(c->part_id, part_of_tri, n_tri * 4, c->stream);
    if (r != CRT_OK) return r;
    CRT_CUDA(c->rest.reserve(n_tri * 36));
    CRT_CUDA(c->part_T.reserve(n_parts * 12));
    const size_t nv = n_tri * 9;
    k_capture_rest<<<(unsigned int)((nv + 255) / 256), 256, 0, c->stream>>>(c->tri_in.as<float>(), nv, c->rest.as<float>());
    CRT_CUDA(cudaGetLastError());
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    c->n_parts = n_parts;
    return CRT_OK;
}

int crt_scene_move_parts(crt_ctx* c, const float* translation, size_t n_parts)
{
    CRT_CHECK_SETTLED(c);
    int r = enqueue_pose(c, CRT_POSE_PARTS, translation, 0, n_parts, c->stream, "crt_scene_move_parts");
    if (r != CRT_OK) return r;
    CRT_CUDA(cudaStreamSynchronize(c->stream));   // the caller's array is free again
    return CRT_OK;
}

int crt_accel_build(crt_ctx* c)
{
    CRT_CHECK_SETTLED(c);
    return lbvh_build(c, false);
}

int crt_accel_refit(crt_ctx* c)
{
    CRT_CHECK_SETTLED(c);
    return lbvh_build(c, true);
}

int crt_trace_device(crt_ctx* c, const crt_ray* d_rays, size_t n, float t_min, float t_max, int accel, crt_hit* d_hits)
{
    CRT_CHECK_SETTLED(c);
    if (n > 0 && (!d_rays || !d_hits)) { set_error("crt_trace_device: null buffer"); return CRT_ERR_INVALID; }
    int r = launch_trace(c, d_rays, n, t_min, t_max, accel, d_hits);
    if (r == CRT_OK) c->host_counters.rays += n;
    return r;
}

int crt_trace(crt_ctx* c, const crt_ray* rays, size_t n, float t_min, float t_max, int accel, crt_hit* hits)
{
    CRT_CHECK_SETTLED(c);
    if (n == 0) return CRT_OK;
    if (!rays || !hits) { set_error("crt_trace: null buffer"); return CRT_ERR_INVALID; }
    CRT_CUDA(c->rays_tmp.reserve(n * sizeof(crt_ray)));
    CRT_CUDA(c->hits_tmp.reserve(n * sizeof(crt_hit)));
    CRT_CUDA(cudaMemcpyAsync(c->rays_tmp.p, rays, n * sizeof(crt_ray), cudaMemcpyHostToDevice, c->stream));
    CRT_CUDA(cudaEventRecord(c->ev0, c->stream));
    int r = launch_trace(c, c->rays_tmp.as<crt_ray>(), n, t_min, t_max, accel, c->hits_tmp.as<crt_hit>());
    if (r != CRT_OK) return r;
    CRT_CUDA(cudaEventRecord(c->ev1, c->stream));
    CRT_CUDA(cudaMemcpyAsync(hits, c->hits_tmp.p, n * sizeof(crt_hit), cudaMemcpyDeviceToHost, c->stream));
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    CRT_CUDA(cudaGetLastError());
    float ms = 0.f;
    CRT_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->host_counters.trace_ms = ms;
    c->host_counters.rays += n;
    return CRT_OK;
}

// no counterpart in the reference (include/crt.h, trace.cu)
int crt_trace_static(crt_ctx* c, const crt_ray* rays, size_t n, float t_min, float t_max, crt_hit* hits)
{
    CRT_CHECK_SETTLED(c);
    return trace_static(c, rays, n, t_min, t_max, hits);
}

int crt_render_accum_device(crt_ctx* c, const crt_camera* cam, int nx, int ny, int sample_begin, int sample_end,
                            int max_depth, int shade_mode, uint64_t seed, int rank, int world, float* d_accum)
{
    CRT_CHECK_CTX(c);
    // a pose queued by crt_scene_pose_async: only the closest-hit launches wait for it (ray generation does not read
    // the scene); render_accum returns after its last launch has finished, so the pose has landed by then
    const bool pending = c->pose_pending;
    const int r = render_accum(c, cam, nx, ny, sample_begin, sample_end, max_depth, shade_mode, seed, rank, world, d_accum,
                               nullptr, pending ? c->ev_frame : nullptr);
    const int s = settle_pose(c);
    return r != CRT_OK ? r : s;
}

int crt_resolve_device(crt_ctx* c, const float* d_accum, size_t n_pixels, int spp, float* d_rgb)
{
    CRT_CHECK_CTX(c);
    if (n_pixels > 0 && (!d_accum || !d_rgb)) { set_error("crt_resolve_device: null buffer"); return CRT_ERR_INVALID; }
    return launch_resolve(c, d_accum, n_pixels, spp, d_rgb);
}

int crt_pack_rgba8_device(crt_ctx* c, const float* d_rgb, int nx, int ny, uint8_t* d_rgba8)
{
    CRT_CHECK_CTX(c);
    if (nx < 0 || ny < 0 || ((size_t)nx * ny > 0 && (!d_rgb || !d_rgba8))) { set_error("crt_pack_rgba8_device: bad arguments"); return CRT_ERR_INVALID; }
    return launch_pack_rgba8(c, d_rgb, nx, ny, d_rgba8);
}

// render + resolve + read-back of one frame; `scene_ready`: event the closest-hit launches wait for (or null)
static int render_to_host(crt_ctx* c, const crt_camera* cam, int nx, int ny, int spp, int max_depth, int shade_mode,
                          uint64_t seed, float* rgb_out, cudaEvent_t scene_ready)
{
    const size_t npix = (size_t)nx * ny;
    CRT_CUDA(c->accum.reserve(npix * 12));
    CRT_CUDA(c->rgb.reserve(npix * 12));
    // every slice of the frame resolves and returns its own rows as soon as they are final (render.cu)
    RenderFinish fin = {c->rgb.as<float>(), rgb_out, spp, false};
    int r = render_accum(c, cam, nx, ny, 0, spp, max_depth, shade_mode, seed, 0, 1, c->accum.as<float>(), &fin, scene_ready);
    if (r != CRT_OK) return r;
    if (!fin.done) {
        if ((r = launch_resolve(c, c->accum.as<float>(), npix, spp, c->rgb.as<float>())) != CRT_OK) return r;
        CRT_CUDA(cudaMemcpyAsync(rgb_out, c->rgb.p, npix * 12, cudaMemcpyDeviceToHost, c->stream));
    }
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    CRT_CUDA(cudaGetLastError());
    return CRT_OK;
}

int crt_render(crt_ctx* c, const crt_camera* cam, int nx, int ny, int spp, int max_depth, int shade_mode,
               uint64_t seed, float* rgb_out)
{
    CRT_CHECK_SETTLED(c);
    if (!rgb_out) { set_error("crt_render: rgb_out is null"); return CRT_ERR_INVALID; }
    if (nx <= 0 || ny <= 0 || spp < 1) { set_error("crt_render: nx, ny and spp must be positive"); return CRT_ERR_INVALID; }
    return render_to_host(c, cam, nx, ny, spp, max_depth, shade_mode, seed, rgb_out, nullptr);
}

// One frame of the reference's animation loop (render.h:206-233: updateFBXObj, Update_BVH, render, copy back) as ONE
// call, so that the steps can overlap: the vertex upload and the refit run on their own stream while the slices
// generate their camera rays, and each slice's rows travel back while the others still render.
int crt_render_frame_pose(crt_ctx* c, int pose_kind, const float* pose, size_t first, size_t n, const crt_camera* cam,
                          int nx, int ny, int spp, int max_depth, int shade_mode, uint64_t seed, float* rgb_out)
{
    CRT_CHECK_SETTLED(c);
    if (!rgb_out) { set_error("crt_render_frame: rgb_out is null"); return CRT_ERR_INVALID; }
    if (nx <= 0 || ny <= 0 || spp < 1) { set_error("crt_render_frame: nx, ny and spp must be positive"); return CRT_ERR_INVALID; }
    if (!c->accel_valid || c->n_prims != (uint32_t)(c->n_tri + c->n_sph + c->n_rect)) {
        set_error("crt_render_frame: call crt_accel_build first (a frame only refits)");
        return CRT_ERR_STATE;
    }
    cudaStream_t up = c->up_stream;
    // behind whatever the context's stream still does with the scene or the staging buffers (an asynchronous
    // crt_trace_device, a render queued by a _device call)
    CRT_CUDA(cudaEventRecord(c->ev_fork, c->stream));
    CRT_CUDA(cudaStreamWaitEvent(up, c->ev_fork, 0));
    int r = enqueue_pose(c, pose_kind, pose, first, n, up, "crt_render_frame");
    if (r != CRT_OK) return r;
    r = lbvh_build(c, true, up);
    if (r != CRT_OK) { cudaStreamSynchronize(up); return r; }
    CRT_CUDA(cudaEventRecord(c->ev_frame, up));
    r = render_to_host(c, cam, nx, ny, spp, max_depth, shade_mode, seed, rgb_out, c->ev_frame);
    CRT_CUDA(cudaStreamSynchronize(up));
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->ev_up0, c->ev_up1) == cudaSuccess) c->host_counters.refit_ms = ms;
    return r;
}

int crt_render_frame(crt_ctx* c, const float* verts, size_t first, size_t n, const crt_camera* cam, int nx, int ny,
                     int spp, int max_depth, int shade_mode, uint64_t seed, float* rgb_out)
{
    return crt_render_frame_pose(c, CRT_POSE_VERTICES, verts, first, n, cam, nx, ny, spp, max_depth, shade_mode, seed, rgb_out);
}

int crt_get_counters(crt_ctx* c, crt_counters* out)
{
    CRT_CHECK_SETTLED(c);
    if (!out) { set_error("crt_get_counters: out is null"); return CRT_ERR_INVALID; }
    unsigned long long h[CNT_COUNT];
    CRT_CUDA(cudaMemcpyAsync(h, c->counters.p, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    if (c->render_ms_pending) {
        c->render_ms_pending = false;
        float ms = 0.f;
        if (cudaEventSynchronize(c->ev_r1) == cudaSuccess && cudaEventElapsedTime(&ms, c->ev_r0, c->ev_r1) == cudaSuccess) c->host_counters.render_ms = ms;
        else (void)cudaGetLastError();
    }
    *out = c->host_counters;
    out->rays += h[CNT_RENDER_RAYS];
    out->paths += h[CNT_RENDER_PATHS];
    out->nodes_visited = h[CNT_NODES];
    out->tris_tested = h[CNT_TRIS];
    out->spheres_tested = h[CNT_SPHERES];
    out->scatters = h[CNT_SCATTERS];
    out->node_fetches = h[CNT_NODE_F];
    out->tri_fetches = h[CNT_TRI_F];
    out->sphere_fetches = h[CNT_SPH_F];
    return CRT_OK;
}

int crt_reset_counters(crt_ctx* c)
{
    CRT_CHECK_SETTLED(c);
    CRT_CUDA(cudaMemsetAsync(c->counters.p, 0, sizeof(unsigned long long) * CNT_COUNT, c->stream));
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    const float b = c->host_counters.build_ms, r = c->host_counters.refit_ms;
    c->host_counters = crt_counters{};
    c->host_counters.build_ms = b;
    c->host_counters.refit_ms = r;
    return CRT_OK;
}

int crt_debug_sort_pairs(crt_ctx* c, uint64_t* d_keys, uint32_t* d_vals, size_t n, int begin_bit, int end_bit)
{
    CRT_CHECK_SETTLED(c);
    if (n == 0) return CRT_OK;
    if (!d_keys || !d_vals || n >= (1ull << 32) || begin_bit < 0 || end_bit > 64 || begin_bit >= end_bit || (begin_bit & 7) || (end_bit & 7)) {
        set_error("crt_debug_sort_pairs: bad arguments");
        return CRT_ERR_INVALID;
    }
    DevBuf k2, v2, hist;
    CRT_CUDA(k2.reserve(n * 8));
    CRT_CUDA(v2.reserve(n * 4));
    CRT_CUDA(hist.reserve(radix_sort_hist_bytes()));
    uint64_t* keys[2] = {d_keys, k2.as<uint64_t>()};
    uint32_t* vals[2] = {d_vals, v2.as<uint32_t>()};
    const int where = radix_sort_pairs(c->stream, keys, vals, (uint32_t)n, begin_bit, end_bit, hist.as<uint32_t>());
    if (where != 0) {
        cudaMemcpyAsync(d_keys, k2.p, n * 8, cudaMemcpyDeviceToDevice, c->stream);
        cudaMemcpyAsync(d_vals, v2.p, n * 4, cudaMemcpyDeviceToDevice, c->stream);
    }
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    k2.release(); v2.release(); hist.release();
    if (e != cudaSuccess) return cuda_fail(e, "crt_debug_sort_pairs", __FILE__, __LINE__);
    return CRT_OK;
}

int crt_debug_read(crt_ctx* c, int which, void* host_out, size_t bytes, int* root_out)
{
    CRT_CHECK_SETTLED(c);
    if (!c->accel_valid) { set_error("crt_debug_read: no acceleration structure"); return CRT_ERR_STATE; }
    if (root_out) *root_out = c->root;
    if (bytes == 0) return CRT_OK;
    if (!host_out) { set_error("crt_debug_read: null buffer"); return CRT_ERR_INVALID; }
    const size_t n = c->n_prims, ni = n > 1 ? n - 1 : 0;
    const DevBuf* b = nullptr;
    size_t have = 0;
    switch (which) {
    case CRT_DBG_NODES: b = &c->nodes; have = ni * 128; break;
    case CRT_DBG_RECS: b = &c->recs; have = n * 64; break;
    case CRT_DBG_KEYS: b = &c->keys[c->builder == CRT_BUILDER_PLOC ? 0 : c->perm_buf]; have = n * 8; break;
    case CRT_DBG_PERM: b = &c->vals[c->perm_buf]; have = n * 4; break;
    case CRT_DBG_CHILD: b = &c->child; have = ni * 8; break;
    case CRT_DBG_RANGE: b = &c->range; have = ni * 8; break;
    case CRT_DBG_LEAF_BOX: b = &c->leaf_box; have = n * 32; break;
    case CRT_DBG_NODE_BOX: b = &c->node_box; have = ni * 32; break;
    default: set_error("crt_debug_read: unknown buffer %d", which); return CRT_ERR_INVALID;
    }
    if (bytes > have || bytes > b->cap) { set_error("crt_debug_read: %zu bytes asked, the buffer holds %zu", bytes, have); return CRT_ERR_INVALID; }
    CRT_CUDA(cudaMemcpyAsync(host_out, b->p, bytes, cudaMemcpyDeviceToHost, c->stream));
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    return CRT_OK;
}

int crt_accel_info(crt_ctx* c, uint64_t* n_prims, uint64_t* n_nodes, uint64_t* bytes, int* max_depth)
{
    CRT_CHECK_SETTLED(c);
    if (!c->accel_valid) { set_error("crt_accel_info: no acceleration structure"); return CRT_ERR_STATE; }
    const uint64_t n = c->n_prims;
    if (n_prims) *n_prims = n;
    if (n_nodes) *n_nodes = n > 1 ? n - 1 : 0;
    if (bytes) *bytes = n * 64 + (n > 1 ? (n - 1) * 128 : 0);
    if (max_depth) *max_depth = c->tree_depth;
    return CRT_OK;
}

}  // extern "C"
