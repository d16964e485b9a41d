// comm.cu — several GPUs behind the C ABI: one context per GPU, one NCCL communicator across them.
//
// The reference drives one GPU from one host thread (kernel.cu:41-111); SURVEY §8e shards its path over the GPUs of
// a box: image tiles per GPU against a replicated scene, one sum-reduce of the partial framebuffers over NVLink.
// This file is that exchange step: communicator setup (one rank per context, ranks may be threads of one process or
// processes), the scene broadcast, the framebuffer reduce, a gather for crt_trace batches split across ranks, and
// crt_render_dist, which strings tile rendering, reduce, resolve and read-back together.
//
// NCCL is loaded at run time (dlopen "libnccl.so.2"): a process that already carries one (PyTorch bundles its own)
// shares it, crt_cli picks up the system's, and libcrt_b200.so itself keeps loading on a machine without NCCL - the
// crt_comm_* calls then fail with a message, nothing else changes. Every collective is issued on crt_stream().
#include "crt_internal.h"

#include <dlfcn.h>
#include <mutex>
#include <nccl.h>

namespace crt {

namespace {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    bool ok = false;
    char why[256] = "";
};

NcclApi g_nccl;
std::once_flag g_nccl_once;

void load_nccl()
{
    void* h = nullptr;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) { snprintf(g_nccl.why, sizeof g_nccl.why, "libnccl.so.2 cannot be loaded (%s)", dlerror()); return; }
    bool all = true;
    auto sym = [&](const char* name) { void* p = dlsym(h, name); if (!p) { all = false; snprintf(g_nccl.why, sizeof g_nccl.why, "%s is missing from libnccl", name); } return p; };
    g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))sym("ncclGetUniqueId");
    g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))sym("ncclCommInitRank");
    g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))sym("ncclCommDestroy");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))sym("ncclGetErrorString");
    g_nccl.Reduce = (decltype(g_nccl.Reduce))sym("ncclReduce");
    g_nccl.Broadcast = (decltype(g_nccl.Broadcast))sym("ncclBroadcast");
    g_nccl.Send = (decltype(g_nccl.Send))sym("ncclSend");
    g_nccl.Recv = (decltype(g_nccl.Recv))sym("ncclRecv");
    g_nccl.GroupStart = (decltype(g_nccl.GroupStart))sym("ncclGroupStart");
    g_nccl.GroupEnd = (decltype(g_nccl.GroupEnd))sym("ncclGroupEnd");
    g_nccl.GetVersion = (decltype(g_nccl.GetVersion))sym("ncclGetVersion");
    g_nccl.ok = all;
}

int need_nccl()
{
    std::call_once(g_nccl_once, load_nccl);
    if (!g_nccl.ok) { set_error("crt_comm: %s", g_nccl.why); return CRT_ERR_STATE; }
    return CRT_OK;
}

int nccl_fail(ncclResult_t r, const char* what)
{
    set_error("NCCL error %d (%s) in %s", (int)r, g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?", what);
    return CRT_ERR_CUDA;
}

#define CRT_NCCL(call)                                              \
    do {                                                            \
        ncclResult_t r__ = (call);                                  \
        if (r__ != ncclSuccess) return nccl_fail(r__, #call);       \
    } while (0)

int need_comm(crt_ctx* c, const char* who)
{
    if (!c) { set_error("null context"); return CRT_ERR_INVALID; }
    if (!c->comm) { set_error("%s: no communicator (crt_comm_init)", who); return CRT_ERR_STATE; }
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice", __FILE__, __LINE__);
    return CRT_OK;
}

int bcast(crt_ctx* c, DevBuf& buf, size_t bytes, int root)
{
    if (bytes == 0) return CRT_OK;
    CRT_CUDA(buf.reserve(bytes));
    CRT_NCCL(g_nccl.Broadcast(buf.p, buf.p, bytes, ncclUint8, root, (ncclComm_t)c->comm, c->stream));
    return CRT_OK;
}

// The tiles rank `part` of `world` owns (tile index % world == part), as a compact array [local tile][32 x 32][3]:
// PACK copies them out of the full frame, UNPACK copies them back in. Pixels beyond the image edge are skipped.
template <bool PACK>
__global__ void __launch_bounds__(256)
k_tiles(float* __restrict__ frame, float* __restrict__ compact, int nx, int ny, int tiles_x, int part, int world, int n_local)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;          // one thread per pixel of a local tile
    const size_t lt = i >> 10;
    if (lt >= (size_t)n_local) return;
    const uint32_t s = (uint32_t)(i & 1023u);
    const size_t t = lt * (size_t)world + (size_t)part;
    const int ty = (int)(t / (size_t)tiles_x), tx = (int)(t - (size_t)ty * tiles_x);
    const int x = tx * CRT_TILE + (int)(s & 31u), y = ty * CRT_TILE + (int)(s >> 5);
    if (x >= nx || y >= ny) return;
    float* f = frame + 3 * ((size_t)y * nx + x);
    float* q = compact + 3 * i;
    if (PACK) { q[0] = f[0]; q[1] = f[1]; q[2] = f[2]; }
    else { f[0] = q[0]; f[1] = q[1]; f[2] = q[2]; }
}

}  // namespace

}  // namespace crt

using namespace crt;

extern "C" {

int crt_comm_unique_id(void* id_out)
{
    if (!id_out) { set_error("crt_comm_unique_id: null buffer"); return CRT_ERR_INVALID; }
    int r = need_nccl();
    if (r != CRT_OK) return r;
    static_assert(sizeof(ncclUniqueId) == CRT_COMM_ID_BYTES, "CRT_COMM_ID_BYTES must be NCCL's id size");
    ncclUniqueId id;
    CRT_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(id_out, &id, sizeof id);
    return CRT_OK;
}

int crt_comm_init(crt_ctx* c, const void* id, int rank, int world)
{
    if (!c) { set_error("null context"); return CRT_ERR_INVALID; }
    if (!id || world < 1 || rank < 0 || rank >= world) { set_error("crt_comm_init: bad id/rank/world (%d/%d)", rank, world); return CRT_ERR_INVALID; }
    if (c->comm) { set_error("crt_comm_init: the context already has a communicator"); return CRT_ERR_STATE; }
    int r = need_nccl();
    if (r != CRT_OK) return r;
    CRT_CUDA(cudaSetDevice(c->device));
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof uid);
    ncclComm_t comm = nullptr;
    CRT_NCCL(g_nccl.CommInitRank(&comm, world, uid, rank));
    c->comm = comm;
    c->comm_rank = rank;
    c->comm_world = world;
    return CRT_OK;
}

int crt_comm_destroy(crt_ctx* c)
{
    if (!c) { set_error("null context"); return CRT_ERR_INVALID; }
    if (!c->comm) return CRT_OK;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    ncclComm_t comm = (ncclComm_t)c->comm;
    c->comm = nullptr;
    c->comm_rank = 0;
    c->comm_world = 1;
    CRT_NCCL(g_nccl.CommDestroy(comm));
    return CRT_OK;
}

int crt_comm_info(crt_ctx* c, int* rank, int* world, int* nccl_version)
{
    if (!c) { set_error("null context"); return CRT_ERR_INVALID; }
    if (rank) *rank = c->comm ? c->comm_rank : 0;
    if (world) *world = c->comm ? c->comm_world : 1;
    if (nccl_version) {
        *nccl_version = 0;
        if (need_nccl() == CRT_OK && g_nccl.GetVersion) g_nccl.GetVersion(nccl_version);
    }
    return CRT_OK;
}

// The scene as uploaded on `root` (materials, triangles, spheres, their material and primitive ids, the index buffer
// of an indexed mesh, the frames) travels to every other rank over NVLink; each rank then builds its own LBVH, which
// is deterministic, so all ranks hold the same tree (SURVEY §8e "scene distribution").
int crt_scene_broadcast(crt_ctx* c, int root)
{
    int r = need_comm(c, "crt_scene_broadcast");
    if (r != CRT_OK) return r;
    if (root < 0 || root >= c->comm_world) { set_error("crt_scene_broadcast: bad root %d", root); return CRT_ERR_INVALID; }
    if (c->pose_pending) { set_error("crt_scene_broadcast: a pose is still queued (crt_sync first)"); return CRT_ERR_STATE; }
    unsigned long long head[10] = {c->n_tri, c->n_sph, c->n_mats, c->idx_first, c->n_idx_tri, c->n_points, c->n_frames, c->n_rect,
                                   c->tri_has_chain ? 1ull : 0ull, c->sph_has_chain ? 1ull : 0ull};
    DevBuf hb;
    CRT_CUDA(hb.reserve(sizeof head));
    if (c->comm_rank == root) CRT_CUDA(cudaMemcpyAsync(hb.p, head, sizeof head, cudaMemcpyHostToDevice, c->stream));
    r = bcast(c, hb, sizeof head, root);
    if (r == CRT_OK) {
        cudaError_t e = cudaMemcpyAsync(head, hb.p, sizeof head, cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) r = cuda_fail(e, "crt_scene_broadcast: header", __FILE__, __LINE__);
    }
    hb.release();
    if (r != CRT_OK) return r;
    const size_t n_tri = head[0], n_sph = head[1], n_mats = head[2], n_idx = head[4], n_frames = head[6], n_rect = head[7];
    if ((r = bcast(c, c->mats, n_mats * sizeof(DMat), root)) != CRT_OK) return r;
    if ((r = bcast(c, c->tri_in, n_tri * 48, root)) != CRT_OK) return r;
    if ((r = bcast(c, c->tri_mat, n_tri * 4, root)) != CRT_OK) return r;
    if ((r = bcast(c, c->tri_id, n_tri * 4, root)) != CRT_OK) return r;
    if ((r = bcast(c, c->sph_in, n_sph * 16, root)) != CRT_OK) return r;
    if ((r = bcast(c, c->sph_mat, n_sph * 4, root)) != CRT_OK) return r;
    if ((r = bcast(c, c->sph_id, n_sph * 4, root)) != CRT_OK) return r;
    if ((r = bcast(c, c->tri_idx, n_idx * 12, root)) != CRT_OK) return r;
    if ((r = bcast(c, c->frames, n_frames * sizeof(DFrame), root)) != CRT_OK) return r;
    if ((r = bcast(c, c->tri_frame, head[8] ? n_tri * 4 : 0, root)) != CRT_OK) return r;
    if ((r = bcast(c, c->sph_frame, head[9] ? n_sph * 4 : 0, root)) != CRT_OK) return r;
    if ((r = bcast(c, c->rect_mat, n_rect * 4, root)) != CRT_OK) return r;
    if ((r = bcast(c, c->rect_id, n_rect * 4, root)) != CRT_OK) return r;
    if ((r = bcast(c, c->rect_frame, n_rect * 4, root)) != CRT_OK) return r;
    CRT_CUDA(cudaStreamSynchronize(c->stream));
    if (c->comm_rank != root) {
        c->n_tri = n_tri; c->n_sph = n_sph; c->n_mats = n_mats;
        c->idx_first = head[3]; c->n_idx_tri = n_idx; c->n_points = head[5];
        c->n_frames = n_frames;
        c->n_rect = n_rect;
        c->tri_has_chain = head[8] != 0; c->sph_has_chain = head[9] != 0;
        c->n_parts = 0;
        c->accel_valid = false;
    }
    return CRT_OK;
}

// Sum of every rank's d_accum onto `root`, in place there (ncclReduce over NVLink). With the tile partition of
// crt_render_accum_device every pixel is non-zero on exactly one rank, and adding zeros is exact.
int crt_reduce_accum_device(crt_ctx* c, float* d_accum, size_t n_floats, int root)
{
    int r = need_comm(c, "crt_reduce_accum_device");
    if (r != CRT_OK) return r;
    if (!d_accum || root < 0 || root >= c->comm_world) { set_error("crt_reduce_accum_device: bad arguments"); return CRT_ERR_INVALID; }
    if (n_floats == 0 || c->comm_world == 1) return CRT_OK;
    CRT_NCCL(g_nccl.Reduce(d_accum, d_accum, n_floats, ncclFloat32, ncclSum, root, (ncclComm_t)c->comm, c->stream));
    return CRT_OK;
}

// The tile partition's own exchange: every pixel of the frame is rendered by exactly one rank, so instead of summing
// N full frames (crt_reduce_accum_device: 24.9 MB through every link for the C3 frame) each rank packs the tiles it
// owns (1/N of the frame) and sends them straight to `root`, which copies them into its frame: 1/N of the bytes per
// link, all ranks into the root at once through the switch, and a copy instead of an addition, so the combined frame
// is the single-GPU frame bit for bit by construction. d_accum: the nx x ny x 3 frame crt_render_accum_device(rank,
// world) wrote on this rank; complete on `root` afterwards.
int crt_gather_tiles_device(crt_ctx* c, float* d_accum, int nx, int ny, int root)
{
    int r = need_comm(c, "crt_gather_tiles_device");
    if (r != CRT_OK) return r;
    const int world = c->comm_world, rank = c->comm_rank;
    if (!d_accum || nx <= 0 || ny <= 0 || root < 0 || root >= world) { set_error("crt_gather_tiles_device: bad arguments"); return CRT_ERR_INVALID; }
    if (world == 1) return CRT_OK;
    const int tiles_x = (nx + CRT_TILE - 1) / CRT_TILE, tiles_y = (ny + CRT_TILE - 1) / CRT_TILE;
    const long long tiles = (long long)tiles_x * tiles_y;
    auto n_local = [&](int p) { return tiles > p ? (tiles - p + world - 1) / world : 0ll; };
    const size_t tile_floats = (size_t)CRT_TILE * CRT_TILE * 3;
    ncclComm_t comm = (ncclComm_t)c->comm;
    cudaStream_t st = c->stream;
    if (rank != root) {
        const long long nl = n_local(rank);
        if (nl == 0) return CRT_OK;
        CRT_CUDA(c->tile_stage.reserve((size_t)nl * tile_floats * 4));
        k_tiles<true><<<(unsigned int)((nl * 1024 + 255) / 256), 256, 0, st>>>(d_accum, c->tile_stage.as<float>(), nx, ny, tiles_x, rank, world, (int)nl);
        CRT_CUDA(cudaGetLastError());
        CRT_NCCL(g_nccl.Send(c->tile_stage.p, (size_t)nl * tile_floats, ncclFloat32, root, comm, st));
        return CRT_OK;
    }
    size_t total = 0;
    for (int p = 0; p < world; p++) if (p != root) total += (size_t)n_local(p) * tile_floats;
    CRT_CUDA(c->tile_stage.reserve(total * 4));
    CRT_NCCL(g_nccl.GroupStart());
    ncclResult_t res = ncclSuccess;
    size_t off = 0;
    for (int p = 0; p < world && res == ncclSuccess; p++) {
        if (p == root || n_local(p) == 0) continue;
        res = g_nccl.Recv(c->tile_stage.as<float>() + off, (size_t)n_local(p) * tile_floats, ncclFloat32, p, comm, st);
        off += (size_t)n_local(p) * tile_floats;
    }
    ncclResult_t end = g_nccl.GroupEnd();
    if (res != ncclSuccess) return nccl_fail(res, "ncclRecv");
    if (end != ncclSuccess) return nccl_fail(end, "ncclGroupEnd");
    off = 0;
    for (int p = 0; p < world; p++) {
        if (p == root || n_local(p) == 0) continue;
        const long long nl = n_local(p);
        k_tiles<false><<<(unsigned int)((nl * 1024 + 255) / 256), 256, 0, st>>>(d_accum, c->tile_stage.as<float>() + off, nx, ny, tiles_x, p, world, (int)nl);
        off += (size_t)nl * tile_floats;
    }
    CRT_CUDA(cudaGetLastError());
    return CRT_OK;
}

// Rank r's `bytes` at d_send arrive at d_recv + r * bytes on `root` (d_recv is only read there).
int crt_gather_device(crt_ctx* c, const void* d_send, size_t bytes, void* d_recv, int root)
{
    int r = need_comm(c, "crt_gather_device");
    if (r != CRT_OK) return r;
    if (root < 0 || root >= c->comm_world || (bytes && !d_send) || (bytes && c->comm_rank == root && !d_recv)) { set_error("crt_gather_device: bad arguments"); return CRT_ERR_INVALID; }
    if (bytes == 0) return CRT_OK;
    ncclComm_t comm = (ncclComm_t)c->comm;
    CRT_NCCL(g_nccl.GroupStart());
    ncclResult_t res = ncclSuccess;
    if (c->comm_rank == root) {
        for (int p = 0; p < c->comm_world && res == ncclSuccess; p++) {
            if (p == root) continue;
            res = g_nccl.Recv((char*)d_recv + (size_t)p * bytes, bytes, ncclUint8, p, comm, c->stream);
        }
    } else {
        res = g_nccl.Send(d_send, bytes, ncclUint8, root, comm, c->stream);
    }
    ncclResult_t end = g_nccl.GroupEnd();
    if (res != ncclSuccess) return nccl_fail(res, "ncclSend/ncclRecv");
    if (end != ncclSuccess) return nccl_fail(end, "ncclGroupEnd");
    if (c->comm_rank == root && (char*)d_recv + (size_t)root * bytes != (const char*)d_send)
        CRT_CUDA(cudaMemcpyAsync((char*)d_recv + (size_t)root * bytes, d_send, bytes, cudaMemcpyDeviceToDevice, c->stream));
    return CRT_OK;
}

// One frame across the communicator's GPUs (collective): every rank renders the tiles it owns of the SAME frame
// (bit-identical to the one-GPU image), the tiles travel to rank 0 (crt_gather_tiles_device), which resolves and copies the image
// to rgb_out (ignored on the other ranks, may be NULL there). Returns when the frame is in rgb_out (rank 0) or when
// this rank's share is queued (others).
int crt_render_dist(crt_ctx* c, const crt_camera* cam, int nx, int ny, int spp, int max_depth, int shade_mode,
                    uint64_t seed, float* rgb_out)
{
    int r = need_comm(c, "crt_render_dist");
    if (r != CRT_OK) return r;
    if (nx <= 0 || ny <= 0 || spp < 1) { set_error("crt_render_dist: nx, ny and spp must be positive"); return CRT_ERR_INVALID; }
    if (c->comm_rank == 0 && !rgb_out) { set_error("crt_render_dist: rgb_out is null on rank 0"); return CRT_ERR_INVALID; }
    const size_t npix = (size_t)nx * ny;
    CRT_CUDA(c->accum.reserve(npix * 12));
    if ((r = crt_render_accum_device(c, cam, nx, ny, 0, spp, max_depth, shade_mode, seed, c->comm_rank, c->comm_world, c->accum.as<float>())) != CRT_OK) return r;
    if ((r = crt_gather_tiles_device(c, c->accum.as<float>(), nx, ny, 0)) != CRT_OK) return r;
    if (c->comm_rank == 0) {
        CRT_CUDA(c->rgb.reserve(npix * 12));
        if ((r = launch_resolve(c, c->accum.as<float>(), npix, spp, c->rgb.as<float>())) != CRT_OK) return r;
        CRT_CUDA(cudaMemcpyAsync(rgb_out, c->rgb.p, npix * 12, cudaMemcpyDeviceToHost, c->stream));
        CRT_CUDA(cudaStreamSynchronize(c->stream));
    }
    return CRT_OK;
}

}  // extern "C"
