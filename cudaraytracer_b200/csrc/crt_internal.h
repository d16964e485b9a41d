// crt_internal.h — context and cross-file declarations of libcrt_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <string>

#include "../../include/crt.h"
#include "crt_device.cuh"

namespace crt {

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define CRT_CUDA(call)                                                          \
    do {                                                                        \
        cudaError_t e__ = (call);                                               \
        if (e__ != cudaSuccess) return crt::cuda_fail(e__, #call, __FILE__, __LINE__); \
    } while (0)

// Grow-only device allocation.
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { (void)cudaGetLastError(); e = cudaMalloc(&p, bytes); want = bytes; }
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T> T* as() const { return (T*)p; }
};

// Device-side work counters, one 64-bit word each.
enum { CNT_RAYS = 0, CNT_NODES, CNT_TRIS, CNT_SPHERES, CNT_SCATTERS, CNT_PATHS, CNT_MAX_STEPS, CNT_NODE_F, CNT_TRI_F, CNT_SPH_F, CNT_RENDER_RAYS, CNT_RENDER_PATHS, CNT_COUNT };

// Wavefront path state, 64 B = 4 x 128-bit words (SoA by word):
//   w0 = {o.x, o.y, o.z, d.x}
//   w1 = {d.y, d.z, path_id, ctr | depth_left << 24}
//   w2, w3 = the material ids of the bounces so far, 16 bits each (<= 16 bounces)
constexpr int CRT_MAX_DEPTH = 16;

}  // namespace crt

// One slice of a frame in flight (render.cu): its own stream, wavefront
// queues, hit and radiance buffers, queue counters and timing events.
constexpr int CRT_MAX_LANES = 4;
#define CRT_TAIL_HINTS 64
struct crt_lane {
    cudaStream_t stream = nullptr;         // lane 0 uses the context's stream
    crt::DevBuf queue[2], qcount, work_ctr, hitbuf, rad, deferred;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;   // bounces of the current job (CRT_OPT_TIMING)
    cudaEvent_t ev_join = nullptr;         // end of the lane's share of a render call, joined into the context's stream
    cudaEvent_t ev_trace[2 * (crt::CRT_MAX_DEPTH + 1)] = {};
    uint32_t* h_counts = nullptr;          // pinned copy of qcount
    // CRT_OPT_TAIL_HINT: a mapped pinned word the slice's k_accumulate writes, (key << 8) | bounce: the first bounce of
    // the last frame with this key whose queue was short enough for k_tail (0: none). Read by the host, unsynchronised:
    // any value it can hold gives the same image
    uint32_t* tail_hint = nullptr;         // CRT_TAIL_HINTS words: one per tile group of a frame cut into several (render.cu)
    uint32_t* tail_hint_dev = nullptr;     // the same words as the device addresses them
};

struct crt_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_fork = nullptr;
    cudaEvent_t ev_r0 = nullptr, ev_r1 = nullptr;   // begin and end of the last render call on the context's stream
    cudaStream_t up_stream = nullptr;      // crt_render_frame: vertex upload + refit, overlapped with ray generation
    cudaEvent_t ev_up0 = nullptr, ev_up1 = nullptr, ev_frame = nullptr;
    bool pose_pending = false;             // crt_scene_pose_async: upload + refit in flight on up_stream, ev_frame marks their end
    crt_lane lane[CRT_MAX_LANES];
    int n_lanes = 3;                       // CRT_OPT_LANES: most slices of a frame in flight (3: profiles/r02_o_ab_registers_lanes.txt)

    // scene as uploaded (caller's order)
    crt::DevBuf tri_in, tri_mat, tri_id;   // n_tri x 12 floats, u32, u32
    crt::DevBuf sph_in, sph_mat, sph_id;   // n_sph x 4 floats
    crt::DevBuf mats;                      // DMat
    size_t n_tri = 0, n_sph = 0, n_mats = 0;
    bool sph_id_default = false;           // sphere ids were numbered by the library, from sph_id_first on
    uint32_t sph_id_first = 0;
    // chains of enclosing Transforms (rotation): the table and, per primitive, (first << 8) | count, 0 = none
    crt::DevBuf frames, tri_frame, sph_frame;
    size_t n_frames = 0;
    bool tri_has_chain = false, sph_has_chain = false;   // tri_frame / sph_frame hold a chain word per primitive
    // rectangles (shapes/rectangle.h): the unit square of a frame; material id | CRT_RECT_FLIP, primitive id, chain
    crt::DevBuf rect_mat, rect_id, rect_frame;
    size_t n_rect = 0;
    // multi-GPU (comm.cu): NCCL communicator of this context's rank
    void* comm = nullptr;
    crt::DevBuf tile_stage;                // crt_gather_tiles_device: packed tiles (own on a sender, everybody else's on the root)
    int comm_rank = 0, comm_world = 1;
    // rigid parts (bones): rest-pose vertices (n_tri x 9 floats), part of each triangle, the current translations
    crt::DevBuf rest, part_id, part_T;
    size_t n_parts = 0;
    // indexed mesh: triangle idx_first + t has the vertices points[tri_idx[3 t + k]]
    crt::DevBuf tri_idx;
    size_t idx_first = 0, n_idx_tri = 0, n_points = 0;

    // LBVH
    crt::DevBuf keys[2], vals[2], hist;
    crt::DevBuf nodes, recs;
    crt::DevBuf ploc_box[2], ploc_nn, ploc_tmp, ploc_blk;   // PLOC rounds: cluster boxes (double-buffered), partners, scratch
    int builder = CRT_BUILDER_LBVH;
    int ploc_radius = 8;                   // positions searched to either side per PLOC round (<= 32)
    crt::DevBuf child, parent, leaf_parent, range, node_box, node_cone, leaf_box, flags, bounds, depth_out;
    bool accel_valid = false;
    int root = 0;
    int perm_buf = 0;                      // which vals[] holds the sorted permutation
    uint32_t n_prims = 0;
    int leaf_max = 2;
    int tree_depth = 0;

    // render
    crt::DevBuf work_ctr, accum, rgb, rays_tmp, hits_tmp;
    crt::DevBuf counters;                  // CNT_COUNT x uint64
    int count_work = 0;
    int leaf_thresh = 8;                   // parked lanes that trigger the primitive-test phase
    int refill = 12;                       // idle lanes that trigger a refill in the persistent traversal
    int packet_primary = 1;                // camera rays through the packet kernel
    bool tail_hint = true;                 // CRT_OPT_TAIL_HINT
    int packet_defer = 192;                // CRT_OPT_PACKET_DEFER: visits after which a camera-ray packet is handed on; 0 = never
    int64_t packet_defer_below = 4000000;  // CRT_OPT_PACKET_DEFER_BELOW: only in slices of at most this many camera samples
    int64_t wide_below = 32768;            // CRT_OPT_WIDE_BELOW: bounces with fewer rays run one warp per ray
    int64_t batch_paths = 1ll << 25;
    int batch_samples = 64;                // CRT_OPT_BATCH_SAMPLES: a frame beyond batch_paths is cut into tile groups of up to this many samples per pixel; 0: into sample batches of all tiles
    int64_t slice_min_samples = 2000000;   // CRT_OPT_SLICE_MIN: a slice (lane) of a frame holds at least this many camera samples
    int timing = 0;                        // CRT_OPT_TIMING: per-launch events, render calls wait for the device
    int grid_trace = 0, grid_shade = 0, grid_packet = 0, grid_tail = 0;   // persistent grids (render.cu:size_grids)
    bool render_ms_pending = false;        // ev0/ev1 of the last render call not read yet

    crt_counters host_counters = {};
};

namespace crt {

// radix_sort.cu
int radix_sort_pairs(cudaStream_t stream, uint64_t* keys[2], uint32_t* vals[2], uint32_t n, int begin_bit,
                     int end_bit, uint32_t* hist);
size_t radix_sort_hist_bytes();

// lbvh.cu
int lbvh_build(crt_ctx* c, bool refit_only, cudaStream_t async = nullptr);

// trace.cu
// anything but plain triangles (spheres, rectangles, chains of Transforms) takes the kernels' general instantiation
inline bool scene_is_general(const crt_ctx* c) { return c->n_sph > 0 || c->n_rect > 0 || c->tri_has_chain; }
int launch_trace(crt_ctx* c, const crt_ray* d_rays, size_t n, float t_min, float t_max, int accel, crt_hit* d_hits);
AccelView accel_view(const crt_ctx* c);
int trace_static(crt_ctx* c, const crt_ray* rays, size_t n, float t_min, float t_max, crt_hit* hits);

// render.cu
// finish (world == 1 only): every slice resolves the rows it owns into d_rgb and copies them to h_rgb as soon as
// its last sample batch is accumulated, on its own stream; *done says whether that happened.
struct RenderFinish { float* d_rgb; float* h_rgb; int spp; bool done; };
// scene_ready: the closest-hit launches (not ray generation) wait for this event
int render_accum(crt_ctx* c, const crt_camera* cam, int nx, int ny, int s0, int s1, int max_depth, int mode,
                 uint64_t seed, int rank, int world, float* d_accum, RenderFinish* finish = nullptr,
                 cudaEvent_t scene_ready = nullptr);
int launch_resolve(crt_ctx* c, const float* d_accum, size_t n_pixels, int spp, float* d_rgb);
int launch_pack_rgba8(crt_ctx* c, const float* d_rgb, int nx, int ny, uint8_t* d_rgba8);

}  // namespace crt
