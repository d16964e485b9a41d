/* crt.h — C ABI of the B200-native render path for CudaRaytracer scenes.
 *
 * This is the drop-in boundary. The reference has no FFI; the
 * surface a replacement sits behind is the set of host functions and the one
 * kernel launch that CudaTest/src/kernel.cu uses. Each entry point below names
 * the reference call it replaces. All paths are relative to
 * /root/reference/CudaTest/src/.
 *
 * Conventions
 *   - plain pointers and sizes; no C++ or torch types;
 *   - every function returns 0 on success and a negative crt_status on
 *     failure, never calls exit() (the reference's checkCudaErrors/CHECK do,
 *     core/deviceManage.h:3-17); crt_last_error() gives the message of the
 *     last failure on the calling thread;
 *   - the caller owns every host array it passes; uploads are complete when
 *     the call returns;
 *   - one host thread per context, contexts are independent, one context
 *     drives one GPU (one process per GPU under torchrun);
 *   - there is no CPU fallback: without a CUDA device crt_create fails.
 */
#ifndef CRT_H
#define CRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CRT_ABI_VERSION 1

typedef struct crt_ctx crt_ctx;

typedef enum {
    CRT_OK = 0,
    CRT_ERR_INVALID = -1,   /* bad argument */
    CRT_ERR_CUDA = -2,      /* CUDA runtime error, see crt_last_error() */
    CRT_ERR_STATE = -3,     /* call out of order (e.g. trace before accel build) */
    CRT_ERR_NOMEM = -4
} crt_status;

#define CRT_MISS 0xFFFFFFFFu

/* A ray as the reference's Ray (core/ray.h:7-23) minus the unused time:
 * origin and a direction of any non-zero length. */
typedef struct { float o[3]; float d[3]; } crt_ray;

/* Closest hit. t is measured along the unit direction, as in the reference
 * where every Hitable::hit() re-normalises the direction
 * (hitable/transform.h:50-54). prim_id is the caller's primitive id
 * (position in the reference's HitableList) or CRT_MISS. u, v are the
 * triangle barycentrics (the reference leaves rec.u/rec.v unset);
 * 0 for spheres. */
typedef struct { float t; uint32_t prim_id; float u; float v; } crt_hit;

/* material/material.h:57-161. param: Metal fuzz (clamped to 1 like
 * material.h:80), Dielectric ref_idx; unused otherwise. albedo: Lambertian
 * texture colour (material/texture.h:12-22), Metal albedo, DiffuseLight
 * emission. texture (Lambertian and DiffuseLight only): CRT_TEX_CONSTANT uses
 * albedo; CRT_TEX_CHECKER is CheckerTexture(even = albedo, odd = albedo2) of two
 * ConstantTextures (material/texture.h:25-42): odd where
 * sin(10 x) sin(10 y) sin(10 z) < 0 at the hit point. A table may hold 32768
 * materials. */
typedef enum {
    CRT_MAT_LAMBERTIAN = 0,
    CRT_MAT_METAL = 1,
    CRT_MAT_DIELECTRIC = 2,
    CRT_MAT_DIFFUSE_LIGHT = 3
} crt_material_type;

typedef enum {
    CRT_TEX_CONSTANT = 0,
    CRT_TEX_CHECKER = 1
} crt_texture_type;

typedef struct { uint32_t type; float albedo[3]; float param; uint32_t texture; float albedo2[3]; } crt_material;

/* The members of the reference Camera after its constructor ran
 * (core/camera.h:46-53), computed once on the host and passed by value so
 * that host and device agree on tan(). */
typedef struct {
    float lower_left_corner[3];
    float horizontal[3];
    float vertical[3];
    float origin[3];
    float x[3], y[3], z[3];
    float lens_radius;
} crt_camera;

/* Which of the reference's three shaders runs per camera sample. */
typedef enum {
    CRT_SHADE_PATH = 0,     /* shade()        core/render.h:48-67  (north-star path tracer) */
    CRT_SHADE_LAMBERT = 1,  /* LambertShade() core/render.h:70-87  (reference's active default) */
    CRT_SHADE_NORMAL = 2    /* shade_normal() core/render.h:90-103 */
} crt_shade_mode;

typedef enum {
    CRT_ACCEL_LBVH = 0,     /* Morton LBVH traversal (replaces hitable/bvh.h:160-190) */
    CRT_ACCEL_BRUTE = 1,    /* linear scan over all primitives (hitable/hitable_list.h:76-91) */
    CRT_ACCEL_LBVH_PACKET = 2 /* LBVH, every 32 consecutive rays walk the tree together: for batches whose neighbours are
                             * nearly parallel (camera rays in pixel order); same hits as CRT_ACCEL_LBVH for any batch */
} crt_accel_kind;

typedef enum {
    CRT_OPT_LEAF_MAX = 1,   /* max primitives per LBVH leaf, 1..8 (default 2); takes effect at the next build */
    CRT_OPT_COUNTERS = 2,   /* 1: run the instrumented kernels that count nodes/primitives visited */
    CRT_OPT_BATCH_PATHS = 3,/* max camera samples in flight per wavefront batch (default 1<<25) */
    CRT_OPT_LEAF_THRESH = 5,/* tuning: lanes parked on a leaf that trigger the primitive-test phase, 1..32 (default 8) */
    CRT_OPT_PACKET_PRIMARY = 6, /* 1 (default): camera rays are traced as one packet per warp; 0: like every other bounce;
                             * 2: packets without the shared interval box test (development aid) */
    CRT_OPT_LANES = 7,      /* render: at most 1..4 slices of the frame in flight on their own streams (default 3; a slice is
                             * never smaller than 2 M camera samples), so one slice's
                             * latency-bound last bounces overlap the other's wide launches; images do not depend on it */
    CRT_OPT_WIDE_BELOW = 9, /* render: once a bounce's queue holds fewer paths than this (default 32768) one launch carries them all to
                             * termination, one WARP per path (closest hits with the lanes working through the ray's stack 32 entries
                             * abreast, then the material, then the next ray), because launches that small are bound by their longest
                             * ray and by launch latency; 0: every bounce is a wavefront launch. Images do not depend on it */
    CRT_OPT_BUILDER = 8,    /* tree topology of crt_accel_build: CRT_BUILDER_LBVH (default) or CRT_BUILDER_PLOC; hits and images do not depend on it */
    CRT_OPT_TIMING = 10,    /* 1: crt_render* records an event pair around every closest-hit launch and waits for the device, so that
                             * crt_counters.trace_queue_ms / bounce_ms are maintained (the roofline's denominator); 0 (default): render
                             * calls only queue work */
    CRT_OPT_SLICE_MIN = 11, /* tuning: fewest camera samples a slice of a frame may hold (default 2,000,000; see CRT_OPT_LANES) */
    CRT_OPT_TAIL_HINT = 12, /* 1 (default): a render call remembers, per slice, the bounce at which the previous frame of the same
                             * shape handed its queue to the one-warp-per-path launch (CRT_OPT_WIDE_BELOW) and queues no launch
                             * beyond it (a frame of depth 8 otherwise queues ~17 launches per slice that find an empty queue);
                             * a wrong guess costs time, never a different image; 0: always queue every bounce */
    CRT_OPT_PACKET_DEFER = 13, /* render: a camera-ray packet that has visited this many nodes + leaves (default 192; the average
                             * packet of C3 visits ~35, the longest ~500) stops, and its rays are traced one WARP per ray by a launch
                             * queued right behind: a launch ends with its longest packet, and ~500 dependent visits take longer
                             * than the rest of a rank's share of a frame split over 8 GPUs. 0: never. Images do not depend on it */
    CRT_OPT_PACKET_DEFER_BELOW = 14, /* ... in slices of at most this many camera samples only (default 4,000,000): in a larger slice
                             * the longest packet hides behind the others and the bookkeeping costs 1-2 % */
    CRT_OPT_BATCH_SAMPLES = 15, /* render: how a frame with more camera samples than CRT_OPT_BATCH_PATHS is cut into jobs. n > 0 (default 64):
                             * into GROUPS OF TILES, each rendered with up to n samples per pixel at once, so that a warp of camera
                             * rays still holds the samples of one or two pixels (a 4K x 64 spp frame on one GPU: 16 tile groups x 64
                             * samples instead of 16 batches of 4 samples over every tile); 0: into sample batches over all of the
                             * rank's tiles. A pixel's samples are accumulated in sample order either way: images do not depend on it */
    CRT_OPT_REFILL = 4      /* tuning: idle lanes per warp that trigger a refill from the ray queue, 1..32 (default 12) */
} crt_option;

typedef enum {
    CRT_BUILDER_LBVH = 0,   /* Karras radix tree over the Morton codes: fastest build */
    CRT_BUILDER_PLOC = 1    /* parallel locally-ordered clustering over the Morton order: 4x the build time; fewer node visits on
                             * scenes of mixed scale (C4 instanced grid: -17 % frame time), none on uniform meshes (C3) */
} crt_builder;

/* Work counters for the roofline and the reference's CSV columns
 * "rendering, update, build" (kernel.cu:61, core/render.h:230). The integer
 * counters other than rays are only maintained while CRT_OPT_COUNTERS is 1. */
typedef struct {
    uint64_t rays;            /* closest-hit queries issued */
    uint64_t nodes_visited;   /* 4-wide node tests executed: one per ray and visit, or one per PACKET and visit where a
                               * packet of camera rays with a common origin takes one interval test (crt_traverse.cuh) */
    uint64_t tris_tested;
    uint64_t spheres_tested;
    uint64_t scatters;        /* material scatter evaluations */
    uint64_t paths;           /* camera samples started */
    float    build_ms;        /* last crt_accel_build */
    float    refit_ms;        /* last crt_accel_refit */
    float    render_ms;       /* last crt_render* call, device time */
    float    trace_ms;        /* last crt_trace (host-buffer form), device time of the kernel */
    float    bounce_ms;       /* CRT_OPT_TIMING only: device time of all trace+shade launches since crt_reset_counters */
    float    trace_queue_ms;  /* CRT_OPT_TIMING only: of which the closest-hit kernels (k_trace_packet, k_trace_queue) alone */
    /* records fetched, the memory side of the roofline: equal to the tests above where
     * every ray walks its own stack; in the packet kernel (camera rays) one fetch serves
     * the <= 32 rays of a packet, while each ray still executes the test */
    uint64_t node_fetches;
    uint64_t tri_fetches;
    uint64_t sphere_fetches;
    uint64_t launches;        /* kernels launched by crt_render*, crt_resolve_device, crt_pack_rgba8_device since crt_reset_counters */
    uint64_t lanes;           /* slices the last crt_render* call kept in flight (CRT_OPT_LANES, capped by the frame size) */
} crt_counters;

/* Image partition for one-process-per-GPU rendering. The image is cut into
 * CRT_TILE x CRT_TILE tiles numbered row-major; part `rank` of `world` owns
 * the tiles with index % world == rank. Pixels a part does not own are
 * written as 0, so parts combine with a plain sum (ncclReduce) bit-exactly. */
#define CRT_TILE 32

/* ---- context: replaces main()'s setup and CudaPointerList (kernel.cu:56-81,
 *      core/deviceManage.h:39-105) ---- */
int  crt_abi_version(void);
int  crt_create(int device, crt_ctx** out);
void crt_destroy(crt_ctx* ctx);
const char* crt_last_error(void);
int  crt_set_option(crt_ctx* ctx, int option, int64_t value);
int  crt_sync(crt_ctx* ctx);
/* The CUDA stream (cudaStream_t) every call of this context is issued on. */
void* crt_stream(crt_ctx* ctx);

/* ---- scene upload: replaces add_mesh_withNormal / create_List
 *      (createScene.h:27-38,175-210) ----
 * tri: n x 12 floats = vertices[0..2] and the stored normal of the reference
 *      Triangle (shapes/triangle.h:49-50), already in world space.
 * sph: n x 4 floats = centre (the Transform position, hitable/transform.h:33)
 *      and radius (shapes/sphere.h:22).
 * mat_id indexes the table given to crt_scene_set_materials.
 * prim_id: the id reported in crt_hit and used to break ties (lowest wins, as
 *      the first list entry wins in hitable_list.h:84); NULL numbers triangles
 *      0..n_tri-1 and spheres n_tri..n_tri+n_sph-1. */
int crt_scene_set_materials(crt_ctx* ctx, const crt_material* mats, size_t n);
int crt_scene_set_triangles(crt_ctx* ctx, const float* tri, const uint32_t* mat_id,
                            const uint32_t* prim_id, size_t n);
int crt_scene_set_spheres(crt_ctx* ctx, const float* sph, const uint32_t* mat_id,
                          const uint32_t* prim_id, size_t n);
/* ---- rotated and nested Hitables, rectangles: the reference's Transform (hitable/transform.h:7-55) ----
 * Every Hitable::hit maps the ray into the object's own frame before the test and rotates the normal on the way
 * back (hitable/hitable.h:29-38): direction re-normalised, origin and direction rotated by the Euler angles
 * (core/vec3.h:200-217), position subtracted. The test then runs IN that frame and reports the hit point there
 * (SURVEY 9 Q6). To reproduce that bit for bit a primitive may carry a CHAIN: the run of frames - one per Hitable
 * from the world list down to the primitive itself, outermost first - its ray passes through. A chain word is
 * (first << 8) | count into the table of crt_scene_set_frames; 0 is the plain case (a primitive directly in the world
 * list under the identity, or a sphere whose only transform is its own translation, given as its centre): vertices
 * and centres are then world-space values and no frame is read. With a chain the triangle vertices, the sphere
 * centre (normally 0) and the rectangle live in the innermost frame. rows: the three rows of the matrix rotate()
 * builds, computed by the caller (they need libm's sinf/cosf; host/scene.h:rotation_rows); has_rotation = 0 skips
 * the multiplication (rotate() by zero angles is the identity). Scale other than 1 is not expressible (the
 * reference's ScaleRay divides only the direction).
 * crt_scene_set_frames drops every chain and rectangle set before it; set it first. */
typedef struct { float rows[9]; float position[3]; uint32_t has_rotation; } crt_frame;
int crt_scene_set_frames(crt_ctx* ctx, const crt_frame* frames, size_t n);
int crt_scene_set_triangle_chains(crt_ctx* ctx, const uint32_t* chain, size_t n_tri);   /* after crt_scene_set_triangles */
int crt_scene_set_sphere_chains(crt_ctx* ctx, const uint32_t* chain, size_t n_sph);     /* after crt_scene_set_spheres */
/* shapes/rectangle.h:4-44: the unit square [-0.5, 0.5]^2 of its frame's z = 0 plane, normal +z (flip_normal: -z),
 * answered only by rays that travel against the normal. chain and prim_id are required. */
int crt_scene_set_rectangles(crt_ctx* ctx, const uint32_t* chain, const uint8_t* flip_normal, const uint32_t* mat_id,
                             const uint32_t* prim_id, size_t n);

/* Replaces updateFBXObj / update_pose (createScene.h:99-140): overwrite the
 * vertices of triangles [first, first+n) with n x 9 floats. Like
 * Triangle::SetVertices (shapes/triangle.h:40-45) the stored normals stay. */
int crt_scene_update_vertices(crt_ctx* ctx, const float* verts, size_t first, size_t n);

/* Indexed mesh pose: the reference's per-frame update is a point array plus the
 * face index array (updateFBXObj uploads newPos and idxVertices, update_pose
 * writes tris[i]->SetVertices({newPos[idx[2]], newPos[idx[1]], newPos[idx[0]]}),
 * createScene.h:99-140). crt_scene_set_indices stores, once, the three point
 * numbers of triangles [first, first + n_tri) in the order the vertices are to be
 * stored (the caller applies the reference's reversal); crt_scene_update_points
 * uploads n_points x 3 floats (12 bytes per point instead of 36 per triangle:
 * 6 MB instead of 36 MB for the 1 M-triangle mesh) and gathers them into the
 * triangles on the device. Stored normals stay, as with SetVertices.
 * crt_scene_set_triangles clears the index buffer. */
int crt_scene_set_indices(crt_ctx* ctx, const uint32_t* idx, size_t first, size_t n_tri, size_t n_points);
int crt_scene_update_points(crt_ctx* ctx, const float* points, size_t n_points);

/* Bone-partitioned geometry: replaces createBoneBVH and Update_BVH(HitableList**,
 * FBXObject*) (createScene.h:221-307, 69-97; hitable/BoneBVH.h:5-173). The reference
 * gives every bone its own BoneBVHNode tree over the triangles skinned to it and
 * a per-bone translation nowTransform, and refits all of them each frame with one
 * thread per bone. Here a part is a set of triangles that moves rigidly:
 * crt_scene_set_parts names the part of every triangle (n_tri must equal the
 * uploaded triangle count) and records the current vertices as the rest pose;
 * crt_scene_move_parts uploads one translation per part (n_parts x 3 floats,
 * 12 bytes per bone instead of 36 bytes per triangle) and writes
 * vertex = rest vertex + translation[part] in one device pass (single rounded
 * FP32 adds; stored normals stay, a translation does not turn them). Follow it
 * with crt_accel_refit, whose bottom-up fit is parallel over all primitives of all
 * parts at once. Hits are those of the flat list over the moved triangles, which is
 * also what the reference's bone trees return. crt_scene_set_triangles clears the
 * parts. */
int crt_scene_set_parts(crt_ctx* ctx, const uint32_t* part_of_tri, size_t n_tri, size_t n_parts);
int crt_scene_move_parts(crt_ctx* ctx, const float* translation, size_t n_parts);

/* ---- acceleration structure: replaces create_BVHfromList and Update_BVH
 *      (createScene.h:212-219, 46-58) ---- */
int crt_accel_build(crt_ctx* ctx);
int crt_accel_refit(crt_ctx* ctx);

/* ---- closest hit: replaces (*world)->hit(r, t_min, t_max, rec, frame)
 *      (core/render.h:53, hitable/hitable.h:29-38) for n rays ----
 * Host-buffer form copies in and out; the _device form takes device pointers,
 * is asynchronous on crt_stream() and is what the C5 sweep times. */
int crt_trace(crt_ctx* ctx, const crt_ray* rays, size_t n, float t_min, float t_max,
              int accel, crt_hit* hits);
int crt_trace_device(crt_ctx* ctx, const crt_ray* d_rays, size_t n, float t_min, float t_max,
                     int accel, crt_hit* d_hits);

/* crt_trace for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc). It has NO counterpart in the reference.
 * Same contract and same hits as crt_trace(..., CRT_ACCEL_LBVH, ...) - host arrays in and out - but every byte of
 * device scratch it needs (the staged rays, the hit records, the persistent kernel's queue cursor) lives in one
 * file-scope `static __device__` array of libcrt_b200.so, CRT_TRACE_STATIC_SCRATCH_BYTES long (about 2.6 MB, held
 * by every GPU that loads the library until the process exits). The call makes no cudaMalloc / cudaFree and calls no
 * allocating routine; it takes at most CRT_TRACE_STATIC_MAX_RAYS rays and fails with CRT_ERR_INVALID beyond that
 * (split the batch). ALL calls on a GPU share the one array, which keeps the previous call's contents: the call
 * clears what must start at zero and waits for its own work before it returns, but two calls that overlap on the same
 * GPU (two host threads, two contexts on one device) race on it - serialise them. Needs crt_accel_build. */
#define CRT_TRACE_STATIC_MAX_RAYS 65536
#define CRT_TRACE_STATIC_SCRATCH_BYTES (256 + CRT_TRACE_STATIC_MAX_RAYS * (sizeof(crt_ray) + sizeof(crt_hit)))
int crt_trace_static(crt_ctx* ctx, const crt_ray* rays, size_t n, float t_min, float t_max, crt_hit* hits);

/* ---- render: replaces render<<<>>> + the D2H copy (core/render.h:105-129,
 *      222-226) ----
 * rgb_out: caller-owned host array of nx*ny*3 floats in the reference's
 * colorBuffer layout: row 0 at the bottom, values after /spp, sqrt and clip. */
int crt_render(crt_ctx* ctx, const crt_camera* cam, int nx, int ny, int spp, int max_depth,
               int shade_mode, uint64_t seed, float* rgb_out);

/* One iteration of the reference's animation loop as a single call:
 * updateFBXObj + Update_BVH + render<<<>>> + the D2H copy
 * (core/render.h:206-233, createScene.h:46-58,125-140). Same results as
 * crt_scene_update_vertices(verts, first, n) + crt_accel_refit + crt_render,
 * but the steps overlap: upload and refit run on their own stream while the
 * camera rays are generated, and finished rows are copied back while the rest
 * of the frame still renders. n may be 0 (no geometry change). verts and
 * rgb_out must stay untouched until the call returns (it returns when the frame
 * is in rgb_out); page-locked buffers are what makes the copies overlap.
 * Needs a built acceleration structure (crt_accel_build). */
int crt_render_frame(crt_ctx* ctx, const float* verts, size_t first, size_t n,
                     const crt_camera* cam, int nx, int ny, int spp, int max_depth,
                     int shade_mode, uint64_t seed, float* rgb_out);

/* crt_render_frame for any of the three pose forms. pose_kind says what `pose` holds:
 *   CRT_POSE_VERTICES  n x 9 floats for triangles [first, first + n)      (crt_scene_update_vertices)
 *   CRT_POSE_POINTS    n x 3 floats, n = n_points, first = 0              (crt_scene_update_points)
 *   CRT_POSE_PARTS     n x 3 floats, n = n_parts, first = 0               (crt_scene_move_parts)
 * crt_render_frame(verts, ...) is crt_render_frame_pose(CRT_POSE_VERTICES, verts, ...). */
typedef enum { CRT_POSE_VERTICES = 0, CRT_POSE_POINTS = 1, CRT_POSE_PARTS = 2 } crt_pose_kind;
int crt_render_frame_pose(crt_ctx* ctx, int pose_kind, const float* pose, size_t first, size_t n,
                          const crt_camera* cam, int nx, int ny, int spp, int max_depth,
                          int shade_mode, uint64_t seed, float* rgb_out);

/* The pose + refit half of crt_render_frame_pose on its own, for callers that render through the _device entry
 * points (one process per GPU): queues the upload of `pose` (forms as above), the pose kernel and the refit on the
 * context's upload stream and returns at once. The next crt_render_accum_device generates its camera rays meanwhile
 * and only its closest-hit launches wait for the refit; every other entry point first waits for the queued pose.
 * `pose` must stay untouched until the next call on this context returns (page-locked memory makes the copy
 * asynchronous). Needs a built acceleration structure. */
int crt_scene_pose_async(crt_ctx* ctx, int pose_kind, const float* pose, size_t first, size_t n);

/* Multi-GPU building blocks (device pointers, asynchronous on crt_stream(): the calls return once their launches
 * are queued, and d_accum / d_rgb are valid in stream order there).
 * crt_render_accum_device adds nothing to d_accum: it overwrites it with, per
 * pixel this part owns, the sum over samples [sample_begin, sample_end) in
 * sample order, and 0 elsewhere. crt_resolve_device applies
 * render.h:123-128 (divide by spp, sqrt, clip). */
int crt_render_accum_device(crt_ctx* ctx, const crt_camera* cam, int nx, int ny,
                            int sample_begin, int sample_end, int max_depth, int shade_mode,
                            uint64_t seed, int rank, int world, float* d_accum);
int crt_resolve_device(crt_ctx* ctx, const float* d_accum, size_t n_pixels, int spp, float* d_rgb);
/* WritePng's conversion (core/render.h:137-146) on the GPU: vertical flip,
 * char(255.99*c), alpha 255. d_rgba8: nx*ny*4 bytes, row 0 at the top. */
int crt_pack_rgba8_device(crt_ctx* ctx, const float* d_rgb, int nx, int ny, uint8_t* d_rgba8);

/* ---- several GPUs: one context per GPU, one communicator across them ----
 * The reference drives a single GPU (kernel.cu:41-111); this is the exchange step of the sharded path. Ranks may be
 * threads of one process (crt_cli --gpus N) or one process per GPU (torchrun). NCCL is loaded at run time
 * (libnccl.so.2); without it these calls fail with CRT_ERR_STATE and nothing else changes. All collectives are
 * issued on crt_stream() and every rank of the communicator must make the same call.
 *   crt_comm_unique_id   rank 0 creates the 128-byte id and hands it to every rank (any transport)
 *   crt_comm_init        collective; gives the context its rank
 *   crt_scene_broadcast  the scene uploaded on `root` travels to all ranks over NVLink (ncclBroadcast per SoA
 *                        buffer); each rank then calls crt_accel_build - the build is deterministic, so all trees agree
 *   crt_reduce_accum_device  sum of the ranks' accumulators onto `root`, in place there. With the tile partition
 *                        (crt_render_accum_device(rank, world)) each pixel is non-zero on one rank: the sum is exact
 *   crt_gather_tiles_device  the tile partition's own exchange: each rank sends only the tiles it rendered (1/N of the
 *                        frame) to `root`, which copies them into its d_accum - the frame there is then the single-GPU
 *                        frame bit for bit; what crt_render_dist uses
 *   crt_gather_device    rank r's `bytes` land at d_recv + r * bytes on `root` (hit records of a crt_trace batch
 *                        split across ranks)
 *   crt_render_dist      one frame: every rank renders its tiles, tile gather, rank 0 resolves and copies the image to
 *                        rgb_out (may be NULL elsewhere); bit-identical to crt_render on one GPU */
#define CRT_COMM_ID_BYTES 128
int crt_comm_unique_id(void* id_out);
int crt_comm_init(crt_ctx* ctx, const void* id, int rank, int world);
int crt_comm_destroy(crt_ctx* ctx);
int crt_comm_info(crt_ctx* ctx, int* rank, int* world, int* nccl_version);
int crt_scene_broadcast(crt_ctx* ctx, int root);
int crt_reduce_accum_device(crt_ctx* ctx, float* d_accum, size_t n_floats, int root);
int crt_gather_tiles_device(crt_ctx* ctx, float* d_accum, int nx, int ny, int root);
int crt_gather_device(crt_ctx* ctx, const void* d_send, size_t bytes, void* d_recv, int root);
int crt_render_dist(crt_ctx* ctx, const crt_camera* cam, int nx, int ny, int spp, int max_depth, int shade_mode,
                    uint64_t seed, float* rgb_out);

/* ---- timing and work counters: replaces StopWatch + CSV
 *      (swatch.h:4-18, Loader/CSVWriter.h:8-32) ---- */
int crt_get_counters(crt_ctx* ctx, crt_counters* out);
int crt_reset_counters(crt_ctx* ctx);

/* ---- test hooks: the builder's stages on their own (no counterpart in the reference, whose builder is one recursive
 *      function, hitable/bvh.h:76-125) ----
 * crt_debug_sort_pairs: the hand-written LSD radix sort of the LBVH build on caller-owned DEVICE arrays: n (64-bit
 *      key, 32-bit value) pairs, stably sorted by key bits [begin_bit, end_bit) (multiples of 8), result back in
 *      d_keys / d_vals; synchronous.
 * crt_debug_read: copies an internal device buffer of the last crt_accel_build to host memory: CRT_DBG_NODES (one
 *      128-byte 4-wide node per binary node, layout in csrc/crt_device.cuh), CRT_DBG_RECS (64 bytes per sorted
 *      primitive), CRT_DBG_KEYS (sorted 64-bit Morton keys), CRT_DBG_PERM (sorted position -> input index),
 *      CRT_DBG_CHILD / CRT_DBG_RANGE (int2 per binary node: children, covered range), CRT_DBG_LEAF_BOX /
 *      CRT_DBG_NODE_BOX (two float4 per entry). `bytes` must not exceed the buffer; *root_out gets the root reference. */
typedef enum { CRT_DBG_NODES = 0, CRT_DBG_RECS = 1, CRT_DBG_KEYS = 2, CRT_DBG_PERM = 3, CRT_DBG_CHILD = 4, CRT_DBG_RANGE = 5,
               CRT_DBG_LEAF_BOX = 6, CRT_DBG_NODE_BOX = 7 } crt_debug_buffer;
int crt_debug_sort_pairs(crt_ctx* ctx, uint64_t* d_keys, uint32_t* d_vals, size_t n, int begin_bit, int end_bit);
int crt_debug_read(crt_ctx* ctx, int which, void* host_out, size_t bytes, int* root_out);

/* Sizes of the acceleration structure, for reports. */
int crt_accel_info(crt_ctx* ctx, uint64_t* n_prims, uint64_t* n_nodes, uint64_t* bytes, int* max_depth);

#ifdef __cplusplus
}
#endif
#endif /* CRT_H */
