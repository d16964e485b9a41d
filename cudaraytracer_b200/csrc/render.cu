// render.cu — wavefront path tracer behind crt_render*.
//
// Replaces the reference's render<<<>>> kernel (core/render.h:105-129: one
// thread per pixel, a serial spp loop, recursive shade() with virtual scatter
// calls on a 16 KiB per-thread stack) with a wavefront over camera samples:
//
//   k_raygen      one thread per (pixel, sample): jitter + Camera::get_ray
//                 (render.h:116-118, camera.h:40-44) -> path queue
//   per bounce:
//   k_trace_packet (camera rays) / k_trace_queue (every other bounce): LBVH
//                 closest hit for every queue entry (crt_traverse.cuh) -> hit buffer
//                 k_trace_deferred (small slices): the rays of camera packets that stopped
//                 after CRT_OPT_PACKET_DEFER visits, one warp per ray
//   k_shade       material scatter (material.h:57-143): either a new ray into
//                 the next queue (ballot compaction, one atomic per block) or
//                 the path's radiance
//   k_tail        once a bounce's queue is shorter than CRT_OPT_WIDE_BELOW: one warp per
//                 path, closest hit + scatter until the path ends, in one launch
//   k_accumulate  per pixel, adds the batch's samples IN SAMPLE ORDER, as the
//                 reference's "col += shade(...)" loop does (render.h:115-122)
//   k_resolve     col /= ns; sqrt; clip (render.h:123-128)
//   k_pack_rgba8  WritePng's float -> RGBA8 + vertical flip (render.h:137-146)
//
// shade() is a recursion "emitted + attenuation * shade(next)" (render.h:58).
// A path therefore records the material id of every bounce (16 bits each in
// its queue entry) and, when it ends, folds the product from the innermost
// term outwards, which gives the same rounding as the recursion.
//
// Random numbers: include/crt_rng.h, keyed by (seed, pixel, sample); a path
// carries only its draw counter.
//
// A frame is rendered as up to CRT_OPT_LANES interleaved slices, each running
// this loop on its own stream and buffers (render_accum below).
#include "crt_internal.h"
#include "crt_traverse.cuh"
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>

namespace crt {

// CRT_QUEUE_LOCAL_DIR: a queue entry carries the direction the primitive tests see (ray_dir_local of the ray's direction:
// two normalisations, 2 square roots and 6 divisions) beside the raw direction, computed ONCE by the kernel that
// writes the entry (k_raygen, k_shade), instead of once in the closest-hit kernel - at a refill, with a dozen of 32
// lanes active - and once more in k_shade for the hit point. Words of an entry (SoA, `cap` apart):
//   0: {o.xyz, dl.x}  1: {dl.yz, path_id, ctr|depth}  2, 3: material ids  4: {d.xyz raw, 0}
// The closest-hit kernels still read words 0 and 1 only; k_shade reads word 4 as well (sky, reflect, refract use
// the raw direction). 0: words 0 and 1 hold the raw direction and there is no word 4.
// Measured (profiles/r02_zz_*): the closest-hit launches of a C3 frame 0.561 -> 0.533 ms each, k_raygen and k_shade
// pay for what they save, the frame stays at 6.8 ms (end to end 7.47 -> 7.42 ms); queues grow from 64 to 80 B a path.
#ifndef CRT_QUEUE_LOCAL_DIR
#define CRT_QUEUE_LOCAL_DIR 1
#endif
#define CRT_QUEUE_WORDS (CRT_QUEUE_LOCAL_DIR ? 5 : 4)
#ifndef CRT_RB
#define CRT_RB 128
#endif
constexpr int RB = CRT_RB;   // threads per block of the wavefront kernels

struct RenderParams {
    DCamera cam;
    int nx, ny;
    int tiles_x;
    int rank, world;      // the job renders tiles rank, rank + world, ... (a lane of a rank is a sub-rank: render_accum)
    int band;             // 1: the unit dealt out is a whole tile ROW (rows rank, rank + world, ...), so a job's pixels
                          // are bands of full image rows that can be read back with one strided copy
    uint32_t n_slots;     // slots of this job: its tiles * CRT_TILE^2
    uint32_t tile_base;   // the job's first tile, counted in the (sub-)rank's own tiles (a frame beyond the batch size is cut into tile groups)
    int s_begin;          // first sample index of this batch
    int s_count;          // samples in this batch
    int max_depth;
    int mode;
    uint64_t seed;
    uint64_t seed_key;    // crt_rng_seed_key(seed): the seed's share of every stream key, mixed once on the host
    uint32_t cap;         // queue capacity (entries per SoA word)
};

// slot -> pixel. 10 low bits address a pixel inside a 32x32 tile so that 32
// consecutive slots (one warp) cover an 8x4 block.
__device__ __forceinline__ bool slot_to_pixel(const RenderParams& P, uint32_t slot, int& x, int& y)
{
    const uint32_t lt = (slot >> 10) + P.tile_base, s = slot & 1023u;
    uint32_t tx, ty;
    if (P.band) {
        const uint32_t r = lt / (uint32_t)P.tiles_x;
        tx = lt - r * (uint32_t)P.tiles_x;
        ty = r * (uint32_t)P.world + (uint32_t)P.rank;
    } else {
        const uint32_t t = lt * (uint32_t)P.world + (uint32_t)P.rank;
        ty = t / (uint32_t)P.tiles_x;
        tx = t - ty * (uint32_t)P.tiles_x;
    }
    const uint32_t px = (s & 7u) | (((s >> 5) & 3u) << 3);
    const uint32_t py = ((s >> 3) & 3u) | (((s >> 7) & 7u) << 2);
    x = (int)(tx * CRT_TILE + px);
    y = (int)(ty * CRT_TILE + py);
    return x < P.nx && y < P.ny;
}

// Path id <-> (slot, sample of the batch). Sample-minor: the s_count samples of a
// pixel are consecutive, so a warp of primary rays is 32 / s_count pixels x all
// their samples - nearly identical rays that walk the same nodes.
#ifndef CRT_SAMPLE_MINOR
#define CRT_SAMPLE_MINOR 1
#endif
__device__ __forceinline__ void split_path_id(const RenderParams& P, uint32_t id, uint32_t& slot, uint32_t& s_local)
{
#if CRT_SAMPLE_MINOR
    slot = id / (uint32_t)P.s_count;
    s_local = id - slot * (uint32_t)P.s_count;
#else
    s_local = id / P.n_slots;
    slot = id - s_local * P.n_slots;
#endif
}
__device__ __forceinline__ size_t make_path_id(const RenderParams& P, uint32_t slot, uint32_t s_local)
{
#if CRT_SAMPLE_MINOR
    return (size_t)slot * (uint32_t)P.s_count + s_local;
#else
    return (size_t)s_local * P.n_slots + slot;
#endif
}

// Warp-aggregated append: every lane of the warp must call this.
__device__ __forceinline__ uint32_t warp_append(bool want, uint32_t* counter)
{
    const uint32_t m = __ballot_sync(0xffffffffu, want);
    if (m == 0u) return 0u;
    const uint32_t lane = threadIdx.x & 31u;
    const int leader = __ffs(m) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + (uint32_t)__popc(m & ((1u << lane) - 1u));
}

// Block-aggregated append: ONE atomic on the queue counter per block and iteration
// instead of one per warp (a million same-address atomics per bounce serialise in L2).
// Every thread of the block must call this the same number of times.
// The scratch words are double-buffered by the parity of the call (`it`), so that a call does not have to wait for
// the previous call's readers: two barriers per call instead of three (measured at v10: no difference in the C3
// frame time, 8.92 ms either way - the barrier waits are hidden behind the other resident blocks).
#ifndef CRT_APPEND_BARRIERS
#define CRT_APPEND_BARRIERS 2
#endif
template <int THREADS>
__device__ __forceinline__ uint32_t block_append(bool want, uint32_t* counter, uint32_t* s_cnt /* 2 x (THREADS/32 + 1) */, uint32_t it)
{
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const uint32_t m = __ballot_sync(0xffffffffu, want);
#if CRT_APPEND_BARRIERS == 3
    __syncthreads();                                   // the previous call's readers are done with s_cnt
#else
    s_cnt += (it & 1u) * (THREADS / 32 + 1);
#endif
    if (lane == 0) s_cnt[w] = (uint32_t)__popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t tot = 0;
#pragma unroll
        for (int i = 0; i < THREADS / 32; i++) { const uint32_t c = s_cnt[i]; s_cnt[i] = tot; tot += c; }
        s_cnt[THREADS / 32] = tot ? atomicAdd(counter, tot) : 0u;
    }
    __syncthreads();
    return s_cnt[THREADS / 32] + s_cnt[w] + (uint32_t)__popc(m & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(RB)
k_raygen(RenderParams P, float4* __restrict__ q, uint32_t* __restrict__ q_count)
{
    __shared__ uint32_t s_cnt[2 * (RB / 32 + 1)];
    const uint32_t total = P.n_slots * (uint32_t)P.s_count;
    uint32_t it = 0;
    for (uint32_t bbase = blockIdx.x * RB; bbase < total; bbase += gridDim.x * RB, it++) {   // block-uniform trip count
        const uint32_t id = bbase + threadIdx.x;
        bool live = id < total;
        float4 w0 = make_float4(0, 0, 0, 0), w1 = w0;
#if CRT_QUEUE_LOCAL_DIR
        float4 w4 = w0;
#endif
        if (live) {
            uint32_t slot, s_local;
            split_path_id(P, id, slot, s_local);
            int x, y;
            live = slot_to_pixel(P, slot, x, y);
            if (live) {
                const uint32_t pixel = (uint32_t)y * (uint32_t)P.nx + (uint32_t)x;
                Rng rng;
                rng.key = crt_rng_key_from(P.seed_key, pixel, (uint32_t)(P.s_begin + (int)s_local));
                rng.ctr = 0;
                const float u = x::div(x::add((float)x, rnd(rng)), (float)P.nx);   // render.h:116
                const float v = x::div(x::add((float)y, rnd(rng)), (float)P.ny);   // render.h:117
                v3 o, d;
                camera_get_ray(P.cam, u, v, rng, o, d);                            // render.h:118
#if CRT_QUEUE_LOCAL_DIR
                const v3 dl = ray_dir_local(d);
                w0 = make_float4(o.x, o.y, o.z, dl.x);
                w1 = make_float4(dl.y, dl.z, __uint_as_float(id), __uint_as_float(rng.ctr | ((uint32_t)P.max_depth << 24)));
                w4 = make_float4(d.x, d.y, d.z, 0.0f);
#else
                w0 = make_float4(o.x, o.y, o.z, d.x);
                w1 = make_float4(d.y, d.z, __uint_as_float(id), __uint_as_float(rng.ctr | ((uint32_t)P.max_depth << 24)));
#endif
            }
        }
        const uint32_t pos = block_append<RB>(live, q_count, s_cnt, it);
        if (live) {   // the material-id words (2, 3) are first written by k_shade and never read at bounce 0
            __stcs(q + pos, w0);
            __stcs(q + (size_t)P.cap + pos, w1);
#if CRT_QUEUE_LOCAL_DIR
            __stcs(q + 4 * (size_t)P.cap + pos, w4);
#endif
        }
    }
}

// attenuation a finished bounce contributed, by material (material.h:69,90,111)
__device__ __forceinline__ v3 bounce_attenuation(const DMat& m, bool odd)
{
    return m.type == 2u ? V(1.0f, 1.0f, 1.0f) : texture_value(m, odd);
}

__device__ __forceinline__ DMat load_mat(const DMat* __restrict__ mats, uint32_t id)
{
    const float4* p = reinterpret_cast<const float4*>(mats + id);
    const float4 a = __ldg(p);
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(p + 1));
    DMat m;
    m.r = a.x; m.g = a.y; m.b = a.z; m.param = a.w; m.type = b.x;
    m.r2 = __uint_as_float(b.y); m.g2 = __uint_as_float(b.z); m.b2 = __uint_as_float(b.w);
    return m;
}

// ---- closest hit for every queue entry (persistent warps, crt_traverse.cuh) ----
struct QueueSrc {
    const float4* q;
    size_t cap;
    __device__ __forceinline__ void load(uint32_t i, v3& o, v3& d) const
    {
        // queue entries and hit records stream through once per launch (GB per bounce): evict-first loads and
        // stores keep them from pushing the node and record working set (which does fit) out of the 126 MB L2
        const float4 w0 = __ldcs(q + i);
#if CRT_QUEUE_LOCAL_DIR
        const float4 w4 = __ldcs(q + 4 * cap + i);      // the raw direction (chained primitives normalise per level)
        o = V(w0.x, w0.y, w0.z);
        d = V(w4.x, w4.y, w4.z);
#else
        const float4 w1 = __ldcs(q + cap + i);
        o = V(w0.x, w0.y, w0.z);
        d = V(w0.w, w1.x, w1.y);
#endif
    }
#if CRT_QUEUE_LOCAL_DIR
    // origin and the direction the primitive tests see, as the entry's writer computed it
    __device__ __forceinline__ void load_local(uint32_t i, v3& o, v3& dl) const
    {
        const float4 w0 = __ldcs(q + i);
        const float4 w1 = __ldcs(q + cap + i);
        o = V(w0.x, w0.y, w0.z);
        dl = V(w0.w, w1.x, w1.y);
    }
#endif
};
struct QueueSink {
    float4* hits;   // {t, sorted primitive position, u, v} per queue entry
    __device__ __forceinline__ void store(uint32_t i, const Hit& h) const
    {
        __stcs(hits + i, make_float4(h.t, __uint_as_float(h.k), h.u, h.v));
    }
};

template <bool COUNT, bool HAS_SPH>
__global__ void __launch_bounds__(TP_BLOCK, CRT_TRACE_MIN_BLOCKS)
k_trace_queue(AccelView A, const float4* __restrict__ q, uint32_t cap, const uint32_t* __restrict__ n_ptr,
              uint32_t* __restrict__ work_ctr, float4* __restrict__ hits, unsigned long long* __restrict__ counters,
              uint32_t refill, uint32_t leaf_thresh, uint32_t wide_below)
{
    __shared__ int s_stack[CRT_STACK_WORDS * CRT_STACK_SMEM * TP_BLOCK];
    const uint32_t n = *n_ptr;
    if (n < wide_below) return;                     // a queue this small belongs to k_tail (queued right behind)
    TraceCounters cnt = {0u, 0u, 0u, 0u, 0u, 0u, 0u};
    QueueSrc src = {q, cap};
    QueueSink sink = {hits};
    traverse_persistent<COUNT, HAS_SPH>(A, n, work_ctr, 0.001f, FLT_MAX, src, sink, s_stack, cnt, refill, leaf_thresh);   // render.h:53
    if (COUNT) {
        const uint32_t nodes = __reduce_add_sync(0xffffffffu, cnt.nodes);
        const uint32_t tris = __reduce_add_sync(0xffffffffu, cnt.tris);
        const uint32_t sph = __reduce_add_sync(0xffffffffu, cnt.spheres);
        if ((threadIdx.x & 31u) == 0u) {
            atomicAdd(counters + CNT_NODES, (unsigned long long)nodes);
            atomicAdd(counters + CNT_TRIS, (unsigned long long)tris);
            atomicAdd(counters + CNT_SPHERES, (unsigned long long)sph);
            atomicAdd(counters + CNT_NODE_F, (unsigned long long)nodes);
            atomicAdd(counters + CNT_TRI_F, (unsigned long long)tris);
            atomicAdd(counters + CNT_SPH_F, (unsigned long long)sph);
        }
        atomicMax(counters + CNT_MAX_STEPS, (unsigned long long)cnt.max_steps);
    }
}

// ---- closest hit for camera rays: one packet per warp (crt_traverse.cuh traverse_packets) ----
template <bool COUNT, bool HAS_SPH, bool DEFER>
__global__ void __launch_bounds__(TP_BLOCK, CRT_PACKET_MIN_BLOCKS)
k_trace_packet(AccelView A, const float4* __restrict__ q, uint32_t cap, const uint32_t* __restrict__ n_ptr,
               uint32_t* __restrict__ work_ctr, float4* __restrict__ hits, unsigned long long* __restrict__ counters, int interval,
               PacketDefer defer)
{
    TraceCounters cnt = {0u, 0u, 0u, 0u, 0u, 0u, 0u};
    QueueSrc src = {q, cap};
    QueueSink sink = {hits};
    traverse_packets<COUNT, HAS_SPH, DEFER>(A, *n_ptr, work_ctr, 0.001f, FLT_MAX, src, sink, cnt, interval != 0, defer);   // render.h:53
    if (COUNT) {
        const uint32_t nodes = __reduce_add_sync(0xffffffffu, cnt.nodes);
        const uint32_t tris = __reduce_add_sync(0xffffffffu, cnt.tris);
        const uint32_t sph = __reduce_add_sync(0xffffffffu, cnt.spheres);
        if ((threadIdx.x & 31u) == 0u) {
            atomicAdd(counters + CNT_NODES, (unsigned long long)nodes);
            atomicAdd(counters + CNT_TRIS, (unsigned long long)tris);
            atomicAdd(counters + CNT_SPHERES, (unsigned long long)sph);
            atomicAdd(counters + CNT_NODE_F, (unsigned long long)cnt.node_f);
            atomicAdd(counters + CNT_TRI_F, (unsigned long long)cnt.tri_f);
            atomicAdd(counters + CNT_SPH_F, (unsigned long long)cnt.sph_f);
        }
    }
}

// ---- the rays of the packets that were handed on (crt_traverse.cuh PacketDefer): one warp per ray ----
template <bool COUNT, bool HAS_SPH>
__global__ void __launch_bounds__(TP_BLOCK)
k_trace_deferred(AccelView A, const float4* __restrict__ q, uint32_t cap, const uint32_t* __restrict__ list,
                 const uint32_t* __restrict__ n_ptr, uint32_t list_cap, uint32_t* __restrict__ cursor, float4* __restrict__ hits,
                 unsigned long long* __restrict__ counters)
{
    __shared__ uint2 s_stack[(TP_BLOCK / 32) * CRT_WIDE_STACK];
    const uint32_t n = min(*n_ptr, list_cap);
    if (n == 0u) return;
    const uint32_t lane = threadIdx.x & 31u;
    uint2* stack = s_stack + (threadIdx.x >> 5) * CRT_WIDE_STACK;
    TraceCounters cnt = {0u, 0u, 0u, 0u, 0u, 0u, 0u};
    QueueSrc src = {q, cap};
    QueueSink sink = {hits};
    for (;;) {
        uint32_t i = 0;
        if (lane == 0) i = atomicAdd(cursor, 1u);
        i = __shfl_sync(0xffffffffu, i, 0);
        if (i >= n) break;
        const uint32_t ray_i = list[i];
        if (ray_i == 0xFFFFFFFFu) continue;             // a lane of the packet past the end of the queue
        v3 o, dr;
        src.load(ray_i, o, dr);
        const Hit h = trace_wide_one<COUNT, HAS_SPH>(A, o, dr, 0.001f, FLT_MAX, stack, cnt);   // render.h:53
        if (lane == 0) sink.store(ray_i, h);
    }
    if (COUNT) {
        const uint32_t nodes = __reduce_add_sync(0xffffffffu, cnt.nodes);
        const uint32_t tris = __reduce_add_sync(0xffffffffu, cnt.tris);
        const uint32_t sph = __reduce_add_sync(0xffffffffu, cnt.spheres);
        if (lane == 0) {
            atomicAdd(counters + CNT_NODES, (unsigned long long)nodes);
            atomicAdd(counters + CNT_TRIS, (unsigned long long)tris);
            atomicAdd(counters + CNT_SPHERES, (unsigned long long)sph);
            atomicAdd(counters + CNT_NODE_F, (unsigned long long)nodes);
            atomicAdd(counters + CNT_TRI_F, (unsigned long long)tris);
            atomicAdd(counters + CNT_SPH_F, (unsigned long long)sph);
        }
    }
}

// ---- shade(): the body of the reference's recursion for one path segment ----
struct PathState {
    v3 o, d;              // the ray that was traced (d as shade() holds it: not normalised)
    v3 dl;                // CRT_QUEUE_LOCAL_DIR: ray_dir_local(d), carried by the queue entry
    uint32_t path_id;     // slot * s_count + sample of the batch
    uint32_t pack;        // draw counter | depth_left << 24
    uint32_t ms[8];       // material ids of the bounces so far, 16 bits each
};

// One evaluation of shade()'s body (render.h:48-67) for a path whose ray has the closest hit (ht, hk = sorted
// primitive position or ~0u). Returns true when the path goes on: S then holds the scattered ray, the bounce's
// material id, the new draw counter and depth. Otherwise L is the path's radiance, unwound through every recorded
// bounce innermost first.
template <bool COUNT, bool HAS_SPH>
__device__ __forceinline__ bool shade_step(const RenderParams& P, const AccelView& A, const DMat* __restrict__ mats,
                                           PathState& S, const float ht, const uint32_t hk, v3& L, uint32_t& n_scatter)
{
    const v3 o = S.o, d = S.d;
    const int depth_left = (int)(S.pack >> 24);
    const int bounces = P.max_depth - depth_left;
    bool next = false;
    if (hk == 0xFFFFFFFFu) {
        L = background_sky(d);                                             // render.h:65
    } else {
        Rng rng;
        {
            uint32_t slot, s_local;
            split_path_id(P, S.path_id, slot, s_local);
            int px, py;
            slot_to_pixel(P, slot, px, py);
            rng.key = crt_rng_key_from(P.seed_key, (uint32_t)py * (uint32_t)P.nx + (uint32_t)px, (uint32_t)(P.s_begin + (int)s_local));
            rng.ctr = S.pack & 0xFFFFFFu;
        }
        float4 r0, r1, r2, r3;
        ldg256(A.recs + 4 * (size_t)hk, r0, r1);
        ldg256(A.recs + 4 * (size_t)hk + 2, r2, r3);
        const uint32_t inf_y = __float_as_uint(r3.y), chain = __float_as_uint(r3.z);
        // the ray in the hit primitive's own frame: hit point and normal are reported there
        v3 ol = o, dl;
        if (HAS_SPH && chain != 0u) chain_ray(A.frames, chain, o, d, ol, dl);
#if CRT_QUEUE_LOCAL_DIR
        else dl = S.dl;
#else
        else dl = ray_dir_local(d);
#endif
        v3 p, n;
        if (HAS_SPH && (inf_y & CRT_KIND_SPHERE)) {
            p = vadd(vsub(ol, V(r0.x, r0.y, r0.z)), vscale(ht, dl));       // sphere.h:40 (local frame)
            n = vdivf(p, r0.w);                                            // sphere.h:41
        } else if (HAS_SPH && (inf_y & CRT_KIND_RECT)) {
            p = vadd(ol, vscale(ht, dl));                                  // rectangle.h:41
            n = V(0.0f, 0.0f, (inf_y & CRT_RECT_FLIP) ? -1.0f : 1.0f);     // rectangle.h:23,42
        } else {
            p = vadd(ol, vscale(ht, dl));                                  // triangle.h:95
            n = V(r0.w, r1.w, r2.w);                                       // triangle.h:96
        }
        if (HAS_SPH && chain != 0u) n = chain_normal(A.frames, chain, n);  // hitable.h:36, every level on the way up
        const uint32_t mat_id = inf_y & CRT_MAT_MASK;
        const DMat m = load_mat(mats, mat_id);
        const bool odd = checker_odd(m, p);                                // texture.h:31
        const v3 e = emitted(m, odd);                                      // render.h:56
        if (P.mode == CRT_SHADE_PATH) {
            v3 att, out_d;
            if (COUNT && depth_left > 0) n_scatter++;
            if (depth_left > 0 && scatter(m, d, p, n, rng, att, out_d, odd)) {  // render.h:57
                next = true;
                const uint32_t hist = (mat_id & 0x7FFFu) | (odd ? 0x8000u : 0u);   // 15 bits of id + which checker colour
#pragma unroll
                for (int j = 0; j < 8; j++)
                    if (j == (bounces >> 1))
                        S.ms[j] = (bounces & 1) ? ((S.ms[j] & 0xFFFFu) | (hist << 16)) : hist;
                S.o = p;
                S.d = out_d;
#if CRT_QUEUE_LOCAL_DIR
                S.dl = ray_dir_local(out_d);
#endif
                S.pack = (rng.ctr & 0xFFFFFFu) | ((uint32_t)(depth_left - 1) << 24);
            } else {
                L = vadd(e, V(0.1f, 0.1f, 0.1f));                          // render.h:61
            }
        } else if (P.mode == CRT_SHADE_LAMBERT) {                          // render.h:70-87
            v3 att, out_d;
            if (COUNT) n_scatter++;
            scatter(m, d, p, n, rng, att, out_d, odd);
            float c = vdot(d, n);
            if (c < 0.0f) c = 0.0f;
            L = vadd(vscale(0.2f, vmul(vscale(c, att), background_sky(d))), e);
        } else {                                                           // render.h:90-103
            L = n;
        }
    }
    if (!next) {
        // unwind "emitted + attenuation * shade(next)" (render.h:58), innermost first
#pragma unroll
        for (int j = CRT_MAX_DEPTH - 1; j >= 0; j--) {
            if (j < bounces) {
                const uint32_t id = (S.ms[j >> 1] >> ((j & 1) * 16)) & 0xFFFFu;
                const DMat mj = load_mat(mats, id & 0x7FFFu);
                // a material that scattered is not a light (material.h:149-155): its emitted() is 0
                L = vadd(V(0.0f, 0.0f, 0.0f), vmul(bounce_attenuation(mj, (id & 0x8000u) != 0u), L));
            }
        }
    }
    return next;
}

// path state of queue entry i; the material-id words exist once a bounce has been recorded
__device__ __forceinline__ void load_path(const RenderParams& P, const float4* __restrict__ q, uint32_t i, PathState& S)
{
    const float4 w0 = __ldcs(q + i);
    const float4 w1 = __ldcs(q + (size_t)P.cap + i);
    S.o = V(w0.x, w0.y, w0.z);
#if CRT_QUEUE_LOCAL_DIR
    const float4 w4 = __ldcs(q + 4 * (size_t)P.cap + i);
    S.dl = V(w0.w, w1.x, w1.y);
    S.d = V(w4.x, w4.y, w4.z);
#else
    S.d = V(w0.w, w1.x, w1.y);
    S.dl = S.d;           // unused
#endif
    S.path_id = __float_as_uint(w1.z);
    S.pack = __float_as_uint(w1.w);
#pragma unroll
    for (int j = 0; j < 8; j++) S.ms[j] = 0u;
    const int bounces = P.max_depth - (int)(S.pack >> 24);
    if (bounces > 0) {
        const uint4 a2 = __ldcs(reinterpret_cast<const uint4*>(q) + 2 * (size_t)P.cap + i);
        S.ms[0] = a2.x; S.ms[1] = a2.y; S.ms[2] = a2.z; S.ms[3] = a2.w;
        if (bounces > 8) {
            const uint4 a3 = __ldcs(reinterpret_cast<const uint4*>(q) + 3 * (size_t)P.cap + i);
            S.ms[4] = a3.x; S.ms[5] = a3.y; S.ms[6] = a3.z; S.ms[7] = a3.w;
        }
    }
}

// one thread per queue entry, fully converged. A queue of fewer than `tail_below` entries belongs to k_tail.
#ifndef CRT_SHADE_PREFETCH
#define CRT_SHADE_PREFETCH 0
#endif
#ifndef CRT_SHADE_MIN_BLOCKS
#define CRT_SHADE_MIN_BLOCKS 10  // 48 registers instead of 56, no spills: 10 resident blocks instead of 9 hide more of the record and
                                 // queue latency (C3 frame 6.95 -> 6.76 ms, profiles/r02_s_ab.txt)
#endif
template <bool COUNT, bool HAS_SPH>
__global__ void __launch_bounds__(RB, CRT_SHADE_MIN_BLOCKS * 128 / RB)
k_shade(RenderParams P, AccelView A, const DMat* __restrict__ mats, const float4* __restrict__ q_in,
        const float4* __restrict__ hits, const uint32_t* __restrict__ n_in_ptr, float4* __restrict__ q_out,
        uint32_t* __restrict__ n_out_ptr, float4* __restrict__ rad, unsigned long long* __restrict__ counters,
        uint32_t tail_below)
{
    __shared__ uint32_t s_cnt[2 * (RB / 32 + 1)];
    const uint32_t n_in = *n_in_ptr;
    if (n_in < tail_below) return;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t n_scatter = 0;

    uint32_t it = 0;
    for (uint32_t bbase = blockIdx.x * RB; bbase < n_in; bbase += gridDim.x * RB, it++) {   // block-uniform trip count
        const uint32_t i = bbase + threadIdx.x;
        const bool live = i < n_in;
        bool next = false;
        PathState S;
#if CRT_SHADE_PREFETCH
        {   // the next iteration's queue entry and hit record on their way into L2 while this one is shaded (development switch)
            const uint32_t i_next = i + gridDim.x * RB;
            if (i_next < n_in) {
                asm volatile("prefetch.global.L2 [%0];" ::"l"(q_in + i_next));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(q_in + (size_t)P.cap + i_next));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(hits + i_next));
            }
        }
#endif
        if (live) {
            load_path(P, q_in, i, S);
            const float4 hr = __ldcs(hits + i);
            v3 L;
            next = shade_step<COUNT, HAS_SPH>(P, A, mats, S, hr.x, __float_as_uint(hr.y), L, n_scatter);
            if (!next) __stcs(rad + S.path_id, make_float4(L.x, L.y, L.z, 0.0f));
        }
        const uint32_t pos = block_append<RB>(next, n_out_ptr, s_cnt, it);
        if (next) {
#if CRT_QUEUE_LOCAL_DIR
            __stcs(q_out + pos, make_float4(S.o.x, S.o.y, S.o.z, S.dl.x));
            __stcs(q_out + (size_t)P.cap + pos, make_float4(S.dl.y, S.dl.z, __uint_as_float(S.path_id), __uint_as_float(S.pack)));
            __stcs(q_out + 4 * (size_t)P.cap + pos, make_float4(S.d.x, S.d.y, S.d.z, 0.0f));
#else
            __stcs(q_out + pos, make_float4(S.o.x, S.o.y, S.o.z, S.d.x));
            __stcs(q_out + (size_t)P.cap + pos, make_float4(S.d.y, S.d.z, __uint_as_float(S.path_id), __uint_as_float(S.pack)));
#endif
            __stcs(reinterpret_cast<uint4*>(q_out) + 2 * (size_t)P.cap + pos, make_uint4(S.ms[0], S.ms[1], S.ms[2], S.ms[3]));
            if (P.max_depth > 8)
                __stcs(reinterpret_cast<uint4*>(q_out) + 3 * (size_t)P.cap + pos, make_uint4(S.ms[4], S.ms[5], S.ms[6], S.ms[7]));
        }
    }
    if (COUNT) {
        const uint32_t sc = __reduce_add_sync(0xffffffffu, n_scatter);
        if (lane == 0) atomicAdd(counters + CNT_SCATTERS, (unsigned long long)sc);
    }
}

// ---- the tail of a slice: a queue of FEW paths is carried to termination by one launch ----
// The last bounces of a frame hold 20,000 paths down to a handful; as wavefront launches each of them is bound by
// its longest ray and by launch latency (bounces 3-8 of a C3 slice: 18 launches for 0.04 % of the rays). Here one
// WARP takes a path of the queue and runs shade()'s whole remaining recursion: closest hit with the 32 lanes
// working through the ray's stack abreast (crt_traverse.cuh trace_wide_one), then the material evaluation, which is
// warp-uniform (every lane computes the same values from the same draw counter), then the next ray, until the path
// ends. Nothing goes back to a queue. The arithmetic per path segment is shade_step and the primitive tests of the
// wavefront kernels, so radiance values are the same bit for bit.
template <bool COUNT, bool HAS_SPH>
__global__ void __launch_bounds__(TP_BLOCK)
k_tail(RenderParams P, AccelView A, const DMat* __restrict__ mats, const float4* __restrict__ q_in,
       const uint32_t* __restrict__ n_in_ptr, uint32_t* __restrict__ work_ctr, float4* __restrict__ rad,
       unsigned long long* __restrict__ counters, uint32_t tail_below)
{
    __shared__ uint2 s_stack[(TP_BLOCK / 32) * CRT_WIDE_STACK];
    const uint32_t n = *n_in_ptr;
    if (n == 0u || n >= tail_below) return;         // the wavefront kernels took (or take) this queue
    const uint32_t lane = threadIdx.x & 31u;
    uint2* stack = s_stack + (threadIdx.x >> 5) * CRT_WIDE_STACK;
    TraceCounters cnt = {0u, 0u, 0u, 0u, 0u, 0u, 0u};
    uint32_t n_scatter = 0, n_more = 0;             // warp-uniform: scatter evaluations, rays beyond each path's queued one
    for (;;) {
        uint32_t i = 0;
        if (lane == 0) i = atomicAdd(work_ctr, 1u);
        i = __shfl_sync(0xffffffffu, i, 0);
        if (i >= n) break;
        PathState S;
        load_path(P, q_in, i, S);                   // one address for the warp: broadcast loads
        for (;;) {
            const Hit h = trace_wide_one<COUNT, HAS_SPH>(A, S.o, S.d, 0.001f, FLT_MAX, stack, cnt);   // render.h:53
            v3 L;
            if (!shade_step<COUNT, HAS_SPH>(P, A, mats, S, h.t, h.k, L, n_scatter)) {
                if (lane == 0) __stcs(rad + S.path_id, make_float4(L.x, L.y, L.z, 0.0f));
                break;
            }
            n_more++;
        }
    }
    if (lane == 0 && n_more) atomicAdd(counters + CNT_RENDER_RAYS, (unsigned long long)n_more);
    if (COUNT) {
        const uint32_t nodes = __reduce_add_sync(0xffffffffu, cnt.nodes);
        const uint32_t tris = __reduce_add_sync(0xffffffffu, cnt.tris);
        const uint32_t sph = __reduce_add_sync(0xffffffffu, cnt.spheres);
        if (lane == 0) {
            atomicAdd(counters + CNT_NODES, (unsigned long long)nodes);
            atomicAdd(counters + CNT_TRIS, (unsigned long long)tris);
            atomicAdd(counters + CNT_SPHERES, (unsigned long long)sph);
            atomicAdd(counters + CNT_NODE_F, (unsigned long long)nodes);
            atomicAdd(counters + CNT_TRI_F, (unsigned long long)tris);
            atomicAdd(counters + CNT_SPH_F, (unsigned long long)sph);
            atomicAdd(counters + CNT_SCATTERS, (unsigned long long)n_scatter);
        }
    }
}

// Also the slice's tally: every queue entry was one closest-hit query, queue 0 holds the camera samples (the host
// never reads the queue lengths, so a render call does not have to wait for its launches).
__global__ void __launch_bounds__(256)
k_accumulate(RenderParams P, const float4* __restrict__ rad, float* __restrict__ accum, const uint32_t* __restrict__ qcount,
             int bounces, unsigned long long* __restrict__ counters, uint32_t wide_below, int forced_at, uint32_t hint_key,
             uint32_t* __restrict__ hint)
{
    if (blockIdx.x == 0 && (int)threadIdx.x < bounces) {
        const uint32_t q = qcount[threadIdx.x];
        if (q) atomicAdd(counters + CNT_RENDER_RAYS, (unsigned long long)q);
        if (threadIdx.x == 0 && q) atomicAdd(counters + CNT_RENDER_PATHS, (unsigned long long)q);
    }
    if (hint && blockIdx.x == 0 && threadIdx.x == 0) {
        // CRT_OPT_TAIL_HINT, for the next frame of this shape: the first bounce whose queue k_tail takes. This frame's
        // queue lengths are known up to the bounce where its own k_tail took over (forced_at, when the host guessed it)
        const int last = forced_at ? forced_at : bounces - 1;
        uint32_t h = 0u;
        for (int b = 1; b <= last && !h; b++)
            if (qcount[b] < wide_below) h = (uint32_t)b;
        // guessed too early: up to twice the threshold the guess stands, beyond that the next frame queues every bounce
        if (!h && forced_at && qcount[forced_at] < 2u * wide_below) h = (uint32_t)forced_at;
        *reinterpret_cast<volatile uint32_t*>(hint) = (hint_key << 8) | h;
    }
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= P.n_slots) return;
    int x, y;
    if (!slot_to_pixel(P, slot, x, y)) return;
    const size_t pixel = (size_t)y * P.nx + x;
    v3 col = V(accum[3 * pixel], accum[3 * pixel + 1], accum[3 * pixel + 2]);
    for (int s = 0; s < P.s_count; s++) {
        const float4 r = rad[make_path_id(P, slot, (uint32_t)s)];
        col = vadd(col, V(r.x, r.y, r.z));                                         // render.h:119-121
    }
    accum[3 * pixel] = col.x;
    accum[3 * pixel + 1] = col.y;
    accum[3 * pixel + 2] = col.z;
}

__global__ void __launch_bounds__(256)
k_resolve(const float* __restrict__ accum, size_t n_values, float inv_ns, float* __restrict__ rgb)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_values) return;
    rgb[i] = clip01(x::sqrt(x::mul(accum[i], inv_ns)));                            // render.h:123-128, vec3.h:182-189
}

// k_resolve over the image rows of the tile rows rank, rank + world, ... (the pixels of one slice in band mode)
__global__ void __launch_bounds__(256)
k_resolve_bands(const float* __restrict__ accum, int nx, int ny, int rank, int world, int n_bands, float inv_ns,
                float* __restrict__ rgb)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t row_values = (size_t)nx * 3;
    const size_t lr = i / row_values;
    if (lr >= (size_t)n_bands * CRT_TILE) return;
    const size_t row = ((lr / CRT_TILE) * (size_t)world + (size_t)rank) * CRT_TILE + lr % CRT_TILE;
    if (row >= (size_t)ny) return;
    const size_t g = row * row_values + (i - lr * row_values);
    rgb[g] = clip01(x::sqrt(x::mul(accum[g], inv_ns)));                            // render.h:123-128
}

__global__ void __launch_bounds__(256)
k_pack_rgba8(const float* __restrict__ rgb, int nx, int ny, uchar4* __restrict__ out)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)nx * ny) return;
    const int row = (int)(i / nx), col = (int)(i - (size_t)row * nx);
    const size_t src = (size_t)(ny - 1 - row) * nx + col;                          // render.h:140
    uchar4 o;
    o.x = (unsigned char)(int)(255.99 * (double)rgb[3 * src]);                     // render.h:141-143
    o.y = (unsigned char)(int)(255.99 * (double)rgb[3 * src + 1]);
    o.z = (unsigned char)(int)(255.99 * (double)rgb[3 * src + 2]);
    o.w = 255;
    out[i] = o;
}

static DCamera to_dcamera(const crt_camera* c)
{
    DCamera d;
    d.llc = V(c->lower_left_corner[0], c->lower_left_corner[1], c->lower_left_corner[2]);
    d.horizontal = V(c->horizontal[0], c->horizontal[1], c->horizontal[2]);
    d.vertical = V(c->vertical[0], c->vertical[1], c->vertical[2]);
    d.origin = V(c->origin[0], c->origin[1], c->origin[2]);
    d.x = V(c->x[0], c->x[1], c->x[2]);
    d.y = V(c->y[0], c->y[1], c->y[2]);
    d.z = V(c->z[0], c->z[1], c->z[2]);
    d.lens_radius = c->lens_radius;
    return d;
}

// persistent grid sizes, from the occupancy of one instantiation each (they have near-identical resource needs)
static void size_grids(crt_ctx* c)
{
    if (c->grid_trace) return;
    int bps = 1, bps_shade = 1, bps_packet = 1, bps_tail = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_trace_queue<false, false>, TP_BLOCK, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps_shade, k_shade<false, false>, RB, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps_packet, k_trace_packet<false, false, false>, TP_BLOCK, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps_tail, k_tail<false, false>, TP_BLOCK, 0);
    c->grid_trace = c->sm_count * std::max(1, bps);
    c->grid_shade = c->sm_count * std::max(1, bps_shade);
    c->grid_packet = c->sm_count * std::max(1, bps_packet);
    c->grid_tail = c->sm_count * std::max(1, bps_tail);
}

template <bool COUNT, bool HAS_SPH>
static void launch_bounce(crt_ctx* c, crt_lane& L, int b, const RenderParams& P, const AccelView& A,
                          const float4* q_in, const uint32_t* n_in, uint32_t* work_ctr, float4* q_out, uint32_t* n_out, int tail_at)
{
    unsigned long long* counters = c->counters.as<unsigned long long>();
    const bool packet = b == 0 && c->packet_primary;
    if (tail_at && b == tail_at) {   // CRT_OPT_TAIL_HINT: this queue goes to k_tail whatever its length
        k_tail<COUNT, HAS_SPH><<<c->grid_tail, TP_BLOCK, 0, L.stream>>>(P, A, c->mats.as<DMat>(), q_in, n_in, work_ctr + 32,
                                                                        L.rad.as<float4>(), counters, 0xffffffffu);
        c->host_counters.launches++;
        return;
    }
    // the queue length is only known on the device: from the second bounce on, a queue of fewer than tail_below
    // paths is left alone by the wavefront kernels (they return at once) and finished by k_tail. With a guess of the
    // bounce where that happens (tail_at) the bounces before it are plain wavefront launches
    const uint32_t tail_below = (b == 0 || tail_at) ? 0u : (uint32_t)c->wide_below;
    if (c->timing) cudaEventRecord(L.ev_trace[2 * b], L.stream);
    if (packet) { // camera rays: a warp holds the samples of one or two pixels
        // CRT_OPT_PACKET_DEFER: a packet stops after that many visits and its rays are traced one warp per ray by the
        // launch behind it. Only for slices of at most CRT_OPT_PACKET_DEFER_BELOW camera samples: there the launch is
        // as short as its longest packet; in a large slice the counting and the second launch cost more than they save
        PacketDefer defer = {nullptr, nullptr, 0u, 0u};
        const uint32_t list_cap = 1u << 18;
        if (c->packet_defer > 0 && (long long)P.n_slots * P.s_count <= c->packet_defer_below && L.deferred.reserve((size_t)list_cap * 4) == cudaSuccess)
            defer = PacketDefer{L.deferred.as<uint32_t>(), work_ctr + 64, list_cap, (uint32_t)c->packet_defer};
        if (defer.after) {
            k_trace_packet<COUNT, HAS_SPH, true><<<c->grid_packet, TP_BLOCK, 0, L.stream>>>(A, q_in, P.cap, n_in, work_ctr, L.hitbuf.as<float4>(),
                                                                                           counters, c->packet_primary == 1 ? 1 : 0, defer);
            k_trace_deferred<COUNT, HAS_SPH><<<c->grid_tail, TP_BLOCK, 0, L.stream>>>(A, q_in, P.cap, defer.list, defer.n, list_cap, work_ctr + 96,
                                                                                     L.hitbuf.as<float4>(), counters);
            c->host_counters.launches++;
        } else {
            k_trace_packet<COUNT, HAS_SPH, false><<<c->grid_packet, TP_BLOCK, 0, L.stream>>>(A, q_in, P.cap, n_in, work_ctr, L.hitbuf.as<float4>(),
                                                                                            counters, c->packet_primary == 1 ? 1 : 0, defer);
        }
    } else
        k_trace_queue<COUNT, HAS_SPH><<<c->grid_trace, TP_BLOCK, 0, L.stream>>>(A, q_in, P.cap, n_in, work_ctr, L.hitbuf.as<float4>(),
                                                                                counters, (uint32_t)c->refill, (uint32_t)c->leaf_thresh, tail_below);
    if (c->timing) cudaEventRecord(L.ev_trace[2 * b + 1], L.stream);
    if (COUNT && c->timing && getenv("CRT_TRACE_TIMING")) {   // development aid: longest dependent chain of the launch
        unsigned long long ms = 0;
        cudaMemcpyAsync(&ms, counters + CNT_MAX_STEPS, 8, cudaMemcpyDeviceToHost, L.stream);
        cudaMemsetAsync(counters + CNT_MAX_STEPS, 0, 8, L.stream);
        cudaStreamSynchronize(L.stream);
        fprintf(stderr, "[crt] bounce %d: longest ray took %llu traversal steps\n", b, ms);
    }
    k_shade<COUNT, HAS_SPH><<<c->grid_shade, RB, 0, L.stream>>>(P, A, c->mats.as<DMat>(), q_in, L.hitbuf.as<float4>(), n_in, q_out,
                                                                n_out, L.rad.as<float4>(), counters, tail_below);
    if (tail_below > 0u) {
        k_tail<COUNT, HAS_SPH><<<c->grid_tail, TP_BLOCK, 0, L.stream>>>(P, A, c->mats.as<DMat>(), q_in, n_in, work_ctr + 32,
                                                                        L.rad.as<float4>(), counters, tail_below);
        c->host_counters.launches++;
    }
}

// A frame is cut into jobs: sample batches (when all samples of the rank's
// pixels do not fit CRT_OPT_BATCH_PATHS) x n_lanes slices of the rank's tiles.
// The slices of one sample batch run concurrently, one per lane (stream): each
// is the full wavefront loop raygen -> (trace, shade) x bounces -> accumulate
// over its own queues; slices interleave tile by tile, so one slice's
// latency-bound last bounces and the drain phase of every persistent launch
// overlap the other slices' wide launches. Slices own disjoint pixels and a
// pixel's sample batches stay on one lane in order, so d_accum is summed in
// the same order for any lane count and the image does not depend on it.
//
// Nothing here waits for the device: queue lengths stay on the device (every
// kernel reads its own), the ray and path totals are tallied by k_accumulate,
// and the lanes' streams are joined into the context's stream with events, so
// the call returns as soon as its launches are queued and d_accum is valid in
// stream order on crt_stream(). Only CRT_OPT_TIMING (per-launch events for the
// roofline) makes it wait.
int render_accum(crt_ctx* c, const crt_camera* cam, int nx, int ny, int s0, int s1, int max_depth, int mode,
                 uint64_t seed, int rank, int world, float* d_accum, RenderFinish* finish, cudaEvent_t scene_ready)
{
    if (finish) finish->done = false;
    if (!cam || nx <= 0 || ny <= 0 || s1 < s0 || !d_accum) { set_error("crt_render: bad arguments"); return CRT_ERR_INVALID; }
    if (max_depth < 0 || max_depth > CRT_MAX_DEPTH) { set_error("crt_render: max_depth must be in [0, %d]", CRT_MAX_DEPTH); return CRT_ERR_INVALID; }
    if (mode < 0 || mode > 2) { set_error("crt_render: unknown shade mode %d", mode); return CRT_ERR_INVALID; }
    if (world < 1 || rank < 0 || rank >= world) { set_error("crt_render: bad rank/world %d/%d", rank, world); return CRT_ERR_INVALID; }
    if (!c->accel_valid) { set_error("crt_render: call crt_accel_build first"); return CRT_ERR_STATE; }
    if (c->n_prims > 0 && c->n_mats == 0) { set_error("crt_render: no materials set"); return CRT_ERR_STATE; }
    cudaStream_t st = c->stream;

    RenderParams P;
    P.cam = to_dcamera(cam);
    P.nx = nx; P.ny = ny;
    P.tiles_x = (nx + CRT_TILE - 1) / CRT_TILE;
    const int tiles_y = (ny + CRT_TILE - 1) / CRT_TILE;
    const long long tiles = (long long)P.tiles_x * tiles_y;
    // a frame that goes back to the caller's host buffer: its slices own whole tile rows, so that each can hand over
    // its finished rows while the others still render. Otherwise slices and ranks interleave tile by tile (balances
    // a little better: 9.28 against 9.36 ms on C3)
    P.band = (finish && world == 1) ? 1 : 0;
    const long long units = P.band ? tiles_y : tiles;              // what is dealt out to ranks and slices
    const long long unit_tiles = P.band ? P.tiles_x : 1;
    const long long local_tiles = units > rank ? (units - rank + world - 1) / world * unit_tiles : 0;
    P.rank = rank; P.world = world; P.tile_base = 0u;
    if (local_tiles * CRT_TILE * CRT_TILE >= (1ll << 31)) { set_error("crt_render: image too large"); return CRT_ERR_INVALID; }
    P.max_depth = max_depth; P.mode = mode; P.seed = seed; P.seed_key = crt_rng_seed_key(seed);

    CRT_CUDA(cudaEventRecord(c->ev_r0, st));
    CRT_CUDA(cudaMemsetAsync(d_accum, 0, (size_t)nx * ny * 3 * sizeof(float), st));
    const int n_samples = s1 - s0;
    float bounce_ms = 0.f, trace_q_ms = 0.f;
    if (local_tiles > 0 && n_samples > 0) {
        // lane l of rank r renders tiles r + world * (l + n_lanes * j): the slices interleave, so they are balanced
        // at most CRT_OPT_LANES, at least 2 M camera samples per slice: below that a frame is launch bound and more
        // launches cost more than the overlap gains (C1, 2.1 M samples in 0.8 ms: 1 slice 0.79 ms, 4 slices 0.91 ms;
        // C3 at 960x540x16, 8.3 M samples: 3.91 ms -> 3.35 ms)
        const long long want = (local_tiles * CRT_TILE * CRT_TILE * n_samples) / c->slice_min_samples;
        const int n_lanes = (int)std::max(1ll, std::min((long long)c->n_lanes, std::min(local_tiles / unit_tiles, want)));
        c->host_counters.lanes = (uint64_t)n_lanes;
        const int sub_world = world * n_lanes;
        auto sub_units = [&](int l) {
            const long long sr = (long long)l * world + rank;
            return units > sr ? (units - sr + sub_world - 1) / sub_world : 0ll;
        };
        auto sub_tiles = [&](int l) { return sub_units(l) * unit_tiles; };
        // A frame with more camera samples than batch_paths runs as several jobs per slice. CRT_OPT_BATCH_SAMPLES > 0:
        // the slice's tiles are cut into groups, and a group is rendered with up to that many samples per pixel at once,
        // so a warp of camera rays still holds the samples of one or two pixels and walks the tree as one packet (a 4K x
        // 64 spp frame on one GPU used to run as 16 batches of FOUR samples over every tile: eight pixels per warp).
        // 0: sample batches over all of the slice's tiles. Either way a pixel's jobs follow each other in sample order on
        // one stream, so d_accum is summed in the same order and the image does not depend on the cut.
        const long long lane_tiles0 = sub_tiles(0);                                      // lane 0 has the most tiles
        const long long slot_sz = (long long)CRT_TILE * CRT_TILE;
        const long long lane_paths = std::max(1ll, (long long)(c->batch_paths / n_lanes));
        long long per_batch, tiles_job;                                                  // samples per job, tiles per job and slice
        if (c->batch_samples > 0 && lane_tiles0 * slot_sz * n_samples > lane_paths) {
            per_batch = std::min((long long)n_samples, (long long)c->batch_samples);
            per_batch = std::max(1ll, std::min(per_batch, lane_paths / slot_sz));
            tiles_job = std::max(1ll, std::min(lane_tiles0, lane_paths / (slot_sz * per_batch)));
        } else {
            per_batch = c->batch_paths / (local_tiles * slot_sz);                        // samples of all slices together
            if (per_batch < 1) per_batch = 1;
            if (per_batch > n_samples) per_batch = n_samples;
            tiles_job = lane_tiles0;
        }
        while (tiles_job * slot_sz * per_batch >= (1ll << 31) && per_batch > 1) per_batch--;
        while (tiles_job * slot_sz * per_batch >= (1ll << 31) && tiles_job > 1) tiles_job--;
        const size_t cap = (size_t)(tiles_job * slot_sz) * (size_t)per_batch;
        P.cap = (uint32_t)cap;
        const int n_levels = CRT_MAX_DEPTH + 3;
        for (int l = 0; l < n_lanes; l++) {
            crt_lane& L = c->lane[l];
            CRT_CUDA(L.queue[0].reserve(cap * 16 * CRT_QUEUE_WORDS));
            CRT_CUDA(L.queue[1].reserve(cap * 16 * CRT_QUEUE_WORDS));
            CRT_CUDA(L.rad.reserve(cap * 16));
            CRT_CUDA(L.qcount.reserve(sizeof(uint32_t) * n_levels));
            CRT_CUDA(L.work_ctr.reserve(sizeof(uint32_t) * 128));
            CRT_CUDA(L.hitbuf.reserve(cap * 16));
        }
        size_grids(c);
        const AccelView A = accel_view(c);
        const bool sph = scene_is_general(c);
        const int bounces = mode == CRT_SHADE_PATH ? max_depth + 1 : 1;
        static const bool verbose = getenv("CRT_TRACE_TIMING") != nullptr;   // development aid

        // lanes 1.. start after what the caller queued on the context's stream (uploads, refit, the memset above)
        CRT_CUDA(cudaEventRecord(c->ev_fork, st));
        for (int l = 1; l < n_lanes; l++) CRT_CUDA(cudaStreamWaitEvent(c->lane[l].stream, c->ev_fork, 0));

        for (int sb = s0; sb < s1; sb += (int)per_batch) {
          P.s_begin = sb;
          P.s_count = (s1 - sb) < per_batch ? (s1 - sb) : (int)per_batch;
          for (long long g0 = 0; g0 < lane_tiles0; g0 += tiles_job) {
            for (int l = 0; l < n_lanes; l++) {
                crt_lane& L = c->lane[l];
                L.h_counts[0] = 0;
                if (sub_tiles(l) <= g0) continue;
                const bool first_job = sb == s0 && g0 == 0;
                const bool last_job = sb + P.s_count >= s1 && g0 + tiles_job >= sub_tiles(l);   // of this slice
                RenderParams PL = P;
                PL.world = sub_world;
                PL.rank = l * world + rank;
                PL.tile_base = (uint32_t)g0;
                PL.n_slots = (uint32_t)(std::min(tiles_job, sub_tiles(l) - g0) * slot_sz);
                CRT_CUDA(cudaMemsetAsync(L.qcount.p, 0, sizeof(uint32_t) * n_levels, L.stream));
                CRT_CUDA(cudaMemsetAsync(L.work_ctr.p, 0, sizeof(uint32_t) * 128, L.stream));   // [b] queue cursors, [32 + b] tail, [64 + b] / [96 + b] deferred rays: count, cursor
                const uint32_t total = PL.n_slots * (uint32_t)PL.s_count;
                const unsigned int rg = (unsigned int)min((size_t)(total + RB - 1) / RB, (size_t)c->sm_count * 32);
                k_raygen<<<rg, RB, 0, L.stream>>>(PL, L.queue[0].as<float4>(), L.qcount.as<uint32_t>());
                // CRT_OPT_TAIL_HINT: where the last frame of this shape handed over to k_tail (crt_lane::tail_hint)
                const uint32_t hint_key = ((uint32_t)nx * 0x9e3779b1u ^ (uint32_t)ny * 0x85ebca6bu ^ (uint32_t)PL.s_count * 0xc2b2ae35u ^
                                           (uint32_t)max_depth * 0x27d4eb2fu ^ (uint32_t)PL.world * 0x165667b1u ^ (uint32_t)PL.rank * 0xd3a2646cu ^
                                           (uint32_t)c->n_prims * 0xfd7046c5u ^ (uint32_t)c->wide_below * 0xb55a4f09u ^ PL.tile_base * 0x2545f491u ^
                                           PL.n_slots * 0x9fb21c65u) & 0xffffffu;
                // a frame cut into tile groups: the groups show different parts of the scene, so each keeps its own hint
                // (one group's sky would send the next group's million bounce rays to the one-warp-per-path launch)
                const uint32_t hint_slot = (uint32_t)((g0 / tiles_job) % CRT_TAIL_HINTS);
                int tail_at = 0;
                const bool hinting = c->tail_hint && !c->timing && c->wide_below > 0 && bounces > 2;
                if (hinting) {
                    const uint32_t hv = *reinterpret_cast<volatile uint32_t*>(L.tail_hint + hint_slot);
                    if ((hv >> 8) == hint_key && (int)(hv & 0xffu) < bounces) tail_at = (int)(hv & 0xffu);
                }
                const int n_bounces = tail_at ? tail_at + 1 : bounces;
                c->host_counters.launches += 2 + 2 * (uint64_t)(tail_at ? tail_at : bounces);   // raygen, (trace, shade) per bounce, accumulate
                if (c->timing) CRT_CUDA(cudaEventRecord(L.ev_begin, L.stream));
                // ray generation does not read the scene: only the closest-hit launches wait for an upload and refit
                // queued on another stream (crt_render_frame)
                if (scene_ready && first_job) CRT_CUDA(cudaStreamWaitEvent(L.stream, scene_ready, 0));
                for (int b = 0; b < n_bounces; b++) {
                    const float4* qi = L.queue[b & 1].as<float4>();
                    float4* qo = L.queue[(b + 1) & 1].as<float4>();
                    const uint32_t* ni = L.qcount.as<uint32_t>() + b;
                    uint32_t* no = L.qcount.as<uint32_t>() + b + 1;
                    uint32_t* wc = L.work_ctr.as<uint32_t>() + b;
                    if (c->count_work) {
                        if (sph) launch_bounce<true, true>(c, L, b, PL, A, qi, ni, wc, qo, no, tail_at);
                        else launch_bounce<true, false>(c, L, b, PL, A, qi, ni, wc, qo, no, tail_at);
                    } else {
                        if (sph) launch_bounce<false, true>(c, L, b, PL, A, qi, ni, wc, qo, no, tail_at);
                        else launch_bounce<false, false>(c, L, b, PL, A, qi, ni, wc, qo, no, tail_at);
                    }
                }
                if (c->timing) CRT_CUDA(cudaEventRecord(L.ev_end, L.stream));
                k_accumulate<<<(PL.n_slots + 255) / 256, 256, 0, L.stream>>>(PL, L.rad.as<float4>(), d_accum, L.qcount.as<uint32_t>(), bounces,
                                                                            c->counters.as<unsigned long long>(), (uint32_t)c->wide_below, tail_at,
                                                                            hint_key, hinting ? L.tail_hint_dev + hint_slot : nullptr);
                if (finish && P.band && last_job) {
                    // the slice's pixels are final: resolve its rows and start their copy to the caller's buffer while
                    // the other slices still render (rows of tile rows PL.rank, PL.rank + sub_world, ...)
                    const int n_bands = (int)sub_units(l);
                    const size_t band_values = (size_t)CRT_TILE * nx * 3, band_bytes = band_values * sizeof(float);
                    const size_t nv = (size_t)n_bands * band_values;
                    k_resolve_bands<<<(unsigned int)((nv + 255) / 256), 256, 0, L.stream>>>(d_accum, nx, ny, PL.rank, sub_world, n_bands,
                                                                                          1.0f / (float)finish->spp, finish->d_rgb);
                    c->host_counters.launches++;
                    int n_full = 0;                                                // bands with all CRT_TILE rows inside the image
                    while (n_full < n_bands && ((long long)n_full * sub_world + PL.rank + 1) * CRT_TILE <= ny) n_full++;
                    const size_t off = (size_t)PL.rank * band_values;
                    if (n_full > 0)
                        CRT_CUDA(cudaMemcpy2DAsync(finish->h_rgb + off, band_bytes * sub_world, finish->d_rgb + off, band_bytes * sub_world,
                                                   band_bytes, (size_t)n_full, cudaMemcpyDeviceToHost, L.stream));
                    if (n_full < n_bands) {                                        // the image's last rows
                        const size_t row0 = ((size_t)n_full * sub_world + PL.rank) * CRT_TILE;
                        if (row0 < (size_t)ny)
                            CRT_CUDA(cudaMemcpyAsync(finish->h_rgb + row0 * nx * 3, finish->d_rgb + row0 * nx * 3,
                                                     ((size_t)ny - row0) * nx * 3 * sizeof(float), cudaMemcpyDeviceToHost, L.stream));
                    }
                    finish->done = true;
                }
                if (c->timing) CRT_CUDA(cudaMemcpyAsync(L.h_counts, L.qcount.p, sizeof(uint32_t) * n_levels, cudaMemcpyDeviceToHost, L.stream));
            }
            if (!c->timing) continue;
            // CRT_OPT_TIMING: per-launch durations of this batch (the roofline's denominator)
            float batch_ms = 0.f;
            for (int l = 0; l < n_lanes; l++) {
                crt_lane& L = c->lane[l];
                CRT_CUDA(cudaStreamSynchronize(L.stream));
                if (L.h_counts[0] == 0) continue;
                float bms = 0.f;
                CRT_CUDA(cudaEventElapsedTime(&bms, L.ev_begin, L.ev_end));
                batch_ms = bms > batch_ms ? bms : batch_ms;
                for (int b = 0; b < bounces; b++) {
                    float tms = 0.f;
                    CRT_CUDA(cudaEventElapsedTime(&tms, L.ev_trace[2 * b], L.ev_trace[2 * b + 1]));
                    trace_q_ms += tms;
                    if (verbose) fprintf(stderr, "[crt] lane %d bounce %d: %u rays, trace %.3f ms\n", l, b, L.h_counts[b], tms);
                }
                if (verbose) fprintf(stderr, "[crt] lane %d: bounces %.3f ms\n", l, bms);
            }
            bounce_ms += batch_ms;
          }
        }
        // the other lanes' work joins the context's stream: whatever the caller queues there next (a reduce, a
        // resolve, a copy) is ordered behind the whole frame
        for (int l = 1; l < n_lanes; l++) {
            CRT_CUDA(cudaEventRecord(c->lane[l].ev_join, c->lane[l].stream));
            CRT_CUDA(cudaStreamWaitEvent(st, c->lane[l].ev_join, 0));
        }
    }
    CRT_CUDA(cudaEventRecord(c->ev_r1, st));
    CRT_CUDA(cudaGetLastError());
    c->render_ms_pending = true;
    c->host_counters.bounce_ms += bounce_ms;
    c->host_counters.trace_queue_ms += trace_q_ms;
    return CRT_OK;
}

int launch_resolve(crt_ctx* c, const float* d_accum, size_t n_pixels, int spp, float* d_rgb)
{
    if (spp < 1) { set_error("crt_resolve: spp must be >= 1"); return CRT_ERR_INVALID; }
    const size_t nv = n_pixels * 3;
    if (nv == 0) return CRT_OK;
    const float inv_ns = 1.0f / (float)spp;   // float(1.0 / t), vec3.h:183 (correctly rounded either way)
    k_resolve<<<(unsigned int)((nv + 255) / 256), 256, 0, c->stream>>>(d_accum, nv, inv_ns, d_rgb);
    c->host_counters.launches++;
    CRT_CUDA(cudaGetLastError());
    return CRT_OK;
}

int launch_pack_rgba8(crt_ctx* c, const float* d_rgb, int nx, int ny, uint8_t* d_rgba8)
{
    const size_t n = (size_t)nx * ny;
    if (n == 0) return CRT_OK;
    k_pack_rgba8<<<(unsigned int)((n + 255) / 256), 256, 0, c->stream>>>(d_rgb, nx, ny, reinterpret_cast<uchar4*>(d_rgba8));
    c->host_counters.launches++;
    CRT_CUDA(cudaGetLastError());
    return CRT_OK;
}

}  // namespace crt
