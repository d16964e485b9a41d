// trace.cu — closest-hit kernels behind crt_trace / crt_trace_device:
//   k_trace_lbvh  : one ray per thread through the LBVH (crt_traverse.cuh)
//   k_trace_brute : the reference's linear scan (hitable/hitable_list.h:76-91)
//                   over the primitives exactly as uploaded; primitives are
//                   staged through shared memory and every thread of the block
//                   tests them against its own ray. It shares no code with the
//                   LBVH build or traversal, which makes it the on-GPU cross
//                   check of both, and it serves scenes before crt_accel_build.
#include "crt_internal.h"
#include "crt_traverse.cuh"

namespace crt {

constexpr int TRACE_BLOCK = TP_BLOCK;

AccelView accel_view(const crt_ctx* c)
{
    AccelView a;
    a.nodes = c->nodes.as<float4>();
    a.recs = c->recs.as<float4>();
    a.frames = c->frames.as<DFrame>();
    a.root = c->root;
    a.n_prims = c->n_prims;
    return a;
}

__device__ __forceinline__ void load_ray(const crt_ray* __restrict__ rays, size_t i, v3& o, v3& d)
{
    const float2* p = reinterpret_cast<const float2*>(rays + i);   // 24-byte records, 8-byte aligned
    const float2 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    o = V(a.x, a.y, b.x);
    d = V(b.y, c.x, c.y);
}

__device__ __forceinline__ void flush_counters(unsigned long long* counters, const TraceCounters& cnt, uint32_t rays)
{
    const unsigned int act = __activemask();
    const uint32_t nodes = __reduce_add_sync(act, cnt.nodes);
    const uint32_t tris = __reduce_add_sync(act, cnt.tris);
    const uint32_t sph = __reduce_add_sync(act, cnt.spheres);
    const uint32_t r = __reduce_add_sync(act, rays);
    if ((threadIdx.x & 31) == (unsigned int)(__ffs(act) - 1)) {
        atomicAdd(counters + CNT_NODES, (unsigned long long)nodes);
        atomicAdd(counters + CNT_TRIS, (unsigned long long)tris);
        atomicAdd(counters + CNT_SPHERES, (unsigned long long)sph);
        atomicAdd(counters + CNT_NODE_F, (unsigned long long)nodes);
        atomicAdd(counters + CNT_TRI_F, (unsigned long long)tris);
        atomicAdd(counters + CNT_SPH_F, (unsigned long long)sph);
        atomicAdd(counters + CNT_RAYS, (unsigned long long)r);
    }
}

struct RaySrc {
    const crt_ray* rays;
    __device__ __forceinline__ void load(uint32_t i, v3& o, v3& d) const { load_ray(rays, i, o, d); }
};
struct HitSink {
    crt_hit* hits;
    __device__ __forceinline__ void store(uint32_t i, const Hit& h) const
    {
        const bool hit = h.k != 0xFFFFFFFFu;
        reinterpret_cast<float4*>(hits)[i] =
            make_float4(hit ? h.t : 0.0f, __uint_as_float(hit ? h.id : CRT_MISS_ID), hit ? h.u : 0.0f, hit ? h.v : 0.0f);
    }
};

template <bool COUNT, bool HAS_SPH>
__global__ void __launch_bounds__(TP_BLOCK, CRT_TRACE_MIN_BLOCKS)
k_trace_lbvh(AccelView A, const crt_ray* __restrict__ rays, uint32_t n, uint32_t* __restrict__ work_ctr, float t_min,
             float t_max, crt_hit* __restrict__ hits, unsigned long long* __restrict__ counters, uint32_t refill, uint32_t leaf_thresh)
{
    __shared__ int s_stack[CRT_STACK_WORDS * CRT_STACK_SMEM * TP_BLOCK];
    TraceCounters cnt = {0u, 0u, 0u, 0u, 0u, 0u, 0u};
    RaySrc src = {rays};
    HitSink sink = {hits};
    traverse_persistent<COUNT, HAS_SPH>(A, n, work_ctr, t_min, t_max, src, sink, s_stack, cnt, refill, leaf_thresh);
    if (COUNT) flush_counters(counters, cnt, 0u);
}

template <bool COUNT, bool HAS_SPH>
__global__ void __launch_bounds__(TP_BLOCK, CRT_PACKET_MIN_BLOCKS)
k_trace_lbvh_packet(AccelView A, const crt_ray* __restrict__ rays, uint32_t n, uint32_t* __restrict__ work_ctr, float t_min,
                    float t_max, crt_hit* __restrict__ hits, unsigned long long* __restrict__ counters)
{
    TraceCounters cnt = {0u, 0u, 0u, 0u, 0u, 0u, 0u};
    RaySrc src = {rays};
    HitSink sink = {hits};
    traverse_packets<COUNT, HAS_SPH>(A, n, work_ctr, t_min, t_max, src, sink, cnt);
    if (COUNT) {   // tests per ray, fetches per packet (kept by lane 0)
        const uint32_t nodes = __reduce_add_sync(0xffffffffu, cnt.nodes), tris = __reduce_add_sync(0xffffffffu, cnt.tris);
        const uint32_t sph = __reduce_add_sync(0xffffffffu, cnt.spheres);
        const uint32_t nf = __reduce_add_sync(0xffffffffu, cnt.node_f), tf = __reduce_add_sync(0xffffffffu, cnt.tri_f);
        const uint32_t sf = __reduce_add_sync(0xffffffffu, cnt.sph_f);
        if ((threadIdx.x & 31u) == 0u) {
            atomicAdd(counters + CNT_NODES, (unsigned long long)nodes);
            atomicAdd(counters + CNT_TRIS, (unsigned long long)tris);
            atomicAdd(counters + CNT_SPHERES, (unsigned long long)sph);
            atomicAdd(counters + CNT_NODE_F, (unsigned long long)nf);
            atomicAdd(counters + CNT_TRI_F, (unsigned long long)tf);
            atomicAdd(counters + CNT_SPH_F, (unsigned long long)sf);
        }
    }
}

constexpr int BRUTE_TILE = 64;

// chains / rectangles (BruteExtra): per-primitive chain words in input order (null = none of that kind) and the
// rectangles' material word (CRT_RECT_FLIP), id and chain
struct BruteExtra {
    const DFrame* frames;
    const uint32_t *tri_chain, *sph_chain, *rect_chain, *rect_mat, *rect_id;
    uint32_t n_rect;
};

template <bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK)
k_trace_brute(const float* __restrict__ tri_in, const uint32_t* __restrict__ tri_id, uint32_t n_tri,
              const float* __restrict__ sph_in, const uint32_t* __restrict__ sph_id, uint32_t n_sph, BruteExtra X,
              const crt_ray* __restrict__ rays, size_t n, float t_min, float t_max, crt_hit* __restrict__ hits,
              unsigned long long* __restrict__ counters)
{
    __shared__ float s_prim[BRUTE_TILE * 12];
    __shared__ uint32_t s_id[BRUTE_TILE];
    __shared__ uint32_t s_chain[BRUTE_TILE];
    const size_t i = (size_t)blockIdx.x * TRACE_BLOCK + threadIdx.x;
    const bool live = i < n;
    v3 o = V(0.f, 0.f, 0.f), d = V(0.f, 0.f, 1.f), d_raw = d;
    if (live) { load_ray(rays, i, o, d_raw); d = ray_dir_local(d_raw); }
    float best_t = t_max, best_u = 0.f, best_v = 0.f;
    uint32_t best_id = CRT_MISS_ID;
    TraceCounters cnt = {0u, 0u, 0u, 0u, 0u, 0u, 0u};

    for (uint32_t base = 0; base < n_tri; base += BRUTE_TILE) {
        const uint32_t m = min((uint32_t)BRUTE_TILE, n_tri - base);
        __syncthreads();
        for (uint32_t j = threadIdx.x; j < m * 12; j += TRACE_BLOCK) s_prim[j] = tri_in[(size_t)base * 12 + j];
        for (uint32_t j = threadIdx.x; j < m; j += TRACE_BLOCK) { s_id[j] = tri_id[base + j]; s_chain[j] = X.tri_chain ? X.tri_chain[base + j] : 0u; }
        __syncthreads();
        if (!live) continue;
        for (uint32_t j = 0; j < m; j++) {
            const float* t = s_prim + 12 * j;
            const v3 v0 = V(t[0], t[1], t[2]);
            const v3 e1 = vsub(V(t[3], t[4], t[5]), v0);
            const v3 e2 = vsub(V(t[6], t[7], t[8]), v0);
            float tt, u, v;
            if (COUNT) cnt.tris++;
            v3 ol = o, dl = d;
            if (s_chain[j]) chain_ray(X.frames, s_chain[j], o, d_raw, ol, dl);
            if (!triangle_test(ol, dl, v0, e1, e2, V(t[9], t[10], t[11]), t_min, best_t, tt, u, v)) continue;
            if (tt == best_t && (best_id == CRT_MISS_ID || s_id[j] >= best_id)) continue;
            best_t = tt; best_u = u; best_v = v; best_id = s_id[j];
        }
    }
    for (uint32_t base = 0; base < n_sph; base += BRUTE_TILE) {
        const uint32_t m = min((uint32_t)BRUTE_TILE, n_sph - base);
        __syncthreads();
        for (uint32_t j = threadIdx.x; j < m * 4; j += TRACE_BLOCK) s_prim[j] = sph_in[(size_t)base * 4 + j];
        for (uint32_t j = threadIdx.x; j < m; j += TRACE_BLOCK) { s_id[j] = sph_id[base + j]; s_chain[j] = X.sph_chain ? X.sph_chain[base + j] : 0u; }
        __syncthreads();
        if (!live) continue;
        for (uint32_t j = 0; j < m; j++) {
            const float* s = s_prim + 4 * j;
            float tt;
            if (COUNT) cnt.spheres++;
            v3 ol = o, dl = d;
            if (s_chain[j]) chain_ray(X.frames, s_chain[j], o, d_raw, ol, dl);
            if (!sphere_test(vsub(ol, V(s[0], s[1], s[2])), dl, s[3], t_min, best_t, tt)) continue;
            if (tt == best_t && (best_id == CRT_MISS_ID || s_id[j] >= best_id)) continue;
            best_t = tt; best_u = 0.f; best_v = 0.f; best_id = s_id[j];
        }
    }
    if (live) {
        for (uint32_t j = 0; j < X.n_rect; j++) {             // shapes/rectangle.h: few, read straight from global memory
            float tt, u, v;
            if (COUNT) cnt.spheres++;
            v3 ol, dl;
            chain_ray(X.frames, X.rect_chain[j], o, d_raw, ol, dl);
            if (!rect_test(ol, dl, (X.rect_mat[j] & CRT_RECT_FLIP) != 0u, t_min, best_t, tt, u, v)) continue;
            if (tt == best_t && (best_id == CRT_MISS_ID || X.rect_id[j] >= best_id)) continue;
            best_t = tt; best_u = u; best_v = v; best_id = X.rect_id[j];
        }
        const bool hit = best_id != CRT_MISS_ID;
        reinterpret_cast<float4*>(hits)[i] =
            make_float4(hit ? best_t : 0.0f, __uint_as_float(best_id), hit ? best_u : 0.0f, hit ? best_v : 0.0f);
    }
    if (COUNT) flush_counters(counters, cnt, live ? 1u : 0u);
}

int launch_trace(crt_ctx* c, const crt_ray* d_rays, size_t n, float t_min, float t_max, int accel, crt_hit* d_hits)
{
    if (n == 0) return CRT_OK;
    unsigned long long* counters = c->counters.as<unsigned long long>();
    cudaStream_t st = c->stream;
    if (accel == CRT_ACCEL_BRUTE) {
        const unsigned int grid = (unsigned int)((n + TRACE_BLOCK - 1) / TRACE_BLOCK);
        BruteExtra X;
        X.frames = c->frames.as<DFrame>();
        X.tri_chain = c->tri_has_chain ? c->tri_frame.as<uint32_t>() : nullptr;
        X.sph_chain = c->sph_has_chain ? c->sph_frame.as<uint32_t>() : nullptr;
        X.rect_chain = c->rect_frame.as<uint32_t>(); X.rect_mat = c->rect_mat.as<uint32_t>(); X.rect_id = c->rect_id.as<uint32_t>();
        X.n_rect = (uint32_t)c->n_rect;
        if (c->count_work)
            k_trace_brute<true><<<grid, TRACE_BLOCK, 0, st>>>(c->tri_in.as<float>(), c->tri_id.as<uint32_t>(), (uint32_t)c->n_tri,
                                                              c->sph_in.as<float>(), c->sph_id.as<uint32_t>(), (uint32_t)c->n_sph, X,
                                                              d_rays, n, t_min, t_max, d_hits, counters);
        else
            k_trace_brute<false><<<grid, TRACE_BLOCK, 0, st>>>(c->tri_in.as<float>(), c->tri_id.as<uint32_t>(), (uint32_t)c->n_tri,
                                                               c->sph_in.as<float>(), c->sph_id.as<uint32_t>(), (uint32_t)c->n_sph, X,
                                                               d_rays, n, t_min, t_max, d_hits, counters);
    } else if (accel == CRT_ACCEL_LBVH || accel == CRT_ACCEL_LBVH_PACKET) {
        if (!c->accel_valid) { set_error("crt_trace: call crt_accel_build first"); return CRT_ERR_STATE; }
        if (n >= (1ull << 32) - TP_CHUNK * 4096ull) { set_error("crt_trace: at most 2^32 - 2^20 rays per call"); return CRT_ERR_INVALID; }
        const AccelView A = accel_view(c);
        CRT_CUDA(c->work_ctr.reserve(sizeof(uint32_t) * 64));
        CRT_CUDA(cudaMemsetAsync(c->work_ctr.p, 0, sizeof(uint32_t), st));
        int bps = 1;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_trace_lbvh<false, false>, TP_BLOCK, 0);
        if (bps < 1) bps = 1;
        const size_t want = (n + 31) / 32;                           // warps that can get a (smallest) chunk
        size_t blocks = (want + (TP_BLOCK / 32) - 1) / (TP_BLOCK / 32);
        const size_t cap = (size_t)c->sm_count * bps;
        const unsigned int grid = (unsigned int)(blocks < cap ? blocks : cap);
        const bool sph = scene_is_general(c);
        uint32_t* wc = c->work_ctr.as<uint32_t>();
        const uint32_t un = (uint32_t)n;
        if (accel == CRT_ACCEL_LBVH_PACKET) {
            // the packet's farthest closest hit is reduced over the raw bits of positive distances
            if (t_min < 0.0f) { set_error("crt_trace: CRT_ACCEL_LBVH_PACKET needs t_min >= 0"); return CRT_ERR_INVALID; }
            int bpp = 1;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bpp, k_trace_lbvh_packet<false, false>, TP_BLOCK, 0);
            const size_t capp = (size_t)c->sm_count * (bpp < 1 ? 1 : bpp);
            const unsigned int grid = (unsigned int)(blocks < capp ? blocks : capp);
            if (c->count_work) {
                if (sph) k_trace_lbvh_packet<true, true><<<grid, TP_BLOCK, 0, st>>>(A, d_rays, un, wc, t_min, t_max, d_hits, counters);
                else k_trace_lbvh_packet<true, false><<<grid, TP_BLOCK, 0, st>>>(A, d_rays, un, wc, t_min, t_max, d_hits, counters);
            } else {
                if (sph) k_trace_lbvh_packet<false, true><<<grid, TP_BLOCK, 0, st>>>(A, d_rays, un, wc, t_min, t_max, d_hits, counters);
                else k_trace_lbvh_packet<false, false><<<grid, TP_BLOCK, 0, st>>>(A, d_rays, un, wc, t_min, t_max, d_hits, counters);
            }
        } else if (c->count_work) {
            if (sph) k_trace_lbvh<true, true><<<grid, TP_BLOCK, 0, st>>>(A, d_rays, un, wc, t_min, t_max, d_hits, counters, (uint32_t)c->refill, (uint32_t)c->leaf_thresh);
            else k_trace_lbvh<true, false><<<grid, TP_BLOCK, 0, st>>>(A, d_rays, un, wc, t_min, t_max, d_hits, counters, (uint32_t)c->refill, (uint32_t)c->leaf_thresh);
        } else {
            if (sph) k_trace_lbvh<false, true><<<grid, TP_BLOCK, 0, st>>>(A, d_rays, un, wc, t_min, t_max, d_hits, counters, (uint32_t)c->refill, (uint32_t)c->leaf_thresh);
            else k_trace_lbvh<false, false><<<grid, TP_BLOCK, 0, st>>>(A, d_rays, un, wc, t_min, t_max, d_hits, counters, (uint32_t)c->refill, (uint32_t)c->leaf_thresh);
        }
    } else {
        set_error("crt_trace: unknown accel kind %d", accel);
        return CRT_ERR_INVALID;
    }
    CRT_CUDA(cudaGetLastError());
    return CRT_OK;
}

// ---- crt_trace_static: the host-buffer closest-hit call with its device scratch in a static array ----
// (no counterpart in the reference, see include/crt.h)
// Layout of the array: [0, 256) the persistent kernel's queue cursor (one uint32 in use), then MAX_RAYS staged rays
// (24 B each), then MAX_RAYS hit records (16 B each); 256 + 65536 * 24 is a multiple of 16, so the float4 stores of
// the hit records stay aligned.
static __device__ __align__(256) unsigned char g_trace_static_scratch[CRT_TRACE_STATIC_SCRATCH_BYTES];

// In host code the array's name is a host-side stand-in, not a device address: the address every GPU gives its own
// copy comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int TRACE_STATIC_MAX_DEVICES = 64;
static unsigned char* g_trace_static_ptr[TRACE_STATIC_MAX_DEVICES] = {};

int trace_static(crt_ctx* c, const crt_ray* rays, size_t n, float t_min, float t_max, crt_hit* hits)
{
    if (n == 0) return CRT_OK;
    if (!rays || !hits) { set_error("crt_trace_static: null buffer"); return CRT_ERR_INVALID; }
    if (n > (size_t)CRT_TRACE_STATIC_MAX_RAYS) {
        set_error("crt_trace_static: %zu rays, but the static scratch holds at most %d per call (split the batch, or use crt_trace)",
                  n, (int)CRT_TRACE_STATIC_MAX_RAYS);
        return CRT_ERR_INVALID;
    }
    if (!c->accel_valid) { set_error("crt_trace_static: call crt_accel_build first"); return CRT_ERR_STATE; }
    if (c->device < 0 || c->device >= TRACE_STATIC_MAX_DEVICES) { set_error("crt_trace_static: device %d beyond %d", c->device, TRACE_STATIC_MAX_DEVICES); return CRT_ERR_INVALID; }
    if (!g_trace_static_ptr[c->device]) {            // set-up, once per device; the caller made c->device current
        void* p = nullptr;
        CRT_CUDA(cudaGetSymbolAddress(&p, g_trace_static_scratch));
        if (!p) { set_error("crt_trace_static: no device address for the scratch array"); return CRT_ERR_CUDA; }
        g_trace_static_ptr[c->device] = (unsigned char*)p;
    }
    unsigned char* base = g_trace_static_ptr[c->device];
    uint32_t* cursor = reinterpret_cast<uint32_t*>(base);
    crt_ray* d_rays = reinterpret_cast<crt_ray*>(base + 256);
    crt_hit* d_hits = reinterpret_cast<crt_hit*>(base + 256 + (size_t)CRT_TRACE_STATIC_MAX_RAYS * sizeof(crt_ray));
    cudaStream_t st = c->stream;
    CRT_CUDA(cudaMemsetAsync(cursor, 0, 256, st));   // the array keeps the last call's cursor: it must start at zero
    CRT_CUDA(cudaMemcpyAsync(d_rays, rays, n * sizeof(crt_ray), cudaMemcpyHostToDevice, st));
    int bps = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_trace_lbvh<false, false>, TP_BLOCK, 0);
    if (bps < 1) bps = 1;
    const size_t want = (n + 31) / 32;
    const size_t blocks = (want + (TP_BLOCK / 32) - 1) / (TP_BLOCK / 32);
    const size_t cap = (size_t)c->sm_count * bps;
    const unsigned int grid = (unsigned int)(blocks < cap ? blocks : cap);
    const AccelView A = accel_view(c);
    unsigned long long* counters = c->counters.as<unsigned long long>();
    if (scene_is_general(c))
        k_trace_lbvh<false, true><<<grid, TP_BLOCK, 0, st>>>(A, d_rays, (uint32_t)n, cursor, t_min, t_max, d_hits, counters, (uint32_t)c->refill, (uint32_t)c->leaf_thresh);
    else
        k_trace_lbvh<false, false><<<grid, TP_BLOCK, 0, st>>>(A, d_rays, (uint32_t)n, cursor, t_min, t_max, d_hits, counters, (uint32_t)c->refill, (uint32_t)c->leaf_thresh);
    CRT_CUDA(cudaGetLastError());
    CRT_CUDA(cudaMemcpyAsync(hits, d_hits, n * sizeof(crt_hit), cudaMemcpyDeviceToHost, st));
    CRT_CUDA(cudaStreamSynchronize(st));
    CRT_CUDA(cudaGetLastError());
    c->host_counters.rays += n;
    return CRT_OK;
}

}  // namespace crt
