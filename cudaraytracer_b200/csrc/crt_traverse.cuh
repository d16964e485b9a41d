// crt_traverse.cuh — closest-hit query over the flat LBVH, persistent-warp form.
//
// Replaces HitableList::collision_detection (hitable/hitable_list.h:76-91, the
// O(N) scan the oracle runs) and BVHNode::collision_detection
// (hitable/bvh.h:160-190, both children always visited, no ordering, no
// t_max shrink).
//
// Execution model (chosen from the ncu capture in profiles/r01_v1_summary.md:
// the one-ray-per-thread kernel ran with 6 of 32 lanes active on secondary rays):
//   - persistent warps: a warp grabs chunks of TP_CHUNK rays from the queue with
//     one atomic and keeps its lanes busy - whenever at least TP_REFILL lanes
//     have finished their ray they take the next rays of the chunk (ballot +
//     popc ranks), so a long ray no longer idles the other 31 lanes;
//   - phased traversal: per iteration the warp runs EITHER one inner-node step
//     (lanes that reached a leaf park) OR, once leaf_thresh lanes are parked or
//     no lane is at a node, one primitive test for every parked lane. Measured
//     alternatives (profiles/r01_*): while-while ran the node loop with 5 of 32
//     lanes, if-if ran the primitive test with 1.7 of 32 lanes;
//   - one 64-byte node = both children's boxes, four 128-bit loads; near child
//     first, far child pushed; the ray interval shrinks with every hit;
//   - the first CRT_STACK_SMEM stack levels live in shared memory laid out
//     [level][thread] (bank-conflict free), deeper levels spill to local memory;
//   - primitives are tested with the reference's exact arithmetic
//     (crt_device.cuh), so which primitive wins never depends on the tree;
//     equal t resolves to the lower primitive id, as the first list entry wins
//     in hitable_list.h:84.
// Box tests are conservative: slabs are (plane - origin) * inv_dir (a few ulp
// of relative error) and the exit distance is widened by 4 ulp.
#pragma once
#include "crt_device.cuh"

namespace crt {

#define CRT_STACK_SMEM 32     // levels kept in shared memory
#define CRT_STACK_LOCAL 72    // spill levels (63 Morton bits + 32 index bits bound the depth)
#define CRT_NODE_DONE ((int)0x80000000)

constexpr int TP_BLOCK = 128;    // threads per block of every traversal kernel
constexpr int TP_CHUNK = 256;    // most rays a warp takes from the queue per atomic
constexpr int TP_REFILL = 8;     // default: refill once this many lanes are idle

struct TraceCounters { uint32_t nodes, tris, spheres; };

__device__ __forceinline__ float safe_rcp(float d)
{
    return fabsf(d) > 1e-18f ? __frcp_rn(d) : copysignf(1e18f, d);
}

__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

// Src: bool load(uint32_t i, v3& o, v3& d)  -> origin and raw direction of ray i
// Sink: void store(uint32_t i, const Hit& h) -> h.k = sorted primitive position or ~0u
// n_ptr: number of rays (device memory); work_ctr: zeroed queue cursor.
template <bool COUNT, bool HAS_SPH, class Src, class Sink>
__device__ __forceinline__ void traverse_persistent(const AccelView& A, const uint32_t n, uint32_t* work_ctr,
                                                    const float t_min, const float t_max, const Src& src,
                                                    const Sink& sink, int* s_stack, TraceCounters& cnt,
                                                    const uint32_t refill, const uint32_t leaf_thresh)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    int* const my_stack = s_stack + threadIdx.x;
    int lstack[CRT_STACK_LOCAL];

    int node = CRT_NODE_DONE;
    int sp = 0;
    uint32_t ray_i = 0;
    v3 o = V(0.f, 0.f, 0.f), d = V(0.f, 0.f, 1.f);
    float idx = 0.f, idy = 0.f, idz = 0.f;
    Hit best;
    best.t = t_max; best.k = 0xFFFFFFFFu; best.id = CRT_MISS_ID; best.u = best.v = 0.f;

    // rays per grab: large enough to amortise the atomic, small enough that every
    // warp of the grid gets several grabs (a 2M-ray launch has only ~1.4 full chunks per warp)
    uint32_t chunk = (n / (gridDim.x * (TP_BLOCK / 32u) * 4u)) & ~31u;
    chunk = chunk < 32u ? 32u : (chunk > (uint32_t)TP_CHUNK ? (uint32_t)TP_CHUNK : chunk);
    uint32_t chunk_next = 0, chunk_end = 0;   // warp-uniform
    bool more = true;                         // warp-uniform

    for (;;) {
        const uint32_t idle_mask = __ballot_sync(0xffffffffu, node == CRT_NODE_DONE);
        if (idle_mask) {
            const uint32_t n_idle = (uint32_t)__popc(idle_mask);
            if (more && (n_idle >= refill || idle_mask == 0xffffffffu)) {
                if (chunk_next >= chunk_end) {
                    uint32_t b = 0;
                    if (lane == 0) b = atomicAdd(work_ctr, chunk);
                    b = __shfl_sync(0xffffffffu, b, 0);
                    if (b >= n) { more = false; chunk_next = chunk_end = 0; }
                    else { chunk_next = b; chunk_end = min(b + chunk, n); }
                }
                if (more) {
                    const uint32_t avail = chunk_end - chunk_next;
                    const uint32_t rank = (uint32_t)__popc(idle_mask & lt_mask);
                    if (node == CRT_NODE_DONE && rank < avail) {
                        ray_i = chunk_next + rank;
                        v3 dr;
                        src.load(ray_i, o, dr);
                        d = ray_dir_local(dr);
                        idx = safe_rcp(d.x); idy = safe_rcp(d.y); idz = safe_rcp(d.z);
                        best.t = t_max; best.k = 0xFFFFFFFFu; best.u = best.v = 0.f;
                        sp = 0;
                        node = A.root;
                        if (A.n_prims == 0u) { sink.store(ray_i, best); node = CRT_NODE_DONE; }
                    }
                    chunk_next += min(n_idle, avail);
                }
            }
            if (__ballot_sync(0xffffffffu, node != CRT_NODE_DONE) == 0u) {
                if (!more) break;
                continue;
            }
        }

        // ---- one phase per iteration: an inner-node step for the lanes at a node,
        //      or - once enough lanes are parked on a leaf, or nobody is at a node -
        //      one primitive test for every parked lane ----
        const uint32_t leaf_mask = __ballot_sync(0xffffffffu, node < 0 && node != CRT_NODE_DONE);
        const uint32_t int_mask = __ballot_sync(0xffffffffu, node >= 0);
        if (int_mask != 0u && (uint32_t)__popc(leaf_mask) < leaf_thresh) {
            if (node >= 0) {
                const float4* np = A.nodes + 4 * (size_t)node;
                const float4 n0 = ldg4(np), n1 = ldg4(np + 1), n2 = ldg4(np + 2), n3 = ldg4(np + 3);
                if (COUNT) cnt.nodes++;
                const float c0lox = (n0.x - o.x) * idx, c0hix = (n0.y - o.x) * idx;
                const float c0loy = (n0.z - o.y) * idy, c0hiy = (n0.w - o.y) * idy;
                const float c1lox = (n1.x - o.x) * idx, c1hix = (n1.y - o.x) * idx;
                const float c1loy = (n1.z - o.y) * idy, c1hiy = (n1.w - o.y) * idy;
                const float c0loz = (n2.x - o.z) * idz, c0hiz = (n2.y - o.z) * idz;
                const float c1loz = (n2.z - o.z) * idz, c1hiz = (n2.w - o.z) * idz;
                const float tn0 = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), t_min));
                const float tf0 = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), best.t));
                const float tn1 = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), t_min));
                const float tf1 = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), best.t));
                const bool h0 = tn0 <= tf0 * 1.0000005f;
                const bool h1 = tn1 <= tf1 * 1.0000005f;
                const int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
                if (h0 && h1) {
                    const bool swap = tn1 < tn0;
                    const int far_c = swap ? c0 : c1;
                    node = swap ? c1 : c0;
                    if (sp < CRT_STACK_SMEM) my_stack[sp * TP_BLOCK] = far_c;
                    else if (sp < CRT_STACK_SMEM + CRT_STACK_LOCAL) lstack[sp - CRT_STACK_SMEM] = far_c;
                    sp++;
                } else if (h0) {
                    node = c0;
                } else if (h1) {
                    node = c1;
                } else if (sp > 0) {
                    sp--;
                    node = sp < CRT_STACK_SMEM ? my_stack[sp * TP_BLOCK] : lstack[sp - CRT_STACK_SMEM];
                } else {
                    sink.store(ray_i, best);
                    node = CRT_NODE_DONE;
                }
            }
        } else if (node < 0 && node != CRT_NODE_DONE) {
            // the leaf reference itself is the cursor: test its first primitive, then shrink it
            const uint32_t ref = ~(uint32_t)node;
            const uint32_t k = ref >> 3;
            const uint32_t rest = ref & 7u;
            {
                const float4* rp = A.recs + 3 * (size_t)k;
                const float4 r0 = ldg4(rp);
                bool is_sphere = false;
                if (HAS_SPH) is_sphere = (__ldg(&A.info[k]).y & CRT_KIND_SPHERE) != 0u;
                float t, u = 0.0f, v = 0.0f;
                bool ok;
                if (HAS_SPH && is_sphere) {
                    if (COUNT) cnt.spheres++;
                    ok = sphere_test(vsub(o, V(r0.x, r0.y, r0.z)), d, r0.w, t_min, best.t, t);
                } else {
                    const float4 r1 = ldg4(rp + 1), r2 = ldg4(rp + 2);
                    if (COUNT) cnt.tris++;
                    ok = triangle_test(o, d, V(r0.x, r0.y, r0.z), V(r1.x, r1.y, r1.z), V(r2.x, r2.y, r2.z),
                                       V(r0.w, r1.w, r2.w), t_min, best.t, t, u, v);
                }
                if (ok && t == best.t) {
                    // equal distance: lowest primitive id wins; t == t_max itself never counts
                    if (best.k == 0xFFFFFFFFu) ok = false;
                    else if (__ldg(&A.info[k]).x >= __ldg(&A.info[best.k]).x) ok = false;
                }
                if (ok) { best.t = t; best.k = k; best.u = u; best.v = v; }
            }
            if (rest != 0u) {
                node = (int)~(((k + 1u) << 3) | (rest - 1u));
            } else if (sp > 0) {
                sp--;
                node = sp < CRT_STACK_SMEM ? my_stack[sp * TP_BLOCK] : lstack[sp - CRT_STACK_SMEM];
            } else {
                sink.store(ray_i, best);
                node = CRT_NODE_DONE;
            }
        }
    }
}

}  // namespace crt
