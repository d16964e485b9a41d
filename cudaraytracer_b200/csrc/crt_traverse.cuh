// crt_traverse.cuh — closest-hit query over the flat LBVH.
//
// Replaces HitableList::collision_detection (hitable/hitable_list.h:76-91, the
// O(N) scan the oracle runs) and BVHNode::collision_detection
// (hitable/bvh.h:160-190, both children always visited, no ordering, no
// t_max shrink) with an ordered, stack-based traversal:
//   - one 64-byte node = both children's boxes, fetched with four 128-bit loads;
//   - near child first, far child pushed, the ray interval shrinks with every hit;
//   - the first STACK_SMEM stack levels live in shared memory laid out
//     [level][thread] (bank-conflict free), deeper levels spill to local memory;
//   - primitives are tested with the reference's exact arithmetic
//     (crt_device.cuh), so which primitive wins never depends on the tree;
//     equal t is resolved towards the lower primitive id, as the first list
//     entry wins in hitable_list.h:84.
// Box tests are conservative: slabs are evaluated as (plane - origin) * inv_dir
// (relative error of a few ulp) and the exit distance is widened by 4 ulp.
#pragma once
#include "crt_device.cuh"

namespace crt {

#define CRT_STACK_SMEM 32     // levels kept in shared memory
#define CRT_STACK_LOCAL 72    // spill levels (63 Morton bits + 32 index bits bound the depth)

struct TraceCounters { uint32_t nodes, tris, spheres; };

__device__ __forceinline__ float safe_rcp(float d)
{
    return fabsf(d) > 1e-18f ? __frcp_rn(d) : copysignf(1e18f, d);
}

__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

// o, d: origin and the direction the primitive tests see (ray_dir_local).
// s_stack points at this thread's column of the shared stack; consecutive
// levels are `stride` ints apart.
template <bool COUNT, bool HAS_SPH>
__device__ __forceinline__ void traverse(const AccelView& A, v3 o, v3 d, float t_min, float t_max,
                                         Hit& best, int* s_stack, int stride, TraceCounters& cnt)
{
    best.t = t_max;
    best.k = 0xFFFFFFFFu;
    best.id = CRT_MISS_ID;
    best.u = best.v = 0.0f;
    if (A.n_prims == 0) return;

    const float idx = safe_rcp(d.x), idy = safe_rcp(d.y), idz = safe_rcp(d.z);
    int lstack[CRT_STACK_LOCAL];
    int sp = 0;
    int node = A.root;

    while (true) {
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
                if (sp < CRT_STACK_SMEM) s_stack[sp * stride] = far_c;
                else if (sp < CRT_STACK_SMEM + CRT_STACK_LOCAL) lstack[sp - CRT_STACK_SMEM] = far_c;
                sp++;
                continue;
            }
            if (h0) { node = c0; continue; }
            if (h1) { node = c1; continue; }
        } else {
            const uint32_t ref = ~(uint32_t)node;
            const uint32_t first = ref >> 3;
            const uint32_t count = (ref & 7u) + 1u;
            for (uint32_t k = first; k < first + count; k++) {
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
                if (!ok) continue;
                if (t == best.t) {
                    // equal distance: lowest primitive id wins; t == t_max itself never counts
                    if (best.k == 0xFFFFFFFFu) continue;
                    if (__ldg(&A.info[k]).x >= __ldg(&A.info[best.k]).x) continue;
                }
                best.t = t; best.k = k; best.u = u; best.v = v;
            }
        }
        if (sp == 0) break;
        sp--;
        node = sp < CRT_STACK_SMEM ? s_stack[sp * stride] : lstack[sp - CRT_STACK_SMEM];
    }
    if (best.k != 0xFFFFFFFFu) best.id = __ldg(&A.info[best.k]).x;
}

}  // namespace crt
