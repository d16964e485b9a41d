// crt_traverse.cuh — closest-hit query over the flat LBVH, persistent-warp form.
//
// Replaces HitableList::collision_detection (hitable/hitable_list.h:76-91, the
// O(N) scan the oracle runs) and BVHNode::collision_detection
// (hitable/bvh.h:160-190, both children always visited, no ordering, no
// t_max shrink).
//
// Execution model (each choice from an ncu capture, profiles/r01_v*_summary.md):
//   - persistent warps: a warp grabs chunks of rays from the queue with one
//     atomic and keeps its lanes busy - whenever at least `refill` lanes have
//     finished their ray they take the next rays of the chunk (ballot + popc
//     ranks), so a long ray no longer idles the other 31 lanes (v1: 6 of 32
//     lanes active on secondary rays);
//   - phased traversal: per iteration the warp runs EITHER one inner-node step
//     (lanes that reached a leaf park) OR, once leaf_thresh lanes are parked or
//     no lane is at a node, one primitive test for every parked lane. Measured
//     alternatives: while-while ran the node loop with 5 of 32 lanes, if-if ran
//     the primitive test with 1.7 of 32 lanes. Two ballots per iteration
//     classify every lane as node / leaf / idle;
//   - one 128-byte node (one cache line) = the boxes of up to four children,
//     fetched with three 256-bit loads (LDG.E.256, sm_100) + one 128-bit load of
//     the child references; a primitive record is 64 bytes = two 256-bit loads.
//     v2 used four 128-bit loads per binary node and was L1-wavefront bound on
//     incoherent rays (87 % L1/TEX, one wavefront per lane per load
//     instruction); v3 (binary, 256-bit loads) was bound by the per-iteration
//     bookkeeping and the dependent node fetch, which a 4-wide node halves;
//   - slab distances are one FMA per plane, plane * inv_dir - origin * inv_dir,
//     with a per-axis slack folded into the constant so the test stays
//     conservative (see RaySlabs); the hit children are sorted by entry
//     distance with a 5-exchange network, the nearest is entered and the others
//     are pushed farthest first; the ray interval shrinks with every hit;
//   - the first CRT_STACK_SMEM stack levels live in shared memory laid out
//     [level][thread] (bank-conflict free) and addressed through a byte cursor
//     kept in a register; deeper levels spill to local memory;
//   - primitives are tested with the reference's exact arithmetic
//     (crt_device.cuh), so which primitive wins never depends on the tree;
//     equal t resolves to the lower primitive id, as the first list entry wins
//     in hitable_list.h:84.
#pragma once
#include "crt_device.cuh"
#include <cuda_fp16.h>

namespace crt {

// CRT_STACK_T (development switch, off): a per-ray stack entry also carries the entry distance of its box, and an
// entry popped after a closer hit was found is dropped without fetching its node (the rule of traverse_wide); the
// shared-memory levels are halved so that the block's shared memory stays 16 KB. Measured at v9 on C3: camera rays
// through the per-ray kernel visit 13 % fewer nodes and test 22 % fewer triangles in the same time (0.504 -> 0.507 ms),
// and the frame is slower (8.90 -> 9.16 ms): the second stack word and the pop loop cost what the skipped visits save.
#ifndef CRT_STACK_T
#define CRT_STACK_T 0
#endif
#if CRT_STACK_T
#define CRT_STACK_SMEM 16     // levels kept in shared memory (reference + entry distance each)
#define CRT_STACK_WORDS 2
#else
#ifndef CRT_STACK_SMEM
#define CRT_STACK_SMEM 32     // levels kept in shared memory (16 KB a block: half of an SM's 256 KB at 8 blocks, the other half is L1)
#endif
#define CRT_STACK_WORDS 1
#endif
#define CRT_STACK_LOCAL (288 - CRT_STACK_SMEM)   // spill levels. A descent pushes <= 3 references per inner node, so a tree of depth D needs
                              // <= 3 (D - 1) entries; the radix tree has D <= 96 (63 Morton bits + 32 index bits + leaf),
                              // and lbvh_build keeps any other topology within CRT_STACK_ENTRIES
#define CRT_STACK_ENTRIES (CRT_STACK_SMEM + CRT_STACK_LOCAL)
#define CRT_NODE_DONE ((int)0x80000000)

// Slack of every "is this box entered before the closest hit so far" comparison, relative to the distance: 2^-14,
// twice the tolerance of deviation D3 (crt_device.cuh:triangle_test), so that a triangle hit D3 accepts (its distance
// lies within 2^-15 of its own box's interval, and every ancestor's box contains that box) is never in a subtree the
// traversal skipped unless a strictly closer hit exists.
// The box distances are accurate to a few 2^-24, but the distance a PRIMITIVE test reports is not: the reference's
// Moeller-Trumbore arithmetic (shapes/triangle.h:69-92) loses k 2^-24 t / cos(incidence) to cancellation in
// s = o - v0 and the two cross products, so a triangle hit at a grazing angle can report a t that lies outside its
// own box's interval by more than the 2^-21 the boxes are widened by. Which hit wins must not depend on the order
// in which a traversal meets two such near-ties (the list scan meets them all), so a subtree is only skipped when
// its entry distance exceeds the closest hit by this margin. Measured at C4's size (camera 2,700 units from the
// scene, 5.3e8 camera rays): with 2^-21 and without D3 the packet and the per-ray traversal disagreed on 14 pixels
// of the frame (profiles/r02_e_c4_invariance.txt).
#ifndef CRT_T_SLACK
#define CRT_T_SLACK 6.103515625e-5f             // 2^-14
#define CRT_T_SLACK1 1.00006103515625f          // 1 + 2^-14
#endif

constexpr int TP_BLOCK = 128;    // threads per block of every traversal kernel
#ifndef CRT_TRACE_MIN_BLOCKS
#define CRT_TRACE_MIN_BLOCKS 8   // resident blocks per SM the register allocation must allow
#endif
#ifndef CRT_PACKET_MIN_BLOCKS
#define CRT_PACKET_MIN_BLOCKS 8  // 10 and 12 resident blocks (51 and 40 registers) measured slower
#endif
constexpr int TP_CHUNK = 256;    // most rays a warp takes from the queue per atomic
constexpr int TP_REFILL = 12;    // default: refill once this many lanes are idle (sweeps at v9 and v11: 12 <= 8 < 4, 16)

// nodes/tris/spheres: tests executed per ray (work, flops). *_f: records FETCHED - equal to the tests in the
// per-ray kernels, one per packet visit in the packet kernel (a packet of <= 32 rays shares each fetch).
struct TraceCounters { uint32_t nodes, tris, spheres, max_steps, node_f, tri_f, sph_f; };

__device__ __forceinline__ float safe_rcp(float d)
{
    return fabsf(d) > 1e-18f ? __frcp_rn(d) : copysignf(1e18f, d);
}

__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

// L2 policy for the node and record loads: evict-last, so that the GB of queue traffic that streams through
// every launch (evict-first, render.cu) does not displace the tree, which fits the 126 MB L2.
__device__ __forceinline__ uint64_t l2_keep_policy()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

// 32 bytes with one LDG.E.256 through the read-only path; p must be 32-byte aligned.
__device__ __forceinline__ void ldg256(const float4* p, float4& a, float4& b)
{
    const uint64_t pol = l2_keep_policy();       // a few uniform-datapath instructions, hoisted out of loops
    asm volatile("ld.global.nc.L2::cache_hint.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8], %9;"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p), "l"(pol));
}

__device__ __forceinline__ void sts32(uint32_t addr, int v) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v)); }
__device__ __forceinline__ int lds32(uint32_t addr)
{
    int v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

// The same with a "memory" clobber, for the packet traversal where lane 0 writes
// the warp's stack and every lane reads it after __syncwarp(): without the
// clobber the compiler may move the load across the barrier.
__device__ __forceinline__ void sts32_fenced(uint32_t addr, int v) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ int lds32_fenced(uint32_t addr)
{
    int v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}

// Per-ray constants of the slab test. For a plane at coordinate p the exact
// distance is (p - o) / d. The kernel evaluates fma(p, inv, c) with
// inv = rn(1 / d) and c = -(o * inv) moved by a slack s = 2^-22 * |o * inv|
// TOWARDS the conservative side: the plane that bounds the interval from below
// (lo for inv >= 0, hi otherwise) gets c - s, the other c + s. s covers the
// rounding of o * inv (half an ulp of it); the rounding of the FMA itself and
// of inv are relative to the result and covered by the 2^-21 widening of the
// boxes at build time plus the factor on the exit distance below. The slack is
// per axis, so an axis with a huge |o * inv| (direction component near zero)
// does not switch off culling on the other two.
struct RaySlabs {
    float ix, iy, iz;
    float cx_lo, cx_hi, cy_lo, cy_hi, cz_lo, cz_hi;
};

__device__ __forceinline__ void slab_axis(float o, float d, float& inv, float& c_lo, float& c_hi)
{
    inv = safe_rcp(d);
    const float oi = o * inv;
    const float s = fabsf(oi) * 2.384185791015625e-7f + 1e-37f;   // 2^-22
    const float enter = -oi - s, leave = -oi + s;
    c_lo = inv >= 0.0f ? enter : leave;
    c_hi = inv >= 0.0f ? leave : enter;
}

__device__ __forceinline__ RaySlabs make_slabs(v3 o, v3 d)
{
    RaySlabs s;
    slab_axis(o.x, d.x, s.ix, s.cx_lo, s.cx_hi);
    slab_axis(o.y, d.y, s.iy, s.cy_lo, s.cy_hi);
    slab_axis(o.z, d.z, s.iz, s.cz_lo, s.cz_hi);
    return s;
}

// A node's 24 words with three 256-bit loads (the whole record: boxes AND child references, crt_device.cuh):
// w[5k..5k+4] = {lo.x, lo.y, lo.z, half2(ext.x, ext.y), half2(ext.z, 0)} of child k, w[20..23] = child references.
__device__ __forceinline__ void load_node(const float4* __restrict__ np, float (&w)[24])
{
    float4 a, b;
    ldg256(np, a, b);
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    ldg256(np + 2, a, b);
    w[8] = a.x; w[9] = a.y; w[10] = a.z; w[11] = a.w; w[12] = b.x; w[13] = b.y; w[14] = b.z; w[15] = b.w;
    ldg256(np + 4, a, b);
    w[16] = a.x; w[17] = a.y; w[18] = a.z; w[19] = a.w; w[20] = b.x; w[21] = b.y; w[22] = b.z; w[23] = b.w;
}

// hi = lo + ext with ext stored as a half rounded UP at build time so that this very addition gives a plane at
// or beyond the true one (lbvh.cu:k_emit_nodes checks it with the same operation).
__device__ __forceinline__ void child_box(const float lox, const float loy, const float loz, const float e_xy, const float e_z,
                                          float& hix, float& hiy, float& hiz)
{
    const uint32_t u = __float_as_uint(e_xy), v = __float_as_uint(e_z);
    hix = __fadd_rn(lox, __half2float(__ushort_as_half((unsigned short)(u & 0xFFFFu))));
    hiy = __fadd_rn(loy, __half2float(__ushort_as_half((unsigned short)(u >> 16))));
    hiz = __fadd_rn(loz, __half2float(__ushort_as_half((unsigned short)(v & 0xFFFFu))));
}

// Node-level version of the reference's `if (dot(direction, normal) < 0) return false` (shapes/triangle.h:61): the
// spare halves of children 0-2 hold a vector A built from the normal cone of the node's whole subtree
// (lbvh.cu:cone_to_halves) such that dot(d, A) < -1 for the unit direction d means every triangle below would fail
// that test; such a node is left without testing its boxes. A = 0 (never true) for subtrees with spheres or normals
// too spread out. CRT_NORMAL_CONES=0 compiles the test out (development switch).
#ifndef CRT_NORMAL_CONES
#define CRT_NORMAL_CONES 1
#endif
__device__ __forceinline__ bool node_backfacing(const float (&w)[24], const v3& d)
{
#if CRT_NORMAL_CONES
    const float ax = __half2float(__ushort_as_half((unsigned short)(__float_as_uint(w[4]) >> 16)));
    const float ay = __half2float(__ushort_as_half((unsigned short)(__float_as_uint(w[9]) >> 16)));
    const float az = __half2float(__ushort_as_half((unsigned short)(__float_as_uint(w[14]) >> 16)));
    return fmaf(d.x, ax, fmaf(d.y, ay, d.z * az)) < -1.0f;
#else
    return false;
#endif
}

// A primitive under a chain of Transforms (rotated / nested Hitables, rectangles): the rare path, kept out of line so
// that the plain triangle and sphere tests keep their registers. tuv: t, u, v out.
static __device__ __noinline__ bool framed_test_dev(const DFrame* __restrict__ frames, uint32_t chain, uint32_t info, float4 r0, float4 r1,
                                             float4 r2, v3 o, v3 d_raw, float t_min, float t_best, float* tuv)
{
    float t = 0.f, u = 0.f, v = 0.f;
    const bool ok = framed_test(frames, chain, info, r0, r1, r2, o, d_raw, t_min, t_best, t, u, v);
    tuv[0] = t; tuv[1] = u; tuv[2] = v;
    return ok;
}

// order-preserving float <-> uint32 (for REDUX min/max over a warp)
__device__ __forceinline__ uint32_t f2sortable(float f)
{
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float sortable2f(uint32_t s)
{
    return __uint_as_float((s & 0x80000000u) ? (s & 0x7fffffffu) : ~s);
}
__device__ __forceinline__ float warp_min_f(float v) { return sortable2f(__reduce_min_sync(0xffffffffu, f2sortable(v))); }
__device__ __forceinline__ float warp_max_f(float v) { return sortable2f(__reduce_max_sync(0xffffffffu, f2sortable(v))); }

// The direction the primitive tests see (ray_dir_local of the raw direction): read from the source when it keeps it
// (Src::load_local - the render queues, whose writer computed it once for this kernel and for the shading kernel
// behind it), else computed here from the raw direction.
template <class Src>
__device__ __forceinline__ auto load_ray_local(const Src& s, uint32_t i, v3& o, v3& d, int) -> decltype(s.load_local(i, o, d), void())
{
    s.load_local(i, o, d);
}
template <class Src>
__device__ __forceinline__ void load_ray_local(const Src& s, uint32_t i, v3& o, v3& d, long)
{
    v3 dr;
    s.load(i, o, dr);
    d = ray_dir_local(dr);
}

// Src: bool load(uint32_t i, v3& o, v3& d)  -> origin and raw direction of ray i
// Sink: void store(uint32_t i, const Hit& h) -> h.k = sorted primitive position or ~0u, h.id = caller's id
// work_ctr: zeroed queue cursor.
template <bool COUNT, bool HAS_SPH, class Src, class Sink>
__device__ __forceinline__ void traverse_persistent(const AccelView& A, const uint32_t n, uint32_t* work_ctr,
                                                    const float t_min, const float t_max, const Src& src,
                                                    const Sink& sink, int* s_stack, TraceCounters& cnt,
                                                    const uint32_t refill, const uint32_t leaf_thresh)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    // shared-memory stack: level L of this thread lives at stack_base + L * (TP_BLOCK * 4)
    const uint32_t stack_base = (uint32_t)__cvta_generic_to_shared(s_stack + threadIdx.x);
    constexpr uint32_t LEVEL = TP_BLOCK * 4u;
    int lstack[CRT_STACK_LOCAL];
    lstack[0] = 0;
#if CRT_STACK_T
    const uint32_t tstack_base = stack_base + CRT_STACK_SMEM * LEVEL;   // entry distances: the second half of s_stack
    float lstack_t[CRT_STACK_LOCAL];
    lstack_t[0] = 0.f;
#endif

    int node = CRT_NODE_DONE;
    uint32_t sp = 0;                          // stack depth
    uint32_t ray_i = 0;
    uint32_t ray_steps = 0;                   // COUNT only: loop iterations spent on the current ray
    v3 o = V(0.f, 0.f, 0.f), d = V(0.f, 0.f, 1.f);
    RaySlabs S = make_slabs(o, d);
    Hit best;
    best.t = t_max; best.k = 0xFFFFFFFFu; best.id = CRT_MISS_ID; best.u = best.v = 0.f;

    // rays per grab: large enough to amortise the atomic, small enough that every
    // warp of the grid gets several grabs (a 2M-ray launch has only ~1.4 full chunks per warp)
    uint32_t chunk = (n / (gridDim.x * (TP_BLOCK / 32u) * 4u)) & ~31u;
    chunk = chunk < 32u ? 32u : (chunk > (uint32_t)TP_CHUNK ? (uint32_t)TP_CHUNK : chunk);
    uint32_t chunk_next = 0, chunk_end = 0;   // warp-uniform
    bool more = true;                         // warp-uniform

#if CRT_STACK_T
    auto push = [&](int r, float tn) {
        if (sp < CRT_STACK_SMEM) { sts32(stack_base + sp * LEVEL, r); sts32(tstack_base + sp * LEVEL, __float_as_int(tn)); }
        else if (sp < CRT_STACK_SMEM + CRT_STACK_LOCAL) { lstack[sp - CRT_STACK_SMEM] = r; lstack_t[sp - CRT_STACK_SMEM] = tn; }
        sp++;
    };
    auto pop_or_finish = [&]() {
        for (;;) {
            if (sp == 0u) {
                sink.store(ray_i, best);
                node = CRT_NODE_DONE;
                break;
            }
            sp--;
            float tn;
            if (sp < CRT_STACK_SMEM) { node = lds32(stack_base + sp * LEVEL); tn = __int_as_float(lds32(tstack_base + sp * LEVEL)); }
            else { node = lstack[sp - CRT_STACK_SMEM]; tn = lstack_t[sp - CRT_STACK_SMEM]; }
            // still entered before the current closest hit? (the slack of the test that pushed it)
            if (tn <= best.t * CRT_T_SLACK1) break;
        }
    };
#else
    auto push = [&](int r, float) {
        if (sp < CRT_STACK_SMEM) sts32(stack_base + sp * LEVEL, r);
        else if (sp < CRT_STACK_SMEM + CRT_STACK_LOCAL) lstack[sp - CRT_STACK_SMEM] = r;
        sp++;
    };
    auto pop_or_finish = [&]() {
        if (sp > 0u) {
            sp--;
            node = sp < CRT_STACK_SMEM ? lds32(stack_base + sp * LEVEL) : lstack[sp - CRT_STACK_SMEM];
        } else {
            sink.store(ray_i, best);
            node = CRT_NODE_DONE;
        }
    };
#endif

    for (;;) {
        const uint32_t int_mask = __ballot_sync(0xffffffffu, node >= 0);
        const uint32_t leaf_mask = __ballot_sync(0xffffffffu, node < 0 && node != CRT_NODE_DONE);
        const uint32_t busy = int_mask | leaf_mask;
        if (busy != 0xffffffffu) {
            const uint32_t idle_mask = ~busy;
            const uint32_t n_idle = (uint32_t)__popc(idle_mask);
            if (more && (n_idle >= refill || busy == 0u)) {
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
                        load_ray_local(src, ray_i, o, d, 0);
                        S = make_slabs(o, d);
                        best.t = t_max; best.k = 0xFFFFFFFFu; best.id = CRT_MISS_ID; best.u = best.v = 0.f;
                        sp = 0;
                        ray_steps = 0;
                        node = A.root;
                        if (A.n_prims == 0u) { sink.store(ray_i, best); node = CRT_NODE_DONE; }
                    }
                    chunk_next += min(n_idle, avail);
                }
                continue;                      // classify the lanes again
            }
            if (busy == 0u) break;             // nothing in flight and the queue is empty
        }

        // ---- one phase per iteration: an inner-node step for the lanes at a node,
        //      or - once enough lanes are parked on a leaf, or nobody is at a node -
        //      one primitive test for every parked lane ----
        if (int_mask != 0u && (uint32_t)__popc(leaf_mask) < leaf_thresh) {
            if (node >= 0) {
                const float4* np = A.nodes + 8 * (size_t)(uint32_t)node;
                float f[24];
                load_node(np, f);
                if (COUNT) { cnt.nodes++; ray_steps++; cnt.max_steps = max(cnt.max_steps, ray_steps); }
                float key[4];
                int ref[4] = {__float_as_int(f[20]), __float_as_int(f[21]), __float_as_int(f[22]), __float_as_int(f[23])};
                int n_hit = 0;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    float hx, hy, hz;
                    child_box(f[5 * k], f[5 * k + 1], f[5 * k + 2], f[5 * k + 3], f[5 * k + 4], hx, hy, hz);
                    const float ax = fmaf(f[5 * k], S.ix, S.cx_lo), bx = fmaf(hx, S.ix, S.cx_hi);
                    const float ay = fmaf(f[5 * k + 1], S.iy, S.cy_lo), by = fmaf(hy, S.iy, S.cy_hi);
                    const float az = fmaf(f[5 * k + 2], S.iz, S.cz_lo), bz = fmaf(hz, S.iz, S.cz_hi);
                    const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), t_min));
                    const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), best.t));
                    const bool h = tn <= tf * CRT_T_SLACK1 && ref[k] != CRT_SLOT_EMPTY;
                    key[k] = h ? tn : INFINITY;
                    n_hit += h ? 1 : 0;
                }
                if (node_backfacing(f, d)) n_hit = 0;   // every triangle below faces away from this ray (Q1)
                // order the four slots by entry distance (misses last): 5 compare-exchanges
#define CRT_CSWAP(a, b)                                                                    \
                {                                                                          \
                    const bool sw = key[b] < key[a];                                       \
                    const float ka = key[a], kb = key[b];                                  \
                    const int ra = ref[a], rb = ref[b];                                    \
                    key[a] = sw ? kb : ka; key[b] = sw ? ka : kb;                          \
                    ref[a] = sw ? rb : ra; ref[b] = sw ? ra : rb;                          \
                }
                CRT_CSWAP(0, 1) CRT_CSWAP(2, 3) CRT_CSWAP(0, 2) CRT_CSWAP(1, 3) CRT_CSWAP(1, 2)
#undef CRT_CSWAP
                // farthest first, so the nearest pending child is popped first
                if (n_hit > 3) push(ref[3], key[3]);
                if (n_hit > 2) push(ref[2], key[2]);
                if (n_hit > 1) push(ref[1], key[1]);
                if (n_hit > 0) node = ref[0];
                else pop_or_finish();
            }
        } else if (node < 0 && node != CRT_NODE_DONE) {
            // the leaf reference itself is the cursor: test its first primitive, then shrink it
            const uint32_t ref = ~(uint32_t)node;
            const uint32_t k = ref >> 3;
            const uint32_t rest = ref & 7u;
            if (COUNT) { ray_steps++; cnt.max_steps = max(cnt.max_steps, ray_steps); }
            {
                const float4* rp = A.recs + 4 * (size_t)k;
                float4 r0, r1, r2, r3;
                ldg256(rp, r0, r1);
                ldg256(rp + 2, r2, r3);
                const uint32_t pid = __float_as_uint(r3.x);
                float t, u = 0.0f, v = 0.0f;
                bool ok;
                if (HAS_SPH && __float_as_uint(r3.z) != 0u) {
                    if (COUNT) cnt.spheres++;
                    v3 oo, dr;
                    src.load(ray_i, oo, dr);            // the direction as the caller gave it: the chain normalises per level
                    float tuv[3];
                    ok = framed_test_dev(A.frames, __float_as_uint(r3.z), __float_as_uint(r3.y), r0, r1, r2, o, dr, t_min, best.t, tuv);
                    t = tuv[0]; u = tuv[1]; v = tuv[2];
                } else if (HAS_SPH && (__float_as_uint(r3.y) & CRT_KIND_SPHERE)) {
                    if (COUNT) cnt.spheres++;
                    ok = sphere_test(vsub(o, V(r0.x, r0.y, r0.z)), d, r0.w, t_min, best.t, t);
                } else {
                    if (COUNT) cnt.tris++;
                    ok = triangle_test(o, d, V(r0.x, r0.y, r0.z), V(r1.x, r1.y, r1.z), V(r2.x, r2.y, r2.z),
                                       V(r0.w, r1.w, r2.w), t_min, best.t, t, u, v);
                }
                // equal distance: lowest primitive id wins; t == t_max itself never counts
                if (ok && t == best.t && (best.k == 0xFFFFFFFFu || pid >= best.id)) ok = false;
                if (ok) { best.t = t; best.k = k; best.id = pid; best.u = u; best.v = v; }
            }
            if (rest != 0u) node = (int)~(((k + 1u) << 3) | (rest - 1u));
            else pop_or_finish();
        }
    }
}


// ---------------------------------------------------------------------------
// Packet form for COHERENT rays (camera rays: with the sample-minor path order
// of render.cu a warp holds all samples of one or two pixels, i.e. 32 nearly
// identical rays). The warp walks ONE stack; control flow is warp-uniform, so
// there is no lane refill, no node/leaf phase vote and no per-lane stack, and
// node and record addresses are uniform, so a load is one L1 wavefront for the
// whole warp. Every lane tests the primitives of a visited leaf against its own
// ray and keeps its own closest hit; visiting a node another lane needed never
// changes a lane's result (same primitive tests and tie-break as
// traverse_persistent), so the node test only has to be conservative for the
// packet as a whole:
//   - rays with one common origin (a pinhole camera): ONE interval test per
//     child for the whole packet. With inv = 1/d in [imin, imax] over the lanes,
//     a plane at p is crossed at (p - o) * inv in [min, max of the two products];
//     lane k (k & 3) evaluates child k from three 64-bit loads of the child-major
//     node, so the four children cost one pass of ~35 instructions instead of
//     four per-lane slab tests, the entry bound of a child orders the descent
//     and the farthest closest hit of the packet prunes it. The bounds are
//     widened by 2^-21 relative, which covers the three roundings involved;
//   - anything else (finite aperture, arbitrary crt_trace batches): every lane
//     tests the four boxes against its own ray, a child is entered when any
//     lane's ray enters it, children are ordered by the first entering lane.
#define CRT_PACKET_STACK CRT_STACK_ENTRIES   // same bound as the per-ray stack
#ifndef CRT_PACKET_STACK_SMEM
#define CRT_PACKET_STACK_SMEM 0              // entries of a warp's packet stack in shared memory (see traverse_packets)
#endif

// Long packets are handed on. A camera-ray launch ends with its longest packet: two pixels on the limb of the mesh
// walk ~500 dependent nodes and leaves where the average packet walks ~35, at ~1.3 us per visit while eight warps
// share a scheduler - 0.5 ms on its own, more than the rest of a rank's share of a frame split over 8 GPUs takes, and
// neither starting it first nor prefetching for it helps (profiles/r02_l_ab_packet_sched.txt). A packet that has made
// `after` visits therefore stops: its rays' queue indices go on a list and a launch queued right behind (render.cu
// k_trace_deferred) traces each listed ray with a whole WARP (trace_wide_one: the stack 32 entries abreast, ~15 rounds
// instead of hundreds of steps), all listed rays at once. Nothing of the abandoned walk is kept, so the hit is the
// closest hit of the ray as any other kernel finds it.
struct PacketDefer {
    uint32_t* list;      // queue indices of the rays handed on
    uint32_t* n;         // entries reserved (may run past cap: then the packet walks on instead)
    uint32_t cap;
    uint32_t after;      // 0: never
};

template <bool COUNT, bool HAS_SPH, bool DEFER = false, class Src, class Sink>
__device__ __forceinline__ void traverse_packets(const AccelView& A, const uint32_t n, uint32_t* work_ctr,
                                                 const float t_min, const float t_max, const Src& src,
                                                 const Sink& sink, TraceCounters& cnt, const bool interval = true,
                                                 const PacketDefer defer = PacketDefer{nullptr, nullptr, 0u, 0u})
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t kid = lane & 3u;
#ifndef CRT_PACKET_STACK_T
#define CRT_PACKET_STACK_T 1
#endif
    // The packet's stack is warp-uniform: one sequence of {child reference, entry distance} for the 32 lanes.
    // CRT_PACKET_STACK_SMEM > 0: its first that-many entries live in shared memory, one copy per WARP (8 bytes an entry);
    // every lane stores the same value to the same address and later reads what it stored itself, so no lane depends
    // on another lane's store, and every push follows a __ballot_sync / __shfl_sync of the same node step, which all
    // lanes reach only after their pops of the step before (no lane can overwrite an entry another still has to pop).
    // Deeper entries, and with CRT_PACKET_STACK_SMEM == 0 all of them, are per-lane local arrays in which every lane
    // keeps the same values (slot X of 32 lanes is one 128-byte line: a warp with 20 entries holds 5 KB of L1).
    // The entry distance (interval test only): an entry beyond the packet's farthest closest hit is dropped when
    // popped, without fetching its node.
#if CRT_PACKET_STACK_SMEM
    __shared__ int2 s_pstack[(TP_BLOCK / 32) * CRT_PACKET_STACK_SMEM];
    int2* const pstack = s_pstack + (threadIdx.x >> 5) * CRT_PACKET_STACK_SMEM;
#endif
#if CRT_PACKET_STACK_T
    float wstack_t[CRT_PACKET_STACK - CRT_PACKET_STACK_SMEM];
#endif
    int wstack[CRT_PACKET_STACK - CRT_PACKET_STACK_SMEM];
    constexpr uint32_t FULL = 0xffffffffu;
    // packets per grab
    uint32_t chunk = (n / (gridDim.x * (TP_BLOCK / 32u) * 8u)) & ~31u;
    chunk = chunk < 32u ? 32u : (chunk > 1024u ? 1024u : chunk);

    for (;;) {
        uint32_t begin = 0;
        if (lane == 0) begin = atomicAdd(work_ctr, chunk);
        begin = __shfl_sync(FULL, begin, 0);
        if (begin >= n) break;
        const uint32_t end = min(begin + chunk, n);
        for (uint32_t first = begin; first < end; first += 32u) {
            const uint32_t ray_i = first + lane;
            const bool live = ray_i < end;
            v3 o = V(0.f, 0.f, 0.f), d = V(0.f, 0.f, 1.f);
            if (live) load_ray_local(src, ray_i, o, d, 0);
            const RaySlabs S = make_slabs(o, d);
            // lane 0 is always live; a dead lane takes its ray so that it widens nothing
            const float o0x = __shfl_sync(FULL, o.x, 0), o0y = __shfl_sync(FULL, o.y, 0), o0z = __shfl_sync(FULL, o.z, 0);
            const float i0x = __shfl_sync(FULL, S.ix, 0), i0y = __shfl_sync(FULL, S.iy, 0), i0z = __shfl_sync(FULL, S.iz, 0);
            const float l_ix = live ? S.ix : i0x, l_iy = live ? S.iy : i0y, l_iz = live ? S.iz : i0z;
            const float ix_lo = warp_min_f(l_ix), ix_hi = warp_max_f(l_ix);
            const float iy_lo = warp_min_f(l_iy), iy_hi = warp_max_f(l_iy);
            const float iz_lo = warp_min_f(l_iz), iz_hi = warp_max_f(l_iz);
            // the interval test is only as tight as the interval of 1/d: on an axis where the packet's directions
            // change sign (or nearly do) it cannot cull, and the packet would walk a whole slab of the scene
            auto narrow = [](float lo, float hi) { return lo > 0.0f ? hi <= 2.0f * lo : (hi < 0.0f && lo >= 2.0f * hi); };
            const bool uni = interval && narrow(ix_lo, ix_hi) && narrow(iy_lo, iy_hi) && narrow(iz_lo, iz_hi) &&
                             __all_sync(FULL, !live || (o.x == o0x && o.y == o0y && o.z == o0z));
            float t_far = t_max;                 // warp-uniform: the largest closest-hit distance of the packet
            Hit best;
            best.t = t_max; best.k = 0xFFFFFFFFu; best.id = CRT_MISS_ID; best.u = best.v = 0.f;
            bool handed_on = false;              // warp-uniform
            if (A.n_prims != 0u) {
                int node = A.root;               // warp-uniform
                uint32_t sp = 0;                 // warp-uniform
                uint32_t visits = 0;             // warp-uniform
                auto push = [&](int r, float tn) {
#if CRT_PACKET_STACK_SMEM
                    if (sp < (uint32_t)CRT_PACKET_STACK_SMEM) { pstack[sp++] = make_int2(r, __float_as_int(tn)); return; }
#endif
#if CRT_PACKET_STACK_T
                    wstack_t[sp - CRT_PACKET_STACK_SMEM] = tn;
#endif
                    wstack[sp - CRT_PACKET_STACK_SMEM] = r;
                    sp++;
                };
                // the nearest stacked child that is still entered before the packet's farthest closest hit -> node
                auto pop_next = [&]() -> bool {
                    while (sp > 0u) {
                        --sp;
                        float tn = 0.0f;
#if CRT_PACKET_STACK_SMEM
                        if (sp < (uint32_t)CRT_PACKET_STACK_SMEM) {
                            const int2 e = pstack[sp];
                            node = e.x;
                            tn = __int_as_float(e.y);
                        } else
#endif
                        {
                            node = wstack[sp - CRT_PACKET_STACK_SMEM];
#if CRT_PACKET_STACK_T
                            tn = wstack_t[sp - CRT_PACKET_STACK_SMEM];
#endif
                        }
#if CRT_PACKET_STACK_T
                        if (tn > t_far) continue;
#endif
                        (void)tn;
                        return true;
                    }
                    return false;
                };
                for (;;) {
                    if (DEFER && ++visits > defer.after) {
                        uint32_t base = 0;
                        if (lane == 0) base = atomicAdd(defer.n, 32u);
                        base = __shfl_sync(FULL, base, 0);
                        if (base + 32u <= defer.cap) {
                            defer.list[base + lane] = live ? ray_i : 0xFFFFFFFFu;
                            handed_on = true;
                            break;
                        }
                        visits = 0;              // the list is full: walk on (and ask again after another `after` visits)
                    }
                    if (node >= 0) {
                        const float4* np = A.nodes + 8 * (size_t)(uint32_t)node;
                        const int4 refs = __ldg(reinterpret_cast<const int4*>(np + 5));
                        float key[4];
                        int ref[4] = {refs.x, refs.y, refs.z, refs.w};
                        int n_hit = 0;
                        if (COUNT && lane == 0) cnt.node_f++;
                        if (uni) {
                            if (COUNT && lane == 0) cnt.nodes++;       // one interval test serves the packet
                            const float* cp = reinterpret_cast<const float*>(np) + 5u * kid;
                            const float clx = __ldg(cp), cly = __ldg(cp + 1), clz = __ldg(cp + 2);
                            float chx, chy, chz;
                            child_box(clx, cly, clz, __ldg(cp + 3), __ldg(cp + 4), chx, chy, chz);
                            const float ax0 = clx - o0x, ax1 = chx - o0x;
                            const float ay0 = cly - o0y, ay1 = chy - o0y;
                            const float az0 = clz - o0z, az1 = chz - o0z;
                            const float x00 = ax0 * ix_lo, x01 = ax0 * ix_hi, x10 = ax1 * ix_lo, x11 = ax1 * ix_hi;
                            const float y00 = ay0 * iy_lo, y01 = ay0 * iy_hi, y10 = ay1 * iy_lo, y11 = ay1 * iy_hi;
                            const float z00 = az0 * iz_lo, z01 = az0 * iz_hi, z10 = az1 * iz_lo, z11 = az1 * iz_hi;
                            const float nx = fminf(fminf(x00, x01), fminf(x10, x11)), fx = fmaxf(fmaxf(x00, x01), fmaxf(x10, x11));
                            const float ny = fminf(fminf(y00, y01), fminf(y10, y11)), fy = fmaxf(fmaxf(y00, y01), fmaxf(y10, y11));
                            const float nz = fminf(fminf(z00, z01), fminf(z10, z11)), fz = fmaxf(fmaxf(z00, z01), fmaxf(z10, z11));
                            float tn = fmaxf(fmaxf(nx, ny), nz), tf = fminf(fminf(fx, fy), fz);
                            tn = fmaf(-fabsf(tn), CRT_T_SLACK, tn);                // towards "enters earlier"
                            tf = fmaf(fabsf(tf), CRT_T_SLACK, tf);                 // and "leaves later"
                            tn = fmaxf(tn, t_min);
                            tf = fminf(tf, t_far);
                            // two-level select on the bits of kid (a chain of ?: compiles to divergent branches)
                            const int ref_lo = (kid & 1u) ? refs.y : refs.x, ref_hi = (kid & 1u) ? refs.w : refs.z;
                            const int my_ref = (kid & 2u) ? ref_hi : ref_lo;
                            const bool enter = tn <= tf && my_ref != CRT_SLOT_EMPTY;
                            const uint32_t m4 = __ballot_sync(FULL, enter) & 0xFu;
                            if ((m4 & (m4 - 1u)) == 0u) {
                                // none or one child entered (the common case): no ordering needed
                                if (m4 == 0u) {
                                    if (!pop_next()) break;
                                } else {
                                    const int e_lo = (m4 & 1u) ? refs.x : refs.y, e_hi = (m4 & 4u) ? refs.z : refs.w;
                                    node = (m4 & 3u) ? e_lo : e_hi;
                                }
                                continue;
                            }
#pragma unroll
                            for (int k = 0; k < 4; k++) {
                                const float tk = __shfl_sync(FULL, tn, k);
                                const bool h = (m4 >> k) & 1u;
                                key[k] = h ? tk : INFINITY;
                                n_hit += h ? 1 : 0;
                            }
                        } else {
                            float f[24];
                            load_node(np, f);
                            if (COUNT && live) cnt.nodes++;
#pragma unroll
                            for (int k = 0; k < 4; k++) {
                                float hx, hy, hz;
                                child_box(f[5 * k], f[5 * k + 1], f[5 * k + 2], f[5 * k + 3], f[5 * k + 4], hx, hy, hz);
                                const float ax = fmaf(f[5 * k], S.ix, S.cx_lo), bx = fmaf(hx, S.ix, S.cx_hi);
                                const float ay = fmaf(f[5 * k + 1], S.iy, S.cy_lo), by = fmaf(hy, S.iy, S.cy_hi);
                                const float az = fmaf(f[5 * k + 2], S.iz, S.cz_lo), bz = fmaf(hz, S.iz, S.cz_hi);
                                const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), t_min));
                                const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), best.t));
                                const bool h = live && tn <= tf * CRT_T_SLACK1;
                                const uint32_t m = ref[k] != CRT_SLOT_EMPTY ? __ballot_sync(FULL, h) : 0u;   // ref is uniform
                                // entry distance of the first lane that enters the child; uniform
                                key[k] = m ? __shfl_sync(FULL, tn, __ffs(m) - 1) : INFINITY;
                                n_hit += m ? 1 : 0;
                            }
                        }
#define CRT_CSWAP(a, b)                                                                    \
                        {                                                                  \
                            const bool sw = key[b] < key[a];                               \
                            const float ka = key[a], kb = key[b];                          \
                            const int ra = ref[a], rb = ref[b];                            \
                            key[a] = sw ? kb : ka; key[b] = sw ? ka : kb;                  \
                            ref[a] = sw ? rb : ra; ref[b] = sw ? ra : rb;                  \
                        }
                        CRT_CSWAP(0, 1) CRT_CSWAP(2, 3) CRT_CSWAP(0, 2) CRT_CSWAP(1, 3) CRT_CSWAP(1, 2)
#undef CRT_CSWAP
                        // the per-lane test orders children by the first entering lane: not a bound for the packet
                        if (n_hit > 3) push(ref[3], uni ? key[3] : 0.0f);
                        if (n_hit > 2) push(ref[2], uni ? key[2] : 0.0f);
                        if (n_hit > 1) push(ref[1], uni ? key[1] : 0.0f);
                        if (n_hit > 0) { node = ref[0]; continue; }
                    } else {
                        const uint32_t ref = ~(uint32_t)node;
                        const uint32_t k0 = ref >> 3, count = (ref & 7u) + 1u;
                        for (uint32_t j = 0; j < count; j++) {
                            const uint32_t k = k0 + j;
                            const float4* rp = A.recs + 4 * (size_t)k;
                            float4 r0, r1, r2, r3;
                            ldg256(rp, r0, r1);
                            ldg256(rp + 2, r2, r3);
                            const uint32_t pid = __float_as_uint(r3.x);
                            float t, u = 0.0f, v = 0.0f;
                            bool ok;
                            if (HAS_SPH && __float_as_uint(r3.z) != 0u) {
                                if (COUNT && live) cnt.spheres++;
                                if (COUNT && lane == 0) cnt.sph_f++;
                                v3 oo, dr;
                                src.load(live ? ray_i : first, oo, dr);
                                float tuv[3];
                                ok = framed_test_dev(A.frames, __float_as_uint(r3.z), __float_as_uint(r3.y), r0, r1, r2, o, dr, t_min, best.t, tuv);
                                t = tuv[0]; u = tuv[1]; v = tuv[2];
                            } else if (HAS_SPH && (__float_as_uint(r3.y) & CRT_KIND_SPHERE)) {
                                if (COUNT && live) cnt.spheres++;
                                if (COUNT && lane == 0) cnt.sph_f++;
                                ok = sphere_test(vsub(o, V(r0.x, r0.y, r0.z)), d, r0.w, t_min, best.t, t);
                            } else {
                                if (COUNT && live) cnt.tris++;
                                if (COUNT && lane == 0) cnt.tri_f++;
                                ok = triangle_test(o, d, V(r0.x, r0.y, r0.z), V(r1.x, r1.y, r1.z), V(r2.x, r2.y, r2.z),
                                                   V(r0.w, r1.w, r2.w), t_min, best.t, t, u, v);
                            }
                            if (ok && t == best.t && (best.k == 0xFFFFFFFFu || pid >= best.id)) ok = false;
                            if (ok && live) { best.t = t; best.k = k; best.id = pid; best.u = u; best.v = v; }
                        }
                        // distances are positive, so their bit patterns order like the values
                        if (uni) t_far = __uint_as_float(__reduce_max_sync(FULL, live ? __float_as_uint(best.t) : 0u));
                    }
                    if (!pop_next()) break;
                }
            }
            if (live && !handed_on) sink.store(ray_i, best);
            __syncwarp();
        }
    }
}

// ---------------------------------------------------------------------------
// Wide form for launches with FEW rays (the last bounces of a frame: 20,000 rays
// down to a handful). There the per-ray kernel is bound by its longest ray: a
// grazing ray walks 200-700 DEPENDENT node visits while 99.9 % of the GPU idles
// (bounces 3-8 of a C3 frame: 0.04 % of the rays, 7 % of the frame). Here a whole
// warp works on ONE ray: the stack lives in shared memory and every round the 32
// lanes pop up to 32 entries at once - each lane tests the four children of its
// node (or the primitives of its leaf) against the same ray - and push what was
// entered, so a ray's node visits are processed 32 abreast and only the depth of
// the tree is left as a dependent chain. Closest hits are merged once per round
// (minimum t, then lowest id). The primitive tests and the tie rule are those of
// traverse_persistent, and the result of a closest-hit query does not depend on
// the visiting order, so the hits are the same bit for bit.
//
// Stack bound: a round that pops p entries pushes <= 4 p. Wide rounds keep the
// stack below CRT_WIDE_STACK - CRT_STACK_ENTRIES; past that the warp pops one
// entry per round, which is a depth-first walk and adds <= CRT_STACK_ENTRIES
// (3 per level, see above) to what was there.
#define CRT_WIDE_STACK 1024   // entries {reference, entry distance} per warp: 8 KB

// One ray, the whole warp: o and d (= ray_dir_local of the raw direction) are warp-uniform, and so is the result.
template <bool COUNT, bool HAS_SPH>
__device__ __forceinline__ Hit trace_wide_one(const AccelView& A, const v3 o, const v3 d_raw, const float t_min, const float t_max,
                                              uint2* s_stack, TraceCounters& cnt)
{
    constexpr uint32_t FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    constexpr uint32_t WIDE_LIMIT = CRT_WIDE_STACK - CRT_STACK_ENTRIES;
    {
        const v3 d = ray_dir_local(d_raw);
        const RaySlabs S = make_slabs(o, d);
        Hit best;                                    // warp-uniform
        best.t = t_max; best.k = 0xFFFFFFFFu; best.id = CRT_MISS_ID; best.u = best.v = 0.f;
        uint32_t sp = 0;                             // warp-uniform
        if (A.n_prims != 0u) {
            if (lane == 0) s_stack[0] = make_uint2((uint32_t)A.root, __float_as_uint(t_min));
            sp = 1;
        }
        __syncwarp();
        while (sp > 0u) {
            uint32_t p = sp < 32u ? sp : 32u;
            if (sp + 3u * p > WIDE_LIMIT) p = sp + 3u <= WIDE_LIMIT ? (WIDE_LIMIT - sp) / 3u : 1u;
            const bool have = lane < p;
            uint2 e = make_uint2(0u, 0u);
            if (have) e = s_stack[sp - 1u - lane];   // lane 0 takes the top (the nearest child pushed last)
            __syncwarp();
            sp -= p;
            const int ref = (int)e.x;
            // still entered before the current closest hit? (same slack as the test that pushed it)
            const bool alive = have && __uint_as_float(e.y) <= best.t * CRT_T_SLACK1;

            // ---- inner node: the four children against the ray ----
            float key[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
            int cref[4] = {0, 0, 0, 0};
            int n_hit = 0;
            if (alive && ref >= 0) {
                const float4* np = A.nodes + 8 * (size_t)(uint32_t)ref;
                float f[24];
                load_node(np, f);
                if (COUNT) { cnt.nodes++; cnt.node_f++; }
                cref[0] = __float_as_int(f[20]); cref[1] = __float_as_int(f[21]);
                cref[2] = __float_as_int(f[22]); cref[3] = __float_as_int(f[23]);
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    float hx, hy, hz;
                    child_box(f[5 * k], f[5 * k + 1], f[5 * k + 2], f[5 * k + 3], f[5 * k + 4], hx, hy, hz);
                    const float ax = fmaf(f[5 * k], S.ix, S.cx_lo), bx = fmaf(hx, S.ix, S.cx_hi);
                    const float ay = fmaf(f[5 * k + 1], S.iy, S.cy_lo), by = fmaf(hy, S.iy, S.cy_hi);
                    const float az = fmaf(f[5 * k + 2], S.iz, S.cz_lo), bz = fmaf(hz, S.iz, S.cz_hi);
                    const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), t_min));
                    const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), best.t));
                    const bool h = tn <= tf * CRT_T_SLACK1 && cref[k] != CRT_SLOT_EMPTY;
                    key[k] = h ? tn : INFINITY;
                    n_hit += h ? 1 : 0;
                }
                if (node_backfacing(f, d)) n_hit = 0;
#define CRT_CSWAP(a, b)                                                                    \
                {                                                                          \
                    const bool sw = key[b] < key[a];                                       \
                    const float ka = key[a], kb = key[b];                                  \
                    const int ra = cref[a], rb = cref[b];                                  \
                    key[a] = sw ? kb : ka; key[b] = sw ? ka : kb;                          \
                    cref[a] = sw ? rb : ra; cref[b] = sw ? ra : rb;                        \
                }
                CRT_CSWAP(0, 1) CRT_CSWAP(2, 3) CRT_CSWAP(0, 2) CRT_CSWAP(1, 3) CRT_CSWAP(1, 2)
#undef CRT_CSWAP
            }

            // ---- leaf: its primitives against the ray, this lane's closest candidate ----
            float ct = INFINITY, cu = 0.f, cv = 0.f;
            uint32_t ck = 0xFFFFFFFFu, cid = CRT_MISS_ID;
            if (alive && ref < 0) {
                const uint32_t lr = ~(uint32_t)ref;
                const uint32_t k0 = lr >> 3, count = (lr & 7u) + 1u;
                float bound = best.t;
                for (uint32_t j = 0; j < count; j++) {
                    const uint32_t k = k0 + j;
                    const float4* rp = A.recs + 4 * (size_t)k;
                    float4 r0, r1, r2, r3;
                    ldg256(rp, r0, r1);
                    ldg256(rp + 2, r2, r3);
                    const uint32_t pid = __float_as_uint(r3.x);
                    float t, u = 0.0f, v = 0.0f;
                    bool ok;
                    if (HAS_SPH && __float_as_uint(r3.z) != 0u) {
                        if (COUNT) { cnt.spheres++; cnt.sph_f++; }
                        float tuv[3];
                        ok = framed_test_dev(A.frames, __float_as_uint(r3.z), __float_as_uint(r3.y), r0, r1, r2, o, d_raw, t_min, bound, tuv);
                        t = tuv[0]; u = tuv[1]; v = tuv[2];
                    } else if (HAS_SPH && (__float_as_uint(r3.y) & CRT_KIND_SPHERE)) {
                        if (COUNT) { cnt.spheres++; cnt.sph_f++; }
                        ok = sphere_test(vsub(o, V(r0.x, r0.y, r0.z)), d, r0.w, t_min, bound, t);
                    } else {
                        if (COUNT) { cnt.tris++; cnt.tri_f++; }
                        ok = triangle_test(o, d, V(r0.x, r0.y, r0.z), V(r1.x, r1.y, r1.z), V(r2.x, r2.y, r2.z),
                                           V(r0.w, r1.w, r2.w), t_min, bound, t, u, v);
                    }
                    // against this lane's candidate: nearer, or as near with a lower id
                    if (ok && (t < ct || (t == ct && pid < cid))) { ct = t; ck = k; cid = pid; cu = u; cv = v; bound = t; }
                }
            }

            // ---- push the entered children, farthest first: 4 passes, pass c holds every lane's c-th farthest ----
#pragma unroll
            for (int c = 3; c >= 0; c--) {
                const bool has = n_hit > c;
                const uint32_t m = __ballot_sync(FULL, has);
                // within a pass the lane that popped the top of the stack writes last, so its children end on top
                if (has) s_stack[sp + (uint32_t)__popc(m & ~lt_mask & ~(1u << lane))] = make_uint2((uint32_t)cref[c], __float_as_uint(key[c]));
                sp += (uint32_t)__popc(m);
            }

            // ---- merge the round's candidates into the closest hit: minimum t, then lowest id ----
            const uint32_t any = __ballot_sync(FULL, ck != 0xFFFFFFFFu);
            if (any) {
                const uint32_t tbits = __float_as_uint(ct);                   // t > 0: bit patterns order like values
                const uint32_t tmin = __reduce_min_sync(FULL, ck != 0xFFFFFFFFu ? tbits : 0xFFFFFFFFu);
                const bool at_min = ck != 0xFFFFFFFFu && tbits == tmin;
                const uint32_t idmin = __reduce_min_sync(FULL, at_min ? cid : 0xFFFFFFFFu);
                const uint32_t w = (uint32_t)__ffs(__ballot_sync(FULL, at_min && cid == idmin)) - 1u;
                const float wt = __uint_as_float(tmin);
                const uint32_t wk = __shfl_sync(FULL, ck, w);
                const float wu = __shfl_sync(FULL, cu, w), wv = __shfl_sync(FULL, cv, w);
                // the rule of traverse_persistent: t == t_max itself never counts, equal t keeps the lower id
                bool take = wt < best.t || (wt == best.t && best.k != 0xFFFFFFFFu && idmin < best.id);
                if (take) { best.t = wt; best.k = wk; best.id = idmin; best.u = wu; best.v = wv; }
            }
            __syncwarp();
        }
        return best;
    }
}

template <bool COUNT, bool HAS_SPH, class Src, class Sink>
__device__ __forceinline__ void traverse_wide(const AccelView& A, const uint32_t n, uint32_t* work_ctr,
                                              const float t_min, const float t_max, const Src& src,
                                              const Sink& sink, uint2* s_stack, TraceCounters& cnt)
{
    const uint32_t lane = threadIdx.x & 31u;
    for (;;) {
        uint32_t ray_i = 0;
        if (lane == 0) ray_i = atomicAdd(work_ctr, 1u);
        ray_i = __shfl_sync(0xffffffffu, ray_i, 0);
        if (ray_i >= n) break;
        v3 o, dr;
        src.load(ray_i, o, dr);                     // the same address in every lane: one broadcast load
        const Hit best = trace_wide_one<COUNT, HAS_SPH>(A, o, dr, t_min, t_max, s_stack, cnt);
        if (lane == 0) sink.store(ray_i, best);
    }
}

}  // namespace crt
