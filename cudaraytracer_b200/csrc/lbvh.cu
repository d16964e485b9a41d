// lbvh.cu — GPU LBVH build and refit.
//
// Replaces create_BVH<<<1,1>>> (createScene.h:42-44 -> hitable/bvh.h:76-125:
// recursive median split with a sequential sort per node on ONE device thread)
// and UpdateBVH<<<1,1>>> (createScene.h:46-58 -> hitable/bvh.h:135-158:
// recursive refit on one thread) with data-parallel passes:
//   1. k_scene_bounds : primitive boxes -> scene box (warp shuffles + float atomics)
//   2. k_morton       : 63-bit Morton code of each box centre (21 bits per axis);
//                       bit 63 marks oversize primitives (see is_oversize)
//   3. radix_sort_pairs (radix_sort.cu): (code, primitive) pairs
//   4. k_hierarchy    : Karras 2012 binary radix tree, one thread per internal
//                       node, ties between equal codes broken by position
//   5. k_fit          : one thread per sorted primitive writes its 64-byte record
//                       and box, then climbs; the second thread to reach a node
//                       (atomic flag) merges the children's boxes
//   6. k_emit_nodes   : writes the 128-byte 4-wide traversal nodes (the boxes of
//                       a binary node's grandchildren); subtrees covering
//                       <= leaf_max primitives collapse into one leaf (a Karras
//                       node always covers a contiguous range of sorted primitives)
// Refit = steps 5 and 6 on the existing topology.
#include "crt_internal.h"
#include <cuda_fp16.h>
#include "crt_traverse.cuh"

namespace crt {

struct Box { float lx, ly, lz, hx, hy, hz; };

__device__ __forceinline__ Box box_union(const Box& a, const Box& b)
{
    Box r;
    r.lx = fminf(a.lx, b.lx); r.ly = fminf(a.ly, b.ly); r.lz = fminf(a.lz, b.lz);
    r.hx = fmaxf(a.hx, b.hx); r.hy = fmaxf(a.hy, b.hy); r.hz = fmaxf(a.hz, b.hz);
    return r;
}

// The primitives as uploaded: triangles first, then spheres, then rectangles; a primitive may sit under a chain of
// Transforms (crt_device.cuh:chain_ray), rectangles always do.
struct PrimSrc {
    const float* tri_in;
    const float* sph_in;
    const uint32_t *tri_mat, *tri_id, *sph_mat, *sph_id, *rect_mat, *rect_id;
    const uint32_t *tri_chain, *sph_chain, *rect_chain;   // null: no chains of that kind
    const DFrame* frames;
    const float* pad_bounds;                               // scene box that scales the padding of chained boxes, or null
    uint32_t n_tri, n_sph, n_rect;
};

__device__ __forceinline__ uint32_t prim_chain(const PrimSrc& S, uint32_t p)
{
    if (p < S.n_tri) return S.tri_chain ? S.tri_chain[p] : 0u;
    if (p < S.n_tri + S.n_sph) return S.sph_chain ? S.sph_chain[p - S.n_tri] : 0u;
    return S.rect_chain[p - S.n_tri - S.n_sph];
}

// A point of a chained primitive's own frame in world space: the inverse of chain_ray's map, innermost level first.
// The rotation matrices are orthonormal up to rounding, so the transpose inverts them to a few 1e-7.
__device__ __forceinline__ void chain_point_to_world(const DFrame* __restrict__ frames, uint32_t chain, float& x, float& y, float& z)
{
    const DFrame* f = frames + (chain >> 8);
    for (int i = (int)(chain & 255u) - 1; i >= 0; i--) {
        x += f[i].px; y += f[i].py; z += f[i].pz;
        if (f[i].has_rot) {
            const float* m = f[i].m;
            const float wx = m[0] * x + m[3] * y + m[6] * z, wy = m[1] * x + m[4] * y + m[7] * z, wz = m[2] * x + m[5] * y + m[8] * z;
            x = wx; y = wy; z = wz;
        }
    }
}

// Box of input primitive p: the reference's Triangle::bounding_box / Sphere::bounding_box / Rectangle::bounding_box
// (shapes/triangle.h:102-115, shapes/sphere.h:58-62, shapes/rectangle.h:11-14) in WORLD space, widened per axis by
// 2^-21 of its largest coordinate (see the end of this function).
// A primitive under a chain is tested in its own frame (crt_device.cuh:framed_test) on a ray that was rotated and
// translated in fp32, so where its world-space image lies is only known up to the rounding of that map, which grows
// with the distance of the ray origin: its box is padded by 2^-16 of the scene's largest coordinate per level
// (rays that start within 16 scene sizes are covered). The reference's own TransformAABB rotates only two corners
// (hitable/transform.h:22-31, marked "wrong?" by its author) and is not followed: the list scan, which is the
// behaviour reproduced, uses no boxes.
__device__ __forceinline__ Box prim_box(const PrimSrc& S, uint32_t p)
{
    Box b;
    const uint32_t chain = prim_chain(S, p);
    if (chain == 0u) {
        if (p < S.n_tri) {
            const float* t = S.tri_in + 12 * (size_t)p;
            const float x0 = t[0], y0 = t[1], z0 = t[2], x1 = t[3], y1 = t[4], z1 = t[5], x2 = t[6], y2 = t[7], z2 = t[8];
            b.lx = fminf(x0, fminf(x1, x2)); b.ly = fminf(y0, fminf(y1, y2)); b.lz = fminf(z0, fminf(z1, z2));
            b.hx = fmaxf(x0, fmaxf(x1, x2)); b.hy = fmaxf(y0, fmaxf(y1, y2)); b.hz = fmaxf(z0, fmaxf(z1, z2));
        } else {
            const float* s = S.sph_in + 4 * (size_t)(p - S.n_tri);
            const float r = fabsf(s[3]);
            b.lx = s[0] - r; b.ly = s[1] - r; b.lz = s[2] - r;
            b.hx = s[0] + r; b.hy = s[1] + r; b.hz = s[2] + r;
        }
    } else {
        float c[4][3];
        int nc;
        float r = 0.0f;
        if (p < S.n_tri) {
            const float* t = S.tri_in + 12 * (size_t)p;
            for (int k = 0; k < 3; k++) { c[k][0] = t[3 * k]; c[k][1] = t[3 * k + 1]; c[k][2] = t[3 * k + 2]; }
            nc = 3;
        } else if (p < S.n_tri + S.n_sph) {
            const float* s = S.sph_in + 4 * (size_t)(p - S.n_tri);
            c[0][0] = s[0]; c[0][1] = s[1]; c[0][2] = s[2];
            r = fabsf(s[3]);
            nc = 1;
        } else {
            for (int k = 0; k < 4; k++) { c[k][0] = (k & 1) ? 0.5f : -0.5f; c[k][1] = (k & 2) ? 0.5f : -0.5f; c[k][2] = 0.0f; }
            nc = 4;
        }
        b = Box{INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
        float local_max = 0.0f;                     // D3 pads the triangle's box in its OWN frame, by its coordinates there
        for (int k = 0; k < nc; k++) local_max = fmaxf(local_max, fmaxf(fabsf(c[k][0]), fmaxf(fabsf(c[k][1]), fabsf(c[k][2]))));
        for (int k = 0; k < nc; k++) {
            chain_point_to_world(S.frames, chain, c[k][0], c[k][1], c[k][2]);
            b.lx = fminf(b.lx, c[k][0]); b.ly = fminf(b.ly, c[k][1]); b.lz = fminf(b.lz, c[k][2]);
            b.hx = fmaxf(b.hx, c[k][0]); b.hy = fmaxf(b.hy, c[k][1]); b.hz = fmaxf(b.hz, c[k][2]);
        }
        float scale = fmaxf(fmaxf(fabsf(b.lx), fabsf(b.hx)), fmaxf(fmaxf(fabsf(b.ly), fabsf(b.hy)), fmaxf(fabsf(b.lz), fabsf(b.hz)))) + r;
        scale = fmaxf(scale, local_max);
        if (S.pad_bounds)
            for (int k = 0; k < 6; k++) scale = fmaxf(scale, fabsf(S.pad_bounds[k]));
        const float pad = r + scale * 1.52587890625e-5f * (float)(chain & 255u) + 1e-30f;   // 2^-16 per level
        if (!(scale < INFINITY)) b = Box{-INFINITY, -INFINITY, -INFINITY, INFINITY, INFINITY, INFINITY};   // NaN / inf geometry: never culled
        b.lx -= pad; b.ly -= pad; b.lz -= pad; b.hx += pad; b.hy += pad; b.hz += pad;
    }
    // every axis widened by 2^-21 of its largest coordinate (a few ulps: enough for the roundings between a hit point
    // and the planes, little enough that the boxes of a bounce ray's neighbouring triangles stay clear of its origin -
    // 2^-15 cost 25 % more triangle tests per ray on C3, profiles/r02_ab_d3.txt): twice what deviation D3 (crt_device.cuh:triangle_test)
    // pads the triangle's own box by, so that a hit D3 accepts is always inside every box on the way down
#ifndef CRT_BOX_PAD
#define CRT_BOX_PAD 4.76837158203125e-7f   // 2^-21
#endif
    const float k = CRT_BOX_PAD;
    const float tiny = 1e-29f;
    const float mx = fmaxf(fabsf(b.lx), fabsf(b.hx)) * k + tiny, my = fmaxf(fabsf(b.ly), fabsf(b.hy)) * k + tiny,
                mz = fmaxf(fabsf(b.lz), fabsf(b.hz)) * k + tiny;
    b.lx -= mx; b.ly -= my; b.lz -= mz;
    b.hx += mx; b.hy += my; b.hz += mz;
    return b;
}

__device__ __forceinline__ void atomic_min_f(float* a, float v)
{
    if (v >= 0.0f) atomicMin((int*)a, __float_as_int(v));
    else atomicMax((unsigned int*)a, __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_f(float* a, float v)
{
    if (v >= 0.0f) atomicMax((int*)a, __float_as_int(v));
    else atomicMin((unsigned int*)a, __float_as_uint(v));
}

// bounds[0..5]: box of all primitives; bounds[6..11]: box of the primitives that are not oversize.
__global__ void k_init_bounds(float* bounds)
{
    if (threadIdx.x < 12) bounds[threadIdx.x] = (threadIdx.x % 6) < 3 ? INFINITY : -INFINITY;
}

// A primitive is OVERSIZE when its box is longer than 1/8 of the scene box along
// the scene's longest axis (a ground quad under a finely tessellated mesh, the
// ground sphere of the random-spheres scene). In a Morton-ordered tree such a
// primitive sits as a leaf deep inside some spatial cell and inflates the box of
// every ancestor up to the root, so every ray walks that chain first. Oversize
// primitives get key bit 63 set: they sort behind everything else and form
// their own subtree directly under the root, and the Morton grid is laid over
// the remaining primitives only.
__device__ __forceinline__ bool is_oversize(const Box& b, const float* __restrict__ bounds)
{
    const float scene = fmaxf(bounds[3] - bounds[0], fmaxf(bounds[4] - bounds[1], bounds[5] - bounds[2]));
    const float prim = fmaxf(b.hx - b.lx, fmaxf(b.hy - b.ly, b.hz - b.lz));
    return prim > 0.125f * scene;
}

__global__ void __launch_bounds__(256)
k_scene_bounds(PrimSrc S, uint32_t n, const float* __restrict__ all_bounds, float* __restrict__ bounds)
{
    // all_bounds == nullptr: pass 1, every primitive; otherwise pass 2, only those that are not oversize
    Box b = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        const Box pb = prim_box(S, p);
        if (all_bounds == nullptr || !is_oversize(pb, all_bounds)) b = box_union(b, pb);
    }
    for (int o = 16; o > 0; o >>= 1) {
        b.lx = fminf(b.lx, __shfl_xor_sync(0xffffffffu, b.lx, o));
        b.ly = fminf(b.ly, __shfl_xor_sync(0xffffffffu, b.ly, o));
        b.lz = fminf(b.lz, __shfl_xor_sync(0xffffffffu, b.lz, o));
        b.hx = fmaxf(b.hx, __shfl_xor_sync(0xffffffffu, b.hx, o));
        b.hy = fmaxf(b.hy, __shfl_xor_sync(0xffffffffu, b.hy, o));
        b.hz = fmaxf(b.hz, __shfl_xor_sync(0xffffffffu, b.hz, o));
    }
    if ((threadIdx.x & 31) == 0 && b.lx <= b.hx) {
        atomic_min_f(bounds + 0, b.lx); atomic_min_f(bounds + 1, b.ly); atomic_min_f(bounds + 2, b.lz);
        atomic_max_f(bounds + 3, b.hx); atomic_max_f(bounds + 4, b.hy); atomic_max_f(bounds + 5, b.hz);
    }
}

__device__ __forceinline__ uint64_t expand21(uint64_t v)
{
    v &= 0x1fffffull;
    v = (v | v << 32) & 0x1f00000000ffffull;
    v = (v | v << 16) & 0x1f0000ff0000ffull;
    v = (v | v << 8) & 0x100f00f00f00f00full;
    v = (v | v << 4) & 0x10c30c30c30c30c3ull;
    v = (v | v << 2) & 0x1249249249249249ull;
    return v;
}

__global__ void __launch_bounds__(256)
k_morton(PrimSrc S, uint32_t n, const float* __restrict__ bounds, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const Box b = prim_box(S, p);
    const bool oversize = is_oversize(b, bounds);
    // grid over the regular primitives; if every primitive is oversize, over all of them
    const float* gb = bounds[6] <= bounds[9] ? bounds + 6 : bounds;
    const float lx = gb[0], ly = gb[1], lz = gb[2];
    const float ex = gb[3] - lx, ey = gb[4] - ly, ez = gb[5] - lz;
    const float scale = 2097152.0f;   // 2^21
    const float fx = ex > 0.0f ? (0.5f * (b.lx + b.hx) - lx) / ex : 0.0f;
    const float fy = ey > 0.0f ? (0.5f * (b.ly + b.hy) - ly) / ey : 0.0f;
    const float fz = ez > 0.0f ? (0.5f * (b.lz + b.hz) - lz) / ez : 0.0f;
    const uint64_t qx = (uint64_t)fminf(fmaxf(fx * scale, 0.0f), scale - 1.0f);
    const uint64_t qy = (uint64_t)fminf(fmaxf(fy * scale, 0.0f), scale - 1.0f);
    const uint64_t qz = (uint64_t)fminf(fmaxf(fz * scale, 0.0f), scale - 1.0f);
    keys[p] = (expand21(qx) << 2) | (expand21(qy) << 1) | expand21(qz) | (oversize ? (1ull << 63) : 0ull);
    vals[p] = p;
}

// Common-prefix length of sorted codes i and j; equal codes fall back to the
// positions so the tree stays a proper binary tree with duplicate codes.
__device__ __forceinline__ int delta(const uint64_t* __restrict__ keys, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    const uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz(i ^ j);
    return __clzll((long long)(a ^ b));
}

// Karras, "Maximizing Parallelism in the Construction of BVHs, Octrees, and
// k-d Trees" (HPG 2012), Algorithm 1. child: >= 0 internal node, < 0: ~leaf position.
__global__ void __launch_bounds__(256)
k_hierarchy(const uint64_t* __restrict__ keys, int n, int2* __restrict__ child, int2* __restrict__ range,
            int* __restrict__ parent, int* __restrict__ leaf_parent)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    const int left = (first == gamma) ? ~gamma : gamma;
    const int right = (last == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    child[i] = make_int2(left, right);
    range[i] = make_int2(first, last);
    if (left < 0) leaf_parent[gamma] = i; else parent[left] = i;
    if (right < 0) leaf_parent[gamma + 1] = i; else parent[right] = i;
    if (i == 0) parent[0] = -1;
}

__device__ __forceinline__ Box load_box(const float4* __restrict__ base, size_t idx)
{
    const float4 a = __ldcg(base + 2 * idx), b = __ldcg(base + 2 * idx + 1);
    Box r = {a.x, a.y, a.z, b.x, b.y, b.z};
    return r;
}
__device__ __forceinline__ void store_box(float4* __restrict__ base, size_t idx, const Box& b)
{
    base[2 * idx] = make_float4(b.lx, b.ly, b.lz, 0.0f);
    base[2 * idx + 1] = make_float4(b.hx, b.hy, b.hz, 0.0f);
}

// ---- normal cones -----------------------------------------------------------
// The reference's triangle test starts with `if (dot(direction, normal) < 0) return false` (shapes/triangle.h:61,
// SURVEY 9 Q1): a triangle whose STORED normal points against the ray is never hit. Every subtree therefore gets a
// cone {unit axis, cos of the half-angle} that contains the stored normals of all its triangles (a sphere anywhere
// below makes it "none": c = -2), merged bottom-up beside the boxes; k_emit_nodes turns the cone of a wide node's own
// subtree into a vector A with  dot(unit direction, A) < -1  =>  dot(direction, n) < 0 for every normal n below, and
// the per-ray traversals drop such a node before testing its children (crt_traverse.cuh:node_backfacing). The cones
// depend on the stored normals only, which no pose update changes (Triangle::SetVertices keeps them).
struct Cone { float x, y, z, c; };
__device__ __forceinline__ Cone cone_none() { return Cone{0.f, 0.f, 0.f, -2.f}; }

__device__ __forceinline__ Cone cone_of_normal(float nx, float ny, float nz)
{
    const float l2 = nx * nx + ny * ny + nz * nz;
    if (!(l2 > 1e-30f) || !(l2 < 1e30f)) return cone_none();    // zero, huge or NaN: never culled
    const float r = rsqrtf(l2);
    return Cone{nx * r, ny * r, nz * r, 1.0f};
}

// Smallest cone that contains both cones: the wider one if it already contains the other, else half-angle
// (theta + da + db) / 2 around the axis of a turned towards the axis of b by (half-angle - da); inflated by 5e-4 rad
// per merge for the rounding of the construction; "none" beyond a half-angle of 1.5 rad (nothing can be culled by a
// cone that wide).
__device__ __forceinline__ Cone cone_merge(const Cone& a, const Cone& b)
{
    if (a.c < -1.5f || b.c < -1.5f) return cone_none();
    // angle between the axes from the cross product (acos of a dot near 1 loses half the bits)
    const float cx = a.y * b.z - a.z * b.y, cy = a.z * b.x - a.x * b.z, cz = a.x * b.y - a.y * b.x;
    const float sin_t = sqrtf(cx * cx + cy * cy + cz * cz), cos_t = a.x * b.x + a.y * b.y + a.z * b.z;
    const float theta = atan2f(sin_t, cos_t);
    const float da = acosf(fminf(a.c, 1.0f)), db = acosf(fminf(b.c, 1.0f));
    if (theta + db <= da) return Cone{a.x, a.y, a.z, cosf(fminf(da + 5e-4f, 3.0f))};
    if (theta + da <= db) return Cone{b.x, b.y, b.z, cosf(fminf(db + 5e-4f, 3.0f))};
    if (!(sin_t > 1e-6f)) {                                     // parallel axes (antiparallel ones give theta = pi below)
        if (cos_t < 0.0f) return cone_none();
        return Cone{a.x, a.y, a.z, cosf(fminf(fmaxf(da, db) + theta + 5e-4f, 3.0f))};
    }
    const float delta = 0.5f * (theta + da + db) + 5e-4f;
    if (!(delta < 1.5f)) return cone_none();
    // spherical interpolation from a towards b by phi = delta - da (0 <= phi <= theta)
    const float phi = fminf(fmaxf(delta - da, 0.0f), theta);
    const float wa = sinf(theta - phi) / sin_t, wb = sinf(phi) / sin_t;
    float mx = wa * a.x + wb * b.x, my = wa * a.y + wb * b.y, mz = wa * a.z + wb * b.z;
    const float r = rsqrtf(mx * mx + my * my + mz * mz);
    return Cone{mx * r, my * r, mz * r, cosf(delta)};
}

// Cone of the sorted primitive at position k, from its record (written by the thread that owns the leaf).
__device__ __forceinline__ Cone leaf_cone(const float4* __restrict__ recs, size_t k)
{
    const float4 r0 = __ldcg(recs + 4 * k), r1 = __ldcg(recs + 4 * k + 1), r2 = __ldcg(recs + 4 * k + 2), r3 = __ldcg(recs + 4 * k + 3);
    // spheres and rectangles have no stored normal to cull by; a chained triangle's normal test runs in its own frame
    if ((__float_as_uint(r3.y) & (CRT_KIND_SPHERE | CRT_KIND_RECT)) || __float_as_uint(r3.z) != 0u) return cone_none();
    return cone_of_normal(r0.w, r1.w, r2.w);
}

__device__ __forceinline__ Cone load_cone(const float4* __restrict__ base, size_t idx)
{
    const float4 v = __ldcg(base + idx);
    return Cone{v.x, v.y, v.z, v.w};
}

// The cone as the vector the traversal tests, three halves: A = axis / (sin(half-angle) + 0.01). For a unit direction d,
// dot(d, A) < -1 puts d within 90 deg - half-angle - 0.5 deg of -axis, so dot(d, n) < -0.008 for every unit normal n in
// the cone: far below the few 1e-7 by which the reference's float evaluation of dot(direction, normal) can differ
// from the true value, and the 0.01 also covers the rounding of A to halves (direction 7e-4 rad, length 5e-4
// relative) and of the traversal's three FMAs. All zero = never culled.
__device__ __forceinline__ void cone_to_halves(const Cone& c, uint32_t& hx, uint32_t& hy, uint32_t& hz)
{
    hx = hy = hz = 0u;
    if (c.c < -1.5f || !(c.c > 0.0707f)) return;                // none, or wider than 1.5 rad
    const float s = sqrtf(fmaxf(0.0f, 1.0f - c.c * c.c));
    const float L = 1.0f / (s + 0.01f);
    hx = (uint32_t)__half_as_ushort(__float2half_rn(c.x * L));
    hy = (uint32_t)__half_as_ushort(__float2half_rn(c.y * L));
    hz = (uint32_t)__half_as_ushort(__float2half_rn(c.z * L));
}

// CONES: also merge the normal cones (a build; a refit leaves them alone: no pose update changes a stored normal)
template <bool CONES>
__global__ void __launch_bounds__(256)
k_fit(PrimSrc S, uint32_t n, const uint32_t* __restrict__ perm, const int2* __restrict__ child,
      const int* __restrict__ parent, const int* __restrict__ leaf_parent, int* __restrict__ flags,
      float4* __restrict__ recs, float4* __restrict__ leaf_box, float4* __restrict__ node_box,
      float4* __restrict__ node_cone)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t p = perm[k];
    const uint32_t n_tri = S.n_tri;
    const float chain_bits = __uint_as_float(prim_chain(S, p));
    if (p < n_tri) {
        const float* t = S.tri_in + 12 * (size_t)p;
        const v3 v0 = V(t[0], t[1], t[2]);
        const v3 e1 = vsub(V(t[3], t[4], t[5]), v0);       // triangle.h:69
        const v3 e2 = vsub(V(t[6], t[7], t[8]), v0);       // triangle.h:70
        recs[4 * (size_t)k] = make_float4(v0.x, v0.y, v0.z, t[9]);
        recs[4 * (size_t)k + 1] = make_float4(e1.x, e1.y, e1.z, t[10]);
        recs[4 * (size_t)k + 2] = make_float4(e2.x, e2.y, e2.z, t[11]);
        recs[4 * (size_t)k + 3] = make_float4(__uint_as_float(S.tri_id[p]), __uint_as_float(S.tri_mat[p]), chain_bits, 0.f);
    } else if (p < n_tri + S.n_sph) {
        const uint32_t q = p - n_tri;
        const float* s = S.sph_in + 4 * (size_t)q;
        recs[4 * (size_t)k] = make_float4(s[0], s[1], s[2], s[3]);
        recs[4 * (size_t)k + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
        recs[4 * (size_t)k + 2] = make_float4(0.f, 0.f, 0.f, 0.f);
        recs[4 * (size_t)k + 3] = make_float4(__uint_as_float(S.sph_id[q]), __uint_as_float(S.sph_mat[q] | CRT_KIND_SPHERE), chain_bits, 0.f);
    } else {
        const uint32_t q = p - n_tri - S.n_sph;                  // rect_mat carries CRT_RECT_FLIP
        recs[4 * (size_t)k] = make_float4(0.f, 0.f, 0.f, 0.f);
        recs[4 * (size_t)k + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
        recs[4 * (size_t)k + 2] = make_float4(0.f, 0.f, 0.f, 0.f);
        recs[4 * (size_t)k + 3] = make_float4(__uint_as_float(S.rect_id[q]), __uint_as_float(S.rect_mat[q] | CRT_KIND_RECT), chain_bits, 0.f);
    }
    store_box(leaf_box, k, prim_box(S, p));
    if (n < 2) return;
    __threadfence();
    int node = leaf_parent[k];
    while (node >= 0) {
        if (atomicAdd(&flags[node], 1) == 0) return;       // first arrival: the sibling is not done
        __threadfence();
        const int2 ch = child[node];
        const Box a = ch.x < 0 ? load_box(leaf_box, (size_t)(~ch.x)) : load_box(node_box, (size_t)ch.x);
        const Box b = ch.y < 0 ? load_box(leaf_box, (size_t)(~ch.y)) : load_box(node_box, (size_t)ch.y);
        store_box(node_box, (size_t)node, box_union(a, b));
        if (CONES) {
            const Cone ca = ch.x < 0 ? leaf_cone(recs, (size_t)(~ch.x)) : load_cone(node_cone, (size_t)ch.x);
            const Cone cb = ch.y < 0 ? leaf_cone(recs, (size_t)(~ch.y)) : load_cone(node_cone, (size_t)ch.y);
            const Cone cm = cone_merge(ca, cb);
            node_cone[node] = make_float4(cm.x, cm.y, cm.z, cm.c);
        }
        __threadfence();
        node = parent[node];
    }
}

// One 4-wide traversal node per binary node i that is not collapsed into a
// leaf: its slots start as i's two children, and the inner slot with the
// largest surface area is replaced by its own two children until four slots
// are in use (a leaf, or a subtree of <= leaf_max primitives, stays a slot). Every binary node gets its
// wide node at the same index, so no renumbering pass is needed; only the
// nodes reachable from the root through slot references are ever fetched.
struct Slot { int ref; Box box; };

// The extent hi - lo as a half, rounded up until the traversal's own reconstruction lo + half (one fp32 addition,
// crt_traverse.cuh:child_box) reaches hi; +inf when the extent is beyond the half range (the box is then open
// upwards, which is conservative).
__device__ __forceinline__ uint32_t half_extent_up(float lo, float hi)
{
    unsigned short h = __half_as_ushort(__float2half_ru(hi - lo));
    while (h < 0x7C00u && __fadd_rn(lo, __half2float(__ushort_as_half(h))) < hi) h++;
    return (uint32_t)h;
}

__device__ __forceinline__ bool resolve_child(int c, int leaf_max, const int2* __restrict__ range,
                                              const float4* __restrict__ leaf_box, const float4* __restrict__ node_box,
                                              Slot& out)
{
    if (c < 0) {
        const uint32_t pos = (uint32_t)(~c);
        out.ref = ~(int)(pos << 3);
        out.box = load_box(leaf_box, pos);
        return true;                                       // a leaf
    }
    const int2 cr = range[c];
    const int size = cr.y - cr.x + 1;
    out.box = load_box(node_box, (size_t)c);
    if (size <= leaf_max) {
        out.ref = ~(int)(((uint32_t)cr.x << 3) | (uint32_t)(size - 1));
        return true;
    }
    out.ref = c;
    return false;                                          // an inner node
}

__global__ void __launch_bounds__(256)
k_emit_nodes(int n, int leaf_max, const int2* __restrict__ child, const int2* __restrict__ range,
             const float4* __restrict__ leaf_box, const float4* __restrict__ node_box,
             const float4* __restrict__ node_cone, float4* __restrict__ nodes)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int2 rg = range[i];
    if (rg.y - rg.x + 1 <= leaf_max) return;               // collapsed into a leaf of an ancestor
    const int2 ch = child[i];
    Slot slot[4];
    bool inner[4] = {false, false, false, false};
    int ns = 2;
    inner[0] = !resolve_child(ch.x, leaf_max, range, leaf_box, node_box, slot[0]);
    inner[1] = !resolve_child(ch.y, leaf_max, range, leaf_box, node_box, slot[1]);
    // open the inner slot with the largest surface area until the node is full: a large
    // child is the one most rays enter anyway, so its two children take its place
    while (ns < 4) {
        int pick = -1;
        float best = -1.0f;
        for (int s = 0; s < ns; s++) {
            if (!inner[s]) continue;
            const Box& b = slot[s].box;
            const float dx = b.hx - b.lx, dy = b.hy - b.ly, dz = b.hz - b.lz;
            const float area = dx * dy + dy * dz + dz * dx;
            if (area > best) { best = area; pick = s; }
        }
        if (pick < 0) break;
        const int2 gc = child[slot[pick].ref];
        inner[pick] = !resolve_child(gc.x, leaf_max, range, leaf_box, node_box, slot[pick]);
        inner[ns] = !resolve_child(gc.y, leaf_max, range, leaf_box, node_box, slot[ns]);
        ns++;
    }
    for (int s = ns; s < 4; s++) {
        slot[s].ref = CRT_SLOT_EMPTY;
        slot[s].box = Box{0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    }
    float4* np = nodes + 8 * (size_t)i;
    uint32_t w[24];                                           // layout: crt_device.cuh
    for (int s = 0; s < 4; s++) {
        const Box& b = slot[s].box;
        w[5 * s] = __float_as_uint(b.lx); w[5 * s + 1] = __float_as_uint(b.ly); w[5 * s + 2] = __float_as_uint(b.lz);
        const uint32_t ex = half_extent_up(b.lx, b.hx), ey = half_extent_up(b.ly, b.hy), ez = half_extent_up(b.lz, b.hz);
        w[5 * s + 3] = ex | (ey << 16);
        w[5 * s + 4] = ez;
        w[20 + s] = (uint32_t)slot[s].ref;
    }
    // the normal cone of this node's own subtree rides in the spare halves of children 0-2 (layout: crt_device.cuh)
    uint32_t ax, ay, az;
    cone_to_halves(load_cone(node_cone, (size_t)i), ax, ay, az);
    w[4] |= ax << 16; w[9] |= ay << 16; w[14] |= az << 16;
    uint4* nw = reinterpret_cast<uint4*>(np);
    for (int q = 0; q < 6; q++) nw[q] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
}

__global__ void __launch_bounds__(256)
k_depth(int n, const int* __restrict__ parent, const int* __restrict__ leaf_parent, int* __restrict__ out)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int d = 1;
    for (int node = leaf_parent[k]; node >= 0; node = parent[node]) d++;
    const unsigned int act = __activemask();
    d = __reduce_max_sync(act, d);
    if ((threadIdx.x & 31) == (unsigned int)(__ffs(act) - 1)) atomicMax(out, d);
}

// ---------------------------------------------------------------------------
// PLOC topology (Meister & Bittner, "Parallel Locally-Ordered Clustering for
// BVH Construction", TVCG 2018) as the alternative to k_hierarchy: the
// Morton-sorted primitives start as one cluster each; per round every cluster
// looks PLOC_R positions to either side for the partner whose union with it has
// the smallest surface area, mutual partners merge into a new node that takes
// the left partner's place, and the array is compacted in order. Unlike the
// radix tree, whose splits follow the bits of the code, a merge looks at the
// boxes themselves, so large primitives stay unmerged until the clusters around
// them have grown and the tree costs fewer node visits per ray.
// Candidate order: (area, |i - j|, parity of min(i, j), min(i, j)) is a strict
// total order on unordered pairs, so the smallest pair of a round is always
// mutual (every round merges something) and n identical boxes pair up as
// (0,1)(2,3).. instead of chaining.
// Node ids are handed out from n-2 downwards, in array order inside a round:
// the last merge is node 0, the root, as in the radix tree, and the nodes of a
// level lie in Morton order. A final pass renumbers the leaves in depth-first
// order so that every subtree covers a contiguous range of records again
// (k_emit_nodes folds small subtrees into leaves by range).
constexpr int PLOC_R = 32;   // largest search radius (shared-memory halo)
constexpr int PLOC_B = 256;

__global__ void __launch_bounds__(256)
k_ploc_init(PrimSrc S, uint32_t n, const uint32_t* __restrict__ perm, float4* __restrict__ lo, float4* __restrict__ hi)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Box b = prim_box(S, perm[i]);
    lo[i] = make_float4(b.lx, b.ly, b.lz, __int_as_float(~(int)i));   // .w: reference of the cluster's subtree
    hi[i] = make_float4(b.hx, b.hy, b.hz, __int_as_float(1));         // .w: primitives below it
}

__global__ void __launch_bounds__(PLOC_B)
k_ploc_nn(int nc, int radius, const float4* __restrict__ lo, const float4* __restrict__ hi, int* __restrict__ nn)
{
    __shared__ float4 s_lo[PLOC_B + 2 * PLOC_R], s_hi[PLOC_B + 2 * PLOC_R];
    const int base = (int)blockIdx.x * PLOC_B - PLOC_R;
    for (int t = threadIdx.x; t < PLOC_B + 2 * PLOC_R; t += PLOC_B) {
        const int g = base + t;
        if (g >= 0 && g < nc) { s_lo[t] = lo[g]; s_hi[t] = hi[g]; }
    }
    __syncthreads();
    const int i = (int)blockIdx.x * PLOC_B + (int)threadIdx.x;
    if (i >= nc) return;
    const float4 al = s_lo[threadIdx.x + PLOC_R], ah = s_hi[threadIdx.x + PLOC_R];
    unsigned long long best = ~0ull;
    int best_j = -1;
#pragma unroll 4
    for (int o = -radius; o <= radius; o++) {
        const int j = i + o;
        if (o == 0 || j < 0 || j >= nc) continue;
        const float4 bl = s_lo[threadIdx.x + PLOC_R + o], bh = s_hi[threadIdx.x + PLOC_R + o];
        const float dx = fmaxf(ah.x, bh.x) - fminf(al.x, bl.x);
        const float dy = fmaxf(ah.y, bh.y) - fminf(al.y, bl.y);
        const float dz = fmaxf(ah.z, bh.z) - fminf(al.z, bl.z);
        const float area = dx * dy + dy * dz + dz * dx;            // >= 0: its bits order like the value
        const uint32_t g = (uint32_t)(o < 0 ? -o : o);
        const uint32_t m = (uint32_t)(o < 0 ? j : i);
        const unsigned long long key = ((unsigned long long)__float_as_uint(area) << 32) |
                                       (unsigned long long)((g << 2) | ((m & 1u) << 1) | (o > 0 ? 1u : 0u));
        if (key < best) { best = key; best_j = j; }
    }
    nn[i] = best_j;
}

// exclusive scan of one int per thread over a PLOC_B-thread block; *total = block sum
__device__ __forceinline__ int block_exscan(int v, int* s_warp, int* total)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int x = v;
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    __syncthreads();
    if (lane == 31) s_warp[w] = x;
    __syncthreads();
    if (w == 0) {
        int t = lane < PLOC_B / 32 ? s_warp[lane] : 0;
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, t, o);
            if (lane >= o) t += y;
        }
        if (lane < PLOC_B / 32) s_warp[lane] = t;                  // inclusive warp sums
    }
    __syncthreads();
    *total = s_warp[PLOC_B / 32 - 1];
    return x - v + (w > 0 ? s_warp[w - 1] : 0);
}

// role of cluster i in this round: 0 = stays, 1 = left partner (becomes the merged node), 2 = right partner (leaves)
__device__ __forceinline__ int ploc_role(int nc, const int* __restrict__ nn, int i)
{
    if (i >= nc) return 2;
    const int j = nn[i];
    if (j < 0 || nn[j] != i) return 0;
    return i < j ? 1 : 2;
}

__global__ void __launch_bounds__(PLOC_B)
k_ploc_count(int nc, const int* __restrict__ nn, int2* __restrict__ blk)
{
    __shared__ int s_a[PLOC_B / 32], s_b[PLOC_B / 32];
    const int i = (int)blockIdx.x * PLOC_B + (int)threadIdx.x;
    const int role = ploc_role(nc, nn, i);
    const unsigned int keep = __ballot_sync(0xffffffffu, role != 2), merge = __ballot_sync(0xffffffffu, role == 1);
    if ((threadIdx.x & 31) == 0) { s_a[threadIdx.x >> 5] = __popc(keep); s_b[threadIdx.x >> 5] = __popc(merge); }
    __syncthreads();
    if (threadIdx.x == 0) {
        int a = 0, b = 0;
        for (int w = 0; w < PLOC_B / 32; w++) { a += s_a[w]; b += s_b[w]; }
        blk[blockIdx.x] = make_int2(a, b);
    }
}

// one block: blk[] -> exclusive prefix sums in place, the totals to *total
__global__ void __launch_bounds__(1024)
k_ploc_scan(int nb, int2* __restrict__ blk, int2* __restrict__ total)
{
    __shared__ int2 s_w[32];
    __shared__ int2 s_carry;
    if (threadIdx.x == 0) s_carry = make_int2(0, 0);
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int b0 = 0; b0 < nb; b0 += 1024) {
        const int b = b0 + (int)threadIdx.x;
        const int2 v = b < nb ? blk[b] : make_int2(0, 0);
        int2 x = v;
        for (int o = 1; o < 32; o <<= 1) {
            const int yx = __shfl_up_sync(0xffffffffu, x.x, o), yy = __shfl_up_sync(0xffffffffu, x.y, o);
            if (lane >= o) { x.x += yx; x.y += yy; }
        }
        if (lane == 31) s_w[w] = x;
        __syncthreads();
        if (w == 0) {
            int2 t = s_w[lane];
            for (int o = 1; o < 32; o <<= 1) {
                const int yx = __shfl_up_sync(0xffffffffu, t.x, o), yy = __shfl_up_sync(0xffffffffu, t.y, o);
                if (lane >= o) { t.x += yx; t.y += yy; }
            }
            s_w[lane] = t;
        }
        __syncthreads();
        const int2 carry = s_carry;
        const int2 wp = w > 0 ? s_w[w - 1] : make_int2(0, 0);
        if (b < nb) blk[b] = make_int2(carry.x + wp.x + x.x - v.x, carry.y + wp.y + x.y - v.y);
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = make_int2(carry.x + wp.x + x.x, carry.y + wp.y + x.y);
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = s_carry;
}

__global__ void __launch_bounds__(PLOC_B)
k_ploc_merge(int nc, const int* __restrict__ nn, const float4* __restrict__ lo_in, const float4* __restrict__ hi_in,
             float4* __restrict__ lo_out, float4* __restrict__ hi_out, const int2* __restrict__ blk, int id_top,
             int2* __restrict__ child, int* __restrict__ parent, int* __restrict__ leaf_parent, int* __restrict__ size)
{
    __shared__ int s_warp[PLOC_B / 32];
    const int i = (int)blockIdx.x * PLOC_B + (int)threadIdx.x;
    const int role = ploc_role(nc, nn, i);
    int tot;
    const int k_keep = block_exscan(role != 2 ? 1 : 0, s_warp, &tot);
    const int k_merge = block_exscan(role == 1 ? 1 : 0, s_warp, &tot);
    if (role == 2) return;
    const int2 off = blk[blockIdx.x];
    const int pos = off.x + k_keep;
    float4 l = lo_in[i], h = hi_in[i];
    if (role == 1) {
        const int j = nn[i];
        const float4 bl = lo_in[j], bh = hi_in[j];
        const int id = id_top - (off.y + k_merge);
        const int ra = __float_as_int(l.w), rb = __float_as_int(bl.w);
        child[id] = make_int2(ra, rb);
        if (ra < 0) leaf_parent[~ra] = id; else parent[ra] = id;
        if (rb < 0) leaf_parent[~rb] = id; else parent[rb] = id;
        const int sz = __float_as_int(h.w) + __float_as_int(bh.w);
        size[id] = sz;
        l = make_float4(fminf(l.x, bl.x), fminf(l.y, bl.y), fminf(l.z, bl.z), __int_as_float(id));
        h = make_float4(fmaxf(h.x, bh.x), fmaxf(h.y, bh.y), fmaxf(h.z, bh.z), __int_as_float(sz));
    }
    lo_out[pos] = l;
    hi_out[pos] = h;
}

// Depth-first position of sorted primitive i: the primitives in the left
// subtrees of the ancestors it reaches from the right. The same walk gives the
// first position of an inner node.
__device__ __forceinline__ int ploc_first_pos(int ref, int up, const int2* __restrict__ child,
                                              const int* __restrict__ parent, const int* __restrict__ size)
{
    int off = 0;
    for (int node = up; node >= 0; node = parent[node]) {
        const int2 ch = child[node];
        if (ch.y == ref) off += ch.x < 0 ? 1 : size[ch.x];
        ref = node;
    }
    return off;
}

__global__ void __launch_bounds__(256)
k_ploc_leaf_pos(int n, const int2* __restrict__ child, const int* __restrict__ parent, const int* __restrict__ leaf_parent,
                const int* __restrict__ size, const uint32_t* __restrict__ perm_in, int* __restrict__ newpos,
                uint32_t* __restrict__ perm_out, int* __restrict__ leaf_parent_out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int p = ploc_first_pos(~i, leaf_parent[i], child, parent, size);
    newpos[i] = p;
    perm_out[p] = perm_in[i];
    leaf_parent_out[p] = leaf_parent[i];
}

__global__ void __launch_bounds__(256)
k_ploc_node_range(int n, int2* __restrict__ child, const int* __restrict__ parent, const int* __restrict__ size,
                  const int* __restrict__ newpos, int2* __restrict__ range)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int first = ploc_first_pos(i, parent[i], child, parent, size);
    range[i] = make_int2(first, first + size[i] - 1);
}

// after every walk is done: leaf references switch to the depth-first numbering
__global__ void __launch_bounds__(256)
k_ploc_relabel(int n, int2* __restrict__ child, const int* __restrict__ newpos)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int2 ch = child[i];
    if (ch.x < 0) ch.x = ~newpos[~ch.x];
    if (ch.y < 0) ch.y = ~newpos[~ch.y];
    child[i] = ch;
}

static inline unsigned int grid_for(size_t n, int block) { return (unsigned int)((n + block - 1) / block); }

// PLOC rounds on the Morton-sorted primitives, then the depth-first renumbering.
// Leaves child / parent / leaf_parent / range and the permutation in the state
// k_hierarchy leaves them in (root = node 0, leaf references = record positions).
static int ploc_topology(crt_ctx* c, const PrimSrc& S, uint32_t n)
{
    cudaStream_t st = c->stream;
    CRT_CUDA(c->ploc_box[0].reserve((size_t)n * 32));
    CRT_CUDA(c->ploc_box[1].reserve((size_t)n * 32));
    CRT_CUDA(c->ploc_nn.reserve((size_t)n * 4));
    CRT_CUDA(c->ploc_tmp.reserve((size_t)n * 4));
    const unsigned int nb_max = grid_for(n, PLOC_B);
    CRT_CUDA(c->ploc_blk.reserve((size_t)(nb_max + 1) * sizeof(int2)));
    int2* blk = c->ploc_blk.as<int2>();
    int2* d_total = blk + nb_max;
    float4* lo[2] = {c->ploc_box[0].as<float4>(), c->ploc_box[1].as<float4>()};
    float4* hi[2] = {lo[0] + n, lo[1] + n};
    int* size = c->flags.as<int>();                         // free until k_fit: primitives below each inner node
    k_ploc_init<<<grid_for(n, 256), 256, 0, st>>>(S, n, c->vals[c->perm_buf].as<uint32_t>(), lo[0], hi[0]);
    int nc = (int)n, created = 0, cur = 0;
    while (nc > 1) {
        const unsigned int nb = grid_for((size_t)nc, PLOC_B);
        k_ploc_nn<<<nb, PLOC_B, 0, st>>>(nc, c->ploc_radius, lo[cur], hi[cur], c->ploc_nn.as<int>());
        k_ploc_count<<<nb, PLOC_B, 0, st>>>(nc, c->ploc_nn.as<int>(), blk);
        k_ploc_scan<<<1, 1024, 0, st>>>((int)nb, blk, d_total);
        k_ploc_merge<<<nb, PLOC_B, 0, st>>>(nc, c->ploc_nn.as<int>(), lo[cur], hi[cur], lo[cur ^ 1], hi[cur ^ 1], blk,
                                            (int)n - 2 - created, c->child.as<int2>(), c->parent.as<int>(),
                                            c->leaf_parent.as<int>(), size);
        int2 total = make_int2(0, 0);
        CRT_CUDA(cudaMemcpyAsync(&total, d_total, sizeof(int2), cudaMemcpyDeviceToHost, st));
        CRT_CUDA(cudaStreamSynchronize(st));
        if (total.y <= 0 || total.x != nc - total.y) { set_error("PLOC build: round made no progress (%d clusters, %d merges)", nc, total.y); return CRT_ERR_CUDA; }
        created += total.y;
        nc = total.x;
        cur ^= 1;
    }
    const int root_parent = -1;
    CRT_CUDA(cudaMemcpyAsync(c->parent.p, &root_parent, sizeof(int), cudaMemcpyHostToDevice, st));
    const int other = c->perm_buf ^ 1;
    k_ploc_leaf_pos<<<grid_for(n, 256), 256, 0, st>>>((int)n, c->child.as<int2>(), c->parent.as<int>(), c->leaf_parent.as<int>(), size,
                                                       c->vals[c->perm_buf].as<uint32_t>(), c->ploc_nn.as<int>(),
                                                       c->vals[other].as<uint32_t>(), c->ploc_tmp.as<int>());
    k_ploc_node_range<<<grid_for(n - 1, 256), 256, 0, st>>>((int)n, c->child.as<int2>(), c->parent.as<int>(), size,
                                                             c->ploc_nn.as<int>(), c->range.as<int2>());
    k_ploc_relabel<<<grid_for(n - 1, 256), 256, 0, st>>>((int)n, c->child.as<int2>(), c->ploc_nn.as<int>());
    CRT_CUDA(cudaMemcpyAsync(c->leaf_parent.p, c->ploc_tmp.p, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    c->perm_buf = other;
    return CRT_OK;
}

// `async` (refit only): enqueue on that stream and return without waiting; the caller orders its readers behind
// c->ev_up1 and reads the duration from (ev_up0, ev_up1) once it has synchronised (crt_render_frame).
int lbvh_build(crt_ctx* c, bool refit_only, cudaStream_t async)
{
    const size_t n = c->n_tri + c->n_sph + c->n_rect;
    if (n >= (1ull << 28)) { set_error("too many primitives (%zu), limit is 2^28", n); return CRT_ERR_INVALID; }
    if (async && !refit_only) { set_error("lbvh_build: only a refit can be queued without waiting"); return CRT_ERR_INVALID; }
    cudaStream_t st = async ? async : c->stream;
    cudaEvent_t ev_begin = async ? c->ev_up0 : c->ev0, ev_end = async ? c->ev_up1 : c->ev1;
    if (refit_only && (!c->accel_valid || c->n_prims != (uint32_t)n)) {
        set_error("crt_accel_refit: no acceleration structure for the current primitive count");
        return CRT_ERR_STATE;
    }
    c->n_prims = (uint32_t)n;
    CRT_CUDA(cudaEventRecord(ev_begin, st));
    if (n == 0) {
        c->accel_valid = true;
        c->root = -1;
        c->tree_depth = 0;
        CRT_CUDA(cudaEventRecord(ev_end, st));
        if (!async) CRT_CUDA(cudaEventSynchronize(ev_end));
        return CRT_OK;
    }
    PrimSrc S;
    S.tri_in = c->tri_in.as<float>(); S.sph_in = c->sph_in.as<float>();
    S.tri_mat = c->tri_mat.as<uint32_t>(); S.tri_id = c->tri_id.as<uint32_t>();
    S.sph_mat = c->sph_mat.as<uint32_t>(); S.sph_id = c->sph_id.as<uint32_t>();
    S.rect_mat = c->rect_mat.as<uint32_t>(); S.rect_id = c->rect_id.as<uint32_t>();
    S.tri_chain = c->tri_has_chain ? c->tri_frame.as<uint32_t>() : nullptr;
    S.sph_chain = c->sph_has_chain ? c->sph_frame.as<uint32_t>() : nullptr;
    S.rect_chain = c->rect_frame.as<uint32_t>();
    S.frames = c->frames.as<DFrame>();
    S.pad_bounds = nullptr;
    S.n_tri = (uint32_t)c->n_tri; S.n_sph = (uint32_t)c->n_sph; S.n_rect = (uint32_t)c->n_rect;
    const uint32_t un = (uint32_t)n;

    if (!refit_only) {
        CRT_CUDA(c->keys[0].reserve(n * 8)); CRT_CUDA(c->keys[1].reserve(n * 8));
        CRT_CUDA(c->vals[0].reserve(n * 4)); CRT_CUDA(c->vals[1].reserve(n * 4));
        CRT_CUDA(c->hist.reserve(radix_sort_hist_bytes()));
        CRT_CUDA(c->bounds.reserve(12 * sizeof(float)));
        CRT_CUDA(c->recs.reserve(n * 64));
        CRT_CUDA(c->leaf_box.reserve(n * 32));
        CRT_CUDA(c->leaf_parent.reserve(n * 4));
        CRT_CUDA(c->depth_out.reserve(sizeof(int)));
        const size_t ni = n > 1 ? n - 1 : 1;
        CRT_CUDA(c->nodes.reserve(ni * 128));
        CRT_CUDA(c->child.reserve(ni * 8));
        CRT_CUDA(c->range.reserve(ni * 8));
        CRT_CUDA(c->parent.reserve(ni * 4));
        CRT_CUDA(c->node_box.reserve(ni * 32));
        CRT_CUDA(c->node_cone.reserve(ni * 16));
        CRT_CUDA(c->flags.reserve(ni * 4));

        k_init_bounds<<<1, 32, 0, st>>>(c->bounds.as<float>());
        const unsigned int gb = min(grid_for(n, 256), (unsigned int)(c->sm_count * 8));
        k_scene_bounds<<<gb, 256, 0, st>>>(S, un, nullptr, c->bounds.as<float>());
        k_scene_bounds<<<gb, 256, 0, st>>>(S, un, c->bounds.as<float>(), c->bounds.as<float>() + 6);
        k_morton<<<grid_for(n, 256), 256, 0, st>>>(S, un, c->bounds.as<float>(), c->keys[0].as<uint64_t>(), c->vals[0].as<uint32_t>());
        uint64_t* keys[2] = {c->keys[0].as<uint64_t>(), c->keys[1].as<uint64_t>()};
        uint32_t* vals[2] = {c->vals[0].as<uint32_t>(), c->vals[1].as<uint32_t>()};
        c->perm_buf = radix_sort_pairs(st, keys, vals, un, 0, 64, c->hist.as<uint32_t>());
        if (n > 1 && c->builder == CRT_BUILDER_LBVH)
            k_hierarchy<<<grid_for(n - 1, 256), 256, 0, st>>>(keys[c->perm_buf], (int)n, c->child.as<int2>(),
                                                              c->range.as<int2>(), c->parent.as<int>(),
                                                              c->leaf_parent.as<int>());
        if (n > 1 && c->builder == CRT_BUILDER_PLOC) {
            const int r = ploc_topology(c, S, un);
            if (r != CRT_OK) return r;
        }
    }
    if (n > 1) CRT_CUDA(cudaMemsetAsync(c->flags.p, 0, (n - 1) * 4, st));
    S.pad_bounds = c->bounds.as<float>();       // the final boxes of chained primitives are padded by the scene's scale
    if (refit_only)
        k_fit<false><<<grid_for(n, 256), 256, 0, st>>>(S, un, c->vals[c->perm_buf].as<uint32_t>(), c->child.as<int2>(),
                                            c->parent.as<int>(), c->leaf_parent.as<int>(), c->flags.as<int>(),
                                            c->recs.as<float4>(), c->leaf_box.as<float4>(),
                                            c->node_box.as<float4>(), c->node_cone.as<float4>());
    else
        k_fit<true><<<grid_for(n, 256), 256, 0, st>>>(S, un, c->vals[c->perm_buf].as<uint32_t>(), c->child.as<int2>(),
                                            c->parent.as<int>(), c->leaf_parent.as<int>(), c->flags.as<int>(),
                                            c->recs.as<float4>(), c->leaf_box.as<float4>(),
                                            c->node_box.as<float4>(), c->node_cone.as<float4>());
    const int leaf_max = c->leaf_max;
    if (n > 1)
        k_emit_nodes<<<grid_for(n - 1, 256), 256, 0, st>>>((int)n, leaf_max, c->child.as<int2>(), c->range.as<int2>(),
                                                           c->leaf_box.as<float4>(), c->node_box.as<float4>(),
                                                           c->node_cone.as<float4>(), c->nodes.as<float4>());
    c->root = (n <= (size_t)leaf_max) ? ~(int)(n - 1) : 0;
    CRT_CUDA(cudaEventRecord(ev_end, st));
    if (async) { CRT_CUDA(cudaGetLastError()); return CRT_OK; }
    if (!refit_only) {
        CRT_CUDA(cudaMemsetAsync(c->depth_out.p, 0, sizeof(int), st));
        if (n > 1)
            k_depth<<<grid_for(n, 256), 256, 0, st>>>((int)n, c->parent.as<int>(), c->leaf_parent.as<int>(),
                                                      c->depth_out.as<int>());
        int depth = 1;
        CRT_CUDA(cudaMemcpyAsync(&depth, c->depth_out.p, sizeof(int), cudaMemcpyDeviceToHost, st));
        CRT_CUDA(cudaStreamSynchronize(st));
        c->tree_depth = n > 1 ? depth : 1;
        // the traversal stacks hold CRT_STACK_ENTRIES references (crt_traverse.cuh); a clustered tree is not bounded
        // by the key length like the radix tree, so a pathologically deep one is replaced by the radix tree
        if (c->builder == CRT_BUILDER_PLOC && 3 * (c->tree_depth - 1) > CRT_STACK_ENTRIES) {
            c->builder = CRT_BUILDER_LBVH;
            const int r = lbvh_build(c, false);
            c->builder = CRT_BUILDER_PLOC;
            return r;
        }
    }
    CRT_CUDA(cudaEventSynchronize(c->ev1));
    CRT_CUDA(cudaGetLastError());
    float ms = 0.f;
    CRT_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    if (refit_only) c->host_counters.refit_ms = ms; else c->host_counters.build_ms = ms;
    c->accel_valid = true;
    return CRT_OK;
}

}  // namespace crt
