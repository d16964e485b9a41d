// crt_device.cuh — per-ray device functions of the render hot path:
// primitive tests, camera, materials, sky. Each follows the reference lines it
// cites (paths relative to /root/reference/CudaTest/src/) operation for
// operation, in the non-fused arithmetic of crt_math.cuh.
#pragma once
#include "crt_math.cuh"

namespace crt {

#ifndef CRT_MISS_ID
#define CRT_MISS_ID 0xFFFFFFFFu
#endif

// ---- HBM layouts -----------------------------------------------------------
// Sorted primitive record, 64 B = 4 x float4 (two 256-bit loads):
//   triangle: r0 = {v0.xyz, n.x}  r1 = {e1.xyz, n.y}  r2 = {e2.xyz, n.z}
//             e1 = v1 - v0, e2 = v2 - v0 (the subtraction triangle.h:69-70
//             performs per test, done once; same float result)
//   sphere:   r0 = {centre.xyz, radius}, r1 = r2 = 0
//   rectangle (shapes/rectangle.h: the unit square of its frame's z = 0 plane): r0 = r1 = r2 = 0
//   all:      r3 = {prim_id, mat_id (15 bits) | CRT_KIND_* flags, chain, 0} as uint bits; chain = (first << 8) | count
//             names the primitive's run of DFrame records (0 = the plain case: no frame, see chain_ray)
// 4-wide node: 96 B of payload at a 128 B stride (one cache line), read with three 256-bit loads:
//   words 5k..5k+4 = child k: {lo.x, lo.y, lo.z, half2(ext.x, ext.y), half2(ext.z, A_k)}; A = (A_0, A_1, A_2) is the
//        normal-cone vector of the NODE's own subtree (0 = none; lbvh.cu:cone_to_halves, crt_traverse.cuh:node_backfacing),
//        the fourth child's spare half is 0
//        hi = lo + ext; ext is a half rounded up at build time until lo + ext >= hi in fp32, so the box only
//        grows (by < 2^-10 of its own extent) and the child references fit the same three loads - a fourth
//        load per node visit was a quarter of the L1 wavefronts of the incoherent bounces
//   words 20..23  = child[0..3] as int bits; child >= 0: node index, child < 0: leaf,
//        ~child = (first sorted primitive << 3) | (count - 1);
//        CRT_SLOT_EMPTY: unused slot (a node has 2..4 children)
//   words 24..31  = padding
#define CRT_SLOT_EMPTY ((int)0x80000000)
#define CRT_KIND_SPHERE 0x80000000u
#define CRT_KIND_RECT 0x40000000u
#define CRT_RECT_FLIP 0x20000000u
#define CRT_MAT_MASK 0x7FFFu

// type: crt_material_type, or CRT_DMAT_CHECKER + it for a Lambertian / DiffuseLight whose texture is a
// CheckerTexture(even = r g b, odd = r2 g2 b2)
#define CRT_DMAT_CHECKER 4u
struct DMat { float r, g, b, param; uint32_t type; float r2, g2, b2; };  // 32 B

// One level of a primitive's chain of enclosing Transforms (hitable/transform.h:7-14): the rows of the matrix
// rotate() builds from the Euler angles (core/vec3.h:200-217), computed on the host, and the position. 64 B.
struct DFrame { float m[9]; float px, py, pz; uint32_t has_rot; uint32_t pad[3]; };

struct AccelView {
    const float4* nodes;
    const float4* recs;
    const DFrame* frames;  // chains of enclosing Transforms (null when the scene has none)
    int root;            // child-style reference of the root
    uint32_t n_prims;
};

struct DCamera {         // core/camera.h:46-53
    v3 llc, horizontal, vertical, origin, x, y, z;
    float lens_radius;
};

struct Hit {
    float t;
    uint32_t k;          // sorted primitive position (traversal) or input index (brute force)
    uint32_t id;         // caller's primitive id, CRT_MISS_ID on miss
    float u, v;
};

// The direction every primitive test sees: the world list's Hitable::hit
// normalises once and each primitive's normalises again
// (hitable/hitable.h:34, hitable/transform.h:50-54).
CRT_HD v3 ray_dir_local(v3 d) { return vunit(vunit(d)); }

// Deviation D3: a triangle hit counts only if its distance t lies where the ray is inside the triangle's own bounding
// box - the box of v0, v0 + e1, v0 + e2 with its planes moved out by 2^-22 of its largest coordinate - within 2^-15
// relative. The reference's Moeller-Trumbore arithmetic loses all accuracy on sliver triangles and then reports, with
// u and v in range, a t that is tens of units off (measured on C4's pole slivers, profiles/r02_g_c4_walk.txt): the
// list scan takes such hits, any bounding-volume hierarchy - the reference's own BVHNode included, whose leaves test
// their box first (hitable/bvh.h:165) - cannot reach them, and which of two traversals stumbles on one would decide
// the pixel. With D3 the answer is a function of the ray and the triangle alone, and an accepted hit lies inside the
// triangle's padded box, hence inside every box above it (lbvh.cu:prim_box widens by twice as much, and the
// traversals skip a box only beyond 2^-14 of the closest hit, crt_traverse.cuh:CRT_T_SLACK).
//
// d3_interval is the definition, in plain rounded float operations that every side (kernels, host emulation, both
// oracles) evaluates bit for bit. d3_agrees is a shortcut only this header takes: when the point along the ray,
// o + t d, agrees with the point the barycentrics give, v0 + u e1 + v e2 (which lies in the box by construction), to
// within HALF of D3's tolerances, d3_interval is certain to accept, and its three divisions are skipped. Almost every
// real hit takes the shortcut; grazing hits, whose u and v are themselves inaccurate, fall through to the definition.
CRT_HD void d3_axis(float o, float d, float a, float e1, float e2, float& t_near, float& t_far)
{
    const float b = x::add(a, e1), c = x::add(a, e2);
    float lo = fminf(a, fminf(b, c)), hi = fmaxf(a, fmaxf(b, c));
    const float pad = x::add(x::mul(fmaxf(fabsf(lo), fabsf(hi)), 2.384185791015625e-7f), 1e-30f);   // 2^-22
    lo = x::sub(lo, pad);
    hi = x::add(hi, pad);
    const float inv = fabsf(d) > 1e-18f ? x::div(1.0f, d) : copysignf(1e18f, d);
    const float t0 = x::mul(x::sub(lo, o), inv), t1 = x::mul(x::sub(hi, o), inv);
    t_near = fmaxf(t_near, fminf(t0, t1));
    t_far = fminf(t_far, fmaxf(t0, t1));
}
#define CRT_D3_EPS 3.0517578125e-5f   // 2^-15
#ifndef CRT_D3
#define CRT_D3 1                      // 0: development switch for A/B timing only (parity with the oracles is lost)
#endif

CRT_HD bool d3_interval(v3 o, v3 d, v3 v0, v3 e1, v3 e2, float t)
{
    float t_near = -INFINITY, t_far = INFINITY;
    d3_axis(o.x, d.x, v0.x, e1.x, e2.x, t_near, t_far);
    d3_axis(o.y, d.y, v0.y, e1.y, e2.y, t_near, t_far);
    d3_axis(o.z, d.z, v0.z, e1.z, e2.z, t_near, t_far);
    return t >= x::sub(t_near, x::mul(fabsf(t_near), CRT_D3_EPS)) && t <= x::add(t_far, x::mul(fabsf(t_far), CRT_D3_EPS));
}

CRT_HD bool d3_agrees_axis(float o, float d, float a, float e1, float e2, float t, float t_eps, float u, float v)
{
    const float pp = fmaf(v, e2, fmaf(u, e1, a));
    const float p = fmaf(t, d, o);
    return fabsf(p - pp) <= fmaf(fabsf(pp), 1.1920928955078125e-7f, t_eps * fabsf(d));   // 2^-23 of the coordinate + 2^-16 of t |d|
}
CRT_HD bool d3_agrees(v3 o, v3 d, v3 v0, v3 e1, v3 e2, float t, float u, float v)
{
    const float t_eps = t * 1.52587890625e-5f;
    return d3_agrees_axis(o.x, d.x, v0.x, e1.x, e2.x, t, t_eps, u, v) && d3_agrees_axis(o.y, d.y, v0.y, e1.y, e2.y, t, t_eps, u, v) &&
           d3_agrees_axis(o.z, d.z, v0.z, e1.z, e2.z, t, t_eps, u, v);
}

// shapes/triangle.h:57-100 plus deviation D1 (accept only t_min < t < t_max) and deviation D3 (above).
// v0/e1/e2/n as stored in the record. A candidate with t equal to the current best is still returned (t <= t_best);
// the caller breaks the tie by primitive id.
CRT_HD bool triangle_candidate(v3 o, v3 d, v3 v0, v3 e1, v3 e2, v3 n, float t_min, float t_best,
                               float& t_out, float& u_out, float& v_out)
{
    const float EPSILON = 0.000001f;                       // triangle.h:9,22
    if (vdot(d, n) < 0.0f) return false;                   // triangle.h:61
    v3 h = vcross(d, e2);
    float a = vdot(e1, h);
    if (a > -EPSILON && a < EPSILON) return false;         // triangle.h:77
    float f = x::rcp(a);                                   // triangle.h:80
    v3 s = vsub(o, v0);
    float u = x::mul(f, vdot(s, h));
    if (u < 0.0f || u > 1.0f) return false;                // triangle.h:84
    v3 q = vcross(s, e1);
    float v = x::mul(f, vdot(d, q));
    if (v < 0.0f || x::add(u, v) > 1.0f) return false;     // triangle.h:89
    float t = x::mul(f, vdot(e2, q));                      // triangle.h:92
    if (!(t > t_min && t <= t_best)) return false;         // D1
    t_out = t; u_out = u; v_out = v;
    return true;
}

CRT_HD bool triangle_test(v3 o, v3 d, v3 v0, v3 e1, v3 e2, v3 n, float t_min, float t_best,
                          float& t_out, float& u_out, float& v_out)
{
    if (!triangle_candidate(o, d, v0, e1, e2, n, t_min, t_best, t_out, u_out, v_out)) return false;
#if CRT_D3
    return d3_agrees(o, d, v0, e1, e2, t_out, u_out, v_out) || d3_interval(o, d, v0, e1, e2, t_out);   // D3
#else
    return true;
#endif
}

// shapes/sphere.h:27-55; oc = origin in the sphere's frame (origin - centre,
// hitable/transform.h:38-41).
CRT_HD bool sphere_test(v3 oc, v3 d, float radius, float t_min, float t_best, float& t_out)
{
    float a = vdot(d, d);
    float b = vdot(oc, d);
    float c = x::sub(vdot(oc, oc), x::mul(radius, radius));
    float disc = x::sub(x::mul(b, b), x::mul(a, c));
    if (disc > 0.0f) {
        float sq = x::sqrt(disc);
        float tmp = x::div(x::sub(-b, sq), a);
        if (tmp <= t_best && tmp > t_min) { t_out = tmp; return true; }
        tmp = x::div(x::add(-b, sq), a);
        if (tmp <= t_best && tmp > t_min) { t_out = tmp; return true; }
    }
    return false;
}

// shapes/rectangle.h:21-44 in the rectangle's own frame: the unit square of the plane z = 0, answered only by rays
// that travel AGAINST its normal (the opposite of Triangle's rule), bounds inclusive. The list keeps a hit only if
// t < closest (hitable/hitable_list.h:84), hence t <= t_best here and the caller's tie rule; a NaN t (direction and
// origin both in the plane) passes every reference comparison but not that last one.
CRT_HD bool rect_test(v3 o, v3 d, bool flip, float t_min, float t_best, float& t_out, float& u_out, float& v_out)
{
    const v3 normal = V(0.0f, 0.0f, flip ? -1.0f : 1.0f);
    if (vdot(d, normal) > 0.0f) return false;              // rectangle.h:24
    const float t = x::div(-o.z, d.z);                     // rectangle.h:25
    if (t < t_min || t > t_best) return false;             // rectangle.h:26
    const float px = x::add(o.x, x::mul(t, d.x));          // rectangle.h:33-34
    const float py = x::add(o.y, x::mul(t, d.y));
    if (px < -0.5f || px > 0.5f || py < -0.5f || py > 0.5f) return false;
    if (!(t == t)) return false;
    t_out = t;
    u_out = x::div(x::sub(px, -0.5f), x::sub(0.5f, -0.5f));   // rectangle.h:37-38
    v_out = x::div(x::sub(py, -0.5f), x::sub(0.5f, -0.5f));
    return true;
}

// ---- chains of enclosing Transforms ---------------------------------------------
// rotate(v, rotation) (core/vec3.h:200-217) with the matrix rows precomputed on the host (they need libm's
// sin/cos): component k = (v * row_k).sum() = v.x m[3k] + v.y m[3k+1] + v.z m[3k+2], summed left to right.
CRT_HD v3 vrot(const float* m, v3 v)
{
    return V(x::add(x::add(x::mul(v.x, m[0]), x::mul(v.y, m[1])), x::mul(v.z, m[2])),
             x::add(x::add(x::mul(v.x, m[3]), x::mul(v.y, m[4])), x::mul(v.z, m[5])),
             x::add(x::add(x::mul(v.x, m[6]), x::mul(v.y, m[7])), x::mul(v.z, m[8])));
}

// The ray as a primitive's collision_detection receives it. Every Hitable on the way down - the world list, nested
// lists, the primitive itself - maps the ray in its hit() by TransformRay = Translate(Rotate(Scale(r)))
// (hitable/hitable.h:34, hitable/transform.h:11-14,38-54): direction re-normalised (scale is 1), origin and
// direction rotated, position subtracted from the origin. A chain lists those levels outermost first. The plain
// case (chain 0) is two levels without rotation: d = unit(unit(d)) = ray_dir_local, and the one translation a
// sphere may carry is its centre.
CRT_HD void chain_ray(const DFrame* __restrict__ frames, uint32_t chain, v3 o, v3 d_raw, v3& ol, v3& dl)
{
    const DFrame* f = frames + (chain >> 8);
    const uint32_t n = chain & 255u;
    for (uint32_t i = 0; i < n; i++) {
        d_raw = vunit(d_raw);                                            // transform.h:50-54 (ScaleRay, scale 1)
        if (f[i].has_rot) { o = vrot(f[i].m, o); d_raw = vrot(f[i].m, d_raw); }   // transform.h:43-48
        o = vsub(o, V(f[i].px, f[i].py, f[i].pz));                       // transform.h:38-41
    }
    ol = o;
    dl = d_raw;
}

// The normal on its way back up: every level's hit() rotates it by +rotation (hitable/hitable.h:36; not the inverse,
// SURVEY 9 Q6), innermost first.
CRT_HD v3 chain_normal(const DFrame* __restrict__ frames, uint32_t chain, v3 n)
{
    const DFrame* f = frames + (chain >> 8);
    for (int i = (int)(chain & 255u) - 1; i >= 0; i--)
        if (f[i].has_rot) n = vrot(f[i].m, n);
    return n;
}

// A primitive under a chain against the world ray (o, d_raw = the direction as shade() holds it): the test runs in
// the primitive's own frame, as the reference's does. info = record word r3.y.
CRT_HD bool framed_test(const DFrame* __restrict__ frames, uint32_t chain, uint32_t info, float4 r0, float4 r1, float4 r2,
                        v3 o, v3 d_raw, float t_min, float t_best, float& t, float& u, float& v)
{
    v3 ol, dl;
    chain_ray(frames, chain, o, d_raw, ol, dl);
    u = v = 0.0f;
    if (info & CRT_KIND_SPHERE) return sphere_test(vsub(ol, V(r0.x, r0.y, r0.z)), dl, r0.w, t_min, t_best, t);
    if (info & CRT_KIND_RECT) return rect_test(ol, dl, (info & CRT_RECT_FLIP) != 0u, t_min, t_best, t, u, v);
    return triangle_test(ol, dl, V(r0.x, r0.y, r0.z), V(r1.x, r1.y, r1.z), V(r2.x, r2.y, r2.z), V(r0.w, r1.w, r2.w),
                         t_min, t_best, t, u, v);
}

// ---- camera ----------------------------------------------------------------
struct Rng { uint64_t key; uint32_t ctr; };
CRT_HD float rnd(Rng& r) { return crt_rng_uniform(r.key, r.ctr++); }

// core/camera.h:6-13. C++ leaves the evaluation order of the two constructor
// arguments unspecified; the reference as compiled for oracle/_ref draws right
// to left, so the first draw is y.
CRT_HD v3 random_in_unit_disk(Rng& r)
{
    v3 p;
    do {
        float b = rnd(r), a = rnd(r);
        p = vsub(vscale(2.0f, V(a, b, 0.0f)), V(1.0f, 1.0f, 0.0f));
    } while (vdot(p, p) >= 1.0f);
    return p;
}

// core/camera.h:40-44
CRT_HD void camera_get_ray(const DCamera& c, float s, float t, Rng& r, v3& o, v3& d)
{
    v3 rd = vscale(c.lens_radius, random_in_unit_disk(r));
    v3 offset = vadd(vscale(rd.x, c.x), vscale(rd.y, c.y));
    o = vadd(c.origin, offset);
    d = vsub(vsub(vadd(vadd(c.llc, vscale(s, c.horizontal)), vscale(t, c.vertical)), c.origin), offset);
}

// ---- materials -------------------------------------------------------------
CRT_HD v3 reflect(v3 v, v3 n) { return vsub(v, vscale(x::mul(2.0f, vdot(v, n)), n)); }   // material.h:6-8

// material.h:11-19; right-to-left draws, see random_in_unit_disk.
CRT_HD v3 random_in_unit_sphere(Rng& r)
{
    v3 p;
    do {
        float c = rnd(r), b = rnd(r), a = rnd(r);
        p = vsub(vscale(2.0f, V(a, b, c)), V(1.0f, 1.0f, 1.0f));
    } while (vlen2(p) >= 1.0f);
    return p;
}

// material.h:22-26. powf(x, 5) is evaluated as a double product rounded once,
// which agrees with glibc's (double-precision-internal) powf.
CRT_HD float schlick(float cosine, float ref_idx)
{
    float r0 = x::div(x::sub(1.0f, ref_idx), x::add(1.0f, ref_idx));
    r0 = x::mul(r0, r0);
    double b = (double)x::sub(1.0f, cosine);
    float p5 = (float)(b * b * b * b * b);
    return x::add(r0, x::mul(x::sub(1.0f, r0), p5));
}

// material.h:29-39
CRT_HD bool refract(v3 v, v3 n, float ni_over_nt, v3& refracted)
{
    v3 uv = vunit(v);
    float dt = vdot(uv, n);
    float disc = x::sub(1.0f, x::mul(x::mul(ni_over_nt, ni_over_nt), x::sub(1.0f, x::mul(dt, dt))));
    if (disc > 0.0f) {
        refracted = vsub(vscale(ni_over_nt, vsub(uv, vscale(dt, n))), vscale(x::sqrt(disc), n));
        return true;
    }
    return false;
}

// Material::scatter, material.h:57-155. in_d: direction of the incoming ray as
// shade() holds it (not normalised). Returns the bool of scatter(); att and
// out_d are written as the reference writes them.
// CheckerTexture::value, material/texture.h:30-38: which of the two colours the hit point selects. Only the sign
// of the product is used, so the last bits of sinf do not matter.
CRT_HD bool checker_odd(const DMat& m, v3 p)
{
    if (m.type < CRT_DMAT_CHECKER) return false;
    const float sines = sinf(x::mul(10.0f, p.x)) * sinf(x::mul(10.0f, p.y)) * sinf(x::mul(10.0f, p.z));
    return sines < 0.0f;
}
CRT_HD v3 texture_value(const DMat& m, bool odd) { return odd ? V(m.r2, m.g2, m.b2) : V(m.r, m.g, m.b); }

CRT_HD bool scatter(const DMat& m, v3 in_d, v3 p, v3 n, Rng& r, v3& att, v3& out_d, bool odd = false)
{
    if (m.type == 0u || m.type == CRT_DMAT_CHECKER) {     // Lambertian, material.h:61-71
        v3 target = vadd(vadd(p, n), random_in_unit_sphere(r));
        out_d = vsub(target, p);
        att = texture_value(m, odd);
        return true;
    }
    if (m.type == 1u) {                                    // Metal, material.h:83-92 (fuzz clamp :80)
        float fuzz = m.param < 1.0f ? m.param : 1.0f;
        v3 reflected = reflect(vunit(in_d), n);
        out_d = vadd(reflected, vscale(fuzz, random_in_unit_sphere(r)));
        att = V(m.r, m.g, m.b);
        return vdot(out_d, n) > 0.0f;
    }
    if (m.type == 2u) {                                    // Dielectric, material.h:103-141
        float ref_idx = m.param;
        v3 outward, refracted = V(0.0f, 0.0f, 0.0f);
        v3 reflected = reflect(in_d, n);
        float ni_over_nt, reflect_prob, cosine;
        att = V(1.0f, 1.0f, 1.0f);
        float dn = vdot(in_d, n);
        if (dn > 0.0f) {
            outward = vneg(n);
            ni_over_nt = ref_idx;
            cosine = x::div(dn, vlen(in_d));
            cosine = x::sqrt(x::sub(1.0f, x::mul(x::mul(ref_idx, ref_idx), x::sub(1.0f, x::mul(cosine, cosine)))));
        } else {
            outward = n;
            ni_over_nt = x::div(1.0f, ref_idx);
            cosine = x::div(-dn, vlen(in_d));
        }
        if (refract(in_d, outward, ni_over_nt, refracted)) reflect_prob = schlick(cosine, ref_idx);
        else reflect_prob = 1.0f;
        if (rnd(r) < reflect_prob) out_d = reflected;
        else out_d = refracted;
        return true;
    }
    att = V(0.0f, 0.0f, 0.0f);                             // DiffuseLight, material.h:149-155
    out_d = V(0.0f, 0.0f, 0.0f);
    return false;
}

CRT_HD v3 emitted(const DMat& m, bool odd = false)         // material.h:49-53,156-158
{
    return (m.type == 3u || m.type == CRT_DMAT_CHECKER + 3u) ? texture_value(m, odd) : V(0.0f, 0.0f, 0.0f);
}

// core/render.h:41-46
CRT_HD v3 background_sky(v3 d)
{
    v3 n = vunit(d);
    float t = x::mul(0.5f, x::add(n.y, 1.0f));
    return vlerp(t, V(1.0f, 1.0f, 1.0f), V(0.5f, 0.7f, 1.0f));
}

}  // namespace crt
