/* crt_oracle.c — TEST INFRASTRUCTURE: plain-C restatement of the reference's
 * render hot path. Parity is pinned: tests/test_oracle.py
 * checks every function here against the unmodified reference headers
 * compiled in oracle/_ref (when present) and against the golden vectors under
 * tests/golden/ that were generated from them (tests/golden/make_golden.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library. The product
 * (cudaraytracer_b200/) never does; it has no CPU path.
 *
 * All paths cited are relative to /root/reference/CudaTest/src/.
 * Compile with -ffp-contract=off: the reference arithmetic is unfused.
 *
 * Deviation D1, applied when the scene was created with d1=1:
 * a triangle hit counts only if t_min < t < t_max. d1=0 keeps the raw
 * behaviour of shapes/triangle.h:92-99 (no range test).
 */
#include "crt_oracle.h"
#include "crt_rng.h"

#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ------------------------------------------------------------------ vec3 */
/* core/vec3.h:7-240 */
typedef struct { float x, y, z; } v3;

static inline v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }          /* vec3.h:73 */
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }          /* vec3.h:77 */
static inline v3 vmul(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }          /* vec3.h:89 */
static inline v3 vscale(float t, v3 a) { return V(t * a.x, t * a.y, t * a.z); }           /* vec3.h:93-99 */
static inline v3 vdivv(v3 a, v3 b) { return V(a.x / b.x, a.y / b.y, a.z / b.z); }         /* vec3.h:101 */
static inline v3 vdivf(v3 a, float t) { return V(a.x / t, a.y / t, a.z / t); }            /* vec3.h:105 */
static inline v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }                               /* vec3.h:24 */
static inline float vdot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }        /* vec3.h:109 */
static inline v3 vcross(v3 a, v3 b)                                                        /* vec3.h:113 */
{
    return V(a.y * b.z - a.z * b.y, -(a.x * b.z - a.z * b.x), a.x * b.y - a.y * b.x);
}
static inline float vlen(v3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }       /* vec3.h:38 */
static inline float vlen2(v3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }             /* vec3.h:42 */
static inline v3 vunit(v3 a) { return vdivf(a, vlen(a)); }                                /* vec3.h:191 */
static inline v3 vlerp(float t, v3 from, v3 to) { return vadd(from, vscale(t, vsub(to, from))); } /* vec3.h:196 */
static inline float vsum(v3 a) { return a.x + a.y + a.z; }                                /* vec3.h:68 */

/* vec3.h:200-217. Euler rotation, degrees; Z angle negated. */
static v3 vrotate(v3 p, v3 rot)
{
    float rx = (float)((M_PI / 180.) * rot.x);
    float sx = sinf(rx), cx = cosf(rx);
    float ry = (float)((M_PI / 180.) * rot.y);
    float sy = sinf(ry), cy = cosf(ry);
    float rz = (float)(-(M_PI / 180.) * rot.z);
    float sz = sinf(rz), cz = cosf(rz);
    v3 r0 = V(cy * cz, -cy * sz, sy);
    v3 r1 = V(sx * sy * cz + cx * sz, -sx * sy * sz + cx * cz, -sx * cy);
    v3 r2 = V(-cx * sy * cz + sx * sz, cx * sy * sz + sx * cz, cx * cy);
    return V(vsum(vmul(p, r0)), vsum(vmul(p, r1)), vsum(vmul(p, r2)));
}

/* vec3.h:119-131 */
static inline float clip1(float f, int lo, int hi)
{
    if (f > hi) return (float)hi;
    else if (f < lo) return (float)lo;
    return f;
}

/* ------------------------------------------------------------------ types */
typedef struct { v3 o, d; } ray_t;                       /* core/ray.h:7-23 (time unused) */
typedef struct { v3 position, rotation, scale; } xform_t; /* hitable/transform.h:33-35 */
typedef struct { float t; v3 p, normal; uint32_t mat; uint32_t id; float u, v; } hitrec_t; /* hitable/hitable.h:11-18 */

typedef struct {
    int kind;          /* 0 triangle, 1 sphere, 2 rectangle (shapes/rectangle.h) */
    int flip;          /* rectangle: flipNormal */
    int n_outer;       /* enclosing HitableLists between the world list and this primitive */
    xform_t* outer;    /* their Transforms, outermost first */
    v3 v0, v1, v2, n;  /* triangle: shapes/triangle.h:49-50 */
    float radius;      /* sphere: shapes/sphere.h:22 */
    xform_t xf;
    uint32_t mat, id;
} prim_t;

typedef struct { uint32_t type; v3 albedo; float param; int checker; v3 odd; } mat_t;

struct oracle_scene {
    prim_t* prims; size_t n_prims, cap_prims;
    mat_t* mats; size_t n_mats, cap_mats;
    int d1;
    int finalized;
    xform_t world_xf;
};

typedef struct { uint64_t key; uint32_t ctr; const float* inject; int inject_n, inject_pos; } rng_t;

static inline float rnd(rng_t* s)
{
    if (s->inject) { float v = s->inject[s->inject_pos % s->inject_n]; s->inject_pos++; return v; }
    return crt_rng_uniform(s->key, s->ctr++);
}

static const xform_t XF_IDENTITY = {{0, 0, 0}, {0, 0, 0}, {1, 1, 1}};

/* hitable/transform.h:11-14,38-54: Translate(Rotate(Scale(r))) */
static ray_t transform_ray(const xform_t* xf, ray_t r)
{
    v3 dir = vdivv(r.d, xf->scale);           /* ScaleRay: only the direction, then re-normalised */
    r.d = vunit(dir);
    r.o = vrotate(r.o, xf->rotation);         /* RotateRay */
    r.d = vrotate(r.d, xf->rotation);
    r.o = vsub(r.o, xf->position);            /* TranslateRay */
    return r;
}

/* Deviation D3 (see cudaraytracer_b200/csrc/crt_device.cuh:triangle_test): one axis of the interval in which the ray
 * is inside the box of v0, v0 + e1, v0 + e2; plain rounded float operations in this order. */
static void d3_axis(float o, float d, float a, float e1, float e2, float* t_near, float* t_far)
{
    float b = a + e1, c = a + e2;
    float lo = fminf(a, fminf(b, c)), hi = fmaxf(a, fmaxf(b, c));
    float pad = fmaxf(fabsf(lo), fabsf(hi)) * 2.384185791015625e-7f + 1e-30f;   /* 2^-22 of the largest coordinate */
    lo = lo - pad;
    hi = hi + pad;
    float inv = fabsf(d) > 1e-18f ? 1.0f / d : copysignf(1e18f, d);
    float t0 = (lo - o) * inv, t1 = (hi - o) * inv;
    *t_near = fmaxf(*t_near, fminf(t0, t1));
    *t_far = fminf(*t_far, fmaxf(t0, t1));
}

/* shapes/triangle.h:57-100 */
static int triangle_test(const prim_t* p, ray_t r, float t_min, float t_max, int d1, hitrec_t* rec)
{
    const float EPSILON = 0.000001;           /* triangle.h:9,22 */
    if (vdot(r.d, p->n) < 0) return 0;
    v3 e1 = vsub(p->v1, p->v0);
    v3 e2 = vsub(p->v2, p->v0);
    v3 h = vcross(r.d, e2);
    float a = vdot(e1, h);
    if (a > -EPSILON && a < EPSILON) return 0;
    float f = (float)(1.0 / a);
    v3 s = vsub(r.o, p->v0);
    float u = f * vdot(s, h);
    if (u < 0.0 || u > 1.0) return 0;
    v3 q = vcross(s, e1);
    float v = f * vdot(r.d, q);
    if (v < 0.0 || u + v > 1.0) return 0;
    float t = f * vdot(e2, q);
    if (d1 && !(t > t_min && t < t_max)) return 0;   /* deviation D1 */
    if (d1) {                                         /* deviation D3, applied together with D1 */
        const float E = 3.0517578125e-5f;             /* 2^-15 */
        float t_near = -INFINITY, t_far = INFINITY, m;
        d3_axis(r.o.x, r.d.x, p->v0.x, e1.x, e2.x, &t_near, &t_far);
        d3_axis(r.o.y, r.d.y, p->v0.y, e1.y, e2.y, &t_near, &t_far);
        d3_axis(r.o.z, r.d.z, p->v0.z, e1.z, e2.z, &t_near, &t_far);
        m = fabsf(t_near) * E;
        if (!(t >= t_near - m)) return 0;
        m = fabsf(t_far) * E;
        if (!(t <= t_far + m)) return 0;
    }
    rec->t = t;
    rec->p = vadd(r.o, vscale(t, r.d));       /* ray.h:18 */
    rec->normal = p->n;
    rec->mat = p->mat;
    rec->id = p->id;
    rec->u = u;
    rec->v = v;
    return 1;
}

/* shapes/sphere.h:27-55; the ray is already in the sphere's local frame */
static int sphere_test(const prim_t* p, ray_t r, float t_min, float t_max, hitrec_t* rec)
{
    v3 oc = r.o;
    float a = vdot(r.d, r.d);
    float b = vdot(oc, r.d);
    float c = vdot(oc, oc) - p->radius * p->radius;
    float disc = b * b - a * c;
    if (disc > 0) {
        float tmp = (-b - sqrtf(disc)) / a;
        if (tmp < t_max && tmp > t_min) goto found;
        tmp = (-b + sqrtf(disc)) / a;
        if (tmp < t_max && tmp > t_min) goto found;
        return 0;
    found:
        rec->t = tmp;
        rec->p = vadd(r.o, vscale(tmp, r.d));
        rec->normal = vdivf(rec->p, p->radius);
        rec->mat = p->mat;
        rec->id = p->id;
        rec->u = rec->v = 0;
        return 1;
    }
    return 0;
}

/* shapes/rectangle.h:21-44 */
static int rect_test(const prim_t* p, ray_t r, float t_min, float t_max, hitrec_t* rec)
{
    v3 normal = p->flip ? V(0, 0, -1) : V(0, 0, 1);
    if (vdot(r.d, normal) > 0) return 0;
    float t = -r.o.z / r.d.z;
    if (t < t_min || t > t_max) return 0;
    float x0 = -0.5f, x1 = 0.5f, y0 = -0.5f, y1 = 0.5f;
    float x = r.o.x + t * r.d.x;
    float y = r.o.y + t * r.d.y;
    if (x < x0 || x > x1 || y < y0 || y > y1) return 0;
    rec->u = (x - x0) / (x1 - x0);
    rec->v = (y - y0) / (y1 - y0);
    rec->t = t;
    rec->mat = p->mat;
    rec->id = p->id;
    rec->p = vadd(r.o, vscale(t, r.d));
    rec->normal = normal;
    return 1;
}

/* hitable/hitable.h:29-38 for one primitive, below the HitableLists that enclose it (each of them a Hitable whose
 * hit() transforms the ray on the way down and rotates the normal on the way up; a list of one entry returns its
 * entry's hit when t < t_max, hitable_list.h:84) */
static int prim_hit(const prim_t* p, ray_t r, float t_min, float t_max, int d1, hitrec_t* rec)
{
    for (int l = 0; l < p->n_outer; l++) r = transform_ray(&p->outer[l], r);
    ray_t lr = transform_ray(&p->xf, r);
    int flag = p->kind == 0 ? triangle_test(p, lr, t_min, t_max, d1, rec)
             : p->kind == 1 ? sphere_test(p, lr, t_min, t_max, rec) : rect_test(p, lr, t_min, t_max, rec);
    if (flag) rec->normal = vrotate(rec->normal, p->xf.rotation);
    if (flag && p->n_outer > 0 && !(rec->t < t_max)) flag = 0;
    if (flag) for (int l = p->n_outer - 1; l >= 0; l--) rec->normal = vrotate(rec->normal, p->outer[l].rotation);
    return flag;
}

/* (*world)->hit(): hitable.h:29-38 on the list + hitable_list.h:76-91 */
static int world_hit(const struct oracle_scene* s, ray_t r, float t_min, float t_max, hitrec_t* rec)
{
    ray_t wr = transform_ray(&s->world_xf, r);
    hitrec_t tmp;
    int hit_anything = 0;
    double closest = t_max;
    for (size_t i = 0; i < s->n_prims; i++) {
        if (prim_hit(&s->prims[i], wr, t_min, (float)closest, s->d1, &tmp) && tmp.t < closest) {
            hit_anything = 1;
            closest = tmp.t;
            *rec = tmp;
        }
    }
    if (hit_anything) rec->normal = vrotate(rec->normal, s->world_xf.rotation);
    return hit_anything;
}

/* ------------------------------------------------------------- materials */
static v3 reflect(v3 v, v3 n) { return vsub(v, vscale(2.0f * vdot(v, n), n)); }   /* material.h:6-8 */

static v3 random_in_unit_sphere(rng_t* s)                                          /* material.h:11-19 */
{
    v3 p;
    do {
        /* C++ leaves the evaluation order of the three constructor arguments
         * unspecified; the reference as compiled for oracle/_ref (g++) draws
         * right to left, so the first draw lands in z. The kernels follow. */
        float c = rnd(s), b = rnd(s), a = rnd(s);
        p = vsub(vscale(2.0f, V(a, b, c)), V(1, 1, 1));
    } while (vlen2(p) >= 1.0f);
    return p;
}

static float schlick(float cosine, float ref_idx)                                  /* material.h:22-26 */
{
    float r0 = (1.0f - ref_idx) / (1.0f + ref_idx);
    r0 = r0 * r0;
    return r0 + (1.0f - r0) * powf(1.0f - cosine, 5.0f);
}

static int refract(v3 v, v3 n, float ni_over_nt, v3* refracted)                    /* material.h:29-39 */
{
    v3 uv = vunit(v);
    float dt = vdot(uv, n);
    float disc = 1.0f - ni_over_nt * ni_over_nt * (1 - dt * dt);
    if (disc > 0) {
        *refracted = vsub(vscale(ni_over_nt, vsub(uv, vscale(dt, n))), vscale(sqrtf(disc), n));
        return 1;
    }
    return 0;
}

/* material.h:57-161. Returns scatter()'s bool; att/out always written the way
 * the reference writes them. */
/* ConstantTexture::value / CheckerTexture::value, material/texture.h:17-19,30-38 */
static v3 texture_value(const mat_t* m, v3 p)
{
    if (m->checker) {
        float sines = sinf(10 * p.x) * sinf(10 * p.y) * sinf(10 * p.z);
        if (sines < 0) return m->odd;
    }
    return m->albedo;
}

static int scatter(const mat_t* m, ray_t in, const hitrec_t* rec, v3* att, ray_t* out, rng_t* s)
{
    switch (m->type) {
    case 0: { /* Lambertian, material.h:61-71 */
        v3 target = vadd(vadd(rec->p, rec->normal), random_in_unit_sphere(s));
        out->o = rec->p;
        out->d = vsub(target, rec->p);
        *att = texture_value(m, rec->p);                                            /* material.h:69 */
        return 1;
    }
    case 1: { /* Metal, material.h:83-92; fuzz clamped at construction (material.h:80) */
        float fuzz = m->param < 1 ? m->param : 1;
        v3 reflected = reflect(vunit(in.d), rec->normal);
        out->o = rec->p;
        out->d = vadd(reflected, vscale(fuzz, random_in_unit_sphere(s)));
        *att = m->albedo;
        return vdot(out->d, rec->normal) > 0.0f;
    }
    case 2: { /* Dielectric, material.h:103-141 */
        float ref_idx = m->param;
        v3 outward, refracted = V(0, 0, 0);
        v3 reflected = reflect(in.d, rec->normal);
        float ni_over_nt, reflect_prob, cosine;
        *att = V(1.0f, 1.0f, 1.0f);
        if (vdot(in.d, rec->normal) > 0.0f) {
            outward = vneg(rec->normal);
            ni_over_nt = ref_idx;
            cosine = vdot(in.d, rec->normal) / vlen(in.d);
            cosine = sqrtf(1 - ref_idx * ref_idx * (1 - cosine * cosine));
        } else {
            outward = rec->normal;
            ni_over_nt = 1.0f / ref_idx;
            cosine = -vdot(in.d, rec->normal) / vlen(in.d);
        }
        if (refract(in.d, outward, ni_over_nt, &refracted)) reflect_prob = schlick(cosine, ref_idx);
        else reflect_prob = 1.0f;
        out->o = rec->p;
        if (rnd(s) < reflect_prob) out->d = reflected;
        else out->d = refracted;
        return 1;
    }
    default: /* DiffuseLight, material.h:149-155 */
        return 0;
    }
}

static v3 emitted(const mat_t* m, v3 p)                                            /* material.h:49-53,156-158 */
{
    return m->type == 3 ? texture_value(m, p) : V(0, 0, 0);
}

/* ---------------------------------------------------------------- camera */
typedef struct { v3 llc, horizontal, vertical, origin, x, y, z; float lens_radius; } camera_t;

static camera_t camera_from22(const float* c)
{
    camera_t k;
    k.llc = V(c[0], c[1], c[2]); k.horizontal = V(c[3], c[4], c[5]); k.vertical = V(c[6], c[7], c[8]);
    k.origin = V(c[9], c[10], c[11]); k.x = V(c[12], c[13], c[14]); k.y = V(c[15], c[16], c[17]);
    k.z = V(c[18], c[19], c[20]); k.lens_radius = c[21];
    return k;
}

static v3 random_in_unit_disk(rng_t* s)                                            /* camera.h:6-13 */
{
    v3 p;
    do {
        float b = rnd(s), a = rnd(s);   /* right-to-left, see random_in_unit_sphere */
        p = vsub(vscale(2.0f, V(a, b, 0)), V(1, 1, 0));
    } while (vdot(p, p) >= 1.0f);
    return p;
}

static ray_t get_ray(const camera_t* c, float s, float t, rng_t* st)               /* camera.h:40-44 */
{
    v3 rd = vscale(c->lens_radius, random_in_unit_disk(st));
    v3 offset = vadd(vscale(rd.x, c->x), vscale(rd.y, c->y));
    ray_t r;
    r.o = vadd(c->origin, offset);
    r.d = vsub(vsub(vadd(vadd(c->llc, vscale(s, c->horizontal)), vscale(t, c->vertical)), c->origin), offset);
    return r;
}

/* --------------------------------------------------------------- shading */
static v3 sky(v3 d)                                                                /* render.h:41-46 */
{
    v3 n = vunit(d);
    float t = 0.5f * (n.y + 1.0f);
    return vlerp(t, V(1, 1, 1), V(0.5f, 0.7f, 1.0f));
}

static v3 shade_path(const struct oracle_scene* s, ray_t r, int depth, rng_t* st, uint64_t* rays)  /* render.h:48-67 */
{
    hitrec_t rec;
    (*rays)++;
    if (world_hit(s, r, 0.001f, FLT_MAX, &rec)) {
        ray_t next;
        v3 att;
        const mat_t* m = &s->mats[rec.mat];
        v3 e = emitted(m, rec.p);
        if (depth > 0 && scatter(m, r, &rec, &att, &next, st))
            return vadd(e, vmul(att, shade_path(s, next, depth - 1, st, rays)));
        return vadd(e, V(0.1f, 0.1f, 0.1f));
    }
    return sky(r.d);
}

static v3 shade_lambert(const struct oracle_scene* s, ray_t r, rng_t* st, uint64_t* rays)          /* render.h:70-87 */
{
    hitrec_t rec;
    (*rays)++;
    if (world_hit(s, r, 0.001f, FLT_MAX, &rec)) {
        ray_t next;
        v3 att = V(0, 0, 0);
        const mat_t* m = &s->mats[rec.mat];
        v3 e = emitted(m, rec.p);
        scatter(m, r, &rec, &att, &next, st);
        float c = vdot(r.d, rec.normal);
        if (c < 0) c = 0;
        return vadd(vscale(0.2f, vmul(vscale(c, att), sky(r.d))), e);
    }
    return sky(r.d);
}

static v3 shade_normal(const struct oracle_scene* s, ray_t r, uint64_t* rays)                      /* render.h:90-103 */
{
    hitrec_t rec;
    (*rays)++;
    if (world_hit(s, r, 0.001f, FLT_MAX, &rec)) return rec.normal;
    return sky(r.d);
}

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

/* ------------------------------------------------------------------- API */
int oracle_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void* oracle_scene_create(int d1)
{
    struct oracle_scene* s = (struct oracle_scene*)calloc(1, sizeof *s);
    s->d1 = d1;
    s->world_xf = XF_IDENTITY;
    return s;
}

void oracle_scene_destroy(void* h)
{
    struct oracle_scene* s = (struct oracle_scene*)h;
    if (!s) return;
    free(s->prims);
    free(s->mats);
    free(s);
}

int oracle_scene_add_material(void* h, uint32_t type, const float* albedo, float param)
{
    struct oracle_scene* s = (struct oracle_scene*)h;
    if (type > 3) return -1;
    if (s->n_mats == s->cap_mats) {
        s->cap_mats = s->cap_mats ? 2 * s->cap_mats : 16;
        s->mats = (mat_t*)realloc(s->mats, s->cap_mats * sizeof(mat_t));
    }
    mat_t* m = &s->mats[s->n_mats];
    m->type = type;
    m->albedo = V(albedo[0], albedo[1], albedo[2]);
    m->param = param;
    m->checker = 0;
    m->odd = V(0, 0, 0);
    return (int)s->n_mats++;
}

void oracle_checker_odd(const float* p, size_t n, uint8_t* out)
{
    mat_t m = {0, V(0, 0, 0), 0.0f, 1, V(1, 1, 1)};
    for (size_t i = 0; i < n; i++) out[i] = texture_value(&m, V(p[3 * i], p[3 * i + 1], p[3 * i + 2])).x > 0.5f ? 1 : 0;
}

int oracle_scene_add_checker_material(void* h, uint32_t type, const float* even, const float* odd)
{
    if (type != 0 && type != 3) return -1;
    int id = oracle_scene_add_material(h, type, even, 0.0f);
    if (id < 0) return id;
    struct oracle_scene* s = (struct oracle_scene*)h;
    s->mats[id].checker = 1;
    s->mats[id].odd = V(odd[0], odd[1], odd[2]);
    return id;
}

static prim_t* push_prims(struct oracle_scene* s, size_t n)
{
    if (s->n_prims + n > s->cap_prims) {
        s->cap_prims = 2 * (s->n_prims + n);
        s->prims = (prim_t*)realloc(s->prims, s->cap_prims * sizeof(prim_t));
    }
    prim_t* p = s->prims + s->n_prims;
    s->n_prims += n;
    return p;
}

void oracle_scene_add_triangles(void* h, const float* tri, const uint32_t* mat, const uint32_t* ids, size_t n)
{
    struct oracle_scene* s = (struct oracle_scene*)h;
    size_t base = s->n_prims;
    prim_t* p = push_prims(s, n);
    for (size_t i = 0; i < n; i++) {
        const float* t = tri + 12 * i;
        memset(&p[i], 0, sizeof(prim_t));
        p[i].kind = 0;
        p[i].v0 = V(t[0], t[1], t[2]); p[i].v1 = V(t[3], t[4], t[5]); p[i].v2 = V(t[6], t[7], t[8]);
        p[i].n = V(t[9], t[10], t[11]);
        p[i].xf = XF_IDENTITY;
        p[i].mat = mat ? mat[i] : 0;
        p[i].id = ids ? ids[i] : (uint32_t)(base + i);
    }
}

void oracle_scene_add_spheres(void* h, const float* sph, const uint32_t* mat, const uint32_t* ids, size_t n)
{
    struct oracle_scene* s = (struct oracle_scene*)h;
    size_t base = s->n_prims;
    prim_t* p = push_prims(s, n);
    for (size_t i = 0; i < n; i++) {
        memset(&p[i], 0, sizeof(prim_t));
        p[i].kind = 1;
        p[i].xf = XF_IDENTITY;
        p[i].xf.position = V(sph[4 * i], sph[4 * i + 1], sph[4 * i + 2]);
        p[i].radius = sph[4 * i + 3];
        p[i].mat = mat ? mat[i] : 0;
        p[i].id = ids ? ids[i] : (uint32_t)(base + i);
    }
}

/* same contract as ref_scene_add_chained (oracle/ref_harness.cpp) */
int oracle_scene_add_chained(void* h, int kind, const float* data, const uint32_t* mat, const uint32_t* ids, const uint8_t* flip,
                             size_t n, const float* levels, int n_levels)
{
    struct oracle_scene* s = (struct oracle_scene*)h;
    if (kind < 0 || kind > 2 || n_levels < 1 || !levels || !ids) return -1;
    prim_t* p = push_prims(s, n);
    for (size_t i = 0; i < n; i++) {
        memset(&p[i], 0, sizeof(prim_t));
        p[i].kind = kind;
        if (kind == 0) {
            const float* t = data + 12 * i;
            p[i].v0 = V(t[0], t[1], t[2]); p[i].v1 = V(t[3], t[4], t[5]); p[i].v2 = V(t[6], t[7], t[8]);
            p[i].n = V(t[9], t[10], t[11]);
        }
        if (kind == 1) p[i].radius = data[i];
        p[i].flip = flip ? flip[i] : 0;
        p[i].n_outer = n_levels - 1;
        p[i].outer = p[i].n_outer ? (xform_t*)malloc(sizeof(xform_t) * (size_t)p[i].n_outer) : NULL;
        for (int l = 0; l < n_levels; l++) {
            const float* q = levels + 6 * l;
            xform_t x = {{q[0], q[1], q[2]}, {q[3], q[4], q[5]}, {1, 1, 1}};
            if (l < n_levels - 1) p[i].outer[l] = x; else p[i].xf = x;
        }
        p[i].mat = mat ? mat[i] : 0;
        p[i].id = ids[i];
    }
    return 0;
}

/* the rows rotate() builds (vec3.h:211-213), as ref_rotation_rows */
void oracle_rotation_rows(const float* rot3, float* out9)
{
    for (int j = 0; j < 3; j++) {
        v3 v = vrotate(V(j == 0 ? 1.f : 0.f, j == 1 ? 1.f : 0.f, j == 2 ? 1.f : 0.f), V(rot3[0], rot3[1], rot3[2]));
        out9[j] = v.x; out9[3 + j] = v.y; out9[6 + j] = v.z;
    }
}

static int cmp_prim_id(const void* a, const void* b)
{
    uint32_t x = ((const prim_t*)a)->id, y = ((const prim_t*)b)->id;
    return x < y ? -1 : x > y;
}

int oracle_scene_finalize(void* h)
{
    struct oracle_scene* s = (struct oracle_scene*)h;
    for (size_t i = 0; i < s->n_prims; i++)
        if (s->prims[i].mat >= s->n_mats) return -1;
    qsort(s->prims, s->n_prims, sizeof(prim_t), cmp_prim_id);   /* ids are unique: order is total */
    s->finalized = 1;
    return 0;
}

double oracle_trace(void* h, const float* rays, size_t n, float t_min, float t_max, int use_bvh,
                    float* out_t, uint32_t* out_id, float* out_p, float* out_n)
{
    struct oracle_scene* s = (struct oracle_scene*)h;
    (void)use_bvh;
    if (!s->finalized) return -1;
    double t0 = now_s();
#pragma omp parallel for schedule(dynamic, 16)
    for (long long i = 0; i < (long long)n; i++) {
        const float* q = rays + 6 * i;
        ray_t r = {V(q[0], q[1], q[2]), V(q[3], q[4], q[5])};
        hitrec_t rec;
        int hit = world_hit(s, r, t_min, t_max, &rec);
        if (out_t) out_t[i] = hit ? rec.t : 0.f;
        if (out_id) out_id[i] = hit ? rec.id : 0xFFFFFFFFu;
        if (out_p && hit) { out_p[3 * i] = rec.p.x; out_p[3 * i + 1] = rec.p.y; out_p[3 * i + 2] = rec.p.z; }
        if (out_n && hit) { out_n[3 * i] = rec.normal.x; out_n[3 * i + 1] = rec.normal.y; out_n[3 * i + 2] = rec.normal.z; }
    }
    return now_s() - t0;
}

int oracle_scatter(uint32_t type, const float* albedo, float param, const float* ray, const float* recpn,
                   const float* randoms, int n_randoms, float* out, int* n_draws)
{
    if (type > 3) return -1;
    mat_t m = {type, V(albedo[0], albedo[1], albedo[2]), param, 0, V(0, 0, 0)};
    rng_t st = {0, 0, randoms, n_randoms, 0};
    ray_t in = {V(ray[0], ray[1], ray[2]), V(ray[3], ray[4], ray[5])};
    hitrec_t rec;
    memset(&rec, 0, sizeof rec);
    rec.t = 1.f;
    rec.p = V(recpn[0], recpn[1], recpn[2]);
    rec.normal = V(recpn[3], recpn[4], recpn[5]);
    v3 att = V(0, 0, 0);
    ray_t sc = {V(0, 0, 0), V(0, 0, 0)};
    int ok = scatter(&m, in, &rec, &att, &sc, &st);
    out[0] = att.x; out[1] = att.y; out[2] = att.z;
    out[3] = sc.o.x; out[4] = sc.o.y; out[5] = sc.o.z;
    out[6] = sc.d.x; out[7] = sc.d.y; out[8] = sc.d.z;
    if (n_draws) *n_draws = st.inject_pos;
    return ok;
}

/* core/camera.h:18-38 */
void oracle_camera(const float* lookfrom, const float* lookat, const float* vup, float vfov, float aspect,
                   float aperture, float focus_dist, float* out22)
{
    float lens_radius = aperture / 2.0f;
    float theta = (float)(vfov * M_PI / 180.0f);
    float half_height = tanf(theta / 2.0f);
    float half_width = half_height * aspect;
    v3 origin = V(lookfrom[0], lookfrom[1], lookfrom[2]);
    v3 z = vunit(vsub(origin, V(lookat[0], lookat[1], lookat[2])));
    v3 x = vunit(vcross(V(vup[0], vup[1], vup[2]), z));
    v3 y = vcross(z, x);
    v3 llc = vsub(vsub(vsub(origin, vscale(half_width * focus_dist, x)), vscale(half_height * focus_dist, y)),
                  vscale(focus_dist, z));
    v3 horizontal = vscale(2.0f * half_width * focus_dist, x);
    v3 vertical = vscale(2.0f * half_height * focus_dist, y);
    const v3* v[7] = {&llc, &horizontal, &vertical, &origin, &x, &y, &z};
    for (int i = 0; i < 7; i++) { out22[3 * i] = v[i]->x; out22[3 * i + 1] = v[i]->y; out22[3 * i + 2] = v[i]->z; }
    out22[21] = lens_radius;
}

uint32_t oracle_get_ray(const float* cam22, float s, float t, uint64_t seed, uint32_t pixel, uint32_t sample,
                        uint32_t ctr, float* out6)
{
    camera_t cam = camera_from22(cam22);
    rng_t st = {crt_rng_key(seed, pixel, sample), ctr, NULL, 0, 0};
    ray_t r = get_ray(&cam, s, t, &st);
    out6[0] = r.o.x; out6[1] = r.o.y; out6[2] = r.o.z;
    out6[3] = r.d.x; out6[4] = r.d.y; out6[5] = r.d.z;
    return st.ctr;
}

/* core/render.h:105-129 */
double oracle_render(void* h, const float* cam22, int nx, int ny, int ns, int max_depth, int mode,
                     uint64_t seed, int y0, int y1, int y_step, float* rgb, float* accum, uint64_t* rays_out)
{
    struct oracle_scene* s = (struct oracle_scene*)h;
    if (!s->finalized) return -1;
    camera_t cam = camera_from22(cam22);
    if (y_step < 1) y_step = 1;
    uint64_t total_rays = 0;
    double t0 = now_s();
    /* one parallel loop over the pixels of the selected rows (a bounded sample may have fewer rows than cores) */
    const long long n_rows = y1 > y0 ? (y1 - y0 + y_step - 1) / y_step : 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : total_rays)
    for (long long q = 0; q < n_rows * nx; q++) {
        uint64_t rays = 0;
        {
            const int y = y0 + (int)(q / nx) * y_step, x = (int)(q % nx);
            int pixel_index = y * nx + x;
            v3 col = V(0, 0, 0);
            for (int i = 0; i < ns; i++) {
                rng_t st = {crt_rng_key(seed, (uint32_t)pixel_index, (uint32_t)i), 0, NULL, 0, 0};
                float u = (float)(x + rnd(&st)) / (float)nx;
                float v = (float)(y + rnd(&st)) / (float)ny;
                ray_t r = get_ray(&cam, u, v, &st);
                v3 c;
                if (mode == 0) c = shade_path(s, r, max_depth, &st, &rays);
                else if (mode == 1) c = shade_lambert(s, r, &st, &rays);
                else c = shade_normal(s, r, &rays);
                col = vadd(col, c);
            }
            if (accum) { accum[3 * pixel_index] = col.x; accum[3 * pixel_index + 1] = col.y; accum[3 * pixel_index + 2] = col.z; }
            float k = (float)(1.0 / (float)ns);             /* vec3.h:182-189 */
            col = V(col.x * k, col.y * k, col.z * k);
            col = V(sqrtf(col.x), sqrtf(col.y), sqrtf(col.z));
            col = V(clip1(col.x, 0, 1), clip1(col.y, 0, 1), clip1(col.z, 0, 1));
            if (rgb) { rgb[3 * pixel_index] = col.x; rgb[3 * pixel_index + 1] = col.y; rgb[3 * pixel_index + 2] = col.z; }
        }
        total_rays += rays;
    }
    if (rays_out) *rays_out = total_rays;
    return now_s() - t0;
}

/* core/render.h:137-146 */
void oracle_pack_rgba8(const float* rgb, int nx, int ny, uint8_t* rgba)
{
    for (int i = 0; i < ny; i++)
        for (int j = 0; j < nx; j++) {
            size_t src = (size_t)(ny - 1 - i) * nx + j;
            size_t dst = (size_t)i * nx + j;
            for (int k = 0; k < 3; k++) rgba[4 * dst + k] = (uint8_t)(int)(255.99 * rgb[3 * src + k]);
            rgba[4 * dst + 3] = 255;
        }
}

/* Loader/obj_loader.h:20-133: "v x y z" and "f a b c" lines only; indices stay
 * 1-based floats; points are mean-centred. */
int oracle_parse_obj(const char* path, float* points, float* idx, int max_points, int max_tris, int* np_out, int* nt_out)
{
    FILE* f = fopen(path, "r");
    if (!f) return -1;
    char line[1024];
    int np = 0, nt = 0;
    while (fgets(line, sizeof line, f)) {
        char label[64];
        int off = 0;
        if (sscanf(line, "%63s%n", label, &off) != 1) continue;
        if (strcmp(label, "v") == 0) {
            float a = 0, b = 0, c = 0;
            sscanf(line + off, "%f %f %f", &a, &b, &c);
            if (np >= max_points) { fclose(f); return -2; }
            points[3 * np] = a; points[3 * np + 1] = b; points[3 * np + 2] = c;
            np++;
        } else if (strcmp(label, "f") == 0) {
            float a = 0, b = 0, c = 0;
            sscanf(line + off, "%f %f %f", &a, &b, &c);
            if (nt >= max_tris) { fclose(f); return -2; }
            idx[3 * nt] = a; idx[3 * nt + 1] = b; idx[3 * nt + 2] = c;
            nt++;
        }
    }
    fclose(f);
    /* computeMean / centering, obj_loader.h:120-133 (mean / float(np) is a true divide, vec3.h:105) */
    v3 mean = V(0, 0, 0);
    for (int i = 0; i < np; i++) mean = vadd(mean, V(points[3 * i], points[3 * i + 1], points[3 * i + 2]));
    mean = vdivf(mean, (float)np);
    for (int i = 0; i < np; i++) { points[3 * i] -= mean.x; points[3 * i + 1] -= mean.y; points[3 * i + 2] -= mean.z; }
    *np_out = np;
    *nt_out = nt;
    return 0;
}

/* shapes/triangle.h:14-17 */
void oracle_triangle_normal(const float* v9, float* n3)
{
    v3 v0 = V(v9[0], v9[1], v9[2]), v1 = V(v9[3], v9[4], v9[5]), v2 = V(v9[6], v9[7], v9[8]);
    v3 n = vunit(vcross(vsub(v1, v0), vsub(v2, v0)));
    n3[0] = n.x; n3[1] = n.y; n3[2] = n.z;
}
