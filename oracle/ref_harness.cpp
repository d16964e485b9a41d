// ref_harness.cpp — TEST INFRASTRUCTURE, "parity pinned by the reference itself".
//
// Compiles the UNMODIFIED reference headers under /root/reference/CudaTest/src
// as host C++ (through oracle/shim) and exports a small C ABI around them so
// that tests/ and bench.py's reference arm can run the reference's own
// hit / scatter / camera / BVH / OBJ code on the host cores.
//
// Nothing under cudaraytracer_b200/ may load this library; only tests/,
// __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) do.
//
// What comes from the reference, unmodified: vec3, Ray, AABB, Camera,
// Transform, Hitable::hit, HitableList, BVHNode, Sphere, Triangle, Lambertian,
// Metal, Dielectric, DiffuseLight, ConstantTexture, parseObjByName.
// What is restated here because core/render.h cannot be included (it pulls in
// createScene.h's kernel launches, the FBX SDK and <windows.h>):
// backgroundSky, shade, LambertShade, shade_normal, the render pixel loop and
// WritePng's float->RGBA8 conversion (core/render.h:41-157). Each restated
// function cites the lines it follows.
//
// Deviation D1: Triangle::collision_detection never tests t
// against [t_min, t_max] (shapes/triangle.h:92-99), so the raw list scan
// accepts hits behind the ray origin. With d1 != 0 the harness wraps the
// unmodified test in a filter t_min < t < t_max; with d1 == 0 the raw
// behaviour is kept so tests can count how many rays D1 changes.

#include "cuda_shim.h"

#include <cstdio>
#include <vector>
#include <string>
#include <chrono>
#ifdef _OPENMP
#include <omp.h>
#endif

// Reference headers, in dependency order (they do not include what they use).
#include "core/vec3.h"
#include "core/ray.h"
#include "core/aabb.h"
#include "core/camera.h"
#include "hitable/hitable.h"
#include "hitable/hitable_list.h"
#include "hitable/bvh.h"
#include "shapes/sphere.h"
#include "shapes/triangle.h"
#include "shapes/rectangle.h"
#include "material/material.h"
#include "Loader/obj_loader.h"

namespace {

// The primitive id rides in HitRecord::u / ::v, which neither Triangle nor
// Sphere writes. Two 16-bit halves stay exact in a float.
inline void tag(HitRecord& rec, uint32_t id)
{
    rec.u = float(id & 0xFFFFu);
    rec.v = float(id >> 16);
}
inline uint32_t untag(const HitRecord& rec)
{
    return uint32_t(rec.u) | (uint32_t(rec.v) << 16);
}

// Deviation D3 (cudaraytracer_b200/csrc/crt_device.cuh:triangle_test): one axis of the interval in which the ray is
// inside the box of v0, v0 + e1, v0 + e2, in plain rounded float operations in this order.
inline void d3_axis(float o, float d, float a, float e1, float e2, float& t_near, float& t_far)
{
    float b = a + e1, c = a + e2;
    float lo = fminf(a, fminf(b, c)), hi = fmaxf(a, fmaxf(b, c));
    float pad = fmaxf(fabsf(lo), fabsf(hi)) * 2.384185791015625e-7f + 1e-30f;   /* 2^-22 of the largest coordinate */
    lo = lo - pad;
    hi = hi + pad;
    float inv = fabsf(d) > 1e-18f ? 1.0f / d : copysignf(1e18f, d);
    float t0 = (lo - o) * inv, t1 = (hi - o) * inv;
    t_near = fmaxf(t_near, fminf(t0, t1));
    t_far = fminf(t_far, fmaxf(t0, t1));
}

class TaggedTriangle : public Triangle {
public:
    TaggedTriangle(vec3 vs[3], vec3 n, Material* m, Transform* t, uint32_t id_, bool d1_)
        : Triangle(vs, n, m, false, t, false), id(id_), d1(d1_) {}
    bool collision_detection(const Ray& r, float t_min, float t_max, HitRecord& rec,
                             int frameIndex) const override
    {
        if (!Triangle::collision_detection(r, t_min, t_max, rec, frameIndex)) return false;
        if (d1 && !(rec.t > t_min && rec.t < t_max)) return false;
        if (d1) {                                   // deviation D3, applied together with D1
            const vec3 e1 = vertices[1] - vertices[0], e2 = vertices[2] - vertices[0];
            const vec3 o = r.origin(), d = r.direction();
            float t_near = -INFINITY, t_far = INFINITY;
            for (int k = 0; k < 3; k++) d3_axis(o[k], d[k], vertices[0][k], e1[k], e2[k], t_near, t_far);
            const float E = 3.0517578125e-5f;        // 2^-15
            float m = fabsf(t_near) * E;
            if (!(rec.t >= t_near - m)) return false;
            m = fabsf(t_far) * E;
            if (!(rec.t <= t_far + m)) return false;
        }
        tag(rec, id);
        return true;
    }
    uint32_t id;
    bool d1;
};

class TaggedSphere : public Sphere {
public:
    TaggedSphere(Transform* t, float r, Material* m, uint32_t id_) : Sphere(t, r, m), id(id_) {}
    bool collision_detection(const Ray& r, float t_min, float t_max, HitRecord& rec,
                             int frameIndex) const override
    {
        if (!Sphere::collision_detection(r, t_min, t_max, rec, frameIndex)) return false;
        tag(rec, id);
        return true;
    }
    uint32_t id;
};

class TaggedRectangle : public Rectangle {
public:
    TaggedRectangle(Material* m, bool flip, Transform* t, uint32_t id_) : Rectangle(m, flip, t), id(id_) {}
    bool collision_detection(const Ray& r, float t_min, float t_max, HitRecord& rec,
                             int frameIndex) const override
    {
        if (!Rectangle::collision_detection(r, t_min, t_max, rec, frameIndex)) return false;
        tag(rec, id);                               // Rectangle writes rec.u / rec.v (rectangle.h:37-38); the id replaces them
        return true;
    }
    uint32_t id;
};

struct PendingPrim {
    int kind;  // 0 triangle, 1 sphere, 2 rectangle
    float data[12];
    uint32_t mat;
    uint32_t id;
    int flip = 0;
    // Transforms of the Hitables between the world list and this primitive, outermost first; the last one is the
    // primitive's own. Empty: the plain case (identity for a triangle, position = centre for a sphere).
    std::vector<float> levels;   // 6 floats each: position, rotation in degrees
};

struct RefScene {
    std::vector<Material*> mats;
    std::vector<PendingPrim> pending;
    std::vector<Hitable*> prims;  // list order == ascending prim id order given by the caller
    HitableList* world = nullptr;
    BVHNode* bvh = nullptr;
    Hitable** bvh_storage = nullptr;
    int d1 = 1;
    int use_bvh = 0;   // ref_render goes through the reference's BVHNode instead of the list scan
};

Transform* identity_transform() { return new Transform(vec3(0), vec3(0), vec3(1)); }

// core/render.h:41-46
vec3 sky(const vec3& d)
{
    vec3 n = unit_vector(d);
    float s = 0.5f * (n[1] + 1.0f);
    return lerp(s, vec3(1), vec3(0.5f, 0.7f, 1.0f));
}

// core/render.h:48-67
vec3 shade_path(const Ray& r, Hitable** world, int depth, curandState* st, uint64_t* rays)
{
    HitRecord rec;
    if (rays) (*rays)++;
    if ((*world)->hit(r, 0.001, FLT_MAX, rec, 0)) {
        Ray next;
        vec3 att;
        vec3 e = rec.mat_ptr->emitted(rec.u, rec.v, rec.p);
        if (depth > 0 && rec.mat_ptr->scatter(r, rec, att, next, st))
            return e + att * shade_path(next, world, depth - 1, st, rays);
        return e + vec3(0.1, 0.1, 0.1);
    }
    return sky(r.direction());
}

// core/render.h:70-87
vec3 shade_lambert(const Ray& r, Hitable** world, curandState* st, uint64_t* rays)
{
    HitRecord rec;
    if (rays) (*rays)++;
    if ((*world)->hit(r, 0.001, FLT_MAX, rec, 0)) {
        Ray next;
        vec3 att;
        vec3 e = rec.mat_ptr->emitted(rec.u, rec.v, rec.p);
        rec.mat_ptr->scatter(r, rec, att, next, st);
        float c = dot(r.direction(), rec.normal);
        if (c < 0) c = 0;
        return att * c * sky(r.direction()) * 0.2 + e;
    }
    return sky(r.direction());
}

// core/render.h:90-103
vec3 shade_nrm(const Ray& r, Hitable** world, uint64_t* rays)
{
    HitRecord rec;
    if (rays) (*rays)++;
    if ((*world)->hit(r, 0.001, FLT_MAX, rec, 0)) return rec.normal;
    return sky(r.direction());
}

Camera camera_from(const float* c)
{
    // Rebuild a reference Camera object from its 22 members without running the
    // constructor's trigonometry again.
    Camera cam(vec3(0, 0, 1), vec3(0, 0, 0), vec3(0, 1, 0), 40.f, 1.f, 0.f, 1.f);
    cam.lower_left_corner = vec3(c[0], c[1], c[2]);
    cam.horizontal = vec3(c[3], c[4], c[5]);
    cam.vertical = vec3(c[6], c[7], c[8]);
    cam.origin = vec3(c[9], c[10], c[11]);
    cam.x = vec3(c[12], c[13], c[14]);
    cam.y = vec3(c[15], c[16], c[17]);
    cam.z = vec3(c[18], c[19], c[20]);
    cam.lens_radius = c[21];
    return cam;
}

Material* make_material(uint32_t type, const float* albedo, float param)
{
    switch (type) {
    case 0: return new Lambertian(new ConstantTexture(vec3(albedo[0], albedo[1], albedo[2])));
    case 1: return new Metal(vec3(albedo[0], albedo[1], albedo[2]), param);
    case 2: return new Dielectric(param);
    case 3: return new DiffuseLight(new ConstantTexture(vec3(albedo[0], albedo[1], albedo[2])));
    }
    return nullptr;
}

}  // namespace

extern "C" {

int ref_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void* ref_scene_create(int d1)
{
    RefScene* s = new RefScene();
    s->d1 = d1;
    return s;
}

// CheckerTexture::value of n points x 3 floats: 1 where it returns the odd texture (material/texture.h:30-38)
void ref_checker_odd(const float* p, size_t n, uint8_t* out)
{
    ConstantTexture even(vec3(0, 0, 0)), odd(vec3(1, 1, 1));
    CheckerTexture c(&even, &odd);
    for (size_t i = 0; i < n; i++) out[i] = c.value(0, 0, vec3(p[3 * i], p[3 * i + 1], p[3 * i + 2]))[0] > 0.5f ? 1 : 0;
}

// Lambertian / DiffuseLight over CheckerTexture(ConstantTexture(even), ConstantTexture(odd)), material/texture.h:25-42
int ref_scene_add_checker_material(void* h, uint32_t type, const float* even, const float* odd)
{
    RefScene* s = (RefScene*)h;
    Texture* t = new CheckerTexture(new ConstantTexture(vec3(even[0], even[1], even[2])),
                                    new ConstantTexture(vec3(odd[0], odd[1], odd[2])));
    Material* m = type == 0 ? (Material*)new Lambertian(t) : type == 3 ? (Material*)new DiffuseLight(t) : nullptr;
    if (!m) return -1;
    s->mats.push_back(m);
    return int(s->mats.size()) - 1;
}

int ref_scene_add_material(void* h, uint32_t type, const float* albedo, float param)
{
    RefScene* s = (RefScene*)h;
    Material* m = make_material(type, albedo, param);
    if (!m) return -1;
    s->mats.push_back(m);
    return int(s->mats.size()) - 1;
}

// tri: v0 v1 v2 normal (12 floats), as include/crt.h crt_scene_set_triangles.
void ref_scene_add_triangles(void* h, const float* tri, const uint32_t* mat, const uint32_t* ids, size_t n)
{
    RefScene* s = (RefScene*)h;
    size_t base = s->pending.size();
    for (size_t i = 0; i < n; i++) {
        PendingPrim p;
        p.kind = 0;
        memcpy(p.data, tri + 12 * i, 48);
        p.mat = mat ? mat[i] : 0;
        p.id = ids ? ids[i] : uint32_t(base + i);
        s->pending.push_back(p);
    }
}

// sph: centre xyz + radius.
void ref_scene_add_spheres(void* h, const float* sph, const uint32_t* mat, const uint32_t* ids, size_t n)
{
    RefScene* s = (RefScene*)h;
    size_t base = s->pending.size();
    for (size_t i = 0; i < n; i++) {
        PendingPrim p;
        p.kind = 1;
        memset(p.data, 0, sizeof p.data);
        memcpy(p.data, sph + 4 * i, 16);
        p.mat = mat ? mat[i] : 0;
        p.id = ids ? ids[i] : uint32_t(base + i);
        s->pending.push_back(p);
    }
}

// Primitives under a chain of Transforms: `levels` (n_levels x 6 floats: position, rotation in degrees) are the
// Transforms of the Hitables between the world list and the primitive, outermost first, the last being the
// primitive's own. kind 0: data = n x 12 floats (v0 v1 v2 normal in the primitive's frame); kind 1: data = n radii;
// kind 2 (Rectangle, shapes/rectangle.h): no data, flip = flipNormal per rectangle. Every primitive gets its own
// nest of single-entry HitableLists, which the list scan cannot tell from shared ones.
int ref_scene_add_chained(void* h, int kind, const float* data, const uint32_t* mat, const uint32_t* ids, const uint8_t* flip,
                          size_t n, const float* levels, int n_levels)
{
    RefScene* s = (RefScene*)h;
    if (kind < 0 || kind > 2 || n_levels < 1 || !levels || !ids) return -1;
    for (size_t i = 0; i < n; i++) {
        PendingPrim p;
        p.kind = kind;
        memset(p.data, 0, sizeof p.data);
        if (kind == 0) memcpy(p.data, data + 12 * i, 48);
        if (kind == 1) p.data[3] = data[i];
        p.mat = mat ? mat[i] : 0;
        p.id = ids[i];
        p.flip = flip ? flip[i] : 0;
        p.levels.assign(levels, levels + 6 * (size_t)n_levels);
        s->pending.push_back(p);
    }
    return 0;
}

// rotate(v, rotation) (core/vec3.h:200-217) applied to the three unit vectors: out9[3 k + j] = rotate(e_j)[k], i.e.
// the rows rotate0..2 the function builds, as this host's libm evaluates them.
void ref_rotation_rows(const float* rot3, float* out9)
{
    const vec3 r(rot3[0], rot3[1], rot3[2]);
    for (int j = 0; j < 3; j++) {
        vec3 e(j == 0 ? 1.f : 0.f, j == 1 ? 1.f : 0.f, j == 2 ? 1.f : 0.f);
        vec3 v = rotate(e, r);
        for (int k = 0; k < 3; k++) out9[3 * k + j] = v[k];
    }
}

// Builds the reference objects. The world is a HitableList whose entry order
// is ascending prim id, i.e. the caller's list order, with an explicit
// identity Transform (the reference's default-constructed one is
// uninitialised, hitable/transform.h:8 + core/vec3.h:9).
int ref_scene_finalize(void* h)
{
    RefScene* s = (RefScene*)h;
    std::stable_sort(s->pending.begin(), s->pending.end(),
                     [](const PendingPrim& a, const PendingPrim& b) { return a.id < b.id; });
    for (const PendingPrim& p : s->pending) {
        if (p.mat >= s->mats.size()) return -1;
        Material* m = s->mats[p.mat];
        if (!p.levels.empty()) {
            // one nested HitableList per enclosing level, then the primitive under its own Transform
            const size_t nl = p.levels.size() / 6;
            auto xf = [&](size_t l) {
                const float* q = p.levels.data() + 6 * l;
                return new Transform(vec3(q[0], q[1], q[2]), vec3(q[3], q[4], q[5]), vec3(1));
            };
            Hitable* leaf;
            if (p.kind == 0) {
                vec3 vs[3] = {vec3(p.data[0], p.data[1], p.data[2]), vec3(p.data[3], p.data[4], p.data[5]),
                              vec3(p.data[6], p.data[7], p.data[8])};
                leaf = new TaggedTriangle(vs, vec3(p.data[9], p.data[10], p.data[11]), m, xf(nl - 1), p.id, s->d1 != 0);
            } else if (p.kind == 1) {
                leaf = new TaggedSphere(xf(nl - 1), p.data[3], m, p.id);
            } else {
                leaf = new TaggedRectangle(m, p.flip != 0, xf(nl - 1), p.id);
            }
            for (size_t l = nl - 1; l-- > 0;) {
                Hitable** one = (Hitable**)malloc(sizeof(Hitable*));
                one[0] = leaf;
                leaf = new HitableList(one, 1, xf(l));
            }
            s->prims.push_back(leaf);
            continue;
        }
        if (p.kind == 0) {
            vec3 vs[3] = {vec3(p.data[0], p.data[1], p.data[2]), vec3(p.data[3], p.data[4], p.data[5]),
                          vec3(p.data[6], p.data[7], p.data[8])};
            vec3 n(p.data[9], p.data[10], p.data[11]);
            s->prims.push_back(new TaggedTriangle(vs, n, m, identity_transform(), p.id, s->d1 != 0));
        } else {
            Transform* t = new Transform(vec3(p.data[0], p.data[1], p.data[2]), vec3(0), vec3(1));
            s->prims.push_back(new TaggedSphere(t, p.data[3], m, p.id));
        }
    }
    s->world = new HitableList(s->prims.data(), int(s->prims.size()), identity_transform());
    return 0;
}

// The reference's own accelerator (hitable/bvh.h:76-125) built host-side; the
// axis draws come from the counter-based stream keyed by `seed`.
double ref_bvh_build(void* h, uint64_t seed)
{
    RefScene* s = (RefScene*)h;
    auto t0 = std::chrono::steady_clock::now();
    curandState st;
    st.key = crt_rng_key(seed, 0, 0);
    size_t n = s->prims.size();
    s->bvh_storage = (Hitable**)malloc(sizeof(Hitable*) * n);
    for (size_t i = 0; i < n; i++) s->bvh_storage[i] = s->prims[i];
    s->bvh = new BVHNode(s->bvh_storage, int(n), 0, 1, &st);
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

// hitable/bvh.h:135-158
double ref_bvh_refit(void* h)
{
    RefScene* s = (RefScene*)h;
    if (!s->bvh) return -1;
    auto t0 = std::chrono::steady_clock::now();
    s->bvh->UpdateBVH();
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

// Which accelerator ref_render uses: 0 = HitableList scan (the parity oracle), 1 = the reference's shipping BVHNode
// (kernel.cu:24-39 renders through it; needs ref_bvh_build first). Returns the setting in force.
int ref_scene_use_bvh(void* h, int flag)
{
    RefScene* s = (RefScene*)h;
    s->use_bvh = (flag && s->bvh) ? 1 : 0;
    return s->use_bvh;
}

// Statistics of the reference's BVHNode traversal (hitable/bvh.h:160-190 with Hitable::hit, hitable/hitable.h:29-38)
// for n rays: nodes whose box was tested and primitives handed to a leaf's list scan. The walk below restates the
// control flow only to COUNT; it calls the reference's own Transform::TransformRay and AABB::hit, and like the
// reference it descends into both children and never shrinks t_max.
static void bvh_count(BVHNode* node, const Ray& r, float t_min, float t_max, uint64_t& nodes, uint64_t& prims)
{
    Ray tr = node->transform->TransformRay(r);
    nodes++;
    if (!node->box.hit(tr, t_min, t_max)) return;
    if (node->isLeaf) { prims += (uint64_t)node->childList->list_size; return; }
    bvh_count(node->left, tr, t_min, t_max, nodes, prims);
    bvh_count(node->right, tr, t_min, t_max, nodes, prims);
}

int ref_bvh_stats(void* h, const float* rays, size_t n, float t_min, float t_max, uint64_t* nodes_out, uint64_t* prims_out)
{
    RefScene* s = (RefScene*)h;
    if (!s->bvh) return -1;
    uint64_t nodes = 0, prims = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : nodes, prims)
    for (long long i = 0; i < (long long)n; i++) {
        const float* q = rays + 6 * i;
        Ray r(vec3(q[0], q[1], q[2]), vec3(q[3], q[4], q[5]), 0.f);
        bvh_count(s->bvh, r, t_min, t_max, nodes, prims);
    }
    if (nodes_out) *nodes_out = nodes;
    if (prims_out) *prims_out = prims;
    return 0;
}

// Closest hit for n rays through (*world)->hit (core/render.h:53). rays: n x 6
// floats (origin, direction). use_bvh selects the reference BVHNode instead of
// the list scan. Outputs may be NULL. Returns seconds spent.
double ref_trace(void* h, const float* rays, size_t n, float t_min, float t_max, int use_bvh,
                 float* out_t, uint32_t* out_id, float* out_p, float* out_n)
{
    RefScene* s = (RefScene*)h;
    Hitable* world = use_bvh ? (Hitable*)s->bvh : (Hitable*)s->world;
    if (!world) return -1;
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 16)
    for (long long i = 0; i < (long long)n; i++) {
        const float* q = rays + 6 * i;
        Ray r(vec3(q[0], q[1], q[2]), vec3(q[3], q[4], q[5]), 0.f);
        HitRecord rec;
        rec.u = rec.v = 0;
        bool hit = world->hit(r, t_min, t_max, rec, 0);
        if (out_t) out_t[i] = hit ? rec.t : 0.f;
        if (out_id) out_id[i] = hit ? untag(rec) : 0xFFFFFFFFu;
        if (out_p && hit) { out_p[3 * i] = rec.p[0]; out_p[3 * i + 1] = rec.p[1]; out_p[3 * i + 2] = rec.p[2]; }
        if (out_n && hit) { out_n[3 * i] = rec.normal[0]; out_n[3 * i + 1] = rec.normal[1]; out_n[3 * i + 2] = rec.normal[2]; }
    }
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

// One Material::scatter call (material/material.h:57-143) with an injected
// random sequence. in: ray origin+direction (6), rec: p (3), normal (3).
// out: attenuation (3), scattered origin (3), scattered direction (3).
// Returns the bool of scatter(); *n_draws = uniforms consumed.
int ref_scatter(uint32_t type, const float* albedo, float param, const float* ray, const float* recpn,
                const float* randoms, int n_randoms, float* out, int* n_draws)
{
    Material* m = make_material(type, albedo, param);
    if (!m) return -1;
    curandState st;
    st.inject = randoms;
    st.inject_n = n_randoms;
    Ray r(vec3(ray[0], ray[1], ray[2]), vec3(ray[3], ray[4], ray[5]), 0.f);
    HitRecord rec;
    rec.t = 1.f;
    rec.u = rec.v = 0.f;
    rec.p = vec3(recpn[0], recpn[1], recpn[2]);
    rec.normal = vec3(recpn[3], recpn[4], recpn[5]);
    rec.mat_ptr = m;
    vec3 att(0, 0, 0);
    Ray sc;
    bool ok = m->scatter(r, rec, att, sc, &st);
    for (int k = 0; k < 3; k++) {
        out[k] = att[k];
        out[3 + k] = sc.origin()[k];
        out[6 + k] = sc.direction()[k];
    }
    if (n_draws) *n_draws = st.inject_pos;
    return ok ? 1 : 0;
}

// Reference Camera constructor (core/camera.h:18-38) -> the 22 floats of
// crt_camera.
void ref_camera(const float* lookfrom, const float* lookat, const float* vup, float vfov, float aspect,
                float aperture, float focus_dist, float* out22)
{
    Camera c(vec3(lookfrom[0], lookfrom[1], lookfrom[2]), vec3(lookat[0], lookat[1], lookat[2]),
             vec3(vup[0], vup[1], vup[2]), vfov, aspect, aperture, focus_dist);
    const vec3* v[7] = {&c.lower_left_corner, &c.horizontal, &c.vertical, &c.origin, &c.x, &c.y, &c.z};
    for (int i = 0; i < 7; i++)
        for (int k = 0; k < 3; k++) out22[3 * i + k] = (*v[i])[k];
    out22[21] = c.lens_radius;
}

// Camera::get_ray (core/camera.h:40-44) with the stream of (seed,pixel,sample)
// positioned at draw `ctr`. out: origin+direction; returns the new counter.
uint32_t ref_get_ray(const float* cam22, float s, float t, uint64_t seed, uint32_t pixel, uint32_t sample,
                     uint32_t ctr, float* out6)
{
    Camera cam = camera_from(cam22);
    curandState st;
    st.key = crt_rng_key(seed, pixel, sample);
    st.ctr = ctr;
    Ray r = cam.get_ray(s, t, &st);
    for (int k = 0; k < 3; k++) { out6[k] = r.origin()[k]; out6[3 + k] = r.direction()[k]; }
    return st.ctr;
}

// The render pixel loop (core/render.h:105-129) over rows [y0, y1) with the
// row stride/sample subsetting the bench's bounded sample needs.
// mode: 0 shade, 1 LambertShade, 2 shade_normal. rgb: nx*ny*3, row 0 bottom.
// accum (optional): the per-pixel sum before /ns, sqrt, clip.
// Per Q11 the per-pixel stream is handed to get_ray (the reference passes
// state[0], a data race). Returns seconds; *rays_out = closest-hit queries.
double ref_render(void* h, const float* cam22, int nx, int ny, int ns, int max_depth, int mode,
                  uint64_t seed, int y0, int y1, int y_step, float* rgb, float* accum, uint64_t* rays_out)
{
    RefScene* s = (RefScene*)h;
    if (!s->world) return -1;
    Hitable* world_ptr = (s->use_bvh && s->bvh) ? (Hitable*)s->bvh : (Hitable*)s->world;
    Hitable** world = &world_ptr;
    Camera cam = camera_from(cam22);
    if (y_step < 1) y_step = 1;
    uint64_t total_rays = 0;
    auto t0 = std::chrono::steady_clock::now();
    // one parallel loop over the pixels of the selected rows (a bounded sample may have fewer rows than cores)
    const long long n_rows = y1 > y0 ? (y1 - y0 + y_step - 1) / y_step : 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : total_rays)
    for (long long k = 0; k < n_rows * nx; k++) {
        uint64_t rays = 0;
        {
            const int y = y0 + (int)(k / nx) * y_step, x = (int)(k % nx);
            int pixel_index = y * nx + x;
            vec3 col(0, 0, 0);
            for (int i = 0; i < ns; i++) {
                curandState st;
                st.key = crt_rng_key(seed, uint32_t(pixel_index), uint32_t(i));
                st.ctr = 0;
                float u = float(x + curand_uniform(&st)) / float(nx);
                float v = float(y + curand_uniform(&st)) / float(ny);
                Ray r = cam.get_ray(u, v, &st);
                if (mode == 0) col += shade_path(r, world, max_depth, &st, &rays);
                else if (mode == 1) col += shade_lambert(r, world, &st, &rays);
                else col += shade_nrm(r, world, &rays);
            }
            if (accum) { accum[3 * pixel_index] = col[0]; accum[3 * pixel_index + 1] = col[1]; accum[3 * pixel_index + 2] = col[2]; }
            col /= float(ns);
            col[0] = sqrt(col[0]);
            col[1] = sqrt(col[1]);
            col[2] = sqrt(col[2]);
            col = clip(col);
            if (rgb) { rgb[3 * pixel_index] = col[0]; rgb[3 * pixel_index + 1] = col[1]; rgb[3 * pixel_index + 2] = col[2]; }
        }
        total_rays += rays;
    }
    auto t1 = std::chrono::steady_clock::now();
    if (rays_out) *rays_out = total_rays;
    return std::chrono::duration<double>(t1 - t0).count();
}

// WritePng's conversion (core/render.h:137-146): vertical flip, char(255.99*c), alpha 255.
void ref_pack_rgba8(const float* rgb, int nx, int ny, uint8_t* rgba)
{
    for (int i = 0; i < ny; i++)
        for (int j = 0; j < nx; j++) {
            size_t src = size_t(ny - 1 - i) * nx + j;
            size_t dst = size_t(i) * nx + j;
            for (int k = 0; k < 3; k++) rgba[4 * dst + k] = (uint8_t)(int)(255.99 * rgb[3 * src + k]);
            rgba[4 * dst + 3] = 255;
        }
}

// Loader/obj_loader.h:20-111. points/idx must hold max_points / max_tris vec3.
int ref_parse_obj(const char* path, float* points, float* idx, int max_points, int max_tris, int* np, int* nt)
{
    // Count first so the caller-allocated arrays cannot overflow (the
    // reference trusts the caller, obj_loader.h:154-155).
    FILE* f = fopen(path, "r");
    if (!f) return -1;
    char line[512];
    int cv = 0, cf = 0;
    while (fgets(line, sizeof line, f)) {
        if (line[0] == 'v' && line[1] == ' ') cv++;
        else if (line[0] == 'f' && line[1] == ' ') cf++;
    }
    fclose(f);
    if (cv > max_points || cf > max_tris) return -2;
    std::vector<vec3> p(cv > 0 ? cv : 1), t(cf > 0 ? cf : 1);
    int n_p = 0, n_t = 0;
    parseObjByName(path, p.data(), t.data(), n_p, n_t);
    for (int i = 0; i < n_p; i++) for (int k = 0; k < 3; k++) points[3 * i + k] = p[i][k];
    for (int i = 0; i < n_t; i++) for (int k = 0; k < 3; k++) idx[3 * i + k] = t[i][k];
    *np = n_p;
    *nt = n_t;
    return 0;
}

// Triangle constructor normal (shapes/triangle.h:14-17).
void ref_triangle_normal(const float* v9, float* n3)
{
    vec3 vs[3] = {vec3(v9[0], v9[1], v9[2]), vec3(v9[3], v9[4], v9[5]), vec3(v9[6], v9[7], v9[8])};
    Triangle t(vs, nullptr, false, identity_transform(), false);
    n3[0] = t.normal[0];
    n3[1] = t.normal[1];
    n3[2] = t.normal[2];
}

}  // extern "C"
