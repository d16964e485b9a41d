// scene.h — host-side scene surface with the reference's class names and
// constructor signatures, plus the flattener that turns a HitableList into the
// SoA arrays of include/crt.h.
//
// The reference builds these objects with device-side `new` inside <<<1,1>>>
// kernels and traverses them with virtual calls on the GPU
// (CudaTest/src/createScene.h:27-38,142-210). Here they are plain host
// DESCRIPTIONS: no hit or scatter code lives in them (that is hand-written
// CUDA behind the C ABI, and there is no CPU render path); the only thing a
// Hitable can do is describe itself to a FlatScene.
//
//   reference class                     file:line                     here
//   Ray                                 core/ray.h:7-23               Ray
//   Camera                              core/camera.h:16-54           Camera (+ to_crt())
//   Transform                           hitable/transform.h:6-57      Transform
//   Hitable / HitRecord                 hitable/hitable.h:11-58       Hitable (flatten instead of hit)
//   HitableList                         hitable/hitable_list.h:6-74   HitableList
//   Sphere                              shapes/sphere.h:6-24          Sphere
//   Triangle                            shapes/triangle.h:5-54        Triangle
//   Rectangle                           shapes/rectangle.h:4-44       Rectangle
//   Texture / ConstantTexture           material/texture.h:6-22       same
//   Material / Lambertian / Metal /     material/material.h:42-161    same (describe() instead of scatter())
//   Dielectric / DiffuseLight
//   MeshData                            shapes/MeshObject.h:8-15      MeshData
#pragma once
#include <cstdint>
#include <cstdio>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/crt.h"
#include "vec3.h"

// ---------------------------------------------------------------- core types
class Ray {                                                          // core/ray.h:7-23
public:
    Ray() : _time(0.f), _origin(vec3(0.f)), _direction(vec3(0.f)) {}
    Ray(const vec3& o, const vec3& d, float t = 0.f) : _time(t), _origin(o), _direction(d) {}
    float time() const { return _time; }
    vec3 origin() const { return _origin; }
    vec3 direction() const { return _direction; }
    vec3 point_at_t(float t) const { return _origin + t * _direction; }
    crt_ray to_crt() const
    {
        crt_ray r = {{_origin[0], _origin[1], _origin[2]}, {_direction[0], _direction[1], _direction[2]}};
        return r;
    }
    float _time;                                                     // member order of core/ray.h:20-22
    vec3 _origin;
    vec3 _direction;
};

class Transform {                                                    // hitable/transform.h:6-36
public:
    Transform() : position(vec3(0.f)), rotation(vec3(0.f)), scale(vec3(1.f)) {}
    Transform(vec3 p, vec3 r, vec3 s) : position(p), rotation(r), scale(s) {}
    void ResetTransform() { position = vec3(0.f); rotation = vec3(0.f); scale = vec3(1.f); }
    bool has_rotation() const { return rotation[0] != 0.f || rotation[1] != 0.f || rotation[2] != 0.f; }
    bool has_position() const { return position[0] != 0.f || position[1] != 0.f || position[2] != 0.f; }
    bool unit_scale() const { return scale[0] == 1.f && scale[1] == 1.f && scale[2] == 1.f; }
    bool is_pure_translation() const { return !has_rotation() && unit_scale(); }
    bool is_identity() const { return is_pure_translation() && !has_position(); }
    vec3 position, rotation, scale;
};

// --------------------------------------------------------------- materials
class Texture {                                                      // material/texture.h:6-9
public:
    virtual ~Texture() {}
    virtual vec3 value(float u, float v, const vec3& p) const = 0;
    virtual bool is_constant() const { return false; }
};

class ConstantTexture : public Texture {                             // material/texture.h:12-22
public:
    ConstantTexture() : color(vec3(0.f)) {}
    ConstantTexture(vec3 c) : color(c) {}
    vec3 value(float, float, const vec3&) const override { return color; }
    bool is_constant() const override { return true; }
    vec3 color;
};

class CheckerTexture : public Texture {                               // material/texture.h:25-42
public:
    CheckerTexture() : odd(nullptr), even(nullptr) {}
    CheckerTexture(Texture* t0, Texture* t1) : odd(t1), even(t0) {}
    vec3 value(float u, float v, const vec3& p) const override
    {
        float sines = std::sin(10 * p.x()) * std::sin(10 * p.y()) * std::sin(10 * p.z());
        return sines < 0 ? odd->value(u, v, p) : even->value(u, v, p);
    }
    Texture* odd;
    Texture* even;
};

class Material {                                                     // material/material.h:42-54
public:
    virtual ~Material() {}
    // The record the GPU scatter kernel switches on (csrc/crt_device.cuh scatter()).
    virtual crt_material describe() const = 0;
};

namespace crt_host {
inline crt_material make_material(uint32_t type, const vec3& a, float param)
{
    crt_material m;
    m.type = type;
    m.albedo[0] = a[0]; m.albedo[1] = a[1]; m.albedo[2] = a[2];
    m.param = param;
    m.texture = CRT_TEX_CONSTANT;
    m.albedo2[0] = m.albedo2[1] = m.albedo2[2] = 0.f;
    return m;
}
inline vec3 constant_colour(const Texture* t, const char* who)
{
    if (!t || !t->is_constant())
        throw std::invalid_argument(std::string(who) + ": only ConstantTexture, or a CheckerTexture of two of them, is supported on the render path (material/texture.h:12-42)");
    return t->value(0.f, 0.f, vec3(0.f));
}
// The record of a material whose texture is a ConstantTexture or a CheckerTexture of two ConstantTextures.
inline crt_material textured_material(uint32_t type, const Texture* t, const char* who)
{
    if (const CheckerTexture* c = dynamic_cast<const CheckerTexture*>(t)) {
        crt_material m = make_material(type, constant_colour(c->even, who), 0.f);
        const vec3 o = constant_colour(c->odd, who);
        m.texture = CRT_TEX_CHECKER;
        m.albedo2[0] = o[0]; m.albedo2[1] = o[1]; m.albedo2[2] = o[2];
        return m;
    }
    return make_material(type, constant_colour(t, who), 0.f);
}
}  // namespace crt_host

class Lambertian : public Material {                                 // material/material.h:57-74
public:
    Lambertian(Texture* a) : albedo(a) {}
    crt_material describe() const override
    {
        return crt_host::textured_material(CRT_MAT_LAMBERTIAN, albedo, "Lambertian");
    }
    Texture* albedo;
};

class Metal : public Material {                                      // material/material.h:77-96
public:
    Metal(const vec3& a, float f) : albedo(a) { fuzz = f < 1 ? f : 1; }
    crt_material describe() const override { return crt_host::make_material(CRT_MAT_METAL, albedo, fuzz); }
    vec3 albedo;
    float fuzz;
};

class Dielectric : public Material {                                 // material/material.h:99-143
public:
    Dielectric(float ri) : ref_idx(ri) {}
    crt_material describe() const override { return crt_host::make_material(CRT_MAT_DIELECTRIC, vec3(1.f), ref_idx); }
    float ref_idx;
};

class DiffuseLight : public Material {                               // material/material.h:146-161
public:
    DiffuseLight(Texture* texture) : emit(texture) {}
    crt_material describe() const override
    {
        return crt_host::textured_material(CRT_MAT_DIFFUSE_LIGHT, emit, "DiffuseLight");
    }
    Texture* emit;
};

// ------------------------------------------------------------- flat scene
// What crt_scene_set_materials / _triangles / _spheres take. Primitive ids are
// positions in a depth-first walk of the list, i.e. the order in which the
// reference's HitableList scan would visit them (first wins ties,
// hitable/hitable_list.h:84).
//
// Transforms. Every Hitable::hit maps the ray into the object's frame (hitable/hitable.h:29-38,
// hitable/transform.h:11-14): direction re-normalised, origin and direction rotated, position subtracted - once per
// Hitable on the way from the world list down to the primitive - and reports hit point and normal in the innermost
// frame. A primitive is PLAIN when that walk is the world list under the identity and then the
// primitive itself without rotation (a triangle also without position; a sphere's position is its centre): the
// arrays then hold world-space values and the kernels take their fast path. Anything else - a rotation, a
// translated triangle, a list inside a list - keeps the reference's arithmetic through a CHAIN: the primitive's run
// of crt_frame records, outermost first (crt_scene_set_frames; include/crt.h), with its vertices left in its own
// frame. Chains of primitives that sit in the same nest of lists share the frames of the enclosing levels.
struct FlatScene {
    std::vector<float> tri;          // 12 floats per triangle: v0, v1, v2, stored normal
    std::vector<uint32_t> tri_mat, tri_id, tri_chain;
    std::vector<float> sph;          // 4 floats per sphere: centre, radius
    std::vector<uint32_t> sph_mat, sph_id, sph_chain;
    std::vector<uint32_t> rect_chain, rect_mat, rect_id;   // shapes/rectangle.h: the unit square of a frame
    std::vector<uint8_t> rect_flip;
    std::vector<crt_frame> frames;
    std::vector<crt_material> mats;
    std::map<const Material*, uint32_t> mat_index;
    uint32_t next_id = 0;

    size_t n_triangles() const { return tri_mat.size(); }
    size_t n_spheres() const { return sph_mat.size(); }
    size_t n_rectangles() const { return rect_mat.size(); }
    bool has_chains() const { return !frames.empty(); }

    // The chain word of a primitive whose ray passes the Transforms `levels` (world list first, its own last):
    // (first << 8) | count. Each primitive gets its own run of frames (a run must be contiguous).
    uint32_t add_chain(const std::vector<const Transform*>& levels)
    {
        if (levels.size() > 255) throw std::invalid_argument("flatten: more than 255 nested Hitables");
        const uint32_t first = (uint32_t)frames.size();
        for (const Transform* t : levels) {
            crt_frame f;
            const bool rot = t->has_rotation();
            // the rows rotate() builds (core/vec3.h:211-213): rotate(e_j)[k] is entry j of row k
            const vec3 c0 = rot ? rotate(vec3(1.f, 0.f, 0.f), t->rotation) : vec3(1.f, 0.f, 0.f);
            const vec3 c1 = rot ? rotate(vec3(0.f, 1.f, 0.f), t->rotation) : vec3(0.f, 1.f, 0.f);
            const vec3 c2 = rot ? rotate(vec3(0.f, 0.f, 1.f), t->rotation) : vec3(0.f, 0.f, 1.f);
            for (int k = 0; k < 3; k++) { f.rows[3 * k] = c0[k]; f.rows[3 * k + 1] = c1[k]; f.rows[3 * k + 2] = c2[k]; }
            f.position[0] = t->position[0]; f.position[1] = t->position[1]; f.position[2] = t->position[2];
            f.has_rotation = rot ? 1u : 0u;
            frames.push_back(f);
        }
        if (first + levels.size() >= (1u << 24)) throw std::invalid_argument("flatten: more than 2^24 frames");
        return (first << 8) | (uint32_t)levels.size();
    }

    uint32_t material(const Material* m)
    {
        if (!m) throw std::invalid_argument("flatten: primitive without a material");
        auto it = mat_index.find(m);
        if (it != mat_index.end()) return it->second;
        const uint32_t id = (uint32_t)mats.size();
        mats.push_back(m->describe());
        mat_index[m] = id;
        return id;
    }
};

// ---------------------------------------------------------------- hitables
class Hitable {                                                      // hitable/hitable.h:22-58
public:
    Hitable(Transform* t) : transform(t) {}
    Hitable() { transform = new Transform(); }                       // hitable.h:27 allocates too
    virtual ~Hitable() {}

    // Replaces hit()/collision_detection(): append this object's primitives to `out`. `above` holds the Transforms of
    // the Hitables the ray passed on its way here, the world list's first.
    virtual void flatten(FlatScene& out, std::vector<const Transform*>& above) const = 0;

    Transform* transform;

protected:
    const Transform* own_transform(const char* who) const
    {
        if (!transform) throw std::invalid_argument(std::string(who) + ": null Transform");
        if (!transform->unit_scale())
            throw std::invalid_argument(std::string(who) + ": scale must be 1; the reference's ScaleRay divides only the direction "
                                        "and does not scale the origin (hitable/transform.h:50-54; SURVEY 9 Q7)");
        return transform;
    }
    // plain = the world list under the identity, then this primitive without rotation (and, for a triangle, without
    // position): nothing but the sphere's centre is needed to reproduce the reference's arithmetic
    static bool is_plain(const std::vector<const Transform*>& above, const Transform* own, bool allow_position)
    {
        return above.size() == 1 && above[0]->is_identity() && !own->has_rotation() && (allow_position || !own->has_position());
    }
};

class HitableList : public Hitable {                                 // hitable/hitable_list.h:6-74
public:
    HitableList() : list(nullptr), list_size(0) {}
    HitableList(Hitable** l, int n) : list(nullptr), list_size(0) { assign(l, n); }
    HitableList(int n) : list((Hitable**)std::calloc((size_t)(n > 0 ? n : 1), sizeof(Hitable*))), list_size(n), cap_(n > 0 ? n : 1) {}
    HitableList(Hitable** l, int n, Transform* t) : Hitable(t), list(nullptr), list_size(0) { assign(l, n); }
    ~HitableList() override { std::free(list); }
    HitableList(const HitableList&) = delete;
    HitableList& operator=(const HitableList&) = delete;

    void append(Hitable* data)                                       // hitable_list.h:17-39 (O(n) there, amortised here)
    {
        if (list_size + 1 > cap_) {
            cap_ = cap_ < 16 ? 16 : cap_ * 2;
            list = (Hitable**)std::realloc(list, sizeof(Hitable*) * (size_t)cap_);
        }
        list[list_size++] = data;
    }
    void Reseize(int n)                                              // hitable_list.h:41-64 (sic)
    {
        cap_ = n > 0 ? n : 1;
        list = (Hitable**)std::realloc(list, sizeof(Hitable*) * (size_t)cap_);
        for (int i = list_size; i < n; i++) list[i] = nullptr;
        list_size = n;
    }
    void freeMemory() { std::free(list); list = nullptr; list_size = 0; cap_ = 0; }

    void flatten(FlatScene& out, std::vector<const Transform*>& above) const override
    {
        above.push_back(own_transform("HitableList"));
        for (int i = 0; i < list_size; i++) {
            if (!list[i]) throw std::invalid_argument("HitableList: null entry");
            list[i]->flatten(out, above);
        }
        above.pop_back();
    }

    Hitable** list;
    int list_size;

private:
    int cap_ = 0;
    void assign(Hitable** l, int n)
    {
        cap_ = n > 0 ? n : 1;
        list = (Hitable**)std::malloc(sizeof(Hitable*) * (size_t)cap_);
        for (int i = 0; i < n; i++) list[i] = l[i];
        list_size = n;
    }
};

class Sphere : public Hitable {                                      // shapes/sphere.h:6-24
public:
    Sphere(Transform* t) : Hitable(t), radius(0.f), material(nullptr) {}
    Sphere(Transform* t, float r, Material* mat) : Hitable(t), radius(r), material(mat) {}
    void flatten(FlatScene& out, std::vector<const Transform*>& above) const override
    {
        // The sphere sits at its frame's origin (sphere.h:31); the hit point and normal stay in that frame
        // (SURVEY 9 Q6). Plain: the kernel reproduces that from centre = the sphere's own position.
        const Transform* own = own_transform("Sphere");
        if (is_plain(above, own, true)) {
            out.sph.insert(out.sph.end(), {own->position[0], own->position[1], own->position[2], radius});
            out.sph_chain.push_back(0u);
        } else {
            above.push_back(own);
            out.sph.insert(out.sph.end(), {0.f, 0.f, 0.f, radius});
            out.sph_chain.push_back(out.add_chain(above));
            above.pop_back();
        }
        out.sph_mat.push_back(out.material(material));
        out.sph_id.push_back(out.next_id++);
    }
    float radius;
    Material* material;
};

class Triangle : public Hitable {                                    // shapes/triangle.h:5-54
public:
    Triangle() : EPSILON(0.00001f), flipNormal(false), backCulling(false), material(nullptr) {}
    Triangle(vec3 vs[3], Material* mat, bool flip, Transform* t, const bool cull = false)
        : Hitable(t), EPSILON(0.000001f), flipNormal(flip), backCulling(cull), material(mat)
    {
        for (int i = 0; i < 3; i++) vertices[i] = vs[i];
        normal = unit_vector(cross(vertices[1] - vertices[0], vertices[2] - vertices[0]));   // triangle.h:14-17
    }
    Triangle(vec3 vs[3], vec3 triNormal, Material* mat, bool flip, Transform* t, const bool cull = false)
        : Hitable(t), EPSILON(0.000001f), normal(triNormal), flipNormal(flip), backCulling(cull), material(mat)
    {
        for (int i = 0; i < 3; i++) vertices[i] = vs[i];
    }
    void SetVertices(vec3 vs[3]) { for (int i = 0; i < 3; i++) vertices[i] = vs[i]; }   // triangle.h:40-45 (normal stays)

    void flatten(FlatScene& out, std::vector<const Transform*>& above) const override
    {
        if (backCulling) throw std::invalid_argument("Triangle: backCulling is never set by live reference code and is not supported");
        // Mesh triangles carry identity transforms (createScene.h:184): plain. A triangle that is translated, rotated
        // or nested keeps its vertices in its own frame and gets a chain, so that hit point and normal come out in
        // that frame as the reference reports them.
        const Transform* own = own_transform("Triangle");
        for (int i = 0; i < 3; i++) out.tri.insert(out.tri.end(), {vertices[i][0], vertices[i][1], vertices[i][2]});
        out.tri.insert(out.tri.end(), {normal[0], normal[1], normal[2]});
        if (is_plain(above, own, false)) {
            out.tri_chain.push_back(0u);
        } else {
            above.push_back(own);
            out.tri_chain.push_back(out.add_chain(above));
            above.pop_back();
        }
        out.tri_mat.push_back(out.material(material));
        out.tri_id.push_back(out.next_id++);
    }

    const float EPSILON;
    vec3 vertices[3];
    vec3 normal;
    bool flipNormal;
    bool backCulling;
    Material* material;
};

class Rectangle : public Hitable {                                   // shapes/rectangle.h:4-19
public:
    Rectangle() : mat_ptr(nullptr), flipNormal(false) {}
    Rectangle(Material* mat, bool flip) : mat_ptr(mat), flipNormal(flip) {}
    Rectangle(Material* mat, bool flip, Transform* t) : Hitable(t), mat_ptr(mat), flipNormal(flip) {}
    void flatten(FlatScene& out, std::vector<const Transform*>& above) const override
    {
        // the unit square of its frame's z = 0 plane (rectangle.h:28-31): it only becomes a wall through its
        // Transform, so it always has a chain
        above.push_back(own_transform("Rectangle"));
        out.rect_chain.push_back(out.add_chain(above));
        above.pop_back();
        out.rect_flip.push_back(flipNormal ? 1 : 0);
        out.rect_mat.push_back(out.material(mat_ptr));
        out.rect_id.push_back(out.next_id++);
    }
    Material* mat_ptr;
    bool flipNormal;
};

inline FlatScene flatten_scene(const Hitable& world)
{
    FlatScene f;
    std::vector<const Transform*> above;
    world.flatten(f, above);
    return f;
}

// ------------------------------------------------------------------ camera
class Camera {                                                       // core/camera.h:16-54
public:
    Camera(vec3 lookfrom, vec3 lookat, vec3 vup, float vfov, float aspect, float aperture, float focus_dist)
    {
        lens_radius = aperture / 2.0f;
        float theta = vfov * M_PI / 180.0f;                          // double product, rounded to float (camera.h:26)
        float half_height = std::tan(theta / 2.0f);
        float half_width = half_height * aspect;
        origin = lookfrom;
        z = unit_vector(lookfrom - lookat);
        x = unit_vector(cross(vup, z));
        y = cross(z, x);
        lower_left_corner = origin - half_width * focus_dist * x - half_height * focus_dist * y - focus_dist * z;
        horizontal = 2.0f * half_width * focus_dist * x;
        vertical = 2.0f * half_height * focus_dist * y;
    }
    // get_ray (camera.h:40-44) runs in k_raygen on the GPU; the host passes the
    // constructed members by value so both sides share one tan().
    crt_camera to_crt() const
    {
        crt_camera c;
        for (int i = 0; i < 3; i++) {
            c.lower_left_corner[i] = lower_left_corner[i];
            c.horizontal[i] = horizontal[i];
            c.vertical[i] = vertical[i];
            c.origin[i] = origin[i];
            c.x[i] = x[i]; c.y[i] = y[i]; c.z[i] = z[i];
        }
        c.lens_radius = lens_radius;
        return c;
    }
    vec3 lower_left_corner, horizontal, vertical, origin;
    vec3 x, y, z;
    float lens_radius;
};

// -------------------------------------------------------------------- mesh
class MeshData {                                                     // shapes/MeshObject.h:8-15
public:
    int nPoints = 0;
    int nTriangles = 0;
    vec3* points = nullptr;
    vec3* idxVertex = nullptr;
    vec3* normals = nullptr;
};

// add_mesh_withNormal's loop (createScene.h:175-190) as a host function: one
// shared material, an identity Transform per triangle, winding reversed to
// idx[2], idx[1], idx[0] (createScene.h:183). index_base is 0 for FBX data
// (Loader/FbxLoader.h:54-57) and 1 for parseObjByName's 1-based floats
// (Loader/obj_loader.h:62-64). With normals == nullptr the
// Triangle constructor computes them (triangle.h:14-17).
// `triangles`, when given, receives the Triangle* of every face like the
// reference's d_triangleData.
inline void add_mesh(HitableList* list, const MeshData& mesh, Material* mat, int index_base = 0,
                     std::vector<Triangle*>* triangles = nullptr)
{
    for (int i = 0; i < mesh.nTriangles; i++) {
        const vec3 idx = mesh.idxVertex[i];
        int k[3] = {int(idx[2]) - index_base, int(idx[1]) - index_base, int(idx[0]) - index_base};
        for (int j = 0; j < 3; j++)
            if (k[j] < 0 || k[j] >= mesh.nPoints) {
                char msg[128];
                std::snprintf(msg, sizeof msg, "add_mesh: face %d references vertex %d of %d", i, k[j], mesh.nPoints);
                throw std::out_of_range(msg);
            }
        vec3 v[3] = {mesh.points[k[0]], mesh.points[k[1]], mesh.points[k[2]]};
        Transform* t = new Transform(vec3(0.f), vec3(0.f), vec3(1.f));
        Triangle* tri = mesh.normals ? new Triangle(v, mesh.normals[i], mat, false, t, false) : new Triangle(v, mat, false, t, false);
        list->append(tri);
        if (triangles) triangles->push_back(tri);
    }
}
