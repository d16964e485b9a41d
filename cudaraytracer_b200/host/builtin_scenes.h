// builtin_scenes.h — the scenes of BASELINE.json's configs built with the
// reference-style host classes of scene.h (C1 bunny, C2 random spheres, C3
// displaced-sphere mesh). They produce the same primitives, in the
// same order, as cudaraytracer_b200/scenes.py, which the parity tests use.
#pragma once
#include <cmath>
#include <string>
#include <vector>

#include "../../include/crt_rng.h"
#include "obj_loader.h"
#include "scene.h"

namespace crt_host {

struct SceneOptions {
    std::string obj = "objects/small_bunny.obj", camera = "framed";
    int n_theta = 1000, n_phi = 500;
    uint64_t seed = 1;
};

// Sequential (0,1] draws from the hash of include/crt_rng.h, the same stream
// cudaraytracer_b200/scenes.py uses for its synthetic scenes.
struct Stream {
    uint64_t key; uint32_t ctr = 0;
    Stream(uint64_t seed, uint32_t stream) : key(crt_rng_key(seed, stream, 0xC0FFEEu)) {}
    double draw() { return (double)crt_rng_uniform(key, ctr++); }
};

struct World {
    HitableList list;
    std::vector<Triangle*> triangles;                                 // the reference's d_triangleData
    ObjMesh obj;
    vec3 lookfrom, lookat;
    float vfov = 40.f, aperture = 0.f, focus = 10.f;
};

// C1: the bundled mesh as the reference's (dead) OBJ caller set it up: points
// x100 (obj_loader.h:164), one Lambertian(0.65, 0.05, 0.05) (createScene.h:180).
inline void build_bunny(World& w, const SceneOptions& o)
{
    if (!loadObj(o.obj, w.obj)) throw std::runtime_error("can't open the file " + o.obj);
    for (int i = 0; i < w.obj.mesh.nPoints; i++) w.obj.mesh.points[i] *= 100.0;
    Material* mat = new Lambertian(new ConstantTexture(vec3(0.65, 0.05, 0.05)));
    add_mesh(&w.list, w.obj.mesh, mat, 1, &w.triangles);
    if (o.camera == "reference") { w.lookfrom = vec3(0, 100, 1000); w.lookat = vec3(0, 150, 0); }   // createScene.h:160
    else { w.lookfrom = vec3(0, 0, 30); w.lookat = vec3(0, 0, 0); }
}

// C2: the random-spheres layout the reference's Sphere/material classes were
// written for; same draws as scenes.c2_spheres.
inline void build_spheres(World& w, const SceneOptions& o)
{
    Stream st(o.seed, 2);
    auto place = [&](double x, double y, double z, double r, Material* m) {
        w.list.append(new Sphere(new Transform(vec3((float)x, (float)y, (float)z), vec3(0.f), vec3(1.f)), (float)r, m));
    };
    place(0, -1000, 0, 1000, new Lambertian(new ConstantTexture(vec3(0.5, 0.5, 0.5))));
    for (int a = -11; a < 11; a++)
        for (int b = -11; b < 11; b++) {
            const double choose = st.draw();
            const double cx = a + 0.9 * st.draw();
            const double cz = b + 0.9 * st.draw();
            double r[6];
            for (double& v : r) v = st.draw();
            if (std::sqrt((cx - 4.0) * (cx - 4.0) + cz * cz) <= 0.9) continue;
            Material* m;
            if (choose < 0.8) m = new Lambertian(new ConstantTexture(vec3((float)(r[0] * r[1]), (float)(r[2] * r[3]), (float)(r[4] * r[5]))));
            else if (choose < 0.95) m = new Metal(vec3((float)(0.5 * (1 + r[0])), (float)(0.5 * (1 + r[1])), (float)(0.5 * (1 + r[2]))), (float)(0.5 * r[3]));
            else m = new Dielectric(1.5f);
            place(cx, 0.2, cz, 0.2, m);
        }
    place(0, 1, 0, 1.0, new Dielectric(1.5f));
    place(-4, 1, 0, 1.0, new Lambertian(new ConstantTexture(vec3(0.4, 0.2, 0.1))));
    place(4, 1, 0, 1.0, new Metal(vec3(0.7, 0.6, 0.5), 0.0f));
    w.lookfrom = vec3(13, 2, 3); w.lookat = vec3(0, 0, 0);
    w.vfov = 20.f; w.aperture = 0.1f; w.focus = 10.f;
}

inline vec3 displaced_point(int i, int j, int n_theta, int n_phi, double phase)
{
    const double th = (double)(i % n_theta) * (2.0 * M_PI / n_theta), ph = (double)j * (M_PI / n_phi);
    const double R = 100.0 + 5.0 * std::sin(8.0 * th + phase) * std::sin(6.0 * ph);
    return vec3((float)(R * std::sin(ph) * std::cos(th)), (float)(R * std::cos(ph)), (float)(R * std::sin(ph) * std::sin(th)));
}

// The vertices of the C3 mesh (displaced UV sphere) in the order
// scenes.displaced_sphere emits them: upper triangles (a, b, d) of bands
// 1..n_phi-1, then lower triangles (a, d, c) of bands 0..n_phi-2.
inline void mesh_vertices(int n_theta, int n_phi, double phase, std::vector<float>& verts, std::vector<int>* bands)
{
    verts.clear();
    if (bands) bands->clear();
    auto emit = [&](const vec3& p, const vec3& q, const vec3& r, int band) {
        for (const vec3* v : {&p, &q, &r}) { verts.push_back((*v)[0]); verts.push_back((*v)[1]); verts.push_back((*v)[2]); }
        if (bands) bands->push_back(band);
    };
    for (int j = 1; j < n_phi; j++)
        for (int i = 0; i < n_theta; i++)
            emit(displaced_point(i, j, n_theta, n_phi, phase), displaced_point(i + 1, j, n_theta, n_phi, phase),
                 displaced_point(i + 1, j + 1, n_theta, n_phi, phase), j);
    for (int j = 0; j < n_phi - 1; j++)
        for (int i = 0; i < n_theta; i++)
            emit(displaced_point(i, j, n_theta, n_phi, phase), displaced_point(i + 1, j + 1, n_theta, n_phi, phase),
                 displaced_point(i, j + 1, n_theta, n_phi, phase), j);
}

// The same mesh in the indexed form the reference animates (mesh->points and
// mesh->idxVertex, createScene.h:99-140): (n_phi+1) x (n_theta+1) points, row
// major, and for every triangle of mesh_vertices() the three point numbers in
// stored-vertex order. points[idx] reproduces mesh_vertices() exactly.
inline void mesh_points(int n_theta, int n_phi, double phase, std::vector<float>& points)
{
    points.clear();
    for (int j = 0; j <= n_phi; j++)
        for (int i = 0; i <= n_theta; i++) {
            const vec3 p = displaced_point(i, j, n_theta, n_phi, phase);
            points.push_back(p[0]); points.push_back(p[1]); points.push_back(p[2]);
        }
}

inline void mesh_indices(int n_theta, int n_phi, std::vector<uint32_t>& idx)
{
    idx.clear();
    auto num = [&](int i, int j) { return (uint32_t)(j * (n_theta + 1) + i); };
    for (int j = 1; j < n_phi; j++)
        for (int i = 0; i < n_theta; i++) { idx.push_back(num(i, j)); idx.push_back(num(i + 1, j)); idx.push_back(num(i + 1, j + 1)); }
    for (int j = 0; j < n_phi - 1; j++)
        for (int i = 0; i < n_theta; i++) { idx.push_back(num(i, j)); idx.push_back(num(i + 1, j + 1)); idx.push_back(num(i, j + 1)); }
}

// C3: stored normals point INTO the sphere and the ground's DOWN, because the
// reference triangle test only reports hits with dot(direction, normal) >= 0
// (shapes/triangle.h:61).
inline void build_mesh(World& w, const SceneOptions& o)
{
    std::vector<float> v;
    std::vector<int> bands;
    mesh_vertices(o.n_theta, o.n_phi, 0.0, v, &bands);
    Material* even = new Lambertian(new ConstantTexture(vec3(0.7, 0.3, 0.2)));
    Material* odd = new Metal(vec3(0.8, 0.8, 0.8), 0.1f);
    Material* ground = new Lambertian(new ConstantTexture(vec3(0.5, 0.5, 0.5)));
    for (size_t t = 0; t < bands.size(); t++) {
        vec3 p[3] = {vec3(v[9 * t], v[9 * t + 1], v[9 * t + 2]), vec3(v[9 * t + 3], v[9 * t + 4], v[9 * t + 5]), vec3(v[9 * t + 6], v[9 * t + 7], v[9 * t + 8])};
        vec3 n = -unit_vector(cross(p[1] - p[0], p[2] - p[0]));
        if (!(n[0] == n[0]) || !(n[1] == n[1]) || !(n[2] == n[2])) n = vec3(0, 1, 0);
        Triangle* tri = new Triangle(p, n, (bands[t] & 1) ? odd : even, false, new Transform(), false);
        w.list.append(tri);
        w.triangles.push_back(tri);
    }
    const float h = 2000.f, y = -110.f;
    vec3 q[4] = {vec3(-h, y, -h), vec3(h, y, -h), vec3(h, y, h), vec3(-h, y, h)};
    vec3 t0[3] = {q[0], q[1], q[2]}, t1[3] = {q[0], q[2], q[3]};
    w.list.append(new Triangle(t0, vec3(0, -1, 0), ground, false, new Transform(), false));
    w.list.append(new Triangle(t1, vec3(0, -1, 0), ground, false, new Transform(), false));
    w.lookfrom = vec3(0, 60, 320); w.lookat = vec3(0, 0, 0);
}

// C4: grid x grid instances of a ~100,000-triangle displaced sphere (pitch 250) over a ground quad, flattened: every
// instance's translation is added to its vertices (one fp32 addition each, as scenes.c4_instanced does), so the
// triangles carry identity Transforms and take the plain path of the traversal. The reference's own instancing
// device - one HitableList per instance under a translated Transform (hitable_list.h:11) - works too (scene.h turns
// it into a chain of frames) and gives the reference's local-frame hit points instead.
inline void build_instanced(World& w, const SceneOptions& o, int grid = 10, float pitch = 250.f)
{
    const int n_theta = o.n_theta == 1000 ? 500 : o.n_theta, n_phi = o.n_theta == 1000 ? 101 : o.n_phi;
    std::vector<float> v;
    std::vector<int> bands;
    mesh_vertices(n_theta, n_phi, 0.0, v, &bands);
    Material* even = new Lambertian(new ConstantTexture(vec3(0.7, 0.3, 0.2)));
    Material* odd = new Metal(vec3(0.8, 0.8, 0.8), 0.1f);
    Material* ground = new Lambertian(new ConstantTexture(vec3(0.5, 0.5, 0.5)));
    std::vector<vec3> normals(bands.size());
    for (size_t t = 0; t < bands.size(); t++) {
        const vec3 p0(v[9 * t], v[9 * t + 1], v[9 * t + 2]), p1(v[9 * t + 3], v[9 * t + 4], v[9 * t + 5]), p2(v[9 * t + 6], v[9 * t + 7], v[9 * t + 8]);
        vec3 n = -unit_vector(cross(p1 - p0, p2 - p0));
        if (!(n[0] == n[0]) || !(n[1] == n[1]) || !(n[2] == n[2])) n = vec3(0, 1, 0);
        normals[t] = n;
    }
    const float off0 = (float)(-(grid - 1) * (double)pitch / 2.0);
    for (int a = 0; a < grid; a++)
        for (int b = 0; b < grid; b++) {
            const vec3 shift((float)(off0 + a * (double)pitch), 0.f, (float)(off0 + b * (double)pitch));
            for (size_t t = 0; t < bands.size(); t++) {
                vec3 p[3];
                for (int k = 0; k < 3; k++) p[k] = vec3(v[9 * t + 3 * k], v[9 * t + 3 * k + 1], v[9 * t + 3 * k + 2]) + shift;
                w.list.append(new Triangle(p, normals[t], (bands[t] & 1) ? odd : even, false, new Transform(), false));
            }
        }
    const float h = 20000.f, y = -110.f;
    vec3 q[4] = {vec3(-h, y, -h), vec3(h, y, -h), vec3(h, y, h), vec3(-h, y, h)};
    vec3 t0[3] = {q[0], q[1], q[2]}, t1[3] = {q[0], q[2], q[3]};
    w.list.append(new Triangle(t0, vec3(0, -1, 0), ground, false, new Transform(), false));
    w.list.append(new Triangle(t1, vec3(0, -1, 0), ground, false, new Transform(), false));
    w.lookfrom = vec3(0, 900, 2600); w.lookat = vec3(0, 0, 0);
}

// A Cornell box out of the reference's own parts, the same scene as scenes.cornell_box: five walls
// and a luminous sheet that are Rectangles - the unit square of shapes/rectangle.h turned and placed by Transform
// rotations (scale stays 1, SURVEY 9 Q7, so the box is one unit wide) - the block of objects/cbox_smallbox.obj
// ("f a//n" faces; scaled by 1/555) as Triangles inside two nested HitableLists, a glass sphere under a rotated
// Transform, a CheckerTexture floor. A frame maps world to local: local = rotate(world) - position
// (hitable/transform.h:38-48); a wall with inward normal N needs the rotation that turns N into +z, and its
// position is then (0, 0, -0.5).
inline void build_cornell(World& w, const SceneOptions& o)
{
    Material* white = new Lambertian(new ConstantTexture(vec3(0.73f, 0.73f, 0.73f)));
    Material* red = new Lambertian(new ConstantTexture(vec3(0.65f, 0.05f, 0.05f)));
    Material* green = new Lambertian(new ConstantTexture(vec3(0.12f, 0.45f, 0.15f)));
    Material* light = new DiffuseLight(new ConstantTexture(vec3(1.2f, 1.2f, 1.1f)));
    Material* floor_m = new Lambertian(new CheckerTexture(new ConstantTexture(vec3(0.73f, 0.73f, 0.73f)), new ConstantTexture(vec3(0.2f, 0.2f, 0.6f))));
    Material* glass = new Dielectric(1.5f);
    Material* metal = new Metal(vec3(0.8f, 0.85f, 0.88f), 0.0f);
    auto wall = [&](float z, vec3 rot, Material* m) {
        w.list.append(new Rectangle(m, false, new Transform(vec3(0.f, 0.f, z), rot, vec3(1.f))));
    };
    wall(-0.5f, vec3(0.f, 0.f, 0.f), white);            // back wall, N = +z
    wall(-0.5f, vec3(90.f, 0.f, 0.f), floor_m);         // floor, N = +y
    wall(-0.5f, vec3(-90.f, 0.f, 0.f), white);          // ceiling, N = -y
    wall(-0.5f, vec3(0.f, -90.f, 0.f), red);            // wall at x = -0.5
    wall(-0.5f, vec3(0.f, 90.f, 0.f), green);           // wall at x = +0.5
    wall(-0.49f, vec3(-90.f, 0.f, 0.f), light);         // a luminous sheet just below the ceiling
    const std::string obj = o.obj == "objects/small_bunny.obj" ? std::string("objects/cbox_smallbox.obj") : o.obj;
    if (!loadObj(obj, w.obj)) throw std::runtime_error("can't open the file " + obj);
    HitableList* turned = new HitableList(0);
    turned->transform = new Transform(vec3(0.f), vec3(0.f, 20.f, 0.f), vec3(1.f));
    for (int i = 0; i < w.obj.mesh.nTriangles; i++) {
        const vec3 idx = w.obj.mesh.idxVertex[i];
        const int k[3] = {int(idx[2]) - 1, int(idx[1]) - 1, int(idx[0]) - 1};      // 1-based, winding reversed (createScene.h:183)
        vec3 p[3];
        for (int j = 0; j < 3; j++) {
            if (k[j] < 0 || k[j] >= w.obj.mesh.nPoints) throw std::out_of_range("build_cornell: face index out of range");
            const vec3 q = w.obj.mesh.points[k[j]];
            p[j] = vec3(q[0] / 555.f, q[1] / 555.f, q[2] / 555.f);
        }
        const vec3 n = -unit_vector(cross(p[1] - p[0], p[2] - p[0]));           // stored normals point into the block
        Triangle* t = new Triangle(p, n, metal, false, new Transform(vec3(0.02f, 0.f, -0.03f), vec3(0.f), vec3(1.f)), false);
        turned->append(t);
        w.triangles.push_back(t);
    }
    Hitable* inner[1] = {turned};
    w.list.append(new HitableList(inner, 1, new Transform(vec3(0.17f, -0.351f, 0.1f), vec3(0.f), vec3(1.f))));
    const vec3 rot(0.f, 30.f, 0.f);
    w.list.append(new Sphere(new Transform(rotate(vec3(-0.22f, -0.33f, 0.18f), rot), rot, vec3(1.f)), 0.17f, glass));
    w.lookfrom = vec3(0.f, 0.f, 2.4f); w.lookat = vec3(0.f, 0.f, 0.f);
}

}  // namespace crt_host
