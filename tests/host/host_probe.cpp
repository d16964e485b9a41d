// host_probe.cpp — TEST INFRASTRUCTURE. Exercises the header-only host scene
// surface (cudaraytracer_b200/host) without a GPU and dumps what it produces
// so tests/test_host_cpp.py can compare it with scenes.py and the goldens.
//   host_probe flat   <bunny|spheres|mesh|nested> <out.bin> [obj] [segments]
//   host_probe camera <nx> <ny> <lfx lfy lfz> <lax lay laz> <vfov> <aperture> <focus> <out.bin>
//   host_probe obj    <file.obj> <out.bin>
//   host_probe png    <w> <h> <comp> <in.raw> <out.png>
//   host_probe writepng <nx> <ny> <in_rgb_f32.raw> <prefix>
//   host_probe reject <scale|texture|nomat>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../cudaraytracer_b200/host/builtin_scenes.h"
#include "../../cudaraytracer_b200/host/obj_loader.h"
#include "../../cudaraytracer_b200/host/png_out.h"
#include "../../cudaraytracer_b200/host/scene.h"

static void put(FILE* f, const void* p, size_t n) { if (n && std::fwrite(p, 1, n, f) != n) std::exit(3); }

static std::vector<uint8_t> slurp(const char* path)
{
    FILE* f = std::fopen(path, "rb");
    if (!f) { std::perror(path); std::exit(2); }
    std::vector<uint8_t> v;
    uint8_t buf[65536];
    size_t n;
    while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) v.insert(v.end(), buf, buf + n);
    std::fclose(f);
    return v;
}

static void dump_flat(const FlatScene& s, const char* path)
{
    FILE* f = std::fopen(path, "wb");
    if (!f) { std::perror(path); std::exit(2); }
    const uint64_t hdr[5] = {s.n_triangles(), s.n_spheres(), s.mats.size(), s.n_rectangles(), s.frames.size()};
    put(f, hdr, sizeof hdr);
    put(f, s.tri.data(), s.tri.size() * 4);
    put(f, s.tri_mat.data(), s.tri_mat.size() * 4);
    put(f, s.tri_id.data(), s.tri_id.size() * 4);
    put(f, s.sph.data(), s.sph.size() * 4);
    put(f, s.sph_mat.data(), s.sph_mat.size() * 4);
    put(f, s.sph_id.data(), s.sph_id.size() * 4);
    put(f, s.mats.data(), s.mats.size() * sizeof(crt_material));
    put(f, s.tri_chain.data(), s.tri_chain.size() * 4);
    put(f, s.sph_chain.data(), s.sph_chain.size() * 4);
    put(f, s.rect_chain.data(), s.rect_chain.size() * 4);
    put(f, s.rect_mat.data(), s.rect_mat.size() * 4);
    put(f, s.rect_id.data(), s.rect_id.size() * 4);
    put(f, s.rect_flip.data(), s.rect_flip.size());
    put(f, s.frames.data(), s.frames.size() * sizeof(crt_frame));
    std::fclose(f);
}

int main(int argc, char** argv)
{
    if (argc < 2) return 64;
    const std::string cmd = argv[1];
    try {
        if (cmd == "flat" && argc >= 4) {
            crt_host::World w;
            crt_host::SceneOptions o;
            const std::string which = argv[2];
            if (argc > 4) o.obj = argv[4];
            if (argc > 5) { o.n_theta = std::atoi(argv[5]); o.n_phi = o.n_theta / 2; }
            if (which == "bunny") crt_host::build_bunny(w, o);
            else if (which == "spheres") crt_host::build_spheres(w, o);
            else if (which == "mesh") crt_host::build_mesh(w, o);
            else if (which == "cornell") crt_host::build_cornell(w, o);
            else if (which == "nested") {
                // a list inside a translated list (hitable_list.h:11): offsets add up
                Material* m = new Metal(vec3(0.8f, 0.6f, 0.2f), 3.0f);          // fuzz clamps to 1
                HitableList* inner = new HitableList();
                inner->append(new Sphere(new Transform(vec3(1, 2, 3), vec3(0.f), vec3(1.f)), 0.5f, m));
                vec3 vs[3] = {vec3(0, 0, 0), vec3(1, 0, 0), vec3(0, 1, 0)};
                inner->append(new Triangle(vs, m, false, new Transform(), false));
                Hitable* arr[1] = {inner};
                HitableList* outer = new HitableList(arr, 1, new Transform(vec3(10, 20, 30), vec3(0.f), vec3(1.f)));
                w.list.append(outer);
                w.list.append(new Sphere(new Transform(), 2.0f, new Dielectric(1.5f)));
                // CheckerTexture(t0 = even, t1 = odd), texture.h:28, under a Lambertian and under a light
                Texture* chk = new CheckerTexture(new ConstantTexture(vec3(0.2f, 0.3f, 0.1f)), new ConstantTexture(vec3(0.9f, 0.9f, 0.9f)));
                w.list.append(new Sphere(new Transform(vec3(0, -1000, 0), vec3(0.f), vec3(1.f)), 1000.0f, new Lambertian(chk)));
                w.list.append(new Sphere(new Transform(vec3(0, 9, 0), vec3(0.f), vec3(1.f)), 1.0f, new DiffuseLight(chk)));
            } else return 64;
            dump_flat(flatten_scene(w.list), argv[3]);
            return 0;
        }
        if (cmd == "camera" && argc == 14) {
            const int nx = std::atoi(argv[2]), ny = std::atoi(argv[3]);
            Camera c(vec3((float)std::atof(argv[4]), (float)std::atof(argv[5]), (float)std::atof(argv[6])),
                     vec3((float)std::atof(argv[7]), (float)std::atof(argv[8]), (float)std::atof(argv[9])), vec3(0, 1, 0),
                     (float)std::atof(argv[10]), float(nx) / float(ny), (float)std::atof(argv[11]), (float)std::atof(argv[12]));
            const crt_camera cc = c.to_crt();
            FILE* f = std::fopen(argv[13], "wb");
            put(f, &cc, sizeof cc);
            std::fclose(f);
            return 0;
        }
        if (cmd == "obj" && argc == 4) {
            ObjMesh m;
            if (!loadObj(argv[2], m)) return 2;
            FILE* f = std::fopen(argv[3], "wb");
            const uint64_t hdr[2] = {(uint64_t)m.mesh.nPoints, (uint64_t)m.mesh.nTriangles};
            put(f, hdr, sizeof hdr);
            put(f, m.mesh.points, (size_t)m.mesh.nPoints * 12);
            put(f, m.mesh.idxVertex, (size_t)m.mesh.nTriangles * 12);
            std::fclose(f);
            return 0;
        }
        if (cmd == "png" && argc == 7) {
            const int w = std::atoi(argv[2]), h = std::atoi(argv[3]), comp = std::atoi(argv[4]);
            const std::vector<uint8_t> raw = slurp(argv[5]);
            if (raw.size() != (size_t)w * h * comp) return 65;
            return stbi_write_png(argv[6], w, h, comp, raw.data(), 0) ? 0 : 1;
        }
        if (cmd == "writepng" && argc == 6) {
            const int nx = std::atoi(argv[2]), ny = std::atoi(argv[3]);
            const std::vector<uint8_t> raw = slurp(argv[4]);
            if (raw.size() != (size_t)nx * ny * 12) return 65;
            return WritePng(nx, ny, 7, reinterpret_cast<const vec3*>(raw.data()), argv[5]) ? 0 : 1;
        }
        if (cmd == "indexed" && argc == 5) {
            // host_probe indexed SEGMENTS PHASE OUT: the indexed form of the C3 mesh (mesh_points + mesh_indices) must
            // reproduce mesh_vertices() bit for bit; OUT gets {n_points, n_tri, points, idx}
            const int nt = std::atoi(argv[2]), np = nt / 2;
            const double phase = std::atof(argv[3]);
            std::vector<float> verts, points;
            std::vector<uint32_t> idx;
            crt_host::mesh_vertices(nt, np, phase, verts, nullptr);
            crt_host::mesh_points(nt, np, phase, points);
            crt_host::mesh_indices(nt, np, idx);
            if (idx.size() * 3 != verts.size()) return 2;
            for (size_t i = 0; i < idx.size(); i++)
                for (int k = 0; k < 3; k++)
                    if (std::memcmp(&points[(size_t)idx[i] * 3 + k], &verts[i * 3 + k], 4) != 0) return 3;
            FILE* f = std::fopen(argv[4], "wb");
            if (!f) return 1;
            const uint32_t hdr[2] = {(uint32_t)(points.size() / 3), (uint32_t)(idx.size() / 3)};
            put(f, hdr, sizeof hdr);
            put(f, points.data(), points.size() * 4);
            put(f, idx.data(), idx.size() * 4);
            std::fclose(f);
            return 0;
        }
        if (cmd == "reject" && argc == 3) {
            const std::string what = argv[2];
            HitableList l;
            Material* m = new Lambertian(new ConstantTexture(vec3(0.5f)));
            if (what == "scale") l.append(new Sphere(new Transform(vec3(0.f), vec3(0.f), vec3(2.f)), 1.f, m));
            else if (what == "nomat") l.append(new Sphere(new Transform(), 1.f, nullptr));
            else if (what == "texture") {
                struct Odd : Texture { vec3 value(float, float, const vec3& p) const override { return p; } };
                l.append(new Sphere(new Transform(), 1.f, new Lambertian(new Odd())));
            } else return 64;
            flatten_scene(l);
            return 0;   // the tests expect a throw, i.e. exit code 9
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "host_probe: %s\n", e.what());
        return 9;
    }
    return 64;
}
