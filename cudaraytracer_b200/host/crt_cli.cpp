// crt_cli.cpp — host driver of the render path: the role of the reference's
// main() (CudaTest/src/kernel.cu:41-111), with its hard-coded literals turned
// into flags ("Config / flags").
//
//   crt_cli --scene bunny --obj objects/small_bunny.obj --out out/picture_
//   crt_cli --scene spheres --nx 1200 --ny 800 --spp 64 --mode path
//   crt_cli --scene mesh --frames 4 --animate --csv output.csv
//   crt_cli --scene instanced --gpus 8 --nx 3840 --ny 2160 --spp 64 --mode path      (one host thread per GPU)
//
// Per frame it does what renderBVHAnimation does (core/render.h:191-237):
// update the vertices, refit the BVH, render, copy back, write
// "<out><frame>.png", and it records the reference's three timings
// (build once; update and rendering per frame) into a CSV with the reference's
// header "frame,rendering,update,build" (kernel.cu:61).
// Defaults are the reference's: 1024x512, 4 spp, depth 8, LambertShade
// (kernel.cu:44-49, render.h:120).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "builtin_scenes.h"
#include "obj_loader.h"
#include "png_out.h"
#include "renderer.h"
#include "scene.h"

namespace {

struct Options {
    std::string scene = "bunny", out = "picture_", csv, mode = "lambert";
    crt_host::SceneOptions sc;
    int nx = 1024, ny = 512, spp = 4, depth = 8, frames = 1, device = 0, gpus = 1;
    bool animate = false, quiet = false, overlap = false;
    std::string pose = "vertices";   // --pose vertices|points|bones: the form the per-frame geometry update takes
};

double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

[[noreturn]] void die(const std::string& msg)
{
    std::fprintf(stderr, "crt_cli: %s\n", msg.c_str());               // fail fast like deviceManage.h:3-17
    std::exit(1);
}

Options parse(int argc, char** argv)
{
    Options o;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto val = [&]() -> const char* { if (i + 1 >= argc) die("missing value after " + a); return argv[++i]; };
        if (a == "--scene") o.scene = val();
        else if (a == "--obj") o.sc.obj = val();
        else if (a == "--out") o.out = val();
        else if (a == "--csv") o.csv = val();
        else if (a == "--mode") o.mode = val();
        else if (a == "--camera") o.sc.camera = val();
        else if (a == "--nx") o.nx = std::atoi(val());
        else if (a == "--ny") o.ny = std::atoi(val());
        else if (a == "--spp") o.spp = std::atoi(val());
        else if (a == "--depth") o.depth = std::atoi(val());
        else if (a == "--frames") o.frames = std::atoi(val());
        else if (a == "--device") o.device = std::atoi(val());
        else if (a == "--gpus") o.gpus = std::atoi(val());
        else if (a == "--segments") { o.sc.n_theta = std::atoi(val()); o.sc.n_phi = o.sc.n_theta / 2; }
        else if (a == "--seed") o.sc.seed = std::strtoull(val(), nullptr, 10);
        else if (a == "--animate") o.animate = true;
        else if (a == "--overlap") o.overlap = true;
        else if (a == "--pose") o.pose = val();
        else if (a == "--quiet") o.quiet = true;
        else if (a == "--help" || a == "-h") {
            std::printf("usage: crt_cli [--scene bunny|spheres|mesh|instanced|cornell] [--obj FILE] [--camera framed|reference] [--nx N --ny N]\n"
                        "               [--spp N] [--depth N] [--mode path|lambert|normal] [--seed S] [--frames N]\n"
                        "               [--animate [--overlap] [--pose vertices|points|bones]]\n"
                        "               [--segments N] [--out PREFIX] [--csv FILE] [--device D] [--gpus N] [--quiet]\n");
            std::exit(0);
        } else die("unknown flag " + a);
    }
    if (o.nx <= 0 || o.ny <= 0 || o.spp <= 0 || o.frames <= 0 || o.gpus <= 0) die("nx, ny, spp, frames and gpus must be positive");
    return o;
}

void build_world(crt_host::World& w, const Options& o)
{
    if (o.scene == "bunny") crt_host::build_bunny(w, o.sc);
    else if (o.scene == "spheres") crt_host::build_spheres(w, o.sc);
    else if (o.scene == "mesh") crt_host::build_mesh(w, o.sc);
    else if (o.scene == "instanced") crt_host::build_instanced(w, o.sc);
    else if (o.scene == "cornell") crt_host::build_cornell(w, o.sc);
    else die("unknown --scene " + o.scene);
}

// --gpus N: one host thread per GPU (devices device .. device + N - 1), each with its own context; thread 0 plays
// main(): it builds the scene, the others receive it over NVLink (crt_scene_broadcast), every GPU builds the same
// LBVH and renders the tiles it owns of each frame, and the partial framebuffers are summed onto GPU 0
// (crt_render_dist), which writes the PNG. With --animate every thread computes the frame's pose itself.
int run_multi(const Options& o, int mode)
{
    const int world = o.gpus;
    unsigned char id[CRT_COMM_ID_BYTES];
    crt_host::check(crt_comm_unique_id(id), "crt_comm_unique_id");
    std::vector<std::string> errors((size_t)world);
    std::vector<crt_host::FrameTiming> rows;
    auto rank_main = [&](int rank) {
        try {
            crt_host::Renderer r(o.device + rank);
            crt_host::World w;
            build_world(w, o);                                        // the camera literals; only rank 0 uploads the geometry
            if (rank == 0) r.set_world(w.list);
            r.comm_init(id, rank, world);
            r.scene_broadcast(0);
            double t0 = now_s();
            r.create_BVHfromList();
            const double build_s = now_s() - t0;
            Camera cam(w.lookfrom, w.lookat, vec3(0, 1, 0), w.vfov, float(o.nx) / float(o.ny), w.aperture, w.focus);
            std::vector<vec3> colorBuffer(rank == 0 ? (size_t)o.nx * o.ny : 0);
            std::vector<float> verts;
            const bool moving = o.animate && o.scene == "mesh";
            for (int frame = 0; frame < o.frames; frame++) {
                double update_s = 0.0;
                if (moving) {
                    t0 = now_s();
                    crt_host::mesh_vertices(o.sc.n_theta, o.sc.n_phi, 0.3 * frame, verts, nullptr);
                    r.update_vertices(verts.data(), 0, verts.size() / 9);
                    r.Update_BVH();
                    update_s = now_s() - t0;
                }
                t0 = now_s();
                r.render_dist(rank == 0 ? colorBuffer.data() : nullptr, cam, o.nx, o.ny, o.spp, o.depth, mode, o.sc.seed + (uint64_t)frame);
                const double render_s = now_s() - t0;
                if (rank == 0) {
                    if (!WritePng(o.nx, o.ny, frame, colorBuffer.data(), o.out.c_str())) throw std::runtime_error("cannot write " + o.out + std::to_string(frame) + ".png");
                    rows.push_back({frame, render_s, update_s, frame == 0 ? build_s : 0.0});
                    if (!o.quiet) std::printf("frame %d on %d GPUs: rendering %g s, update %g s\n", frame, world, render_s, update_s);
                }
            }
        } catch (const std::exception& e) {
            errors[(size_t)rank] = e.what();
        }
    };
    std::vector<std::thread> threads;
    for (int rank = 1; rank < world; rank++) threads.emplace_back(rank_main, rank);
    rank_main(0);
    for (auto& t : threads) t.join();
    for (int rank = 0; rank < world; rank++)
        if (!errors[(size_t)rank].empty()) die("GPU " + std::to_string(o.device + rank) + ": " + errors[(size_t)rank]);
    if (!o.csv.empty() && !crt_host::write_csv(o.csv, rows)) die("cannot write " + o.csv);
    return 0;
}

}  // namespace

int main(int argc, char** argv)
{
    const Options o = parse(argc, argv);
    int mode = CRT_SHADE_LAMBERT;
    if (o.mode == "path") mode = CRT_SHADE_PATH;
    else if (o.mode == "normal") mode = CRT_SHADE_NORMAL;
    else if (o.mode != "lambert") die("unknown --mode " + o.mode);
    if (o.gpus > 1) {
        try { return run_multi(o, mode); } catch (const std::exception& e) { die(e.what()); }
    }
    try {
        crt_host::World w;
        build_world(w, o);

        crt_host::Renderer r(o.device);
        r.set_world(w.list);
        if (!o.quiet) std::printf("scene %s: %zu triangles, %zu spheres, %zu materials\n", o.scene.c_str(), r.flat().n_triangles(), r.flat().n_spheres(), r.flat().mats.size());
        double t0 = now_s();
        r.create_BVHfromList();                                       // kernel.cu:29-37 ("build")
        const double build_s = now_s() - t0;
        if (!o.quiet) std::printf("BVH build: %g s (device %g ms)\n", build_s, r.counters().build_ms);

        Camera cam(w.lookfrom, w.lookat, vec3(0, 1, 0), w.vfov, float(o.nx) / float(o.ny), w.aperture, w.focus);   // createScene.h:142-154
        std::vector<vec3> colorBuffer((size_t)o.nx * o.ny);
        std::vector<crt_host::FrameTiming> rows;
        std::vector<float> verts;
        // --pose: how a frame's geometry reaches the device. vertices: 9 floats per triangle; points: the mesh's point
        // array through the index buffer (what updateFBXObj uploads, createScene.h:125-140); bones: one translation
        // per rigid part (Update_BVH over bones, createScene.h:80-97)
        const bool moving = o.animate && o.scene == "mesh";
        int pose_kind = CRT_POSE_VERTICES;
        const size_t n_bones = 8;
        std::vector<vec3> nowTransform(n_bones, vec3(0.f, 0.f, 0.f));   // vec3() leaves its members unset, like the reference's (core/vec3.h:9)
        if (o.pose == "points") pose_kind = CRT_POSE_POINTS;
        else if (o.pose == "bones") pose_kind = CRT_POSE_PARTS;
        else if (o.pose != "vertices") die("unknown --pose " + o.pose);
        if (moving && pose_kind == CRT_POSE_POINTS) {
            std::vector<uint32_t> idx;
            crt_host::mesh_indices(o.sc.n_theta, o.sc.n_phi, idx);
            r.set_indices(idx.data(), 0, idx.size() / 3, (size_t)(o.sc.n_theta + 1) * (o.sc.n_phi + 1));
        }
        if (moving && pose_kind == CRT_POSE_PARTS) {                  // bones = groups of latitude bands; the ground is bone 0
            std::vector<int> bands;
            crt_host::mesh_vertices(o.sc.n_theta, o.sc.n_phi, 0.0, verts, &bands);
            std::vector<uint32_t> part(r.flat().n_triangles(), 0u);
            for (size_t t = 0; t < bands.size(); t++) part[t] = (uint32_t)((size_t)bands[t] * n_bones / (size_t)o.sc.n_phi);
            r.createBoneBVH(part.data(), part.size(), n_bones);
        }
        auto make_pose = [&](int frame) -> size_t {                   // fills verts with the frame's pose, returns its rows
            if (pose_kind == CRT_POSE_POINTS) { crt_host::mesh_points(o.sc.n_theta, o.sc.n_phi, 0.3 * frame, verts); return verts.size() / 3; }
            if (pose_kind == CRT_POSE_PARTS) {
                for (size_t b = 1; b < n_bones; b++) nowTransform[b] = vec3(0.f, 3.f * std::sin(0.3f * frame + (float)b), 0.f);
                verts.assign(&nowTransform[0].e[0], &nowTransform[0].e[0] + 3 * n_bones);
                return n_bones;
            }
            crt_host::mesh_vertices(o.sc.n_theta, o.sc.n_phi, 0.3 * frame, verts, nullptr);
            return verts.size() / 9;
        };
        for (int frame = 0; frame < o.frames; frame++) {
            double update_s = 0.0;
            if (moving && o.overlap) {                                // the whole loop body as one overlapped call
                const size_t rows_n = make_pose(frame);
                t0 = now_s();
                r.render_frame_pose(pose_kind, verts.data(), rows_n, colorBuffer.data(), cam, o.nx, o.ny, o.spp, o.depth, mode,
                                    o.sc.seed + (uint64_t)frame);
                const double frame_s = now_s() - t0;
                if (!WritePng(o.nx, o.ny, frame, colorBuffer.data(), o.out.c_str())) die("cannot write " + o.out + std::to_string(frame) + ".png");
                rows.push_back({frame, frame_s, 0.0, frame == 0 ? build_s : 0.0});
                if (!o.quiet) std::printf("frame %d: update + refit + rendering %g s (refit %g ms on the device)\n", frame, frame_s, r.counters().refit_ms);
                continue;
            }
            if (moving) {                                             // render.h:207-217: new pose, then refit
                t0 = now_s();
                const size_t rows_n = make_pose(frame);
                if (pose_kind == CRT_POSE_POINTS) { r.update_points(verts.data(), rows_n); r.Update_BVH(); }
                else if (pose_kind == CRT_POSE_PARTS) r.Update_BVH(nowTransform.data(), n_bones);
                else { r.update_vertices(verts.data(), 0, rows_n); r.Update_BVH(); }
                update_s = now_s() - t0;
            }
            t0 = now_s();
            r.render(colorBuffer.data(), cam, o.nx, o.ny, o.spp, o.depth, mode, o.sc.seed + (uint64_t)frame);   // render.h:219-228
            const double render_s = now_s() - t0;
            if (!WritePng(o.nx, o.ny, frame, colorBuffer.data(), o.out.c_str())) die("cannot write " + o.out + std::to_string(frame) + ".png");
            rows.push_back({frame, render_s, update_s, frame == 0 ? build_s : 0.0});
            if (!o.quiet) {
                const crt_counters c = r.counters();
                std::printf("frame %d: rendering %g s (device %g ms), update %g s; %llu rays so far\n", frame, render_s, c.render_ms, update_s, (unsigned long long)c.rays);
            }
        }
        if (!o.csv.empty() && !crt_host::write_csv(o.csv, rows)) die("cannot write " + o.csv);
    } catch (const std::exception& e) {
        die(e.what());
    }
    return 0;
}
