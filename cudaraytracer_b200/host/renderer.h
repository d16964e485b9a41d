// renderer.h — the reference's host call sequence on top of the C ABI.
//
// kernel.cu:41-111 drives the reference through a handful of host functions
// (SURVEY 8b). Renderer keeps their names and order and forwards each to the
// entry point of include/crt.h that replaces it:
//
//   reference (file:line)                              here                        C ABI
//   create_FBXMesh / init_List  createScene.h:167,192  set_world(list)             crt_scene_set_*
//   Transform rotations, Rectangle hitable/transform.h  set_world (chains of frames)  crt_scene_set_frames / _chains / _rectangles
//   create_BVHfromList          createScene.h:212      create_BVHfromList()        crt_accel_build
//   updateFBXObj                createScene.h:125      update_vertices(...)        crt_scene_update_vertices
//   update_pose(newPos, idx)    createScene.h:99,125   set_indices / update_points crt_scene_set_indices / _update_points
//   createBoneBVH               createScene.h:256      createBoneBVH(part, n)      crt_scene_set_parts
//   Update_BVH(boneBvh, obj)    createScene.h:80       Update_BVH(nowTransform, n) crt_scene_move_parts + crt_accel_refit
//   Update_BVH                  createScene.h:53       Update_BVH()                crt_accel_refit
//   render<<<>>> + D2H copy     render.h:105,222-226   render(colorBuffer, ...)    crt_render
//   the loop body as a whole    render.h:206-233       render_frame(...)           crt_render_frame (steps overlapped)
//   WritePng                    render.h:135           WritePng (png_out.h)        -
//   (one GPU only in the reference)                    comm_init / scene_broadcast crt_comm_init / crt_scene_broadcast
//                                                      render_dist(colorBuffer...) crt_render_dist (tiles per GPU + ncclReduce)
//   StopWatch + writeCSV        swatch.h, CSVWriter.h  counters(), write_csv()     crt_get_counters
//
// Errors: the reference prints and exits (core/deviceManage.h:3-17). The C ABI
// returns codes; this wrapper turns them into std::runtime_error carrying
// crt_last_error(), and the CLI prints and exits like the reference.
// There is no CPU path: constructing a Renderer without a CUDA device throws.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/crt.h"
#include "scene.h"

namespace crt_host {

inline void check(int rc, const char* what)
{
    if (rc != CRT_OK) throw std::runtime_error(std::string(what) + ": " + crt_last_error());
}

struct FrameTiming { int frame; double rendering, update, build; };  // CSV columns of kernel.cu:61

class Renderer {
public:
    explicit Renderer(int device = 0) { check(crt_create(device, &ctx_), "crt_create"); }
    ~Renderer() { crt_destroy(ctx_); }
    Renderer(const Renderer&) = delete;
    Renderer& operator=(const Renderer&) = delete;

    crt_ctx* ctx() const { return ctx_; }

    // Flatten a world list and upload it (replaces add_mesh_withNormal / create_List).
    void set_world(const Hitable& world)
    {
        flat_ = flatten_scene(world);
        upload();
    }
    void set_flat(const FlatScene& f) { flat_ = f; upload(); }
    const FlatScene& flat() const { return flat_; }

    void create_BVHfromList() { check(crt_accel_build(ctx_), "crt_accel_build"); }
    void Update_BVH() { check(crt_accel_refit(ctx_), "crt_accel_refit"); }

    // New vertex positions of triangles [first, first + n): 9 floats each
    // (Triangle::SetVertices for every face, shapes/triangle.h:40-45).
    void update_vertices(const float* verts, size_t first, size_t n)
    {
        check(crt_scene_update_vertices(ctx_, verts, first, n), "crt_scene_update_vertices");
    }

    // Indexed pose, the form updateFBXObj / update_pose use (createScene.h:99-140): the face indices once
    // (idx[3 t + k] = point of stored vertex k of triangle first + t), then one point array per frame.
    void set_indices(const uint32_t* idx, size_t first, size_t n_tri, size_t n_points)
    {
        check(crt_scene_set_indices(ctx_, idx, first, n_tri, n_points), "crt_scene_set_indices");
    }
    void update_points(const float* points, size_t n_points)
    {
        check(crt_scene_update_points(ctx_, points, n_points), "crt_scene_update_points");
    }

    // Bones (createScene.h:256-307, hitable/BoneBVH.h): the part every triangle moves with; the current vertices
    // become the rest pose. Update_BVH(nowTransform, n) is the reference's per-frame bone update: one translation per
    // bone (BoneBVHNode::UpdateBVH(nowT), BoneBVH.h:113-137), then the refit.
    void createBoneBVH(const uint32_t* part_of_tri, size_t n_tri, size_t n_parts)
    {
        check(crt_scene_set_parts(ctx_, part_of_tri, n_tri, n_parts), "crt_scene_set_parts");
    }
    void Update_BVH(const vec3* nowTransform, size_t n_parts)
    {
        check(crt_scene_move_parts(ctx_, &nowTransform[0].e[0], n_parts), "crt_scene_move_parts");
        check(crt_accel_refit(ctx_), "crt_accel_refit");
    }

    // render<<<>>> + cudaMemcpy D2H (render.h:222-226): colorBuffer is the
    // caller's nx*ny vec3 array, row 0 at the bottom, values after gamma and clip.
    void render(vec3* colorBuffer, const Camera& camera, int nx, int ny, int samples, int max_depth,
                int shade_mode = CRT_SHADE_LAMBERT, uint64_t seed = 0)
    {
        static_assert(sizeof(vec3) == 3 * sizeof(float), "vec3 must be three packed floats");
        const crt_camera cam = camera.to_crt();
        check(crt_render(ctx_, &cam, nx, ny, samples, max_depth, shade_mode, seed, &colorBuffer[0].e[0]), "crt_render");
    }

    // One iteration of the animation loop (render.h:206-233): new pose, refit, render, copy back - one call, so the
    // upload and refit overlap ray generation and finished rows come back while the rest renders. Same image as
    // update_vertices + Update_BVH + render.
    void render_frame(const float* verts, size_t first, size_t n, vec3* colorBuffer, const Camera& camera, int nx, int ny,
                      int samples, int max_depth, int shade_mode = CRT_SHADE_LAMBERT, uint64_t seed = 0)
    {
        const crt_camera cam = camera.to_crt();
        check(crt_render_frame(ctx_, verts, first, n, &cam, nx, ny, samples, max_depth, shade_mode, seed, &colorBuffer[0].e[0]),
              "crt_render_frame");
    }

    // render_frame with the pose as points (CRT_POSE_POINTS) or bone translations (CRT_POSE_PARTS).
    void render_frame_pose(int pose_kind, const float* pose, size_t n, vec3* colorBuffer, const Camera& camera, int nx, int ny,
                           int samples, int max_depth, int shade_mode = CRT_SHADE_LAMBERT, uint64_t seed = 0)
    {
        const crt_camera cam = camera.to_crt();
        check(crt_render_frame_pose(ctx_, pose_kind, pose, 0, n, &cam, nx, ny, samples, max_depth, shade_mode, seed,
                                    &colorBuffer[0].e[0]), "crt_render_frame_pose");
    }

    // The pose + refit half of render_frame_pose alone (crt_scene_pose_async): returns at once; the next
    // crt_render_accum_device on ctx() overlaps its ray generation with the upload and the refit.
    void pose_async(int pose_kind, const float* pose, size_t first, size_t n)
    {
        check(crt_scene_pose_async(ctx_, pose_kind, pose, first, n), "crt_scene_pose_async");
    }

    // Several GPUs, one Renderer (context) per GPU, each driven by its own host thread or process: rank 0 draws the
    // id (crt_comm_unique_id) and hands it to the others; comm_init is collective. scene_broadcast ships rank 0's
    // uploaded scene to the other GPUs over NVLink; every rank then calls create_BVHfromList. render_dist renders
    // this rank's tiles of the frame, sums the partial framebuffers onto rank 0 (ncclReduce) and, there, resolves and
    // copies the image into colorBuffer (ignored, may be null, on the other ranks).
    void comm_init(const void* id, int rank, int world) { check(crt_comm_init(ctx_, id, rank, world), "crt_comm_init"); }
    void scene_broadcast(int root = 0) { check(crt_scene_broadcast(ctx_, root), "crt_scene_broadcast"); }
    void render_dist(vec3* colorBuffer, const Camera& camera, int nx, int ny, int samples, int max_depth,
                     int shade_mode = CRT_SHADE_LAMBERT, uint64_t seed = 0)
    {
        const crt_camera cam = camera.to_crt();
        check(crt_render_dist(ctx_, &cam, nx, ny, samples, max_depth, shade_mode, seed, colorBuffer ? &colorBuffer[0].e[0] : nullptr),
              "crt_render_dist");
    }

    // (*world)->hit for a batch of rays (render.h:53).
    void trace(const std::vector<Ray>& rays, std::vector<crt_hit>& hits, float t_min = 0.001f, float t_max = 3.402823466e38f,
               int accel = CRT_ACCEL_LBVH)
    {
        std::vector<crt_ray> r(rays.size());
        for (size_t i = 0; i < rays.size(); i++) r[i] = rays[i].to_crt();
        hits.resize(rays.size());
        check(crt_trace(ctx_, r.data(), r.size(), t_min, t_max, accel, hits.data()), "crt_trace");
    }

    crt_counters counters() const
    {
        crt_counters c;
        check(crt_get_counters(ctx_, &c), "crt_get_counters");
        return c;
    }
    void set_option(int option, int64_t value) { check(crt_set_option(ctx_, option, value), "crt_set_option"); }

private:
    void upload()
    {
        check(crt_scene_set_materials(ctx_, flat_.mats.data(), flat_.mats.size()), "crt_scene_set_materials");
        check(crt_scene_set_frames(ctx_, flat_.frames.data(), flat_.frames.size()), "crt_scene_set_frames");   // drops old chains
        check(crt_scene_set_triangles(ctx_, flat_.tri.data(), flat_.tri_mat.data(), flat_.tri_id.data(), flat_.n_triangles()),
              "crt_scene_set_triangles");
        check(crt_scene_set_spheres(ctx_, flat_.sph.data(), flat_.sph_mat.data(), flat_.sph_id.data(), flat_.n_spheres()),
              "crt_scene_set_spheres");
        if (flat_.has_chains()) {
            check(crt_scene_set_triangle_chains(ctx_, flat_.tri_chain.data(), flat_.n_triangles()), "crt_scene_set_triangle_chains");
            check(crt_scene_set_sphere_chains(ctx_, flat_.sph_chain.data(), flat_.n_spheres()), "crt_scene_set_sphere_chains");
        }
        check(crt_scene_set_rectangles(ctx_, flat_.rect_chain.data(), flat_.rect_flip.data(), flat_.rect_mat.data(), flat_.rect_id.data(),
                                       flat_.n_rectangles()), "crt_scene_set_rectangles");
    }
    crt_ctx* ctx_ = nullptr;
    FlatScene flat_;
};

// Loader/CSVWriter.h:8-32: header line, then one row per frame, "\n" endings.
inline bool write_csv(const std::string& path, const std::vector<FrameTiming>& rows)
{
    FILE* f = std::fopen(path.c_str(), "w");
    if (!f) return false;
    std::fprintf(f, "frame,rendering,update,build\n");
    for (const FrameTiming& r : rows) std::fprintf(f, "%d,%g,%g,%g\n", r.frame, r.rendering, r.update, r.build);
    return std::fclose(f) == 0;
}

}  // namespace crt_host
