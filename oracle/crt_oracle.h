/* crt_oracle.h — TEST INFRASTRUCTURE. C ABI of the plain-C restatement of the
 * reference's render hot path; mirrors the ref_* entry points of
 * oracle/ref_harness.cpp one for one so the same Python wrapper drives both.
 * See crt_oracle.c for the reference file:line each function follows. */
#ifndef CRT_ORACLE_H
#define CRT_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
int    oracle_num_threads(void);
void*  oracle_scene_create(int d1);
void   oracle_scene_destroy(void* scene);
int    oracle_scene_add_material(void* scene, uint32_t type, const float* albedo, float param);
void   oracle_checker_odd(const float* points, size_t n, uint8_t* out);
int    oracle_scene_add_checker_material(void* scene, uint32_t type, const float* even, const float* odd);
void   oracle_scene_add_triangles(void* scene, const float* tri, const uint32_t* mat, const uint32_t* ids, size_t n);
void   oracle_scene_add_spheres(void* scene, const float* sph, const uint32_t* mat, const uint32_t* ids, size_t n);
int    oracle_scene_add_chained(void* scene, int kind, const float* data, const uint32_t* mat, const uint32_t* ids,
                                const uint8_t* flip, size_t n, const float* levels, int n_levels);
void   oracle_rotation_rows(const float* rot3, float* out9);
int    oracle_scene_finalize(void* scene);
double oracle_trace(void* scene, const float* rays, size_t n, float t_min, float t_max, int use_bvh,
                    float* out_t, uint32_t* out_id, float* out_p, float* out_n);
int    oracle_scatter(uint32_t type, const float* albedo, float param, const float* ray, const float* recpn,
                      const float* randoms, int n_randoms, float* out, int* n_draws);
void   oracle_camera(const float* lookfrom, const float* lookat, const float* vup, float vfov, float aspect,
                     float aperture, float focus_dist, float* out22);
uint32_t oracle_get_ray(const float* cam22, float s, float t, uint64_t seed, uint32_t pixel, uint32_t sample,
                        uint32_t ctr, float* out6);
double oracle_render(void* scene, const float* cam22, int nx, int ny, int ns, int max_depth, int mode,
                     uint64_t seed, int y0, int y1, int y_step, float* rgb, float* accum, uint64_t* rays_out);
void   oracle_pack_rgba8(const float* rgb, int nx, int ny, uint8_t* rgba);
int    oracle_parse_obj(const char* path, float* points, float* idx, int max_points, int max_tris, int* np, int* nt);
void   oracle_triangle_normal(const float* v9, float* n3);
#ifdef __cplusplus
}
#endif
#endif
