// emu.cu — TEST INFRASTRUCTURE. Compiles the per-ray device functions of
// cudaraytracer_b200/csrc/crt_device.cuh for the HOST (they are
// __host__ __device__) so the CPU-only test tier can compare the exact code the
// kernels inline against the oracle before any GPU is involved. Not part of
// the product; nothing in cudaraytracer_b200/ loads it.
#include "../../cudaraytracer_b200/csrc/crt_device.cuh"
#include <string.h>

using namespace crt;

extern "C" {

// same contract as oracle_scatter / ref_scatter but with the counter stream
int emu_scatter(uint32_t type, const float* albedo, float param, const float* ray, const float* recpn,
                uint64_t key, uint32_t ctr, float* out, uint32_t* ctr_out)
{
    DMat m;
    m.r = albedo[0]; m.g = albedo[1]; m.b = albedo[2]; m.param = param; m.type = type;
    m.r2 = m.g2 = m.b2 = 0.0f;
    Rng r; r.key = key; r.ctr = ctr;
    v3 att, od;
    bool ok = scatter(m, V(ray[3], ray[4], ray[5]), V(recpn[0], recpn[1], recpn[2]), V(recpn[3], recpn[4], recpn[5]), r, att, od);
    out[0] = att.x; out[1] = att.y; out[2] = att.z;
    out[3] = recpn[0]; out[4] = recpn[1]; out[5] = recpn[2];
    out[6] = od.x; out[7] = od.y; out[8] = od.z;
    *ctr_out = r.ctr;
    return ok ? 1 : 0;
}

// which colour of a CheckerTexture the points select (1 = odd), n points x 3 floats
void emu_checker_odd(const float* p, size_t n, uint8_t* out)
{
    DMat m;
    m.r = m.g = m.b = 0.0f; m.param = 0.0f; m.type = CRT_DMAT_CHECKER; m.r2 = m.g2 = m.b2 = 1.0f;
    for (size_t i = 0; i < n; i++) out[i] = checker_odd(m, V(p[3 * i], p[3 * i + 1], p[3 * i + 2])) ? 1 : 0;
}

uint32_t emu_get_ray(const float* c, float s, float t, uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t ctr,
                     float* out6)
{
    DCamera cam;
    cam.llc = V(c[0], c[1], c[2]); cam.horizontal = V(c[3], c[4], c[5]); cam.vertical = V(c[6], c[7], c[8]);
    cam.origin = V(c[9], c[10], c[11]); cam.x = V(c[12], c[13], c[14]); cam.y = V(c[15], c[16], c[17]);
    cam.z = V(c[18], c[19], c[20]); cam.lens_radius = c[21];
    Rng r; r.key = crt_rng_key(seed, pixel, sample); r.ctr = ctr;
    v3 o, d;
    camera_get_ray(cam, s, t, r, o, d);
    out6[0] = o.x; out6[1] = o.y; out6[2] = o.z; out6[3] = d.x; out6[4] = d.y; out6[5] = d.z;
    return r.ctr;
}

// Linear scan with the device primitive tests and tie rule: the host twin of
// k_trace_brute. tri: n x 12 (v0 v1 v2 n), sph: m x 4.
void emu_trace(const float* tri, const uint32_t* tri_id, uint32_t n_tri, const float* sph, const uint32_t* sph_id,
               uint32_t n_sph, const float* rays, size_t n, float t_min, float t_max, float* out_t, uint32_t* out_id)
{
    for (size_t i = 0; i < n; i++) {
        const float* q = rays + 6 * i;
        v3 o = V(q[0], q[1], q[2]);
        v3 d = ray_dir_local(V(q[3], q[4], q[5]));
        float best = t_max;
        uint32_t best_id = CRT_MISS_ID;
        for (uint32_t j = 0; j < n_tri; j++) {
            const float* t = tri + 12 * (size_t)j;
            v3 v0 = V(t[0], t[1], t[2]);
            float tt, u, v;
            if (!triangle_test(o, d, v0, vsub(V(t[3], t[4], t[5]), v0), vsub(V(t[6], t[7], t[8]), v0),
                               V(t[9], t[10], t[11]), t_min, best, tt, u, v)) continue;
            if (tt == best && (best_id == CRT_MISS_ID || tri_id[j] >= best_id)) continue;
            best = tt; best_id = tri_id[j];
        }
        for (uint32_t j = 0; j < n_sph; j++) {
            const float* s = sph + 4 * (size_t)j;
            float tt;
            if (!sphere_test(vsub(o, V(s[0], s[1], s[2])), d, s[3], t_min, best, tt)) continue;
            if (tt == best && (best_id == CRT_MISS_ID || sph_id[j] >= best_id)) continue;
            best = tt; best_id = sph_id[j];
        }
        out_t[i] = best_id != CRT_MISS_ID ? best : 0.f;
        out_id[i] = best_id;
    }
}

// The same scan over primitives under chains of Transforms (crt_device.cuh:framed_test), the host twin of what the
// traversal kernels run for them: frames = n x 13 floats (9 rows, position, has_rotation), kind / chain / ids per
// primitive in id order, tri12 / sph4 indexed by `slot` (position in its kind's array). Also returns the hit point
// and the normal as shade() would see them (local frame, normal rotated up the chain).
void emu_trace_chained(const float* frames13, uint32_t n_frames, const uint32_t* kind, const uint32_t* chain, const uint32_t* slot,
                       const uint32_t* ids, const uint8_t* flip, uint32_t n_prims, const float* tri12, const float* sph4,
                       const float* rays, size_t n, float t_min, float t_max, float* out_t, uint32_t* out_id, float* out_p, float* out_n)
{
    DFrame* fr = new DFrame[n_frames ? n_frames : 1];
    for (uint32_t i = 0; i < n_frames; i++) {
        memcpy(fr[i].m, frames13 + 13 * i, 36);
        fr[i].px = frames13[13 * i + 9]; fr[i].py = frames13[13 * i + 10]; fr[i].pz = frames13[13 * i + 11];
        fr[i].has_rot = frames13[13 * i + 12] != 0.f;
    }
    for (size_t i = 0; i < n; i++) {
        const float* q = rays + 6 * i;
        const v3 o = V(q[0], q[1], q[2]), d_raw = V(q[3], q[4], q[5]);
        const v3 d = ray_dir_local(d_raw);
        float best = t_max;
        uint32_t best_id = CRT_MISS_ID, best_j = 0;
        for (uint32_t j = 0; j < n_prims; j++) {
            float4 r0 = make_float4(0, 0, 0, 0), r1 = r0, r2 = r0;
            uint32_t info = 0;
            if (kind[j] == 0) {
                const float* t = tri12 + 12 * (size_t)slot[j];
                const v3 v0 = V(t[0], t[1], t[2]), e1 = vsub(V(t[3], t[4], t[5]), v0), e2 = vsub(V(t[6], t[7], t[8]), v0);
                r0 = make_float4(v0.x, v0.y, v0.z, t[9]); r1 = make_float4(e1.x, e1.y, e1.z, t[10]); r2 = make_float4(e2.x, e2.y, e2.z, t[11]);
            } else if (kind[j] == 1) {
                const float* s = sph4 + 4 * (size_t)slot[j];
                r0 = make_float4(s[0], s[1], s[2], s[3]);
                info = CRT_KIND_SPHERE;
            } else {
                info = CRT_KIND_RECT | (flip[j] ? CRT_RECT_FLIP : 0u);
            }
            float tt, u, v;
            bool ok;
            if (chain[j]) ok = framed_test(fr, chain[j], info, r0, r1, r2, o, d_raw, t_min, best, tt, u, v);
            else if (kind[j] == 1) ok = sphere_test(vsub(o, V(r0.x, r0.y, r0.z)), d, r0.w, t_min, best, tt);
            else ok = triangle_test(o, d, V(r0.x, r0.y, r0.z), V(r1.x, r1.y, r1.z), V(r2.x, r2.y, r2.z), V(r0.w, r1.w, r2.w), t_min, best, tt, u, v);
            if (!ok) continue;
            if (tt == best && (best_id == CRT_MISS_ID || ids[j] >= best_id)) continue;
            best = tt; best_id = ids[j]; best_j = j;
        }
        out_t[i] = best_id != CRT_MISS_ID ? best : 0.f;
        out_id[i] = best_id;
        if (best_id == CRT_MISS_ID) continue;
        // hit point and normal as render.cu:shade_step derives them
        const uint32_t j = best_j;
        v3 ol = o, dl;
        if (chain[j]) chain_ray(fr, chain[j], o, d_raw, ol, dl); else dl = d;
        v3 p, nn;
        if (kind[j] == 1) { const float* s = sph4 + 4 * (size_t)slot[j]; p = vadd(vsub(ol, V(s[0], s[1], s[2])), vscale(best, dl)); nn = vdivf(p, s[3]); }
        else if (kind[j] == 2) { p = vadd(ol, vscale(best, dl)); nn = V(0.f, 0.f, flip[j] ? -1.f : 1.f); }
        else { const float* t = tri12 + 12 * (size_t)slot[j]; p = vadd(ol, vscale(best, dl)); nn = V(t[9], t[10], t[11]); }
        if (chain[j]) nn = chain_normal(fr, chain[j], nn);
        out_p[3 * i] = p.x; out_p[3 * i + 1] = p.y; out_p[3 * i + 2] = p.z;
        out_n[3 * i] = nn.x; out_n[3 * i + 1] = nn.y; out_n[3 * i + 2] = nn.z;
    }
    delete[] fr;
}

// Deviation D3's two forms on n (ray, triangle) pairs: rays n x 6 (origin, unit direction), tris n x 12. out[i]: bit 0
// = the triangle test up to D1 gives a candidate, bit 1 = the shortcut d3_agrees accepts it, bit 2 = the definition
// d3_interval accepts it. The shortcut must imply the definition.
void emu_d3(const float* rays, const float* tris, size_t n, uint8_t* out)
{
    for (size_t i = 0; i < n; i++) {
        const float* q = rays + 6 * i;
        const float* t = tris + 12 * i;
        const v3 o = V(q[0], q[1], q[2]), d = V(q[3], q[4], q[5]);
        const v3 v0 = V(t[0], t[1], t[2]), e1 = vsub(V(t[3], t[4], t[5]), v0), e2 = vsub(V(t[6], t[7], t[8]), v0);
        float tt, u, v;
        out[i] = 0;
        if (!triangle_candidate(o, d, v0, e1, e2, V(t[9], t[10], t[11]), 0.001f, 3.4e38f, tt, u, v)) continue;
        out[i] = 1 | (d3_agrees(o, d, v0, e1, e2, tt, u, v) ? 2 : 0) | (d3_interval(o, d, v0, e1, e2, tt) ? 4 : 0);
    }
}

void emu_sky(const float* d, float* out)
{
    v3 s = background_sky(V(d[0], d[1], d[2]));
    out[0] = s.x; out[1] = s.y; out[2] = s.z;
}

float emu_uniform(uint64_t key, uint32_t ctr) { return crt_rng_uniform(key, ctr); }
uint64_t emu_key(uint64_t seed, uint32_t pixel, uint32_t sample) { return crt_rng_key(seed, pixel, sample); }

}  // extern "C"
