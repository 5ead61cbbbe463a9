// Camera / frustum arithmetic shared by the host producers (pm_camera.cpp) and the on-device shadow fit
// (pm_kernels.cu, shadow_fit_kernel): one source text, compiled for both sides without contraction
// (gcc x86-64 without -mfma on the host, nvcc -fmad=false on the device; IEEE sqrt and division on both),
// so a frustum fitted on the device carries the bits the host call would have produced.
// Mirrors core/put/source/camera.cpp:131-172 (camera_update_frustum), :362-416 (get_aabb_corners,
// camera_update_shadow_frustum); conventions as described in pm_camera.cpp.
#pragma once
#include "../../include/pmtech_b200.h"

#include <cfloat>
#include <cmath>
#include <cstring>

#if defined(__CUDACC__)
#define PMB_HD __host__ __device__
#else
#define PMB_HD
#endif

namespace pmcam
{


struct v3
{
    float x, y, z;
};

PMB_HD inline v3    operator-(v3 a, v3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
PMB_HD inline v3    operator-(v3 a) { return {-a.x, -a.y, -a.z}; }
PMB_HD inline float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
PMB_HD inline v3    cross(v3 a, v3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
PMB_HD inline v3    normalize(v3 a)
{
    float m = sqrtf(dot(a, a));
    return {a.x / m, a.y / m, a.z / m};
}
PMB_HD inline v3   load3(const float* p) { return {p[0], p[1], p[2]}; }
PMB_HD inline void store3(float* p, v3 a) { p[0] = a.x, p[1] = a.y, p[2] = a.z; }

struct m4
{
    float m[16];
};

PMB_HD inline m4 identity()
{
    m4 r;
    memset(&r, 0, sizeof(r));
    r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1.0f;
    return r;
}

PMB_HD inline m4 mul(const m4& a, const m4& b)
{
    m4 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
        {
            float acc = 0.0f;
            for (int k = 0; k < 4; ++k)
                acc += a.m[i * 4 + k] * b.m[k * 4 + j];
            r.m[i * 4 + j] = acc;
        }
    return r;
}

PMB_HD inline m4 rows(v3 r0, v3 r1, v3 r2)
{
    m4 r = identity();
    r.m[0] = r0.x, r.m[1] = r0.y, r.m[2] = r0.z;
    r.m[4] = r1.x, r.m[5] = r1.y, r.m[6] = r1.z;
    r.m[8] = r2.x, r.m[9] = r2.y, r.m[10] = r2.z;
    return r;
}

PMB_HD inline m4 translation(v3 t)
{
    m4 r = identity();
    r.m[3] = t.x, r.m[7] = t.y, r.m[11] = t.z;
    return r;
}

PMB_HD inline v3 transform_point(const m4& m, v3 p)
{
    return {m.m[0] * p.x + m.m[1] * p.y + m.m[2] * p.z + m.m[3], m.m[4] * p.x + m.m[5] * p.y + m.m[6] * p.z + m.m[7],
            m.m[8] * p.x + m.m[9] * p.y + m.m[10] * p.z + m.m[11]};
}

// mat::inverse4x4 of the maths library: Laplace expansion over the twelve 2x2 sub-determinants in binary32, the odd
// cofactors negated after their sum — the same operation order as inverse4x4_rows (pm_kernels.cu), which pmtech's
// stored draw-call records pin bit for bit (tests/test_pmtech_scene_files.py).  maths::unproject_sc inverts the
// view-projection with it (camera.cpp:146-147), so the frustum corners carry its rounding, not a binary64 inverse's.
PMB_HD inline bool invert(const m4& in, m4& out)
{
    const float* m = in.m;
    const float  s0 = m[0] * m[5] - m[4] * m[1];
    const float  s1 = m[0] * m[6] - m[4] * m[2];
    const float  s2 = m[0] * m[7] - m[4] * m[3];
    const float  s3 = m[1] * m[6] - m[5] * m[2];
    const float  s4 = m[1] * m[7] - m[5] * m[3];
    const float  s5 = m[2] * m[7] - m[6] * m[3];
    const float  c5 = m[10] * m[15] - m[14] * m[11];
    const float  c4 = m[9] * m[15] - m[13] * m[11];
    const float  c3 = m[9] * m[14] - m[13] * m[10];
    const float  c2 = m[8] * m[15] - m[12] * m[11];
    const float  c1 = m[8] * m[14] - m[12] * m[10];
    const float  c0 = m[8] * m[13] - m[12] * m[9];
    const float  det = s0 * c5 - s1 * c4 + s2 * c3 + s3 * c2 - s4 * c1 + s5 * c0;
    const float  inv = 1.0f / det;
    float*       o = out.m;
    o[0] = (m[5] * c5 - m[6] * c4 + m[7] * c3) * inv;
    o[1] = -(m[1] * c5 - m[2] * c4 + m[3] * c3) * inv;
    o[2] = (m[13] * s5 - m[14] * s4 + m[15] * s3) * inv;
    o[3] = -(m[9] * s5 - m[10] * s4 + m[11] * s3) * inv;
    o[4] = -(m[4] * c5 - m[6] * c2 + m[7] * c1) * inv;
    o[5] = (m[0] * c5 - m[2] * c2 + m[3] * c1) * inv;
    o[6] = -(m[12] * s5 - m[14] * s2 + m[15] * s1) * inv;
    o[7] = (m[8] * s5 - m[10] * s2 + m[11] * s1) * inv;
    o[8] = (m[4] * c4 - m[5] * c2 + m[7] * c0) * inv;
    o[9] = -(m[0] * c4 - m[1] * c2 + m[3] * c0) * inv;
    o[10] = (m[12] * s4 - m[13] * s2 + m[15] * s0) * inv;
    o[11] = -(m[8] * s4 - m[9] * s2 + m[11] * s0) * inv;
    o[12] = -(m[4] * c3 - m[5] * c1 + m[6] * c0) * inv;
    o[13] = (m[0] * c3 - m[1] * c1 + m[2] * c0) * inv;
    o[14] = -(m[12] * s3 - m[13] * s1 + m[14] * s0) * inv;
    o[15] = (m[8] * s3 - m[9] * s1 + m[10] * s0) * inv;
    return det != 0.0f;
}

PMB_HD inline m4 frustum_projection(float l, float r, float b, float t, float n, float f)
{
    m4 o;
    memset(&o, 0, sizeof(o));
    o.m[0] = (2.0f * n) / (r - l);
    o.m[2] = (r + l) / (r - l);
    o.m[5] = (2.0f * n) / (t - b);
    o.m[6] = (t + b) / (t - b);
    o.m[10] = (-f - n) / (f - n);
    o.m[11] = (-(2.0f * n) * f) / (f - n);
    o.m[14] = -1.0f;
    return o;
}

PMB_HD inline m4 ortho_projection(float l, float r, float b, float t, float n, float f)
{
    m4 o;
    memset(&o, 0, sizeof(o));
    o.m[0] = 2.0f / (r - l);
    o.m[3] = -(r + l) / (r - l);
    o.m[5] = 2.0f / (t - b);
    o.m[7] = -(t + b) / (t - b);
    o.m[10] = -2.0f / (f - n);
    o.m[11] = -(f + n) / (f - n);
    o.m[15] = 1.0f;
    return o;
}

// maths::unproject_sc with a (1,1) viewport: [0,1]^3 -> NDC -> world
PMB_HD inline v3 unproject(float sx, float sy, float sz, const m4& inv_view_proj)
{
    float        ndc[4] = {sx * 2.0f - 1.0f, sy * 2.0f - 1.0f, sz * 2.0f - 1.0f, 1.0f};
    float        r[4];
    const float* m = inv_view_proj.m;
    for (int i = 0; i < 4; ++i)
        r[i] = m[i * 4] * ndc[0] + m[i * 4 + 1] * ndc[1] + m[i * 4 + 2] * ndc[2] + m[i * 4 + 3] * ndc[3];
    return {r[0] / r[3], r[1] / r[3], r[2] / r[3]};
}

PMB_HD inline m4& view_of(pmb_camera* c) { return *reinterpret_cast<m4*>(c->view); }
PMB_HD inline m4& proj_of(pmb_camera* c) { return *reinterpret_cast<m4*>(c->proj); }

// camera.cpp:131-172
PMB_HD inline void update_frustum(pmb_camera* cam)
{
    const float ndc[4][2] = {{0.0f, 1.0f}, {1.0f, 1.0f}, {0.0f, 0.0f}, {1.0f, 0.0f}};
    m4          inv;
    invert(mul(proj_of(cam), view_of(cam)), inv);
    v3 c[2][4];
    for (int i = 0; i < 4; ++i)
    {
        c[0][i] = unproject(ndc[i][0], ndc[i][1], 0.0f, inv);
        c[1][i] = unproject(ndc[i][0], ndc[i][1], 1.0f, inv);
    }
    for (int d = 0; d < 2; ++d)
        for (int i = 0; i < 4; ++i)
            store3(cam->camera_frustum.corners[d][i], c[d][i]);
    // left, top, right, bottom, near, far (:152-161)
    const v3 tri[6][3] = {{c[0][0], c[1][0], c[0][2]}, {c[0][0], c[0][1], c[1][0]}, {c[0][1], c[0][3], c[1][1]},
                          {c[0][2], c[1][2], c[0][3]}, {c[0][0], c[0][2], c[0][1]}, {c[1][0], c[1][1], c[1][2]}};
    for (int i = 0; i < 6; ++i)
    {
        v3 v1 = normalize(tri[i][1] - tri[i][0]);
        v3 v2 = normalize(tri[i][2] - tri[i][0]);
        store3(cam->camera_frustum.n[i], cross(v1, v2)); // not renormalised (:168)
        store3(cam->camera_frustum.p[i], tri[i][0]);
    }
}

// camera.cpp:384-416: light-space basis, fit an ortho box around the 8 corners of [mn, mx]
PMB_HD inline void update_shadow_frustum(pmb_camera* cam, const float light_dir[3], const float mn[3], const float mx[3])
{
    v3 dir = load3(light_dir);
    v3 rgt = cross(dir, v3{0.0f, 1.0f, 0.0f});
    v3 up = cross(rgt, dir);
    m4 sv = rows(rgt, up, -dir);
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    // get_aabb_corners (camera.cpp:362-382): corner = min + offset * size, in the reference's corner order
    const float offsets[8][3] = {{0, 0, 0}, {1, 1, 1}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {1, 0, 1}, {1, 1, 0}, {0, 1, 1}};
    const float size[3] = {mx[0] - mn[0], mx[1] - mn[1], mx[2] - mn[2]};
    for (int i = 0; i < 8; ++i)
    {
        v3 corner = {mn[0] + offsets[i][0] * size[0], mn[1] + offsets[i][1] * size[1], mn[2] + offsets[i][2] * size[2]};
        v3 p = transform_point(sv, corner);
        float q[3] = {p.x, p.y, p.z * -1.0f};
        for (int k = 0; k < 3; ++k)
        {
            lo[k] = lo[k] < q[k] ? lo[k] : q[k]; // min_union(cmin, p)
            hi[k] = hi[k] > q[k] ? hi[k] : q[k]; // max_union(cmax, p)
        }
    }
    view_of(cam) = sv;
    proj_of(cam) = ortho_projection(lo[0], hi[0], lo[1], hi[1], lo[2], hi[2]);
    update_frustum(cam);
}

// shadow_camera_from_entity, spot branch (ecs_scene.cpp:556-570): `proj` is camera_create_perspective(100, 1, 0.1,
// 500) — the same for every spot light, made once on the host (tanf); the view comes from the light's world matrix.
PMB_HD inline void spot_shadow(pmb_camera* cam, const m4& proj, const float w[16])
{
    const v3 r0 = normalize(v3{w[2], w[6], w[10]}); // row 0 = normalize(column 2)
    const v3 r1 = normalize(v3{w[0], w[4], w[8]});  // row 1 = normalize(column 0)
    const v3 r2 = normalize(v3{w[1], w[5], w[9]});  // row 2 = normalize(column 1)
    proj_of(cam) = proj;
    view_of(cam) = mul(rows(r0, r1, r2), translation(-v3{w[3], w[7], w[11]}));
    update_frustum(cam);
}

} // namespace pmcam
