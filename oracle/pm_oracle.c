/*
 * pm_oracle.c — CPU ORACLE (test infrastructure only; see pm_oracle.h header comment).
 * Pinned by pmtech's own saved scenes for loops A-C, stage 0 and the draw-call records; the plane
 * tests remain declared (parity unpinned for ecs_cull.cpp) — see the header comment of pm_oracle.h.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -fopenmp -fPIC -shared   (oracle/Makefile)
 * Every float operation below is written as one C operation per IEEE operation, left to
 * right, so that with contraction off the rounding sequence is exactly what the source says.
 *
 * Citations are relative to /root/reference/.
 */
#include "pm_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* torchrun exports OMP_NUM_THREADS=1 to its workers; the timed all-cores baseline asks for its threads itself */
void orc_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0)
        omp_set_num_threads(n);
#else
    (void)n;
#endif
}


/* ------------------------------------------------------------------------------------------
 * maths:: restatements (third_party/maths, absent).  Call sites that fix the semantics:
 *   mat4 * mat4             core/put/source/ecs/ecs_scene.cpp:1470
 *   transform_vector        ecs_scene.cpp:1505   (affine point transform, w = 1)
 *   get_translation         ecs_scene.cpp:1496   (m[3], m[7], m[11])
 *   min_union / max_union   ecs_scene.cpp:1507-1508,1524-1525,1548-1554
 *   mag                     ecs_scene.cpp:1512
 *   sgn, plane_distance, dot    core/put/source/ecs/ecs_cull.cpp:36-38
 *   point_plane_distance    ecs_cull.cpp:68
 * ---------------------------------------------------------------------------------------- */

/* generic Vec<N,T> dot: T r = 0; for i: r += a[i] * b[i]; */
static inline float dotn(const float* a, const float* b, int n)
{
    float r = 0.0f;
    for (int i = 0; i < n; ++i)
    {
        float t = a[i] * b[i];
        r = r + t;
    }
    return r;
}

float orc_dot3(const float a[3], const float b[3])
{
    return dotn(a, b, 3);
}

/* Mat<4,4>::operator*: result(r,c) = dot(lhs.row(r), rhs.column(c)) */
void orc_mat4_mul(orc_mat4* out, const orc_mat4* a, const orc_mat4* b)
{
    orc_mat4 res;
    for (int r = 0; r < 4; ++r)
    {
        for (int c = 0; c < 4; ++c)
        {
            float row[4] = {a->m[r * 4 + 0], a->m[r * 4 + 1], a->m[r * 4 + 2], a->m[r * 4 + 3]};
            float col[4] = {b->m[0 * 4 + c], b->m[1 * 4 + c], b->m[2 * 4 + c], b->m[3 * 4 + c]};
            res.m[r * 4 + c] = dotn(row, col, 4);
        }
    }
    *out = res;
}

/* Mat<4,4>::transform_vector(vec3): (v, 1) dotted with rows 0..2 */
void orc_transform_vector(float out[3], const orc_mat4* m, const float v[3])
{
    float v4[4] = {v[0], v[1], v[2], 1.0f};
    float res[3];
    for (int r = 0; r < 3; ++r)
        res[r] = dotn(v4, &m->m[r * 4], 4);
    out[0] = res[0];
    out[1] = res[1];
    out[2] = res[2];
}

static inline float fmin_union(float a, float b) { return a < b ? a : b; } /* min(a,b) */
static inline float fmax_union(float a, float b) { return a > b ? a : b; } /* max(a,b) */

/* sgn: (0 < v) - (v < 0) per component */
static inline float fsgn(float v) { return (float)((0.0f < v) - (v < 0.0f)); }

/* maths::plane_distance(x0, xN) = dot(xN, x0) * -1  (so that dot(x, n) + pd is the signed
 * distance, as the SIMD variants use it: ecs_cull.cpp:149,191-193) */
float orc_plane_distance(const float x0[3], const float xN[3])
{
    float d = dotn(xN, x0, 3);
    return d * -1.0f;
}

/* maths::point_plane_distance(p0, x0, xN) = dot(p0, xN) + plane_distance(x0, xN) */
float orc_point_plane_distance(const float p0[3], const float x0[3], const float xN[3])
{
    float d = orc_plane_distance(x0, xN);
    float pp = dotn(p0, xN, 3);
    return pp + d;
}

/* quat::normalise + quat::get_matrix (third_party/maths, absent).  get_matrix is a non-const member
 * that normalises the quaternion ITSELF and then expands it: every call site hands it a mutable
 * lvalue (`cmp_transform& t; t.rotation.get_matrix`, ecs_scene.cpp:1416-1421; `quat q = …;
 * q.get_matrix(rot_mat); … rotation = q`, :1291-1300).  What the reference's own scene file pins
 * (tests/golden/pmtech_sdf_scene.npz, examples/assets/scene/sdf.pms): the EXPANSION of each stored
 * rotation reproduces the stored local matrix bit for bit, and re-normalising a stored rotation on a
 * copy does not (2 of 7 rotated entities move by 1-2 ulp) — so the stored rotation is the one the
 * expansion ran on, i.e. the normalisation happens in place (or not at all; the file cannot tell these
 * apart, gizmo-accumulated rotations that stay within 1e-7 of unit length speak for in place). */
void orc_quat_normalise(float q[4])
{
    float x = q[0], y = q[1], z = q[2], w = q[3];
    float mag2 = ((w * w + x * x) + y * y) + z * z;
    float mag = sqrtf(mag2);
    if (mag != 0.0f && mag != 1.0f)
    {
        q[0] = x / mag;
        q[1] = y / mag;
        q[2] = z / mag;
        q[3] = w / mag;
    }
}

void orc_quat_expand(orc_mat4* out, const float q[4])
{
    float x = q[0], y = q[1], z = q[2], w = q[3];
    float* m = out->m;
    m[0] = 1.0f - 2.0f * y * y - 2.0f * z * z;
    m[1] = 2.0f * x * y - 2.0f * z * w;
    m[2] = 2.0f * x * z + 2.0f * y * w;
    m[3] = 0.0f;
    m[4] = 2.0f * x * y + 2.0f * z * w;
    m[5] = 1.0f - 2.0f * x * x - 2.0f * z * z;
    m[6] = 2.0f * y * z - 2.0f * x * w;
    m[7] = 0.0f;
    m[8] = 2.0f * x * z - 2.0f * y * w;
    m[9] = 2.0f * y * z + 2.0f * x * w;
    m[10] = 1.0f - 2.0f * x * x - 2.0f * y * y;
    m[11] = 0.0f;
    m[12] = 0.0f;
    m[13] = 0.0f;
    m[14] = 0.0f;
    m[15] = 1.0f;
}

void orc_quat_get_matrix(orc_mat4* out, float q[4])
{
    orc_quat_normalise(q);
    orc_quat_expand(out, q);
}

static void mat4_identity(orc_mat4* m)
{
    memset(m, 0, sizeof(*m));
    m->m[0] = m->m[5] = m->m[10] = m->m[15] = 1.0f;
}

/* ------------------------------------------------------------------------------------------
 * update_scene loop A — core/put/source/ecs/ecs_scene.cpp:1405-1471
 * ---------------------------------------------------------------------------------------- */
void orc_update_transforms(orc_scene* scene)
{
    for (size_t n = 0; n < scene->num_entities; ++n)
    {
        /* :1408-1412 force physics entity to sync and ignore controlled transform */
        if (scene->state_flags[n] & ORC_STATE_SYNC_PHYSICS_TRANSFORM)
        {
            scene->state_flags[n] &= ~ORC_STATE_SYNC_PHYSICS_TRANSFORM;
            scene->entities[n] &= ~ORC_CMP_TRANSFORM;
        }

        /* :1415-1442 controlled transform: local = T * R * S, then clear the flag.
         * The bake itself runs only when the caller supplies transforms[] (stage 0, SURVEY
         * §8f-1); with transforms == NULL local_matrices[n] is taken as already baked by the
         * host.  The physics set_transform commands (:1429-1438) are host-side, out of scope. */
        if (scene->entities[n] & ORC_CMP_TRANSFORM)
        {
            if (scene->transforms)
            {
                orc_transform* t = &scene->transforms[n];                  /* :1416 cmp_transform& t */
                orc_mat4 rot_mat, translation_mat, scale_mat, tr;
                orc_quat_get_matrix(&rot_mat, t->rotation);                /* :1421 normalises t.rotation */
                mat4_identity(&translation_mat);
                translation_mat.m[3] = t->translation[0];
                translation_mat.m[7] = t->translation[1];
                translation_mat.m[11] = t->translation[2];
                mat4_identity(&scale_mat);
                scale_mat.m[0] = t->scale[0];
                scale_mat.m[5] = t->scale[1];
                scale_mat.m[10] = t->scale[2];
                orc_mat4_mul(&tr, &translation_mat, &rot_mat);             /* (T * R) */
                orc_mat4_mul(&scene->local_matrices[n], &tr, &scale_mat);  /* * S     */
            }
            scene->entities[n] &= ~ORC_CMP_TRANSFORM;                      /* :1441 */
        }
        else if (scene->entities[n] & ORC_CMP_PHYSICS)
        {
            /* :1443-1446 — no rigid-body matrix yet: `continue`, world matrix left untouched.
             * (The has-matrix branch :1448-1462 rebuilds local from the physics thread's
             * transform on the host, then falls through to the hierarchy step.) */
            int has = scene->has_rb_matrix ? scene->has_rb_matrix[n] : 0;
            if (!has)
                continue;
        }

        /* :1465-1470 hierarchical scene transform */
        uint32_t parent = scene->parents[n];
        if (parent == n)
            scene->world_matrices[n] = scene->local_matrices[n];
        else
            orc_mat4_mul(&scene->world_matrices[n], &scene->world_matrices[parent], &scene->local_matrices[n]);
    }
}

/* ------------------------------------------------------------------------------------------
 * draw-call packing — ecs_scene.cpp:1962-1991
 * ---------------------------------------------------------------------------------------- */
void orc_inverse4x4(orc_mat4* out, const orc_mat4* mat)
{
    const float* m = mat->m;
    /* 2x2 sub-determinants of the upper two and the lower two rows */
    float s0 = m[0] * m[5] - m[4] * m[1];
    float s1 = m[0] * m[6] - m[4] * m[2];
    float s2 = m[0] * m[7] - m[4] * m[3];
    float s3 = m[1] * m[6] - m[5] * m[2];
    float s4 = m[1] * m[7] - m[5] * m[3];
    float s5 = m[2] * m[7] - m[6] * m[3];
    float c5 = m[10] * m[15] - m[14] * m[11];
    float c4 = m[9] * m[15] - m[13] * m[11];
    float c3 = m[9] * m[14] - m[13] * m[10];
    float c2 = m[8] * m[15] - m[12] * m[11];
    float c1 = m[8] * m[14] - m[12] * m[10];
    float c0 = m[8] * m[13] - m[12] * m[9];
    float det = s0 * c5 - s1 * c4 + s2 * c3 + s3 * c2 - s4 * c1 + s5 * c0;
    float invdet = 1.0f / det;
    float* o = out->m;
    o[0] = (m[5] * c5 - m[6] * c4 + m[7] * c3) * invdet;
    o[1] = -(m[1] * c5 - m[2] * c4 + m[3] * c3) * invdet;
    o[2] = (m[13] * s5 - m[14] * s4 + m[15] * s3) * invdet;
    o[3] = -(m[9] * s5 - m[10] * s4 + m[11] * s3) * invdet;
    o[4] = -(m[4] * c5 - m[6] * c2 + m[7] * c1) * invdet;
    o[5] = (m[0] * c5 - m[2] * c2 + m[3] * c1) * invdet;
    o[6] = -(m[12] * s5 - m[14] * s2 + m[15] * s1) * invdet;
    o[7] = (m[8] * s5 - m[10] * s2 + m[11] * s1) * invdet;
    o[8] = (m[4] * c4 - m[5] * c2 + m[7] * c0) * invdet;
    o[9] = -(m[0] * c4 - m[1] * c2 + m[3] * c0) * invdet;
    o[10] = (m[12] * s4 - m[13] * s2 + m[15] * s0) * invdet;
    o[11] = -(m[8] * s4 - m[9] * s2 + m[11] * s0) * invdet;
    o[12] = -(m[4] * c3 - m[5] * c1 + m[6] * c0) * invdet;
    o[13] = (m[0] * c3 - m[1] * c1 + m[2] * c0) * invdet;
    o[14] = -(m[12] * s3 - m[13] * s1 + m[14] * s0) * invdet;
    o[15] = (m[8] * s3 - m[9] * s1 + m[10] * s0) * invdet;
}

void orc_update_draw_calls(const orc_scene* scene, const uint32_t* cbuffer, orc_draw_call* dc, float time_ms)
{
    for (size_t n = 0; n < scene->num_entities; ++n)
    {
        /* :1964-1970 material cbuffer upload: renderer side, not part of the data path */
        dc[n].world_matrix = scene->world_matrices[n];                     /* :1972 */
        dc[n].v1[0] = (float)n;                                            /* :1975 */
        dc[n].v1[1] = time_ms;                                             /* :1976 pen::get_time_ms() */
        if (cbuffer[n] == 0xFFFFFFFFu || cbuffer[n] == 0u)                 /* :1978 is_invalid_or_null */
            continue;
        if (scene->entities[n] & ORC_CMP_SUB_INSTANCE)                     /* :1981 */
            continue;
        if ((scene->entities[n] & ORC_CMP_SKINNED) || (scene->entities[n] & ORC_CMP_PRE_SKINNED)) /* :1985 */
            mat4_identity(&dc[n].world_matrix);
        orc_mat4 invt;
        const float* w = scene->world_matrices[n].m;
        for (int r = 0; r < 4; ++r)                                        /* :1990 transposed() */
            for (int c = 0; c < 4; ++c)
                invt.m[r * 4 + c] = w[c * 4 + r];
        orc_mat4 inv;
        orc_inverse4x4(&inv, &invt);                                       /* :1991 */
        dc[n].world_matrix_inv_transpose = inv;                            /* :1993 */
    }
}

/* ------------------------------------------------------------------------------------------
 * root motion — ecs_scene.cpp:1286-1304 (the part of the controller update that feeds stage 0)
 * ---------------------------------------------------------------------------------------- */
void orc_apply_root_motion(orc_scene* scene, const uint32_t* targets, const float* rotations, const float* deltas,
                           uint32_t count)
{
    for (uint32_t i = 0; i < count; ++i)
    {
        uint32_t p = targets[i];
        if (p >= scene->num_entities)
            continue;
        orc_mat4 rot_mat;
        float q[4] = {rotations[i * 4], rotations[i * 4 + 1], rotations[i * 4 + 2], rotations[i * 4 + 3]}; /* :1291 quat q = … */
        orc_quat_get_matrix(&rot_mat, q);                                   /* :1292 (q normalised) */
        float tt[3];
        orc_transform_vector(tt, &rot_mat, &deltas[i * 3]);                 /* :1296 */
        orc_transform* t = &scene->transforms[p];
        for (int k = 0; k < 4; ++k)
            t->rotation[k] = q[k];                                          /* :1300 rotation = q */
        for (int k = 0; k < 3; ++k)
            t->translation[k] = t->translation[k] + tt[k];                  /* :1303 */
        scene->entities[p] |= ORC_CMP_TRANSFORM;                            /* :1304 */
    }
}

/* ------------------------------------------------------------------------------------------
 * skin palettes — ecs_scene.cpp:1917-1959 (same product in the pre-skin path, :1872-1880)
 * ---------------------------------------------------------------------------------------- */
void orc_skin_palettes(const orc_scene* scene, const orc_skin* skins, uint32_t num_skins, const orc_mat4* bind, orc_mat4* out)
{
    for (uint32_t k = 0; k < num_skins; ++k)
    {
        const orc_skin* sk = &skins[k];
        for (uint32_t i = 0; i < sk->num_joints; ++i)                       /* :1950-1955 */
        {
            uint32_t joint = sk->joints_offset + i;
            if (joint >= scene->num_entities)
                continue;
            orc_mat4_mul(&out[sk->palette_offset + i], &scene->world_matrices[joint], &bind[sk->bind_offset + i]);
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * update_scene loop B — ecs_scene.cpp:1474-1526
 * ---------------------------------------------------------------------------------------- */
static const float k_corners[8][3] = {
    /* :1474-1480 */
    {0.0f, 0.0f, 0.0f},
    {1.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f}, {0.0f, 0.0f, 1.0f},
    {1.0f, 1.0f, 0.0f}, {0.0f, 1.0f, 1.0f}, {1.0f, 0.0f, 1.0f},
    {1.0f, 1.0f, 1.0f}};

static void bounds_one(orc_scene* scene, size_t n, float ext_min[3], float ext_max[3])
{
    orc_bounding_volume* bv = &scene->bounding_volumes[n];
    float min[3], max[3];
    for (int i = 0; i < 3; ++i)
    {
        min[i] = bv->min_extents[i];                       /* :1488 */
        max[i] = bv->max_extents[i] - min[i];              /* :1489 (max holds the size) */
    }
    float* tmin = bv->transformed_min_extents;
    float* tmax = bv->transformed_max_extents;

    if (scene->entities[n] & ORC_CMP_BONE)                 /* :1494-1498 */
    {
        const orc_mat4* w = &scene->world_matrices[n];
        tmin[0] = tmax[0] = w->m[3];
        tmin[1] = tmax[1] = w->m[7];
        tmin[2] = tmax[2] = w->m[11];
        return;
    }

    for (int i = 0; i < 3; ++i)                            /* :1500-1501 */
    {
        tmax[i] = -FLT_MAX;
        tmin[i] = FLT_MAX;
    }

    for (int c = 0; c < 8; ++c)                            /* :1503-1509 */
    {
        float q[3], p[3];
        for (int i = 0; i < 3; ++i)
        {
            float s = max[i] * k_corners[c][i];
            q[i] = min[i] + s;
        }
        orc_transform_vector(p, &scene->world_matrices[n], q);
        for (int i = 0; i < 3; ++i)
        {
            tmax[i] = fmax_union(tmax[i], p[i]);
            tmin[i] = fmin_union(tmin[i], p[i]);
        }
    }

    float d[3];
    for (int i = 0; i < 3; ++i)
        d[i] = tmax[i] - tmin[i];
    float trad = sqrtf(dotn(d, d, 3)) * 0.5f;              /* :1512 mag(tmax - tmin) * 0.5f */
    bv->radius = trad;

    orc_pos_extent* pe = &scene->pos_extent[n];            /* :1515-1518 */
    for (int i = 0; i < 3; ++i)
    {
        float h = d[i] * 0.5f;
        pe->pos[i] = tmin[i] + h;
    }
    for (int i = 0; i < 3; ++i)
        pe->extent[i] = tmax[i] - pe->pos[i];
    pe->extent[3] = trad;

    if (!(scene->entities[n] & ORC_CMP_GEOMETRY))          /* :1520-1521 */
        return;

    for (int i = 0; i < 3; ++i)                            /* :1524-1525 */
    {
        ext_min[i] = fmin_union(tmin[i], ext_min[i]);
        ext_max[i] = fmax_union(tmax[i], ext_max[i]);
    }
}

void orc_update_bounds(orc_scene* scene)
{
    float* emin = &scene->renderable_extents[0];
    float* emax = &scene->renderable_extents[3];
    for (int i = 0; i < 3; ++i)                            /* :1482-1483 */
    {
        emin[i] = FLT_MAX;
        emax[i] = -FLT_MAX;
    }
    for (size_t n = 0; n < scene->num_entities; ++n)
        bounds_one(scene, n, emin, emax);
}

/* ------------------------------------------------------------------------------------------
 * update_scene loop C — ecs_scene.cpp:1529-1556
 * ---------------------------------------------------------------------------------------- */
void orc_expand_parent_extents(orc_scene* scene)
{
    for (intptr_t n = (intptr_t)scene->num_entities - 1; n > 0; --n)
    {
        if (!(scene->entities[n] & ORC_CMP_ALLOCATED))     /* :1531-1532 */
            continue;
        uint32_t p = scene->parents[n];
        if (p == (uint32_t)n)                              /* :1535-1536 */
            continue;

        float* parent_tmin = scene->bounding_volumes[p].transformed_min_extents;
        float* parent_tmax = scene->bounding_volumes[p].transformed_max_extents;
        float* tmin = scene->bounding_volumes[n].transformed_min_extents;
        float* tmax = scene->bounding_volumes[n].transformed_max_extents;

        if (scene->entities[p] & ORC_CMP_ANIM_CONTROLLER)  /* :1544-1550, pad = vec3f(0) */
        {
            for (int i = 0; i < 3; ++i)
            {
                float a = tmin[i] - 0.0f;
                float b = tmax[i] + 0.0f;
                parent_tmin[i] = fmin_union(parent_tmin[i], a);
                parent_tmax[i] = fmax_union(parent_tmax[i], b);
            }
        }
        else                                               /* :1551-1555 */
        {
            for (int i = 0; i < 3; ++i)
            {
                parent_tmin[i] = fmin_union(parent_tmin[i], tmin[i]);
                parent_tmax[i] = fmax_union(parent_tmax[i], tmax[i]);
            }
        }
    }
}

void orc_update_scene(orc_scene* scene)
{
    orc_update_transforms(scene);
    orc_update_bounds(scene);
    orc_expand_parent_extents(scene);
}

/* ------------------------------------------------------------------------------------------
 * stretchy buffer — core/pen/include/data_struct.h:22-72.  Header = {int cap, int count}
 * 8 bytes before the data pointer; grows to max(2*cap, count+n); new memory zeroed.
 * ---------------------------------------------------------------------------------------- */
int orc_sb_count(const uint32_t* sb)
{
    return sb ? ((const int*)sb - 2)[1] : 0;
}

void orc_sb_free(uint32_t* sb)
{
    if (sb)
        free((int*)sb - 2);
}

static void* sb_growf(void* arr, int increment, int itemsize)
{
    int start = arr ? ((int*)arr - 2)[1] : 0;
    int dbl_cur = arr ? 2 * ((int*)arr - 2)[0] : 0;
    int min_needed = start + increment;
    int m = dbl_cur > min_needed ? dbl_cur : min_needed;
    size_t total_size = (size_t)itemsize * (size_t)m + sizeof(int) * 2;
    int* p = (int*)realloc(arr ? (int*)arr - 2 : NULL, total_size);
    if (!p)
        abort();
    size_t preserve = sizeof(int) * 2 + (size_t)itemsize * (size_t)start;
    memset((char*)p + preserve, 0, total_size - preserve);
    if (!arr)
        p[1] = 0;
    p[0] = m;
    return p + 2;
}

void orc_sb_push(uint32_t** sb, uint32_t v)
{
    uint32_t* a = *sb;
    /* stb__sbneedgrow: a == 0 || count + 1 >= cap */
    if (a == NULL || ((int*)a - 2)[1] + 1 >= ((int*)a - 2)[0])
        a = (uint32_t*)sb_growf(a, 1, (int)sizeof(uint32_t));
    a[((int*)a - 2)[1]++] = v;
    *sb = a;
}

/* test helper: build a stretchy buffer holding a copy of src[0..n) in one step */
uint32_t* orc_sb_from_array(const uint32_t* src, uint32_t n)
{
    if (n == 0)
        return NULL;
    uint32_t* a = (uint32_t*)sb_growf(NULL, (int)n + 1, (int)sizeof(uint32_t));
    memcpy(a, src, sizeof(uint32_t) * (size_t)n);
    ((int*)a - 2)[1] = (int)n;
    return a;
}

/* ------------------------------------------------------------------------------------------
 * filter + cull — core/put/source/ecs/ecs_cull.cpp:21-105 (scalar: the live path,
 * ecs_scene.cpp:959-960)
 * ---------------------------------------------------------------------------------------- */
void orc_filter_entities_scalar(const orc_scene* scene, uint32_t** entities_out)
{
    uint32_t accept_entities = (uint32_t)(ORC_CMP_GEOMETRY | ORC_CMP_MATERIAL); /* :86 */
    uint32_t reject_entities = (uint32_t)ORC_CMP_SUB_INSTANCE;                  /* :87 */

    for (uint32_t i = 0; i < scene->num_entities; ++i)
    {
        if ((scene->entities[i] & accept_entities) != accept_entities)  /* :92-93 */
            continue;
        if (scene->state_flags[i] & ORC_STATE_HIDDEN)                   /* :95-96 */
            continue;
        if (reject_entities)                                            /* :99-101 */
            if (scene->entities[i] & reject_entities)
                continue;
        orc_sb_push(entities_out, i);                                   /* :103 */
    }
}

void orc_frustum_cull_aabb_scalar(const orc_scene* scene, const orc_frustum* frust,
                                  uint32_t* entities_in, uint32_t** entities_out)
{
    uint32_t n = (uint32_t)orc_sb_count(entities_in);
    for (uint32_t i = 0; i < n; ++i)
    {
        uint32_t e = entities_in[i];
        const float* pos = scene->pos_extent[e].pos;       /* :30 */
        const float* extent = scene->pos_extent[e].extent; /* :31 */

        int inside = 1;
        for (int p = 0; p < 6; ++p)                        /* :34-44, all six, no early out */
        {
            float sign_flip[3], c[3];
            for (int k = 0; k < 3; ++k)
                sign_flip[k] = fsgn(frust->n[p][k]) * -1.0f;          /* :36 */
            float pd = orc_plane_distance(frust->p[p], frust->n[p]);  /* :37 */
            for (int k = 0; k < 3; ++k)
            {
                float s = extent[k] * sign_flip[k];
                c[k] = pos[k] + s;
            }
            float d2 = dotn(c, frust->n[p], 3);                       /* :38 */
            if (d2 > -pd)                                             /* :40 */
                inside = 0;
        }
        if (inside)
            orc_sb_push(entities_out, e);                             /* :46-49 */
    }
}

void orc_frustum_cull_sphere_scalar(const orc_scene* scene, const orc_frustum* frust,
                                    uint32_t* entities_in, uint32_t** entities_out)
{
    uint32_t n = (uint32_t)orc_sb_count(entities_in);
    for (uint32_t i = 0; i < n; ++i)
    {
        uint32_t e = entities_in[i];
        const float* pos = scene->pos_extent[e].pos;       /* :62 */
        float radius = scene->pos_extent[e].extent[3];     /* :63 */

        int inside = 1;
        for (int p = 0; p < 6; ++p)                        /* :66-75 */
        {
            float d = orc_point_plane_distance(pos, frust->p[p], frust->n[p]);
            if (d > radius)
            {
                inside = 0;
                break;
            }
        }
        if (inside)
            orc_sb_push(entities_out, e);
    }
}

void orc_frame(orc_scene* scene, const orc_frustum* views, uint32_t num_views, int mode,
               uint32_t** lists_out)
{
    orc_update_scene(scene);
    for (uint32_t v = 0; v < num_views; ++v)
    {
        /* ecs_scene.cpp:957-960 + :1085-1093: per view, filter all entities, cull, free */
        uint32_t* filtered = NULL;
        uint32_t* culled = NULL;
        orc_filter_entities_scalar(scene, &filtered);
        if (mode == 0)
            orc_frustum_cull_aabb_scalar(scene, &views[v], filtered, &culled);
        else
            orc_frustum_cull_sphere_scalar(scene, &views[v], filtered, &culled);
        orc_sb_free(filtered);
        lists_out[v] = culled;
    }
}

/* ------------------------------------------------------------------------------------------
 * all-cores variant — same per-entity arithmetic; hierarchy by depth level (each level is
 * data-parallel because parents[n] <= n puts every parent at a strictly smaller depth);
 * min/max reductions are exact and order-independent; lists are built as per-thread
 * contiguous segments and concatenated in thread order = ascending entity order.
 * ---------------------------------------------------------------------------------------- */
int orc_frame_omp(orc_scene* scene, const orc_frustum* views, uint32_t num_views, int mode,
                  uint32_t** lists_out)
{
#ifndef _OPENMP
    orc_frame(scene, views, num_views, mode, lists_out);
    return 1;
#else
    const size_t N = scene->num_entities;
    int nthreads = omp_get_max_threads();

    /* depth levels (serial O(N) prefix: depth[n] = depth[parent] + 1) */
    uint32_t* depth = (uint32_t*)malloc(sizeof(uint32_t) * (N ? N : 1));
    uint32_t max_depth = 0;
    for (size_t n = 0; n < N; ++n)
    {
        uint32_t p = scene->parents[n];
        depth[n] = (p == n) ? 0u : depth[p] + 1u;
        if (depth[n] > max_depth)
            max_depth = depth[n];
    }
    size_t* level_start = (size_t*)calloc((size_t)max_depth + 2, sizeof(size_t));
    for (size_t n = 0; n < N; ++n)
        level_start[depth[n] + 1]++;
    for (uint32_t l = 0; l <= max_depth; ++l)
        level_start[l + 1] += level_start[l];
    uint32_t* order = (uint32_t*)malloc(sizeof(uint32_t) * (N ? N : 1));
    {
        size_t* cursor = (size_t*)malloc(sizeof(size_t) * ((size_t)max_depth + 1));
        memcpy(cursor, level_start, sizeof(size_t) * ((size_t)max_depth + 1));
        for (size_t n = 0; n < N; ++n)
            order[cursor[depth[n]]++] = (uint32_t)n;
        free(cursor);
    }

    /* loop A by level (stage-0 / flag side effects as in the serial version) */
    for (uint32_t l = 0; l <= max_depth; ++l)
    {
        size_t b = level_start[l], e = level_start[l + 1];
#pragma omp parallel for schedule(static)
        for (size_t i = b; i < e; ++i)
        {
            size_t n = order[i];
            if (scene->state_flags[n] & ORC_STATE_SYNC_PHYSICS_TRANSFORM)
            {
                scene->state_flags[n] &= ~ORC_STATE_SYNC_PHYSICS_TRANSFORM;
                scene->entities[n] &= ~ORC_CMP_TRANSFORM;
            }
            if (scene->entities[n] & ORC_CMP_TRANSFORM)
            {
                scene->entities[n] &= ~ORC_CMP_TRANSFORM; /* local taken as baked (no stage 0 here) */
            }
            else if (scene->entities[n] & ORC_CMP_PHYSICS)
            {
                int has = scene->has_rb_matrix ? scene->has_rb_matrix[n] : 0;
                if (!has)
                    continue;
            }
            uint32_t parent = scene->parents[n];
            if (parent == n)
                scene->world_matrices[n] = scene->local_matrices[n];
            else
                orc_mat4_mul(&scene->world_matrices[n], &scene->world_matrices[parent], &scene->local_matrices[n]);
        }
    }

    /* loop B with per-thread extents */
    float gmin[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, gmax[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
#pragma omp parallel
    {
        float lmin[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, lmax[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
#pragma omp for schedule(static) nowait
        for (size_t n = 0; n < N; ++n)
            bounds_one(scene, n, lmin, lmax);
#pragma omp critical
        for (int i = 0; i < 3; ++i)
        {
            gmin[i] = fmin_union(lmin[i], gmin[i]);
            gmax[i] = fmax_union(lmax[i], gmax[i]);
        }
    }
    for (int i = 0; i < 3; ++i)
    {
        scene->renderable_extents[i] = gmin[i];
        scene->renderable_extents[3 + i] = gmax[i];
    }

    /* loop C: deepest level first; children of one level fold into parents of shallower
     * levels; min/max exact, so per-parent order is irrelevant — but two children of one
     * parent may race, so this stays serial per level group (it is a small share). */
    orc_expand_parent_extents(scene);

    /* per view: filter + cull in parallel segments */
    for (uint32_t v = 0; v < num_views; ++v)
    {
        uint32_t** seg = (uint32_t**)calloc((size_t)nthreads, sizeof(uint32_t*));
#pragma omp parallel num_threads(nthreads)
        {
            int t = omp_get_thread_num();
            int nt = omp_get_num_threads();
            size_t b = N * (size_t)t / (size_t)nt, e = N * (size_t)(t + 1) / (size_t)nt;
            orc_scene sub = *scene;
            uint32_t* filtered = NULL;
            uint32_t accept = (uint32_t)(ORC_CMP_GEOMETRY | ORC_CMP_MATERIAL);
            for (size_t i = b; i < e; ++i)
            {
                if ((scene->entities[i] & accept) != accept) continue;
                if (scene->state_flags[i] & ORC_STATE_HIDDEN) continue;
                if (scene->entities[i] & ORC_CMP_SUB_INSTANCE) continue;
                orc_sb_push(&filtered, (uint32_t)i);
            }
            uint32_t* culled = NULL;
            if (mode == 0)
                orc_frustum_cull_aabb_scalar(&sub, &views[v], filtered, &culled);
            else
                orc_frustum_cull_sphere_scalar(&sub, &views[v], filtered, &culled);
            orc_sb_free(filtered);
            seg[t] = culled;
        }
        uint32_t* out = NULL;
        size_t total = 0;
        for (int t = 0; t < nthreads; ++t)
            total += (size_t)orc_sb_count(seg[t]);
        if (total)
        {
            out = (uint32_t*)sb_growf(NULL, (int)total + 1, (int)sizeof(uint32_t));
            size_t o = 0;
            for (int t = 0; t < nthreads; ++t)
            {
                int c = orc_sb_count(seg[t]);
                if (c)
                    memcpy(out + o, seg[t], sizeof(uint32_t) * (size_t)c);
                o += (size_t)c;
            }
            ((int*)out - 2)[1] = (int)total;
        }
        for (int t = 0; t < nthreads; ++t)
            orc_sb_free(seg[t]);
        free(seg);
        lists_out[v] = out;
    }

    free(order);
    free(level_start);
    free(depth);
    return nthreads;
#endif
}

/* ------------------------------------------------------------------------------------------
 * view / frustum producers — core/put/source/camera.cpp
 * The matrix constructors and unproject live in the absent maths repo; these follow the
 * GL-style right-handed conventions the tree implies (proj * view, NDC z in [-1,1]).  Their
 * rounding does not enter parity: a frustum is an *input* to both the oracle and the device.
 * ---------------------------------------------------------------------------------------- */
static void mat4_perspective(orc_mat4* o, float l, float r, float b, float t, float n, float f)
{
    memset(o, 0, sizeof(*o));
    o->m[0] = (2.0f * n) / (r - l);
    o->m[2] = (r + l) / (r - l);
    o->m[5] = (2.0f * n) / (t - b);
    o->m[6] = (t + b) / (t - b);
    o->m[10] = (-f - n) / (f - n);
    o->m[11] = (-(2.0f * n) * f) / (f - n);
    o->m[14] = -1.0f;
}

static void mat4_ortho(orc_mat4* o, float l, float r, float b, float t, float n, float f)
{
    memset(o, 0, sizeof(*o));
    o->m[0] = 2.0f / (r - l);
    o->m[3] = -(r + l) / (r - l);
    o->m[5] = 2.0f / (t - b);
    o->m[7] = -(t + b) / (t - b);
    o->m[10] = -2.0f / (f - n);
    o->m[11] = -(f + n) / (f - n);
    o->m[15] = 1.0f;
}

/* maths::unproject_sc inverts the view-projection with mat::inverse4x4 (binary32 Laplace expansion): the same
 * restatement the draw-call packing uses, which pmtech's stored records pin bit for bit */
static int mat4_inverse(orc_mat4* out, const orc_mat4* in)
{
    orc_inverse4x4(out, in);
    return 1;
}

static void v3_sub(float o[3], const float a[3], const float b[3])
{
    o[0] = a[0] - b[0]; o[1] = a[1] - b[1]; o[2] = a[2] - b[2];
}
static void v3_normalize(float o[3], const float a[3])
{
    float m = sqrtf(dotn(a, a, 3));
    o[0] = a[0] / m; o[1] = a[1] / m; o[2] = a[2] / m;
}
static void v3_cross(float o[3], const float a[3], const float b[3])
{
    float x = a[1] * b[2] - a[2] * b[1];
    float y = a[2] * b[0] - a[0] * b[2];
    float z = a[0] * b[1] - a[1] * b[0];
    o[0] = x; o[1] = y; o[2] = z;
}

/* camera.cpp:19-49 */
void orc_camera_create_perspective(orc_camera* cam, float fov_degrees, float aspect, float near_plane, float far_plane)
{
    float half = (fov_degrees * 3.14159265358979323846f / 180.0f) / 2.0f;
    float ny = 2.0f * tanf(half) * near_plane;
    float nx = ny * aspect;
    mat4_perspective(&cam->proj, -nx * 0.5f, nx * 0.5f, -ny * 0.5f, ny * 0.5f, near_plane, far_plane);
}

/* camera.cpp:51-57 */
void orc_camera_create_orthographic(orc_camera* cam, float left, float right, float bottom, float top, float znear, float zfar)
{
    mat4_ortho(&cam->proj, left, right, bottom, top, znear, zfar);
}

/* camera.cpp:316-319 */
void orc_camera_create_cubemap(orc_camera* cam, float near_plane, float far_plane)
{
    orc_camera_create_perspective(cam, 90.0f, 1.0f, near_plane, far_plane);
}

/* camera.cpp:321-360 */
void orc_camera_set_cubemap_face(orc_camera* cam, uint32_t face)
{
    static const float at[6][3] = {{-1, 0, 0}, {1, 0, 0}, {0, -1, 0}, {0, 1, 0}, {0, 0, 1}, {0, 0, -1}};
    static const float right[6][3] = {{0, 0, 1}, {0, 0, -1}, {1, 0, 0}, {1, 0, 0}, {1, 0, 0}, {-1, 0, -0.0f}};
    static const float up[6][3] = {{0, 1, 0}, {0, 1, 0}, {0, 0, 1}, {0, 0, -1}, {0, 1, 0}, {0, 1, 0}};
    orc_mat4 v, t;
    memset(&v, 0, sizeof(v));
    for (int i = 0; i < 3; ++i)
    {
        v.m[0 + i] = right[face][i];
        v.m[4 + i] = up[face][i];
        v.m[8 + i] = at[face][i];
    }
    v.m[15] = 1.0f;
    mat4_identity(&t);
    t.m[3] = -cam->pos[0];
    t.m[7] = -cam->pos[1];
    t.m[11] = -cam->pos[2];
    orc_mat4_mul(&cam->view, &v, &t);
}

/* a conventional right-handed look-at (camera_update_look_at, camera.cpp:236-262, builds
 * the same rows: right, up, -forward, then translates by -pos) */
void orc_camera_look_at(orc_camera* cam, const float pos[3], const float at[3], const float up_hint[3])
{
    float f[3], r[3], u[3], back[3];
    v3_sub(f, at, pos);
    v3_normalize(f, f);
    v3_cross(r, f, up_hint);
    v3_normalize(r, r);
    v3_cross(u, r, f);
    back[0] = -f[0]; back[1] = -f[1]; back[2] = -f[2];
    orc_mat4 v, t;
    memset(&v, 0, sizeof(v));
    for (int i = 0; i < 3; ++i)
    {
        v.m[0 + i] = r[i];
        v.m[4 + i] = u[i];
        v.m[8 + i] = back[i];
    }
    v.m[15] = 1.0f;
    mat4_identity(&t);
    t.m[3] = -pos[0];
    t.m[7] = -pos[1];
    t.m[11] = -pos[2];
    cam->pos[0] = pos[0]; cam->pos[1] = pos[1]; cam->pos[2] = pos[2];
    orc_mat4_mul(&cam->view, &v, &t);
}

/* maths::unproject_sc with viewport (1,1): screen [0,1]^2 x depth [0,1] -> NDC [-1,1]^3,
 * times inverse(view_proj), perspective divide */
static void unproject_sc(float out[3], const float scr[3], const orc_mat4* inv_view_proj)
{
    float ndc[4] = {scr[0] * 2.0f - 1.0f, scr[1] * 2.0f - 1.0f, scr[2] * 2.0f - 1.0f, 1.0f};
    float r[4];
    for (int i = 0; i < 4; ++i)
        r[i] = dotn(&inv_view_proj->m[i * 4], ndc, 4);
    out[0] = r[0] / r[3];
    out[1] = r[1] / r[3];
    out[2] = r[2] / r[3];
}

/* camera.cpp:131-172 */
void orc_camera_update_frustum(orc_camera* cam)
{
    static const float ndc_coords[4][2] = {{0.0f, 1.0f}, {1.0f, 1.0f}, {0.0f, 0.0f}, {1.0f, 0.0f}}; /* :133-138 */
    orc_mat4 view_proj, inv;
    orc_mat4_mul(&view_proj, &cam->proj, &cam->view);                                               /* :142 */
    mat4_inverse(&inv, &view_proj);
    orc_frustum* f = &cam->camera_frustum;
    for (int i = 0; i < 4; ++i)                                                                     /* :144-148 */
    {
        float s0[3] = {ndc_coords[i][0], ndc_coords[i][1], 0.0f};
        float s1[3] = {ndc_coords[i][0], ndc_coords[i][1], 1.0f};
        unproject_sc(f->corners[0][i], s0, &inv);
        unproject_sc(f->corners[1][i], s1, &inv);
    }
    const float* pv[18] = {
        /* :152-161 */
        f->corners[0][0], f->corners[1][0], f->corners[0][2], /* left   */
        f->corners[0][0], f->corners[0][1], f->corners[1][0], /* top    */
        f->corners[0][1], f->corners[0][3], f->corners[1][1], /* right  */
        f->corners[0][2], f->corners[1][2], f->corners[0][3], /* bottom */
        f->corners[0][0], f->corners[0][2], f->corners[0][1], /* near   */
        f->corners[1][0], f->corners[1][1], f->corners[1][2]  /* far    */
    };
    for (int i = 0; i < 6; ++i)                                                                     /* :163-171 */
    {
        float v1[3], v2[3];
        v3_sub(v1, pv[i * 3 + 1], pv[i * 3 + 0]);
        v3_normalize(v1, v1);
        v3_sub(v2, pv[i * 3 + 2], pv[i * 3 + 0]);
        v3_normalize(v2, v2);
        v3_cross(f->n[i], v1, v2);      /* not renormalised, :168 */
        f->p[i][0] = pv[i * 3][0];
        f->p[i][1] = pv[i * 3][1];
        f->p[i][2] = pv[i * 3][2];
    }
}

/* shadow_camera_from_entity, spot branch — ecs_scene.cpp:556-570 */
void orc_camera_spot_shadow(orc_camera* cam, const orc_mat4* world)
{
    orc_camera_create_perspective(cam, 100.0f, 1.0f, 0.1f, 500.0f);          /* :558 */
    const float* w = world->m;
    const float  c2[3] = {w[2], w[6], w[10]}, c0[3] = {w[0], w[4], w[8]}, c1[3] = {w[1], w[5], w[9]};
    float        r0[3], r1[3], r2[3];
    v3_normalize(r0, c2);                                                    /* :560 row 0 = normalize(column 2) */
    v3_normalize(r1, c0);                                                    /* :561 row 1 = normalize(column 0) */
    v3_normalize(r2, c1);                                                    /* :562 row 2 = normalize(column 1) */
    orc_mat4 v, t;
    memset(&v, 0, sizeof(v));
    for (int i = 0; i < 3; ++i)
    {
        v.m[0 + i] = r0[i];
        v.m[4 + i] = r1[i];
        v.m[8 + i] = r2[i];
    }
    v.m[15] = 1.0f;                                                          /* :563 */
    mat4_identity(&t);                                                       /* :565 create_translation(-get_translation()) */
    t.m[3] = -w[3];
    t.m[7] = -w[7];
    t.m[11] = -w[11];
    orc_mat4_mul(&cam->view, &v, &t);                                        /* :567 */
    orc_camera_update_frustum(cam);                                          /* :569 */
}

/* camera.cpp:362-416 */
void orc_camera_update_shadow_frustum(orc_camera* cam, const float light_dir[3], const float mn[3], const float mx[3])
{
    static const float offsets[8][3] = {{0, 0, 0}, {1, 1, 1}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {1, 0, 1}, {1, 1, 0}, {0, 1, 1}}; /* :365-374 */
    float unit_y[3] = {0.0f, 1.0f, 0.0f};
    float right[3], up[3];
    v3_cross(right, light_dir, unit_y);                 /* :387 */
    v3_cross(up, right, light_dir);                     /* :388 */
    orc_mat4 sv;                                        /* :390-391 set_vectors(right, up, -light_dir, 0) */
    memset(&sv, 0, sizeof(sv));
    for (int i = 0; i < 3; ++i)
    {
        sv.m[0 + i] = right[i];
        sv.m[4 + i] = up[i];
        sv.m[8 + i] = -light_dir[i];
    }
    sv.m[15] = 1.0f;
    float cmin[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, cmax[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = 0; i < 8; ++i)                         /* :394-409 */
    {
        float c[3], p[3];
        for (int k = 0; k < 3; ++k)
            c[k] = mn[k] + offsets[i][k] * (mx[k] - mn[k]);
        orc_transform_vector(p, &sv, c);
        p[2] *= -1.0f;
        for (int k = 0; k < 3; ++k)
        {
            cmin[k] = fmin_union(cmin[k], p[k]);
            cmax[k] = fmax_union(cmax[k], p[k]);
        }
    }
    cam->view = sv;                                     /* :412-414 */
    mat4_ortho(&cam->proj, cmin[0], cmax[0], cmin[1], cmax[1], cmin[2], cmax[2]);
    orc_camera_update_frustum(cam);                     /* :416 */
}
