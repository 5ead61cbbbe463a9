/*
 * pm_oracle.h — CPU ORACLE for the pmtech ECS scene-update hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may build, load or call anything
 * under oracle/.  The product path (pmtech_b200/) never links or imports it.
 *
 * PINNING: the reference's arithmetic lives in the un-vendored submodule third_party/maths
 * (github.com/polymonster/maths @ ce54dd670760a191fee629a138ee43260ba57940), absent from
 * /root/reference, and its tests are golden *images* through a graphics API.  What it does hold
 * are two saved scenes (examples/assets/scene/sdf.pms, lucy.obj.pms): save_scene dumps the
 * component arrays as pmtech's last update_scene left them, inputs and outputs alike.  Extracted
 * into tests/golden/pmtech_{sdf,lucy}_scene.npz (tests/golden/make_pms_fixture.py), they pin —
 * bit for bit, tests/test_pmtech_scene_files.py — loop A (mat4 * mat4), loop B (corner transform,
 * min/max, radius), loop C (parent expansion), stage 0's T * R * S with the quaternion expansion,
 * and the draw-call records with mat::inverse4x4's -0 pattern.
 * STILL UNPINNED (declared): the plane tests of ecs_cull.cpp:21-82 (dot = `r = 0; r += a[i]*b[i]`
 * left to right, plane_distance = dot * -1, no FMA contraction), the expansion for rotations
 * about x and z, whether quat::get_matrix normalises in place or not at all (see pm_oracle.c),
 * and the camera producers.  No reference-computed vector reaches those.
 *
 * Citations are relative to /root/reference/.
 */
#ifndef PM_ORACLE_H
#define PM_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* core/put/source/ecs/ecs_scene.h:144-171 (e_cmp) — the bits the path reads */
#define ORC_CMP_ALLOCATED       (1ull << 0)
#define ORC_CMP_GEOMETRY        (1ull << 1)
#define ORC_CMP_PHYSICS         (1ull << 2)
#define ORC_CMP_MATERIAL        (1ull << 4)
#define ORC_CMP_BONE            (1ull << 7)
#define ORC_CMP_ANIM_CONTROLLER (1ull << 9)
#define ORC_CMP_TRANSFORM       (1ull << 12)
#define ORC_CMP_SUB_INSTANCE    (1ull << 14)
/* core/put/source/ecs/ecs_scene.h:69-83 (e_state) */
#define ORC_STATE_HIDDEN                 (1ull << 2)
#define ORC_STATE_SYNC_PHYSICS_TRANSFORM (1ull << 7)

/* mat4: 16 contiguous f32, row-major storage, column-vector convention
 * (physics/physics_bullet.cpp:310-315; translation = m[3],m[7],m[11]). */
typedef struct orc_mat4 { float m[16]; } orc_mat4;

/* maths::transform as used by cmp_transform (ecs_scene.h:347): translation, rotation
 * (quat x,y,z,w), scale = 40 B. */
typedef struct orc_transform {
    float translation[3];
    float rotation[4];
    float scale[3];
} orc_transform;

/* ecs_scene.h:248-255 — 52 B AoS record, inputs and outputs interleaved */
typedef struct orc_bounding_volume {
    float min_extents[3];
    float max_extents[3];
    float transformed_min_extents[3];
    float transformed_max_extents[3];
    float radius;
} orc_bounding_volume;

/* ecs_scene.h:257-261 */
typedef struct orc_pos_extent {
    float pos[4];    /* w unused */
    float extent[4]; /* w = radius */
} orc_pos_extent;

/* camera.h:41-46 */
typedef struct orc_frustum {
    float n[6][3];
    float p[6][3];
    float corners[2][4][3];
} orc_frustum;

/* The slice of put::ecs::ecs_scene (ecs_scene.h:460-543) the path touches. */
typedef struct orc_scene {
    size_t               num_entities;
    uint64_t*            entities;
    uint64_t*            state_flags;
    uint32_t*            parents;
    orc_transform*       transforms;        /* NULL: stage 0 (TRS bake) not exercised */
    orc_mat4*            local_matrices;
    orc_mat4*            world_matrices;
    orc_bounding_volume* bounding_volumes;
    orc_pos_extent*      pos_extent;
    float                renderable_extents[6]; /* min.xyz, max.xyz (ecs_scene.h:263-267,536) */
    const uint8_t*       has_rb_matrix;     /* stand-in for physics::has_rb_matrix(n); NULL = all 0 */
} orc_scene;

/* ---- maths restatements (exposed so tests can pin them on exact-arithmetic cases) ---- */
void  orc_mat4_mul(orc_mat4* out, const orc_mat4* a, const orc_mat4* b);
void  orc_transform_vector(float out[3], const orc_mat4* m, const float v[3]);
float orc_dot3(const float a[3], const float b[3]);
float orc_plane_distance(const float x0[3], const float xN[3]);
float orc_point_plane_distance(const float p0[3], const float x0[3], const float xN[3]);
void  orc_quat_normalise(float q[4]);
void  orc_quat_expand(orc_mat4* out, const float q[4]);
void  orc_quat_get_matrix(orc_mat4* out, float q[4]); /* normalises q in place */

/* ---- update_scene: loops A, B, C (ecs_scene.cpp:1405-1556) ---- */
void orc_update_scene(orc_scene* scene);
/* individual stages, same order as the reference runs them */
void orc_update_transforms(orc_scene* scene);        /* :1405-1471 */
void orc_update_bounds(orc_scene* scene);            /* :1474-1526 */
void orc_expand_parent_extents(orc_scene* scene);    /* :1529-1556 */

/* ---- animation hand-off around update_scene ----
 * root motion, the tail of the animation controller (ecs_scene.cpp:1286-1304): for rig i,
 *   rot_mat = get_matrix(rotations[i]); transforms[p].rotation = rotations[i];
 *   transforms[p].translation += rot_mat.transform_vector(deltas[i]); entities[p] |= e_cmp::transform
 * with p = targets[i].  (Whether quat::get_matrix normalises its quaternion in place is not provable
 * from the tree: the rotation is stored as given.) */
void orc_apply_root_motion(orc_scene* scene, const uint32_t* targets, const float* rotations, const float* deltas,
                           uint32_t count);
/* skin palettes (ecs_scene.cpp:1917-1959): bb[i] = world_matrices[joints_offset + i] * joint_bind_matrices[i] */
typedef struct orc_skin {
    uint32_t joints_offset; /* get_index_from_ref(root_joint_ref) + p_skin->bone_offset */
    uint32_t num_joints;
    uint32_t bind_offset;    /* first of this skin's joint_bind_matrices in `bind` */
    uint32_t palette_offset; /* first matrix written in `out` */
} orc_skin;
void orc_skin_palettes(const orc_scene* scene, const orc_skin* skins, uint32_t num_skins, const orc_mat4* bind, orc_mat4* out);

/* ---- draw-call packing (ecs_scene.cpp:1962-1991), the consumer of world_matrices ----
 * cmp_draw_call (ecs_scene.h:184-190): 160 B */
typedef struct orc_draw_call {
    orc_mat4 world_matrix;
    float    v1[4]; /* v1.x = entity index, v1.y = time in ms; z, w user data */
    float    v2[4]; /* user data */
    orc_mat4 world_matrix_inv_transpose;
} orc_draw_call;
#define ORC_CMP_SKINNED     (1ull << 6)
#define ORC_CMP_PRE_SKINNED (1ull << 16)
/* mat::inverse4x4 of the absent maths library: Laplace expansion over 2x2 sub-determinants.
 * UNVERIFIED recollection (operation order declared here). */
void orc_inverse4x4(orc_mat4* out, const orc_mat4* m);
/* cbuffer: scene->cbuffer handles (ecs_scene.h:493); invalid (-1) or null (0) entries stop after v1. */
void orc_update_draw_calls(const orc_scene* scene, const uint32_t* cbuffer, orc_draw_call* draw_call_data, float time_ms);

/* ---- stretchy buffer (core/pen/include/data_struct.h:22-72) ---- */
int      orc_sb_count(const uint32_t* sb);
void     orc_sb_free(uint32_t* sb);
void     orc_sb_push(uint32_t** sb, uint32_t v);
uint32_t* orc_sb_from_array(const uint32_t* src, uint32_t n); /* test helper */

/* ---- culling (ecs_cull.cpp:21-105, scalar = the live path) ---- */
void orc_filter_entities_scalar(const orc_scene* scene, uint32_t** filtered_entities_out);
void orc_frustum_cull_aabb_scalar(const orc_scene* scene, const orc_frustum* frust,
                                  uint32_t* entities_in, uint32_t** entities_out);
void orc_frustum_cull_sphere_scalar(const orc_scene* scene, const orc_frustum* frust,
                                    uint32_t* entities_in, uint32_t** entities_out);

/* ---- one frame as the reference runs it: update_scene then, per view, filter + cull
 * (ecs_scene.cpp:957-960 is reached once per view, pmfx_renderer.cpp:2953-3043).
 * lists_out[v] receives a stretchy buffer (caller orc_sb_free's).  mode 0 = aabb, 1 = sphere. */
void orc_frame(orc_scene* scene, const orc_frustum* views, uint32_t num_views, int mode,
               uint32_t** lists_out);

/* ---- all-cores variant of the same loops, same arithmetic (hierarchy by depth level,
 * per-thread list segments concatenated in order).  Not the reference's code path: the
 * reference is single threaded.  Used only as the "fair host" line in bench.py.
 * Returns the number of threads used. */
int orc_frame_omp(orc_scene* scene, const orc_frustum* views, uint32_t num_views, int mode,
                  uint32_t** lists_out);

/* ---- view / frustum producers (camera.cpp:19-57,131-172,321-360,384-416) ---- */
typedef struct orc_camera {
    float    pos[3];
    orc_mat4 view;
    orc_mat4 proj;
    orc_frustum camera_frustum;
} orc_camera;

void orc_set_num_threads(int n); /* OpenMP threads of orc_frame_omp */
void orc_camera_spot_shadow(orc_camera* cam, const orc_mat4* world); /* ecs_scene.cpp:556-570 */
void orc_camera_create_perspective(orc_camera* cam, float fov_degrees, float aspect, float near_plane, float far_plane);
void orc_camera_create_orthographic(orc_camera* cam, float left, float right, float bottom, float top, float znear, float zfar);
void orc_camera_create_cubemap(orc_camera* cam, float near_plane, float far_plane);
void orc_camera_set_cubemap_face(orc_camera* cam, uint32_t face);
void orc_camera_look_at(orc_camera* cam, const float pos[3], const float at[3], const float up[3]);
void orc_camera_update_frustum(orc_camera* cam);
void orc_camera_update_shadow_frustum(orc_camera* cam, const float light_dir[3], const float mn[3], const float mx[3]);

#ifdef __cplusplus
}
#endif
#endif
