// pm_camera.cpp — host-side view / frustum producers of the cull path (per view, not per entity).
//
// Mirrors core/put/source/camera.cpp: camera_create_perspective :19-49, camera_create_orthographic
// :51-57, camera_create_cubemap :316-319, camera_set_cubemap_face :321-360, camera_update_look_at
// :236-262, camera_update_frustum :131-172, camera_update_shadow_frustum :384-416.
// Conventions (row-major storage, column vectors, proj * view, GL clip space) are the ones the
// reference tree implies; the matrix constructors themselves live in its absent maths submodule.
// A frustum is an *input* of the device path, so none of this arithmetic enters kernel parity.
#include "../../include/pmtech_b200.h"

#include <cfloat>
#include <cmath>
#include <cstring>

#include "pm_camera_math.h"

using namespace pmcam;


extern "C" void pmb_camera_create_perspective(pmb_camera* cam, float fov_degrees, float aspect_ratio, float near_plane, float far_plane)
{
    // camera.cpp:19-49: near-plane rectangle from the vertical fov, then a frustum matrix
    float half_fov = (fov_degrees * 3.14159265358979323846f / 180.0f) / 2.0f;
    float ny = 2.0f * std::tan(half_fov) * near_plane;
    float nx = ny * aspect_ratio;
    proj_of(cam) = frustum_projection(-nx * 0.5f, nx * 0.5f, -ny * 0.5f, ny * 0.5f, near_plane, far_plane);
}

extern "C" void pmb_camera_create_orthographic(pmb_camera* cam, float left, float right, float bottom, float top, float znear, float zfar)
{
    proj_of(cam) = ortho_projection(left, right, bottom, top, znear, zfar); // camera.cpp:51-57
}

extern "C" void pmb_camera_create_cubemap(pmb_camera* cam, float near_plane, float far_plane)
{
    pmb_camera_create_perspective(cam, 90.0f, 1.0f, near_plane, far_plane); // camera.cpp:316-319
}

extern "C" void pmb_camera_set_cubemap_face(pmb_camera* cam, uint32_t face)
{
    // camera.cpp:324-349: +x -x +y -y +z -z
    static const v3 at[6] = {{-1, 0, 0}, {1, 0, 0}, {0, -1, 0}, {0, 1, 0}, {0, 0, 1}, {0, 0, -1}};
    static const v3 right[6] = {{0, 0, 1}, {0, 0, -1}, {1, 0, 0}, {1, 0, 0}, {1, 0, 0}, {-1, 0, -0.0f}};
    static const v3 up[6] = {{0, 1, 0}, {0, 1, 0}, {0, 0, 1}, {0, 0, -1}, {0, 1, 0}, {0, 1, 0}};
    face %= 6;
    view_of(cam) = mul(rows(right[face], up[face], at[face]), translation(-load3(cam->pos))); // :351-359
}

extern "C" void pmb_camera_look_at(pmb_camera* cam, const float pos[3], const float at[3], const float up_hint[3])
{
    v3 eye = load3(pos);
    v3 fwd = normalize(load3(at) - eye);
    v3 rgt = normalize(cross(fwd, load3(up_hint)));
    v3 up = cross(rgt, fwd);
    store3(cam->pos, eye);
    view_of(cam) = mul(rows(rgt, up, -fwd), translation(-eye));
}

extern "C" void pmb_camera_update_frustum(pmb_camera* cam) { pmcam::update_frustum(cam); } // camera.cpp:131-172

extern "C" void pmb_camera_update_shadow_frustum(pmb_camera* cam, const float light_dir[3], const float mn[3], const float mx[3])
{
    pmcam::update_shadow_frustum(cam, light_dir, mn, mx); // camera.cpp:384-416
}

extern "C" void pmb_camera_spot_shadow(pmb_camera* cam, const float world_matrix[16])
{
    pmb_camera_create_perspective(cam, 100.0f, 1.0f, 0.1f, 500.0f); // ecs_scene.cpp:558
    const m4 proj = proj_of(cam);
    pmcam::spot_shadow(cam, proj, world_matrix);                    // :560-569
}
