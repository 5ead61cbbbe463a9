// shim_driver.cpp — calls include/put_ecs_b200.hpp the way pmtech's own call sites do
// (ecs::update -> update_scene, ecs_scene.cpp:1322-1339; render_scene_view -> filter + cull,
// ecs_scene.cpp:957-960,1085-1093) on a mock ecs_scene built over caller arrays.  Exposed as C
// functions so tests/test_shim_gpu.py can drive it with numpy arrays and compare with the oracle.
#include <stdlib.h>
#include <string.h>

#include "mock_pmtech.h"

#include "../../include/put_ecs_b200.hpp"

using namespace put;
using namespace put::ecs;

static ecs_scene g_scene;

extern "C" void shim_bind(u32 soa_size, u32 num_entities, u64* entities, u64* state_flags, u32* parents, f32* local,
                          f32* world, f32* bounding_volumes, f32* pos_extent)
{
    g_scene.soa_size = soa_size;
    g_scene.num_entities = num_entities;
    g_scene.entities.data = entities;
    g_scene.state_flags.data = state_flags;
    g_scene.parents.data = parents;
    g_scene.local_matrices.data = (mat4*)local;
    g_scene.world_matrices.data = (mat4*)world;
    g_scene.bounding_volumes.data = (cmp_bounding_volume*)bounding_volumes;
    g_scene.pos_extent.data = (cmp_pos_extent*)pos_extent;
}

extern "C" void shim_update(f32* renderable_extents_out)
{
    b200::update_scene(&g_scene, 1.0f / 60.0f);
    memcpy(renderable_extents_out, &g_scene.renderable_extents, 6 * sizeof(f32));
}

// the resident form: what a pmtech build that knows its edits passes (include/put_ecs_b200.hpp, frame_options)
extern "C" void shim_update_opts(f32* renderable_extents_out, const uint8_t* has_rb_matrix, int full_upload, const u32* dirty_first_count,
                                 u32 num_dirty, const f32* transforms, int mark_dirty_on_device, int download_mask)
{
    b200::frame_options opt;
    opt.has_rb_matrix = has_rb_matrix;
    opt.full_upload = full_upload != 0;
    opt.dirty = (const b200::dirty_range*)dirty_first_count;
    opt.num_dirty = num_dirty;
    opt.transforms = transforms;
    opt.mark_dirty_on_device = mark_dirty_on_device != 0;
    opt.download_world = (download_mask & 1) != 0;
    opt.download_bounds = (download_mask & 2) != 0;
    opt.download_pos_extent = (download_mask & 4) != 0;
    opt.download_flags = (download_mask & 8) != 0;
    b200::update_scene(&g_scene, 1.0f / 60.0f, opt);
    memcpy(renderable_extents_out, &g_scene.renderable_extents, 6 * sizeof(f32));
}

// all views of the frame in one device pass (pmfx knows them before it walks them)
extern "C" void shim_prepare_views(const f32* frusta60, u32 num_views, int sphere)
{
    camera*        cams = (camera*)calloc(num_views ? num_views : 1, sizeof(camera));
    const camera** ptrs = (const camera**)malloc(sizeof(camera*) * (num_views ? num_views : 1));
    for (u32 v = 0; v < num_views; ++v)
    {
        memcpy(&cams[v].camera_frustum, frusta60 + (size_t)v * 60, sizeof(frustum));
        ptrs[v] = &cams[v];
    }
    b200::prepare_views(&g_scene, ptrs, num_views, sphere ? PMB_CULL_SPHERE : PMB_CULL_AABB);
    free(ptrs);
    free(cams);
}

// one render_scene_view: filter, cull, copy the list out, free both lists like the reference does
extern "C" int shim_render_scene_view(const f32* frustum60, int sphere, u32* list_out, int capacity)
{
    camera cam;
    memset(&cam, 0, sizeof(cam));
    memcpy(&cam.camera_frustum, frustum60, sizeof(frustum));
    u32* filtered_entities = nullptr;
    b200::filter_entities_scalar(&g_scene, &filtered_entities);
    u32* culled_entities = nullptr;
    if (sphere)
        b200::frustum_cull_sphere(&g_scene, &cam, filtered_entities, &culled_entities);
    else
        b200::frustum_cull_aabb(&g_scene, &cam, filtered_entities, &culled_entities);
    int n = sb_count(culled_entities);
    if (n <= capacity)
        for (int i = 0; i < n; ++i)
            list_out[i] = culled_entities[i];
    sb_free(filtered_entities);
    sb_free(culled_entities);
    return n;
}

// render_scene_view's filter + cull as one call (b200::culled_entities): no per-view copy of the filter list
extern "C" int shim_view_list(const f32* frustum60, int sphere, u32* list_out, int capacity)
{
    camera cam;
    memset(&cam, 0, sizeof(cam));
    memcpy(&cam.camera_frustum, frustum60, sizeof(frustum));
    u32* culled = nullptr;
    b200::culled_entities(&g_scene, &cam, &culled, sphere ? PMB_CULL_SPHERE : PMB_CULL_AABB);
    int n = sb_count(culled);
    if (n <= capacity && list_out)
        memcpy(list_out, culled, sizeof(u32) * (size_t)n);
    sb_free(culled);
    return n;
}

// a caller-supplied (not ascending) list, culled in the caller's order (ecs_cull.cpp:29-49)
extern "C" int shim_cull_list(const f32* frustum60, const u32* in, int nin, u32* list_out)
{
    camera cam;
    memset(&cam, 0, sizeof(cam));
    memcpy(&cam.camera_frustum, frustum60, sizeof(frustum));
    int* raw = (int*)malloc(sizeof(int) * 2 + sizeof(u32) * (size_t)(nin + 1));
    raw[0] = nin + 1;
    raw[1] = nin;
    u32* sb_in = (u32*)(raw + 2);
    memcpy(sb_in, in, sizeof(u32) * (size_t)nin);
    u32* out = nullptr;
    b200::frustum_cull_aabb_scalar(&g_scene, &cam, sb_in, &out);
    int n = sb_count(out);
    for (int i = 0; i < n; ++i)
        list_out[i] = out[i];
    sb_free(sb_in);
    sb_free(out);
    return n;
}

// render_shadow_views for `num` directional lights: cameras fitted on the device, lists out (ecs_scene.cpp:538-554, :573-663)
extern "C" int shim_dir_shadow_view(const f32* direction3, u32* list_out, int capacity, f32* view_proj_frustum92)
{
    camera cam;
    memset(&cam, 0, sizeof(cam));
    vec3f  dir = {direction3[0], direction3[1], direction3[2]};
    u32*   culled_entities = nullptr;
    b200::cull_dir_shadow_views(&g_scene, &dir, 1, (const extents*)nullptr, &cam, &culled_entities);
    memcpy(view_proj_frustum92, &cam.view, 64);
    memcpy(view_proj_frustum92 + 16, &cam.proj, 64);
    memcpy(view_proj_frustum92 + 32, &cam.camera_frustum, sizeof(frustum));
    int n = sb_count(culled_entities);
    if (n <= capacity)
        for (int i = 0; i < n; ++i)
            list_out[i] = culled_entities[i];
    sb_free(culled_entities);
    return n;
}

extern "C" void shim_release() { b200::release(&g_scene); }
