// mock_pmtech.h — the slice of pmtech's types that include/put_ecs_b200.hpp touches, with the
// reference's member names and record layouts (core/put/source/ecs/ecs_scene.h:248-267,405-412,
// 460-543; core/put/source/camera.h:41-73; core/pen/include/types.h:14-30).  Test scaffolding: the
// real headers need third_party/maths, which is absent from this sandbox.
#pragma once
#include <stddef.h>
#include <stdint.h>

typedef uint32_t u32;
typedef uint64_t u64;
typedef float    f32;

struct vec2f { f32 x, y; };
struct vec3f { f32 x, y, z; };
struct vec4f { f32 x, y, z, w; };
struct mat4  { f32 m[16]; };

namespace put
{
    struct frustum // camera.h:41-46
    {
        vec3f n[6];
        vec3f p[6];
        vec3f corners[2][4];
    };

    struct camera // camera.h:55-73, the reference's members in the reference's order (name, a Str, left out)
    {
        vec3f   pos;
        vec3f   focus;
        vec2f   rot;
        mat4    view;
        mat4    proj;
        mat4    view_projection;
        vec2f   jitter;
        f32     fov;
        f32     aspect;
        f32     near_plane;
        f32     far_plane;
        f32     zoom;
        u32     cbuffer;
        u32     flags;
        frustum camera_frustum;
    };

    namespace ecs
    {
        template <typename T>
        struct cmp_array // ecs_scene.h:405-412
        {
            u32 size = sizeof(T);
            T*  data = nullptr;
        };

        struct cmp_bounding_volume // ecs_scene.h:248-255
        {
            vec3f min_extents;
            vec3f max_extents;
            vec3f transformed_min_extents;
            vec3f transformed_max_extents;
            f32   radius;
        };

        struct cmp_pos_extent // ecs_scene.h:257-261
        {
            vec4f pos;
            vec4f extent;
        };

        struct extents // ecs_scene.h:263-267
        {
            vec3f min;
            vec3f max;
        };

        struct ecs_scene // ecs_scene.h:460-543, the members of this path
        {
            cmp_array<u64>                 entities;
            cmp_array<u64>                 state_flags;
            cmp_array<u32>                 parents;
            cmp_array<mat4>                local_matrices;
            cmp_array<mat4>                world_matrices;
            cmp_array<cmp_bounding_volume> bounding_volumes;
            cmp_array<cmp_pos_extent>      pos_extent;
            size_t                         num_entities = 0;
            u32                            soa_size = 0;
            extents                        renderable_extents;
        };
    } // namespace ecs
} // namespace put

// core/pen/include/data_struct.h:22-32
#define sb_free(a) ((a) ? free(((int*)(void*)(a)) - 2), 0 : 0)
#define sb_count(a) ((a) ? (((int*)(void*)(a)) - 2)[1] : 0)
