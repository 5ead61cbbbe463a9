// put_ecs_b200.hpp — the reference-side binding: put::ecs entry points with pmtech's own
// signatures, implemented over the C ABI of pmtech_b200.h.  Header-only C++11; a pmtech maintainer
// includes it from core/put/source/ecs/ecs_scene.cpp / ecs_cull.cpp and links
// libpmtech_b200.so.  Templated on the scene / camera types only so that the same text also
// compiles against the small mock of those types used by tests/shim/ (the real headers need the
// maths submodule, which is absent from this sandbox).
//
//   put::ecs::b200::update_scene(scene, dt)                      replaces loops A-C of
//                                                                put::ecs::update_scene, ecs_scene.cpp:1405-1556
//   put::ecs::b200::filter_entities_scalar(scene, &out)          ecs_cull.h:16, ecs_cull.cpp:84-105
//   put::ecs::b200::frustum_cull_aabb_scalar(scene, cam, in, &out)    ecs_cull.h:17, ecs_cull.cpp:21-51
//   put::ecs::b200::frustum_cull_sphere_scalar(scene, cam, in, &out)  ecs_cull.h:18, ecs_cull.cpp:53-82
//   put::ecs::b200::frustum_cull_aabb / frustum_cull_sphere      ecs_cull.h:21-22 (dispatch to scalar, :533-543)
//   put::ecs::b200::cull_views(scene, cams, n, outs)             batched form: every view of the frame at once
//
// Semantics kept from the reference: void returns, failures assert (PEN_ASSERT = assert,
// console.h:120); output lists are stretchy buffers the caller sb_free's, left null when nothing is
// pushed (data_struct.h:22-72); host pointers are re-read from the cmp_arrays on every call because
// resize_scene_buffers may move them (ecs_scene.cpp:122-152); single user thread.
#ifndef PUT_ECS_B200_HPP
#define PUT_ECS_B200_HPP

#include <assert.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <map>

#include "pmtech_b200.h"

namespace put
{
    namespace ecs
    {
        namespace b200
        {
            struct scene_context
            {
                pmb_scene* dev = nullptr;
                uint32_t   soa_size = 0;
            };

            inline std::map<const void*, scene_context>& contexts()
            {
                static std::map<const void*, scene_context> s_contexts;
                return s_contexts;
            }

            inline void check(int rc)
            {
                (void)rc;
                assert(rc >= 0 && "pmtech_b200 call failed: see pmb_last_error()");
            }

            // device mirror of one ecs_scene; re-created when the scene's SoA capacity changes
            template <class Scene>
            pmb_scene* device_scene(const Scene* scene, int cuda_device = 0)
            {
                scene_context& c = contexts()[scene];
                if (!c.dev)
                {
                    check(pmb_scene_create(&c.dev, scene->soa_size, cuda_device));
                    c.soa_size = scene->soa_size;
                }
                else if (c.soa_size != scene->soa_size)
                {
                    check(pmb_scene_resize(c.dev, scene->soa_size));
                    c.soa_size = scene->soa_size;
                }
                return c.dev;
            }

            template <class Scene>
            void release(const Scene* scene)
            {
                auto it = contexts().find(scene);
                if (it == contexts().end())
                    return;
                pmb_scene_destroy(it->second.dev);
                contexts().erase(it);
            }

            // loops A-C of put::ecs::update_scene: uploads the arrays the loops read, runs the device
            // stage, reads back what the rest of the frame reads on the host.  (A tighter integration
            // uploads only dirty ranges and leaves world_matrices on the device for draw-call packing.)
            template <class Scene>
            void update_scene(Scene* scene, float dt)
            {
                (void)dt; // unused by these loops, as in the reference
                pmb_scene*     d = device_scene(scene);
                const uint32_t n = (uint32_t)scene->num_entities;
                check(pmb_upload(d, PMB_PARENTS, scene->parents.data, 0, n));
                check(pmb_upload(d, PMB_LOCAL_MATRICES, scene->local_matrices.data, 0, n));
                check(pmb_upload(d, PMB_ENTITIES, scene->entities.data, 0, n));
                check(pmb_upload(d, PMB_STATE_FLAGS, scene->state_flags.data, 0, n));
                check(pmb_upload(d, PMB_BOUNDING_VOLUMES, scene->bounding_volumes.data, 0, n));
                // outputs that loop A / B leave untouched for some entities (physics without an rb matrix keeps
                // its world matrix, bones keep radius and pos_extent) must hold the host's current values
                check(pmb_upload(d, PMB_WORLD_MATRICES, scene->world_matrices.data, 0, n));
                check(pmb_upload(d, PMB_POS_EXTENT, scene->pos_extent.data, 0, n));
                check(pmb_update_scene(d, n));
                check(pmb_download(d, PMB_WORLD_MATRICES, scene->world_matrices.data, 0, n));
                check(pmb_download(d, PMB_BOUNDING_VOLUMES, scene->bounding_volumes.data, 0, n));
                check(pmb_download(d, PMB_POS_EXTENT, scene->pos_extent.data, 0, n));
                check(pmb_download(d, PMB_ENTITIES, scene->entities.data, 0, n));
                check(pmb_download(d, PMB_STATE_FLAGS, scene->state_flags.data, 0, n));
                check(pmb_get_renderable_extents(d, reinterpret_cast<float*>(&scene->renderable_extents)));
            }

            template <class Camera>
            pmb_frustum to_frustum(const Camera* cam)
            {
                // camera.h:41-46: vec3f n[6], p[6], corners[2][4] — the same 240 bytes
                pmb_frustum f;
                static_assert(sizeof(cam->camera_frustum) == sizeof(pmb_frustum), "frustum layout");
                memcpy(&f, &cam->camera_frustum, sizeof(f));
                return f;
            }

            inline int sb_count_u32(const uint32_t* sb)
            {
                return sb ? ((const int*)sb)[-1] : 0; // stb__sbn, data_struct.h:31
            }

            // every view of the frame in one device pass; outs[v] receives a stretchy buffer (or stays
            // null when the view sees nothing), ascending entity order = the reference's sb_push order
            template <class Scene, class Camera>
            void cull_views(const Scene* scene, const Camera* const* cams, uint32_t num_views, uint32_t** outs,
                            int cull_mode = PMB_CULL_AABB)
            {
                pmb_scene*   d = device_scene(scene);
                pmb_frustum* fr = (pmb_frustum*)malloc(sizeof(pmb_frustum) * (num_views ? num_views : 1));
                for (uint32_t v = 0; v < num_views; ++v)
                    fr[v] = to_frustum(cams[v]);
                check(pmb_cull_views(d, (uint32_t)scene->num_entities, fr, num_views, cull_mode));
                check(pmb_get_lists_sb(d, outs, num_views));
                free(fr);
            }

            // render_shadow_views for directional lights (ecs_scene.cpp:573-663): shadow_camera_from_entity (:538-554)
            // fits each light's camera to scene->renderable_extents and the view is culled with it.  Here the fit runs
            // on the device, where the update left the extents: no read-back between update and shadow culls.
            // `light_directions` = cmp_light::direction per light; `constraints` = &scene->shadow_extent_constraints
            // ({min, max}) or null; cams_out[l] receives view, proj and camera_frustum for the shadow cbuffer.
            template <class Scene, class Camera, class Vec3, class Extents>
            void cull_dir_shadow_views(const Scene* scene, const Vec3* light_directions, uint32_t num_lights,
                                       const Extents* constraints, Camera* cams_out, uint32_t** outs, int cull_mode = PMB_CULL_AABB)
            {
                static_assert(sizeof(Vec3) == 12 && sizeof(Extents) == 24, "vec3f / extents layout");
                pmb_scene* d = device_scene(scene);
                check(pmb_cull_shadow_views(d, (uint32_t)scene->num_entities, (const float*)light_directions, num_lights,
                                            (const float*)constraints, cull_mode, 0));
                pmb_camera* pc = (pmb_camera*)malloc(sizeof(pmb_camera) * (num_lights ? num_lights : 1));
                check(pmb_download_shadow_cameras(d, pc, num_lights));
                for (uint32_t l = 0; l < num_lights; ++l)
                {
                    static_assert(sizeof(cams_out[l].view) == 64 && sizeof(cams_out[l].camera_frustum) == sizeof(pmb_frustum),
                                  "camera layout");
                    memcpy(&cams_out[l].view, pc[l].view, 64);
                    memcpy(&cams_out[l].proj, pc[l].proj, 64);
                    memcpy(&cams_out[l].camera_frustum, &pc[l].camera_frustum, sizeof(pmb_frustum));
                }
                free(pc);
                check(pmb_get_lists_sb(d, outs, num_lights));
            }

            template <class Scene>
            void filter_entities_scalar(const Scene* scene, uint32_t** filtered_entities_out)
            {
                pmb_scene*  d = device_scene(scene);
                pmb_frustum none;
                memset(&none, 0, sizeof(none));
                check(pmb_cull_views(d, (uint32_t)scene->num_entities, &none, 1, PMB_CULL_FILTER_ONLY));
                check(pmb_get_lists_sb(d, filtered_entities_out, 1));
            }

            // The reference culls the list it is handed and keeps its order.  The device culls every
            // filtered entity in ascending order; when `entities_in` is filter_entities_scalar's output
            // (the only call site, ecs_scene.cpp:959-960) the two coincide and the device list is
            // returned as is; any other input list is intersected with it, order preserved.
            template <class Scene, class Camera>
            void cull_list(const Scene* scene, const Camera* cam, uint32_t* entities_in, uint32_t** entities_out, int mode)
            {
                uint32_t*     vis = nullptr;
                const Camera* cams[1] = {cam};
                cull_views(scene, cams, 1, &vis, mode);
                const int nin = sb_count_u32(entities_in), nvis = sb_count_u32(vis);
                bool      ascending = true;
                for (int i = 1; i < nin && ascending; ++i)
                    ascending = entities_in[i - 1] < entities_in[i];
                // result = entities_in restricted to vis (vis is ascending: binary search)
                int* raw = (int*)malloc(sizeof(int) * 2 + sizeof(uint32_t) * (size_t)(nin + 1));
                uint32_t* out = (uint32_t*)(raw + 2);
                int       cnt = 0;
                for (int i = 0; i < nin; ++i)
                {
                    int lo = 0, hi = nvis;
                    while (lo < hi)
                    {
                        int mid = (lo + hi) / 2;
                        if (vis[mid] < entities_in[i])
                            lo = mid + 1;
                        else
                            hi = mid;
                    }
                    if (lo < nvis && vis[lo] == entities_in[i])
                        out[cnt++] = entities_in[i];
                }
                if (vis)
                    free((int*)vis - 2); // sb_free
                if (cnt == 0)
                {
                    free(raw);
                    return; // the reference leaves *entities_out untouched when nothing is pushed
                }
                raw[0] = nin + 1;
                raw[1] = cnt;
                *entities_out = out;
            }

            template <class Scene, class Camera>
            void frustum_cull_aabb_scalar(const Scene* scene, const Camera* cam, uint32_t* entities_in, uint32_t** entities_out)
            {
                cull_list(scene, cam, entities_in, entities_out, PMB_CULL_AABB);
            }

            template <class Scene, class Camera>
            void frustum_cull_sphere_scalar(const Scene* scene, const Camera* cam, uint32_t* entities_in, uint32_t** entities_out)
            {
                cull_list(scene, cam, entities_in, entities_out, PMB_CULL_SPHERE);
            }

            // ecs_cull.cpp:533-543: the public dispatchers are hard-wired to the scalar versions
            template <class Scene, class Camera>
            void frustum_cull_aabb(const Scene* scene, const Camera* cam, uint32_t* entities_in, uint32_t** entities_out)
            {
                frustum_cull_aabb_scalar(scene, cam, entities_in, entities_out);
            }

            template <class Scene, class Camera>
            void frustum_cull_sphere(const Scene* scene, const Camera* cam, uint32_t* entities_in, uint32_t** entities_out)
            {
                frustum_cull_sphere_scalar(scene, cam, entities_in, entities_out);
            }

            inline void simd_init() {} // ecs_cull.cpp:530-531: nothing to detect, the device path is the only one
        } // namespace b200
    }     // namespace ecs
} // namespace put

#endif // PUT_ECS_B200_HPP
