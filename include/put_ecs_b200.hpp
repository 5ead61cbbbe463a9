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
//   put::ecs::b200::update_scene(scene, dt, options)            the resident form: mirrors stay on the device, only the
//                                                                ranges the host dirtied go up, only what the host reads
//                                                                comes down, physics::has_rb_matrix as a byte array
//   put::ecs::b200::prepare_views(scene, cams, n)               all views of the frame in ONE device pass; the per-view
//                                                                filter / cull calls of render_scene_view are then served
//                                                                from a cache keyed on (frame, the 240 frustum bytes)
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
            // one cached draw list of the current frame: the frustum it was culled with, the mode, the entries
            struct cached_view
            {
                pmb_frustum frustum;
                int         mode;
                uint32_t    count;
                uint32_t*   entries; // count entries
                bool        owned;   // malloc'ed by this record (else: part of the context's page-locked slab)
            };

            struct scene_context
            {
                pmb_scene*   dev = nullptr;
                uint32_t     soa_size = 0;
                uint64_t     frame = 0;         // bumped by every update_scene: the cache belongs to one frame
                bool         resident = false;  // the mirrors hold this scene's arrays (first update done)
                cached_view* views = nullptr;   // the frame's cached lists (filter-only lists included)
                uint32_t     num_views = 0, cap_views = 0;
                uint32_t*    slab = nullptr;    // page-locked memory the first prepare_views of a frame reads its lists into
                size_t       slab_entries = 0;
                bool         slab_in_use = false;
            };

            struct dirty_range
            {
                uint32_t first, count;
            };

            // what update_scene moves between host and device.  The defaults re-send and read back everything, which is
            // what a binding without any knowledge of the host's edits has to do; a pmtech build passes what it knows.
            struct frame_options
            {
                // physics::has_rb_matrix(n) (physics.cpp:291-305) per entity as bytes, or null when the scene has no
                // rigid bodies: loop A rebuilds the world matrix of a physics entity only when this is set (:1443-1463)
                const uint8_t* has_rb_matrix = nullptr;
                // false: only `dirty` ranges of local_matrices (or transforms), entities and state_flags go up; parents,
                // bounds, world matrices and pos_extent are taken to be unchanged since the last update (the first
                // update of a scene, and any update after its capacity changed, always sends everything)
                bool full_upload = true;
                const dirty_range* dirty = nullptr;
                uint32_t           num_dirty = 0;
                // stage 0 on the device: the 40 B cmp_transform records of the dirty ranges go up instead of the 64 B
                // local matrices and local = T * R * S is baked there (ecs_scene.cpp:1415-1427); `transforms` =
                // scene->transforms.data (nullptr: local matrices travel).  The baked local matrices and the rotations
                // get_matrix normalises in place (:1421) then live in the device mirrors (PMB_LOCAL_MATRICES /
                // PMB_TRANSFORMS, pmb_download when the host wants them).
                const void* transforms = nullptr;
                // e_cmp::transform of the dirty ranges is set on the device instead of sending the 8 B entity words
                bool mark_dirty_on_device = false;
                // outputs the host reads after the update; renderable_extents always comes back
                bool download_world = true, download_bounds = true, download_pos_extent = true, download_flags = true;
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
                    c.resident = false; // resize_scene_buffers moved the host arrays: everything is sent again
                }
                return c.dev;
            }

            inline void clear_view_cache(scene_context& c)
            {
                for (uint32_t v = 0; v < c.num_views; ++v)
                    if (c.views[v].owned)
                        free(c.views[v].entries);
                c.num_views = 0;
                c.slab_in_use = false;
            }

            inline void cache_view(scene_context& c, const pmb_frustum& f, int mode, const uint32_t* entries, uint32_t count)
            {
                if (c.num_views == c.cap_views)
                {
                    c.cap_views = c.cap_views ? c.cap_views * 2 : 16;
                    c.views = (cached_view*)realloc(c.views, sizeof(cached_view) * c.cap_views);
                }
                cached_view& e = c.views[c.num_views++];
                e.frustum = f;
                e.mode = mode;
                e.count = count;
                e.owned = true;
                e.entries = (uint32_t*)malloc(sizeof(uint32_t) * (count ? count : 1));
                if (count && entries)
                    memcpy(e.entries, entries, sizeof(uint32_t) * count);
            }

            // the filter-only list ignores the planes: it is cached under an all-zero frustum
            inline const cached_view* find_view(const scene_context& c, const pmb_frustum& f, int mode)
            {
                for (uint32_t v = 0; v < c.num_views; ++v)
                    if (c.views[v].mode == mode && (mode == PMB_CULL_FILTER_ONLY || memcmp(&c.views[v].frustum, &f, sizeof(f)) == 0))
                        return &c.views[v];
                return nullptr;
            }

            inline int sb_count_u32(const uint32_t* sb)
            {
                return sb ? ((const int*)sb)[-1] : 0; // stb__sbn, data_struct.h:31
            }

            // appends `count` entries to a stretchy buffer the way a run of sb_push calls would (data_struct.h:40-72):
            // an existing list grows, a null one is created, nothing happens for an empty run
            inline void sb_append(uint32_t** list, const uint32_t* entries, uint32_t count)
            {
                if (!count)
                    return;
                const int old = sb_count_u32(*list);
                int*      raw = (int*)malloc(sizeof(int) * 2 + sizeof(uint32_t) * ((size_t)old + count + 1));
                raw[0] = old + (int)count + 1;
                raw[1] = old + (int)count;
                uint32_t* out = (uint32_t*)(raw + 2);
                if (old)
                    memcpy(out, *list, sizeof(uint32_t) * (size_t)old);
                memcpy(out + old, entries, sizeof(uint32_t) * count);
                if (*list)
                    free((int*)*list - 2); // sb_free
                *list = out;
            }

            // a stretchy buffer (data_struct.h:22-72) holding a copy of `entries`; null for an empty list, as the
            // reference leaves its output when nothing is pushed
            inline uint32_t* sb_copy(const uint32_t* entries, uint32_t count)
            {
                if (!count)
                    return nullptr;
                int* raw = (int*)malloc(sizeof(int) * 2 + sizeof(uint32_t) * ((size_t)count + 1));
                raw[0] = (int)count + 1;
                raw[1] = (int)count;
                memcpy(raw + 2, entries, sizeof(uint32_t) * count);
                return (uint32_t*)(raw + 2);
            }

            template <class Scene>
            void release(const Scene* scene)
            {
                auto it = contexts().find(scene);
                if (it == contexts().end())
                    return;
                clear_view_cache(it->second);
                free(it->second.views);
                pmb_host_free(it->second.slab);
                pmb_scene_destroy(it->second.dev);
                contexts().erase(it);
            }

            // loops A-C of put::ecs::update_scene: sends up what the loops read and the host changed, runs the device
            // stage, reads back what the rest of the frame reads on the host (frame_options; the defaults move
            // everything).  Starts a new frame: the cached draw lists of the previous one are dropped.
            template <class Scene>
            void update_scene(Scene* scene, float dt, const frame_options& opt = frame_options())
            {
                (void)dt; // unused by these loops, as in the reference
                pmb_scene*     d = device_scene(scene);
                scene_context& c = contexts()[scene];
                const uint32_t n = (uint32_t)scene->num_entities;
                ++c.frame;
                clear_view_cache(c);
                const uint64_t cmp_transform_bit = 1ull << 12; // e_cmp::transform (ecs_scene.h:144-171)
                if (opt.full_upload || !c.resident)
                {
                    check(pmb_upload(d, PMB_PARENTS, scene->parents.data, 0, n));
                    check(pmb_upload(d, PMB_LOCAL_MATRICES, scene->local_matrices.data, 0, n));
                    check(pmb_upload(d, PMB_ENTITIES, scene->entities.data, 0, n));
                    check(pmb_upload(d, PMB_STATE_FLAGS, scene->state_flags.data, 0, n));
                    check(pmb_upload(d, PMB_BOUNDING_VOLUMES, scene->bounding_volumes.data, 0, n));
                    // outputs that loop A / B leave untouched for some entities (physics without an rb matrix keeps
                    // its world matrix, bones keep radius and pos_extent) must hold the host's current values
                    check(pmb_upload(d, PMB_WORLD_MATRICES, scene->world_matrices.data, 0, n));
                    check(pmb_upload(d, PMB_POS_EXTENT, scene->pos_extent.data, 0, n));
                    if (opt.transforms)
                        check(pmb_upload(d, PMB_TRANSFORMS, opt.transforms, 0, n));
                    c.resident = true;
                }
                else
                {
                    for (uint32_t k = 0; k < opt.num_dirty; ++k)
                    {
                        const uint32_t first = opt.dirty[k].first;
                        uint32_t       count = opt.dirty[k].count;
                        if (first >= n)
                            continue;
                        if (count > n - first)
                            count = n - first;
                        if (opt.transforms)
                            check(pmb_upload(d, PMB_TRANSFORMS, (const char*)opt.transforms + (size_t)first * 40, first, count));
                        else
                            check(pmb_upload(d, PMB_LOCAL_MATRICES, scene->local_matrices.data + first, first, count));
                        if (opt.mark_dirty_on_device)
                            check(pmb_mark_transforms_dirty(d, first, count));
                        else
                        {
                            check(pmb_upload(d, PMB_ENTITIES, scene->entities.data + first, first, count));
                            check(pmb_upload(d, PMB_STATE_FLAGS, scene->state_flags.data + first, first, count));
                        }
                    }
                }
                if (opt.has_rb_matrix)
                    check(pmb_upload(d, PMB_HAS_RB_MATRIX, opt.has_rb_matrix, 0, n));
                check(pmb_update_scene(d, n));
                if (opt.download_world)
                    check(pmb_download(d, PMB_WORLD_MATRICES, scene->world_matrices.data, 0, n));
                if (opt.download_bounds)
                    check(pmb_download(d, PMB_BOUNDING_VOLUMES, scene->bounding_volumes.data, 0, n));
                if (opt.download_pos_extent)
                    check(pmb_download(d, PMB_POS_EXTENT, scene->pos_extent.data, 0, n));
                if (opt.download_flags)
                {
                    check(pmb_download(d, PMB_ENTITIES, scene->entities.data, 0, n));
                    check(pmb_download(d, PMB_STATE_FLAGS, scene->state_flags.data, 0, n));
                }
                else
                {
                    // what loop A does to the host's flags (ecs_scene.cpp:1408-1412, :1441) for the ranges that carried them
                    const uint64_t sync_physics_bit = 1ull << 7; // e_state::sync_physics_transform (ecs_scene.h:69-83)
                    if (opt.full_upload)
                        for (uint32_t i = 0; i < n; ++i)
                            scene->entities.data[i] &= ~cmp_transform_bit, scene->state_flags.data[i] &= ~sync_physics_bit;
                    else
                        for (uint32_t k = 0; k < opt.num_dirty; ++k)
                            for (uint32_t i = opt.dirty[k].first; i < n && i - opt.dirty[k].first < opt.dirty[k].count; ++i)
                                scene->entities.data[i] &= ~cmp_transform_bit, scene->state_flags.data[i] &= ~sync_physics_bit;
                }
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

            // Every view of the frame (pmfx knows them before it walks them, pmfx_renderer.cpp:2953-3043) in ONE device
            // pass, lists fetched once into the frame's cache.  The filter_entities_scalar / frustum_cull_* calls that
            // render_scene_view makes per view (ecs_scene.cpp:957-960) then cost a lookup and a copy instead of two
            // device passes and two synchronisations each.
            template <class Scene, class Camera>
            void prepare_views(const Scene* scene, const Camera* const* cams, uint32_t num_views, int cull_mode = PMB_CULL_AABB)
            {
                if (!num_views)
                    return;
                pmb_scene*     d = device_scene(scene);
                scene_context& c = contexts()[scene];
                const uint32_t n = (uint32_t)scene->num_entities;
                pmb_frustum*   fr = (pmb_frustum*)malloc(sizeof(pmb_frustum) * num_views);
                for (uint32_t v = 0; v < num_views; ++v)
                    fr[v] = to_frustum(cams[v]);
                uint32_t* counts = (uint32_t*)malloc(sizeof(uint32_t) * num_views);
                check(pmb_cull_views(d, n, fr, num_views, cull_mode));
                check(pmb_get_counts(d, counts, num_views));
                size_t total = 0;
                for (uint32_t v = 0; v < num_views; ++v)
                    total += counts[v];
                // The first batch of a frame lands in one page-locked slab (kept between frames, grown when needed): the
                // read-back runs at PCIe speed and is waited for once.  A further batch of the same frame, whose slab is
                // taken, falls back to one malloc'ed buffer per list.
                const bool use_slab = !c.slab_in_use;
                if (use_slab && total > c.slab_entries)
                {
                    pmb_host_free(c.slab);
                    c.slab_entries = total + total / 8 + 4096;
                    c.slab = (uint32_t*)pmb_host_alloc(c.slab_entries * sizeof(uint32_t));
                    assert(c.slab && "page-locked memory for the frame's draw lists");
                }
                uint32_t** dst = (uint32_t**)malloc(sizeof(uint32_t*) * num_views);
                size_t     at = 0;
                for (uint32_t v = 0; v < num_views; ++v)
                {
                    cache_view(c, fr[v], cull_mode, nullptr, counts[v]); // an owned buffer of counts[v] entries, filled below ...
                    cached_view& e = c.views[c.num_views - 1];
                    if (use_slab)
                    {
                        free(e.entries); // ... or a stretch of the slab
                        e.entries = c.slab + at;
                        e.owned = false;
                        at += counts[v];
                    }
                    dst[v] = e.entries;
                }
                if (use_slab)
                    c.slab_in_use = true;
                check(pmb_get_lists(d, dst, counts, nullptr, num_views));
                free(dst), free(counts), free(fr);
            }

            // render_scene_view's pair of calls (ecs_scene.cpp:959-960) as one: the entities of filter_entities_scalar's
            // list that survive the view's frustum, appended to *culled_entities_out like the reference's sb_push run.
            // Saves the host the per-view copy of the filter list, which only exists to be passed on.
            template <class Scene, class Camera>
            void culled_entities(const Scene* scene, const Camera* cam, uint32_t** culled_entities_out, int cull_mode = PMB_CULL_AABB)
            {
                scene_context&     c = contexts()[scene];
                const pmb_frustum  f = to_frustum(cam);
                const cached_view* hit = c.dev ? find_view(c, f, cull_mode) : nullptr;
                if (!hit)
                {
                    const Camera* one[1] = {cam};
                    prepare_views(scene, one, 1, cull_mode);
                    hit = &contexts()[scene].views[contexts()[scene].num_views - 1];
                }
                sb_append(culled_entities_out, hit->entries, hit->count);
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
                scene_context& c = contexts()[scene];
                pmb_frustum    none;
                memset(&none, 0, sizeof(none));
                const cached_view* hit = c.dev ? find_view(c, none, PMB_CULL_FILTER_ONLY) : nullptr;
                if (!hit)
                {
                    pmb_scene*     d = device_scene(scene);
                    const uint32_t n = (uint32_t)scene->num_entities;
                    check(pmb_cull_views(d, n, &none, 1, PMB_CULL_FILTER_ONLY));
                    uint32_t*     buf = (uint32_t*)malloc(sizeof(uint32_t) * ((size_t)n + 1));
                    const int64_t got = pmb_get_list(d, 0, buf, n);
                    check((int)(got < 0 ? got : 0));
                    cache_view(c, none, PMB_CULL_FILTER_ONLY, buf, (uint32_t)got);
                    free(buf);
                    hit = &c.views[c.num_views - 1];
                }
                sb_append(filtered_entities_out, hit->entries, hit->count);
            }

            // The reference culls the list it is handed and keeps its order.  The device culls every entity of
            // filter_entities_scalar's list in ascending order; when `entities_in` is that list (the only call site,
            // ecs_scene.cpp:959-960) the two coincide and the view's list is returned as is; any other input list is
            // intersected with it, order preserved.  Restriction (asserted): every entry of `entities_in` passes
            // filter_entities_scalar — the reference would plane-test an entry that does not, the device never saw it.
            template <class Scene, class Camera>
            void cull_list(const Scene* scene, const Camera* cam, uint32_t* entities_in, uint32_t** entities_out, int mode)
            {
                scene_context&     c = contexts()[scene];
                const pmb_frustum  f = to_frustum(cam);
                const cached_view* hit = c.dev ? find_view(c, f, mode) : nullptr;
                if (!hit)
                {
                    pmb_scene*     d = device_scene(scene);
                    const uint32_t n = (uint32_t)scene->num_entities;
                    check(pmb_cull_views(d, n, &f, 1, mode));
                    uint32_t*     buf = (uint32_t*)malloc(sizeof(uint32_t) * ((size_t)n + 1));
                    const int64_t got = pmb_get_list(d, 0, buf, n);
                    check((int)(got < 0 ? got : 0));
                    cache_view(c, f, mode, buf, (uint32_t)got);
                    free(buf);
                    hit = &c.views[c.num_views - 1];
                }
                const uint32_t* vis = hit->entries;
                const int       nin = sb_count_u32(entities_in), nvis = (int)hit->count;
                // the common case: the input IS the filter list, whose survivors are the view's list
                const cached_view* filt = find_view(c, f, PMB_CULL_FILTER_ONLY);
                if (filt && (int)filt->count == nin && (nin == 0 || memcmp(filt->entries, entities_in, sizeof(uint32_t) * (size_t)nin) == 0))
                {
                    sb_append(entities_out, vis, (uint32_t)nvis);
                    return;
                }
                // result = entities_in restricted to vis (vis is ascending: binary search), in the caller's order
                uint32_t* out = (uint32_t*)malloc(sizeof(uint32_t) * ((size_t)nin + 1));
                uint32_t  cnt = 0;
                for (int i = 0; i < nin; ++i)
                {
                    const uint32_t e = entities_in[i];
                    assert(e < (uint32_t)scene->num_entities && (scene->entities.data[e] & 0x12ull) == 0x12ull &&
                           !(scene->state_flags.data[e] & 0x4ull) && !(scene->entities.data[e] & (1ull << 14)) &&
                           "frustum_cull_*_scalar over the device takes entries of filter_entities_scalar's list only");
                    int lo = 0, hi = nvis;
                    while (lo < hi)
                    {
                        int mid = (lo + hi) / 2;
                        if (vis[mid] < e)
                            lo = mid + 1;
                        else
                            hi = mid;
                    }
                    if (lo < nvis && vis[lo] == e)
                        out[cnt++] = e;
                }
                sb_append(entities_out, out, cnt);
                free(out);
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
