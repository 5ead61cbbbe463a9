/*
 * pmtech_b200.h — C ABI of the B200-native replacement for pmtech's per-frame ECS scene
 * update hot path (put::ecs::update_scene loops A-C and the per-view filter + frustum cull +
 * draw-list build).  This is the drop-in boundary: plain pointers and sizes, no C++/torch
 * types.  The library (libpmtech_b200.so) owns device mirrors of the component arrays; host
 * code uploads the arrays it dirtied, runs the stage, and reads back what it needs.
 *
 * Every entry cites the reference interface it replaces (paths relative to the pmtech tree).
 * Layouts are the reference's own:
 *   parents          u32[N]                         core/put/source/ecs/ecs_scene.h:479
 *   local_matrices   mat4[N], 16 f32 row-major      ecs_scene.h:481
 *   world_matrices   mat4[N]                        ecs_scene.h:482
 *   entities         u64[N]  (e_cmp bits)           ecs_scene.h:144-171, :471
 *   state_flags      u64[N]  (e_state bits)         ecs_scene.h:69-83,  :472
 *   bounding_volumes cmp_bounding_volume[N], 52 B   ecs_scene.h:248-255
 *   pos_extent       cmp_pos_extent[N], 32 B        ecs_scene.h:257-261
 *   frustum          {vec3f n[6], p[6], corners[2][4]}   core/put/source/camera.h:41-46
 *
 * All functions returning int return PMB_OK (0) or a negative pmb_status; pmb_last_error()
 * gives the message for the calling thread.  A handle is single-threaded (the reference calls
 * this path from its one user thread, examples/code/example_common.h:107-127).
 */
#ifndef PMTECH_B200_H
#define PMTECH_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PMB_MAX_VIEWS 16u /* views culled per pass; larger batches are split internally */

typedef enum pmb_status {
    PMB_OK = 0,
    PMB_ERR_INVALID = -1, /* bad argument                                   */
    PMB_ERR_CUDA = -2,    /* a CUDA runtime call failed (see pmb_last_error) */
    PMB_ERR_NO_DEVICE = -3,
    PMB_ERR_OOM = -4,
    PMB_ERR_STATE = -5 /* call out of order (e.g. lists read before a cull)  */
} pmb_status;

/* component arrays mirrored on the device (put::ecs::ecs_scene, ecs_scene.h:460-543) */
typedef enum pmb_array {
    PMB_PARENTS = 0,          /* u32,  in            */
    PMB_LOCAL_MATRICES = 1,   /* 64 B, in            */
    PMB_WORLD_MATRICES = 2,   /* 64 B, out (in for physics entities without an rb matrix) */
    PMB_ENTITIES = 3,         /* u64,  in/out (e_cmp::transform is cleared by the update) */
    PMB_STATE_FLAGS = 4,      /* u64,  in/out (e_state::sync_physics_transform cleared)   */
    PMB_BOUNDING_VOLUMES = 5, /* 52 B AoS, min/max in, transformed min/max + radius out    */
    PMB_POS_EXTENT = 6,       /* 32 B, out (in for a cull without an update)               */
    PMB_HAS_RB_MATRIX = 7,    /* u8,  in: physics::has_rb_matrix(n) (physics.cpp:291-305); never uploaded = all 0 */
    PMB_TRANSFORMS = 8,       /* 40 B cmp_transform {vec3f translation; quat rotation (x,y,z,w); vec3f scale} (ecs_scene.h:347,
                                 :480): in.  Once uploaded, the update bakes local = T * R * S on the device for entities that
                                 carry e_cmp::transform (ecs_scene.cpp:1415-1427); never uploaded = local_matrices arrive baked */
    PMB_CBUFFER = 9,          /* u32 renderer handles (ecs_scene.h:493): in; invalid (-1) or null (0) = no draw-call constants */
    PMB_DRAW_CALLS = 10,      /* 160 B cmp_draw_call {mat4 world; vec4 v1; vec4 v2; mat4 world_inv_transpose}
                                 (ecs_scene.h:184-190): in/out (v1.zw and v2 are the user's); both allocated on first use, zeroed */
    PMB_ARRAY_COUNT = 11
} pmb_array;

/* cull mode: which reference function the lists reproduce */
typedef enum pmb_cull_mode {
    PMB_CULL_AABB = 0,   /* filter_entities_scalar + frustum_cull_aabb_scalar   (ecs_cull.cpp:84-105, :21-51) */
    PMB_CULL_SPHERE = 1, /* filter_entities_scalar + frustum_cull_sphere_scalar (ecs_cull.cpp:84-105, :53-82) */
    PMB_CULL_FILTER_ONLY = 2 /* filter_entities_scalar alone (ecs_cull.cpp:84-105); planes ignored */
} pmb_cull_mode;

/* camera.h:41-46 — 6 plane normals (outward, not renormalised), 6 points on the planes, 8 corners */
typedef struct pmb_frustum {
    float n[6][3];
    float p[6][3];
    float corners[2][4][3];
} pmb_frustum;

/* camera.h:55-73 — only the members the path and the view producers touch */
typedef struct pmb_camera {
    float       pos[3];
    float       view[16];
    float       proj[16];
    pmb_frustum camera_frustum;
} pmb_camera;

typedef struct pmb_scene pmb_scene; /* opaque: device mirrors + scratch + streams */

/* ---- lifetime ------------------------------------------------------------------------------
 * Replaces the host allocation of the component arrays for this path
 * (resize_scene_buffers, ecs_scene.cpp:122-152): `capacity` = scene->soa_size.  New device
 * memory is zeroed like pen::memory_zero does for the host arrays. */
int  pmb_scene_create(pmb_scene** out, uint32_t capacity, int cuda_device);
int  pmb_scene_resize(pmb_scene* s, uint32_t new_capacity); /* keeps contents, zeroes the tail */
void pmb_scene_destroy(pmb_scene* s);
uint32_t pmb_scene_capacity(const pmb_scene* s);
int  pmb_scene_device(const pmb_scene* s);

/* Tuning / test switches of a handle (defaults are what a host application wants).
 * PMB_OPT_HIERARCHY: which form of loop A (ecs_scene.cpp:1465-1470) the streaming pass runs.  AUTO picks by the
 * depth of the hierarchy: up to 4 levels every child re-multiplies its own short chain (nothing is shared between
 * tiles); deeper scenes run a wavefront, one product per entity, from a per-tile schedule built when the parents
 * change (rounds of entities whose parents are final; world matrices needed by later tiles are published).  Both give
 * the reference's bits.  CHAINS is honoured up to 16 levels.
 * PMB_OPT_FUSED_TAIL: 1 (default) = scenes of up to 1 Mi entities run the whole frame as ONE kernel launch
 * (loop C across tiles, flag clears and draw-list compaction behind a grid barrier) unless loop C across tiles is
 * heavy enough to deserve its own launch; 2 = one launch whenever the scene is small enough; 0 = always separate
 * launches. */
typedef enum pmb_option { PMB_OPT_HIERARCHY = 0, PMB_OPT_FUSED_TAIL = 1, PMB_OPT_TRACE = 2 } pmb_option;
typedef enum pmb_hierarchy_mode { PMB_HIERARCHY_AUTO = 0, PMB_HIERARCHY_CHAINS = 1, PMB_HIERARCHY_WAVEFRONT = 2 } pmb_hierarchy_mode;
int pmb_set_option(pmb_scene* s, int option, int value);
/* PMB_OPT_TRACE = 1: the wavefront form records, per 64-entity tile, {loop A start ns, end ns, SM | rounds of the tile's
 * schedule << 32, rounds that need nothing from other tiles | CTA << 32} (4 x u64) — the pen::timer of this path's tiles
 * (cf. ecs_scene.cpp:1401-1402); read back here. */
int pmb_download_trace(pmb_scene* s, uint64_t* host, uint32_t num_tiles);

/* Run on a caller-owned CUDA stream (cudaStream_t passed as void*); NULL = the handle's own. */
int pmb_set_stream(pmb_scene* s, void* cuda_stream);
int pmb_synchronize(pmb_scene* s);

/* ---- array transfer ------------------------------------------------------------------------
 * host points at element `first` of the reference-layout array (i.e. cmp_array<T>::data +
 * first); `count` elements are copied.  Uploading PMB_PARENTS invalidates the cached
 * hierarchy index (depth bands, child list), which the next update rebuilds on the device.
 * Asynchronous on the handle's stream for pinned host memory; pmb_download returns after the
 * data has landed. */
int pmb_upload(pmb_scene* s, int array_id, const void* host, uint32_t first, uint32_t count);
int pmb_download(pmb_scene* s, int array_id, void* host, uint32_t first, uint32_t count);
/* raw device pointer of a mirror for zero-copy consumers (draw-call packing, skinning).
 * PMB_BOUNDING_VOLUMES is split on the device: see pmb_device_bounds(). */
void* pmb_device_ptr(pmb_scene* s, int array_id);
/* device-side split of cmp_bounding_volume: in = f32[N][6] (min,max), out = f32[N][7]
 * (transformed min, transformed max, radius) */
int pmb_device_bounds(pmb_scene* s, const float** bounds_in, float** bounds_out);

/* pinned host memory helpers (page-locked buffers make pmb_upload/pmb_download DMA directly) */
void* pmb_host_alloc(size_t bytes);
void  pmb_host_free(void* p);

/* ---- stages 0-2: put::ecs::update_scene loops A, B, C (ecs_scene.cpp:1405-1556) -------------
 * world matrices through the parent hierarchy, world-space bounds, pos_extent, scene
 * renderable extents, parent bound expansion, and the flag clears of loop A.  The TRS bake
 * (ecs_scene.cpp:1417-1427) runs on the device when PMB_TRANSFORMS has been uploaded; otherwise
 * local_matrices are taken as baked by the host.  The physics branches (:1429-1438, :1448-1462)
 * stay on the host. */
int pmb_update_scene(pmb_scene* s, uint32_t num_entities);
/* scene->renderable_extents (ecs_scene.h:536): min.xyz, max.xyz.  Synchronises. */
int pmb_get_renderable_extents(pmb_scene* s, float out_min_max[6]);

/* ---- stages 3-4: per view filter_entities_scalar + frustum_cull_*_scalar (ecs_cull.cpp:21-105,
 * called once per view from render_scene_view, ecs_scene.cpp:957-960) — batched over views.
 * Builds one ascending index list per view on the device. */
int pmb_cull_views(pmb_scene* s, uint32_t num_entities, const pmb_frustum* views,
                   uint32_t num_views, int cull_mode);

/* ---- the whole frame in one pass: update_scene then the per-view culls, with pos_extent
 * consumed from registers.  Same results as pmb_update_scene + pmb_cull_views. */
int pmb_frame(pmb_scene* s, uint32_t num_entities, const pmb_frustum* views,
              uint32_t num_views, int cull_mode);

/* ---- the same cull for callers that must not allocate GPU memory ---------------------------------
 * (No counterpart in the reference: pmtech culls on the CPU and never meets the question.)
 * pmb_cull_views needs a pmb_scene, and a pmb_scene cudaMalloc's its mirrors, bitmaps, counters and lists.  Three
 * kinds of host cannot go that way: an engine that reserved its whole GPU memory budget in its own pools at
 * start-up (a later cudaMalloc fails, or eats memory the renderer had counted on); code in the middle of a frame
 * (cudaMalloc and cudaFree synchronise the whole device and take an unbounded time); and a plugin admitted into
 * someone else's CUDA context on the condition that it allocates nothing.  For them this entry point takes every
 * buffer from the caller and keeps its scratch — one visibility bitmap row per view plus the chunk counters of the
 * list compaction, about 2 MiB — in a static device array that is part of the library image: every GPU that loads
 * the library holds those 2 MiB until the process exits, used or not.  It calls no cudaMalloc / cudaFree or any
 * other allocating routine.
 *   inputs  (device memory, or any memory the GPU can read): entities u64[n] and state_flags u64[n]
 *           (ecs_scene.h:471-472), pos_extent 32 B records [n] as loop B left them (16-byte aligned; may be NULL for
 *           PMB_CULL_FILTER_ONLY); views = num_views host-side frusta, read before the call returns.
 *   outputs (device memory): lists u32[num_views][list_stride], list_stride >= num_entities, view v's ascending
 *           entity indices at lists + v * list_stride; counts u32[num_views].  Same contents as pmb_cull_views
 *           followed by pmb_get_list / pmb_get_counts on a scene holding the same arrays.
 *   limits  num_entities <= PMB_STATIC_MAX_ENTITIES and num_views <= PMB_MAX_VIEWS; anything larger is refused with
 *           PMB_ERR_INVALID and a message naming the limit (larger scenes: call it per range of entities and add the
 *           range's first index to the entries, or use a pmb_scene).
 * The work is queued on cuda_stream (a cudaStream_t; NULL = the default stream) of the CURRENT device of the calling
 * thread, and the call returns without waiting.  The first call on each device also looks up the device address of
 * the static array (cudaGetSymbolAddress, which may load the module on that device): make that first call during
 * start-up, before any stream capture begins.  There is ONE array per GPU, shared by every caller in the process:
 * two calls whose kernels can overlap on the same GPU (different streams, different threads) race on it and corrupt
 * each other's lists — keep them on one stream or order them with events.  Nothing has to be cleared between calls:
 * a call writes every scratch word before it reads it, so what the previous call left behind is never seen. */
#define PMB_STATIC_MAX_ENTITIES (1u << 20)
int pmb_cull_views_static(const uint64_t* entities_device, const uint64_t* state_flags_device, const float* pos_extent_device,
                          uint32_t num_entities, const pmb_frustum* views, uint32_t num_views, int cull_mode,
                          uint32_t* lists_device, uint64_t list_stride, uint32_t* counts_device, void* cuda_stream);

/* ---- draw lists ------------------------------------------------------------------------------
 * counts_out[v] = sb_count of view v's list.  Synchronises. */
int pmb_get_counts(pmb_scene* s, uint32_t* counts_out, uint32_t num_views);
/* copy view v's list (ascending entity indices) into caller memory; returns the count or <0 */
int64_t pmb_get_list(pmb_scene* s, uint32_t view, uint32_t* host_out, uint32_t capacity);
/* stretchy-buffer form (core/pen/include/data_struct.h:22-72): lists_out[v] is malloc'ed with
 * the {int cap, int count} header 8 bytes before the data so the caller's sb_count / sb_free
 * work; NULL when the view's list is empty (as when the reference never sb_push'es). */
int pmb_get_lists_sb(pmb_scene* s, uint32_t** lists_out, uint32_t num_views);
/* every list of the last cull into caller buffers (page-locked ones move at PCIe speed): lists_out[v] must hold
 * capacities[v] >= the view's count entries (else PMB_ERR_INVALID, nothing copied); counts_out may be NULL; the copies
 * are queued back to back and waited for once */
int pmb_get_lists(pmb_scene* s, uint32_t* const* lists_out, const uint32_t* capacities, uint32_t* counts_out, uint32_t num_views);
/* device pointer + count of view v's list for consumers that stay on the GPU */
const uint32_t* pmb_device_list(pmb_scene* s, uint32_t view);
/* device array u32[num_views] of list lengths (valid after the cull has run on the stream) */
const uint32_t* pmb_device_counts(pmb_scene* s);

/* ---- draw-call packing (ecs_scene.cpp:1962-1991), after the update: world matrix (identity for
 * skinned / pre-skinned meshes), v1.x = entity index, v1.y = time_ms (pen::get_time_ms() upstream),
 * inverse4x4(transposed(world)), skipping entities without a cbuffer handle and sub-instances as the
 * reference does.  The records are a headless device buffer (PMB_DRAW_CALLS; there is no renderer
 * here): a sub-instance block [n+1, n+1+num_instances) is the master's instance buffer (:1994-2011). */
int pmb_update_draw_calls(pmb_scene* s, uint32_t num_entities, float time_ms);
/* The instance-buffer loop behind it (ecs_scene.cpp:1994-2011): every e_cmp::master_instance entity without
 * e_cmp::custom_instance_buffer hands draw_call_data[n + 1 ..] — num_instances * instance_stride bytes — to its
 * instance buffer and the walk strides over its sub-instances.  pmb_instance_blocks is that walk over the host's
 * entity flags and cmp_master_instance records (ecs_scene.h:290-295; host only, no device touched): it fills
 * blocks[0 .. min(count, capacity)) and returns the number of blocks the scene has.  A block's bytes are
 * pmb_device_draw_calls(s) + byte_offset on the device, or come to the host with pmb_download_instance_block. */
typedef struct pmb_master_instance { /* cmp_master_instance */
    uint32_t num_instances;
    uint32_t instance_buffer;
    uint32_t instance_stride;
} pmb_master_instance;
typedef struct pmb_instance_block {
    uint32_t master;          /* entity index n of the master */
    uint32_t instance_buffer; /* renderer handle the bytes belong to */
    uint64_t byte_offset;     /* (n + 1) * sizeof(cmp_draw_call) into the PMB_DRAW_CALLS mirror */
    uint64_t bytes;           /* num_instances * instance_stride */
} pmb_instance_block;
int64_t pmb_instance_blocks(const uint64_t* entities, const pmb_master_instance* master_instances, uint32_t num_entities,
                            pmb_instance_block* blocks, uint32_t capacity);
const void* pmb_device_draw_calls(pmb_scene* s); /* the cmp_draw_call mirror on the device (NULL before its first use) */
int pmb_download_instance_block(pmb_scene* s, const pmb_instance_block* block, void* host_out); /* synchronises */

/* ---- animation hand-off: root motion in, skin palettes out -----------------------------------
 * pmb_apply_root_motion: the tail of the animation controller (ecs_scene.cpp:1286-1304) for `count`
 * rigs: transforms[targets[i]].rotation = rotations[i]; .translation += get_matrix(rotations[i])
 * .transform_vector(deltas[i]); entities[targets[i]] |= e_cmp::transform.  Host arrays: targets
 * u32[count] (distinct), rotations f32[count][4] (x,y,z,w), deltas f32[count][3].  Needs
 * PMB_TRANSFORMS uploaded.  Queued on the handle's stream ahead of the next update. */
int pmb_apply_root_motion(pmb_scene* s, const uint32_t* targets, const float* rotations, const float* deltas, uint32_t count);
/* Skin palettes (ecs_scene.cpp:1917-1959): for skinned mesh k and i < num_joints,
 *   palette[palette_offset + i] = world_matrices[joints_offset + i] * bind[bind_offset + i]
 * joints_offset = get_index_from_ref(root_joint_ref) + cmp_skin::bone_offset, num_joints <= 85 and
 * bind = cmp_skin::joint_bind_matrices (ecs_scene.h:192-199); meshes that share a skin share
 * bind_offset.  The palette buffer stands in for the 85-matrix bone cbuffers (a headless device
 * buffer: there is no renderer here); entries past num_joints are not written. */
typedef struct pmb_skin {
    uint32_t joints_offset;
    uint32_t num_joints;
    uint32_t bind_offset;    /* in matrices */
    uint32_t palette_offset; /* in matrices */
} pmb_skin;
int pmb_skins_set(pmb_scene* s, const pmb_skin* skins, uint32_t num_skins, const float* bind_matrices,
                  uint32_t num_bind_matrices, uint32_t num_palette_matrices);
int pmb_skin_palettes(pmb_scene* s); /* after pmb_update_scene / pmb_frame, on the handle's stream */
int pmb_download_palettes(pmb_scene* s, float* host, uint32_t first_matrix, uint32_t count); /* synchronises */
float* pmb_device_palettes(pmb_scene* s);

/* ---- multi-GPU list exchange helpers (new design: the reference is single-process; a scene too
 * large for one GPU is split into root-subtree ranges, one pmb_scene per GPU, and the per-view
 * lists are concatenated in rank order = ascending global entity index).
 * pmb_copy_segments moves up to PMB_MAX_SEGMENTS runs of u32 between device buffers in one launch
 * on the handle's stream, adding `add` to every value (local -> global entity index).  Offsets are
 * in elements.  pmb_pack_lists is the common case: view v's list (count taken from counts[v]) is
 * copied to dst + sum(counts[0..v)), with `add` applied. */
#define PMB_MAX_SEGMENTS 128u
typedef struct pmb_segment {
    uint64_t src_offset;
    uint64_t dst_offset;
    uint32_t count;
    uint32_t add;
} pmb_segment;
int pmb_copy_segments(pmb_scene* s, const uint32_t* src_device, uint32_t* dst_device, const pmb_segment* segments,
                      uint32_t num_segments);
int pmb_pack_lists(pmb_scene* s, const uint32_t* counts, uint32_t num_views, uint32_t add, uint32_t* dst_device);
/* The same exchange without a host round trip.  pmb_pack_row writes this shard's fixed-size row
 *   row[0 .. num_views) = list lengths, row[num_views ..) = the lists back to back with `add` applied
 * (lengths are read from the device; entries beyond row_stride are dropped, the header keeps the true
 * lengths).  After the rows of all ranks have been all-gathered as world x row_stride u32,
 * pmb_unpack_rows lays the segments out per view in rank order: view v's global list is
 * dst[meta[v] .. meta[v] + meta[num_views + v]); meta[2 * num_views] != 0 reports that a row or dst
 * was too small (nothing is written then: grow the buffers and repeat).  num_views <= PMB_MAX_VIEWS.
 * pmb_unpack_rows runs on cuda_stream (the stream the all-gather ran on; NULL = the handle's). */
int pmb_pack_row(pmb_scene* s, uint32_t num_views, uint32_t add, uint32_t* row_device, uint32_t row_stride);
int pmb_unpack_rows(pmb_scene* s, const uint32_t* rows_device, uint32_t world, uint32_t row_stride, uint32_t num_views,
                    uint32_t* dst_device, uint64_t dst_capacity, uint32_t* meta_device, void* cuda_stream);

/* The exchange as one push over peer memory (NVLink P2P stores; the buffers are the same symmetric
 * allocation on every rank, e.g. torch.distributed._symmetric_memory, their base addresses as seen from
 * THIS device passed in rank order).  Per frame, after pmb_pack_row:
 *   pmb_publish_counts  writes this rank's list lengths into every peer's table u32[world][PMB_MAX_VIEWS]
 *   (cross-GPU barrier: every rank's table is complete)
 *   pmb_push_row        sends each of this rank's view segments to its final place in every peer's
 *                       per-view global lists; meta as for pmb_unpack_rows
 *   (cross-GPU barrier: every rank's lists are complete)
 * world <= PMB_MAX_PEERS.  Both run on cuda_stream (NULL = the handle's). */
#define PMB_MAX_PEERS 16u
int pmb_publish_counts(pmb_scene* s, const uint32_t* row_device, uint32_t num_views, uint32_t rank, uint32_t world,
                       const uint64_t* peer_tables, void* cuda_stream);
int pmb_push_row(pmb_scene* s, const uint32_t* row_device, uint32_t row_stride, uint32_t num_views, uint32_t rank,
                 uint32_t world, const uint32_t* table_device, const uint64_t* peer_lists, uint64_t lists_capacity,
                 uint32_t* meta_device, void* cuda_stream);

/* ---- the exchange step as an object (what a multi-GPU host calls; the helpers above are its building blocks) -----
 * One pmb_exchange per rank (= per pmb_scene / GPU).  It owns a block of device memory {counts table, flags, result,
 * global lists} which the other ranks map: pmb_exchange_ipc_handle + pmb_exchange_connect_ipc between processes (the 64
 * handle bytes travel by whatever the host has: MPI, a pipe, torch.distributed), pmb_exchange_connect_local inside one
 * process (peer access is enabled between the devices).  Per frame, after pmb_frame / pmb_cull_views:
 *   pmb_exchange_launch   enqueues, without blocking: pack (scene stream) and, on the exchange's own stream, the
 *                         publication of counts + renderable_extents into every peer's table.  The rest of the frame's
 *                         exchange — the cross-GPU barrier (stream waits on device flags), the push of this rank's list
 *                         segments into the receivers' global lists (16-byte NVLink stores) and, on a receiver, the wait
 *                         for every rank's push — is enqueued by the NEXT launch behind that frame's publication, or by
 *                         finish: a stream that waits holds back what is enqueued behind it, so the waits are placed
 *                         where the peers have had a whole frame to arrive.
 *   pmb_exchange_finish   completes and waits for the last launch and reports: counts[rank][view], per-view offset / length of the
 *                         global lists, renderable_extents over all shards, and (receivers) the device pointer of the
 *                         global lists: view v = lists_device[view_offset[v] .. + view_total[v]), ascending global
 *                         entity indices = the list a single GPU would have built.  overflow != 0: a row or the global
 *                         lists were too small, nothing was pushed — create a larger exchange.
 * Modes: PMB_XCHG_ROOT lists gathered on rank `root` (default), PMB_XCHG_ALL replicated on every rank,
 * PMB_XCHG_COUNTS counts + extents only (every GPU draws its own shard).  row_capacity = u32 entries of one rank's row
 * (num_views + the sum of its list lengths), list_capacity = entries of the global lists; both must be the same on every
 * rank.  All ranks must call launch the same number of times. */
typedef struct pmb_exchange pmb_exchange;
typedef enum pmb_exchange_mode { PMB_XCHG_ROOT = 0, PMB_XCHG_ALL = 1, PMB_XCHG_COUNTS = 2 } pmb_exchange_mode;
#define PMB_IPC_HANDLE_BYTES 64u
typedef struct pmb_exchange_result {
    uint32_t        num_views, world, overflow, epoch;
    uint32_t        view_offset[PMB_MAX_VIEWS];
    uint32_t        view_total[PMB_MAX_VIEWS];
    uint32_t        counts[PMB_MAX_PEERS][PMB_MAX_VIEWS];
    float           renderable_extents[6];
    float           device_ms;    /* publication .. last push seen, on the exchange's stream (includes waiting for peers) */
    const uint32_t* lists_device; /* NULL on a rank that does not receive lists */
} pmb_exchange_result;
int  pmb_exchange_create(pmb_exchange** out, pmb_scene* s, uint32_t rank, uint32_t world, uint32_t num_views, uint32_t base_index,
                         uint64_t row_capacity, uint64_t list_capacity, int mode, uint32_t root);
void pmb_exchange_destroy(pmb_exchange* x);
int  pmb_exchange_ipc_handle(pmb_exchange* x, void* handle_out /* PMB_IPC_HANDLE_BYTES */);
int  pmb_exchange_connect_ipc(pmb_exchange* x, const void* handles /* world x PMB_IPC_HANDLE_BYTES, rank order */);
int  pmb_exchange_connect_local(pmb_exchange* const* all, uint32_t world);
int  pmb_exchange_launch(pmb_exchange* x);
int  pmb_exchange_flush(pmb_exchange* x); /* enqueue what is left of the last launch now (no waiting on the host) */
int  pmb_exchange_finish(pmb_exchange* x, pmb_exchange_result* out);
/* the global list of one view on the host (receiving ranks; waits for the last launch): entries copied, or < 0 */
int64_t  pmb_exchange_get_list(pmb_exchange* x, uint32_t view, uint32_t* host_out, uint32_t capacity);
/* test / diagnostic switches: wait with one-warp polling kernels although the driver has stream memory operations;
 * the first `words` u32 of this rank's block {table, flags, result} copied out beside whatever the streams are doing */
void pmb_exchange_use_polling_kernels(int on);
int  pmb_exchange_debug_header(pmb_exchange* x, uint32_t* host_out, uint32_t words);
void*    pmb_exchange_stream(pmb_exchange* x);          /* cudaStream_t the exchange runs on */
uint64_t pmb_exchange_launches(const pmb_exchange* x); /* kernels launched by this exchange */

/* One process, several GPUs — the form a pmtech host would use (it is one process on one user thread,
 * examples/code/example_common.h:107-127): the group owns one pmb_scene per device; pmb_group_scene hands them out
 * for uploads (rank r holds the root-aligned entity range that starts at base_index[r], parents rebased).
 * pmb_group_frame enqueues pmb_frame + the exchange on every GPU without waiting in between; pmb_group_finish waits
 * for all of them and reports rank `rank`'s view of the result. */
typedef struct pmb_group pmb_group;
int        pmb_group_create(pmb_group** out, const int* cuda_devices, uint32_t num_devices, const uint32_t* capacities);
void       pmb_group_destroy(pmb_group* g);
uint32_t   pmb_group_size(const pmb_group* g);
pmb_scene* pmb_group_scene(pmb_group* g, uint32_t rank);
pmb_exchange* pmb_group_exchange(pmb_group* g, uint32_t rank);
int pmb_group_exchange_setup(pmb_group* g, uint32_t num_views, const uint32_t* base_index, uint64_t row_capacity,
                             uint64_t list_capacity, int mode, uint32_t root);
int pmb_group_frame(pmb_group* g, const uint32_t* num_entities, const pmb_frustum* views, uint32_t num_views, int cull_mode);
int pmb_group_finish(pmb_group* g, uint32_t rank, pmb_exchange_result* out);

/* ---- host-buffer convenience: what a shimmed update_scene + N render_scene_view calls do -----
 * Uploads the given host arrays (any may be NULL = keep the mirror), runs pmb_frame, copies
 * the per-view lists back (lists_out[v] must hold num_entities u32, counts_out num_views) and
 * optionally the updated arrays.  Everything is inside the call: H2D, kernels, D2H. */
typedef struct pmb_host_arrays {
    const uint32_t* parents;
    const float*    local_matrices;
    const uint64_t* entities;
    const uint64_t* state_flags;
    const void*     bounding_volumes; /* 52 B records */
    const uint8_t*  has_rb_matrix;
    const float*    transforms; /* 40 B cmp_transform records: the stage-0 input (see PMB_TRANSFORMS) */
    /* outputs, any may be NULL */
    float* world_matrices_out;
    void*  bounding_volumes_out; /* 52 B records: transformed min/max + radius written, min/max kept */
    float* pos_extent_out;
    float* renderable_extents_out; /* 6 floats */
} pmb_host_arrays;
int pmb_frame_host(pmb_scene* s, uint32_t num_entities, const pmb_host_arrays* io,
                   const pmb_frustum* views, uint32_t num_views, int cull_mode,
                   uint32_t** lists_out, uint32_t* counts_out);

/* ---- instrumentation (pen::timer around the loops, ecs_scene.cpp:1401-1402,2026-2028) --------
 * With profiling on, each kernel launch is bracketed by CUDA events on the handle's stream;
 * pmb_kernel_stats returns the accumulated device time and launch count per kernel id.
 * enabled: 0 off, 1 every launch group, 2 only the streaming pass (PMB_K_FRAME): one event pair per frame, cheap enough
 * to stay inside a timed region. */
typedef enum pmb_kernel_id {
    PMB_K_FRAME = 0,     /* fused transform + bounds (+ cull masks) streaming pass */
    PMB_K_RESERVED = 1,  /* unused (round 1's band sweeps are gone; the id keeps the others' values) */
    PMB_K_EXPAND = 2,    /* parent bound expansion (loop C)                       */
    PMB_K_CULL = 3,      /* stand-alone cull mask pass                            */
    PMB_K_COMPACT = 4,   /* count + scan + scatter into draw lists                */
    PMB_K_AUX = 5,       /* frame setup, flag clears, hierarchy index             */
    PMB_K_ANIM = 6,      /* TRS bake, root motion, skin palettes                  */
    PMB_K_DRAW = 7,      /* draw-call packing                                     */
    PMB_K_COUNT = 8
} pmb_kernel_id;
int pmb_set_profiling(pmb_scene* s, int enabled);
int pmb_kernel_stats(pmb_scene* s, int kernel_id, double* total_ms, uint64_t* launches);
int pmb_reset_kernel_stats(pmb_scene* s);
/* kernels launched by this handle since creation (for launch accounting) */
uint64_t pmb_total_launches(const pmb_scene* s);

/* ---- view / frustum producers (host side, per view): core/put/source/camera.cpp ---------------
 * camera_create_perspective :19-49, camera_create_orthographic :51-57, camera_create_cubemap
 * :316-319, camera_set_cubemap_face :321-360, camera_update_look_at :236-262,
 * camera_update_frustum :131-172, camera_update_shadow_frustum :384-416. */
void pmb_camera_create_perspective(pmb_camera* cam, float fov_degrees, float aspect_ratio, float near_plane, float far_plane);
void pmb_camera_create_orthographic(pmb_camera* cam, float left, float right, float bottom, float top, float znear, float zfar);
void pmb_camera_create_cubemap(pmb_camera* cam, float near_plane, float far_plane);
void pmb_camera_set_cubemap_face(pmb_camera* cam, uint32_t face);
void pmb_camera_look_at(pmb_camera* cam, const float pos[3], const float at[3], const float up[3]);
void pmb_camera_update_frustum(pmb_camera* cam);
void pmb_camera_update_shadow_frustum(pmb_camera* cam, const float light_dir[3], const float mn[3], const float mx[3]);
void pmb_camera_spot_shadow(pmb_camera* cam, const float world_matrix[16]); /* ecs_scene.cpp:556-570, host-side */

/* `scene->entities[n] |= e_cmp::transform` for n in [first, first + count), done on the device: the host that
 * rewrote a range of cmp_transforms (an animation tick, `entities.cpp:211-216`; the editor, ecs_editor.cpp:2684)
 * uploads the 40 B records with pmb_upload(PMB_TRANSFORMS, …, first, count) and marks the same range here instead
 * of also re-sending the 8 B entity words.  Same result as uploading entities with the bit set. */
int pmb_mark_transforms_dirty(pmb_scene* s, uint32_t first, uint32_t count);

/* ---- shadow views fitted on the device -------------------------------------------
 * For each directional light: shadow_camera_from_entity (ecs_scene.cpp:538-554) — scene
 * renderable_extents, clamped by the optional shadow_extent_constraints box {min, max} when it is not
 * degenerate, grown by 0.1, light_dir = normalize(-direction) — then camera_update_shadow_frustum
 * (camera.cpp:384-416) and the filter + cull of render_scene_view for that frustum
 * (ecs_scene.cpp:957-960).  The extents never leave the device, which removes the device->host
 * sync the host-side pmb_camera_update_shadow_frustum needs between the update and the shadow
 * culls; the fitted cameras (view, proj, frustum — what the renderer puts in its shadow cbuffer)
 * can be read back afterwards, off the critical path.  `light_directions` = cmp_light::direction
 * per light, f32[num_lights][3]; at most PMB_MAX_VIEWS lights per call.  Lists go to views
 * [first_view, first_view + num_lights): to append behind the lists of a pmb_frame, reserve the total
 * with pmb_reserve_views before that frame (growing the list storage does not keep its contents). */
int pmb_reserve_views(pmb_scene* s, uint32_t num_views);
int pmb_cull_shadow_views(pmb_scene* s, uint32_t num_entities, const float* light_directions, uint32_t num_lights,
                          const float* extent_constraints /* f32[6] or NULL */, int cull_mode, uint32_t first_view);
/* Spot lights (the other branch of shadow_camera_from_entity, ecs_scene.cpp:556-570): perspective(100, 1, 0.1, 500) seen
 * from the light's world matrix — rows = normalised columns 2, 0, 1, translated by -translation — which the update
 * has just left on the device.  `light_entities` = the lights' entity indices. */
int pmb_cull_spot_shadow_views(pmb_scene* s, uint32_t num_entities, const uint32_t* light_entities, uint32_t num_lights,
                               int cull_mode, uint32_t first_view);
/* the cameras fitted by the last of the two calls above */
int pmb_download_shadow_cameras(pmb_scene* s, pmb_camera* out, uint32_t count);
const pmb_camera* pmb_device_shadow_cameras(pmb_scene* s);

/* ---- .pms scene files: core/put/source/ecs/ecs_scene.cpp ------------------------
 * scene_header :2031-2044 (64 x s32), save_scene :2162-2346, load_scene :2350-2535.  File =
 * header, u32 component sizes, extension records, lookup strings (u32 length + bytes + hash_id),
 * cameras (56 B each), then every component array back to back (`size * num_nodes` bytes each, in
 * the order of ecs_scene.h:471-513), then the per-node name / geometry / material specialisations
 * (renderer resources: not read here).  A component is taken only when the stored element size is
 * the one this library mirrors — the reference's own rule (`cmp.size == component_sizes[i]`,
 * :2506) — otherwise its offset is 0 and it is skipped. */
typedef enum pmb_pms_component {
    PMB_PMS_ENTITIES = 0,         /* component  0, u64  */
    PMB_PMS_STATE_FLAGS = 1,      /* component  1, u64  */
    PMB_PMS_PARENTS = 2,          /* component  8, u32  */
    PMB_PMS_TRANSFORMS = 3,       /* component  9, 40 B */
    PMB_PMS_LOCAL_MATRICES = 4,   /* component 10, 64 B */
    PMB_PMS_WORLD_MATRICES = 5,   /* component 11, 64 B */
    PMB_PMS_BOUNDING_VOLUMES = 6, /* component 14, 52 B */
    PMB_PMS_CBUFFER = 7,          /* component 22, u32  */
    PMB_PMS_DRAW_CALLS = 8,       /* component 23, 160 B */
    PMB_PMS_POS_EXTENT = 9,       /* component 38, 32 B (version 10 files) */
    PMB_PMS_COMPONENT_COUNT = 10
} pmb_pms_component;

typedef struct pmb_pms_info {
    int32_t  version;
    uint32_t num_nodes;
    int32_t  num_components;
    int32_t  num_base_components; /* = num_components for version < 9 (:2372-2373) */
    int32_t  num_lookup_strings;
    int32_t  num_extensions;
    uint32_t view_flags;
    int32_t  selected_index;
    uint32_t num_cameras;
    uint32_t reserved;
    uint64_t sizes_offset;   /* u32[num_components]                          */
    uint64_t strings_offset; /* first lookup string                          */
    uint64_t cameras_offset; /* first camera record (after the u32 count)    */
    uint64_t data_offset;    /* first component array                        */
    uint64_t names_offset;   /* 3 hash_ids per node follow the last array    */
    uint64_t component_offset[PMB_PMS_COMPONENT_COUNT]; /* byte offset of element 0, 0 = absent / other size */
} pmb_pms_info;

/* Walks the file image; touches no device.  PMB_ERR_INVALID for a truncated or malformed image. */
int pmb_pms_parse(const void* data, uint64_t len, pmb_pms_info* out);
/* load_scene for this path: uploads every path component found in the image to entities
 * [zero_offset, zero_offset + num_nodes) — zero_offset = scene->num_entities for a merge, with
 * parents[n] += zero_offset (:2530-2531), 0 for a plain load.  The handle must already have the
 * capacity (resize_scene_buffers is the caller's, :2399-2400).  Derived outputs are recomputed by
 * the next update, as in the reference. */
int pmb_load_pms(pmb_scene* s, const void* data, uint64_t len, uint32_t zero_offset, pmb_pms_info* info_out);

/* ---- errors / version ----------------------------------------------------------------------- */
const char* pmb_last_error(void);
const char* pmb_version(void);
int         pmb_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* PMTECH_B200_H */
