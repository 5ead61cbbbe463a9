// pm_kernels.h — internal interface between the C ABI (pm_capi.cu) and the kernels
// (pm_kernels.cu).  Not installed; the public boundary is include/pmtech_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pmtech_b200.h"

namespace pm {

// core/put/source/ecs/ecs_scene.h:144-171 (e_cmp) and :69-83 (e_state): the bits the path reads
constexpr uint64_t CMP_ALLOCATED = 1ull << 0;
constexpr uint64_t CMP_GEOMETRY = 1ull << 1;
constexpr uint64_t CMP_PHYSICS = 1ull << 2;
constexpr uint64_t CMP_MATERIAL = 1ull << 4;
constexpr uint64_t CMP_SKINNED = 1ull << 6;
constexpr uint64_t CMP_BONE = 1ull << 7;
constexpr uint64_t CMP_ANIM_CONTROLLER = 1ull << 9;
constexpr uint64_t CMP_TRANSFORM = 1ull << 12;
constexpr uint64_t CMP_SUB_INSTANCE = 1ull << 14;
constexpr uint64_t CMP_PRE_SKINNED = 1ull << 16;
constexpr uint64_t STATE_HIDDEN = 1ull << 2;
constexpr uint64_t STATE_SYNC_PHYSICS_TRANSFORM = 1ull << 7;

constexpr uint32_t TILE = 64;           // entities per tile of the streaming pass: one warp, two entities per lane
constexpr uint32_t WALK = 16;           // most ancestors re-multiplied in-thread by the shallow form of loop A; deeper scenes run the wavefront form
constexpr uint32_t SHALLOW_DEPTH = 4;   // hierarchies up to this depth use the shallow form (every child re-multiplies its own short chain)
constexpr uint32_t TAIL_MAX_ENTITIES = 1u << 20; // scenes up to this size run the whole frame (loop C across tiles, flag clears, draw lists) in ONE launch
constexpr uint32_t COMPACT_WORDS = 1024; // bitmap words (32 entities each) per CTA of the compaction passes
constexpr uint32_t PAD = 1024;          // mirrors are allocated to a multiple of this many entities

// Two frustum planes of one view as the cull consumes them (hoisted out of the per-entity loop of
// ecs_cull.cpp:34-44), laid out so that one packed FMUL2 / FFMA2 evaluates both planes.
// The corner choice `pos + extent * sign_flip` (sign_flip = sgn(n) * -1, ecs_cull.cpp:36-38) is
// folded into the constants: per axis the normal component n sits in `a` (and b = 0) when the
// plane takes the low corner pos - extent (n >= 0), in `b` (and a = 0) when it takes the high
// corner pos + extent (n < 0).  fma(hi, b, lo * a) is then the reference's rounded product
// corner * n bit for bit (the other term is an exact zero), with no select instruction.
// neg_pd = -plane_distance (aabb threshold), pd = plane_distance (sphere).  64 B = four LDCU.128.
struct __align__(16) PlanePair
{
    float ax[2], bx[2], ay[2], by[2], az[2], bz[2];
    float neg_pd[2], pd[2];
};

struct ViewSet
{
    PlanePair pairs[PMB_MAX_VIEWS * 3]; // view v: planes (0,1) (2,3) (4,5)
    float     one[2];                   // (1.0f, 1.0f): see pm_math.cuh add2()
    uint32_t  pad[2];
};

struct TilePlan;

// device mirrors, as the kernels see them
struct SceneArrays
{
    uint32_t        n;          // scene->num_entities
    const uint32_t* parents;    // u32[N]
    const float4*   local;      // 4 x float4 per entity
    float4*         world;      // 4 x float4 per entity
    uint64_t*       entities;   // u64[N]
    uint64_t*       state_flags;
    const uint8_t*  has_rb;     // may be null (= all 0)
    const float*    bv_in;      // f32[N][6]  min, max
    float*          bv_out;     // f32[N][7]  tmin, tmax, radius
    float4*         pos_extent; // 2 x float4 per entity
    uint32_t*       bitmap;     // per-view visibility bitmaps of the current batch: u32[PMB_MAX_VIEWS][bitmap_stride], bit e%32 of word e/32
    uint32_t        bitmap_stride; // words per view
    float*          extents;    // 6 floats: renderable_extents min, max
    float*          extents_next; // the other slot of the double-buffered extents: the last CTA to leave re-arms it for the next frame
    uint32_t*       dirty;      // set when any entity carried e_cmp::transform / sync_physics_transform (double-buffered like extents)
    uint32_t*       dirty_next;
    uint32_t*       ctrl;       // FrameCtrl words (see CTRL_*): tile counters, grid barriers, exit counter; self-resetting
    // wavefront form of loop A (deep hierarchies)
    const TilePlan* plans;        // [tiles]: the static schedule of each tile (built when the hierarchy changes)
    uint64_t*       xbuf;       // u64[N][16]: published world matrices as 8-byte packets {binary32, frame number}
    uint32_t        frame_no;   // this launch's packet tag (never 0)
    uint64_t*       trace;      // diagnostic (PMB_OPT_TRACE), else null: u64[tiles][4] = loop A start ns, end ns, SM | batches << 32, idle polls | CTA << 32
    // one-launch tail (TAIL_MAX_ENTITIES and below)
    const uint32_t* cross;      // children whose parent chain (shallow) / direct parent (wavefront) leaves their tile
    uint32_t        num_cross;
    const uint32_t* anc_off;    // wavefront form: CSR of every boundary child's ancestors, nearest first (anc_off[num_cross + 1];
    const uint32_t* anc;        //   built with the hierarchy index): loop C across tiles without walking parent links
    uint32_t*       tile_counts;  // u32[tiles][PMB_MAX_VIEWS]: visible entities per tile and view
    uint32_t*       group_totals; // u32[tiles / 32][PMB_MAX_VIEWS]: the same summed over groups of 32 tiles (zero between launches)
    uint32_t*       chunk_totals; // u32[tiles / 1024][PMB_MAX_VIEWS]: ... and over chunks of 1024 tiles
    uint32_t*       group_totals_next; // the other set of totals: zeroed behind the barrier for the next launch
    uint32_t*       chunk_totals_next;
    uint32_t*       lists;      // u32[views][list_stride]
    size_t          list_stride;
    uint32_t*       counts;     // u32[views]
};

// Wavefront form of loop A: the schedule of one tile, a function of `parents` alone, built by launch_plan when the
// hierarchy changes and read with the tile (one more bulk copy).  The tile's non-root entities are cut into ROUNDS of
// at most 8 entities whose parents are final before the round starts (roots, entities of earlier rounds, or — last —
// published entities of earlier tiles); a round is (entity, matrix row) work for the 32 lanes.  Loop C inside the tile
// runs the same rounds backwards, each entity folding its box into its direct parent.
struct __align__(16) TilePlan
{
    uint8_t  order[TILE];      // schedule: bits 0-5 tile-relative entity, bit 7 = it has a child in a later tile (publishes)
    uint8_t  rstart[TILE + 2]; // rstart[r] = first position of round r in order[]; rstart[nrounds] = scheduled entities
    uint8_t  nrounds;          // <= TILE
    uint8_t  nlocal;           // rounds that need nothing from earlier tiles (they come first)
    uint8_t  pad[4];
    uint64_t publish;          // bit i = entity i of the tile has a child in a later tile (roots publish outside the rounds)
};
static_assert(sizeof(TilePlan) == 144, "plan layout (bulk copies move multiples of 16 bytes)");

// FrameCtrl: u32 words of the control block (zeroed at creation; the last CTA of every launch zeroes them again)
constexpr uint32_t CTRL_EXIT = 0;       // CTAs that have left
constexpr uint32_t CTRL_BAR0 = 1;       // grid barrier after the tiles (one-launch tail)
constexpr uint32_t CTRL_TILE = 32;      // tile counter c at CTRL_TILE + 32 c (128 B apart)
constexpr uint32_t CTRL_WORDS = 32 + 16 * 32;

enum FrameMode
{
    MODE_UPDATE = 0, // stages 1-2 only
    MODE_AABB = 1,
    MODE_SPHERE = 2,
    MODE_FILTER = 3
};

// host-side hoisting of the per-plane constants (float, no contraction)
void make_view_set(ViewSet* vs, const pmb_frustum* views, uint32_t num_views);

// ---- launches (all asynchronous on `st`) -----------------------------------------------------
// per-device launch geometry of the streaming pass (occupancy of each instantiation), filled once per scene handle
struct FrameGeometry
{
    int sms = 0;
    int ctas_per_sm[4][2][2] = {}; // [mode][deep][tail]
};
cudaError_t frame_geometry(FrameGeometry* g, int device);
void launch_extents_init(float* extents2, cudaStream_t st); // both slots of the double-buffered extents -> +-FLT_MAX
// The streaming pass.  deep: wavefront form of loop A (needs A.plans / A.xbuf).  tail: loop C across tiles, the flag
// clears and the draw-list compaction run inside the same launch behind grid barriers (needs A.cross, A.tile_counts,
// A.lists, A.counts; the grid is co-resident by construction).  Loop C for (child, ancestor) pairs inside a tile is
// always folded into the pass; without `tail` the pairs that cross a tile boundary are left to launch_expand.
void launch_frame(const SceneArrays& A, const ViewSet& vs, uint32_t num_views, int mode, bool deep, bool tail,
                  const FrameGeometry& g, cudaStream_t st);
void launch_expand(const SceneArrays& A, const uint32_t* children, uint32_t num_children, cudaStream_t st);
// the same for the wavefront form, from the ancestor lists (A.cross, A.anc_off, A.anc)
void launch_expand_lists(const SceneArrays& A, cudaStream_t st);
// ancestor lists of the boundary children: counts per child (then an exclusive scan into offsets), then the fill
void launch_ancestor_counts(const uint32_t* parents, const uint32_t* cross, uint32_t num_cross, uint32_t* counts, cudaStream_t st);
cudaError_t scan_ancestor_offsets(uint32_t* counts_then_offsets, uint32_t num_cross, void** temp, size_t* temp_bytes, cudaStream_t st);
void launch_ancestor_fill(const uint32_t* parents, const uint32_t* cross, uint32_t num_cross, const uint32_t* offsets, uint32_t* anc,
                          cudaStream_t st);
void launch_clear_flags(const SceneArrays& A, cudaStream_t st);
void launch_cull(const SceneArrays& A, const ViewSet& vs, uint32_t num_views, int mode, cudaStream_t st);
// shadow views fitted on the device: extents f32[6], directions f32[num_lights][3], constraints f32[6] or null
void launch_shadow_fit(const float* extents, const float* directions, const float* constraints, uint32_t num_lights,
                       pmb_camera* cams, ViewSet* vs, cudaStream_t st);
void launch_spot_fit(const float4* world, const uint32_t* light_entities, uint32_t num_lights, uint32_t n, const float proj[16],
                     pmb_camera* cams, ViewSet* vs, cudaStream_t st);
void launch_cull_device_views(const SceneArrays& A, const ViewSet* vs, uint32_t num_views, int mode, cudaStream_t st);

// compaction of the per-view visibility bitmaps into one ascending u32 list per view
//   chunk_counts: u32[num_views][num_chunks] scratch; lists: u32[num_views][list_stride];
//   counts: u32 totals, written at view_base + v
void launch_compact(const uint32_t* bitmap, uint32_t bitmap_stride, uint32_t n, uint32_t num_views, uint32_t view_base,
                    uint32_t* chunk_counts, uint32_t* lists, size_t list_stride, uint32_t* counts, cudaStream_t st);
uint32_t compact_num_chunks(uint32_t n);
uint32_t compact_num_launches(uint32_t n); // kernels launch_compact issues for n entities

// hierarchy index (rebuilt when parents change): depth per entity (low 16 bits; high 16: links into earlier tiles on the way to the root); the list of children that loop C cannot finish
// inside their own tile (direct_only = false: some ancestor lies in an earlier tile — the shallow form walks whole
// chains in shared memory; true: the direct parent does — the wavefront form folds batch by batch into direct parents);
// and, for the wavefront form, which entities have a child in a later tile (SceneArrays::publish)
void launch_depth(const uint32_t* parents, uint32_t n, uint32_t* depth, uint32_t* max_depth, cudaStream_t st);
void launch_merge_parents(const uint32_t* staged, uint32_t* parents, uint32_t count, uint32_t* changed, cudaStream_t st);
void launch_cross(const uint32_t* parents, uint32_t n, bool direct_only, uint32_t* cross, uint32_t* num_cross, cudaStream_t st);
void launch_publish_mask(const uint32_t* parents, uint32_t n, uint32_t padded, uint64_t* publish, cudaStream_t st);
// wavefront form: the per-tile schedules from the parents and the publish masks (one thread per tile)
void launch_plan(const uint32_t* parents, const uint32_t* depth, uint32_t n, const uint64_t* publish, TilePlan* plans, cudaStream_t st);

// cmp_bounding_volume (52 B AoS) <-> device split
void launch_bv_split(const float* aos, uint32_t count, float* bv_in, float* bv_out, cudaStream_t st);
void launch_bv_merge(const float* bv_in, const float* bv_out, uint32_t count, float* aos, cudaStream_t st);
// multi-GPU list exchange: up to PMB_MAX_SEGMENTS runs of u32 copied (value + add) in one launch
struct SegTable
{
    pmb_segment seg[PMB_MAX_SEGMENTS];
};
void launch_copy_segments(const uint32_t* src, uint32_t* dst, const SegTable& table, uint32_t num_segments,
                          uint32_t max_count, cudaStream_t st);

// draw-call packing: 160 B cmp_draw_call records
void launch_draw_calls(const float4* world, const uint64_t* entities, const uint32_t* cbuffer, float4* dc, uint32_t n, float time_ms,
                       bool user_data, cudaStream_t st);

// the exchange as one push over peer memory: base pointers of the same symmetric buffer on every rank
struct PeerPtrs
{
    uint64_t p[PMB_MAX_PEERS];
};
void launch_publish_counts(const uint32_t* row, uint32_t nv, uint32_t rank, uint32_t world, const PeerPtrs& tabs, cudaStream_t st);
void launch_push_row(const uint32_t* row, uint32_t row_stride, uint32_t nv, uint32_t rank, uint32_t world, const uint32_t* tab,
                     const PeerPtrs& outs, uint64_t out_capacity, uint32_t* meta, cudaStream_t st);

// stage 0 (TRS bake), root motion, skin palettes
typedef pmb_skin SkinJob;
void launch_bake(float* transforms, const uint64_t* entities, const uint64_t* state_flags, float4* local, uint32_t n,
                 cudaStream_t st);
void launch_mark_dirty(uint64_t* entities, uint32_t first, uint32_t count, cudaStream_t st);
void launch_root_motion(float* transforms, uint64_t* entities, const uint32_t* targets, const float* rotations,
                        const float* deltas, uint32_t count, uint32_t n, cudaStream_t st);
void launch_skin(const float4* world, const SkinJob* jobs, uint32_t num_jobs, const float4* bind, float4* palettes, uint32_t n,
                 cudaStream_t st);

// fixed-stride rows {counts[nv], lists back to back} for the all-gather, and their device-side unpacking
// extents (optional): 6 floats copied to row[extents_at .. extents_at + 6) (the row buffer must have room for them)
void launch_pack_row(const uint32_t* lists, size_t list_stride, const uint32_t* counts, uint32_t nv, uint32_t add,
                     uint32_t* row, uint32_t row_stride, cudaStream_t st, const float* extents = nullptr, uint32_t extents_at = 0);
void launch_unpack_rows(const uint32_t* rows, uint32_t world, uint32_t row_stride, uint32_t nv, uint32_t* dst,
                        uint64_t dst_capacity, uint32_t* meta, cudaStream_t st);

// what the exchange (pm_exchange.cu) needs to see of a scene handle (pm_capi.cu)
struct ScenePeek
{
    int             device;
    cudaStream_t    stream;
    const uint32_t* counts;      // u32[views] of the last cull
    const uint32_t* lists;       // u32[views][list_stride]
    size_t          list_stride;
    const float*    extents;     // renderable_extents of the last update (6 floats)
    uint32_t        last_views;
};
bool scene_peek(pmb_scene* s, ScenePeek* out);
int  fail_code(int code, const char* fmt, ...);

} // namespace pm
