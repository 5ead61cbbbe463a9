// pm_kernels.cu — sm_100a kernels of the scene-update path.
//
//   frame_kernel   the streaming pass: one warp per CTA, 64-entity tiles (two entities per lane) handed out in
//                  ascending order by global counters.  A tile's 108 B/entity arrive in one shared-memory stage by
//                  TMA bulk copies and the 64 B world matrices / 28 B bound records leave the same way.  The pass is
//                  bound by instruction issue as much as by HBM (ncu: profiles/), so it is written issue-lean:
//                    * loop A as (entity, matrix-row) work items over all 32 lanes; shallow hierarchies re-multiply
//                      each child's short chain, deep ones run a wavefront from a static per-tile schedule (TilePlan:
//                      rounds of 8 entities x 4 rows, phases by tile-crossing depth) with per-entity publication
//                      between tiles (one multiplication per entity, no per-level launches);
//                    * the 8-corner bound transform uses the monotonicity of rounded addition
//                      (max over corners = the sum of the per-axis maxima, bit for bit);
//                    * the plane tests evaluate two planes per FMUL2 / FFMA2 (packed binary32
//                      pairs, each half rounded on its own);
//                    * visibility leaves the tile as two ballot words per view (per-view bitmaps);
//                    * loop C's parent expansion is folded in for parent links inside the tile
//                      (shared-memory min/max atomics).
//                  For scenes up to TAIL_MAX_ENTITIES the rest of the frame — loop C across tiles, the flag clears,
//                  the compaction of the bitmaps into draw lists — runs in the same launch behind grid barriers.
//   expand_kernel  loop C (parent bound expansion) with exact float min/max atomics, over the
//                  children whose chain crosses a tile boundary (large shallow scenes);
//   expand_lists_kernel  the same for deep hierarchies from flat ancestor lists (no parent-link chasing).
//   cull_kernel    stand-alone per-view plane tests over pos_extent (render_scene_view calls).
//   compact_*      per-view bitmaps -> one ascending u32 index list per view (popc + block scan),
//                  i.e. the sb_push order of the reference (large scenes, and the stand-alone cull).
//
// Reference: core/put/source/ecs/ecs_scene.cpp:1405-1556, core/put/source/ecs/ecs_cull.cpp:21-105.
#include "pm_kernels.h"
#include "pm_math.cuh"
#include "pm_camera_math.h"

#include <float.h>

#include <algorithm>

#include <cub/device/device_scan.cuh>


namespace pm {

// ------------------------------------------------------------------------------------------------
// per-plane constants.  Host: compiled without contraction; device (shadow_fit_kernel): -fmad=false.
// One C operation per IEEE operation on both sides.
// ------------------------------------------------------------------------------------------------
static __host__ __device__ float plane_dot3(const float* a, const float* b)
{
    volatile float r = 0.0f; // volatile: keep every intermediate a rounded binary32
    for (int i = 0; i < 3; ++i)
    {
        volatile float t = a[i] * b[i];
        r = r + t;
    }
    return r;
}

static __host__ __device__ void view_set_planes(ViewSet* vs, uint32_t v, const pmb_frustum& f)
{
    for (int p = 0; p < 6; ++p)
    {
        PlanePair&   k = vs->pairs[v * 3 + p / 2];
        const int    h = p & 1;
        const float* n = f.n[p];
        const float* x0 = f.p[p];
        // maths::plane_distance(x0, xN) = dot(xN, x0) * -1 (ecs_cull.cpp:37); the aabb test
        // compares against -pd (ecs_cull.cpp:40)
        volatile float pd = plane_dot3(n, x0) * -1.0f;
        k.pd[h] = pd;
        k.neg_pd[h] = -pd;
        // sign_flip = sgn(n) * -1 (ecs_cull.cpp:36): n < 0 -> pos + extent (b), n > 0 -> pos - extent
        // (a); n == 0 multiplies either corner by zero, so it shares the n > 0 slot.
        float* a[3] = {k.ax, k.ay, k.az};
        float* b[3] = {k.bx, k.by, k.bz};
        for (int i = 0; i < 3; ++i)
        {
            const bool high = n[i] < 0.0f;
            a[i][h] = high ? 0.0f : n[i];
            b[i][h] = high ? n[i] : 0.0f;
        }
    }
}

void make_view_set(ViewSet* vs, const pmb_frustum* views, uint32_t num_views)
{
    memset(vs, 0, sizeof(*vs));
    vs->one[0] = vs->one[1] = 1.0f;
    for (uint32_t v = 0; v < num_views && v < PMB_MAX_VIEWS; ++v)
        view_set_planes(vs, v, views[v]);
}

// Directional-light shadow views fitted where the scene extents already are:
// shadow_camera_from_entity (ecs_scene.cpp:538-554) -> camera_update_shadow_frustum (camera.cpp:384-416)
// -> camera_update_frustum (:131-172), then the per-plane constants of the cull.  One thread per light; the
// arithmetic is the host producers' own source text (pm_camera_math.h), so the frustum carries the bits
// pmb_camera_update_shadow_frustum would give for the same extents — without the device->host->device trip
// of renderable_extents in the middle of the frame.
__global__ void shadow_fit_kernel(const float* __restrict__ extents, const float* __restrict__ directions,
                                  const float* __restrict__ constraints, uint32_t num_lights, pmb_camera* __restrict__ cams,
                                  ViewSet* __restrict__ vs)
{
    const uint32_t l = threadIdx.x;
    if (l == 0)
    {
        vs->one[0] = vs->one[1] = 1.0f;
        vs->pad[0] = vs->pad[1] = 0u;
    }
    if (l >= num_lights)
        return;
    float emin[3] = {extents[0], extents[1], extents[2]}, emax[3] = {extents[3], extents[4], extents[5]};
    if (constraints) // clamp to the shadow map's max extents when the constraint box is not degenerate (:546-550)
    {
        const float d[3] = {constraints[0] - constraints[3], constraints[1] - constraints[4], constraints[2] - constraints[5]};
        if (plane_dot3(d, d) != 0.0f)
            for (int k = 0; k < 3; ++k)
            {
                emin[k] = constraints[k] > emin[k] ? constraints[k] : emin[k];             // max_union
                emax[k] = constraints[3 + k] < emax[k] ? constraints[3 + k] : emax[k];     // min_union
            }
    }
    const pmcam::v3 dir = pmcam::normalize(-pmcam::load3(directions + l * 3)); // normalize(-lights[n].direction) (:552)
    const float     ld[3] = {dir.x, dir.y, dir.z};
    const float     mn[3] = {emin[0] - 0.1f, emin[1] - 0.1f, emin[2] - 0.1f}; // emin - vec3f(0.1f) (:553)
    const float     mx[3] = {emax[0] + 0.1f, emax[1] + 0.1f, emax[2] + 0.1f};
    pmb_camera      cam;
    memset(&cam, 0, sizeof(cam));
    pmcam::update_shadow_frustum(&cam, ld, mn, mx);
    cams[l] = cam;
    view_set_planes(vs, l, cam.camera_frustum);
}

// Spot-light shadow views (ecs_scene.cpp:556-570): the camera is the light's world matrix, which the update
// has just left on the device.  One thread per light.
__global__ void spot_fit_kernel(const float4* __restrict__ world, const uint32_t* __restrict__ light_entities, uint32_t num_lights,
                                uint32_t n, const pmcam::m4 proj, pmb_camera* __restrict__ cams, ViewSet* __restrict__ vs)
{
    const uint32_t l = threadIdx.x;
    if (l == 0)
    {
        vs->one[0] = vs->one[1] = 1.0f;
        vs->pad[0] = vs->pad[1] = 0u;
    }
    if (l >= num_lights)
        return;
    const uint32_t e = light_entities[l] < n ? light_entities[l] : 0u;
    float          w[16];
    for (int r = 0; r < 4; ++r)
    {
        const float4 row = world[(size_t)e * 4 + r];
        w[r * 4 + 0] = row.x, w[r * 4 + 1] = row.y, w[r * 4 + 2] = row.z, w[r * 4 + 3] = row.w;
    }
    pmb_camera cam;
    memset(&cam, 0, sizeof(cam));
    pmcam::spot_shadow(&cam, proj, w);
    cams[l] = cam;
    view_set_planes(vs, l, cam.camera_frustum);
}

// ------------------------------------------------------------------------------------------------
// per-entity work shared by the streaming and band kernels
// ------------------------------------------------------------------------------------------------
struct Bounds
{
    float tmin[3], tmax[3];
    float radius;
    float pos[3], ext[3];
};

__device__ __forceinline__ float bitsel(uint32_t m, float a_if_set, float b_if_clear)
{
    return __uint_as_float((__float_as_uint(a_if_set) & m) | (__float_as_uint(b_if_clear) & ~m));
}

// loop B of update_scene (ecs_scene.cpp:1486-1526) for one entity.
//
// The reference transforms the 8 corners min + size * {0,1}^3 one by one and folds them with
// min_union / max_union.  Every corner coordinate is
//     p = ((px + py) + pz) + m.w,   px = 0 + x*m.x,  py = y*m.y,  pz = z*m.z   (transform_vector)
// with x, y, z each one of two values.  Rounded addition is monotone in each argument, so the
// largest (smallest) of the 8 rounded results is the one built from the larger (smaller) of the
// two px, of the two py and of the two pz: 3 products, 6 min/max and 3 sums per row give the same
// bits as 8 transforms and 16 min/max.  (px is never -0 because of the leading 0 +, so no corner
// value is -0 and equal values have equal bits.  Holds for finite inputs; with NaN or infinities
// in a matrix or a box the reference's own result depends on the visiting order.)
__device__ __forceinline__ void bounds_from_world(const float4 w0, const float4 w1, const float4 w2, const float bmin[3],
                                                  const float bmax[3], f32x2 one, Bounds& o)
{
    // :1488-1489  min, size; corner coordinate = min + size * {0, 1}
    const f32x2 sel01 = pack2(0.0f, 1.0f);
    f32x2       c[3]; // (lo, hi) per axis
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
        float size = sub(bmax[i], bmin[i]);
        c[i] = add2(pack2(bmin[i], bmin[i]), mul2(pack2(size, size), sel01), one);
    }
    const float4 rows[3] = {w0, w1, w2};
    const f32x2  zero = pack2(0.0f, 0.0f);
#pragma unroll
    for (int r = 0; r < 3; ++r)
    {
        const float4 m = rows[r];
        f32x2        px = add2(zero, mul2(c[0], pack2(m.x, m.x)), one); // 0 + x*m0
        f32x2        py = mul2(c[1], pack2(m.y, m.y));
        f32x2        pz = mul2(c[2], pack2(m.z, m.z));
        float        a0, a1, b0, b1, c0, c1;
        unpack2(px, a0, a1);
        unpack2(py, b0, b1);
        unpack2(pz, c0, c1);
        f32x2 sx = pack2(fminf(a0, a1), fmaxf(a0, a1));
        f32x2 sy = pack2(fminf(b0, b1), fmaxf(b0, b1));
        f32x2 sz = pack2(fminf(c0, c1), fmaxf(c0, c1));
        f32x2 t = add2(add2(add2(sx, sy, one), sz, one), pack2(m.w, m.w), one); // (tmin, tmax) of this row
        unpack2(t, o.tmin[r], o.tmax[r]);
    }
    // :1512 mag(tmax - tmin) * 0.5
    float d[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
        d[i] = sub(o.tmax[i], o.tmin[i]);
    o.radius = mul(__fsqrt_rn(dot3(d[0], d[0], d[1], d[1], d[2], d[2])), 0.5f);
    // :1515-1518
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
        o.pos[i] = add(o.tmin[i], mul(d[i], 0.5f));
        o.ext[i] = sub(o.tmax[i], o.pos[i]);
    }
}

// filter_entities_scalar (ecs_cull.cpp:84-105)
__device__ __forceinline__ bool filter_pass(uint64_t ef, uint64_t sf)
{
    const uint64_t accept = CMP_GEOMETRY | CMP_MATERIAL;
    return ((ef & accept) == accept) && !(sf & STATE_HIDDEN) && !(ef & CMP_SUB_INSTANCE);
}

// what the plane tests of one entity need, prepared once per entity: both box corners (aabb:
// pos - extent, pos + extent, i.e. pos + extent * -1 / * +1 of ecs_cull.cpp:38; sphere: pos twice),
// each duplicated into both halves of a packed pair, and the threshold of the sphere test
struct CullIn
{
    f32x2 lo[3], hi[3];
    float radius;
};

template <int MODE>
__device__ __forceinline__ void cull_prepare(const float pos[3], const float ext[3], float radius, CullIn& ci)
{
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
        const float l = MODE == MODE_AABB ? sub(pos[i], ext[i]) : pos[i];
        const float h = MODE == MODE_AABB ? add(pos[i], ext[i]) : pos[i];
        ci.lo[i] = pack2(l, l);
        ci.hi[i] = pack2(h, h);
    }
    ci.radius = radius;
}

// frustum_cull_aabb_scalar / frustum_cull_sphere_scalar (ecs_cull.cpp:21-82) against view v of the
// batch for NE entities that share the view's constants: all six planes, two per packed operation.
// in[i] = entity i is NOT outside.  v may be a run-time (warp-uniform) index.
template <int MODE, int NE>
__device__ __forceinline__ void cull_view(const ViewSet& VS, uint32_t v, const CullIn (&ci)[NE], bool (&in)[NE])
{
#pragma unroll
    for (int i = 0; i < NE; ++i)
        in[i] = true;
    if (MODE == MODE_FILTER)
        return;
    const f32x2 one = pack2(VS.one[0], VS.one[1]);
#pragma unroll
    for (int j = 0; j < 3; ++j)
    {
        const PlanePair& k = VS.pairs[v * 3 + j];
        const float4*    kq = reinterpret_cast<const float4*>(&k); // four 128-bit constant loads per plane pair
        const float4     q0 = kq[0], q1 = kq[1], q2 = kq[2];
        const f32x2      ax = pack2(q0.x, q0.y), bx = pack2(q0.z, q0.w);
        const f32x2      ay = pack2(q1.x, q1.y), by = pack2(q1.z, q1.w);
        const f32x2      az = pack2(q2.x, q2.y), bz = pack2(q2.z, q2.w);
#pragma unroll
        for (int i = 0; i < NE; ++i)
        {
            // corner.x * n.x etc. (:38): the corner select is in the constants (PlanePair)
            f32x2 tx = fma2(ci[i].hi[0], bx, mul2(ci[i].lo[0], ax));
            f32x2 ty = fma2(ci[i].hi[1], by, mul2(ci[i].lo[1], ay));
            f32x2 tz = fma2(ci[i].hi[2], bz, mul2(ci[i].lo[2], az));
            f32x2 d2 = add2(add2(tx, ty, one), tz, one);
            float da, db;
            if (MODE == MODE_AABB)
            {
                unpack2(d2, da, db);
                in[i] &= !(da > k.neg_pd[0]); // :40
                in[i] &= !(db > k.neg_pd[1]);
            }
            else
            {
                f32x2 d = add2(d2, pack2(k.pd[0], k.pd[1]), one); // point_plane_distance, :68
                unpack2(d, da, db);
                in[i] &= !(da > ci[i].radius); // :70
                in[i] &= !(db > ci[i].radius);
            }
        }
    }
}

// loop A flag logic (ecs_scene.cpp:1408-1446) on a copy of the flags: returns true when the
// entity's world matrix is left untouched this frame (physics entity without an rb matrix)
__device__ __forceinline__ bool world_is_skipped(uint64_t ef, uint64_t sf, const uint8_t* has_rb, uint32_t idx)
{
    if (!(ef & CMP_PHYSICS))
        return false;
    bool transform = (ef & CMP_TRANSFORM) && !(sf & STATE_SYNC_PHYSICS_TRANSFORM);
    if (transform)
        return false;
    return !(has_rb && has_rb[idx]);
}

__device__ __forceinline__ mat4r load_mat_global(const float4* base, uint32_t idx)
{
    mat4r m;
    const float4* p = base + (size_t)idx * 4;
    m.r[0] = p[0];
    m.r[1] = p[1];
    m.r[2] = p[2];
    m.r[3] = p[3];
    return m;
}

constexpr int MAT_STRIDE = 4; // float4 per staged matrix: the tile is a verbatim image of the global array (bulk copies);
                              // the 4-way bank conflict of the per-thread row reads costs LSU cycles, not issue slots

__device__ __forceinline__ mat4r load_mat_shared(const float4* s_mat, uint32_t rel)
{
    mat4r m;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        m.r[k] = s_mat[rel * MAT_STRIDE + k];
    return m;
}

// ------------------------------------------------------------------------------------------------
// shared by frame / band / cull kernels: plane tests of every view of the batch for one warp.
// Must be called by all 32 lanes (it ballots).  Lane v ends up holding the visibility word of
// view v for this warp's 32 entities; `pass` = the entity is in filter_entities_scalar's list.
// ------------------------------------------------------------------------------------------------
template <int MODE>
__device__ __forceinline__ uint32_t cull_warp(const ViewSet& VS, uint32_t nv, bool pass, const CullIn& ci, uint32_t lane)
{
    uint32_t     mine = 0;
    const CullIn c1[1] = {ci};
#pragma unroll 1
    for (uint32_t v = 0; v < nv; ++v) // uniform; rolled: the per-view code stays in the instruction cache
    {
        // straight-line on purpose: the warp runs the plane tests as soon as one lane passes the
        // filter, so a branch around them only adds instructions
        bool in[1];
        cull_view<MODE, 1>(VS, v, c1, in);
        uint32_t b = __ballot_sync(0xFFFFFFFFu, in[0] & pass);
        mine = (lane == v) ? b : mine;
    }
    return mine;
}

// Row r of the world matrix of a non-root entity, shallow form: walk from `parent` towards the root, then
// multiply back down so that the product order is ((top * l1) * l2) ... * local (ecs_scene.cpp:1470 applied in
// ascending entity order).  Row r of a product chain only needs row r of its left-most factor, so the
// four rows of one matrix are four independent work items (four lanes), each doing a quarter of the
// arithmetic with the same rounding sequence as the full product.  `top` is the root's local matrix,
// or the stored world matrix of an ancestor that is not rewritten this frame.  Ancestors inside the
// CTA tile come from shared memory.  Only used for hierarchies of depth <= SHALLOW_DEPTH (never more than WALK
// links); deeper scenes run the wavefront form below, which multiplies every entity exactly once.
__device__ __forceinline__ float4 load_row(const float4* mats, size_t idx, uint32_t r) { return mats[idx * 4 + r]; }

__device__ __forceinline__ float4 chain_row(const SceneArrays& A, uint32_t base, uint32_t parent, uint32_t self_rel, uint32_t r,
                                            const float4* s_mat, const uint32_t* s_par, const uint64_t* s_ent,
                                            const uint64_t* s_sf)
{
    uint32_t chain[WALK];
    int      len = 0;
    uint32_t cur = parent;
    bool     top_world = false;
    for (;;)
    {
        chain[len++] = cur;
        uint32_t rel = cur - base;
        bool     in_tile = rel < TILE;
        uint64_t ce = in_tile ? s_ent[rel] : A.entities[cur];
        if (ce & CMP_PHYSICS)
        {
            uint64_t cs = in_tile ? s_sf[rel] : A.state_flags[cur];
            if (world_is_skipped(ce, cs, A.has_rb, cur))
            {
                top_world = true;
                break;
            }
        }
        uint32_t pp = in_tile ? s_par[rel] : A.parents[cur];
        if (pp >= cur) // root (or an upward link, which a valid scene does not have)
            break;
        if (len == (int)WALK) // cannot happen: the host picks the wavefront form for anything deeper than SHALLOW_DEPTH
            break;
        cur = pp;
    }
    uint32_t top = chain[len - 1];
    float4   acc;
    if (top_world)
        acc = load_row(A.world, top, r);
    else if (top - base < TILE)
        acc = s_mat[(top - base) * MAT_STRIDE + r];
    else
        acc = load_row(A.local, top, r);
    for (int i = len - 2; i >= 0; --i)
    {
        uint32_t c = chain[i];
        mat4r    l = (c - base < TILE) ? load_mat_shared(s_mat, c - base) : load_mat_global(A.local, c);
        acc = mul_row(acc, l);
    }
    return mul_row(acc, load_mat_shared(s_mat, self_rel)); // :1470
}

// ------------------------------------------------------------------------------------------------
// frame_kernel — the streaming pass.  One warp per CTA, persistent (grid = as many CTAs as fit per SM), walking
// 64-entity tiles.  Lane 0 asks for the tile's 108 B/entity with TMA bulk copies (cp.async.bulk + mbarrier) into
// ONE in-place shared-memory stage — with 24+ resident warps per SM the other warps hide a warp's load, and a
// second stage would halve them — and the finished world-matrix and bound tiles leave the same way
// (cp.async.bulk shared -> global), so moving the 232 B/entity costs almost no issue slots: the pass is bound
// by instruction issue as much as by HBM.  Every lane owns TWO entities (tile-relative lane and 32 + lane): the
// per-view constants, the loop and ballot overheads and the per-warp child / parent-expansion passes are
// paid once per 64 entities, and the two independent dependency chains hide each other's latency.  A single
// warp needs no CTA barrier at all.  Tiles are handed out in ascending order by global counters.
//
// Loop A (world = parent * local) comes in two forms, picked by the host per topology:
//   shallow (DEEP = false)  every non-root entity re-multiplies its own chain of at most SHALLOW_DEPTH locals
//                           (chain_row): no dependency between tiles at all;
//   wavefront (DEEP = true) every entity is multiplied exactly once, as soon as its parent is final: the warp keeps
//                           a ready list (parent inside the tile and done, or parent in an earlier tile and
//                           PUBLISHED) and works it off 8 entities x 4 rows at a time.  An entity with a child in
//                           a later tile publishes its world matrix as sixteen 8-byte packets {float, frame
//                           number}: an aligned 8-byte access is single-copy atomic, so every packet validates
//                           itself and neither side needs a fence (the LL idea of NCCL's low-latency protocol).
//                           Tiles are taken in ascending order and a tile only ever waits for lower tiles, which
//                           are finished or running, so the waits cannot deadlock; what remains is each entity's
//                           own chain of tile crossings (a handful), not the depth of the tree times the tiles.
// TAIL = true (scenes up to TAIL_MAX_ENTITIES): what otherwise follows as separate launches — loop C across
// tiles, the flag clears of loop A and the compaction of the visibility bitmaps into draw lists — runs inside
// the same launch behind ONE grid barrier (the grid is co-resident: cooperative launch), so a 64 Ki-entity demo
// scene is one launch instead of seven dependent ones.  The draw-list offsets need no scan pass: every tile adds
// its visible counts to the totals of its group of 32 tiles and of its chunk of 1024 while the tiles run, and
// behind the barrier a tile's offset is three masked 32-wide sums (tiles below it in its group, groups below it
// in its chunk, chunks below it).
// The kernel arms its own control block for the next launch (tile counters, barriers, the other slot of the
// double-buffered extents / dirty marker): there is no set-up launch.
// ------------------------------------------------------------------------------------------------
constexpr int      EPL = TILE / 32; // entities per lane
// measured at 16 Mi entities (frame kernel ms): chunk 16: 0.696, 8: 0.652, 4: 0.634, 2: 0.614 (one counter);
// chunk 1 with one counter 0.93 (the same-address atomic serialises), with 4 counters 0.599, with 8: 0.605
#ifndef PMB_TILE_CHUNK
#define PMB_TILE_CHUNK 1
#endif
#ifndef PMB_TILE_COUNTERS
#define PMB_TILE_COUNTERS 4
#endif
constexpr uint32_t TILE_CHUNK = PMB_TILE_CHUNK;  // consecutive tiles a CTA takes per visit to a tile counter
constexpr uint32_t TILE_COUNTERS = PMB_TILE_COUNTERS; // counters, 128 B apart; counter c hands out the chunks c, c + TILE_COUNTERS, ...
#ifndef PMB_ROLLED_VIEWS
#define PMB_ROLLED_VIEWS 1
#endif
static_assert(EPL == 2, "the streaming pass is written for two entities per lane");
static_assert(TILE_COUNTERS <= 16, "control block layout");

struct __align__(16) TileStage
{
    float4   mat[TILE * MAT_STRIDE]; // local matrices in, world matrices out
    float4   bv[TILE * 7 / 4];       // in: min/max extents (24 B/entity); out: tmin, tmax, radius (28 B/entity)
    uint64_t ent[TILE];
    uint64_t sf[TILE];
    uint32_t par[TILE];
};
constexpr uint32_t STAGE_IN_BYTES = TILE * (64 + 24 + 8 + 8 + 4);
static_assert(sizeof(TileStage) == TILE * (64 + 28 + 8 + 8 + 4), "stage layout");
static_assert(sizeof(TileStage) % 128 == 0, "stage alignment");
// wavefront form: the tile's schedule travels with it
struct __align__(16) TileStageDeep
{
    TileStage tile;
    TilePlan  plan;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// waits for the phase with the given parity; traps instead of hanging if it never completes
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    const uint32_t a = smem_u32(bar);
    for (uint32_t spin = 0;; ++spin)
    {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok)
                     : "r"(a), "r"(parity)
                     : "memory");
        if (ok)
            return;
        if (spin > (1u << 24))
            __trap();
    }
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_done() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); } // writes performed
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ uint64_t global_ns()
{
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ uint32_t ld_acquire(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// wavefront form: a published world-matrix row = four 8-byte packets {binary32, frame number}.  Aligned 8-byte
// accesses are single-copy atomic, so a packet carrying this frame's number carries this frame's value: no fence
// on either side, and packets may land in any order.
__device__ __forceinline__ void publish_row(uint64_t* x, const float4 row, uint32_t frame_no)
{
    const uint64_t f = (uint64_t)frame_no << 32;
    const uint64_t a = f | __float_as_uint(row.x), b = f | __float_as_uint(row.y);
    const uint64_t c = f | __float_as_uint(row.z), d = f | __float_as_uint(row.w);
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(x), "l"(a), "l"(b) : "memory");
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(x + 2), "l"(c), "l"(d) : "memory");
}
__device__ __forceinline__ bool is_published(const uint64_t* x, uint32_t frame_no)
{
    uint64_t v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(x) : "memory");
    return (uint32_t)(v >> 32) == frame_no;
}
// traps instead of hanging the device if the row never arrives (a broken hierarchy index)
__device__ __forceinline__ float4 load_published_row(const uint64_t* x, uint32_t frame_no)
{
    for (uint32_t spin = 0;; ++spin)
    {
        uint64_t a, b, c, d;
        asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(x) : "memory");
        asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(c), "=l"(d) : "l"(x + 2) : "memory");
        if ((uint32_t)(a >> 32) == frame_no && (uint32_t)(b >> 32) == frame_no && (uint32_t)(c >> 32) == frame_no &&
            (uint32_t)(d >> 32) == frame_no)
            return make_float4(__uint_as_float((uint32_t)a), __uint_as_float((uint32_t)b), __uint_as_float((uint32_t)c),
                               __uint_as_float((uint32_t)d));
        __nanosleep(20);
        if (spin > (1u << 25))
            __trap();
    }
}

// every CTA of the (co-resident) grid has arrived; the arrivals' earlier writes are visible afterwards
__device__ __forceinline__ void grid_barrier(uint32_t* counter, uint32_t lane)
{
    __syncwarp();
    if (lane == 0)
    {
        __threadfence();
        atomicAdd(counter, 1u);
        for (uint32_t spin = 0; ld_acquire(counter) < gridDim.x; ++spin)
        {
            __nanosleep(64);
            if (spin > (1u << 24))
                __trap();
        }
        __threadfence();
    }
    __syncwarp();
}

// all input arrays of one tile -> stage, by one thread (mirrors are padded to a multiple of PAD
// entities: full-tile copies are in bounds; every offset is a multiple of 16 bytes because TILE is)
template <bool DEEP>
__device__ __forceinline__ void stage_tile(const SceneArrays& A, uint32_t base, TileStage& S, uint64_t* bar)
{
    mbar_expect_tx(bar, STAGE_IN_BYTES + (DEEP ? (uint32_t)sizeof(TilePlan) : 0u));
    if (DEEP)
        bulk_g2s(&reinterpret_cast<TileStageDeep&>(S).plan, A.plans + base / TILE, sizeof(TilePlan), bar);
    bulk_g2s(S.mat, A.local + (size_t)base * 4, TILE * 64, bar);
    bulk_g2s(S.bv, A.bv_in + (size_t)base * 6, TILE * 24, bar);
    bulk_g2s(S.ent, A.entities + base, TILE * 8, bar);
    bulk_g2s(S.sf, A.state_flags + base, TILE * 8, bar);
    bulk_g2s(S.par, A.parents + base, TILE * 4, bar);
}

// the next chunk of tiles for this CTA: CTAs are spread over TILE_COUNTERS counters so that same-address
// atomics stay far from their serial rate even with one tile per visit
__device__ __forceinline__ uint32_t next_chunk(uint32_t* ctrl)
{
    const uint32_t c = blockIdx.x % TILE_COUNTERS;
    return atomicAdd(ctrl + CTRL_TILE + c * 32, 1u) * TILE_COUNTERS + c;
}

// loop C (ecs_scene.cpp:1529-1556) for the children whose chain leaves their tile: each folds its box into every
// ancestor reachable through allocated links.  The reference's descending sweep produces exactly that union; min /
// max are exact, so the atomics give the same values in any order.  Every lane of the warp calls it (n = ~0u for
// none): lanes whose FIRST fold goes to the same parent — consecutive children of one node, the common case — reduce
// their boxes among themselves and one of them goes on with the union: the 65,536 boxes under the torus demo's
// master node cost 2 K atomics per component instead of 64 K on one address.  Further up every surviving lane
// walks on its own (merging at every level was measured and is slower for deep chains: the warp then moves in
// lockstep through dependent loads, 0.14 ms instead of 0.07 ms for config 2's 260 K boundary children).
__device__ __forceinline__ void expand_up_warp(const SceneArrays& A, uint32_t n, uint32_t lane)
{
    bool     act = n < A.n && n != 0;
    uint32_t p = 0xFFFFFF00u | lane; // lanes with nothing to fold never match anybody
    uint32_t key[6] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u}; // order-preserving keys of the box; NaN = never folds
    if (act)
    {
        const float* own = A.bv_out + (size_t)n * 7; // may already include descendants: a superset that reaches the same ancestors
#pragma unroll
        for (int k = 0; k < 6; ++k)
        {
            const float v = __ldcg(own + k);
            if (v == v)
                key[k] = float_key(v);
        }
        if (!(A.entities[n] & CMP_ALLOCATED)) // :1531
            act = false;
        else
        {
            p = A.parents[n];
            if (p >= n) // :1535 (p > n cannot happen in a valid scene, ecs_utilities.cpp:408-415)
                act = false, p = 0xFFFFFF00u | lane;
        }
    }
    const uint32_t group = __match_any_sync(0xFFFFFFFFu, p);
    if (!act)
        return;
#pragma unroll
    for (int k = 0; k < 3; ++k)
    {
        key[k] = __reduce_min_sync(group, key[k]);
        key[3 + k] = __reduce_max_sync(group, key[3 + k]);
    }
    if (lane != (uint32_t)__ffs(group) - 1u)
        return; // the group's first lane carries the union
    float b[6];
#pragma unroll
    for (int k = 0; k < 6; ++k)
        b[k] = key_float(key[k]); // a component that was NaN everywhere compares false below: never folds
    uint32_t cur = n;
    for (;;)
    {
        // The ancestor's box is read from L2 (never from a line this SM cached before the other CTAs' boxes arrived).
        // A value read here can only be LOOSER than the current one (min / max move one way during the pass), which
        // costs a redundant atomic, never a missed one.
        float* pb = A.bv_out + (size_t)p * 7;
        bool   improved = false;
#pragma unroll
        for (int k = 0; k < 3; ++k)
        {
            if (b[k] < __ldcg(pb + k))
                atomic_min_float(&pb[k], b[k]), improved = true;
            if (b[3 + k] > __ldcg(pb + 3 + k))
                atomic_max_float(&pb[3 + k], b[3 + k]), improved = true;
        }
        // The walk ends where the ancestor's box already holds this one: whoever put those values there — another
        // walker, which is on its way up with values at least as tight, or the streaming pass, whose in-tile folds are
        // carried on by the boundary child above them — takes them to the ancestors further up.
        if (!improved)
            return;
        cur = p;
        if (!(A.entities[cur] & CMP_ALLOCATED)) // :1531
            return;
        p = A.parents[cur];
        if (p >= cur) // :1535
            return;
    }
}

// Loop C across tiles for the wavefront form, from ancestor LISTS instead of parent links.  The streaming pass has
// folded every box into its direct parent inside the tile, so the box of a boundary child (direct parent in an earlier
// tile) holds its whole in-tile subtree, and every descendant of a node lies in the in-tile subtree of exactly one
// boundary child below that node or in the node's own: folding each boundary child's box into ALL its ancestors gives
// the reference's result.  Walking the parent links costs two dependent global loads per level (measured: 0.04-0.07 ms
// of pure latency for 100-260 k children); the hierarchy index therefore keeps each boundary child's ancestors as a
// flat list (nearest first).  A warp takes five children at once: lane = (child g of 5, box component k of 6), so the
// six floats of an ancestor's box are ONE coalesced request to read and one to fold (a lane per (child, ancestor) with
// six separate 4-byte accesses was bound by L2 requests: 0.12 ms for config 2's 1.7 M pairs), and four ancestors of
// each child are in flight together (the list gives their indices up front: three round trips per four levels; a
// warp per child and one level at a time was bound by latency x children: 0.07 ms).
// A link is followed only out of an allocated node (:1531): a child's lanes stop behind its first unallocated ancestor.
constexpr uint32_t XL_CHILDREN = 5, XL_LEVELS = 4;
__device__ __forceinline__ void expand_lists_warp(const SceneArrays& A, uint32_t i0, uint32_t lane)
{
    const uint32_t g = lane / 6u, k = lane - g * 6u;
    const uint32_t i = i0 + g;
    uint32_t       c = 0, o0 = 0, len = 0;
    if (g < XL_CHILDREN && i < A.num_cross)
    {
        c = A.cross[i];
        o0 = A.anc_off[i];
        len = A.anc_off[i + 1] - o0;
        if (!(A.entities[c] & CMP_ALLOCATED)) // :1531
            len = 0;
    }
    const float b = len ? __ldcg(A.bv_out + (size_t)c * 7 + k) : 0.0f;
    bool        dead = false;
#pragma unroll 1
    for (uint32_t j0 = 0; __any_sync(0xFFFFFFFFu, j0 < len && !dead); j0 += XL_LEVELS)
    {
        uint32_t a[XL_LEVELS];
        bool     has[XL_LEVELS], passes[XL_LEVELS];
        float    cur[XL_LEVELS];
#pragma unroll
        for (uint32_t u = 0; u < XL_LEVELS; ++u)
        {
            has[u] = !dead && j0 + u < len;
            a[u] = has[u] ? A.anc[o0 + j0 + u] : 0u;
        }
#pragma unroll
        for (uint32_t u = 0; u < XL_LEVELS; ++u)
        {
            passes[u] = has[u] ? (A.entities[a[u]] & CMP_ALLOCATED) != 0 : true; // the ancestor hands boxes on to ITS parent
            cur[u] = has[u] ? __ldcg(A.bv_out + (size_t)a[u] * 7 + k) : b;
        }
#pragma unroll
        for (uint32_t u = 0; u < XL_LEVELS; ++u)
        {
            if (has[u] && !dead)
            {
                float* pb = A.bv_out + (size_t)a[u] * 7 + k;
                if (k < 3)
                {
                    if (b < cur[u]) // NaN compares false: never folds
                        atomic_min_float(pb, b);
                }
                else if (b > cur[u])
                    atomic_max_float(pb, b);
                dead = !passes[u];
            }
        }
    }
}

__global__ void __launch_bounds__(256) expand_lists_kernel(const SceneArrays A)
{
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i0 = warp * XL_CHILDREN; i0 < A.num_cross; i0 += warps * XL_CHILDREN)
        expand_lists_warp(A, i0, threadIdx.x & 31u);
}

// ancestor lists of the boundary children (index build)
__global__ void ancestor_count_kernel(const uint32_t* __restrict__ parents, const uint32_t* __restrict__ cross, uint32_t num_cross,
                                      uint32_t* __restrict__ counts)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > num_cross)
        return;
    uint32_t cnt = 0;
    if (i < num_cross)
        for (uint32_t cur = cross[i];;)
        {
            const uint32_t p = parents[cur];
            if (p >= cur) // :1535
                break;
            ++cnt;
            cur = p;
        }
    counts[i] = cnt; // counts[num_cross] = 0: the exclusive scan leaves the total there
}

__global__ void ancestor_fill_kernel(const uint32_t* __restrict__ parents, const uint32_t* __restrict__ cross, uint32_t num_cross,
                                     const uint32_t* __restrict__ offsets, uint32_t* __restrict__ anc)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= num_cross)
        return;
    uint32_t at = offsets[i];
    for (uint32_t cur = cross[i];;)
    {
        const uint32_t p = parents[cur];
        if (p >= cur)
            break;
        anc[at++] = p;
        cur = p;
    }
}

template <int MODE, bool DEEP, bool TAIL>
__global__ void __launch_bounds__(32, (DEEP || TAIL) ? 24 : 27)
    frame_kernel(const SceneArrays A, const __grid_constant__ ViewSet VS, uint32_t nv, uint32_t num_tiles)
{
    extern __shared__ __align__(128) unsigned char s_raw[];
    TileStage&          S = *reinterpret_cast<TileStage*>(s_raw);
    __shared__ uint64_t s_bar;
    __shared__ uint8_t  s_child[DEEP ? 4 : TILE]; // shallow: tile-relative indices of the tile's non-root entities, ascending
    __shared__ uint8_t  s_wr[TILE];    // 1: world matrix rewritten this frame, 0: untouched

    const uint32_t lane = threadIdx.x;
    const uint32_t lt = (1u << lane) - 1u;
    const f32x2    one = pack2(VS.one[0], VS.one[1]);

    // scene->renderable_extents (:1524-1525) accumulates per lane over all tiles of this CTA as
    // order-preserving integer keys; min/max are exact, so any order gives the reference's bits
    uint32_t kmin[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, kmax[3] = {0u, 0u, 0u};

    // Tiles are handed out by global counters (TILE_CHUNK tiles per visit, asked for one tile ahead), not by
    // a fixed stride: the CTAs of the whole GPU then work inside one narrow window that moves through the
    // arrays (DRAM page locality: 0.77 ms with the fixed stride, 0.60 ms like this), and a CTA that becomes
    // resident late — or never, when another kernel (the exchange of the previous frame's lists) holds
    // part of an SM — takes fewer tiles instead of leaving a tail.
    uint32_t tile = 0;
    if (lane == 0)
    {
        mbar_init(&s_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        fence_async_smem();
        tile = next_chunk(A.ctrl) * TILE_CHUNK;
    }
    tile = __shfl_sync(0xFFFFFFFFu, tile, 0);

    for (uint32_t it = 0; tile < num_tiles; ++it)
    {
        const uint32_t base = tile * TILE;
        uint32_t       next = 0;
        float*         s_bv = reinterpret_cast<float*>(S.bv);

        if (lane == 0)
        {
            bulk_wait_read(); // the previous tile's outgoing copies have read their shared-memory source
            stage_tile<DEEP>(A, base, S, &s_bar);
            // the next tile: the following one of this chunk, or the head of a fresh chunk (one atomic per
            // TILE_CHUNK tiles keeps the counter off the critical path; consumed at the end of the iteration)
            next = ((tile + 1) % TILE_CHUNK) ? tile + 1 : next_chunk(A.ctrl) * TILE_CHUNK;
        }
        uint64_t trace_t0 = 0;
        if (DEEP && A.trace)
            trace_t0 = global_ns();
        mbar_wait(&s_bar, it & 1);

        uint32_t rel[EPL], e[EPL], parent[EPL];
        uint64_t ef[EPL], sf[EPL];
        bool     active[EPL], skipped[EPL], is_child[EPL];
        bool     any_dirty = false;
#pragma unroll
        for (int g = 0; g < EPL; ++g)
        {
            rel[g] = g * 32 + lane;
            e[g] = base + rel[g];
            active[g] = e[g] < A.n;
            parent[g] = S.par[rel[g]];
            ef[g] = S.ent[rel[g]];
            sf[g] = S.sf[rel[g]];
            skipped[g] = active[g] && world_is_skipped(ef[g], sf[g], A.has_rb, e[g]);
            is_child[g] = active[g] && parent[g] != e[g]; // loop A skips `skipped` ones, loop C does not
            any_dirty |= active[g] && ((ef[g] & CMP_TRANSFORM) || (sf[g] & STATE_SYNC_PHYSICS_TRANSFORM));
            s_wr[rel[g]] = (active[g] && !skipped[g]) ? 1 : 0;
        }
        if (any_dirty)
            *A.dirty = 1u;

        uint32_t       nc = 0;                                         // shallow: children of the tile
        uint32_t       nrounds = 0;                                    // wavefront: rounds of the tile's plan
        if (!DEEP)
        {
            // ---- loop A for the non-roots, as (entity, row) work items spread over all 32 lanes: a tile
            // with 6 children runs 24 lanes for a quarter of the time instead of 6 lanes for all of it.
            // Rounds of 8 children go from the highest index down and write their rows in place: an
            // entity's ancestors have lower indices (ecs_utilities.cpp:408-415), so the locals a later
            // round still needs are never overwritten before it reads them.
            const uint32_t cb0 = __ballot_sync(0xFFFFFFFFu, is_child[0]);
            const uint32_t cb1 = __ballot_sync(0xFFFFFFFFu, is_child[1]);
            const uint32_t nc0 = __popc(cb0);
            nc = nc0 + __popc(cb1);
            if (is_child[0])
                s_child[__popc(cb0 & lt)] = (uint8_t)rel[0];
            if (is_child[1])
                s_child[nc0 + __popc(cb1 & lt)] = (uint8_t)rel[1];
            __syncwarp();
#pragma unroll 1
            for (int hi = (int)nc; hi > 0; hi -= 8)
            {
                const int lo = hi > 8 ? hi - 8 : 0;
                const int j = lo * 4 + (int)lane;
                float4    row;
                uint32_t  dst = 0xFFFFFFFFu; // float4 index into S.mat, or ~0u: nothing to write
                if (j < hi * 4)
                {
                    const uint32_t c = s_child[j >> 2], r = (uint32_t)j & 3u;
                    if (s_wr[c]) // a physics entity without an rb matrix keeps its world matrix (:1445-1446)
                    {
                        row = chain_row(A, base, S.par[c], c, r, S.mat, S.par, S.ent, S.sf);
                        dst = c * MAT_STRIDE + r;
                    }
                }
                __syncwarp(); // every ancestor read of this round is done
                if (dst != 0xFFFFFFFFu)
                    S.mat[dst] = row;
                __syncwarp();
            }
        }
        else
        {
            // ---- loop A, wavefront form, driven by the tile's static schedule (TilePlan, built when the hierarchy
            // changed).  Roots need no arithmetic (their local matrix is their world matrix, already in the stage);
            // an entity that keeps last frame's matrix (:1445-1446) fetches it into the stage so that its children read
            // it there.  Both publish from here when a later tile waits for them.  Everything else runs in rounds of at
            // most 8 entities x 4 matrix rows whose parents are final: rows of the stage, or — in the last rounds of
            // the plan — published rows of earlier tiles, which the lane waits for packet by packet.
            const TilePlan& P = reinterpret_cast<const TileStageDeep*>(s_raw)->plan;
            const uint64_t  pub = P.publish;
#pragma unroll
            for (int g = 0; g < EPL; ++g)
            {
                if (skipped[g])
                {
                    const float4* w = A.world + (size_t)e[g] * 4;
#pragma unroll
                    for (int r = 0; r < 4; ++r)
                        S.mat[rel[g] * MAT_STRIDE + r] = __ldcg(w + r);
                }
                if (active[g] && ((pub >> rel[g]) & 1ull) && (skipped[g] || parent[g] >= e[g]))
                {
#pragma unroll
                    for (int r = 0; r < 4; ++r)
                        publish_row(A.xbuf + (size_t)e[g] * 16 + r * 4, S.mat[rel[g] * MAT_STRIDE + r], A.frame_no);
                }
            }
            __syncwarp();
            nrounds = P.nrounds;
            const uint32_t k = lane >> 2, r = lane & 3u;
#pragma unroll 1
            for (uint32_t rd = 0; rd < nrounds; ++rd)
            {
                const uint32_t start = P.rstart[rd], cnt = P.rstart[rd + 1] - start;
                bool           valid = k < cnt;
                uint32_t       c = 0, item = 0;
                float4         row = make_float4(0.f, 0.f, 0.f, 0.f);
                if (valid)
                {
                    item = P.order[start + k];
                    c = item & 63u;
                    valid = s_wr[c] != 0; // :1445-1446 last frame's world matrix stands (fetched and published above)
                }
                if (valid)
                {
                    const uint32_t par = S.par[c]; // < base + c: roots are not scheduled
                    const float4   prow = par >= base ? S.mat[(par - base) * MAT_STRIDE + r]
                                                      : load_published_row(A.xbuf + (size_t)par * 16 + r * 4, A.frame_no);
                    row = mul_row(prow, load_mat_shared(S.mat, c)); // :1470
                }
                __syncwarp(); // the four lanes of an entity have read its local matrix
                if (valid)
                {
                    S.mat[c * MAT_STRIDE + r] = row;
                    if (item & 0x80u)
                        publish_row(A.xbuf + (size_t)(base + c) * 16 + r * 4, row, A.frame_no);
                }
                __syncwarp();
            }
            if (A.trace && lane == 0) // diagnostic timeline (PMB_OPT_TRACE): when the tile started, when loop A was through
            {
                uint32_t smid;
                asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
                uint64_t* tr = A.trace + (size_t)tile * 4;
                tr[0] = trace_t0;
                tr[1] = global_ns();
                tr[2] = (uint64_t)smid | ((uint64_t)nrounds << 32);
                tr[3] = (uint64_t)P.nlocal | ((uint64_t)blockIdx.x << 32);
            }
        }

        // ---- loop B: bounds, pos_extent, scene extents
        float bmin[EPL][3], bmax[EPL][3];
#pragma unroll
        for (int g = 0; g < EPL; ++g)
#pragma unroll
            for (int i = 0; i < 3; ++i)
            {
                bmin[g][i] = s_bv[rel[g] * 6 + i];
                bmax[g][i] = s_bv[rel[g] * 6 + 3 + i];
            }
        __syncwarp(); // the staged min/max extents are in registers: the buffer becomes the 28 B output records
        float* s_bvout = s_bv;

        bool   pass[EPL];
        CullIn ci[EPL];
#pragma unroll
        for (int g = 0; g < EPL; ++g)
        {
            const bool bone = (ef[g] & CMP_BONE) != 0;
            Bounds     o = {};
            pass[g] = false;
            if (active[g])
            {
                float4 w0, w1, w2;
                if (skipped[g] && !DEEP)
                {
                    // :1445-1446 `continue`: last frame's world matrix stands
                    const float4* p = A.world + (size_t)e[g] * 4;
                    w0 = p[0], w1 = p[1], w2 = p[2];
                }
                else
                {
                    w0 = S.mat[rel[g] * MAT_STRIDE + 0];
                    w1 = S.mat[rel[g] * MAT_STRIDE + 1];
                    w2 = S.mat[rel[g] * MAT_STRIDE + 2];
                }
                if (bone)
                {
                    // :1494-1498 tmin = tmax = translation; radius and pos_extent keep their values
                    o.tmin[0] = o.tmax[0] = w0.w;
                    o.tmin[1] = o.tmax[1] = w1.w;
                    o.tmin[2] = o.tmax[2] = w2.w;
                    o.radius = A.bv_out[(size_t)e[g] * 7 + 6];
                }
                else
                {
                    bounds_from_world(w0, w1, w2, bmin[g], bmax[g], one, o);
                    float4* pe = A.pos_extent + (size_t)e[g] * 2;
                    pe[0] = make_float4(o.pos[0], o.pos[1], o.pos[2], 0.0f); // pos.w is never written by the reference; the arrays start zeroed
                    pe[1] = make_float4(o.ext[0], o.ext[1], o.ext[2], o.radius);
                    if (ef[g] & CMP_GEOMETRY) // :1520
                    {
#pragma unroll
                        for (int i = 0; i < 3; ++i)
                        {
                            // NaN boxes are left out (in the reference a NaN is dropped again by the next comparison)
                            if (o.tmin[i] == o.tmin[i])
                                kmin[i] = min(kmin[i], float_key(o.tmin[i]));
                            if (o.tmax[i] == o.tmax[i])
                                kmax[i] = max(kmax[i], float_key(o.tmax[i]));
                        }
                    }
                }
                // stage the 28 B record so the tile leaves as one bulk copy
#pragma unroll
                for (int i = 0; i < 3; ++i)
                {
                    s_bvout[rel[g] * 7 + i] = o.tmin[i];
                    s_bvout[rel[g] * 7 + 3 + i] = o.tmax[i];
                }
                s_bvout[rel[g] * 7 + 6] = o.radius;
                if (MODE != MODE_UPDATE)
                {
                    pass[g] = filter_pass(ef[g], sf[g]);
                    if (pass[g] && bone)
                    {
                        // a renderable bone is culled against the pos_extent it kept
                        float4 p0 = A.pos_extent[(size_t)e[g] * 2], p1 = A.pos_extent[(size_t)e[g] * 2 + 1];
                        o.pos[0] = p0.x, o.pos[1] = p0.y, o.pos[2] = p0.z;
                        o.ext[0] = p1.x, o.ext[1] = p1.y, o.ext[2] = p1.z;
                        o.radius = p1.w;
                    }
                }
            }
            if (MODE != MODE_UPDATE)
                cull_prepare<MODE>(o.pos, o.ext, o.radius, ci[g]);
        }

        // ---- the plane tests of every view for both entities; lane v keeps view v's two ballot words
        if (MODE != MODE_UPDATE)
        {
            uint32_t mine[EPL] = {0u, 0u};
#if PMB_ROLLED_VIEWS
#pragma unroll 1
            for (uint32_t v = 0; v < nv; ++v) // rolled: the per-view code is fetched once per tile and reused
            {
                bool in[EPL];
                cull_view<MODE, EPL>(VS, v, ci, in);
#pragma unroll
                for (int g = 0; g < EPL; ++g)
                {
                    const uint32_t b = __ballot_sync(0xFFFFFFFFu, in[g] & pass[g]);
                    mine[g] = (lane == v) ? b : mine[g];
                }
            }
#else
#pragma unroll
            for (uint32_t v = 0; v < PMB_MAX_VIEWS; ++v) // unrolled: the view's constants are LDCU loads at fixed offsets
            {
                if (v < nv) // uniform
                {
                    bool in[EPL];
                    cull_view<MODE, EPL>(VS, v, ci, in);
#pragma unroll
                    for (int g = 0; g < EPL; ++g)
                    {
                        const uint32_t b = __ballot_sync(0xFFFFFFFFu, in[g] & pass[g]);
                        mine[g] = (lane == v) ? b : mine[g];
                    }
                }
            }
#endif
            if (lane < nv) // words (base / 32) and (base / 32) + 1 of the view's bitmap row: one 8-byte store
                *reinterpret_cast<uint2*>(A.bitmap + (size_t)lane * A.bitmap_stride + (base >> 5)) = make_uint2(mine[0], mine[1]);
            if (TAIL && lane < PMB_MAX_VIEWS)
            {
                // the tile's visible count per view (one 64 B line), also added to the totals of its group of 32 tiles
                // and of its chunk of 1024: behind the barrier a tile's list offset is three masked 32-wide sums
                const uint32_t cnt = __popc(mine[0]) + __popc(mine[1]);
                A.tile_counts[(size_t)tile * PMB_MAX_VIEWS + lane] = cnt;
                if (cnt)
                {
                    atomicAdd(A.group_totals + (size_t)(tile >> 5) * PMB_MAX_VIEWS + lane, cnt);
                    atomicAdd(A.chunk_totals + (size_t)(tile >> 10) * PMB_MAX_VIEWS + lane, cnt);
                }
            }
        }

        // a tile whose world matrices are all rewritten this frame leaves as one bulk copy
        bool mine_plain = true;
#pragma unroll
        for (int g = 0; g < EPL; ++g)
            mine_plain &= active[g] && !skipped[g];
        const bool plain = __all_sync(0xFFFFFFFFu, mine_plain) != 0;

        // ---- loop C inside the tile (ecs_scene.cpp:1529-1556): every allocated non-root entity folds
        // its box into each ancestor reachable through allocated links; here for the ancestors that
        // live in this tile.  min/max are exact, so the atomics give the same values in any order.
        if (!DEEP)
        {
            // shallow: (entity, box component) work items, six lanes per non-root entity, each walking its chain
            for (uint32_t j = lane; j < nc * 6; j += 32)
            {
                const uint32_t c = s_child[j / 6], k = j % 6;
                const float    v = s_bvout[c * 7 + k]; // may already include descendants: a superset that reaches the same ancestors
                uint32_t       cur = c;
                for (;;)
                {
                    if (!(S.ent[cur] & CMP_ALLOCATED)) // :1531
                        break;
                    uint32_t prel = S.par[cur] - base;
                    if (prel >= cur) // :1535 root; or the link leaves the tile (expand_up's job: p < base wraps
                        break;       // to a huge value); or an upward link, which ecs_utilities.cpp:408-415 rules out
                    float* pb = s_bvout + prel * 7 + k;
                    if (k < 3)
                    {
                        if (v < *pb)
                            atomic_min_float(pb, v);
                    }
                    else if (v > *pb)
                        atomic_max_float(pb, v);
                    cur = prel;
                }
            }
        }
        else
        {
            // wavefront: the plan's rounds in reverse — an entity's parent sits in an earlier round, so when a round
            // runs its boxes already hold everything below them inside the tile, and each entity folds into its DIRECT
            // parent only.  Four lanes per entity: lane q takes component q (tmin.xyz, tmax.x), lanes 0-1 also 4 + q.
            const TilePlan& P = reinterpret_cast<const TileStageDeep*>(s_raw)->plan;
            const uint32_t  k = lane >> 2, q = lane & 3u;
            __syncwarp();
#pragma unroll 1
            for (uint32_t rd = nrounds; rd-- > 0;)
            {
                const uint32_t start = P.rstart[rd], cnt = P.rstart[rd + 1] - start;
                if (k < cnt)
                {
                    const uint32_t c = P.order[start + k] & 63u;
                    const uint32_t par = S.par[c];
                    if (par >= base && (S.ent[c] & CMP_ALLOCATED)) // :1531; links into earlier tiles are expand_up's
                    {
                        const float* cb = s_bvout + c * 7;
                        float*       pb = s_bvout + (par - base) * 7;
                        const float  v = cb[q];
                        if (q < 3)
                        {
                            if (v < pb[q])
                                atomic_min_float(pb + q, v);
                        }
                        else if (v > pb[q])
                            atomic_max_float(pb + q, v);
                        if (q < 2)
                        {
                            const float v2 = cb[4 + q];
                            if (v2 > pb[4 + q])
                                atomic_max_float(pb + 4 + q, v2);
                        }
                    }
                }
                __syncwarp();
            }
        }
        fence_async_smem(); // this lane's shared-memory writes -> visible to the bulk-copy engine
        __syncwarp();

        // ---- tile stores.  The world matrices leave as one bulk copy when every matrix of the tile is in the stage:
        // shallow form, all rewritten; wavefront form, any full tile (an entity that keeps last frame's matrix has
        // fetched it into the stage, so it is rewritten with its own value).
        const bool full = base + TILE <= A.n;
        const bool bulk_world = DEEP ? full : plain;
        if (lane == 0)
        {
            if (bulk_world)
                bulk_s2g(A.world + (size_t)base * 4, S.mat, TILE * 64);
            if (full)
                bulk_s2g(A.bv_out + (size_t)base * 7, s_bvout, TILE * 28);
            bulk_commit();
        }
        if (!bulk_world || !full) // uniform: skipped entities or the ragged last tile
        {
            if (!bulk_world)
            {
                float4* gw = A.world + (size_t)base * 4;
                for (uint32_t q = lane; q < TILE * 4; q += 32)
                    if (s_wr[q >> 2] == 1)
                        gw[q] = S.mat[q];
            }
            if (!full)
            {
                // 28 B records: whole float4 chunks that lie below entity n; the ragged last chunk by scalars
                const uint32_t nfl = (A.n - base) * 7;
                float4*        gb = reinterpret_cast<float4*>(A.bv_out + (size_t)base * 7);
                for (uint32_t q = lane; q * 4 < nfl; q += 32)
                {
                    if (q * 4 + 4 <= nfl)
                        gb[q] = S.bv[q];
                    else
                        for (uint32_t t = q * 4; t < nfl; ++t)
                            A.bv_out[(size_t)base * 7 + t] = s_bvout[t];
                }
            }
            __syncwarp(); // s_wr is rewritten by the next tile
        }
        tile = __shfl_sync(0xFFFFFFFFu, next, 0);
    }
    if (lane == 0)
    {
        if (TAIL)
            bulk_wait_done(); // the bound records are in global memory before the barrier: other CTAs fold into them
        else
            bulk_wait_read(); // shared memory must outlive the outgoing copies
    }

    // ---- renderable extents of this CTA's tiles -> scene
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
        uint32_t a = __reduce_min_sync(0xFFFFFFFFu, kmin[i]);
        uint32_t b = __reduce_max_sync(0xFFFFFFFFu, kmax[i]);
        if (lane == 0)
        {
            if (a != 0xFFFFFFFFu)
            {
                float v = key_float(a);
                if (v < A.extents[i]) // cheap pre-check; the atomic decides
                    atomic_min_float(&A.extents[i], v);
            }
            if (b != 0u)
            {
                float v = key_float(b);
                if (v > A.extents[3 + i])
                    atomic_max_float(&A.extents[3 + i], v);
            }
        }
    }

    if (TAIL)
    {
        // ---- everything that needs the whole scene's tiles: behind a grid barrier instead of in further launches
        grid_barrier(A.ctrl + CTRL_BAR0, lane);
        const uint32_t cta = blockIdx.x, ctas = gridDim.x;
        // (1) the lists' lengths = the sums of the chunk totals
        if (MODE != MODE_UPDATE && cta == 0 && lane < nv)
        {
            uint32_t total = 0;
            for (uint32_t c = 0; c <= ((num_tiles - 1) >> 10); ++c)
                total += __ldcg(A.chunk_totals + (size_t)c * PMB_MAX_VIEWS + lane);
            A.counts[lane] = total;
        }
        // ... and the next cull launch's group / chunk totals start from zero (the two sets alternate between the
        // launches that add to them).  The WHOLE other set is cleared, not just this launch's tile range: the launch
        // that last added to it may have run on more entities than this one (33 KB of stores over the grid).
        if (MODE != MODE_UPDATE)
        {
            constexpr uint32_t groups = TAIL_MAX_ENTITIES / TILE / 32, chunks = TAIL_MAX_ENTITIES / TILE / 1024;
            for (uint32_t i = cta * 32 + lane; i < groups * PMB_MAX_VIEWS; i += ctas * 32)
                A.group_totals_next[i] = 0u;
            for (uint32_t i = cta * 32 + lane; i < chunks * PMB_MAX_VIEWS; i += ctas * 32)
                A.chunk_totals_next[i] = 0u;
        }
        // (2) loop C across tiles: the children whose chain leaves their tile fold into their ancestors
        if (DEEP)
            for (uint32_t i0 = cta * XL_CHILDREN; i0 < A.num_cross; i0 += ctas * XL_CHILDREN)
                expand_lists_warp(A, i0, lane);
        else
            for (uint32_t i0 = cta * 32; i0 < A.num_cross; i0 += ctas * 32)
                expand_up_warp(A, i0 + lane < A.num_cross ? A.cross[i0 + lane] : 0xFFFFFFFFu, lane);
        // (3) loop A's flag clears (ecs_scene.cpp:1408-1412, :1441): every tile has read the pre-update flags
        if (ld_acquire(A.dirty) != 0u)
        {
            for (uint32_t i = cta * 32 + lane; i < A.n; i += ctas * 32)
            {
                const uint64_t ef = A.entities[i], sf = A.state_flags[i];
                if (ef & CMP_TRANSFORM)
                    A.entities[i] = ef & ~CMP_TRANSFORM;
                if (sf & STATE_SYNC_PHYSICS_TRANSFORM)
                    A.state_flags[i] = sf & ~STATE_SYNC_PHYSICS_TRANSFORM;
            }
        }
        // (4) draw lists: tile t's visible entities of view v start where the tiles below it end; ascending inside
        // the tile and across tiles = filter_entities_scalar / frustum_cull_*_scalar's sb_push order
        if (MODE != MODE_UPDATE)
        {
            for (uint32_t t = cta; t < num_tiles; t += ctas)
            {
                uint2 w = make_uint2(0u, 0u);
                if (lane < nv)
                    w = __ldcg(reinterpret_cast<const uint2*>(A.bitmap + (size_t)lane * A.bitmap_stride + (size_t)t * 2));
                if (!__any_sync(0xFFFFFFFFu, (w.x | w.y) != 0u))
                    continue;
                const uint32_t group = t >> 5, chunk = t >> 10;
                const uint32_t tile_l = (t & ~31u) + lane, group_l = (chunk << 5) + lane; // what this lane adds up
#pragma unroll 1
                for (uint32_t v = 0; v < nv; ++v)
                {
                    const uint32_t w0 = __shfl_sync(0xFFFFFFFFu, w.x, v), w1 = __shfl_sync(0xFFFFFFFFu, w.y, v);
                    if ((w0 | w1) == 0u) // uniform
                        continue;
                    uint32_t below = 0;
                    if (tile_l < t)
                        below += __ldcg(A.tile_counts + (size_t)tile_l * PMB_MAX_VIEWS + v);
                    if (group_l < group)
                        below += __ldcg(A.group_totals + (size_t)group_l * PMB_MAX_VIEWS + v);
                    if (lane < chunk)
                        below += __ldcg(A.chunk_totals + (size_t)lane * PMB_MAX_VIEWS + v);
                    uint32_t* out = A.lists + (size_t)v * A.list_stride + __reduce_add_sync(0xFFFFFFFFu, below);
                    if (w0 & (1u << lane))
                        out[__popc(w0 & lt)] = t * TILE + lane;
                    if (w1 & (1u << lane))
                        out[__popc(w0) + __popc(w1 & lt)] = t * TILE + 32 + lane;
                }
            }
        }
    }

    // ---- the last CTA to leave arms the control block for the next launch: tile counters, barrier, and the OTHER
    // slot of the double-buffered extents / dirty marker (this launch's slot is still to be read).  One lane only:
    // warp-wide code here costs the whole kernel its uniform-datapath code generation (1600 -> 1784 SASS
    // instructions, 0.60 -> 0.69 ms at 16 Mi entities).
    __syncwarp();
    if (lane == 0)
    {
        __threadfence();
        if (atomicAdd(A.ctrl + CTRL_EXIT, 1u) == gridDim.x - 1)
        {
            for (uint32_t c = 0; c < TILE_COUNTERS; ++c)
                A.ctrl[CTRL_TILE + c * 32] = 0u;
            A.ctrl[CTRL_BAR0] = 0u;
            A.ctrl[CTRL_EXIT] = 0u;
            // ecs_scene.cpp:1482-1483
            A.extents_next[0] = A.extents_next[1] = A.extents_next[2] = FLT_MAX;
            A.extents_next[3] = A.extents_next[4] = A.extents_next[5] = -FLT_MAX;
            *A.dirty_next = 0u;
        }
    }
}

// both slots of the double-buffered renderable extents (ecs_scene.cpp:1482-1483), once per scene handle
__global__ void extents_init_kernel(float* extents2)
{
    const uint32_t t = threadIdx.x; // 16 floats: slot 0 at [0, 6), slot 1 at [8, 14)
    if (t < 16)
        extents2[t] = (t & 7u) < 3u ? FLT_MAX : ((t & 7u) < 6u ? -FLT_MAX : 0.0f);
}

// loop A's flag clears (ecs_scene.cpp:1408-1412, :1441), applied after the frame so that every
// thread of the frame saw the same pre-update flags.  No-op unless a flag was seen.
__global__ void clear_flags_kernel(uint64_t* entities, uint64_t* state_flags, uint32_t n, const uint32_t* dirty)
{
    if (*dirty == 0u)
        return;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        uint64_t ef = entities[i], sf = state_flags[i];
        if (ef & CMP_TRANSFORM)
            entities[i] = ef & ~CMP_TRANSFORM;
        if (sf & STATE_SYNC_PHYSICS_TRANSFORM)
            state_flags[i] = sf & ~STATE_SYNC_PHYSICS_TRANSFORM;
    }
}

// loop C across tiles as its own launch (scenes above TAIL_MAX_ENTITIES; smaller ones run it inside frame_kernel)
__global__ void expand_kernel(const SceneArrays A, const uint32_t* __restrict__ children, uint32_t num_children)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; // whole warps: the fold is warp-cooperative
    expand_up_warp(A, i < num_children ? children[i] : 0xFFFFFFFFu, threadIdx.x & 31u);
}

// render_scene_view's filter + cull (ecs_scene.cpp:959-960) for a batch of views, from the
// stored pos_extent (48 B read per entity, one bitmap word written per warp and view)
// DEVICE_VIEWS: the plane constants were produced on the device (shadow_fit_kernel) and are read from global
// memory — one warp-uniform load per constant, served by L1 — instead of from the kernel parameters.
template <int MODE, bool DEVICE_VIEWS>
__global__ void __launch_bounds__(256) cull_kernel(const SceneArrays A, const __grid_constant__ ViewSet VSP,
                                                   const ViewSet* __restrict__ VSG, uint32_t nv)
{
    const ViewSet& VS = DEVICE_VIEWS ? *VSG : VSP;
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; // the grid covers whole warps; no early return (ballots)
    const bool     active = e < A.n;
    bool           pass = false;
    float          pos[3] = {0.f, 0.f, 0.f}, ext[3] = {0.f, 0.f, 0.f}, radius = 0.f;
    if (active)
    {
        uint64_t ef = A.entities[e], sf = A.state_flags[e];
        pass = filter_pass(ef, sf);
        if (pass && MODE != MODE_FILTER)
        {
            float4 p0 = A.pos_extent[(size_t)e * 2], p1 = A.pos_extent[(size_t)e * 2 + 1];
            pos[0] = p0.x, pos[1] = p0.y, pos[2] = p0.z;
            ext[0] = p1.x, ext[1] = p1.y, ext[2] = p1.z;
            radius = p1.w;
        }
    }
    CullIn ci;
    cull_prepare<MODE>(pos, ext, radius, ci);
    const uint32_t lane = threadIdx.x & 31;
    uint32_t       mine = cull_warp<MODE>(VS, nv, pass, ci, lane);
    if (lane < nv)
        A.bitmap[(size_t)lane * A.bitmap_stride + (e >> 5)] = mine;
}

// ------------------------------------------------------------------------------------------------
// compaction: per-view bitmaps -> ascending per-view lists.  grid = (chunks, views); a chunk is
// COMPACT_WORDS bitmap words (32 Ki entities), four consecutive words per thread.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 load_words(const uint32_t* __restrict__ row, uint32_t w0, uint32_t num_words)
{
    // rows are 16-byte aligned and w0 is a multiple of 4; words at or beyond num_words were never written
    uint4 w = make_uint4(0u, 0u, 0u, 0u);
    if (w0 + 4 <= num_words)
        w = *reinterpret_cast<const uint4*>(row + w0);
    else
    {
        if (w0 + 0 < num_words) w.x = row[w0 + 0];
        if (w0 + 1 < num_words) w.y = row[w0 + 1];
        if (w0 + 2 < num_words) w.z = row[w0 + 2];
    }
    return w;
}

__global__ void __launch_bounds__(256) compact_count_kernel(const uint32_t* __restrict__ bitmap, uint32_t bitmap_stride,
                                                           uint32_t num_words, uint32_t num_chunks,
                                                           uint32_t* __restrict__ chunk_counts)
{
    __shared__ uint32_t s_warp[8];
    const uint32_t      v = blockIdx.y;
    const uint32_t      w0 = blockIdx.x * (blockDim.x * 4) + threadIdx.x * 4; // a chunk = 4 words per thread of the CTA
    uint4               w = load_words(bitmap + (size_t)v * bitmap_stride, w0, num_words);
    uint32_t            c = __popc(w.x) + __popc(w.y) + __popc(w.z) + __popc(w.w);
    c = __reduce_add_sync(0xFFFFFFFFu, c);
    if ((threadIdx.x & 31) == 0)
        s_warp[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        uint32_t t = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i)
            t += (i < (int)(blockDim.x >> 5)) ? s_warp[i] : 0u;
        chunk_counts[(size_t)v * num_chunks + blockIdx.x] = t;
    }
}

// one CTA per view: exclusive scan of the chunk counts in place, total -> counts[view_base + v]
__global__ void __launch_bounds__(1024) compact_scan_kernel(uint32_t* __restrict__ chunk_counts, uint32_t num_chunks,
                                                           uint32_t* __restrict__ counts, uint32_t view_base)
{
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    uint32_t*           row = chunk_counts + (size_t)blockIdx.x * num_chunks;
    const uint32_t      lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0)
        s_carry = 0;
    __syncthreads();
    for (uint32_t start = 0; start < num_chunks; start += 1024)
    {
        uint32_t i = start + threadIdx.x;
        uint32_t x = i < num_chunks ? row[i] : 0u;
        uint32_t inc = x;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1)
        {
            uint32_t y = __shfl_up_sync(0xFFFFFFFFu, inc, d);
            if ((int)lane >= d)
                inc += y;
        }
        if (lane == 31)
            s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0)
        {
            uint32_t w = s_warp[lane];
            uint32_t winc = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1)
            {
                uint32_t y = __shfl_up_sync(0xFFFFFFFFu, winc, d);
                if ((int)lane >= d)
                    winc += y;
            }
            s_warp[lane] = winc - w; // exclusive warp offsets
        }
        __syncthreads();
        uint32_t carry = s_carry;
        uint32_t excl = carry + s_warp[warp] + inc - x;
        if (i < num_chunks)
            row[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023)
            s_carry = excl + x;
        __syncthreads();
    }
    if (threadIdx.x == 0)
        counts[view_base + blockIdx.x] = s_carry;
}

// One CTA per (32 Ki-entity chunk, view).  Each warp owns 128 consecutive bitmap words and goes through them
// in four rounds of 32 (lane l takes word 32 * round + l: one coalesced load, and the rounds are in list
// order).  In a round every lane drops the entity indices of its word into the warp's shared-memory
// buffer at its scanned offset, then the warp copies the buffer out with consecutive lanes writing
// consecutive list entries: the stores are whole 128 B lines instead of one 4 B word per lane and run.
__global__ void __launch_bounds__(256) compact_scatter_kernel(const uint32_t* __restrict__ bitmap, uint32_t bitmap_stride,
                                                             uint32_t num_words, uint32_t num_chunks,
                                                             const uint32_t* __restrict__ chunk_offsets,
                                                             uint32_t* __restrict__ lists, size_t list_stride)
{
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_buf[8][1024]; // one round of one warp: at most 32 words x 32 bits
    const uint32_t      v = blockIdx.y;
    const uint32_t      lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t*     row = bitmap + (size_t)v * bitmap_stride;
    const uint32_t      first_word = blockIdx.x * (blockDim.x * 4) + warp * 128; // words at or beyond num_words were never written
    uint32_t            w[4], c = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j)
    {
        const uint32_t idx = first_word + j * 32 + lane;
        w[j] = idx < num_words ? row[idx] : 0u;
        c += __popc(w[j]);
    }
    const uint32_t wc = __reduce_add_sync(0xFFFFFFFFu, c);
    if (lane == 0)
        s_warp[warp] = wc;
    __syncthreads();
    if (wc == 0)
        return;
    uint32_t pos = chunk_offsets[(size_t)v * num_chunks + blockIdx.x];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        pos += (i < (int)warp) ? s_warp[i] : 0u;
    uint32_t* out = lists + (size_t)v * list_stride;
    uint32_t* buf = s_buf[warp];
#pragma unroll
    for (int j = 0; j < 4; ++j)
    {
        const uint32_t cj = __popc(w[j]);
        uint32_t       inc = cj;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1)
        {
            uint32_t y = __shfl_up_sync(0xFFFFFFFFu, inc, d);
            if ((int)lane >= d)
                inc += y;
        }
        const uint32_t total = __shfl_sync(0xFFFFFFFFu, inc, 31);
        if (total == 0) // warp-uniform
            continue;
        uint32_t       at = inc - cj, bits = w[j];
        const uint32_t e0 = (first_word + j * 32 + lane) * 32;
        while (bits)
        {
            buf[at++] = e0 + (__ffs(bits) - 1);
            bits &= bits - 1;
        }
        __syncwarp();
        for (uint32_t i = lane; i < total; i += 32)
            out[pos + i] = buf[i];
        __syncwarp();
        pos += total;
    }
}

// ------------------------------------------------------------------------------------------------
// hierarchy index
// ------------------------------------------------------------------------------------------------
__global__ void depth_kernel(const uint32_t* __restrict__ parents, uint32_t n, uint32_t* __restrict__ depth,
                             uint32_t* __restrict__ max_depth)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t d = 0;
    if (i < n)
    {
        uint32_t cur = i, x = 0; // x = parent links on the way up that lead into an earlier tile
        for (;;)
        {
            uint32_t p = parents[cur];
            if (p == cur || p > cur) // roots are self-parented; p > cur violates children-below-parents
                break;
            ++d;
            x += (p / TILE) != (cur / TILE) ? 1u : 0u;
            cur = p;
        }
        depth[i] = min(d, 0xFFFFu) | (min(x, 0xFFFFu) << 16);
    }
    uint32_t wmax = __reduce_max_sync(0xFFFFFFFFu, d);
    if ((threadIdx.x & 31) == 0 && wmax > *max_depth)
        atomicMax(max_depth, wmax);
}

// The children loop C cannot finish inside their own TILE-entity tile.  direct_only = false (shallow form: the
// streaming pass walks whole chains in shared memory): some ancestor lies in an earlier tile.  direct_only = true
// (wavefront form: it folds level by level into direct parents): the direct parent does.
__global__ void cross_kernel(const uint32_t* __restrict__ parents, uint32_t n, uint32_t direct_only, uint32_t* __restrict__ cross,
                             uint32_t* __restrict__ num_cross)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool     crosses = false;
    if (i < n && parents[i] < i)
    {
        const uint32_t tile_base = i - (i % TILE);
        uint32_t       cur = i;
        for (;;)
        {
            uint32_t p = parents[cur];
            if (p >= cur)
                break;
            if (p < tile_base)
            {
                crosses = true;
                break;
            }
            if (direct_only)
                break;
            cur = p;
        }
    }
    uint32_t lane = threadIdx.x & 31;
    uint32_t lt = (1u << lane) - 1u;
    uint32_t bc = __ballot_sync(0xFFFFFFFFu, crosses);
    if (bc)
    {
        uint32_t pos = 0;
        if (lane == 0)
            pos = atomicAdd(num_cross, __popc(bc));
        pos = __shfl_sync(0xFFFFFFFFu, pos, 0);
        if (crosses)
            cross[pos + __popc(bc & lt)] = i;
    }
}

// re-uploaded parents: copy the staged values over the mirror and report whether any of them differs
__global__ void merge_parents_kernel(const uint32_t* __restrict__ staged, uint32_t* __restrict__ parents, uint32_t count,
                                     uint32_t* __restrict__ changed)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool           diff = false;
    if (i < count)
    {
        const uint32_t v = staged[i];
        diff = parents[i] != v;
        if (diff)
            parents[i] = v;
    }
    if (__any_sync(0xFFFFFFFFu, diff) && (threadIdx.x & 31) == 0)
        *changed = 1u;
}

// wavefront form: bit (p % 64) of publish[p / 64] = entity p has a child in a later tile, so its world matrix is
// published to the exchange buffer (one 64-bit word per tile)
__global__ void publish_mask_kernel(const uint32_t* __restrict__ parents, uint32_t n, unsigned long long* __restrict__ publish)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const uint32_t p = parents[i];
    if (p < i && p < i - (i % TILE))
        atomicOr(publish + p / TILE, 1ull << (p % TILE));
}

// wavefront form: the static schedule of every tile (TilePlan, pm_kernels.h).  One thread per tile — this runs when
// the hierarchy changes, not per frame.  List scheduling in phases: phase x holds the entities with x links into
// earlier tiles on their way up to the root (depth_kernel).  Phase 0 needs nothing from other tiles; phase x waits
// only for phase x - 1 entities of earlier tiles, so whatever the order in which tiles block, the waits of the whole
// frame form chains no longer than the deepest crossing count — a warp that blocked on its first foreign parent
// regardless of phase would serialise the tiles one behind the other (measured: 22 ms instead of 0.1 for 1 Mi
// entities).  Inside a phase a round takes the (at most 8) ready entities with the highest priority = leads to an
// entity a later tile waits for, then longest chain below it inside the tile.
__global__ void __launch_bounds__(64) plan_kernel(const uint32_t* __restrict__ parents, const uint32_t* __restrict__ depth, uint32_t n,
                                                 const uint64_t* __restrict__ publish, TilePlan* __restrict__ plans, uint32_t num_tiles)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= num_tiles)
        return;
    const uint32_t base = t * TILE;
    const uint32_t present = base < n ? min(TILE, n - base) : 0u;
    const uint64_t pub = publish[t];
    int8_t         par[TILE];    // tile-relative parent, -1: root (not scheduled), -2: parent in an earlier tile
    uint8_t        key[TILE];    // priority inside a phase
    uint8_t        ext[TILE];    // phase: links into earlier tiles between the entity and its root (capped)
    int8_t         round_of[TILE]; // -1: final before the rounds (roots), 127: not scheduled yet
    for (uint32_t c = 0; c < TILE; ++c)
    {
        par[c] = -1, key[c] = 0, ext[c] = 0, round_of[c] = -1;
        if (c < present)
        {
            const uint32_t p = parents[base + c];
            ext[c] = (uint8_t)min(depth[base + c] >> 16, 255u);
            if (p < base)
                par[c] = -2, round_of[c] = 127;
            else if (p < base + c)
                par[c] = (int8_t)(p - base), round_of[c] = 127;
        }
    }
    // key = 64 * (it or something below it publishes) + height of the subtree below it (children have higher indices)
    for (uint32_t c = present; c-- > 0;)
    {
        if ((pub >> c) & 1ull)
            key[c] |= 64u;
        if (par[c] >= 0)
        {
            const uint32_t p = (uint32_t)par[c];
            const uint32_t h = min((key[c] & 63u) + 1u, 63u);
            key[p] = (uint8_t)((key[p] & 64u) | (key[c] & 64u) | max(key[p] & 63u, h));
        }
    }
    TilePlan P;
    memset(&P, 0, sizeof(P));
    P.publish = pub;
    uint32_t pos = 0, rd = 0;
    for (uint32_t phase = 0;;)
    {
        // the lowest phase that still has entities (an in-tile parent shares its child's phase: capped counts included)
        uint32_t remaining = 0, next_phase = 0xFFFFFFFFu;
        for (uint32_t c = 0; c < present; ++c)
            if (round_of[c] == 127)
                next_phase = min(next_phase, (uint32_t)ext[c]);
        if (next_phase == 0xFFFFFFFFu)
            break;
        phase = next_phase;
        for (uint32_t c = 0; c < present; ++c)
            remaining += (round_of[c] == 127 && ext[c] == phase) ? 1u : 0u;
        while (remaining)
        {
            uint8_t  sel[8];
            uint32_t cnt = 0;
            for (uint32_t c = 0; c < present; ++c)
            {
                if (round_of[c] != 127 || ext[c] != phase)
                    continue;
                if (par[c] >= 0 && round_of[par[c]] >= (int)rd) // its parent is not final before this round
                    continue;
                // keep the 8 best so far, highest key first, lower index first among equals
                uint32_t at = cnt;
                while (at > 0 && key[sel[at - 1]] < key[c])
                    --at;
                if (at >= 8)
                    continue;
                const uint32_t last = cnt < 8 ? cnt : 7;
                for (uint32_t i = last; i > at; --i)
                    sel[i] = sel[i - 1];
                sel[at] = (uint8_t)c;
                if (cnt < 8)
                    ++cnt;
            }
            if (cnt == 0)
            {
                // cannot happen (the lowest unscheduled entity of a phase is always ready); if it ever does, one entity
                // per round in ascending order is always a valid schedule
                for (uint32_t c = 0; c < present; ++c)
                    if (round_of[c] == 127 && ext[c] == phase)
                    {
                        P.rstart[rd] = (uint8_t)pos;
                        P.order[pos++] = (uint8_t)(c | (((pub >> c) & 1ull) ? 0x80u : 0u));
                        round_of[c] = (int8_t)rd++;
                    }
                break;
            }
            P.rstart[rd] = (uint8_t)pos;
            for (uint32_t i = 0; i < cnt; ++i)
            {
                const uint32_t c = sel[i];
                P.order[pos++] = (uint8_t)(c | (((pub >> c) & 1ull) ? 0x80u : 0u));
                round_of[c] = (int8_t)rd;
            }
            remaining -= cnt;
            ++rd;
        }
        if (phase == 0)
            P.nlocal = (uint8_t)rd;
    }
    P.rstart[rd] = (uint8_t)pos;
    P.nrounds = (uint8_t)rd;
    plans[t] = P;
}

// ------------------------------------------------------------------------------------------------
// cmp_bounding_volume AoS (13 floats) <-> split mirrors
// ------------------------------------------------------------------------------------------------
__global__ void bv_split_kernel(const float* __restrict__ aos, uint32_t count, float* __restrict__ bv_in, float* __restrict__ bv_out)
{
    // one thread per float of the AoS stream: coalesced reads, near-coalesced writes
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)count * 13)
        return;
    uint32_t ent = (uint32_t)(i / 13), f = (uint32_t)(i % 13);
    float    v = aos[i];
    if (f < 6)
        bv_in[(size_t)ent * 6 + f] = v;
    else
        bv_out[(size_t)ent * 7 + (f - 6)] = v;
}

__global__ void bv_merge_kernel(const float* __restrict__ bv_in, const float* __restrict__ bv_out, uint32_t count, float* __restrict__ aos)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)count * 13)
        return;
    uint32_t ent = (uint32_t)(i / 13), f = (uint32_t)(i % 13);
    aos[i] = f < 6 ? bv_in[(size_t)ent * 6 + f] : bv_out[(size_t)ent * 7 + (f - 6)];
}

// blockIdx.y = segment, blockIdx.x strides over the run; 128-bit when both ends are aligned
__global__ void __launch_bounds__(256) copy_segments_kernel(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst,
                                                           const __grid_constant__ SegTable T)
{
    const pmb_segment sg = T.seg[blockIdx.y];
    const uint32_t*   s = src + sg.src_offset;
    uint32_t*         d = dst + sg.dst_offset;
    const uint32_t    tid = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    if ((((uintptr_t)s | (uintptr_t)d) & 15u) == 0)
    {
        const uint32_t n4 = sg.count >> 2;
        for (uint32_t i = tid; i < n4; i += stride)
        {
            uint4 v = reinterpret_cast<const uint4*>(s)[i];
            v.x += sg.add, v.y += sg.add, v.z += sg.add, v.w += sg.add;
            reinterpret_cast<uint4*>(d)[i] = v;
        }
        for (uint32_t i = (n4 << 2) + tid; i < sg.count; i += stride)
            d[i] = s[i] + sg.add;
    }
    else
    {
        for (uint32_t i = tid; i < sg.count; i += stride)
            d[i] = s[i] + sg.add;
    }
}

// ------------------------------------------------------------------------------------------------
// multi-GPU list exchange without a host round trip.  A rank's "row" is u32[row_stride]:
//   row[0 .. nv)            the per-view list lengths of this shard
//   row[nv .. nv + sum)     the nv lists back to back, entity indices rebased to global numbering
// (entries beyond row_stride are dropped; the header still holds the true counts, so the reader
// sees the overflow).  Rows of all ranks are all-gathered as fixed-size blocks; unpack_rows lays
// the (rank, view) segments out per view in rank order = the ascending global list of the view.
// Every offset is recomputed from the headers on the device (at most 16 views x 8 ranks).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pack_row_kernel(const uint32_t* __restrict__ lists, size_t list_stride,
                                                      const uint32_t* __restrict__ counts, uint32_t nv, uint32_t add,
                                                      uint32_t* __restrict__ row, uint32_t row_stride,
                                                      const float* __restrict__ extents, uint32_t extents_at)
{
    const uint32_t v = blockIdx.y;
    if (blockIdx.x == 0 && v == 0 && threadIdx.x < nv)
        row[threadIdx.x] = counts[threadIdx.x];
    if (extents && blockIdx.x == 0 && v == 0 && threadIdx.x >= 32 && threadIdx.x < 38) // this frame's renderable extents ride behind the row
        row[extents_at + threadIdx.x - 32] = __float_as_uint(extents[threadIdx.x - 32]);
    uint32_t off = nv;
    for (uint32_t i = 0; i < v; ++i)
        off += counts[i];
    const uint32_t  c = counts[v];
    const uint32_t  room = off < row_stride ? row_stride - off : 0u;
    const uint32_t  m = c < room ? c : room;
    const uint32_t* src = lists + (size_t)v * list_stride;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x)
        row[off + i] = src[i] + add;
}

// meta[0 .. nv) = offset of view v's global list in dst, meta[nv .. 2 nv) = its length,
// meta[2 nv] = 1 when some row or dst overflowed (lists incomplete: grow the buffers and repeat).
// The grid is flat over the OUTPUT: every CTA rebuilds the (view, rank) segment table from the
// gathered headers in shared memory, then copies equal slices of dst whatever the segment sizes
// (one view can hold half of all entries).
constexpr uint32_t UNPACK_SEGS = PMB_MAX_VIEWS * 16; // views x ranks of one box (8) with room to spare
__global__ void __launch_bounds__(256) unpack_rows_kernel(const uint32_t* __restrict__ rows, uint32_t world, uint32_t row_stride,
                                                         uint32_t nv, uint32_t* __restrict__ dst, uint64_t dst_capacity,
                                                         uint32_t* __restrict__ meta)
{
    __shared__ uint32_t s_dst[UNPACK_SEGS + 1]; // first dst index of segment s = v * world + r (dst order), and the total
    __shared__ uint64_t s_src[UNPACK_SEGS];     // first rows[] index of the segment
    __shared__ uint32_t s_overflow;
    const uint32_t      nseg = world * nv;
    if (threadIdx.x == 0)
    {
        uint64_t run = 0;
        bool     overflow = false;
        for (uint32_t r = 0; r < world; ++r)
        {
            uint64_t in_row = nv;
            for (uint32_t v = 0; v < nv; ++v)
            {
                s_src[v * world + r] = (uint64_t)r * row_stride + in_row;
                in_row += rows[(size_t)r * row_stride + v];
            }
            overflow |= in_row > row_stride;
        }
        for (uint32_t v = 0; v < nv; ++v)
            for (uint32_t r = 0; r < world; ++r)
            {
                s_dst[v * world + r] = (uint32_t)run;
                run += rows[(size_t)r * row_stride + v];
            }
        s_dst[nseg] = (uint32_t)run;
        overflow |= run > dst_capacity || run > 0xFFFFFFFFull;
        s_overflow = overflow ? 1u : 0u;
        if (blockIdx.x == 0)
        {
            for (uint32_t v = 0; v < nv; ++v)
            {
                meta[v] = s_dst[v * world];
                meta[nv + v] = s_dst[(v + 1) * world] - s_dst[v * world];
            }
            meta[2 * nv] = s_overflow;
        }
    }
    __syncthreads();
    if (s_overflow) // every CTA reaches the same verdict: all or nothing is written
        return;
    const uint32_t total = s_dst[nseg];
    const uint32_t per = (total + gridDim.x - 1) / gridDim.x;
    const uint32_t lo = blockIdx.x * per, hi = min(total, lo + per);
    for (uint32_t i = lo + threadIdx.x; i < hi; i += blockDim.x)
    {
        // the last segment that starts at or before i (empty segments share a start: take the last)
        uint32_t a = 0, b = nseg;
        while (b - a > 1)
        {
            const uint32_t m = (a + b) >> 1;
            if (s_dst[m] <= i)
                a = m;
            else
                b = m;
        }
        dst[i] = rows[s_src[a] + (i - s_dst[a])];
    }
}

// ------------------------------------------------------------------------------------------------
// the same exchange as ONE push over peer memory (NVLink P2P stores into symmetric buffers): no
// intermediate gathered rows, no unpack pass.  publish_counts writes this rank's list lengths into
// every peer's table; after a cross-GPU barrier every rank knows all lengths, so push_row can send
// each of its (view) segments straight to its final place in every peer's per-view global lists.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) publish_counts_kernel(const uint32_t* __restrict__ row, uint32_t nv, uint32_t rank,
                                                            uint32_t world, const __grid_constant__ PeerPtrs tabs)
{
    const uint32_t q = threadIdx.x / PMB_MAX_VIEWS, v = threadIdx.x % PMB_MAX_VIEWS;
    if (q < world)
        reinterpret_cast<uint32_t*>(tabs.p[q])[rank * PMB_MAX_VIEWS + v] = v < nv ? row[v] : 0u;
    __threadfence_system();
}

__global__ void __launch_bounds__(256) push_row_kernel(const uint32_t* __restrict__ row, uint32_t row_stride, uint32_t nv,
                                                      uint32_t rank, uint32_t world, const uint32_t* __restrict__ tab,
                                                      const __grid_constant__ PeerPtrs outs, uint64_t out_capacity,
                                                      uint32_t* __restrict__ meta)
{
    __shared__ uint32_t s_start[PMB_MAX_VIEWS + 1]; // first entry of my view-v segment in my row's body
    __shared__ uint32_t s_dst[PMB_MAX_VIEWS];       // where that segment starts in the global lists
    __shared__ uint32_t s_overflow;
    if (threadIdx.x == 0)
    {
        uint64_t view_off = 0;
        uint32_t mine = 0;
        bool     overflow = false;
        for (uint32_t v = 0; v < nv; ++v)
        {
            uint64_t total = 0, before = 0;
            for (uint32_t r = 0; r < world; ++r)
            {
                const uint32_t c = tab[r * PMB_MAX_VIEWS + v];
                if (r < rank)
                    before += c;
                total += c;
            }
            s_start[v] = mine;
            s_dst[v] = (uint32_t)(view_off + before);
            if (blockIdx.x == 0)
            {
                meta[v] = (uint32_t)view_off;
                meta[nv + v] = (uint32_t)total;
            }
            mine += tab[rank * PMB_MAX_VIEWS + v];
            view_off += total;
        }
        s_start[nv] = mine;
        for (uint32_t r = 0; r < world; ++r) // a rank whose row was truncated cannot push: every rank sees it the same way
        {
            uint64_t sum = nv;
            for (uint32_t v = 0; v < nv; ++v)
                sum += tab[r * PMB_MAX_VIEWS + v];
            overflow |= sum > row_stride;
        }
        overflow |= view_off > out_capacity || view_off > 0xFFFFFFFFull;
        s_overflow = overflow ? 1u : 0u;
        if (blockIdx.x == 0)
            meta[2 * nv] = s_overflow;
    }
    __syncthreads();
    if (s_overflow)
        return;
    const uint32_t total = s_start[nv];
    const uint32_t per = (total + gridDim.x - 1) / gridDim.x;
    const uint32_t lo = blockIdx.x * per, hi = min(total, lo + per);
    for (uint32_t i = lo + threadIdx.x; i < hi; i += blockDim.x)
    {
        uint32_t v = 0;
        while (v + 1 < nv && s_start[v + 1] <= i)
            ++v;
        const uint32_t val = row[nv + i];
        const uint32_t d = s_dst[v] + (i - s_start[v]);
        for (uint32_t k = 0; k < world; ++k) // start with myself's neighbour so the ranks spread over the links
        {
            uint32_t q = rank + 1 + k;
            q = q >= world ? q - world : q;
            reinterpret_cast<uint32_t*>(outs.p[q])[d] = val;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// stage 0 and the consumers of the bone world matrices
// ------------------------------------------------------------------------------------------------
// loop A's controlled-transform branch (ecs_scene.cpp:1415-1427): local = T * R * S for every entity
// that carries e_cmp::transform and is not forced to sync with physics (:1408-1412).  The flags are
// left for clear_flags_kernel, so the streaming pass sees the same pre-update flags as the reference.
// The two products are the generic mat4 * mat4 of the hierarchy step, zeros and ones included, so
// the bits (signed zeros too) are those of translation_mat * rot_mat * scale_mat.
__global__ void __launch_bounds__(256) bake_kernel(float* __restrict__ transforms, const uint64_t* __restrict__ entities,
                                                  const uint64_t* __restrict__ state_flags, float4* __restrict__ local, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    if (!(entities[i] & CMP_TRANSFORM) || (state_flags[i] & STATE_SYNC_PHYSICS_TRANSFORM))
        return;
    float* t = transforms + (size_t)i * 10; // translation xyz, rotation xyzw, scale xyz (cmp_transform, 40 B)
    mat4r  T, S;
    T.r[0] = make_float4(1.0f, 0.0f, 0.0f, t[0]);
    T.r[1] = make_float4(0.0f, 1.0f, 0.0f, t[1]);
    T.r[2] = make_float4(0.0f, 0.0f, 1.0f, t[2]);
    T.r[3] = make_float4(0.0f, 0.0f, 0.0f, 1.0f);
    float qx = t[3], qy = t[4], qz = t[5], qw = t[6];
    if (quat_normalise(qx, qy, qz, qw)) // t.rotation.get_matrix(rot_mat) normalises t.rotation in place (:1421)
        t[3] = qx, t[4] = qy, t[5] = qz, t[6] = qw;
    const mat4r R = quat_expand(qx, qy, qz, qw);
    S.r[0] = make_float4(t[7], 0.0f, 0.0f, 0.0f);
    S.r[1] = make_float4(0.0f, t[8], 0.0f, 0.0f);
    S.r[2] = make_float4(0.0f, 0.0f, t[9], 0.0f);
    S.r[3] = make_float4(0.0f, 0.0f, 0.0f, 1.0f);
    const mat4r L = mat_mul(mat_mul(T, R), S);
    float4*     o = local + (size_t)i * 4;
    o[0] = L.r[0];
    o[1] = L.r[1];
    o[2] = L.r[2];
    o[3] = L.r[3];
}

// The host's `scene->entities[n] |= e_cmp::transform` for a whole range (what every pmtech site that writes a
// cmp_transform does, e.g. ecs_editor.cpp:2684, ecs_scene.cpp:1302) without sending the 8 B flag words over PCIe:
// the range is OR-ed on the device.  Two entities per thread (16 B accesses).
__global__ void __launch_bounds__(256) mark_dirty_kernel(uint64_t* __restrict__ entities, uint32_t first, uint32_t count)
{
    const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (i >= count)
        return;
    uint64_t* p = entities + first + i;
    if (i + 1 < count && ((first + i) & 1u) == 0)
    {
        ulonglong2 v = *reinterpret_cast<ulonglong2*>(p);
        v.x |= CMP_TRANSFORM;
        v.y |= CMP_TRANSFORM;
        *reinterpret_cast<ulonglong2*>(p) = v;
    }
    else
    {
        p[0] |= CMP_TRANSFORM;
        if (i + 1 < count)
            p[1] |= CMP_TRANSFORM;
    }
}

// root motion (ecs_scene.cpp:1286-1304): the trajectory joint's delta, rotated by the joint's initial
// rotation, moves the rig's parent: transforms[p].rotation = q, .translation += rot_mat * delta,
// entities[p] |= e_cmp::transform.  One thread per (distinct) target.
__global__ void __launch_bounds__(256) root_motion_kernel(float* __restrict__ transforms, uint64_t* __restrict__ entities,
                                                         const uint32_t* __restrict__ targets, const float* __restrict__ rotations,
                                                         const float* __restrict__ deltas, uint32_t count, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count)
        return;
    const uint32_t p = targets[i];
    if (p >= n)
        return;
    const float* q = rotations + (size_t)i * 4;
    const float* d = deltas + (size_t)i * 3;
    float        qx = q[0], qy = q[1], qz = q[2], qw = q[3]; // quat q = …; q.get_matrix(rot_mat) normalises the copy (:1291)
    quat_normalise(qx, qy, qz, qw);
    const mat4r R = quat_expand(qx, qy, qz, qw);
    float*      t = transforms + (size_t)p * 10;
    // rot_mat.transform_vector(delta): dot(row, (delta, 1))
    t[0] = add(t[0], dot4(d[0], R.r[0].x, d[1], R.r[0].y, d[2], R.r[0].z, 1.0f, R.r[0].w));
    t[1] = add(t[1], dot4(d[0], R.r[1].x, d[1], R.r[1].y, d[2], R.r[1].z, 1.0f, R.r[1].w));
    t[2] = add(t[2], dot4(d[0], R.r[2].x, d[1], R.r[2].y, d[2], R.r[2].z, 1.0f, R.r[2].w));
    t[3] = qx, t[4] = qy, t[5] = qz, t[6] = qw; // transforms[p].rotation = q (:1300)
    entities[p] |= CMP_TRANSFORM;
}

// skin palettes (ecs_scene.cpp:1917-1959): bb[i] = world_matrices[joints_offset + i] * joint_bind_matrices[i].
// One CTA per skinned mesh, one thread per (joint, matrix row): the row of the joint's world matrix is
// one coalesced float4, the bind matrices of a shared skin come from L1/L2.
__global__ void __launch_bounds__(128) skin_kernel(const float4* __restrict__ world, const SkinJob* __restrict__ jobs,
                                                  const float4* __restrict__ bind, float4* __restrict__ palettes, uint32_t n)
{
    const SkinJob  job = jobs[blockIdx.x];
    const uint32_t r = threadIdx.x & 3u;
    for (uint32_t j = threadIdx.x >> 2; j < job.num_joints; j += blockDim.x >> 2)
    {
        const uint32_t e = job.joints_offset + j;
        if (e >= n)
            continue;
        mat4r         b;
        const float4* bp = bind + (size_t)(job.bind_offset + j) * 4;
        b.r[0] = bp[0], b.r[1] = bp[1], b.r[2] = bp[2], b.r[3] = bp[3];
        palettes[(size_t)(job.palette_offset + j) * 4 + r] = mul_row(world[(size_t)e * 4 + r], b);
    }
}

// draw-call packing (ecs_scene.cpp:1962-1991): cmp_draw_call{world, v1, v2, inverse transpose} per entity.
// mat::inverse4x4 is the maths library's Laplace expansion (declared order): every
// operation below is one rounded binary32 operation in C's left-to-right order.
__device__ __forceinline__ void inverse4x4_rows(const float m[16], float4 out[4])
{
    const float s0 = sub(mul(m[0], m[5]), mul(m[4], m[1]));
    const float s1 = sub(mul(m[0], m[6]), mul(m[4], m[2]));
    const float s2 = sub(mul(m[0], m[7]), mul(m[4], m[3]));
    const float s3 = sub(mul(m[1], m[6]), mul(m[5], m[2]));
    const float s4 = sub(mul(m[1], m[7]), mul(m[5], m[3]));
    const float s5 = sub(mul(m[2], m[7]), mul(m[6], m[3]));
    const float c5 = sub(mul(m[10], m[15]), mul(m[14], m[11]));
    const float c4 = sub(mul(m[9], m[15]), mul(m[13], m[11]));
    const float c3 = sub(mul(m[9], m[14]), mul(m[13], m[10]));
    const float c2 = sub(mul(m[8], m[15]), mul(m[12], m[11]));
    const float c1 = sub(mul(m[8], m[14]), mul(m[12], m[10]));
    const float c0 = sub(mul(m[8], m[13]), mul(m[12], m[9]));
    const float det = add(sub(add(add(sub(mul(s0, c5), mul(s1, c4)), mul(s2, c3)), mul(s3, c2)), mul(s4, c1)), mul(s5, c0));
    const float inv = __fdiv_rn(1.0f, det);
#define PMB_T3P(a, x, b, y, c, z) mul(add(sub(mul(a, x), mul(b, y)), mul(c, z)), inv)  /* ( a*x - b*y + c*z) * inv */
#define PMB_T3N(a, x, b, y, c, z) mul(-add(sub(mul(a, x), mul(b, y)), mul(c, z)), inv)   /* -(a*x - b*y + c*z) * inv */
    out[0] = make_float4(PMB_T3P(m[5], c5, m[6], c4, m[7], c3), PMB_T3N(m[1], c5, m[2], c4, m[3], c3),
                         PMB_T3P(m[13], s5, m[14], s4, m[15], s3), PMB_T3N(m[9], s5, m[10], s4, m[11], s3));
    out[1] = make_float4(PMB_T3N(m[4], c5, m[6], c2, m[7], c1), PMB_T3P(m[0], c5, m[2], c2, m[3], c1),
                         PMB_T3N(m[12], s5, m[14], s2, m[15], s1), PMB_T3P(m[8], s5, m[10], s2, m[11], s1));
    out[2] = make_float4(PMB_T3P(m[4], c4, m[5], c2, m[7], c0), PMB_T3N(m[0], c4, m[1], c2, m[3], c0),
                         PMB_T3P(m[12], s4, m[13], s2, m[15], s0), PMB_T3N(m[8], s4, m[9], s2, m[11], s0));
    out[3] = make_float4(PMB_T3N(m[4], c3, m[5], c1, m[6], c0), PMB_T3P(m[0], c3, m[1], c1, m[2], c0),
                         PMB_T3N(m[12], s3, m[13], s1, m[14], s0), PMB_T3P(m[8], s3, m[9], s1, m[10], s0));
#undef PMB_T3P
#undef PMB_T3N
}

// One thread per entity for the arithmetic, but every global access goes through shared memory so that
// consecutive lanes touch consecutive 16-byte words.  The 160 B records of a tile are assembled in shared
// memory — the user's v1.zw / v2 (and the old inverse transpose of an entity the reference skips) are read
// first — and leave as ONE contiguous run of whole 128 B lines: a record written piecemeal leaves partial
// 32 B sectors behind (v1.xy is 8 bytes), which the L2 has to complete from DRAM.
constexpr uint32_t DC_TILE = 128;
// USER = false: the host never uploaded PMB_DRAW_CALLS, so v1.zw and v2 are still the zeros the mirror was
// created with and need not be read back (32 B per record, a 64 B DRAM access each).
template <bool USER>
__global__ void __launch_bounds__(DC_TILE) draw_call_kernel(const float4* __restrict__ world, const uint64_t* __restrict__ entities,
                                                           const uint32_t* __restrict__ cbuffer, float4* __restrict__ dc, uint32_t n,
                                                           float time_ms)
{
    __shared__ float4 s_rec[DC_TILE * 10]; // float4 0-3 world, 4 v1, 5 v2, 6-9 inverse transpose
    const uint32_t    base = blockIdx.x * DC_TILE;
    const uint32_t    t = threadIdx.x;
    const uint32_t    cnt = min(DC_TILE, n - base);
    float4*           tile = dc + (size_t)base * 10;
    for (uint32_t q = t; q < cnt * 4; q += DC_TILE) // lane -> consecutive float4 of the world array
        s_rec[(q >> 2) * 10 + (q & 3)] = world[(size_t)base * 4 + q];
    for (uint32_t q = t; q < cnt * 2; q += DC_TILE) // v1, v2 of every record (32 B runs)
        s_rec[(q >> 1) * 10 + 4 + (q & 1)] = USER ? tile[(q >> 1) * 10 + 4 + (q & 1)] : make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    if (t < cnt)
    {
        const uint32_t i = base + t;
        float4*        rec = s_rec + t * 10;
        const float4   w0 = rec[0], w1 = rec[1], w2 = rec[2], w3 = rec[3]; // :1972 world_matrix = world_matrices[n]
        const uint32_t h = cbuffer[i];
        const uint64_t ef = entities[i];
        const bool     live = !(h == 0xFFFFFFFFu || h == 0u) && !(ef & CMP_SUB_INSTANCE); // :1978-1982
        if (live && (ef & (CMP_SKINNED | CMP_PRE_SKINNED))) // :1985-1986 identity (the inverse below still comes from the world matrix)
        {
            rec[0] = make_float4(1.f, 0.f, 0.f, 0.f);
            rec[1] = make_float4(0.f, 1.f, 0.f, 0.f);
            rec[2] = make_float4(0.f, 0.f, 1.f, 0.f);
            rec[3] = make_float4(0.f, 0.f, 0.f, 1.f);
        }
        rec[4].x = (float)i; // :1975 v1.x = entity index
        rec[4].y = time_ms;  // :1976 v1.y = time; z, w and v2 are the user's
        if (live)
        {
            const float m[16] = {w0.x, w1.x, w2.x, w3.x, w0.y, w1.y, w2.y, w3.y, w0.z, w1.z, w2.z, w3.z, w0.w, w1.w, w2.w, w3.w}; // transposed()
            inverse4x4_rows(m, rec + 6);
        }
        else
        {
#pragma unroll
            for (int r = 0; r < 4; ++r)
                rec[6 + r] = tile[t * 10 + 6 + r]; // skipped by the reference: the stored value stands
        }
    }
    __syncthreads();
    for (uint32_t q = t; q < cnt * 10; q += DC_TILE)
        tile[q] = s_rec[q];
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
static inline uint32_t cdiv(uint64_t a, uint32_t b) { return (uint32_t)((a + b - 1) / b); }

constexpr size_t frame_smem(bool deep) { return deep ? sizeof(TileStageDeep) : sizeof(TileStage); }

template <int MODE, bool DEEP, bool TAIL>
static cudaError_t geometry_one(FrameGeometry* g)
{
    cudaError_t e = cudaFuncSetAttribute(frame_kernel<MODE, DEEP, TAIL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)frame_smem(DEEP));
    if (e != cudaSuccess)
        return e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, frame_kernel<MODE, DEEP, TAIL>, 32, frame_smem(DEEP));
    g->ctas_per_sm[MODE][DEEP][TAIL] = occ > 0 ? occ : 1;
    return e;
}

template <int MODE>
static cudaError_t geometry_mode(FrameGeometry* g)
{
    cudaError_t e = geometry_one<MODE, false, false>(g);
    if (e == cudaSuccess) e = geometry_one<MODE, false, true>(g);
    if (e == cudaSuccess) e = geometry_one<MODE, true, false>(g);
    if (e == cudaSuccess) e = geometry_one<MODE, true, true>(g);
    return e;
}

// occupancy of every instantiation of the streaming pass on `device` (the current device); kept in the scene handle
cudaError_t frame_geometry(FrameGeometry* g, int device)
{
    cudaError_t e = cudaDeviceGetAttribute(&g->sms, cudaDevAttrMultiProcessorCount, device);
    if (e == cudaSuccess) e = geometry_mode<MODE_UPDATE>(g);
    if (e == cudaSuccess) e = geometry_mode<MODE_AABB>(g);
    if (e == cudaSuccess) e = geometry_mode<MODE_SPHERE>(g);
    if (e == cudaSuccess) e = geometry_mode<MODE_FILTER>(g);
    return e;
}

void launch_extents_init(float* extents2, cudaStream_t st) { extents_init_kernel<<<1, 32, 0, st>>>(extents2); }

template <int MODE, bool DEEP, bool TAIL>
static void launch_frame_one(const SceneArrays& A, const ViewSet& vs, uint32_t nv, const FrameGeometry& g, cudaStream_t st)
{
    uint32_t tiles = cdiv(A.n, TILE);
    uint32_t grid = (uint32_t)(g.sms * g.ctas_per_sm[MODE][DEEP][TAIL]);
    if (grid > tiles)
        grid = tiles;
    if (TAIL)
    {
        // grid barriers inside: a cooperative launch starts only when the whole grid is resident at once
        void* args[] = {(void*)&A, (void*)&vs, (void*)&nv, (void*)&tiles};
        cudaLaunchCooperativeKernel((const void*)frame_kernel<MODE, DEEP, TAIL>, dim3(grid), dim3(32), args, frame_smem(DEEP), st);
    }
    else
        frame_kernel<MODE, DEEP, TAIL><<<grid, 32, frame_smem(DEEP), st>>>(A, vs, nv, tiles);
}

template <int MODE>
static void launch_frame_mode(const SceneArrays& A, const ViewSet& vs, uint32_t nv, bool deep, bool tail, const FrameGeometry& g,
                              cudaStream_t st)
{
    if (deep)
    {
        if (tail)
            launch_frame_one<MODE, true, true>(A, vs, nv, g, st);
        else
            launch_frame_one<MODE, true, false>(A, vs, nv, g, st);
    }
    else if (tail)
        launch_frame_one<MODE, false, true>(A, vs, nv, g, st);
    else
        launch_frame_one<MODE, false, false>(A, vs, nv, g, st);
}

void launch_frame(const SceneArrays& A, const ViewSet& vs, uint32_t nv, int mode, bool deep, bool tail, const FrameGeometry& g,
                  cudaStream_t st)
{
    if (A.n == 0)
        return;
    switch (mode)
    {
    case MODE_UPDATE: launch_frame_mode<MODE_UPDATE>(A, vs, nv, deep, tail, g, st); break;
    case MODE_AABB: launch_frame_mode<MODE_AABB>(A, vs, nv, deep, tail, g, st); break;
    case MODE_SPHERE: launch_frame_mode<MODE_SPHERE>(A, vs, nv, deep, tail, g, st); break;
    default: launch_frame_mode<MODE_FILTER>(A, vs, nv, deep, tail, g, st); break;
    }
}

void launch_expand(const SceneArrays& A, const uint32_t* children, uint32_t num_children, cudaStream_t st)
{
    if (num_children == 0)
        return;
    expand_kernel<<<cdiv(num_children, 256), 256, 0, st>>>(A, children, num_children);
}

void launch_expand_lists(const SceneArrays& A, cudaStream_t st)
{
    if (A.num_cross == 0)
        return;
    uint32_t grid = cdiv(A.num_cross, 8 * XL_CHILDREN); // five children per warp, eight warps per CTA
    if (grid > 148 * 32)
        grid = 148 * 32;
    expand_lists_kernel<<<grid, 256, 0, st>>>(A);
}

void launch_ancestor_counts(const uint32_t* parents, const uint32_t* cross, uint32_t num_cross, uint32_t* counts, cudaStream_t st)
{
    ancestor_count_kernel<<<cdiv((uint64_t)num_cross + 1, 256), 256, 0, st>>>(parents, cross, num_cross, counts);
}

cudaError_t scan_ancestor_offsets(uint32_t* counts_then_offsets, uint32_t num_cross, void** temp, size_t* temp_bytes, cudaStream_t st)
{
    size_t      need = 0;
    cudaError_t e = cub::DeviceScan::ExclusiveSum(nullptr, need, counts_then_offsets, counts_then_offsets, (int)(num_cross + 1), st);
    if (e != cudaSuccess)
        return e;
    if (need > *temp_bytes)
    {
        if (*temp)
            cudaFree(*temp);
        *temp = nullptr, *temp_bytes = 0;
        e = cudaMalloc(temp, need);
        if (e != cudaSuccess)
            return e;
        *temp_bytes = need;
    }
    return cub::DeviceScan::ExclusiveSum(*temp, need, counts_then_offsets, counts_then_offsets, (int)(num_cross + 1), st);
}

void launch_ancestor_fill(const uint32_t* parents, const uint32_t* cross, uint32_t num_cross, const uint32_t* offsets, uint32_t* anc,
                          cudaStream_t st)
{
    if (num_cross)
        ancestor_fill_kernel<<<cdiv(num_cross, 256), 256, 0, st>>>(parents, cross, num_cross, offsets, anc);
}

void launch_clear_flags(const SceneArrays& A, cudaStream_t st)
{
    if (A.n == 0)
        return;
    uint32_t grid = cdiv(A.n, 256);
    if (grid > 148 * 8)
        grid = 148 * 8;
    clear_flags_kernel<<<grid, 256, 0, st>>>(A.entities, A.state_flags, A.n, A.dirty);
}

void launch_cull(const SceneArrays& A, const ViewSet& vs, uint32_t nv, int mode, cudaStream_t st)
{
    if (A.n == 0)
        return;
    uint32_t grid = cdiv(A.n, 256);
    switch (mode)
    {
    case MODE_AABB: cull_kernel<MODE_AABB, false><<<grid, 256, 0, st>>>(A, vs, nullptr, nv); break;
    case MODE_SPHERE: cull_kernel<MODE_SPHERE, false><<<grid, 256, 0, st>>>(A, vs, nullptr, nv); break;
    default: cull_kernel<MODE_FILTER, false><<<grid, 256, 0, st>>>(A, vs, nullptr, nv); break;
    }
}

void launch_shadow_fit(const float* extents, const float* directions, const float* constraints, uint32_t num_lights,
                       pmb_camera* cams, ViewSet* vs, cudaStream_t st)
{
    shadow_fit_kernel<<<1, PMB_MAX_VIEWS, 0, st>>>(extents, directions, constraints, num_lights, cams, vs);
}

void launch_spot_fit(const float4* world, const uint32_t* light_entities, uint32_t num_lights, uint32_t n, const float proj[16],
                     pmb_camera* cams, ViewSet* vs, cudaStream_t st)
{
    pmcam::m4 p;
    memcpy(p.m, proj, sizeof(p.m));
    spot_fit_kernel<<<1, PMB_MAX_VIEWS, 0, st>>>(world, light_entities, num_lights, n, p, cams, vs);
}

void launch_cull_device_views(const SceneArrays& A, const ViewSet* vs, uint32_t nv, int mode, cudaStream_t st)
{
    if (A.n == 0)
        return;
    static ViewSet dummy; // the by-value parameter of the shared kernel; never read in this instantiation
    uint32_t       grid = cdiv(A.n, 256);
    switch (mode)
    {
    case MODE_AABB: cull_kernel<MODE_AABB, true><<<grid, 256, 0, st>>>(A, dummy, vs, nv); break;
    case MODE_SPHERE: cull_kernel<MODE_SPHERE, true><<<grid, 256, 0, st>>>(A, dummy, vs, nv); break;
    default: cull_kernel<MODE_FILTER, true><<<grid, 256, 0, st>>>(A, dummy, vs, nv); break;
    }
}

constexpr uint32_t SMALL_WORDS = 8192; // scenes up to 256 Ki entities compact in one-warp chunks
// chunk-count slots per view: 1024-word chunks, or up to 64 one-warp chunks for a small scene
uint32_t compact_num_chunks(uint32_t n) { return std::max(cdiv(cdiv(n, 32), COMPACT_WORDS), n ? SMALL_WORDS / 128u : 0u); }
uint32_t compact_num_launches(uint32_t n) { return n == 0 ? 0u : 3u; }

void launch_compact(const uint32_t* bitmap, uint32_t bitmap_stride, uint32_t n, uint32_t nv, uint32_t view_base,
                    uint32_t* chunk_counts, uint32_t* lists, size_t list_stride, uint32_t* counts, cudaStream_t st)
{
    if (nv == 0)
        return;
    uint32_t chunks = compact_num_chunks(n);
    if (chunks == 0)
    {
        cudaMemsetAsync(counts + view_base, 0, sizeof(uint32_t) * nv, st);
        return;
    }
    // only the words the cull passes wrote are read: one per warp that holds an entity below n
    uint32_t num_words = cdiv(n, 32);
    // Small scenes (pmtech's own samples: 64 Ki entities and less) would leave the machine to one or two CTAs per
    // view; they are cut into one-warp chunks of 128 words (4 Ki entities) instead of 1024.
    const uint32_t threads = num_words <= SMALL_WORDS ? 32u : 256u;
    chunks = cdiv(num_words, threads * 4);
    dim3 grid(chunks, nv);
    compact_count_kernel<<<grid, threads, 0, st>>>(bitmap, bitmap_stride, num_words, chunks, chunk_counts);
    compact_scan_kernel<<<nv, 1024, 0, st>>>(chunk_counts, chunks, counts, view_base);
    compact_scatter_kernel<<<grid, threads, 0, st>>>(bitmap, bitmap_stride, num_words, chunks, chunk_counts,
                                                     lists + (size_t)view_base * list_stride, list_stride);
}

void launch_depth(const uint32_t* parents, uint32_t n, uint32_t* depth, uint32_t* max_depth, cudaStream_t st)
{
    cudaMemsetAsync(max_depth, 0, sizeof(uint32_t), st);
    if (n)
        depth_kernel<<<cdiv(n, 256), 256, 0, st>>>(parents, n, depth, max_depth);
}

void launch_merge_parents(const uint32_t* staged, uint32_t* parents, uint32_t count, uint32_t* changed, cudaStream_t st)
{
    if (count)
        merge_parents_kernel<<<cdiv(count, 256), 256, 0, st>>>(staged, parents, count, changed);
}

void launch_cross(const uint32_t* parents, uint32_t n, bool direct_only, uint32_t* cross, uint32_t* num_cross, cudaStream_t st)
{
    cudaMemsetAsync(num_cross, 0, sizeof(uint32_t), st);
    if (n)
        cross_kernel<<<cdiv(n, 256), 256, 0, st>>>(parents, n, direct_only ? 1u : 0u, cross, num_cross);
}

void launch_publish_mask(const uint32_t* parents, uint32_t n, uint32_t padded, uint64_t* publish, cudaStream_t st)
{
    static_assert(TILE == 64, "one 64-bit publish word per tile");
    cudaMemsetAsync(publish, 0, (size_t)(padded / TILE) * 8, st);
    if (n)
        publish_mask_kernel<<<cdiv(n, 256), 256, 0, st>>>(parents, n, reinterpret_cast<unsigned long long*>(publish));
}

void launch_plan(const uint32_t* parents, const uint32_t* depth, uint32_t n, const uint64_t* publish, TilePlan* plans, cudaStream_t st)
{
    const uint32_t tiles = cdiv(n, TILE);
    if (tiles)
        plan_kernel<<<cdiv(tiles, 64), 64, 0, st>>>(parents, depth, n, publish, plans, tiles);
}

void launch_bv_split(const float* aos, uint32_t count, float* bv_in, float* bv_out, cudaStream_t st)
{
    if (count)
        bv_split_kernel<<<cdiv((uint64_t)count * 13, 256), 256, 0, st>>>(aos, count, bv_in, bv_out);
}

void launch_bv_merge(const float* bv_in, const float* bv_out, uint32_t count, float* aos, cudaStream_t st)
{
    if (count)
        bv_merge_kernel<<<cdiv((uint64_t)count * 13, 256), 256, 0, st>>>(bv_in, bv_out, count, aos);
}

void launch_copy_segments(const uint32_t* src, uint32_t* dst, const SegTable& table, uint32_t num_segments,
                          uint32_t max_count, cudaStream_t st)
{
    if (num_segments == 0 || max_count == 0)
        return;
    uint32_t gx = cdiv(cdiv(max_count, 4), 256);
    if (gx > 148 * 4)
        gx = 148 * 4;
    copy_segments_kernel<<<dim3(gx, num_segments), 256, 0, st>>>(src, dst, table);
}

void launch_bake(float* transforms, const uint64_t* entities, const uint64_t* state_flags, float4* local, uint32_t n,
                 cudaStream_t st)
{
    if (n)
        bake_kernel<<<cdiv(n, 256), 256, 0, st>>>(transforms, entities, state_flags, local, n);
}

void launch_mark_dirty(uint64_t* entities, uint32_t first, uint32_t count, cudaStream_t st)
{
    if (count)
        mark_dirty_kernel<<<cdiv(cdiv(count, 2), 256), 256, 0, st>>>(entities, first, count);
}

void launch_root_motion(float* transforms, uint64_t* entities, const uint32_t* targets, const float* rotations,
                        const float* deltas, uint32_t count, uint32_t n, cudaStream_t st)
{
    if (count)
        root_motion_kernel<<<cdiv(count, 256), 256, 0, st>>>(transforms, entities, targets, rotations, deltas, count, n);
}

void launch_skin(const float4* world, const SkinJob* jobs, uint32_t num_jobs, const float4* bind, float4* palettes, uint32_t n,
                 cudaStream_t st)
{
    if (num_jobs)
        skin_kernel<<<num_jobs, 128, 0, st>>>(world, jobs, bind, palettes, n);
}

void launch_publish_counts(const uint32_t* row, uint32_t nv, uint32_t rank, uint32_t world, const PeerPtrs& tabs, cudaStream_t st)
{
    publish_counts_kernel<<<1, 256, 0, st>>>(row, nv, rank, world, tabs);
}

void launch_push_row(const uint32_t* row, uint32_t row_stride, uint32_t nv, uint32_t rank, uint32_t world, const uint32_t* tab,
                     const PeerPtrs& outs, uint64_t out_capacity, uint32_t* meta, cudaStream_t st)
{
    uint32_t gx = cdiv(row_stride, 256 * 16);
    gx = gx < 1 ? 1 : (gx > 148 * 2 ? 148 * 2 : gx);
    push_row_kernel<<<gx, 256, 0, st>>>(row, row_stride, nv, rank, world, tab, outs, out_capacity, meta);
}

void launch_draw_calls(const float4* world, const uint64_t* entities, const uint32_t* cbuffer, float4* dc, uint32_t n, float time_ms,
                       bool user_data, cudaStream_t st)
{
    if (!n)
        return;
    if (user_data)
        draw_call_kernel<true><<<cdiv(n, DC_TILE), DC_TILE, 0, st>>>(world, entities, cbuffer, dc, n, time_ms);
    else
        draw_call_kernel<false><<<cdiv(n, DC_TILE), DC_TILE, 0, st>>>(world, entities, cbuffer, dc, n, time_ms);
}

void launch_pack_row(const uint32_t* lists, size_t list_stride, const uint32_t* counts, uint32_t nv, uint32_t add,
                     uint32_t* row, uint32_t row_stride, cudaStream_t st, const float* extents, uint32_t extents_at)
{
    if (nv == 0)
        return;
    uint32_t gx = cdiv(row_stride, 256 * 4 * (nv ? nv : 1));
    gx = gx < 1 ? 1 : (gx > 148 * 2 ? 148 * 2 : gx);
    pack_row_kernel<<<dim3(gx, nv), 256, 0, st>>>(lists, list_stride, counts, nv, add, row, row_stride, extents, extents ? extents_at : row_stride);
}

void launch_unpack_rows(const uint32_t* rows, uint32_t world, uint32_t row_stride, uint32_t nv, uint32_t* dst,
                        uint64_t dst_capacity, uint32_t* meta, cudaStream_t st)
{
    if (nv == 0 || world == 0)
        return;
    if (world * nv > UNPACK_SEGS)
        return; // the C ABI rejects this before it gets here
    uint32_t gx = cdiv((uint64_t)row_stride * world, 256 * 16);
    gx = gx < 1 ? 1 : (gx > 148 * 4 ? 148 * 4 : gx);
    unpack_rows_kernel<<<gx, 256, 0, st>>>(rows, world, row_stride, nv, dst, dst_capacity, meta);
}

} // namespace pm
