// pm_kernels.cu — sm_100a kernels of the scene-update path.
//
//   frame_kernel   streaming pass, one thread per entity, 256-entity CTA tiles.  The 64 B
//                  matrices and the 24/28 B bound records move between HBM and shared memory as
//                  fully coalesced 128-bit accesses; each thread then reads its own record from
//                  shared memory (80 B padded matrix stride = conflict-free LDS.128).  Ancestor
//                  matrices that fall inside the CTA's tile are taken from the same staging
//                  buffer.  World matrix, bounds, pos_extent and the per-view visibility bits are
//                  produced in registers in one go (loops A, B of update_scene + the plane tests
//                  of frustum_cull_*_scalar), so pos_extent is never re-read.
//   band_kernel    the same per-entity work for entities deeper than WALK levels, run as depth
//                  band sweeps (band b reads the finished world matrix of its ancestor at depth
//                  b*WALK and re-multiplies at most WALK locals below it, in the reference's
//                  left-to-right product order).
//   expand_kernel  loop C (parent bound expansion) with exact float min/max atomics.
//   cull_kernel    stand-alone per-view plane tests over pos_extent (render_scene_view calls).
//   compact_*      u16 visibility masks -> one ascending u32 index list per view
//                  (warp ballot/popc + block scan), i.e. the sb_push order of the reference.
//
// Reference: core/put/source/ecs/ecs_scene.cpp:1405-1556, core/put/source/ecs/ecs_cull.cpp:21-105.
#include "pm_kernels.h"
#include "pm_math.cuh"

#include <float.h>

namespace pm {

// ------------------------------------------------------------------------------------------------
// host: per-plane constants.  Compiled with -ffp-contract=off; one C operation per IEEE operation.
// ------------------------------------------------------------------------------------------------
static float host_dot3(const float* a, const float* b)
{
    volatile float r = 0.0f; // volatile: keep every intermediate a rounded binary32
    for (int i = 0; i < 3; ++i)
    {
        volatile float t = a[i] * b[i];
        r = r + t;
    }
    return r;
}

void make_view_set(ViewSet* vs, const pmb_frustum* views, uint32_t num_views)
{
    for (uint32_t v = 0; v < PMB_MAX_VIEWS; ++v)
    {
        for (int p = 0; p < 6; ++p)
        {
            PlaneK& k = vs->planes[v * 6 + p];
            if (v >= num_views)
            {
                k = PlaneK{0.f, 0.f, 0.f, 0.f, 0u, 0u, 0u, 0.f};
                continue;
            }
            const float* n = views[v].n[p];
            const float* x0 = views[v].p[p];
            // maths::plane_distance(x0, xN) = dot(xN, x0) * -1 (ecs_cull.cpp:37); the aabb test
            // compares against -pd (ecs_cull.cpp:40)
            volatile float pd = host_dot3(n, x0) * -1.0f;
            k.nx = n[0];
            k.ny = n[1];
            k.nz = n[2];
            k.pd = pd;
            k.neg_pd = -pd;
            // sign_flip = sgn(n) * -1 (ecs_cull.cpp:36): n < 0 -> +extent, n > 0 -> -extent;
            // n == 0 multiplies either corner by zero, so it shares the n > 0 select.
            k.mx = n[0] < 0.0f ? 0xFFFFFFFFu : 0u;
            k.my = n[1] < 0.0f ? 0xFFFFFFFFu : 0u;
            k.mz = n[2] < 0.0f ? 0xFFFFFFFFu : 0u;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// per-entity work shared by the streaming and band kernels
// ------------------------------------------------------------------------------------------------
struct EntityOut
{
    float    tmin[3], tmax[3];
    float    radius;
    float    pos[3], ext[3];
    uint32_t mask;
    bool     bone;
    bool     fold_extents; // contributes to scene->renderable_extents
};

__device__ __forceinline__ float bitsel(uint32_t m, float a_if_set, float b_if_clear)
{
    return __uint_as_float((__float_as_uint(a_if_set) & m) | (__float_as_uint(b_if_clear) & ~m));
}

// loop B of update_scene (ecs_scene.cpp:1486-1526) for one entity
__device__ __forceinline__ void bounds_from_world(const float4 w0, const float4 w1, const float4 w2,
                                                  const float bmin[3], const float bmax[3], EntityOut& o)
{
    // :1488-1489  min, size; corner = min + size * {0,1}
    float lo[3], hi[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
        float size = sub(bmax[i], bmin[i]);
        lo[i] = add(bmin[i], mul(size, 0.0f));
        hi[i] = add(bmin[i], mul(size, 1.0f));
    }
    const float4 rows[3] = {w0, w1, w2};
    // :1474-1480 corner order: 000 100 010 001 110 011 101 111 (x,y,z selectors)
    const int sx[8] = {0, 1, 0, 0, 1, 0, 1, 1};
    const int sy[8] = {0, 0, 1, 0, 1, 1, 0, 1};
    const int sz[8] = {0, 0, 0, 1, 0, 1, 1, 1};
#pragma unroll
    for (int r = 0; r < 3; ++r)
    {
        // transform_vector: ((0 + x*m0) + y*m1) + z*m2) + 1*m3 — the 8 corners share partial sums
        const float4 m = rows[r];
        float px[2] = {add(0.0f, mul(lo[0], m.x)), add(0.0f, mul(hi[0], m.x))};
        float py[2] = {mul(lo[1], m.y), mul(hi[1], m.y)};
        float pz[2] = {mul(lo[2], m.z), mul(hi[2], m.z)};
        float pxy[2][2];
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 2; ++b)
                pxy[a][b] = add(px[a], py[b]);
        float tmx = -FLT_MAX, tmn = FLT_MAX; // :1500-1501
#pragma unroll
        for (int c = 0; c < 8; ++c)
        {
            float p = add(add(pxy[sx[c]][sy[c]], pz[sz[c]]), m.w);
            tmx = max_union(tmx, p); // :1507
            tmn = min_union(tmn, p); // :1508
        }
        o.tmin[r] = tmn;
        o.tmax[r] = tmx;
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

// frustum_cull_aabb_scalar / frustum_cull_sphere_scalar (ecs_cull.cpp:21-82) against every view of
// the batch; bit v of the result = entity is in view v's list
template <int MODE>
__device__ __forceinline__ uint32_t cull_views(const ViewSet& VS, uint32_t nv, const float pos[3], const float ext[3],
                                               float radius)
{
    uint32_t mask = 0;
    if (MODE == MODE_FILTER)
        return nv >= 32 ? 0xFFFFFFFFu : ((1u << nv) - 1u);
    float lo[3], hi[3];
    if (MODE == MODE_AABB)
    {
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
            lo[i] = sub(pos[i], ext[i]); // pos + extent * -1
            hi[i] = add(pos[i], ext[i]); // pos + extent * +1
        }
    }
#pragma unroll
    for (int v = 0; v < (int)PMB_MAX_VIEWS; ++v)
    {
        if (v < (int)nv)
        {
            bool outside = false;
#pragma unroll
            for (int p = 0; p < 6; ++p)
            {
                const PlaneK& k = VS.planes[v * 6 + p];
                if (MODE == MODE_AABB)
                {
                    float cx = bitsel(k.mx, hi[0], lo[0]);
                    float cy = bitsel(k.my, hi[1], lo[1]);
                    float cz = bitsel(k.mz, hi[2], lo[2]);
                    float d2 = add(add(mul(cx, k.nx), mul(cy, k.ny)), mul(cz, k.nz));
                    outside |= d2 > k.neg_pd; // :40
                }
                else
                {
                    float pp = add(add(mul(pos[0], k.nx), mul(pos[1], k.ny)), mul(pos[2], k.nz));
                    float d = add(pp, k.pd); // point_plane_distance, :68
                    outside |= d > radius;   // :70
                }
            }
            mask |= outside ? 0u : (1u << v);
        }
    }
    return mask;
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

constexpr int MAT_STRIDE = 5; // float4 per staged matrix (80 B: conflict-free LDS.128 per quarter warp)

// ------------------------------------------------------------------------------------------------
// frame_kernel — the streaming pass
// ------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(TILE) frame_kernel(const SceneArrays A, const __grid_constant__ ViewSet VS, uint32_t nv)
{
    __shared__ float4   s_mat[TILE * MAT_STRIDE]; // local in, world out
    __shared__ float4   s_bvin4[TILE * 6 / 4];
    __shared__ float4   s_bvout4[TILE * 7 / 4];
    __shared__ uint32_t s_par[TILE];
    __shared__ uint64_t s_ent[TILE];
    __shared__ uint64_t s_sf[TILE];
    __shared__ uint8_t  s_wr[TILE];
    __shared__ uint32_t s_extkey[6];

    float* s_bvin = reinterpret_cast<float*>(s_bvin4);
    float* s_bvout = reinterpret_cast<float*>(s_bvout4);

    const uint32_t tid = threadIdx.x;
    const uint32_t base = blockIdx.x * TILE;
    const uint32_t e = base + tid;
    const bool     active = e < A.n;

    // ---- stage the tile (mirrors are padded to a multiple of PAD entities: full-tile loads are in bounds)
    {
        const float4* g = A.local + (size_t)base * 4;
#pragma unroll
        for (int j = 0; j < 4; ++j)
        {
            uint32_t q = tid + TILE * j;
            s_mat[(q >> 2) * MAT_STRIDE + (q & 3)] = __ldg(g + q);
        }
        const float4* gb = reinterpret_cast<const float4*>(A.bv_in + (size_t)base * 6);
        s_bvin4[tid] = __ldg(gb + tid);
        if (tid < TILE * 6 / 4 - TILE)
            s_bvin4[tid + TILE] = __ldg(gb + tid + TILE);
    }
    const uint32_t parent = A.parents[e];
    const uint64_t ef = A.entities[e];
    const uint64_t sf = A.state_flags[e];
    s_par[tid] = parent;
    s_ent[tid] = ef;
    s_sf[tid] = sf;
    if (tid < 3)
    {
        s_extkey[tid] = 0xFFFFFFFFu; // min keys
        s_extkey[3 + tid] = 0u;      // max keys
    }
    __syncthreads();

    // ---- loop A: world matrix
    mat4r W;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        W.r[k] = s_mat[tid * MAT_STRIDE + k];

    const bool skipped = active && world_is_skipped(ef, sf, A.has_rb, e);
    bool       deferred = false;

    if (active && ((ef & CMP_TRANSFORM) || (sf & STATE_SYNC_PHYSICS_TRANSFORM)))
        *A.dirty = 1u;

    if (skipped)
    {
        W = load_mat_global(A.world, e); // :1445-1446 `continue`: last frame's world matrix stands
    }
    else if (active && parent != e)
    {
        // walk to the root (or to an ancestor whose world matrix is not rewritten this frame),
        // then multiply back down so the product order is ((root * l1) * l2) ... * local[e]
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
            if (pp == cur)
                break;
            if (len == (int)WALK)
            {
                deferred = true; // deeper than WALK: a band sweep finishes this entity
                break;
            }
            cur = pp;
        }
        if (!deferred)
        {
            mat4r    acc;
            uint32_t top = chain[len - 1];
            if (top_world)
                acc = load_mat_global(A.world, top);
            else if (top - base < TILE)
            {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    acc.r[k] = s_mat[(top - base) * MAT_STRIDE + k];
            }
            else
                acc = load_mat_global(A.local, top);
            for (int i = len - 2; i >= 0; --i)
            {
                uint32_t c = chain[i];
                mat4r    l;
                if (c - base < TILE)
                {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        l.r[k] = s_mat[(c - base) * MAT_STRIDE + k];
                }
                else
                    l = load_mat_global(A.local, c);
                acc = mat_mul(acc, l);
            }
            W = mat_mul(acc, W); // :1470
        }
    }
    __syncthreads(); // every ancestor read of s_mat is done; the buffer becomes the world tile

#pragma unroll
    for (int k = 0; k < 4; ++k)
        s_mat[tid * MAT_STRIDE + k] = W.r[k];
    s_wr[tid] = (active && !skipped && !deferred) ? 1 : 0;

    // ---- loop B: bounds, pos_extent, scene extents; then the plane tests
    EntityOut o = {};
    o.bone = (ef & CMP_BONE) != 0;
    float old_radius = 0.0f;
    if (active && !deferred)
    {
        if (o.bone)
        {
            // :1494-1498 tmin = tmax = translation; radius and pos_extent keep their values
            o.tmin[0] = o.tmax[0] = W.r[0].w;
            o.tmin[1] = o.tmax[1] = W.r[1].w;
            o.tmin[2] = o.tmax[2] = W.r[2].w;
            old_radius = A.bv_out[(size_t)e * 7 + 6];
        }
        else
        {
            float bmin[3] = {s_bvin[tid * 6 + 0], s_bvin[tid * 6 + 1], s_bvin[tid * 6 + 2]};
            float bmax[3] = {s_bvin[tid * 6 + 3], s_bvin[tid * 6 + 4], s_bvin[tid * 6 + 5]};
            bounds_from_world(W.r[0], W.r[1], W.r[2], bmin, bmax, o);
            o.fold_extents = (ef & CMP_GEOMETRY) != 0; // :1520
            float4* pe = A.pos_extent + (size_t)e * 2;
            pe[0] = make_float4(o.pos[0], o.pos[1], o.pos[2], 0.0f); // pos.w is never written by the reference; the arrays start zeroed
            pe[1] = make_float4(o.ext[0], o.ext[1], o.ext[2], o.radius);
        }
        if (MODE != MODE_UPDATE)
        {
            if (filter_pass(ef, sf))
            {
                if (o.bone)
                {
                    // a renderable bone is culled against the pos_extent it kept
                    float4 p0 = A.pos_extent[(size_t)e * 2], p1 = A.pos_extent[(size_t)e * 2 + 1];
                    o.pos[0] = p0.x, o.pos[1] = p0.y, o.pos[2] = p0.z;
                    o.ext[0] = p1.x, o.ext[1] = p1.y, o.ext[2] = p1.z;
                    o.radius = p1.w;
                }
                o.mask = cull_views<MODE>(VS, nv, o.pos, o.ext, o.radius);
            }
            A.mask[e] = (uint16_t)o.mask;
        }
    }
    // stage the 28 B records so the tile leaves as full 128-bit stores
    if (active && !deferred)
    {
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
            s_bvout[tid * 7 + i] = o.tmin[i];
            s_bvout[tid * 7 + 3 + i] = o.tmax[i];
        }
        s_bvout[tid * 7 + 6] = o.bone ? old_radius : o.radius;
    }
    else if (active)
    {
        // deferred: keep whatever is stored (the band sweep writes the real values later)
#pragma unroll
        for (int i = 0; i < 7; ++i)
            s_bvout[tid * 7 + i] = A.bv_out[(size_t)e * 7 + i];
    }

    // scene->renderable_extents (:1524-1525): exact min/max, so warp REDUX + atomics keep the bits
    {
        uint32_t kmin[3], kmax[3];
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
            bool take_min = o.fold_extents && (o.tmin[i] == o.tmin[i]);
            bool take_max = o.fold_extents && (o.tmax[i] == o.tmax[i]);
            kmin[i] = take_min ? float_key(o.tmin[i]) : 0xFFFFFFFFu;
            kmax[i] = take_max ? float_key(o.tmax[i]) : 0u;
        }
        if (__any_sync(0xFFFFFFFFu, o.fold_extents))
        {
#pragma unroll
            for (int i = 0; i < 3; ++i)
            {
                uint32_t a = __reduce_min_sync(0xFFFFFFFFu, kmin[i]);
                uint32_t b = __reduce_max_sync(0xFFFFFFFFu, kmax[i]);
                if ((tid & 31) == 0)
                {
                    atomicMin(&s_extkey[i], a);
                    atomicMax(&s_extkey[3 + i], b);
                }
            }
        }
    }
    __syncthreads();

    // ---- coalesced tile stores
    {
        float4* g = A.world + (size_t)base * 4;
#pragma unroll
        for (int j = 0; j < 4; ++j)
        {
            uint32_t q = tid + TILE * j;
            uint32_t m = q >> 2;
            if (s_wr[m])
                g[q] = s_mat[m * MAT_STRIDE + (q & 3)];
        }
        // 28 B records: whole float4 chunks that lie below entity n; the ragged last chunk by scalars
        uint32_t cnt = min(TILE, A.n - base);
        uint32_t nfl = cnt * 7;
        float4*  gb = reinterpret_cast<float4*>(A.bv_out + (size_t)base * 7);
#pragma unroll
        for (int j = 0; j < 2; ++j)
        {
            uint32_t q = tid + TILE * j;
            if (q * 4 + 4 <= nfl)
                gb[q] = s_bvout4[q];
            else if (q * 4 < nfl)
            {
                for (uint32_t t = q * 4; t < nfl; ++t)
                    A.bv_out[(size_t)base * 7 + t] = s_bvout[t];
            }
        }
    }
    if (tid < 6)
    {
        uint32_t k = s_extkey[tid];
        if (tid < 3)
        {
            if (k != 0xFFFFFFFFu)
            {
                float v = key_float(k);
                if (v < A.extents[tid]) // cheap pre-check; the atomic decides
                    atomic_min_float(&A.extents[tid], v);
            }
        }
        else if (k != 0u)
        {
            float v = key_float(k);
            if (v > A.extents[tid])
                atomic_max_float(&A.extents[tid], v);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// band_kernel — entities deeper than WALK levels, one depth band per launch (gathered)
// ------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256) band_kernel(const SceneArrays A, const __grid_constant__ ViewSet VS, uint32_t nv,
                                                  const uint32_t* __restrict__ band_list, uint32_t band_count,
                                                  const uint32_t* __restrict__ depth, uint32_t band_index)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= band_count)
        return;
    const uint32_t e = band_list[i];
    if (e >= A.n)
        return;
    const uint64_t ef = A.entities[e];
    const uint64_t sf = A.state_flags[e];
    const bool     skipped = world_is_skipped(ef, sf, A.has_rb, e);
    mat4r          W;
    if (skipped)
        W = load_mat_global(A.world, e);
    else
    {
        // ancestors at depth d-1 ... band_index*WALK; the last one already holds its final world matrix
        const int steps = (int)(depth[e] - band_index * WALK);
        uint32_t  chain[WALK];
        int       len = 0;
        uint32_t  cur = A.parents[e];
        for (;;)
        {
            chain[len++] = cur;
            if (len == steps)
                break;
            uint64_t ce = A.entities[cur];
            if ((ce & CMP_PHYSICS) && world_is_skipped(ce, A.state_flags[cur], A.has_rb, cur))
                break;
            cur = A.parents[cur];
        }
        mat4r acc = load_mat_global(A.world, chain[len - 1]);
        for (int k = len - 2; k >= 0; --k)
            acc = mat_mul(acc, load_mat_global(A.local, chain[k]));
        W = mat_mul(acc, load_mat_global(A.local, e));
        float4* wp = A.world + (size_t)e * 4;
        wp[0] = W.r[0];
        wp[1] = W.r[1];
        wp[2] = W.r[2];
        wp[3] = W.r[3];
    }
    EntityOut o = {};
    o.bone = (ef & CMP_BONE) != 0;
    float* bo = A.bv_out + (size_t)e * 7;
    if (o.bone)
    {
        bo[0] = bo[3] = W.r[0].w;
        bo[1] = bo[4] = W.r[1].w;
        bo[2] = bo[5] = W.r[2].w;
    }
    else
    {
        const float* bi = A.bv_in + (size_t)e * 6;
        float        bmin[3] = {bi[0], bi[1], bi[2]};
        float        bmax[3] = {bi[3], bi[4], bi[5]};
        bounds_from_world(W.r[0], W.r[1], W.r[2], bmin, bmax, o);
#pragma unroll
        for (int k = 0; k < 3; ++k)
        {
            bo[k] = o.tmin[k];
            bo[3 + k] = o.tmax[k];
        }
        bo[6] = o.radius;
        float4* pe = A.pos_extent + (size_t)e * 2;
        pe[0] = make_float4(o.pos[0], o.pos[1], o.pos[2], 0.0f);
        pe[1] = make_float4(o.ext[0], o.ext[1], o.ext[2], o.radius);
        if (ef & CMP_GEOMETRY)
        {
#pragma unroll
            for (int k = 0; k < 3; ++k)
            {
                if (o.tmin[k] < A.extents[k])
                    atomic_min_float(&A.extents[k], o.tmin[k]);
                if (o.tmax[k] > A.extents[3 + k])
                    atomic_max_float(&A.extents[3 + k], o.tmax[k]);
            }
        }
    }
    if (MODE != MODE_UPDATE)
    {
        if (filter_pass(ef, sf))
        {
            if (o.bone)
            {
                float4 p0 = A.pos_extent[(size_t)e * 2], p1 = A.pos_extent[(size_t)e * 2 + 1];
                o.pos[0] = p0.x, o.pos[1] = p0.y, o.pos[2] = p0.z;
                o.ext[0] = p1.x, o.ext[1] = p1.y, o.ext[2] = p1.z;
                o.radius = p1.w;
            }
            o.mask = cull_views<MODE>(VS, nv, o.pos, o.ext, o.radius);
        }
        A.mask[e] = (uint16_t)o.mask;
    }
}

// ------------------------------------------------------------------------------------------------
// small passes
// ------------------------------------------------------------------------------------------------
__global__ void frame_setup_kernel(float* extents, uint32_t* dirty)
{
    // ecs_scene.cpp:1482-1483
    if (threadIdx.x < 3)
        extents[threadIdx.x] = FLT_MAX;
    else if (threadIdx.x < 6)
        extents[threadIdx.x] = -FLT_MAX;
    if (threadIdx.x == 6)
        *dirty = 0u;
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

// loop C (ecs_scene.cpp:1529-1556): every allocated non-root entity folds its box into each
// ancestor reachable through allocated links.  The reference's descending sweep produces exactly
// that union; min/max are exact, so the atomics give the same values in any order.
__global__ void expand_kernel(const SceneArrays A, const uint32_t* __restrict__ children, uint32_t num_children)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= num_children)
        return;
    uint32_t n = children[i];
    if (n >= A.n || n == 0)
        return;
    const float* own = A.bv_out + (size_t)n * 7;
    float        b[6];
#pragma unroll
    for (int k = 0; k < 6; ++k)
        b[k] = own[k]; // may already include descendants: a superset that reaches the same ancestors
    uint32_t cur = n;
    for (;;)
    {
        if (!(A.entities[cur] & CMP_ALLOCATED)) // :1531
            break;
        uint32_t p = A.parents[cur];
        if (p == cur) // :1535
            break;
        float* pb = A.bv_out + (size_t)p * 7;
#pragma unroll
        for (int k = 0; k < 3; ++k)
        {
            if (b[k] < pb[k])
                atomic_min_float(&pb[k], b[k]);
            if (b[3 + k] > pb[3 + k])
                atomic_max_float(&pb[3 + k], b[3 + k]);
        }
        cur = p;
    }
}

// render_scene_view's filter + cull (ecs_scene.cpp:959-960) for a batch of views, from the
// stored pos_extent (48 B read, 2 B written per entity)
template <int MODE>
__global__ void __launch_bounds__(256) cull_kernel(const SceneArrays A, const __grid_constant__ ViewSet VS, uint32_t nv)
{
    uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= A.n)
        return;
    uint64_t ef = A.entities[e], sf = A.state_flags[e];
    uint32_t mask = 0;
    if (filter_pass(ef, sf))
    {
        float4 p0 = A.pos_extent[(size_t)e * 2], p1 = A.pos_extent[(size_t)e * 2 + 1];
        float  pos[3] = {p0.x, p0.y, p0.z};
        float  ext[3] = {p1.x, p1.y, p1.z};
        mask = cull_views<MODE>(VS, nv, pos, ext, p1.w);
    }
    A.mask[e] = (uint16_t)mask;
}

// ------------------------------------------------------------------------------------------------
// compaction: masks -> ascending per-view lists
// ------------------------------------------------------------------------------------------------
constexpr uint32_t CROWS = COMPACT_TILE / 256;

__global__ void __launch_bounds__(256) compact_count_kernel(const uint16_t* __restrict__ mask, uint32_t n, uint32_t nv,
                                                           uint32_t num_tiles, uint32_t* __restrict__ tile_counts)
{
    __shared__ uint32_t s_cnt[PMB_MAX_VIEWS];
    if (threadIdx.x < PMB_MAX_VIEWS)
        s_cnt[threadIdx.x] = 0;
    __syncthreads();
    uint32_t base = blockIdx.x * COMPACT_TILE;
    uint32_t m[CROWS];
#pragma unroll
    for (uint32_t j = 0; j < CROWS; ++j)
    {
        uint32_t e = base + j * 256 + threadIdx.x;
        m[j] = e < n ? mask[e] : 0u;
    }
    for (uint32_t v = 0; v < nv; ++v)
    {
        uint32_t c = 0;
#pragma unroll
        for (uint32_t j = 0; j < CROWS; ++j)
            c += (m[j] >> v) & 1u;
        c = __reduce_add_sync(0xFFFFFFFFu, c);
        if ((threadIdx.x & 31) == 0 && c)
            atomicAdd(&s_cnt[v], c);
    }
    __syncthreads();
    if (threadIdx.x < nv)
        tile_counts[(size_t)threadIdx.x * num_tiles + blockIdx.x] = s_cnt[threadIdx.x];
}

// one CTA per view: exclusive scan of the tile counts in place, total -> counts[view_base + v]
__global__ void __launch_bounds__(1024) compact_scan_kernel(uint32_t* __restrict__ tile_counts, uint32_t num_tiles,
                                                           uint32_t* __restrict__ counts, uint32_t view_base)
{
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    uint32_t*           row = tile_counts + (size_t)blockIdx.x * num_tiles;
    const uint32_t      lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0)
        s_carry = 0;
    __syncthreads();
    for (uint32_t start = 0; start < num_tiles; start += 1024)
    {
        uint32_t i = start + threadIdx.x;
        uint32_t x = i < num_tiles ? row[i] : 0u;
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
        if (i < num_tiles)
            row[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023)
            s_carry = excl + x;
        __syncthreads();
    }
    if (threadIdx.x == 0)
        counts[view_base + blockIdx.x] = s_carry;
}

__global__ void __launch_bounds__(256) compact_scatter_kernel(const uint16_t* __restrict__ mask, uint32_t n, uint32_t nv,
                                                             uint32_t num_tiles, const uint32_t* __restrict__ tile_offsets,
                                                             uint32_t* __restrict__ lists, size_t list_stride)
{
    __shared__ uint32_t s_seg[PMB_MAX_VIEWS][CROWS * 8];
    const uint32_t      lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t      base = blockIdx.x * COMPACT_TILE;
    uint32_t            m[CROWS];
#pragma unroll
    for (uint32_t j = 0; j < CROWS; ++j)
    {
        uint32_t e = base + j * 256 + threadIdx.x;
        m[j] = e < n ? mask[e] : 0u;
    }
    // per (view, row, warp) segment counts
    for (uint32_t v = 0; v < nv; ++v)
    {
#pragma unroll
        for (uint32_t j = 0; j < CROWS; ++j)
        {
            uint32_t b = __ballot_sync(0xFFFFFFFFu, (m[j] >> v) & 1u);
            if (lane == 0)
                s_seg[v][j * 8 + warp] = __popc(b);
        }
    }
    __syncthreads();
    // exclusive scan of the CROWS*8 (= 32) segments of each view: one warp per view
    for (uint32_t v = warp; v < nv; v += 8)
    {
        uint32_t x = s_seg[v][lane];
        uint32_t inc = x;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1)
        {
            uint32_t y = __shfl_up_sync(0xFFFFFFFFu, inc, d);
            if ((int)lane >= d)
                inc += y;
        }
        s_seg[v][lane] = inc - x;
    }
    __syncthreads();
    const uint32_t lt = (1u << lane) - 1u;
    for (uint32_t v = 0; v < nv; ++v)
    {
        uint32_t  tile_off = tile_offsets[(size_t)v * num_tiles + blockIdx.x];
        uint32_t* out = lists + (size_t)v * list_stride;
#pragma unroll
        for (uint32_t j = 0; j < CROWS; ++j)
        {
            uint32_t bit = (m[j] >> v) & 1u;
            uint32_t b = __ballot_sync(0xFFFFFFFFu, bit);
            if (bit)
                out[tile_off + s_seg[v][j * 8 + warp] + __popc(b & lt)] = base + j * 256 + threadIdx.x;
        }
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
        uint32_t cur = i;
        for (;;)
        {
            uint32_t p = parents[cur];
            if (p == cur || p > cur) // roots are self-parented; p > cur violates children-below-parents
                break;
            ++d;
            cur = p;
        }
        depth[i] = d;
    }
    uint32_t wmax = __reduce_max_sync(0xFFFFFFFFu, d);
    if ((threadIdx.x & 31) == 0 && wmax > *max_depth)
        atomicMax(max_depth, wmax);
}

__global__ void children_kernel(const uint32_t* __restrict__ parents, uint32_t n, uint32_t* __restrict__ children,
                                uint32_t* __restrict__ num_children)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool     is_child = i < n && parents[i] != i;
    uint32_t b = __ballot_sync(0xFFFFFFFFu, is_child);
    if (!b)
        return;
    uint32_t lane = threadIdx.x & 31;
    uint32_t pos = 0;
    if (lane == 0)
        pos = atomicAdd(num_children, __popc(b));
    pos = __shfl_sync(0xFFFFFFFFu, pos, 0);
    if (is_child)
        children[pos + __popc(b & ((1u << lane) - 1u))] = i;
}

__global__ void band_count_kernel(const uint32_t* __restrict__ depth, uint32_t n, uint32_t* __restrict__ band_counts)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && depth[i] > WALK)
        atomicAdd(&band_counts[(depth[i] - 1) / WALK], 1u);
}

__global__ void band_fill_kernel(const uint32_t* __restrict__ depth, uint32_t n, const uint32_t* __restrict__ band_offsets,
                                 uint32_t* __restrict__ band_cursor, uint32_t* __restrict__ band_list)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && depth[i] > WALK)
    {
        uint32_t b = (depth[i] - 1) / WALK;
        band_list[band_offsets[b] + atomicAdd(&band_cursor[b], 1u)] = i;
    }
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
// launchers
// ------------------------------------------------------------------------------------------------
static inline uint32_t cdiv(uint64_t a, uint32_t b) { return (uint32_t)((a + b - 1) / b); }

void launch_frame_setup(const SceneArrays& A, cudaStream_t st)
{
    frame_setup_kernel<<<1, 32, 0, st>>>(A.extents, A.dirty);
}

void launch_frame(const SceneArrays& A, const ViewSet& vs, uint32_t nv, int mode, cudaStream_t st)
{
    if (A.n == 0)
        return;
    uint32_t grid = cdiv(A.n, TILE);
    switch (mode)
    {
    case MODE_UPDATE: frame_kernel<MODE_UPDATE><<<grid, TILE, 0, st>>>(A, vs, nv); break;
    case MODE_AABB: frame_kernel<MODE_AABB><<<grid, TILE, 0, st>>>(A, vs, nv); break;
    case MODE_SPHERE: frame_kernel<MODE_SPHERE><<<grid, TILE, 0, st>>>(A, vs, nv); break;
    default: frame_kernel<MODE_FILTER><<<grid, TILE, 0, st>>>(A, vs, nv); break;
    }
}

void launch_band(const SceneArrays& A, const ViewSet& vs, uint32_t nv, int mode, const uint32_t* band_list,
                 uint32_t band_count, const uint32_t* depth, uint32_t band_index, cudaStream_t st)
{
    if (band_count == 0)
        return;
    uint32_t grid = cdiv(band_count, 256);
    switch (mode)
    {
    case MODE_UPDATE: band_kernel<MODE_UPDATE><<<grid, 256, 0, st>>>(A, vs, nv, band_list, band_count, depth, band_index); break;
    case MODE_AABB: band_kernel<MODE_AABB><<<grid, 256, 0, st>>>(A, vs, nv, band_list, band_count, depth, band_index); break;
    case MODE_SPHERE: band_kernel<MODE_SPHERE><<<grid, 256, 0, st>>>(A, vs, nv, band_list, band_count, depth, band_index); break;
    default: band_kernel<MODE_FILTER><<<grid, 256, 0, st>>>(A, vs, nv, band_list, band_count, depth, band_index); break;
    }
}

void launch_expand(const SceneArrays& A, const uint32_t* children, uint32_t num_children, cudaStream_t st)
{
    if (num_children == 0)
        return;
    expand_kernel<<<cdiv(num_children, 256), 256, 0, st>>>(A, children, num_children);
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
    case MODE_AABB: cull_kernel<MODE_AABB><<<grid, 256, 0, st>>>(A, vs, nv); break;
    case MODE_SPHERE: cull_kernel<MODE_SPHERE><<<grid, 256, 0, st>>>(A, vs, nv); break;
    default: cull_kernel<MODE_FILTER><<<grid, 256, 0, st>>>(A, vs, nv); break;
    }
}

uint32_t compact_num_tiles(uint32_t n) { return cdiv(n, COMPACT_TILE); }

void launch_compact(const uint16_t* mask, uint32_t n, uint32_t nv, uint32_t view_base, uint32_t* tile_counts,
                    uint32_t* lists, size_t list_stride, uint32_t* counts, cudaStream_t st)
{
    if (nv == 0)
        return;
    uint32_t tiles = compact_num_tiles(n);
    if (tiles == 0)
    {
        cudaMemsetAsync(counts + view_base, 0, sizeof(uint32_t) * nv, st);
        return;
    }
    compact_count_kernel<<<tiles, 256, 0, st>>>(mask, n, nv, tiles, tile_counts);
    compact_scan_kernel<<<nv, 1024, 0, st>>>(tile_counts, tiles, counts, view_base);
    compact_scatter_kernel<<<tiles, 256, 0, st>>>(mask, n, nv, tiles, tile_counts, lists + (size_t)view_base * list_stride,
                                                  list_stride);
}

void launch_depth(const uint32_t* parents, uint32_t n, uint32_t* depth, uint32_t* max_depth, cudaStream_t st)
{
    cudaMemsetAsync(max_depth, 0, sizeof(uint32_t), st);
    if (n)
        depth_kernel<<<cdiv(n, 256), 256, 0, st>>>(parents, n, depth, max_depth);
}

void launch_children(const uint32_t* parents, uint32_t n, uint32_t* children, uint32_t* num_children, cudaStream_t st)
{
    cudaMemsetAsync(num_children, 0, sizeof(uint32_t), st);
    if (n)
        children_kernel<<<cdiv(n, 256), 256, 0, st>>>(parents, n, children, num_children);
}

void launch_band_count(const uint32_t* depth, uint32_t n, uint32_t* band_counts, cudaStream_t st)
{
    if (n)
        band_count_kernel<<<cdiv(n, 256), 256, 0, st>>>(depth, n, band_counts);
}

void launch_band_fill(const uint32_t* depth, uint32_t n, const uint32_t* band_offsets, uint32_t* band_cursor,
                      uint32_t* band_list, cudaStream_t st)
{
    if (n)
        band_fill_kernel<<<cdiv(n, 256), 256, 0, st>>>(depth, n, band_offsets, band_cursor, band_list);
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

} // namespace pm
