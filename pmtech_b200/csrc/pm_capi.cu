// pm_capi.cu — the C ABI of include/pmtech_b200.h: device mirrors of the component arrays,
// transfers, and the sequencing of the kernels in pm_kernels.cu on one CUDA stream.
//
// There is no CPU fallback anywhere in this file: without a CUDA device pmb_scene_create fails
// with PMB_ERR_NO_DEVICE and nothing else can be called.
#include "pm_kernels.h"

#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

using namespace pm;

// ------------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_err;

static int fail(int code, const char* fmt, ...)
{
    char    buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

// for the host-only translation units of the library (pm_pms.cpp)
int pmb_fail_invalid(const char* fmt, ...)
{
    char    buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return PMB_ERR_INVALID;
}

int pm::fail_code(int code, const char* fmt, ...)
{
    char    buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(expr)                                                                                                 \
    do                                                                                                           \
    {                                                                                                            \
        cudaError_t _e = (expr);                                                                                 \
        if (_e != cudaSuccess)                                                                                   \
            return fail(_e == cudaErrorMemoryAllocation ? PMB_ERR_OOM : PMB_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, \
                        cudaGetErrorString(_e), __FILE__, __LINE__);                                             \
    } while (0)

#define CHECK_LAUNCH() CU(cudaGetLastError())

// ------------------------------------------------------------------------------------------------
// the handle
// ------------------------------------------------------------------------------------------------
struct TimedLaunch
{
    cudaEvent_t a, b;
    int         id;
};

struct pmb_scene
{
    int          device = 0;
    uint32_t     capacity = 0; // scene->soa_size
    uint32_t     padded = 0;   // capacity rounded up to PAD
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    cudaStream_t side = nullptr; // the draw-list compaction of a fused frame runs here, beside loop C and the flag clears
    cudaEvent_t  ev_fork = nullptr, ev_join = nullptr;

    // shadow views fitted on the device: light directions + constraint box in, cameras + plane constants out
    float*      shadow_in = nullptr;   // f32[PMB_MAX_VIEWS * 3 + 6]
    pmb_camera* shadow_cams = nullptr; // [PMB_MAX_VIEWS]
    ViewSet*    shadow_vs = nullptr;
    uint32_t    shadow_count = 0;

    // mirrors (ecs_scene.h:460-543, the arrays of this path)
    uint32_t* parents = nullptr;
    float4*   local = nullptr;
    float4*   world = nullptr;
    uint64_t* entities = nullptr;
    uint64_t* state_flags = nullptr;
    uint8_t*  has_rb = nullptr; // allocated on first upload
    uint32_t* cbuffer = nullptr;    // renderer handles, allocated on first use
    float4*   draw_calls = nullptr; // cmp_draw_call mirror, 10 float4 per entity, allocated on first use
    bool      draw_user_data = false; // the host uploaded records: v1.zw / v2 may hold its data and are preserved
    float*    transforms = nullptr; // cmp_transform mirror, 10 floats per entity; allocated on first upload (enables the bake)
    float*    bv_in = nullptr;
    float*    bv_out = nullptr;
    float4*   pos_extent = nullptr;
    uint32_t* bitmap = nullptr; // u32[PMB_MAX_VIEWS][padded / 32]
    float*    extents = nullptr; // 2 slots of 8 floats (6 used): the streaming pass arms the other slot for the next frame
    uint32_t  extents_slot = 0;  // the armed slot = the one the next frame accumulates into / the one an empty scene reports
    uint32_t  totals_slot = 0;   // which set of group / chunk totals the next one-launch cull adds to: flipped by those launches only
    uint32_t  extents_read = 0;  // the slot the last update wrote
    uint32_t* dirty = nullptr;   // 2 slots, 32 words apart, paired with the extents slots
    uint32_t* ctrl = nullptr;    // CTRL_WORDS: tile counters, grid barriers, exit counter (self-resetting, pm_kernels.h)
    uint32_t  frame_no = 0;      // stamp of the wavefront form; never 0
    FrameGeometry geom;

    // hierarchy index, rebuilt when parents change
    bool      topo_dirty = true;
    bool      topo_check = false;  // parents were re-uploaded: d_scalars[3] says whether any value changed
    uint32_t  topo_n = 0;
    uint32_t* depth = nullptr;     // u32[N]; between index builds: staging for re-uploaded parents
    uint32_t* cross = nullptr;     // children loop C cannot finish inside their own tile (see launch_cross)
    uint32_t* d_scalars = nullptr; // [0] max depth, [2] cross-tile child count, [3] parents-changed marker
    uint32_t  max_depth = 0;
    bool      anc_lists = false;   // wavefront form: loop C across tiles runs from flat ancestor lists (else it walks parent links)
    uint32_t  num_cross = 0;
    bool      deep = false;        // wavefront form of loop A (max_depth > SHALLOW_DEPTH, or forced)
    int       hierarchy_mode = PMB_HIERARCHY_AUTO;
    int       fused_tail = 1;      // PMB_OPT_FUSED_TAIL
    uint64_t* publish = nullptr;   // u64[tiles], wavefront form: scratch of the index build
    TilePlan* plans = nullptr;     // [tiles], wavefront form: the static schedule of every tile
    uint32_t* anc_off = nullptr;   // wavefront form: ancestor lists of the boundary children (CSR offsets, num_cross + 1) ...
    uint32_t* anc = nullptr;       // ... and the ancestors, nearest first
    uint32_t  anc_off_cap = 0, anc_cap = 0;
    void*     scan_temp = nullptr;
    size_t    scan_temp_bytes = 0;
    uint64_t* xbuf = nullptr;      // u64[N][16], wavefront form: allocated when a hierarchy first needs it
    uint64_t* trace = nullptr;     // PMB_OPT_TRACE: u64[tiles][4] timeline of the wavefront form's tiles
    uint32_t* tile_counts = nullptr; // per-tile / per-group / per-chunk visible counts when padded <= TAIL_MAX_ENTITIES

    // draw lists
    uint32_t* lists = nullptr;
    size_t    list_stride = 0;
    uint32_t  lists_views = 0;
    uint32_t* chunk_counts = nullptr;
    uint32_t* d_counts = nullptr;
    uint32_t* h_counts = nullptr; // pinned
    uint32_t  counts_cap = 0;
    uint32_t  last_views = 0;
    bool      counts_on_host = false;

    // animation hand-off: root-motion staging, skins and their palettes
    uint32_t* rm_targets = nullptr;
    float*    rm_rot = nullptr;
    float*    rm_delta = nullptr;
    uint32_t  rm_cap = 0;
    pmb_skin* skins = nullptr;
    uint32_t  num_skins = 0;
    float4*   bind = nullptr;
    float4*   palettes = nullptr;
    uint32_t  num_palette_matrices = 0;

    // cmp_bounding_volume staging (AoS <-> split)
    float*   bv_stage = nullptr;
    uint32_t bv_stage_records = 0;

    // instrumentation
    int                      profiling = 0; // 0 off, 1 every launch group, 2 the streaming pass (PMB_K_FRAME) only
    std::vector<TimedLaunch> timed;
    std::vector<cudaEvent_t> event_pool;
    double                   k_ms[PMB_K_COUNT] = {0};
    uint64_t                 k_launches[PMB_K_COUNT] = {0};
    uint64_t                 total_launches = 0;
};

// the one-launch tail's count tables: per tile, per group of 32 tiles, per chunk of 1024 tiles, PMB_MAX_VIEWS words each
constexpr uint32_t TAIL_TILES = TAIL_MAX_ENTITIES / TILE;
constexpr uint64_t TAIL_MAX_CROSS_WORK = 1u << 20; // boundary children x hierarchy depth above which loop C gets its own launch
constexpr uint32_t TAIL_TOTAL_WORDS = (TAIL_TILES / 32 + TAIL_TILES / 1024) * PMB_MAX_VIEWS; // one set of group + chunk totals
constexpr uint32_t TAIL_COUNT_WORDS = TAIL_TILES * PMB_MAX_VIEWS + 2 * TAIL_TOTAL_WORDS;

static inline uint32_t round_up(uint32_t v, uint32_t m) { return (uint32_t)(((uint64_t)v + m - 1) / m * m); }

struct Timer
{
    // brackets a group of launches with events when profiling is on; always counts launches
    pmb_scene* s;
    int        id;
    cudaEvent_t a = nullptr, b = nullptr;
    cudaStream_t st;
    Timer(pmb_scene* s_, int id_, uint64_t launches, cudaStream_t on = nullptr) : s(s_), id(id_), st(on ? on : s_->stream)
    {
        s->total_launches += launches;
        s->k_launches[id] += launches;
        if (!s->profiling || (s->profiling == 2 && id != PMB_K_FRAME))
            return;
        a = take();
        b = take();
        cudaEventRecord(a, st);
    }
    ~Timer()
    {
        if (!a)
            return;
        cudaEventRecord(b, st);
        s->timed.push_back({a, b, id});
    }
    cudaEvent_t take()
    {
        if (!s->event_pool.empty())
        {
            cudaEvent_t e = s->event_pool.back();
            s->event_pool.pop_back();
            return e;
        }
        cudaEvent_t e;
        cudaEventCreate(&e);
        return e;
    }
};

static void drain_timed(pmb_scene* s)
{
    for (auto& t : s->timed)
    {
        cudaEventSynchronize(t.b);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess)
            s->k_ms[t.id] += ms;
        s->event_pool.push_back(t.a);
        s->event_pool.push_back(t.b);
    }
    s->timed.clear();
}

template <typename T>
static int dev_alloc_zero(T** p, size_t count, cudaStream_t st)
{
    *p = nullptr;
    if (count == 0)
        return PMB_OK;
    CU(cudaMalloc((void**)p, count * sizeof(T)));
    CU(cudaMemsetAsync(*p, 0, count * sizeof(T), st));
    return PMB_OK;
}

template <typename T>
static void dev_free(T*& p)
{
    if (p)
        cudaFree(p);
    p = nullptr;
}

static int alloc_mirrors(pmb_scene* s, uint32_t capacity)
{
    uint32_t padded = round_up(std::max(capacity, 1u), PAD);
    size_t   n = padded;
    int      r;
#define A(call)            \
    if ((r = (call)) != 0) \
    return r
    A(dev_alloc_zero(&s->parents, n, s->stream));
    A(dev_alloc_zero(&s->local, n * 4, s->stream));
    A(dev_alloc_zero(&s->world, n * 4, s->stream));
    A(dev_alloc_zero(&s->entities, n, s->stream));
    A(dev_alloc_zero(&s->state_flags, n, s->stream));
    A(dev_alloc_zero(&s->bv_in, n * 6, s->stream));
    A(dev_alloc_zero(&s->bv_out, n * 7, s->stream));
    A(dev_alloc_zero(&s->pos_extent, n * 2, s->stream));
    A(dev_alloc_zero(&s->bitmap, (size_t)PMB_MAX_VIEWS * (n / 32), s->stream));
    A(dev_alloc_zero(&s->depth, n, s->stream));
    A(dev_alloc_zero(&s->cross, n, s->stream));
    A(dev_alloc_zero(&s->publish, n / TILE, s->stream));
    A(dev_alloc_zero(&s->plans, n / TILE, s->stream));
    if (padded <= TAIL_MAX_ENTITIES)
        A(dev_alloc_zero(&s->tile_counts, (size_t)TAIL_COUNT_WORDS, s->stream));
    A(dev_alloc_zero(&s->chunk_counts, (size_t)PMB_MAX_VIEWS * compact_num_chunks(padded), s->stream));
#undef A
    s->capacity = capacity;
    s->padded = padded;
    s->topo_dirty = true;
    return PMB_OK;
}

static void free_mirrors(pmb_scene* s)
{
    dev_free(s->parents);
    dev_free(s->local);
    dev_free(s->world);
    dev_free(s->entities);
    dev_free(s->state_flags);
    dev_free(s->has_rb);
    dev_free(s->bv_in);
    dev_free(s->bv_out);
    dev_free(s->pos_extent);
    dev_free(s->bitmap);
    dev_free(s->depth);
    dev_free(s->cross);
    dev_free(s->publish);
    dev_free(s->plans);
    dev_free(s->anc_off);
    dev_free(s->anc);
    s->anc_off_cap = s->anc_cap = 0;
    dev_free(s->xbuf);
    dev_free(s->trace);
    dev_free(s->tile_counts);
    dev_free(s->chunk_counts);
    dev_free(s->lists);
    s->lists_views = 0;
    s->list_stride = 0;
}

bool pm::scene_peek(pmb_scene* s, ScenePeek* out)
{
    if (!s || !out)
        return false;
    out->device = s->device;
    out->stream = s->stream;
    out->counts = s->d_counts;
    out->lists = s->lists;
    out->list_stride = s->list_stride;
    out->extents = s->extents + 8 * s->extents_read;
    out->last_views = s->last_views;
    return true;
}

static int bind(pmb_scene* s)
{
    if (!s)
        return fail(PMB_ERR_INVALID, "null scene handle");
    CU(cudaSetDevice(s->device));
    return PMB_OK;
}

#define BIND(s)               \
    do                        \
    {                         \
        int _r = bind(s);     \
        if (_r != PMB_OK)     \
            return _r;        \
    } while (0)

// ------------------------------------------------------------------------------------------------
// lifetime
// ------------------------------------------------------------------------------------------------
extern "C" int pmb_device_count(void)
{
    int         n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess)
    {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" int pmb_scene_create(pmb_scene** out, uint32_t capacity, int cuda_device)
{
    if (!out)
        return fail(PMB_ERR_INVALID, "out is null");
    *out = nullptr;
    int         ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
    {
        cudaGetLastError();
        return fail(PMB_ERR_NO_DEVICE, "no CUDA device: %s (this library has no CPU path)",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    }
    if (cuda_device < 0 || cuda_device >= ndev)
        return fail(PMB_ERR_INVALID, "cuda_device %d out of range [0,%d)", cuda_device, ndev);
    CU(cudaSetDevice(cuda_device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, cuda_device));
    if (prop.major < 10)
        return fail(PMB_ERR_NO_DEVICE, "device %d is sm_%d%d; this build carries sm_100a code only", cuda_device, prop.major,
                    prop.minor);
    pmb_scene* s = new pmb_scene();
    s->device = cuda_device;
    cudaError_t se = cudaStreamCreateWithFlags(&s->own_stream, cudaStreamNonBlocking);
    if (se != cudaSuccess)
    {
        delete s;
        return fail(PMB_ERR_CUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(se));
    }
    s->stream = s->own_stream;
    if (cudaStreamCreateWithFlags(&s->side, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->ev_join, cudaEventDisableTiming) != cudaSuccess)
    {
        pmb_scene_destroy(s);
        return fail(PMB_ERR_CUDA, "side stream / events: %s", cudaGetErrorString(cudaGetLastError()));
    }
    int r = alloc_mirrors(s, capacity);
    if (r == PMB_OK)
        r = dev_alloc_zero(&s->extents, 16, s->stream);
    if (r == PMB_OK)
        r = dev_alloc_zero(&s->dirty, 64, s->stream);
    if (r == PMB_OK)
        r = dev_alloc_zero(&s->ctrl, CTRL_WORDS, s->stream);
    if (r == PMB_OK)
        r = dev_alloc_zero(&s->d_scalars, 4, s->stream);
    if (r == PMB_OK)
    {
        launch_extents_init(s->extents, s->stream); // ecs_scene.cpp:1482-1483 for both slots
        cudaError_t ge = frame_geometry(&s->geom, cuda_device);
        if (ge == cudaSuccess)
            ge = cudaGetLastError();
        if (ge != cudaSuccess)
            r = fail(PMB_ERR_CUDA, "kernel set-up failed: %s", cudaGetErrorString(ge));
    }
    if (r != PMB_OK)
    {
        std::string keep = g_err;
        pmb_scene_destroy(s);
        g_err = keep;
        return r;
    }
    *out = s;
    return PMB_OK;
}

extern "C" void pmb_scene_destroy(pmb_scene* s)
{
    if (!s)
        return;
    cudaSetDevice(s->device);
    cudaStreamSynchronize(s->stream);
    drain_timed(s);
    for (auto e : s->event_pool)
        cudaEventDestroy(e);
    free_mirrors(s);
    dev_free(s->extents);
    dev_free(s->dirty);
    dev_free(s->ctrl);
    dev_free(s->d_scalars);
    if (s->scan_temp)
        cudaFree(s->scan_temp);
    dev_free(s->d_counts);
    dev_free(s->bv_stage);
    dev_free(s->transforms);
    dev_free(s->cbuffer);
    dev_free(s->draw_calls);
    dev_free(s->rm_targets);
    dev_free(s->rm_rot);
    dev_free(s->rm_delta);
    dev_free(s->skins);
    dev_free(s->bind);
    dev_free(s->palettes);
    dev_free(s->shadow_in);
    dev_free(s->shadow_cams);
    dev_free(s->shadow_vs);
    if (s->h_counts)
        cudaFreeHost(s->h_counts);
    if (s->side)
    {
        cudaStreamSynchronize(s->side);
        cudaStreamDestroy(s->side);
    }
    if (s->ev_fork)
        cudaEventDestroy(s->ev_fork);
    if (s->ev_join)
        cudaEventDestroy(s->ev_join);
    if (s->own_stream)
        cudaStreamDestroy(s->own_stream);
    delete s;
}

static int ensure_draw_mirrors(pmb_scene* s)
{
    if (s->cbuffer && s->draw_calls)
        return PMB_OK;
    int r = dev_alloc_zero(&s->cbuffer, (size_t)s->padded, s->stream);
    if (r == PMB_OK)
        r = dev_alloc_zero(&s->draw_calls, (size_t)s->padded * 10, s->stream);
    return r;
}

extern "C" int pmb_scene_resize(pmb_scene* s, uint32_t new_capacity)
{
    BIND(s);
    if (round_up(std::max(new_capacity, 1u), PAD) == s->padded)
    {
        s->capacity = new_capacity; // the padded mirrors already cover it (and are zeroed beyond the old size)
        return PMB_OK;
    }
    CU(cudaStreamSynchronize(s->stream));
    pmb_scene old = *s; // shallow: keeps the old pointers
    s->has_rb = nullptr;
    s->transforms = nullptr;
    s->cbuffer = nullptr;
    s->draw_calls = nullptr;
    s->lists = nullptr;
    s->lists_views = 0;
    s->tile_counts = nullptr;
    s->xbuf = nullptr;
    s->trace = nullptr;
    int r = alloc_mirrors(s, new_capacity);
    if (r != PMB_OK)
        return r;
    if (old.trace)
    {
        // PMB_OPT_TRACE stays on: the timeline is per tile, so it grows with the scene (a stale one would be written past its end)
        cudaFree(old.trace);
        r = dev_alloc_zero(&s->trace, (size_t)(s->padded / TILE) * 4, s->stream);
        if (r != PMB_OK)
            return r;
    }
    size_t keep = std::min(old.padded, s->padded);
#define COPY(field, per)                                                                                              \
    CU(cudaMemcpyAsync(s->field, old.field, keep * (per) * sizeof(*s->field), cudaMemcpyDeviceToDevice, s->stream))
    COPY(parents, 1);
    COPY(local, 4);
    COPY(world, 4);
    COPY(entities, 1);
    COPY(state_flags, 1);
    COPY(bv_in, 6);
    COPY(bv_out, 7);
    COPY(pos_extent, 2);
#undef COPY
    if (old.has_rb)
    {
        r = dev_alloc_zero(&s->has_rb, (size_t)s->padded, s->stream);
        if (r != PMB_OK)
            return r;
        CU(cudaMemcpyAsync(s->has_rb, old.has_rb, keep, cudaMemcpyDeviceToDevice, s->stream));
    }
    if (old.transforms)
    {
        r = dev_alloc_zero(&s->transforms, (size_t)s->padded * 10, s->stream);
        if (r != PMB_OK)
            return r;
        CU(cudaMemcpyAsync(s->transforms, old.transforms, keep * 40, cudaMemcpyDeviceToDevice, s->stream));
    }
    if (old.cbuffer)
    {
        r = ensure_draw_mirrors(s);
        if (r != PMB_OK)
            return r;
        CU(cudaMemcpyAsync(s->cbuffer, old.cbuffer, keep * 4, cudaMemcpyDeviceToDevice, s->stream));
        CU(cudaMemcpyAsync(s->draw_calls, old.draw_calls, keep * 160, cudaMemcpyDeviceToDevice, s->stream));
    }
    CU(cudaStreamSynchronize(s->stream));
    // free the old mirrors
    cudaFree(old.parents), cudaFree(old.local), cudaFree(old.world), cudaFree(old.entities), cudaFree(old.state_flags);
    cudaFree(old.bv_in), cudaFree(old.bv_out), cudaFree(old.pos_extent), cudaFree(old.bitmap), cudaFree(old.depth);
    cudaFree(old.cross), cudaFree(old.chunk_counts), cudaFree(old.publish), cudaFree(old.plans);
    if (old.xbuf)
        cudaFree(old.xbuf);
    if (old.tile_counts)
        cudaFree(old.tile_counts);
    if (old.has_rb)
        cudaFree(old.has_rb);
    if (old.transforms)
        cudaFree(old.transforms);
    if (old.cbuffer)
        cudaFree(old.cbuffer);
    if (old.draw_calls)
        cudaFree(old.draw_calls);
    if (old.lists)
        cudaFree(old.lists);
    return PMB_OK;
}

extern "C" uint32_t pmb_scene_capacity(const pmb_scene* s) { return s ? s->capacity : 0; }
extern "C" int      pmb_scene_device(const pmb_scene* s) { return s ? s->device : -1; }

extern "C" int pmb_set_stream(pmb_scene* s, void* cuda_stream)
{
    BIND(s);
    CU(cudaStreamSynchronize(s->stream));
    s->stream = cuda_stream ? (cudaStream_t)cuda_stream : s->own_stream;
    return PMB_OK;
}

extern "C" int pmb_set_option(pmb_scene* s, int option, int value)
{
    BIND(s);
    switch (option)
    {
    case PMB_OPT_HIERARCHY:
        if (value != PMB_HIERARCHY_AUTO && value != PMB_HIERARCHY_CHAINS && value != PMB_HIERARCHY_WAVEFRONT)
            return fail(PMB_ERR_INVALID, "unknown hierarchy mode %d", value);
        if (s->hierarchy_mode != value)
            s->topo_dirty = true; // the two forms of loop A need different indices
        s->hierarchy_mode = value;
        return PMB_OK;
    case PMB_OPT_FUSED_TAIL: s->fused_tail = value < 0 ? 0 : (value > 2 ? 2 : value); return PMB_OK;
    case PMB_OPT_TRACE:
        CU(cudaStreamSynchronize(s->stream));
        dev_free(s->trace);
        if (value)
            return dev_alloc_zero(&s->trace, (size_t)(s->padded / TILE) * 4, s->stream);
        return PMB_OK;
    default: return fail(PMB_ERR_INVALID, "unknown option %d", option);
    }
}

extern "C" int pmb_download_trace(pmb_scene* s, uint64_t* host, uint32_t num_tiles)
{
    BIND(s);
    if (!s->trace || !host || num_tiles > s->padded / TILE)
        return fail(PMB_ERR_STATE, "no tile trace (PMB_OPT_TRACE off) or bad range");
    CU(cudaMemcpyAsync(host, s->trace, (size_t)num_tiles * 32, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return PMB_OK;
}

extern "C" int pmb_synchronize(pmb_scene* s)
{
    BIND(s);
    CU(cudaStreamSynchronize(s->stream));
    return PMB_OK;
}

// ------------------------------------------------------------------------------------------------
// transfers
// ------------------------------------------------------------------------------------------------
static const uint32_t BV_STAGE_RECORDS = 1u << 21; // 2 Mi records = 104 MiB of AoS per staging pass

static int ensure_bv_stage(pmb_scene* s, uint32_t records)
{
    uint32_t want = std::min(records, BV_STAGE_RECORDS);
    if (s->bv_stage_records >= want)
        return PMB_OK;
    CU(cudaStreamSynchronize(s->stream));
    dev_free(s->bv_stage);
    s->bv_stage_records = 0;
    CU(cudaMalloc((void**)&s->bv_stage, (size_t)want * 13 * sizeof(float)));
    s->bv_stage_records = want;
    return PMB_OK;
}

static int range_ok(pmb_scene* s, uint32_t first, uint32_t count)
{
    if ((uint64_t)first + count > s->capacity)
        return fail(PMB_ERR_INVALID, "range [%u,%u) exceeds capacity %u", first, first + count, s->capacity);
    return PMB_OK;
}

extern "C" int pmb_upload(pmb_scene* s, int array_id, const void* host, uint32_t first, uint32_t count)
{
    BIND(s);
    if (count == 0)
        return PMB_OK;
    if (!host)
        return fail(PMB_ERR_INVALID, "host pointer is null");
    int r = range_ok(s, first, count);
    if (r != PMB_OK)
        return r;
    cudaStream_t st = s->stream;
    switch (array_id)
    {
    case PMB_PARENTS:
        if (s->topo_dirty)
            CU(cudaMemcpyAsync(s->parents + first, host, (size_t)count * 4, cudaMemcpyHostToDevice, st));
        else
        {
            // A binding without dirty tracking re-sends the parents every frame (they rarely change): the values are
            // staged in the depth scratch and merged by a compare-and-copy pass, and the hierarchy index is rebuilt
            // only if one of them really differs (checked by the next update).
            CU(cudaMemcpyAsync(s->depth + first, host, (size_t)count * 4, cudaMemcpyHostToDevice, st));
            Timer t(s, PMB_K_AUX, 1);
            launch_merge_parents(s->depth + first, s->parents + first, count, &s->d_scalars[3], st);
            CHECK_LAUNCH();
            s->topo_check = true;
        }
        break;
    case PMB_LOCAL_MATRICES:
        CU(cudaMemcpyAsync(s->local + (size_t)first * 4, host, (size_t)count * 64, cudaMemcpyHostToDevice, st));
        break;
    case PMB_WORLD_MATRICES:
        CU(cudaMemcpyAsync(s->world + (size_t)first * 4, host, (size_t)count * 64, cudaMemcpyHostToDevice, st));
        break;
    case PMB_ENTITIES:
        CU(cudaMemcpyAsync(s->entities + first, host, (size_t)count * 8, cudaMemcpyHostToDevice, st));
        break;
    case PMB_STATE_FLAGS:
        CU(cudaMemcpyAsync(s->state_flags + first, host, (size_t)count * 8, cudaMemcpyHostToDevice, st));
        break;
    case PMB_POS_EXTENT:
        CU(cudaMemcpyAsync(s->pos_extent + (size_t)first * 2, host, (size_t)count * 32, cudaMemcpyHostToDevice, st));
        break;
    case PMB_HAS_RB_MATRIX:
        if (!s->has_rb)
        {
            r = dev_alloc_zero(&s->has_rb, (size_t)s->padded, st);
            if (r != PMB_OK)
                return r;
        }
        CU(cudaMemcpyAsync(s->has_rb + first, host, (size_t)count, cudaMemcpyHostToDevice, st));
        break;
    case PMB_TRANSFORMS:
        if (!s->transforms)
        {
            r = dev_alloc_zero(&s->transforms, (size_t)s->padded * 10, st);
            if (r != PMB_OK)
                return r;
        }
        CU(cudaMemcpyAsync(s->transforms + (size_t)first * 10, host, (size_t)count * 40, cudaMemcpyHostToDevice, st));
        break;
    case PMB_CBUFFER:
        r = ensure_draw_mirrors(s);
        if (r != PMB_OK)
            return r;
        CU(cudaMemcpyAsync(s->cbuffer + first, host, (size_t)count * 4, cudaMemcpyHostToDevice, st));
        break;
    case PMB_DRAW_CALLS:
        r = ensure_draw_mirrors(s);
        if (r != PMB_OK)
            return r;
        CU(cudaMemcpyAsync(s->draw_calls + (size_t)first * 10, host, (size_t)count * 160, cudaMemcpyHostToDevice, st));
        s->draw_user_data = true;
        break;
    case PMB_BOUNDING_VOLUMES:
    {
        r = ensure_bv_stage(s, count);
        if (r != PMB_OK)
            return r;
        const char* src = (const char*)host;
        for (uint32_t done = 0; done < count;)
        {
            uint32_t c = std::min(count - done, s->bv_stage_records);
            CU(cudaMemcpyAsync(s->bv_stage, src + (size_t)done * 52, (size_t)c * 52, cudaMemcpyHostToDevice, st));
            Timer t(s, PMB_K_AUX, 1);
            launch_bv_split(s->bv_stage, c, s->bv_in + (size_t)(first + done) * 6, s->bv_out + (size_t)(first + done) * 7, st);
            CHECK_LAUNCH();
            done += c;
        }
        break;
    }
    default: return fail(PMB_ERR_INVALID, "unknown array id %d", array_id);
    }
    return PMB_OK;
}

extern "C" int pmb_download(pmb_scene* s, int array_id, void* host, uint32_t first, uint32_t count)
{
    BIND(s);
    if (count == 0)
        return PMB_OK;
    if (!host)
        return fail(PMB_ERR_INVALID, "host pointer is null");
    int r = range_ok(s, first, count);
    if (r != PMB_OK)
        return r;
    cudaStream_t st = s->stream;
    switch (array_id)
    {
    case PMB_PARENTS: CU(cudaMemcpyAsync(host, s->parents + first, (size_t)count * 4, cudaMemcpyDeviceToHost, st)); break;
    case PMB_LOCAL_MATRICES:
        CU(cudaMemcpyAsync(host, s->local + (size_t)first * 4, (size_t)count * 64, cudaMemcpyDeviceToHost, st));
        break;
    case PMB_WORLD_MATRICES:
        CU(cudaMemcpyAsync(host, s->world + (size_t)first * 4, (size_t)count * 64, cudaMemcpyDeviceToHost, st));
        break;
    case PMB_ENTITIES: CU(cudaMemcpyAsync(host, s->entities + first, (size_t)count * 8, cudaMemcpyDeviceToHost, st)); break;
    case PMB_STATE_FLAGS:
        CU(cudaMemcpyAsync(host, s->state_flags + first, (size_t)count * 8, cudaMemcpyDeviceToHost, st));
        break;
    case PMB_POS_EXTENT:
        CU(cudaMemcpyAsync(host, s->pos_extent + (size_t)first * 2, (size_t)count * 32, cudaMemcpyDeviceToHost, st));
        break;
    case PMB_HAS_RB_MATRIX:
        if (!s->has_rb)
            memset(host, 0, count);
        else
            CU(cudaMemcpyAsync(host, s->has_rb + first, (size_t)count, cudaMemcpyDeviceToHost, st));
        break;
    case PMB_TRANSFORMS:
        if (!s->transforms)
            return fail(PMB_ERR_STATE, "transforms were never uploaded");
        CU(cudaMemcpyAsync(host, s->transforms + (size_t)first * 10, (size_t)count * 40, cudaMemcpyDeviceToHost, st));
        break;
    case PMB_CBUFFER:
        r = ensure_draw_mirrors(s);
        if (r != PMB_OK)
            return r;
        CU(cudaMemcpyAsync(host, s->cbuffer + first, (size_t)count * 4, cudaMemcpyDeviceToHost, st));
        break;
    case PMB_DRAW_CALLS:
        r = ensure_draw_mirrors(s);
        if (r != PMB_OK)
            return r;
        CU(cudaMemcpyAsync(host, s->draw_calls + (size_t)first * 10, (size_t)count * 160, cudaMemcpyDeviceToHost, st));
        break;
    case PMB_BOUNDING_VOLUMES:
    {
        r = ensure_bv_stage(s, count);
        if (r != PMB_OK)
            return r;
        char* dst = (char*)host;
        for (uint32_t done = 0; done < count;)
        {
            uint32_t c = std::min(count - done, s->bv_stage_records);
            {
                Timer t(s, PMB_K_AUX, 1);
                launch_bv_merge(s->bv_in + (size_t)(first + done) * 6, s->bv_out + (size_t)(first + done) * 7, c, s->bv_stage, st);
                CHECK_LAUNCH();
            }
            CU(cudaMemcpyAsync(dst + (size_t)done * 52, s->bv_stage, (size_t)c * 52, cudaMemcpyDeviceToHost, st));
            done += c;
        }
        break;
    }
    default: return fail(PMB_ERR_INVALID, "unknown array id %d", array_id);
    }
    CU(cudaStreamSynchronize(st));
    return PMB_OK;
}

extern "C" void* pmb_device_ptr(pmb_scene* s, int array_id)
{
    if (!s)
        return nullptr;
    switch (array_id)
    {
    case PMB_PARENTS: return s->parents;
    case PMB_LOCAL_MATRICES: return s->local;
    case PMB_WORLD_MATRICES: return s->world;
    case PMB_ENTITIES: return s->entities;
    case PMB_STATE_FLAGS: return s->state_flags;
    case PMB_POS_EXTENT: return s->pos_extent;
    case PMB_HAS_RB_MATRIX: return s->has_rb;
    default: return nullptr;
    }
}

extern "C" int pmb_device_bounds(pmb_scene* s, const float** bounds_in, float** bounds_out)
{
    if (!s)
        return fail(PMB_ERR_INVALID, "null scene handle");
    if (bounds_in)
        *bounds_in = s->bv_in;
    if (bounds_out)
        *bounds_out = s->bv_out;
    return PMB_OK;
}

extern "C" void* pmb_host_alloc(size_t bytes)
{
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess)
    {
        cudaGetLastError();
        fail(PMB_ERR_OOM, "cudaHostAlloc(%zu) failed", bytes);
        return nullptr;
    }
    return p;
}

extern "C" void pmb_host_free(void* p)
{
    if (p)
        cudaFreeHost(p);
}

// ------------------------------------------------------------------------------------------------
// hierarchy index
// ------------------------------------------------------------------------------------------------
static int ensure_topology(pmb_scene* s, uint32_t n)
{
    cudaStream_t st = s->stream;
    if (s->topo_check && !s->topo_dirty)
    {
        // parents were re-uploaded since the last build: did any value change?
        uint32_t changed = 0;
        CU(cudaMemcpyAsync(&changed, &s->d_scalars[3], 4, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (changed)
        {
            CU(cudaMemsetAsync(&s->d_scalars[3], 0, 4, st));
            s->topo_dirty = true;
        }
    }
    s->topo_check = false;
    if (!s->topo_dirty && s->topo_n == n)
        return PMB_OK;
    {
        Timer t(s, PMB_K_AUX, 1);
        launch_depth(s->parents, n, s->depth, &s->d_scalars[0], st);
        CHECK_LAUNCH();
    }
    uint32_t sc[3] = {0, 0, 0};
    CU(cudaMemcpyAsync(sc, s->d_scalars, 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    s->max_depth = sc[0];
    // Shallow scenes (instanced props, a few levels): every non-root entity re-multiplies its own short chain inside
    // the streaming pass, which needs nothing from other tiles.  That costs O(depth) products per entity, so deeper
    // scenes (skeleton rigs, config 2's depth-16 trees) run the wavefront form: one product per entity, level by
    // level inside a tile, parents in earlier tiles read once their owner has published them.
    s->deep = s->max_depth > SHALLOW_DEPTH;
    if (s->hierarchy_mode == PMB_HIERARCHY_WAVEFRONT)
        s->deep = true;
    else if (s->hierarchy_mode == PMB_HIERARCHY_CHAINS && s->max_depth <= WALK)
        s->deep = false;
    if (s->deep && !s->xbuf)
    {
        // the exchange buffer of the wavefront form: 128 B per entity, only the published entities' lines are ever touched
        int r = dev_alloc_zero(&s->xbuf, (size_t)s->padded * 16, st);
        if (r != PMB_OK)
            return r;
    }
    {
        Timer t(s, PMB_K_AUX, s->deep ? 3 : 1);
        launch_cross(s->parents, n, s->deep, s->cross, &s->d_scalars[2], st);
        if (s->deep)
        {
            launch_publish_mask(s->parents, n, s->padded, s->publish, st);
            launch_plan(s->parents, s->depth, n, s->publish, s->plans, st);
        }
        CHECK_LAUNCH();
    }
    CU(cudaMemcpyAsync(sc, s->d_scalars, sizeof(sc), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    s->num_cross = sc[2];
    // The lists cost the sum of depth over the boundary children: a few entries per entity for bushy scenes, but about
    // n^2 / 128 for one long chain (300 MB at 100 k entities, and past 2^32 — the u32 offsets — soon after).  num_cross x
    // max_depth bounds the sum; beyond a budget of max(32 n, 4 Mi) entries loop C across tiles walks the parent links
    // (expand_up_warp: any depth, no memory, slower per level) instead.
    const uint64_t anc_bound = (uint64_t)s->num_cross * std::max(s->max_depth, 1u);
    const uint64_t anc_budget = std::min<uint64_t>(std::max<uint64_t>(32ull * n, 1ull << 22), 1ull << 31);
    s->anc_lists = false;
    if (s->deep && s->num_cross && anc_bound <= anc_budget)
    {
        // loop C across tiles works from flat ancestor lists (one per boundary child) instead of chasing parent links
        if (s->num_cross + 1 > s->anc_off_cap)
        {
            dev_free(s->anc_off);
            s->anc_off_cap = 0;
            CU(cudaMalloc((void**)&s->anc_off, ((size_t)s->num_cross + 1) * 4));
            s->anc_off_cap = s->num_cross + 1;
        }
        {
            Timer t(s, PMB_K_AUX, 2);
            launch_ancestor_counts(s->parents, s->cross, s->num_cross, s->anc_off, st);
            CU(scan_ancestor_offsets(s->anc_off, s->num_cross, &s->scan_temp, &s->scan_temp_bytes, st));
        }
        uint32_t total = 0;
        CU(cudaMemcpyAsync(&total, s->anc_off + s->num_cross, 4, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (total > s->anc_cap)
        {
            dev_free(s->anc);
            s->anc_cap = 0;
            CU(cudaMalloc((void**)&s->anc, (size_t)total * 4));
            s->anc_cap = total;
        }
        Timer t(s, PMB_K_AUX, 1);
        launch_ancestor_fill(s->parents, s->cross, s->num_cross, s->anc_off, s->anc, st);
        CHECK_LAUNCH();
        s->anc_lists = true;
    }
    s->topo_dirty = false;
    s->topo_n = n;
    return PMB_OK;
}

static SceneArrays arrays(pmb_scene* s, uint32_t n)
{
    SceneArrays A;
    A.n = n;
    A.parents = s->parents;
    A.local = s->local;
    A.world = s->world;
    A.entities = s->entities;
    A.state_flags = s->state_flags;
    A.has_rb = s->has_rb;
    A.bv_in = s->bv_in;
    A.bv_out = s->bv_out;
    A.pos_extent = s->pos_extent;
    A.bitmap = s->bitmap;
    A.bitmap_stride = s->padded / 32;
    A.extents = s->extents + 8 * s->extents_slot;
    A.extents_next = s->extents + 8 * (s->extents_slot ^ 1u);
    A.dirty = s->dirty + 32 * s->extents_slot;
    A.dirty_next = s->dirty + 32 * (s->extents_slot ^ 1u);
    A.ctrl = s->ctrl;
    A.plans = s->plans;
    A.xbuf = s->xbuf;
    A.frame_no = s->frame_no;
    A.trace = s->trace;
    A.cross = s->cross;
    A.num_cross = s->num_cross;
    A.anc_off = s->anc_off;
    A.anc = s->anc;
    A.tile_counts = s->tile_counts;
    if (s->tile_counts)
    {
        // two sets of group / chunk totals: a one-launch cull adds to one and zeroes the other for the next such launch.
        // They alternate on their own toggle — update-only and multi-launch frames flip the extents slot but neither
        // dirty nor clear a set, so following extents_slot would hand a later cull a set that still holds old totals.
        uint32_t* t0 = s->tile_counts + TAIL_TILES * PMB_MAX_VIEWS;
        uint32_t* t1 = t0 + TAIL_TOTAL_WORDS;
        uint32_t* cur = s->totals_slot ? t1 : t0;
        uint32_t* nxt = s->totals_slot ? t0 : t1;
        A.group_totals = cur;
        A.chunk_totals = cur + (TAIL_TILES / 32) * PMB_MAX_VIEWS;
        A.group_totals_next = nxt;
        A.chunk_totals_next = nxt + (TAIL_TILES / 32) * PMB_MAX_VIEWS;
    }
    else
        A.group_totals = A.chunk_totals = A.group_totals_next = A.chunk_totals_next = nullptr;
    A.lists = s->lists;
    A.list_stride = s->list_stride;
    A.counts = s->d_counts;
    return A;
}

static int ensure_lists(pmb_scene* s, uint32_t num_views)
{
    if (num_views > s->counts_cap)
    {
        CU(cudaStreamSynchronize(s->stream));
        dev_free(s->d_counts);
        if (s->h_counts)
            cudaFreeHost(s->h_counts);
        s->h_counts = nullptr;
        uint32_t cap = std::max(num_views, PMB_MAX_VIEWS);
        int      r = dev_alloc_zero(&s->d_counts, cap, s->stream);
        if (r != PMB_OK)
            return r;
        CU(cudaHostAlloc((void**)&s->h_counts, (size_t)cap * 4, cudaHostAllocDefault));
        s->counts_cap = cap;
    }
    if (num_views > s->lists_views || s->list_stride != s->padded)
    {
        CU(cudaStreamSynchronize(s->stream));
        dev_free(s->lists);
        s->lists_views = 0;
        s->list_stride = s->padded;
        CU(cudaMalloc((void**)&s->lists, (size_t)num_views * s->list_stride * 4));
        s->lists_views = num_views;
    }
    return PMB_OK;
}

static int map_mode(int cull_mode)
{
    switch (cull_mode)
    {
    case PMB_CULL_AABB: return MODE_AABB;
    case PMB_CULL_SPHERE: return MODE_SPHERE;
    case PMB_CULL_FILTER_ONLY: return MODE_FILTER;
    default: return -1;
    }
}

// stages 1-2 (+ the first view batch fused when mode != MODE_UPDATE)
static int run_update(pmb_scene* s, uint32_t n, const ViewSet& vs, uint32_t nv, int mode, bool compact = false)
{
    cudaStream_t st = s->stream;
    if (s->transforms && n)
    {
        // stage 0 (ecs_scene.cpp:1415-1427): local = T * R * S for the entities flagged e_cmp::transform
        Timer t(s, PMB_K_ANIM, 1);
        launch_bake(s->transforms, s->entities, s->state_flags, s->local, n, st);
    }
    s->extents_read = s->extents_slot; // an empty scene reports the armed slot: +-FLT_MAX (ecs_scene.cpp:1482-1483)
    if (n == 0)
    {
        if (compact && nv)
            CU(cudaMemsetAsync(s->d_counts, 0, sizeof(uint32_t) * nv, st));
        return PMB_OK;
    }
    if (++s->frame_no == 0)
        s->frame_no = 1;
    SceneArrays A = arrays(s, n);
    // up to TAIL_MAX_ENTITIES the whole frame is this one launch (pmtech's own demos are 32 Ki and 64 Ki entities)
    // ... unless loop C across tiles is heavy (many boundary children under deep chains): that part is pointer chasing,
    // which the stand-alone expand launch hides with 2048 threads per SM against the streaming pass's 800
    // (measured at config 2, 260 k boundary children, depth 16: 0.40 ms in one launch, 0.30 ms in three)
    const bool light_cross = (uint64_t)s->num_cross * std::max(s->max_depth, 1u) <= TAIL_MAX_CROSS_WORK;
    const bool tail = s->fused_tail && s->tile_counts && n <= TAIL_MAX_ENTITIES && (mode == MODE_UPDATE || compact) &&
                      (light_cross || s->fused_tail > 1) && (!s->deep || s->anc_lists || !s->num_cross);
    {
        Timer t(s, PMB_K_FRAME, 1);
        launch_frame(A, vs, nv, mode, s->deep, tail, s->geom, st);
    }
    s->extents_slot ^= 1u; // the launch armed the other slot for the next frame
    if (tail)
    {
        if (mode != MODE_UPDATE)
            s->totals_slot ^= 1u; // ... and, having added to one set of totals, cleared the other
        CHECK_LAUNCH();
        return PMB_OK;
    }
    // the visibility bitmaps are final here, and the compaction into draw lists touches nothing that loop C and
    // the flag clears touch — it runs beside them on the side stream and joins at the end
    const bool fork = compact && nv;
    if (fork)
    {
        CU(cudaEventRecord(s->ev_fork, st));
        CU(cudaStreamWaitEvent(s->side, s->ev_fork, 0));
        {
            Timer t(s, PMB_K_COMPACT, compact_num_launches(n), s->side);
            launch_compact(s->bitmap, s->padded / 32, n, nv, 0, s->chunk_counts, s->lists, s->list_stride, s->d_counts, s->side);
        }
        CU(cudaEventRecord(s->ev_join, s->side));
    }
    {
        Timer t(s, PMB_K_EXPAND, s->num_cross ? 1 : 0);
        if (s->deep && s->anc_lists)
            launch_expand_lists(A, st);
        else
            launch_expand(A, s->cross, s->num_cross, st);
    }
    {
        Timer t(s, PMB_K_AUX, 1);
        launch_clear_flags(A, st);
    }
    if (fork)
        CU(cudaStreamWaitEvent(st, s->ev_join, 0));
    CHECK_LAUNCH();
    return PMB_OK;
}

static int run_cull_batches(pmb_scene* s, uint32_t n, const pmb_frustum* views, uint32_t num_views, int mode,
                            uint32_t first_view)
{
    SceneArrays  A = arrays(s, n);
    cudaStream_t st = s->stream;
    for (uint32_t v0 = first_view; v0 < num_views; v0 += PMB_MAX_VIEWS)
    {
        uint32_t nv = std::min(num_views - v0, PMB_MAX_VIEWS);
        ViewSet  vs;
        make_view_set(&vs, views + v0, nv);
        {
            Timer t(s, PMB_K_CULL, n ? 1 : 0);
            launch_cull(A, vs, nv, mode, st);
        }
        {
            Timer t(s, PMB_K_COMPACT, compact_num_launches(n));
            launch_compact(s->bitmap, s->padded / 32, n, nv, v0, s->chunk_counts, s->lists, s->list_stride, s->d_counts, st);
        }
        CHECK_LAUNCH();
    }
    return PMB_OK;
}

static int check_n(pmb_scene* s, uint32_t n)
{
    if (n > s->capacity)
        return fail(PMB_ERR_INVALID, "num_entities %u exceeds capacity %u", n, s->capacity);
    return PMB_OK;
}

extern "C" int pmb_update_scene(pmb_scene* s, uint32_t num_entities)
{
    BIND(s);
    int r = check_n(s, num_entities);
    if (r == PMB_OK)
        r = ensure_topology(s, num_entities);
    if (r != PMB_OK)
        return r;
    ViewSet vs;
    make_view_set(&vs, nullptr, 0); // no planes, but the packed arithmetic still needs its `one`
    return run_update(s, num_entities, vs, 0, MODE_UPDATE);
}

extern "C" int pmb_get_renderable_extents(pmb_scene* s, float out[6])
{
    BIND(s);
    if (!out)
        return fail(PMB_ERR_INVALID, "out is null");
    CU(cudaMemcpyAsync(out, s->extents + 8 * s->extents_read, 6 * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return PMB_OK;
}

extern "C" int pmb_cull_views(pmb_scene* s, uint32_t num_entities, const pmb_frustum* views, uint32_t num_views, int cull_mode)
{
    BIND(s);
    int mode = map_mode(cull_mode);
    if (mode < 0)
        return fail(PMB_ERR_INVALID, "unknown cull mode %d", cull_mode);
    if (num_views && !views)
        return fail(PMB_ERR_INVALID, "views is null");
    int r = check_n(s, num_entities);
    if (r == PMB_OK)
        r = ensure_lists(s, std::max(num_views, 1u));
    if (r != PMB_OK)
        return r;
    s->last_views = num_views;
    s->counts_on_host = false;
    return run_cull_batches(s, num_entities, views, num_views, mode, 0);
}

extern "C" int pmb_mark_transforms_dirty(pmb_scene* s, uint32_t first, uint32_t count)
{
    BIND(s);
    if (count == 0)
        return PMB_OK;
    int r = range_ok(s, first, count);
    if (r != PMB_OK)
        return r;
    {
        Timer t(s, PMB_K_ANIM, 1);
        launch_mark_dirty(s->entities, first, count, s->stream);
    }
    CHECK_LAUNCH();
    return PMB_OK;
}

extern "C" int pmb_reserve_views(pmb_scene* s, uint32_t num_views)
{
    BIND(s);
    return ensure_lists(s, std::max(num_views, 1u));
}

// shadow_camera_from_entity + camera_update_(shadow_)frustum + cull, without taking renderable_extents or the lights'
// world matrices to the host.  kind 0: directional (host_in = directions [+ constraint box]); 1: spot (entity indices).
static int run_fitted_views(pmb_scene* s, uint32_t num_entities, uint32_t num_lights, int cull_mode, uint32_t first_view, int kind,
                            const float* directions, const float* constraints, const uint32_t* light_entities)
{
    int mode = map_mode(cull_mode);
    if (mode < 0)
        return fail(PMB_ERR_INVALID, "unknown cull mode %d", cull_mode);
    if (num_lights == 0)
        return PMB_OK;
    if (kind == 0 ? !directions : !light_entities)
        return fail(PMB_ERR_INVALID, kind == 0 ? "light_directions is null" : "light_entities is null");
    if (num_lights > PMB_MAX_VIEWS)
        return fail(PMB_ERR_INVALID, "%u shadow views in one call, at most %u", num_lights, PMB_MAX_VIEWS);
    int r = check_n(s, num_entities);
    if (r != PMB_OK)
        return r;
    if (first_view && (first_view + num_lights > s->lists_views || s->list_stride != s->padded))
        return fail(PMB_ERR_STATE, "lists [%u,%u) not reserved: call pmb_reserve_views before the frame whose lists must survive",
                    first_view, first_view + num_lights);
    r = ensure_lists(s, first_view + num_lights);
    if (r != PMB_OK)
        return r;
    cudaStream_t st = s->stream;
    if (!s->shadow_in)
    {
        r = dev_alloc_zero(&s->shadow_in, PMB_MAX_VIEWS * 3 + 6, st);
        if (r == PMB_OK)
            r = dev_alloc_zero(&s->shadow_cams, PMB_MAX_VIEWS, st);
        if (r == PMB_OK)
            r = dev_alloc_zero(&s->shadow_vs, 1, st);
        if (r != PMB_OK)
            return r;
    }
    float h_in[PMB_MAX_VIEWS * 3 + 6] = {0};
    if (kind == 0)
    {
        memcpy(h_in, directions, (size_t)num_lights * 12);
        if (constraints)
            memcpy(h_in + PMB_MAX_VIEWS * 3, constraints, 24);
    }
    else
        memcpy(h_in, light_entities, (size_t)num_lights * 4);
    CU(cudaMemcpyAsync(s->shadow_in, h_in, sizeof(h_in), cudaMemcpyHostToDevice, st)); // pageable: staged before the call returns
    SceneArrays A = arrays(s, num_entities);
    {
        Timer t(s, PMB_K_AUX, 1);
        if (kind == 0)
            launch_shadow_fit(s->extents + 8 * s->extents_read, s->shadow_in, constraints ? s->shadow_in + PMB_MAX_VIEWS * 3 : nullptr, num_lights,
                              s->shadow_cams, s->shadow_vs, st);
        else
        {
            pmb_camera pc;
            pmb_camera_create_perspective(&pc, 100.0f, 1.0f, 0.1f, 500.0f); // ecs_scene.cpp:558
            launch_spot_fit(s->world, reinterpret_cast<const uint32_t*>(s->shadow_in), num_lights, num_entities, pc.proj, s->shadow_cams,
                            s->shadow_vs, st);
        }
    }
    {
        Timer t(s, PMB_K_CULL, num_entities ? 1 : 0);
        launch_cull_device_views(A, s->shadow_vs, num_lights, mode, st);
    }
    {
        Timer t(s, PMB_K_COMPACT, compact_num_launches(num_entities));
        launch_compact(s->bitmap, s->padded / 32, num_entities, num_lights, first_view, s->chunk_counts, s->lists, s->list_stride,
                       s->d_counts, st);
    }
    CHECK_LAUNCH();
    s->shadow_count = num_lights;
    s->last_views = first_view ? std::max(s->last_views, first_view + num_lights) : num_lights;
    s->counts_on_host = false;
    return PMB_OK;
}

extern "C" int pmb_cull_shadow_views(pmb_scene* s, uint32_t num_entities, const float* light_directions, uint32_t num_lights,
                                     const float* extent_constraints, int cull_mode, uint32_t first_view)
{
    BIND(s);
    return run_fitted_views(s, num_entities, num_lights, cull_mode, first_view, 0, light_directions, extent_constraints, nullptr);
}

extern "C" int pmb_cull_spot_shadow_views(pmb_scene* s, uint32_t num_entities, const uint32_t* light_entities, uint32_t num_lights,
                                          int cull_mode, uint32_t first_view)
{
    BIND(s);
    for (uint32_t i = 0; light_entities && i < num_lights && i < PMB_MAX_VIEWS; ++i)
        if (light_entities[i] >= num_entities)
            return fail(PMB_ERR_INVALID, "light entity %u outside the %u entities", light_entities[i], num_entities);
    return run_fitted_views(s, num_entities, num_lights, cull_mode, first_view, 1, nullptr, nullptr, light_entities);
}

extern "C" int pmb_download_shadow_cameras(pmb_scene* s, pmb_camera* out, uint32_t count)
{
    BIND(s);
    if (count == 0)
        return PMB_OK;
    if (!out || count > s->shadow_count)
        return fail(PMB_ERR_INVALID, "%u shadow cameras asked, %u fitted by the last pmb_cull_shadow_views", count, s->shadow_count);
    CU(cudaMemcpyAsync(out, s->shadow_cams, (size_t)count * sizeof(pmb_camera), cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return PMB_OK;
}

extern "C" const pmb_camera* pmb_device_shadow_cameras(pmb_scene* s) { return s ? s->shadow_cams : nullptr; }

extern "C" int pmb_frame(pmb_scene* s, uint32_t num_entities, const pmb_frustum* views, uint32_t num_views, int cull_mode)
{
    BIND(s);
    int mode = map_mode(cull_mode);
    if (mode < 0)
        return fail(PMB_ERR_INVALID, "unknown cull mode %d", cull_mode);
    if (num_views && !views)
        return fail(PMB_ERR_INVALID, "views is null");
    int r = check_n(s, num_entities);
    if (r == PMB_OK)
        r = ensure_topology(s, num_entities);
    if (r == PMB_OK)
        r = ensure_lists(s, std::max(num_views, 1u));
    if (r != PMB_OK)
        return r;
    s->last_views = num_views;
    s->counts_on_host = false;
    uint32_t nv0 = std::min(num_views, PMB_MAX_VIEWS);
    ViewSet  vs;
    make_view_set(&vs, views, nv0);
    r = run_update(s, num_entities, vs, nv0, nv0 ? mode : MODE_UPDATE, true);
    if (r != PMB_OK)
        return r;
    if (num_views > nv0)
        r = run_cull_batches(s, num_entities, views, num_views, mode, nv0);
    return r;
}

// ------------------------------------------------------------------------------------------------
// draw lists
// ------------------------------------------------------------------------------------------------
static int fetch_counts(pmb_scene* s)
{
    if (s->counts_on_host)
        return PMB_OK;
    if (s->last_views)
    {
        CU(cudaMemcpyAsync(s->h_counts, s->d_counts, (size_t)s->last_views * 4, cudaMemcpyDeviceToHost, s->stream));
        CU(cudaStreamSynchronize(s->stream));
    }
    s->counts_on_host = true;
    return PMB_OK;
}

extern "C" int pmb_get_counts(pmb_scene* s, uint32_t* counts_out, uint32_t num_views)
{
    BIND(s);
    if (num_views > s->last_views)
        return fail(PMB_ERR_STATE, "%u views requested, the last cull built %u", num_views, s->last_views);
    int r = fetch_counts(s);
    if (r != PMB_OK)
        return r;
    for (uint32_t v = 0; v < num_views; ++v)
        counts_out[v] = s->h_counts[v];
    return PMB_OK;
}

extern "C" int64_t pmb_get_list(pmb_scene* s, uint32_t view, uint32_t* host_out, uint32_t capacity)
{
    int r = bind(s);
    if (r != PMB_OK)
        return r;
    if (view >= s->last_views)
        return fail(PMB_ERR_STATE, "view %u requested, the last cull built %u", view, s->last_views);
    r = fetch_counts(s);
    if (r != PMB_OK)
        return r;
    uint32_t c = s->h_counts[view];
    if (c > capacity)
        return fail(PMB_ERR_INVALID, "list of view %u has %u entries, caller capacity is %u", view, c, capacity);
    if (c)
    {
        CU(cudaMemcpyAsync(host_out, s->lists + (size_t)view * s->list_stride, (size_t)c * 4, cudaMemcpyDeviceToHost, s->stream));
        CU(cudaStreamSynchronize(s->stream));
    }
    return (int64_t)c;
}

// every list of the last cull in one go: the copies are queued back to back and waited for once
extern "C" int pmb_get_lists(pmb_scene* s, uint32_t* const* lists_out, const uint32_t* capacities, uint32_t* counts_out, uint32_t num_views)
{
    BIND(s);
    if (num_views > s->last_views)
        return fail(PMB_ERR_STATE, "%u views requested, the last cull built %u", num_views, s->last_views);
    if (num_views && (!lists_out || !capacities))
        return fail(PMB_ERR_INVALID, "null pointer");
    int r = fetch_counts(s);
    if (r != PMB_OK)
        return r;
    for (uint32_t v = 0; v < num_views; ++v)
    {
        const uint32_t c = s->h_counts[v];
        if (counts_out)
            counts_out[v] = c;
        if (c > capacities[v])
            return fail(PMB_ERR_INVALID, "list of view %u has %u entries, caller capacity is %u", v, c, capacities[v]);
    }
    for (uint32_t v = 0; v < num_views; ++v)
        if (s->h_counts[v] && lists_out[v])
            CU(cudaMemcpyAsync(lists_out[v], s->lists + (size_t)v * s->list_stride, (size_t)s->h_counts[v] * 4, cudaMemcpyDeviceToHost,
                               s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return PMB_OK;
}

extern "C" int pmb_get_lists_sb(pmb_scene* s, uint32_t** lists_out, uint32_t num_views)
{
    BIND(s);
    if (num_views > s->last_views)
        return fail(PMB_ERR_STATE, "%u views requested, the last cull built %u", num_views, s->last_views);
    int r = fetch_counts(s);
    if (r != PMB_OK)
        return r;
    for (uint32_t v = 0; v < num_views; ++v)
        lists_out[v] = nullptr;
    int         failed = PMB_OK;
    cudaError_t ce = cudaSuccess;
    for (uint32_t v = 0; v < num_views && failed == PMB_OK; ++v)
    {
        uint32_t c = s->h_counts[v];
        if (!c)
            continue; // the reference leaves the list null when nothing is pushed
        // stretchy buffer header: {int capacity, int count} 8 bytes before the data (data_struct.h:30-32)
        int* raw = (int*)malloc(sizeof(int) * 2 + (size_t)(c + 1) * sizeof(uint32_t));
        if (!raw)
        {
            failed = PMB_ERR_OOM;
            break;
        }
        raw[0] = (int)(c + 1);
        raw[1] = (int)c;
        lists_out[v] = (uint32_t*)(raw + 2);
        lists_out[v][c] = 0;
        ce = cudaMemcpyAsync(lists_out[v], s->lists + (size_t)v * s->list_stride, (size_t)c * 4, cudaMemcpyDeviceToHost, s->stream);
        if (ce != cudaSuccess)
            failed = PMB_ERR_CUDA;
    }
    cudaError_t se = cudaStreamSynchronize(s->stream); // also on failure: no copy may still be writing into a list that is freed
    if (failed == PMB_OK && se != cudaSuccess)
        failed = PMB_ERR_CUDA, ce = se;
    if (failed != PMB_OK)
    {
        for (uint32_t v = 0; v < num_views; ++v)
        {
            if (lists_out[v])
                free((int*)lists_out[v] - 2);
            lists_out[v] = nullptr;
        }
        return failed == PMB_ERR_OOM ? fail(PMB_ERR_OOM, "malloc of a draw list failed")
                                     : fail(PMB_ERR_CUDA, "draw-list read-back failed: %s", cudaGetErrorString(ce));
    }
    return PMB_OK;
}

extern "C" const uint32_t* pmb_device_list(pmb_scene* s, uint32_t view)
{
    if (!s || view >= s->lists_views)
        return nullptr;
    return s->lists + (size_t)view * s->list_stride;
}

extern "C" const uint32_t* pmb_device_counts(pmb_scene* s) { return s ? s->d_counts : nullptr; }

// ------------------------------------------------------------------------------------------------
// multi-GPU list exchange helpers
// ------------------------------------------------------------------------------------------------
extern "C" int pmb_copy_segments(pmb_scene* s, const uint32_t* src, uint32_t* dst, const pmb_segment* segments, uint32_t num_segments)
{
    BIND(s);
    if (num_segments == 0)
        return PMB_OK;
    if (!src || !dst || !segments)
        return fail(PMB_ERR_INVALID, "null pointer");
    for (uint32_t s0 = 0; s0 < num_segments; s0 += PMB_MAX_SEGMENTS)
    {
        uint32_t c = std::min(num_segments - s0, PMB_MAX_SEGMENTS);
        SegTable T;
        memset(&T, 0, sizeof(T));
        uint32_t mx = 0;
        for (uint32_t i = 0; i < c; ++i)
        {
            T.seg[i] = segments[s0 + i];
            mx = std::max(mx, T.seg[i].count);
        }
        Timer t(s, PMB_K_AUX, mx ? 1 : 0);
        launch_copy_segments(src, dst, T, c, mx, s->stream);
        CHECK_LAUNCH();
    }
    return PMB_OK;
}

extern "C" int pmb_pack_lists(pmb_scene* s, const uint32_t* counts, uint32_t num_views, uint32_t add, uint32_t* dst)
{
    BIND(s);
    if (num_views > s->lists_views)
        return fail(PMB_ERR_STATE, "%u views requested, %u lists exist", num_views, s->lists_views);
    std::vector<pmb_segment> segs(num_views);
    uint64_t                 off = 0;
    for (uint32_t v = 0; v < num_views; ++v)
    {
        if (counts[v] > s->list_stride)
            return fail(PMB_ERR_INVALID, "count %u of view %u exceeds the list capacity", counts[v], v);
        segs[v] = pmb_segment{(uint64_t)v * s->list_stride, off, counts[v], add};
        off += counts[v];
    }
    return pmb_copy_segments(s, s->lists, dst, segs.data(), num_views);
}

extern "C" int pmb_pack_row(pmb_scene* s, uint32_t num_views, uint32_t add, uint32_t* row_device, uint32_t row_stride)
{
    BIND(s);
    if (num_views == 0)
        return PMB_OK;
    if (num_views > s->lists_views || num_views > PMB_MAX_VIEWS)
        return fail(PMB_ERR_STATE, "%u views requested, %u lists exist (at most %u per row)", num_views, s->lists_views, PMB_MAX_VIEWS);
    if (!row_device || row_stride < num_views)
        return fail(PMB_ERR_INVALID, "row buffer missing or shorter than its header");
    Timer t(s, PMB_K_AUX, 1);
    launch_pack_row(s->lists, s->list_stride, s->d_counts, num_views, add, row_device, row_stride, s->stream);
    CHECK_LAUNCH();
    return PMB_OK;
}

extern "C" int pmb_unpack_rows(pmb_scene* s, const uint32_t* rows_device, uint32_t world, uint32_t row_stride, uint32_t num_views,
                               uint32_t* dst_device, uint64_t dst_capacity, uint32_t* meta_device, void* cuda_stream)
{
    BIND(s);
    if (num_views == 0 || world == 0)
        return PMB_OK;
    if (num_views > PMB_MAX_VIEWS || world > 16 || !rows_device || !dst_device || !meta_device || row_stride < num_views)
        return fail(PMB_ERR_INVALID, "bad row table (at most %u views x 16 ranks)", PMB_MAX_VIEWS);
    // not bracketed by profiling events: it may run on a stream other than the handle's
    s->total_launches += 1;
    s->k_launches[PMB_K_AUX] += 1;
    launch_unpack_rows(rows_device, world, row_stride, num_views, dst_device, dst_capacity, meta_device,
                       cuda_stream ? (cudaStream_t)cuda_stream : s->stream);
    CHECK_LAUNCH();
    return PMB_OK;
}

extern "C" int pmb_publish_counts(pmb_scene* s, const uint32_t* row_device, uint32_t num_views, uint32_t rank, uint32_t world,
                                  const uint64_t* peer_tables, void* cuda_stream)
{
    BIND(s);
    if (num_views == 0)
        return PMB_OK;
    if (!row_device || !peer_tables || num_views > PMB_MAX_VIEWS || world == 0 || world > PMB_MAX_PEERS || rank >= world)
        return fail(PMB_ERR_INVALID, "bad peer table (at most %u views, %u peers)", PMB_MAX_VIEWS, PMB_MAX_PEERS);
    PeerPtrs P;
    memset(&P, 0, sizeof(P));
    for (uint32_t q = 0; q < world; ++q)
        P.p[q] = peer_tables[q];
    s->total_launches += 1;
    s->k_launches[PMB_K_AUX] += 1;
    launch_publish_counts(row_device, num_views, rank, world, P, cuda_stream ? (cudaStream_t)cuda_stream : s->stream);
    CHECK_LAUNCH();
    return PMB_OK;
}

extern "C" int pmb_push_row(pmb_scene* s, const uint32_t* row_device, uint32_t row_stride, uint32_t num_views, uint32_t rank,
                            uint32_t world, const uint32_t* table_device, const uint64_t* peer_lists, uint64_t lists_capacity,
                            uint32_t* meta_device, void* cuda_stream)
{
    BIND(s);
    if (num_views == 0)
        return PMB_OK;
    if (!row_device || !table_device || !peer_lists || !meta_device || num_views > PMB_MAX_VIEWS || world == 0 ||
        world > PMB_MAX_PEERS || rank >= world || row_stride < num_views)
        return fail(PMB_ERR_INVALID, "bad peer table (at most %u views, %u peers)", PMB_MAX_VIEWS, PMB_MAX_PEERS);
    PeerPtrs P;
    memset(&P, 0, sizeof(P));
    for (uint32_t q = 0; q < world; ++q)
        P.p[q] = peer_lists[q];
    s->total_launches += 1;
    s->k_launches[PMB_K_AUX] += 1;
    launch_push_row(row_device, row_stride, num_views, rank, world, table_device, P, lists_capacity, meta_device,
                    cuda_stream ? (cudaStream_t)cuda_stream : s->stream);
    CHECK_LAUNCH();
    return PMB_OK;
}

extern "C" int pmb_update_draw_calls(pmb_scene* s, uint32_t num_entities, float time_ms)
{
    BIND(s);
    int r = check_n(s, num_entities);
    if (r == PMB_OK)
        r = ensure_draw_mirrors(s);
    if (r != PMB_OK || num_entities == 0)
        return r;
    Timer t(s, PMB_K_DRAW, 1);
    launch_draw_calls(s->world, s->entities, s->cbuffer, s->draw_calls, num_entities, time_ms, s->draw_user_data, s->stream);
    CHECK_LAUNCH();
    return PMB_OK;
}

// ecs_scene.cpp:1994-2011 — the walk over the master instances (host flags, no device)
extern "C" int64_t pmb_instance_blocks(const uint64_t* entities, const pmb_master_instance* master_instances, uint32_t num_entities,
                                       pmb_instance_block* blocks, uint32_t capacity)
{
    if (!entities || !master_instances || (capacity && !blocks))
        return fail(PMB_ERR_INVALID, "null pointer");
    const uint64_t CMP_MASTER_INSTANCE = 1ull << 15, CMP_CUSTOM_INSTANCE_BUFFER = 1ull << 21; // ecs_scene.h:144-171
    int64_t        count = 0;
    for (uint64_t n = 0; n < num_entities; ++n) // 64-bit: `n += num_instances` may step past the end like the reference's size_t
    {
        if (!(entities[n] & CMP_MASTER_INSTANCE))
            continue;
        if (entities[n] & CMP_CUSTOM_INSTANCE_BUFFER)
            continue;
        const pmb_master_instance& m = master_instances[n];
        if ((uint64_t)count < capacity)
        {
            pmb_instance_block& b = blocks[count];
            b.master = (uint32_t)n;
            b.instance_buffer = m.instance_buffer;
            b.byte_offset = (n + 1) * 160ull; // &scene->draw_call_data[n + 1]
            b.bytes = (uint64_t)m.num_instances * m.instance_stride;
        }
        ++count;
        n += m.num_instances; // stride over sub instances
    }
    return count;
}

extern "C" const void* pmb_device_draw_calls(pmb_scene* s) { return s ? s->draw_calls : nullptr; }

extern "C" int pmb_download_instance_block(pmb_scene* s, const pmb_instance_block* block, void* host_out)
{
    BIND(s);
    if (!block || !host_out)
        return fail(PMB_ERR_INVALID, "null pointer");
    if (!s->draw_calls)
        return fail(PMB_ERR_STATE, "no draw-call records yet: call pmb_update_draw_calls first");
    if (block->byte_offset + block->bytes > (uint64_t)s->padded * 160ull)
        return fail(PMB_ERR_INVALID, "instance block [%llu, +%llu) reaches outside the draw-call mirror", (unsigned long long)block->byte_offset,
                    (unsigned long long)block->bytes);
    if (block->bytes)
    {
        CU(cudaMemcpyAsync(host_out, reinterpret_cast<const char*>(s->draw_calls) + block->byte_offset, block->bytes, cudaMemcpyDeviceToHost,
                           s->stream));
        CU(cudaStreamSynchronize(s->stream));
    }
    return PMB_OK;
}

// ------------------------------------------------------------------------------------------------
// animation hand-off: root motion in, skin palettes out
// ------------------------------------------------------------------------------------------------
extern "C" int pmb_apply_root_motion(pmb_scene* s, const uint32_t* targets, const float* rotations, const float* deltas, uint32_t count)
{
    BIND(s);
    if (count == 0)
        return PMB_OK;
    if (!targets || !rotations || !deltas)
        return fail(PMB_ERR_INVALID, "null pointer");
    if (!s->transforms)
        return fail(PMB_ERR_STATE, "root motion needs PMB_TRANSFORMS uploaded");
    cudaStream_t st = s->stream;
    if (count > s->rm_cap)
    {
        CU(cudaStreamSynchronize(st));
        dev_free(s->rm_targets), dev_free(s->rm_rot), dev_free(s->rm_delta);
        s->rm_targets = nullptr, s->rm_rot = nullptr, s->rm_delta = nullptr, s->rm_cap = 0;
        CU(cudaMalloc((void**)&s->rm_targets, (size_t)count * 4));
        CU(cudaMalloc((void**)&s->rm_rot, (size_t)count * 16));
        CU(cudaMalloc((void**)&s->rm_delta, (size_t)count * 12));
        s->rm_cap = count;
    }
    CU(cudaMemcpyAsync(s->rm_targets, targets, (size_t)count * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s->rm_rot, rotations, (size_t)count * 16, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s->rm_delta, deltas, (size_t)count * 12, cudaMemcpyHostToDevice, st));
    Timer t(s, PMB_K_ANIM, 1);
    launch_root_motion(s->transforms, s->entities, s->rm_targets, s->rm_rot, s->rm_delta, count, s->capacity, st);
    CHECK_LAUNCH();
    return PMB_OK;
}

extern "C" int pmb_skins_set(pmb_scene* s, const pmb_skin* skins, uint32_t num_skins, const float* bind_matrices,
                             uint32_t num_bind_matrices, uint32_t num_palette_matrices)
{
    BIND(s);
    if (num_skins && (!skins || !bind_matrices))
        return fail(PMB_ERR_INVALID, "null pointer");
    for (uint32_t k = 0; k < num_skins; ++k)
    {
        const pmb_skin& j = skins[k];
        if (j.num_joints > 85)
            return fail(PMB_ERR_INVALID, "skin %u has %u joints; cmp_skin holds 85 (ecs_scene.h:196)", k, j.num_joints);
        if ((uint64_t)j.joints_offset + j.num_joints > s->capacity || (uint64_t)j.bind_offset + j.num_joints > num_bind_matrices ||
            (uint64_t)j.palette_offset + j.num_joints > num_palette_matrices)
            return fail(PMB_ERR_INVALID, "skin %u reaches outside the scene, the bind matrices or the palette buffer", k);
    }
    CU(cudaStreamSynchronize(s->stream));
    dev_free(s->skins), dev_free(s->bind), dev_free(s->palettes);
    s->skins = nullptr, s->bind = nullptr, s->palettes = nullptr, s->num_skins = 0, s->num_palette_matrices = 0;
    if (num_skins == 0)
        return PMB_OK;
    CU(cudaMalloc((void**)&s->skins, (size_t)num_skins * sizeof(pmb_skin)));
    CU(cudaMalloc((void**)&s->bind, (size_t)num_bind_matrices * 64));
    int r = dev_alloc_zero(&s->palettes, (size_t)num_palette_matrices * 4, s->stream);
    if (r != PMB_OK)
        return r;
    CU(cudaMemcpyAsync(s->skins, skins, (size_t)num_skins * sizeof(pmb_skin), cudaMemcpyHostToDevice, s->stream));
    CU(cudaMemcpyAsync(s->bind, bind_matrices, (size_t)num_bind_matrices * 64, cudaMemcpyHostToDevice, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    s->num_skins = num_skins;
    s->num_palette_matrices = num_palette_matrices;
    return PMB_OK;
}

extern "C" int pmb_skin_palettes(pmb_scene* s)
{
    BIND(s);
    if (s->num_skins == 0)
        return PMB_OK;
    Timer t(s, PMB_K_ANIM, 1);
    launch_skin(s->world, s->skins, s->num_skins, s->bind, s->palettes, s->capacity, s->stream);
    CHECK_LAUNCH();
    return PMB_OK;
}

extern "C" int pmb_download_palettes(pmb_scene* s, float* host, uint32_t first_matrix, uint32_t count)
{
    BIND(s);
    if (count == 0)
        return PMB_OK;
    if (!host || (uint64_t)first_matrix + count > s->num_palette_matrices)
        return fail(PMB_ERR_INVALID, "palette range [%u,%u) outside the %u matrices", first_matrix, first_matrix + count,
                    s->num_palette_matrices);
    CU(cudaMemcpyAsync(host, s->palettes + (size_t)first_matrix * 4, (size_t)count * 64, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return PMB_OK;
}

extern "C" float* pmb_device_palettes(pmb_scene* s) { return s ? reinterpret_cast<float*>(s->palettes) : nullptr; }

// ------------------------------------------------------------------------------------------------
// host-buffer frame: uploads + frame + list read-back in one call
// ------------------------------------------------------------------------------------------------
extern "C" int pmb_frame_host(pmb_scene* s, uint32_t n, const pmb_host_arrays* io, const pmb_frustum* views, uint32_t num_views,
                              int cull_mode, uint32_t** lists_out, uint32_t* counts_out)
{
    BIND(s);
    if (!io)
        return fail(PMB_ERR_INVALID, "io is null");
    int r = PMB_OK;
#define UP(ptr, id)                                     \
    if (r == PMB_OK && io->ptr)                         \
    r = pmb_upload(s, id, io->ptr, 0, n)
    UP(parents, PMB_PARENTS);
    UP(local_matrices, PMB_LOCAL_MATRICES);
    UP(entities, PMB_ENTITIES);
    UP(state_flags, PMB_STATE_FLAGS);
    UP(bounding_volumes, PMB_BOUNDING_VOLUMES);
    UP(has_rb_matrix, PMB_HAS_RB_MATRIX);
    UP(transforms, PMB_TRANSFORMS);
#undef UP
    if (r != PMB_OK)
        return r;
    r = pmb_frame(s, n, views, num_views, cull_mode);
    if (r != PMB_OK)
        return r;
    cudaStream_t st = s->stream;
    // everything below is queued behind the kernels; one synchronise at the end
    if (num_views)
    {
        CU(cudaMemcpyAsync(s->h_counts, s->d_counts, (size_t)num_views * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        s->counts_on_host = true;
        for (uint32_t v = 0; v < num_views; ++v)
        {
            uint32_t c = s->h_counts[v];
            if (counts_out)
                counts_out[v] = c;
            if (lists_out && lists_out[v] && c)
                CU(cudaMemcpyAsync(lists_out[v], s->lists + (size_t)v * s->list_stride, (size_t)c * 4, cudaMemcpyDeviceToHost, st));
        }
    }
    if (io->world_matrices_out)
        CU(cudaMemcpyAsync(io->world_matrices_out, s->world, (size_t)n * 64, cudaMemcpyDeviceToHost, st));
    if (io->pos_extent_out)
        CU(cudaMemcpyAsync(io->pos_extent_out, s->pos_extent, (size_t)n * 32, cudaMemcpyDeviceToHost, st));
    if (io->renderable_extents_out)
        CU(cudaMemcpyAsync(io->renderable_extents_out, s->extents + 8 * s->extents_read, 6 * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (io->bounding_volumes_out)
    {
        r = pmb_download(s, PMB_BOUNDING_VOLUMES, io->bounding_volumes_out, 0, n);
        if (r != PMB_OK)
            return r;
    }
    CU(cudaStreamSynchronize(st));
    return PMB_OK;
}

// ------------------------------------------------------------------------------------------------
// instrumentation
// ------------------------------------------------------------------------------------------------
extern "C" int pmb_set_profiling(pmb_scene* s, int enabled)
{
    BIND(s);
    CU(cudaStreamSynchronize(s->stream));
    drain_timed(s);
    s->profiling = enabled < 0 ? 0 : (enabled > 2 ? 1 : enabled);
    return PMB_OK;
}

extern "C" int pmb_kernel_stats(pmb_scene* s, int kernel_id, double* total_ms, uint64_t* launches)
{
    BIND(s);
    if (kernel_id < 0 || kernel_id >= PMB_K_COUNT)
        return fail(PMB_ERR_INVALID, "unknown kernel id %d", kernel_id);
    drain_timed(s);
    if (total_ms)
        *total_ms = s->k_ms[kernel_id];
    if (launches)
        *launches = s->k_launches[kernel_id];
    return PMB_OK;
}

extern "C" int pmb_reset_kernel_stats(pmb_scene* s)
{
    BIND(s);
    drain_timed(s);
    for (int i = 0; i < PMB_K_COUNT; ++i)
    {
        s->k_ms[i] = 0;
        s->k_launches[i] = 0;
    }
    return PMB_OK;
}

extern "C" uint64_t pmb_total_launches(const pmb_scene* s) { return s ? s->total_launches : 0; }

extern "C" const char* pmb_last_error(void) { return g_err.c_str(); }
extern "C" const char* pmb_version(void) { return "pmtech_b200 0.1 (sm_100a)"; }
