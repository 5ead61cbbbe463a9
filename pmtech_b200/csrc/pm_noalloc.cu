// pm_noalloc.cu — pmb_cull_views_static: the per-view filter + frustum cull + draw-list build for callers that
// cannot allocate GPU memory.
//
// Who that is.  pmb_cull_views works on a pmb_scene, and a pmb_scene owns device memory: mirrors, visibility bitmaps,
// the per-chunk counters of the compaction, the lists — all cudaMalloc'ed when the handle is created, resized or
// given more views.  Some hosts cannot afford that:
//   * an engine that reserves its whole GPU memory budget in its own pools at start-up: a later cudaMalloc either
//     fails or takes memory the renderer had counted on, and the entity data already sits in a buffer the engine
//     carved out of those pools;
//   * code in the middle of a frame: cudaMalloc / cudaFree synchronise the whole device and take an unbounded time,
//     which a 16 ms frame cannot absorb;
//   * a plugin or job running inside somebody else's CUDA context under a rule that it allocates nothing (a fixed
//     memory budget audited at load time).
// Such a caller already has its inputs and its output buffers on the device.  What the cull still needs is scratch —
// one visibility bitmap row per view and the chunk counters of the compaction — and here that scratch is part of the
// library image: a file-scope `static __device__` array.  Every GPU that loads the library carries it (about 2 MiB)
// from the moment the module is loaded until the process exits, whether or not this entry point is ever called.
//
// The kernels are the ones pmb_cull_views launches (cull_kernel, compact_count / scan / scatter in pm_kernels.cu),
// handed pointers into the static array instead of a handle's allocations: same bits, same list order.
//
// The name of a __device__ array is, in host code, only a handle for the runtime — not an address the GPU understands.
// The device address comes from cudaGetSymbolAddress, asked once per device with that device current; that is the
// set-up the first call on each device performs (see scratch_for_current_device).
#include "pm_kernels.h"

#include <mutex>

using namespace pm;

namespace {

constexpr uint32_t STATIC_MAX_ENTITIES = PMB_STATIC_MAX_ENTITIES;
constexpr uint32_t BITMAP_STRIDE = STATIC_MAX_ENTITIES / 32;                 // words per view row (16-byte aligned rows)
constexpr size_t   BITMAP_WORDS = (size_t)PMB_MAX_VIEWS * BITMAP_STRIDE;     // 2 MiB
constexpr uint32_t MAX_CHUNKS = 64;                                          // compact_num_chunks(STATIC_MAX_ENTITIES)
constexpr size_t   CHUNK_WORDS = (size_t)PMB_MAX_VIEWS * MAX_CHUNKS;         // 4 KiB
static_assert(STATIC_MAX_ENTITIES % 256 == 0, "the cull grid writes whole 256-entity blocks of bitmap words");

// The scratch.  Layout: [0, BITMAP_WORDS) visibility bitmaps u32[PMB_MAX_VIEWS][BITMAP_STRIDE];
// [BITMAP_WORDS, +CHUNK_WORDS) chunk counters / offsets of the compaction u32[num_views][chunks].
// Nothing in it has to start at zero: every word a call reads (bitmap words below ceil(n / 32) of the views in use,
// chunk counters below the call's chunk count) is written earlier in the same call, so the contents left by the
// previous call are never observed.
static __device__ __align__(256) uint32_t g_static_scratch[BITMAP_WORDS + CHUNK_WORDS];

constexpr int MAX_DEVICES = 64;
std::mutex    g_setup_lock;
uint32_t*     g_scratch_addr[MAX_DEVICES]; // device address of g_static_scratch on device d, once asked

// Once per device, with that device current: the device address behind the symbol.  Allocates nothing (the array came
// with the module); it may trigger the lazy load of the module on this device, which is why the first call on a
// device belongs outside any stream capture.
int scratch_for_current_device(uint32_t** out)
{
    int         dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return fail_code(e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? PMB_ERR_NO_DEVICE : PMB_ERR_CUDA,
                         "pmb_cull_views_static: no current CUDA device (%s); there is no CPU path", cudaGetErrorString(e));
    if (dev < 0 || dev >= MAX_DEVICES)
        return fail_code(PMB_ERR_INVALID, "pmb_cull_views_static: device ordinal %d outside [0, %d)", dev, MAX_DEVICES);
    std::lock_guard<std::mutex> hold(g_setup_lock);
    if (!g_scratch_addr[dev])
    {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p)
            return fail_code(PMB_ERR_CUDA, "pmb_cull_views_static: cudaGetSymbolAddress on device %d failed: %s", dev,
                             cudaGetErrorString(e));
        g_scratch_addr[dev] = static_cast<uint32_t*>(p);
    }
    *out = g_scratch_addr[dev];
    return PMB_OK;
}

bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

} // namespace

// See include/pmtech_b200.h for the contract.  No cudaMalloc, cudaFree, cudaMallocAsync or other allocating call is
// made here or in the launchers below it (launch_cull, launch_compact: kernel launches and, for an empty scene, one
// cudaMemsetAsync of the counts).
extern "C" int pmb_cull_views_static(const uint64_t* entities_device, const uint64_t* state_flags_device,
                                     const float* pos_extent_device, uint32_t num_entities, const pmb_frustum* views,
                                     uint32_t num_views, int cull_mode, uint32_t* lists_device, uint64_t list_stride,
                                     uint32_t* counts_device, void* cuda_stream)
{
    int mode;
    switch (cull_mode)
    {
    case PMB_CULL_AABB: mode = MODE_AABB; break;
    case PMB_CULL_SPHERE: mode = MODE_SPHERE; break;
    case PMB_CULL_FILTER_ONLY: mode = MODE_FILTER; break;
    default: return fail_code(PMB_ERR_INVALID, "pmb_cull_views_static: unknown cull mode %d", cull_mode);
    }
    if (num_entities > STATIC_MAX_ENTITIES)
        return fail_code(PMB_ERR_INVALID,
                         "pmb_cull_views_static: %u entities, but the built-in scratch holds PMB_STATIC_MAX_ENTITIES = %u; "
                         "split the range or use pmb_cull_views (which allocates)",
                         num_entities, STATIC_MAX_ENTITIES);
    if (num_views > PMB_MAX_VIEWS)
        return fail_code(PMB_ERR_INVALID, "pmb_cull_views_static: %u views in one call, the built-in scratch holds %u", num_views,
                         PMB_MAX_VIEWS);
    if (num_views == 0)
        return PMB_OK;
    if (!views || !counts_device)
        return fail_code(PMB_ERR_INVALID, "pmb_cull_views_static: %s is null", !views ? "views" : "counts_device");
    if (num_entities && (!entities_device || !state_flags_device || !lists_device ||
                         (cull_mode != PMB_CULL_FILTER_ONLY && !pos_extent_device)))
        return fail_code(PMB_ERR_INVALID, "pmb_cull_views_static: a device pointer is null");
    if (num_entities && list_stride < num_entities)
        return fail_code(PMB_ERR_INVALID, "pmb_cull_views_static: list_stride %llu < num_entities %u (a view may keep every entity)",
                         (unsigned long long)list_stride, num_entities);
    if (!aligned(entities_device, 8) || !aligned(state_flags_device, 8) || !aligned(pos_extent_device, 16) ||
        !aligned(lists_device, 4) || !aligned(counts_device, 4))
        return fail_code(PMB_ERR_INVALID, "pmb_cull_views_static: misaligned pointer (flags 8 B, pos_extent 16 B, lists / counts 4 B)");

    uint32_t* scratch = nullptr;
    int       r = scratch_for_current_device(&scratch);
    if (r != PMB_OK)
        return r;
    if ((size_t)compact_num_chunks(num_entities) > MAX_CHUNKS)
        return fail_code(PMB_ERR_STATE, "pmb_cull_views_static: chunk counters do not fit the built-in scratch");

    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    SceneArrays  A = {};
    A.n = num_entities;
    A.entities = const_cast<uint64_t*>(entities_device);       // read only by cull_kernel
    A.state_flags = const_cast<uint64_t*>(state_flags_device); // read only by cull_kernel
    A.pos_extent = reinterpret_cast<float4*>(const_cast<float*>(pos_extent_device));
    A.bitmap = scratch;
    A.bitmap_stride = BITMAP_STRIDE;
    ViewSet vs;
    make_view_set(&vs, views, num_views);
    launch_cull(A, vs, num_views, mode, st);
    launch_compact(scratch, BITMAP_STRIDE, num_entities, num_views, 0, scratch + BITMAP_WORDS, lists_device, (size_t)list_stride,
                   counts_device, st);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess)
        return fail_code(PMB_ERR_CUDA, "pmb_cull_views_static: launch failed: %s", cudaGetErrorString(e));
    return PMB_OK;
}
