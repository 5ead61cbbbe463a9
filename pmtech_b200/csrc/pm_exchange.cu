// pm_exchange.cu — the multi-GPU exchange step of the path (new design: the reference is one process on
// one thread).  A scene too large for one GPU is split into root-subtree ranges, one pmb_scene per GPU; loops A-C need
// no communication.  What is exchanged per frame is
//     counts[rank][view]            every rank gets the whole table (a few hundred bytes)
//     scene->renderable_extents     min / max over the shards, on every rank (exact: min / max commute)
//     the per-view draw lists       rank r's entries of view v go to their final place in the GLOBAL list of view v
//                                   (rank order = ascending entity index = the single-GPU list) — on one consuming
//                                   rank (PMB_XCHG_ROOT, the default: "gathered"), on every rank (PMB_XCHG_ALL) or
//                                   nowhere (PMB_XCHG_COUNTS: every GPU draws its own shard)
// all by NVLink peer stores from kernels of this library: no NCCL kernel competes with the streaming pass for SMs, no
// intermediate rows are received and unpacked, and nothing passes through the host.  Every rank owns one BLOCK of
// device memory {table, flags, result, global lists}; peers map it (cudaIpc between processes, peer access inside one
// process).  Per frame and rank, on the exchange's own stream behind the pack:
//     xchg_publish_kernel   my counts + extents -> every peer's table; release my READY flag at every peer
//     (stream wait)         until all ranks' READY flags have arrived here: the cross-GPU barrier.  The waits are stream
//                           memory operations (cuStreamWaitValue32 on this rank's own flags): no SM spins, so the barrier
//                           never competes with the streaming pass — nor with its cooperative launches, which do not
//                           start beside a kernel that is waiting for them (two shards on one GPU deadlocked that way)
//     xchg_push_kernel      offsets of my segments from the (now complete) table; 16-byte peer stores of my lists into
//                           the receivers' global lists; the last CTA releases my DONE flag at every receiver
//     (stream wait)         until every rank's DONE flag is here: the lists are complete (receivers), and every rank has
//                           read this epoch's table, so the set may be published into again two frames later
// A driver without stream memory operations falls back to one-warp kernels that poll the same flags.
// Two buffer sets alternate between frames, so frame f + 1 can run and exchange while frame f's lists are being read.
#include "pm_kernels.h"

#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <new>

using namespace pm;


// ------------------------------------------------------------------------------------------------
// block layout (u32 words); identical on every rank except for the size of the list area
// ------------------------------------------------------------------------------------------------
namespace {
constexpr uint32_t XP = PMB_MAX_PEERS;
constexpr uint32_t X_SLOT = 32;                      // words per (set, rank) table slot: counts[16], extents[6], pad
constexpr uint32_t X_TAB = 0;                        // [2][XP][X_SLOT]
constexpr uint32_t X_READY = X_TAB + 2 * XP * X_SLOT; // [2][XP] epoch of the last table slot that rank wrote here
constexpr uint32_t X_DONE = X_READY + 2 * XP;        // [2][XP] epoch of the last list push that rank finished here
constexpr uint32_t X_RESULT = X_DONE + 2 * XP;       // [2][64]: view offsets[16], view totals[16], extents[6], overflow, epoch
constexpr uint32_t X_CTR = X_RESULT + 2 * 64;        // [4] CTA counters of the push kernel
constexpr uint32_t X_LISTS = X_CTR + 32;             // [2][list_capacity]
static_assert((X_LISTS * 4) % 16 == 0, "the list area starts on a 16-byte boundary");

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// spins until *p has reached epoch (cyclic compare, like CU_STREAM_WAIT_VALUE_GEQ); a peer that never arrives must not
// hang the device for ever (2^27 x >= 100 ns)
__device__ __forceinline__ void wait_flag(const uint32_t* p, uint32_t epoch)
{
    for (uint32_t spin = 0; (int32_t)(ld_acquire_sys(p) - epoch) < 0; ++spin)
    {
        __nanosleep(100);
        if (spin > (1u << 27))
            __trap();
    }
}

// one warp.  counts: this shard's list lengths (pmb_scene); extents: its renderable extents (6 floats)
template <bool WAIT>
__global__ void __launch_bounds__(32) xchg_publish_kernel(const uint32_t* __restrict__ counts, const float* __restrict__ extents, uint32_t nv,
                                                         uint32_t rank, uint32_t world, uint32_t set, uint32_t epoch,
                                                         const __grid_constant__ PeerPtrs blocks, uint32_t* __restrict__ local)
{
    const uint32_t lane = threadIdx.x;
    uint32_t       val = 0;
    if (lane < PMB_MAX_VIEWS)
        val = lane < nv ? counts[lane] : 0u;
    else if (lane < PMB_MAX_VIEWS + 6)
        val = __float_as_uint(extents[lane - PMB_MAX_VIEWS]);
    for (uint32_t q = 0; q < world; ++q) // one 128-byte store per peer
        reinterpret_cast<uint32_t*>(blocks.p[q])[X_TAB + (set * XP + rank) * X_SLOT + lane] = val;
    __threadfence_system();
    __syncwarp();
    if (lane < world)
    {
        st_release_sys(reinterpret_cast<uint32_t*>(blocks.p[lane]) + X_READY + set * XP + rank, epoch);
        if (WAIT)
            wait_flag(local + X_READY + set * XP + lane, epoch);
    }
}

// grid x 256.  row = {nv counts, the nv lists back to back, global entity indices} (pmb_pack_row)
__global__ void __launch_bounds__(256) xchg_push_kernel(const uint32_t* __restrict__ row, uint32_t row_stride, uint32_t nv, uint32_t rank,
                                                       uint32_t world, uint32_t set, uint32_t epoch, uint32_t* __restrict__ local,
                                                       const __grid_constant__ PeerPtrs blocks, uint32_t receivers,
                                                       uint64_t list_capacity, uint32_t* __restrict__ host_result)
{
    __shared__ uint32_t s_src[PMB_MAX_VIEWS];   // first entry of my view-v segment in my row's body
    __shared__ uint32_t s_dst[PMB_MAX_VIEWS];   // where that segment starts in the global lists
    __shared__ uint32_t s_cnt[PMB_MAX_VIEWS];
    __shared__ uint32_t s_overflow;
    const uint32_t*     tab = local + X_TAB + set * XP * X_SLOT;
    if (threadIdx.x == 0)
    {
        uint64_t  view_off = 0;
        uint32_t  mine = 0;
        bool      overflow = false;
        uint32_t* res = local + X_RESULT + set * 64;
        for (uint32_t v = 0; v < nv; ++v)
        {
            uint64_t total = 0, before = 0;
            for (uint32_t r = 0; r < world; ++r)
            {
                const uint32_t c = tab[r * X_SLOT + v];
                if (r < rank)
                    before += c;
                total += c;
            }
            s_src[v] = mine;
            s_dst[v] = (uint32_t)(view_off + before);
            s_cnt[v] = tab[rank * X_SLOT + v];
            if (blockIdx.x == 0)
            {
                res[v] = (uint32_t)view_off;
                res[PMB_MAX_VIEWS + v] = (uint32_t)total;
            }
            mine += s_cnt[v];
            view_off += total;
        }
        for (uint32_t r = 0; r < world; ++r) // a rank whose row was truncated cannot push: every rank sees it the same way
        {
            uint64_t sum = nv;
            for (uint32_t v = 0; v < nv; ++v)
                sum += tab[r * X_SLOT + v];
            overflow |= receivers != 0u && sum > row_stride;
        }
        overflow |= (receivers != 0u && view_off > list_capacity) || view_off > 0xFFFFFFFFull;
        s_overflow = overflow ? 1u : 0u;
        if (blockIdx.x == 0)
        {
            // scene->renderable_extents over all shards (ecs_scene.cpp:1524-1525 is a min / max fold: any order, same bits)
            for (uint32_t k = 0; k < 3; ++k)
            {
                float mn = __uint_as_float(tab[PMB_MAX_VIEWS + k]), mx = __uint_as_float(tab[PMB_MAX_VIEWS + 3 + k]);
                for (uint32_t r = 1; r < world; ++r)
                {
                    const float a = __uint_as_float(tab[r * X_SLOT + PMB_MAX_VIEWS + k]);
                    const float b = __uint_as_float(tab[r * X_SLOT + PMB_MAX_VIEWS + 3 + k]);
                    mn = a < mn ? a : mn;
                    mx = b > mx ? b : mx;
                }
                res[2 * PMB_MAX_VIEWS + k] = __float_as_uint(mn);
                res[2 * PMB_MAX_VIEWS + 3 + k] = __float_as_uint(mx);
            }
            res[2 * PMB_MAX_VIEWS + 6] = s_overflow;
            res[2 * PMB_MAX_VIEWS + 7] = epoch;
        }
    }
    __syncthreads();
    if (blockIdx.x == 0)
    {
        // The host's copy of the result and of the table goes out as plain stores into mapped page-locked memory: a
        // copy-engine transfer queued behind this stream's waits would sit in front of every later copy of the
        // process on that engine (the next frame's uploads; with two shards on one GPU, the other shard's — a deadlock).
        const uint32_t* res = local + X_RESULT + set * 64;
        for (uint32_t i = threadIdx.x; i < 64 + world * X_SLOT; i += blockDim.x)
            host_result[i] = i < 64 ? res[i] : tab[i - 64];
    }
    if (!s_overflow && receivers)
    {
        const uint32_t  gtid = blockIdx.x * blockDim.x + threadIdx.x, gstride = gridDim.x * blockDim.x;
        const uint32_t* body = row + nv;
        const size_t    lists_word = X_LISTS + (size_t)set * list_capacity;
        for (uint32_t v = 0; v < nv; ++v)
        {
            const uint32_t c = s_cnt[v], s0 = s_src[v], d0 = s_dst[v];
            // the destination decides the alignment: up to 3 leading entries by 4-byte stores, then whole 16-byte stores
            const uint32_t head = min(c, (4u - (d0 & 3u)) & 3u);
            const uint32_t nvec = (c - head) >> 2, tail0 = head + (nvec << 2);
            for (uint32_t j = gtid; j < nvec; j += gstride)
            {
                const uint32_t* sp = body + s0 + head + 4u * j; // local reads, any alignment
                const uint4     val = make_uint4(sp[0], sp[1], sp[2], sp[3]);
                const size_t    at = lists_word + d0 + head + 4u * j;
                for (uint32_t k = 0; k < world; ++k) // start with my neighbour so that the ranks spread over the links
                {
                    uint32_t q = rank + 1 + k;
                    q = q >= world ? q - world : q;
                    if ((receivers >> q) & 1u)
                        *reinterpret_cast<uint4*>(reinterpret_cast<uint32_t*>(blocks.p[q]) + at) = val;
                }
            }
            const uint32_t nscalar = head + (c - tail0);
            for (uint32_t j = gtid; j < nscalar; j += gstride)
            {
                const uint32_t i = j < head ? j : tail0 + (j - head);
                const uint32_t val = body[s0 + i];
                for (uint32_t q = 0; q < world; ++q)
                    if ((receivers >> q) & 1u)
                        reinterpret_cast<uint32_t*>(blocks.p[q])[lists_word + d0 + i] = val;
            }
        }
    }
    // the last CTA to finish tells the peers that everything this rank sends for `epoch` has been written
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0)
    {
        uint32_t* ctr = local + X_CTR + set;
        if (atomicAdd(ctr, 1u) == gridDim.x - 1)
        {
            *ctr = 0u;
            __threadfence_system();
            // ... to EVERY rank, receiver or not: DONE also says "this rank has read the table of `epoch`", which is what
            // allows the peers to publish epoch + 2 into the same set
            for (uint32_t q = 0; q < world; ++q)
                st_release_sys(reinterpret_cast<uint32_t*>(blocks.p[q]) + X_DONE + set * XP + rank, epoch);
        }
    }
}

__global__ void __launch_bounds__(32) xchg_wait_kernel(const uint32_t* __restrict__ flags, uint32_t world, uint32_t set, uint32_t epoch)
{
    if (threadIdx.x < world)
        wait_flag(flags + set * XP + threadIdx.x, epoch);
}
// cuStreamWaitValue32 through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*WaitValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
bool g_force_polling = false; // pmb_exchange_use_polling_kernels: exercise the fallback where stream waits exist

WaitValue32Fn wait_value_fn()
{
    static WaitValue32Fn fn = nullptr;
    static bool          looked = false;
    if (!looked)
    {
        looked = true;
        void*                           p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (WaitValue32Fn)p;
        else
            cudaGetLastError();
    }
    return g_force_polling ? nullptr : fn;
}

// the stream goes on when flags[q] has reached epoch for every q < world
cudaError_t wait_flags(cudaStream_t st, const uint32_t* flags, uint32_t world, uint32_t epoch)
{
    WaitValue32Fn fn = wait_value_fn();
    if (!fn)
        return cudaErrorNotSupported;
    for (uint32_t q = 0; q < world; ++q)
        if (fn((CUstream)st, (CUdeviceptr)(uintptr_t)(flags + q), epoch, CU_STREAM_WAIT_VALUE_GEQ) != CUDA_SUCCESS)
            return cudaErrorUnknown;
    return cudaSuccess;
}
} // namespace

// ------------------------------------------------------------------------------------------------
// the handle
// ------------------------------------------------------------------------------------------------
struct pmb_exchange
{
    pmb_scene*   scene = nullptr;
    int          device = 0;
    uint32_t     rank = 0, world = 1, nv = 0, base = 0, root = 0;
    int          mode = PMB_XCHG_ROOT;
    uint32_t     row_stride = 0;
    uint64_t     list_capacity = 0;
    uint32_t     receivers = 0; // bit q = rank q holds global lists
    uint32_t*    block = nullptr;
    size_t       block_words = 0;
    uint32_t*    send[2] = {nullptr, nullptr};
    uint64_t     peers[PMB_MAX_PEERS] = {0};
    bool         peer_opened[PMB_MAX_PEERS] = {false};
    bool         connected = false;
    cudaStream_t comm = nullptr;
    cudaEvent_t  packed[2] = {nullptr, nullptr}, done[2] = {nullptr, nullptr}, t0[2] = {nullptr, nullptr}, t1[2] = {nullptr, nullptr};
    uint32_t*    h_result = nullptr; // page-locked, mapped: [2][64 + XP * X_SLOT], written by the push kernel
    uint32_t*    d_result = nullptr; // its device address
    uint32_t     epoch = 0;
    int          pending = 0; // phases of the frame `epoch` still to be enqueued: 1 = push + done, 2 = done
    int          last = -1;
    uint64_t     launches = 0;
};

static int xfail(int code, const char* msg)
{
    return fail_code(code, "%s", msg);
}

#define XCU(expr)                                                                                                     \
    do                                                                                                                \
    {                                                                                                                 \
        cudaError_t _e = (expr);                                                                                      \
        if (_e != cudaSuccess)                                                                                        \
            return fail_code(_e == cudaErrorMemoryAllocation ? PMB_ERR_OOM : PMB_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, \
                             cudaGetErrorString(_e), __FILE__, __LINE__);                                             \
    } while (0)

extern "C" int pmb_exchange_create(pmb_exchange** out, pmb_scene* s, uint32_t rank, uint32_t world, uint32_t num_views, uint32_t base_index,
                                   uint64_t row_capacity, uint64_t list_capacity, int mode, uint32_t root)
{
    if (!out)
        return xfail(PMB_ERR_INVALID, "out is null");
    *out = nullptr;
    ScenePeek pk;
    if (!scene_peek(s, &pk))
        return xfail(PMB_ERR_INVALID, "null scene handle");
    if (world == 0 || world > PMB_MAX_PEERS || rank >= world || root >= world || num_views == 0 || num_views > PMB_MAX_VIEWS)
        return fail_code(PMB_ERR_INVALID, "bad exchange shape (at most %u ranks, %u views)", PMB_MAX_PEERS, PMB_MAX_VIEWS);
    if (mode != PMB_XCHG_ROOT && mode != PMB_XCHG_ALL && mode != PMB_XCHG_COUNTS)
        return fail_code(PMB_ERR_INVALID, "unknown exchange mode %d", mode);
    if (row_capacity < num_views || row_capacity > 0xFFFFFFF0ull || list_capacity > 0xFFFFFFF0ull)
        return xfail(PMB_ERR_INVALID, "row / list capacity out of range");
    XCU(cudaSetDevice(pk.device));
    pmb_exchange* x = new (std::nothrow) pmb_exchange();
    if (!x)
        return xfail(PMB_ERR_OOM, "out of host memory");
    x->scene = s, x->device = pk.device, x->rank = rank, x->world = world, x->nv = num_views, x->base = base_index, x->root = root, x->mode = mode;
    x->row_stride = (uint32_t)((row_capacity + 3) & ~3ull);
    x->list_capacity = (list_capacity + 3) & ~3ull; // the same on every rank: it is part of the layout the peers address
    x->receivers = mode == PMB_XCHG_ALL ? (world >= 32 ? 0xFFFFFFFFu : ((1u << world) - 1u)) : (mode == PMB_XCHG_ROOT ? (1u << root) : 0u);
    const bool receives = (x->receivers >> rank) & 1u;
    x->block_words = X_LISTS + (receives ? 2 * (size_t)x->list_capacity : 0);
    cudaError_t e = cudaMalloc((void**)&x->block, x->block_words * 4);
    if (e == cudaSuccess)
        e = cudaMemset(x->block, 0, (size_t)X_LISTS * 4);
    for (int b = 0; b < 2 && e == cudaSuccess; ++b)
    {
        e = cudaMalloc((void**)&x->send[b], ((size_t)x->row_stride + 8) * 4); // the row, then this frame's renderable extents
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&x->packed[b], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&x->done[b], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreate(&x->t0[b]);
        if (e == cudaSuccess) e = cudaEventCreate(&x->t1[b]);
    }
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&x->comm, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaHostAlloc((void**)&x->h_result, 2 * (64 + XP * X_SLOT) * 4, cudaHostAllocMapped);
    if (e == cudaSuccess) e = cudaHostGetDevicePointer((void**)&x->d_result, x->h_result, 0);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess)
    {
        pmb_exchange_destroy(x);
        return fail_code(e == cudaErrorMemoryAllocation ? PMB_ERR_OOM : PMB_ERR_CUDA, "exchange buffers: %s", cudaGetErrorString(e));
    }
    x->peers[rank] = (uint64_t)(uintptr_t)x->block;
    if (world == 1)
        x->connected = true;
    *out = x;
    return PMB_OK;
}

extern "C" void pmb_exchange_destroy(pmb_exchange* x)
{
    if (!x)
        return;
    cudaSetDevice(x->device);
    if (x->comm)
        cudaStreamSynchronize(x->comm);
    for (uint32_t q = 0; q < PMB_MAX_PEERS; ++q)
        if (x->peer_opened[q])
            cudaIpcCloseMemHandle((void*)(uintptr_t)x->peers[q]);
    for (int b = 0; b < 2; ++b)
    {
        if (x->send[b]) cudaFree(x->send[b]);
        if (x->packed[b]) cudaEventDestroy(x->packed[b]);
        if (x->done[b]) cudaEventDestroy(x->done[b]);
        if (x->t0[b]) cudaEventDestroy(x->t0[b]);
        if (x->t1[b]) cudaEventDestroy(x->t1[b]);
    }
    if (x->block) cudaFree(x->block);
    if (x->h_result) cudaFreeHost(x->h_result);
    if (x->comm) cudaStreamDestroy(x->comm);
    delete x;
}

extern "C" int pmb_exchange_ipc_handle(pmb_exchange* x, void* handle_out)
{
    if (!x || !handle_out)
        return xfail(PMB_ERR_INVALID, "null pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) <= PMB_IPC_HANDLE_BYTES, "handle size");
    XCU(cudaSetDevice(x->device));
    cudaIpcMemHandle_t h;
    XCU(cudaIpcGetMemHandle(&h, x->block));
    memset(handle_out, 0, PMB_IPC_HANDLE_BYTES);
    memcpy(handle_out, &h, sizeof(h));
    return PMB_OK;
}

extern "C" int pmb_exchange_connect_ipc(pmb_exchange* x, const void* handles)
{
    if (!x || !handles)
        return xfail(PMB_ERR_INVALID, "null pointer");
    XCU(cudaSetDevice(x->device));
    for (uint32_t q = 0; q < x->world; ++q)
    {
        if (q == x->rank || x->peers[q])
            continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char*)handles + (size_t)q * PMB_IPC_HANDLE_BYTES, sizeof(h));
        void* p = nullptr;
        XCU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        x->peers[q] = (uint64_t)(uintptr_t)p;
        x->peer_opened[q] = true;
    }
    x->connected = true;
    return PMB_OK;
}

extern "C" int pmb_exchange_connect_local(pmb_exchange* const* all, uint32_t world)
{
    if (!all || world == 0 || world > PMB_MAX_PEERS)
        return xfail(PMB_ERR_INVALID, "bad exchange table");
    for (uint32_t r = 0; r < world; ++r)
        if (!all[r] || all[r]->world != world || all[r]->rank != r || all[r]->list_capacity != all[0]->list_capacity ||
            all[r]->mode != all[0]->mode || all[r]->root != all[0]->root || all[r]->nv != all[0]->nv)
            return xfail(PMB_ERR_INVALID, "the exchanges of one group must be created with the same shape, in rank order");
    for (uint32_t r = 0; r < world; ++r)
    {
        XCU(cudaSetDevice(all[r]->device));
        for (uint32_t q = 0; q < world; ++q)
        {
            if (all[q]->device != all[r]->device)
            {
                int can = 0;
                XCU(cudaDeviceCanAccessPeer(&can, all[r]->device, all[q]->device));
                if (!can)
                    return fail_code(PMB_ERR_CUDA, "device %d cannot map the memory of device %d", all[r]->device, all[q]->device);
                cudaError_t e = cudaDeviceEnablePeerAccess(all[q]->device, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled)
                    cudaGetLastError();
                else if (e != cudaSuccess)
                    return fail_code(PMB_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d -> %d): %s", all[r]->device, all[q]->device,
                                     cudaGetErrorString(e));
            }
            all[r]->peers[q] = (uint64_t)(uintptr_t)all[q]->block;
        }
        all[r]->connected = true;
    }
    return PMB_OK;
}

// The exchange of one frame is enqueued in three phases.  On the boxes this was developed on, a stream that waits — for a
// flag, or inside a polling kernel — holds back everything enqueued behind it on the same device, whatever the stream
// (measured: scripts/micro/streams.cu; one hardware queue per process and device).  So the ENQUEUE order has to be a valid
// serial order: a wait is only enqueued when everything that satisfies it could already have been enqueued.
//   phase 1  pack (scene stream) + publication of counts / extents            right after the frame's kernels
//   phase 2  wait for every rank's publication, push the list segments         deferred: behind the NEXT frame's kernels,
//   phase 3  wait for every rank's push                                        so the peers have a whole frame to arrive
// pmb_exchange_launch(e) = phase 1 of e, then phases 2-3 of e - 1; pmb_exchange_finish runs the pending phases of the last
// frame.  A group of ranks in one process (pmb_group) runs each phase for all its ranks before the next phase.
// What a launch can be refused for before it has touched the exchange: a caller's mistake found here leaves the previous
// frame's deferred phases pending, for the next launch / flush / finish to enqueue, and no peer waits in vain.
static int publish_precheck(pmb_exchange* x)
{
    if (!x->connected)
        return xfail(PMB_ERR_STATE, "exchange not connected to its peers");
    ScenePeek pk;
    scene_peek(x->scene, &pk);
    if (pk.last_views < x->nv)
        return fail_code(PMB_ERR_STATE, "the exchange carries %u views, the last cull built %u", x->nv, pk.last_views);
    return PMB_OK;
}

static int phase_publish(pmb_exchange* x)
{
    int rc = publish_precheck(x);
    if (rc != PMB_OK)
        return rc;
    if (x->pending)
        return xfail(PMB_ERR_STATE, "the previous exchange has not been completed");
    ScenePeek pk;
    scene_peek(x->scene, &pk);
    XCU(cudaSetDevice(x->device));
    // 0 is "never written" in the flag words; the buffer set is the epoch's low bit, so the wrap goes 0xFFFFFFFF (set 1)
    // -> 2 (set 0): the sets keep alternating and the cyclic >= of the waits still orders 2 behind 0xFFFFFFFE
    if (++x->epoch == 0)
        x->epoch = 2;
    const uint32_t b = x->epoch & 1u;
    // scene stream: the buffer set is free again; this shard's row = {lengths, lists with global indices}
    XCU(cudaStreamWaitEvent(pk.stream, x->done[b], 0));
    // (the extents slot is re-armed by the frame after next: the values travel with the row)
    // (nobody receives lists: only the header and the extents are packed)
    launch_pack_row(pk.lists, pk.list_stride, pk.counts, x->nv, x->base, x->send[b], x->receivers ? x->row_stride : x->nv, pk.stream, pk.extents,
                    x->row_stride);
    XCU(cudaEventRecord(x->packed[b], pk.stream));
    XCU(cudaStreamWaitEvent(x->comm, x->packed[b], 0));
    PeerPtrs P;
    memset(&P, 0, sizeof(P));
    for (uint32_t q = 0; q < x->world; ++q)
        P.p[q] = x->peers[q];
    xchg_publish_kernel<false><<<1, 32, 0, x->comm>>>(x->send[b], reinterpret_cast<const float*>(x->send[b] + x->row_stride), x->nv, x->rank,
                                                      x->world, b, x->epoch, P, x->block);
    XCU(cudaGetLastError());
    x->launches += 2;
    x->pending = 1;
    x->last = (int)b;
    return PMB_OK;
}

static int phase_push(pmb_exchange* x)
{
    if (x->pending != 1)
        return PMB_OK;
    XCU(cudaSetDevice(x->device));
    const uint32_t b = x->epoch & 1u;
    const bool     stream_waits = wait_value_fn() != nullptr;
    XCU(cudaEventRecord(x->t0[b], x->comm));
    if (stream_waits)
        XCU(wait_flags(x->comm, x->block + X_READY + b * XP, x->world, x->epoch));
    else
    {
        xchg_wait_kernel<<<1, 32, 0, x->comm>>>(x->block + X_READY, x->world, b, x->epoch);
        x->launches += 1;
    }
    PeerPtrs P;
    memset(&P, 0, sizeof(P));
    for (uint32_t q = 0; q < x->world; ++q)
        P.p[q] = x->peers[q];
    uint32_t grid = 1;
    if (x->receivers)
    {
        grid = (x->row_stride + 256 * 16 - 1) / (256 * 16);
        grid = grid < 1 ? 1 : (grid > 148 ? 148 : grid);
    }
    xchg_push_kernel<<<grid, 256, 0, x->comm>>>(x->send[b], x->row_stride, x->nv, x->rank, x->world, b, x->epoch, x->block, P, x->receivers,
                                                x->list_capacity, x->d_result + b * (64 + XP * X_SLOT));
    XCU(cudaGetLastError());
    x->launches += 1;
    x->pending = 2;
    return PMB_OK;
}

static int phase_done(pmb_exchange* x)
{
    if (x->pending != 2)
        return PMB_OK;
    XCU(cudaSetDevice(x->device));
    const uint32_t b = x->epoch & 1u;
    // Every rank, not only the receivers of lists: a rank that went on without it could publish epoch + 2 into this set
    // while a slower peer is still waiting for — or reading — the table of this epoch (seen at 8 GPUs: the rank that
    // receives everybody's lists falls a frame behind, finds epoch + 2 where it waits for epoch, and waits for ever).
    if (wait_value_fn())
        XCU(wait_flags(x->comm, x->block + X_DONE + b * XP, x->world, x->epoch));
    else
    {
        xchg_wait_kernel<<<1, 32, 0, x->comm>>>(x->block + X_DONE, x->world, b, x->epoch);
        x->launches += 1;
    }
    XCU(cudaEventRecord(x->t1[b], x->comm));
    XCU(cudaEventRecord(x->done[b], x->comm));
    XCU(cudaGetLastError());
    x->pending = 0;
    return PMB_OK;
}

extern "C" int pmb_exchange_launch(pmb_exchange* x)
{
    if (!x)
        return xfail(PMB_ERR_INVALID, "null exchange");
    // this frame's publication goes out first; the previous frame's waits line up behind it
    int rc = publish_precheck(x); // nothing has been touched yet: the previous frame's phases stay pending
    if (rc != PMB_OK)
        return rc;
    const bool     had = x->pending != 0;
    const uint32_t prev_epoch = x->epoch;
    const int      prev_pending = x->pending, prev_last = x->last;
    // finish the bookkeeping of the previous frame under its own epoch, but enqueue it behind the new publication:
    // the publication only needs the scene stream's pack and the (other) buffer set
    x->pending = 0;
    rc = phase_publish(x);
    if (rc != PMB_OK)
    {
        // a CUDA call failed half way: put the state back and still enqueue what the peers are waiting for (the previous
        // frame's push and DONE flag).  This frame was not published, so a peer that launched it waits for this rank: the
        // group is poisoned and the hosts must tear it down (the error code is the signal; stream waits have no timeout).
        x->epoch = prev_epoch, x->pending = prev_pending, x->last = prev_last;
        if (had && phase_push(x) == PMB_OK)
            phase_done(x);
        return rc;
    }
    if (!had)
        return PMB_OK;
    const uint32_t cur_epoch = x->epoch;
    const int      cur_last = x->last;
    x->epoch = prev_epoch, x->pending = prev_pending;
    rc = phase_push(x);
    if (rc == PMB_OK)
        rc = phase_done(x);
    x->epoch = cur_epoch, x->pending = 1, x->last = cur_last;
    return rc;
}

extern "C" int pmb_exchange_flush(pmb_exchange* x)
{
    if (!x)
        return xfail(PMB_ERR_INVALID, "null exchange");
    int rc = phase_push(x);
    return rc == PMB_OK ? phase_done(x) : rc;
}

extern "C" int pmb_exchange_finish(pmb_exchange* x, pmb_exchange_result* out)
{
    if (!x || !out)
        return xfail(PMB_ERR_INVALID, "null pointer");
    if (x->last < 0)
        return xfail(PMB_ERR_STATE, "no exchange launched");
    int rc = phase_push(x);
    if (rc == PMB_OK)
        rc = phase_done(x);
    if (rc != PMB_OK)
        return rc;
    XCU(cudaSetDevice(x->device));
    const uint32_t b = (uint32_t)x->last;
    XCU(cudaEventSynchronize(x->done[b]));
    const uint32_t* h = x->h_result + b * (64 + XP * X_SLOT);
    memset(out, 0, sizeof(*out));
    out->num_views = x->nv;
    out->world = x->world;
    out->overflow = h[2 * PMB_MAX_VIEWS + 6];
    out->epoch = h[2 * PMB_MAX_VIEWS + 7];
    for (uint32_t v = 0; v < x->nv; ++v)
    {
        out->view_offset[v] = h[v];
        out->view_total[v] = h[PMB_MAX_VIEWS + v];
        for (uint32_t r = 0; r < x->world; ++r)
            out->counts[r][v] = h[64 + r * X_SLOT + v];
    }
    memcpy(out->renderable_extents, h + 2 * PMB_MAX_VIEWS, 6 * sizeof(float));
    out->lists_device = ((x->receivers >> x->rank) & 1u) ? x->block + X_LISTS + (size_t)b * x->list_capacity : nullptr;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, x->t0[b], x->t1[b]) == cudaSuccess)
        out->device_ms = ms;
    else
        cudaGetLastError();
    return PMB_OK;
}

extern "C" int64_t pmb_exchange_get_list(pmb_exchange* x, uint32_t view, uint32_t* host_out, uint32_t capacity)
{
    if (!x || !host_out)
        return xfail(PMB_ERR_INVALID, "null pointer");
    pmb_exchange_result r;
    int                 rc = pmb_exchange_finish(x, &r);
    if (rc != PMB_OK)
        return rc;
    if (view >= r.num_views)
        return fail_code(PMB_ERR_INVALID, "view %u requested, the exchange carries %u", view, r.num_views);
    if (r.overflow)
        return xfail(PMB_ERR_STATE, "the last exchange overflowed its buffers: nothing was gathered");
    if (!r.lists_device)
        return xfail(PMB_ERR_STATE, "this rank does not hold the global lists");
    if (r.view_total[view] > capacity)
        return fail_code(PMB_ERR_INVALID, "global list of view %u has %u entries, caller capacity is %u", view, r.view_total[view], capacity);
    if (r.view_total[view])
    {
        XCU(cudaMemcpyAsync(host_out, r.lists_device + r.view_offset[view], (size_t)r.view_total[view] * 4, cudaMemcpyDeviceToHost, x->comm));
        XCU(cudaStreamSynchronize(x->comm));
    }
    return (int64_t)r.view_total[view];
}

extern "C" void pmb_exchange_use_polling_kernels(int on) { g_force_polling = on != 0; }

// diagnostic: the first `words` u32 of this rank's block (table, flags, result), read beside whatever the streams are doing
extern "C" int pmb_exchange_debug_header(pmb_exchange* x, uint32_t* host_out, uint32_t words)
{
    if (!x || !host_out || words > X_LISTS)
        return xfail(PMB_ERR_INVALID, "bad header request");
    XCU(cudaSetDevice(x->device));
    cudaStream_t st;
    XCU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    cudaError_t e = cudaMemcpyAsync(host_out, x->block, (size_t)words * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(st);
    cudaStreamDestroy(st);
    XCU(e);
    return PMB_OK;
}

extern "C" void* pmb_exchange_stream(pmb_exchange* x) { return x ? (void*)x->comm : nullptr; }
extern "C" uint64_t pmb_exchange_launches(const pmb_exchange* x) { return x ? x->launches : 0; }

// ------------------------------------------------------------------------------------------------
// one process, several GPUs: pmtech is one C++ process on one user thread (examples/code/example_common.h:107-127), so
// the natural multi-GPU form is a group inside the library: peer access between the devices, one pmb_scene and one
// exchange per GPU, every call enqueues on all of them without waiting in between (the cross-GPU barrier lives in the
// kernels; the host never blocks between two ranks' launches).
// ------------------------------------------------------------------------------------------------
struct pmb_group
{
    uint32_t      world = 0;
    pmb_scene*    scenes[PMB_MAX_PEERS] = {nullptr};
    pmb_exchange* xchg[PMB_MAX_PEERS] = {nullptr};
    uint32_t      nv = 0;
};

extern "C" void pmb_group_destroy(pmb_group* g)
{
    if (!g)
        return;
    for (uint32_t r = 0; r < g->world; ++r)
        pmb_exchange_destroy(g->xchg[r]);
    for (uint32_t r = 0; r < g->world; ++r)
        pmb_scene_destroy(g->scenes[r]);
    delete g;
}

extern "C" int pmb_group_create(pmb_group** out, const int* devices, uint32_t num_devices, const uint32_t* capacities)
{
    if (!out || !devices || !capacities || num_devices == 0 || num_devices > PMB_MAX_PEERS)
        return fail_code(PMB_ERR_INVALID, "a group holds 1..%u devices", PMB_MAX_PEERS);
    *out = nullptr;
    pmb_group* g = new (std::nothrow) pmb_group();
    if (!g)
        return xfail(PMB_ERR_OOM, "out of host memory");
    g->world = num_devices;
    for (uint32_t r = 0; r < num_devices; ++r)
    {
        int rc = pmb_scene_create(&g->scenes[r], capacities[r], devices[r]);
        if (rc != PMB_OK)
        {
            pmb_group_destroy(g);
            return rc;
        }
    }
    *out = g;
    return PMB_OK;
}

extern "C" uint32_t   pmb_group_size(const pmb_group* g) { return g ? g->world : 0; }
extern "C" pmb_scene* pmb_group_scene(pmb_group* g, uint32_t rank) { return (g && rank < g->world) ? g->scenes[rank] : nullptr; }

extern "C" int pmb_group_exchange_setup(pmb_group* g, uint32_t num_views, const uint32_t* base_index, uint64_t row_capacity,
                                        uint64_t list_capacity, int mode, uint32_t root)
{
    if (!g || !base_index)
        return xfail(PMB_ERR_INVALID, "null pointer");
    for (uint32_t r = 0; r < g->world; ++r)
    {
        pmb_exchange_destroy(g->xchg[r]);
        g->xchg[r] = nullptr;
    }
    for (uint32_t r = 0; r < g->world; ++r)
    {
        int rc = pmb_exchange_create(&g->xchg[r], g->scenes[r], r, g->world, num_views, base_index[r], row_capacity, list_capacity, mode, root);
        if (rc != PMB_OK)
            return rc;
    }
    g->nv = num_views;
    return pmb_exchange_connect_local(g->xchg, g->world);
}

extern "C" int pmb_group_frame(pmb_group* g, const uint32_t* num_entities, const pmb_frustum* views, uint32_t num_views, int cull_mode)
{
    if (!g || !num_entities)
        return xfail(PMB_ERR_INVALID, "null pointer");
    if (!g->xchg[0] || num_views != g->nv)
        return xfail(PMB_ERR_STATE, "pmb_group_exchange_setup has not been called for this number of views");
    // Everything below only enqueues: no rank waits on the host for another one.  Phase by phase over all ranks, so that
    // the order of enqueueing is a valid order of execution even where ranks share a device (see phase_publish).
    int rc = PMB_OK;
    uint32_t prev_epoch[PMB_MAX_PEERS];
    int      prev_pending[PMB_MAX_PEERS];
    for (uint32_t r = 0; r < g->world && rc == PMB_OK; ++r)
        rc = pmb_frame(g->scenes[r], num_entities[r], views, num_views, cull_mode);
    for (uint32_t r = 0; r < g->world && rc == PMB_OK; ++r)
        rc = publish_precheck(g->xchg[r]);
    if (rc != PMB_OK)
        return rc; // no exchange has been touched: the previous frame's phases stay pending on every rank
    bool published[PMB_MAX_PEERS] = {};
    int  prev_last[PMB_MAX_PEERS];
    for (uint32_t r = 0; r < g->world; ++r)
        prev_epoch[r] = g->xchg[r]->epoch, prev_pending[r] = g->xchg[r]->pending, prev_last[r] = g->xchg[r]->last;
    for (uint32_t r = 0; r < g->world && rc == PMB_OK; ++r)
    {
        g->xchg[r]->pending = 0;
        rc = phase_publish(g->xchg[r]);
        published[r] = rc == PMB_OK;
    }
    // the previous frame's phases, under its epoch — also when a publication failed on the way (a CUDA error): the ranks
    // that did not publish go back to where they were, and every rank still enqueues the push and the DONE flag its
    // peers wait for.  The frame itself is then lost and the group poisoned (ranks that published wait for those that
    // did not): the error code tells the host to tear it down.
    uint32_t cur_epoch[PMB_MAX_PEERS];
    int      cur_last[PMB_MAX_PEERS];
    for (uint32_t r = 0; r < g->world; ++r)
    {
        pmb_exchange* x = g->xchg[r];
        cur_epoch[r] = x->epoch, cur_last[r] = x->last;
        x->epoch = prev_epoch[r], x->pending = prev_pending[r], x->last = prev_last[r];
    }
    int rc2 = PMB_OK;
    for (uint32_t r = 0; r < g->world && rc2 == PMB_OK; ++r)
        rc2 = phase_push(g->xchg[r]);
    for (uint32_t r = 0; r < g->world && rc2 == PMB_OK; ++r)
        rc2 = phase_done(g->xchg[r]);
    for (uint32_t r = 0; r < g->world; ++r)
        if (published[r])
            g->xchg[r]->epoch = cur_epoch[r], g->xchg[r]->pending = 1, g->xchg[r]->last = cur_last[r];
    return rc != PMB_OK ? rc : rc2;
}

extern "C" int pmb_group_finish(pmb_group* g, uint32_t rank, pmb_exchange_result* out)
{
    if (!g || rank >= g->world || !g->xchg[rank])
        return xfail(PMB_ERR_INVALID, "bad rank or no exchange");
    int rc = PMB_OK;
    for (uint32_t r = 0; r < g->world && rc == PMB_OK; ++r)
        rc = phase_push(g->xchg[r]);
    for (uint32_t r = 0; r < g->world && rc == PMB_OK; ++r)
        rc = phase_done(g->xchg[r]);
    for (uint32_t r = 0; r < g->world && rc == PMB_OK; ++r) // every rank's part of the frame is over before the caller reads anything
    {
        pmb_exchange_result tmp;
        rc = pmb_exchange_finish(g->xchg[r], r == rank ? out : &tmp);
    }
    return rc;
}

extern "C" pmb_exchange* pmb_group_exchange(pmb_group* g, uint32_t rank) { return (g && rank < g->world) ? g->xchg[rank] : nullptr; }
