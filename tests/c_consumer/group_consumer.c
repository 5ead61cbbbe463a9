/* One C process driving several shards of one scene through pmb_group (include/pmtech_b200.h): the form a pmtech host
 * would use for a scene too large for one GPU — no Python, no torch, no NCCL between the caller and the library.
 *
 *   group_consumer <ranks> <entities> [mode: 0 root | 1 all | 2 counts]
 *
 * Builds a synthetic scene (a grid of boxes; every fourth entity the child of the one before it), cuts it into <ranks>
 * root-aligned ranges, gives range r to rank r (device r mod the number of CUDA devices: on a one-GPU box the ranks are
 * shards on the same GPU and run the same protocol), runs two frames of update + 2-view cull + exchange and prints
 * the gathered result.  The output must not depend on <ranks>: tests/test_c_consumer.py compares 2 and 3 ranks with 1. */
#include "pmtech_b200.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define ALLOCATED 1ull
#define GEOMETRY 2ull
#define MATERIAL 16ull

static int die(const char* what, int rc)
{
    fprintf(stderr, "%s failed: %d (%s)\n", what, rc, pmb_last_error());
    return 1;
}

int main(int argc, char** argv)
{
    if (argc < 3)
    {
        fprintf(stderr, "usage: %s ranks entities [mode]\n", argv[0]);
        return 2;
    }
    const uint32_t world = (uint32_t)atoi(argv[1]);
    const uint32_t n = (uint32_t)atoi(argv[2]);
    const int      mode = argc > 3 ? atoi(argv[3]) : PMB_XCHG_ROOT;
    const int      ndev = pmb_device_count();
    if (ndev < 1)
    {
        printf("no_device\n");
        return 3;
    }
    if (world < 1 || world > PMB_MAX_PEERS || n < 4 * world)
        return 2;

    /* the scene: parents, local matrices, flags, bounds (reference layouts) */
    uint32_t* parents = (uint32_t*)malloc(sizeof(uint32_t) * n);
    float*    local = (float*)calloc((size_t)n * 16, sizeof(float));
    uint64_t* entities = (uint64_t*)malloc(sizeof(uint64_t) * n);
    uint64_t* state = (uint64_t*)calloc(n, sizeof(uint64_t));
    float*    bv = (float*)calloc((size_t)n * 13, sizeof(float));
    for (uint32_t i = 0; i < n; ++i)
    {
        const int child = (i % 4u) == 3u;
        parents[i] = child ? i - 1 : i;
        float* m = local + (size_t)i * 16;
        m[0] = m[5] = m[10] = m[15] = 1.0f;
        m[3] = child ? 0.5f : (float)(i % 251u) * 3.0f - 370.0f;
        m[7] = child ? 0.25f : (float)((i / 251u) % 13u) * 2.0f - 12.0f;
        m[11] = child ? -0.5f : (float)(i / 3263u) * 3.0f - 300.0f;
        entities[i] = ALLOCATED | GEOMETRY | ((i % 7u) ? MATERIAL : 0ull);
        for (int k = 0; k < 3; ++k)
        {
            bv[(size_t)i * 13 + k] = -0.5f;
            bv[(size_t)i * 13 + 3 + k] = 0.5f;
        }
    }
    /* root-aligned cuts: multiples of 4 (every group of four is closed under parent links) */
    uint32_t first[PMB_MAX_PEERS + 1], cap[PMB_MAX_PEERS], cnt[PMB_MAX_PEERS];
    int      devices[PMB_MAX_PEERS];
    for (uint32_t r = 0; r <= world; ++r)
        first[r] = r == world ? n : (uint32_t)(((uint64_t)n * r / world) & ~3ull);
    for (uint32_t r = 0; r < world; ++r)
    {
        cnt[r] = first[r + 1] - first[r];
        cap[r] = cnt[r];
        devices[r] = (int)(r % (uint32_t)ndev);
    }
    pmb_group* g = NULL;
    int        rc = pmb_group_create(&g, devices, world, cap);
    if (rc != PMB_OK)
        return die("pmb_group_create", rc);
    for (uint32_t r = 0; r < world; ++r)
    {
        pmb_scene* s = pmb_group_scene(g, r);
        uint32_t*  p = (uint32_t*)malloc(sizeof(uint32_t) * cnt[r]);
        for (uint32_t i = 0; i < cnt[r]; ++i)
            p[i] = parents[first[r] + i] - first[r]; /* rebased to the shard */
        if ((rc = pmb_upload(s, PMB_PARENTS, p, 0, cnt[r])) != PMB_OK) return die("upload parents", rc);
        if ((rc = pmb_upload(s, PMB_LOCAL_MATRICES, local + (size_t)first[r] * 16, 0, cnt[r])) != PMB_OK) return die("upload local", rc);
        if ((rc = pmb_upload(s, PMB_ENTITIES, entities + first[r], 0, cnt[r])) != PMB_OK) return die("upload entities", rc);
        if ((rc = pmb_upload(s, PMB_STATE_FLAGS, state + first[r], 0, cnt[r])) != PMB_OK) return die("upload state", rc);
        if ((rc = pmb_upload(s, PMB_BOUNDING_VOLUMES, bv + (size_t)first[r] * 13, 0, cnt[r])) != PMB_OK) return die("upload bounds", rc);
        if ((rc = pmb_synchronize(s)) != PMB_OK) return die("synchronize", rc);
        free(p);
    }
    /* two views from the library's producers */
    pmb_camera  cam[2];
    pmb_frustum views[2];
    const float pos[3] = {-40.0f, 30.0f, 80.0f}, at[3] = {0.0f, 0.0f, -100.0f}, up[3] = {0.0f, 1.0f, 0.0f};
    pmb_camera_create_perspective(&cam[0], 60.0f, 16.0f / 9.0f, 0.1f, 400.0f);
    pmb_camera_look_at(&cam[0], pos, at, up);
    pmb_camera_update_frustum(&cam[0]);
    const float dir[3] = {0.0f, 1.0f, 0.0f}, mn[3] = {-100.0f, -20.0f, -200.0f}, mx[3] = {100.0f, 20.0f, -50.0f};
    memset(&cam[1], 0, sizeof(cam[1]));
    pmb_camera_update_shadow_frustum(&cam[1], dir, mn, mx);
    views[0] = cam[0].camera_frustum;
    views[1] = cam[1].camera_frustum;

    rc = pmb_group_exchange_setup(g, 2, first, (uint64_t)2 * n / world + 4096, (uint64_t)2 * n + 4096, mode, 0);
    if (rc != PMB_OK)
        return die("pmb_group_exchange_setup", rc);
    pmb_exchange_result res;
    for (int frame = 0; frame < 2; ++frame) /* both buffer sets */
    {
        if ((rc = pmb_group_frame(g, cnt, views, 2, PMB_CULL_AABB)) != PMB_OK) return die("pmb_group_frame", rc);
        if ((rc = pmb_group_finish(g, 0, &res)) != PMB_OK) return die("pmb_group_finish", rc);
    }
    printf("overflow %u\n", res.overflow);
    printf("totals %u %u\n", res.view_total[0], res.view_total[1]);
    printf("extents %.9g %.9g %.9g %.9g %.9g %.9g\n", res.renderable_extents[0], res.renderable_extents[1], res.renderable_extents[2],
           res.renderable_extents[3], res.renderable_extents[4], res.renderable_extents[5]);
    uint32_t sum[2] = {0, 0};
    for (uint32_t r = 0; r < world; ++r)
        sum[0] += res.counts[r][0], sum[1] += res.counts[r][1];
    printf("count_sums %u %u\n", sum[0], sum[1]);
    if (mode != PMB_XCHG_COUNTS)
    {
        uint32_t* list = (uint32_t*)malloc(sizeof(uint32_t) * (n + 1));
        for (uint32_t v = 0; v < 2; ++v)
        {
            const int64_t c = pmb_exchange_get_list(pmb_group_exchange(g, 0), v, list, n);
            if (c < 0)
                return die("pmb_exchange_get_list", (int)c);
            /* FNV-1a over the entries + the first and last few: enough to compare two runs exactly */
            uint64_t h = 1469598103934665603ull;
            int      ascending = 1;
            for (int64_t i = 0; i < c; ++i)
            {
                h = (h ^ list[i]) * 1099511628211ull;
                if (i && list[i] <= list[i - 1])
                    ascending = 0;
            }
            printf("list%u %lld %llu %d", v, (long long)c, (unsigned long long)h, ascending);
            for (int64_t i = 0; i < c && i < 4; ++i)
                printf(" %u", list[i]);
            if (c)
                printf(" %u", list[c - 1]);
            printf("\n");
        }
        free(list);
    }
    pmb_group_destroy(g);
    free(parents), free(local), free(entities), free(state), free(bv);
    return 0;
}
