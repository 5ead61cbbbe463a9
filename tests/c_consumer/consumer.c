/* A C99 consumer of include/pmtech_b200.h — what a pmtech-side shim does, without Python in between.
 *
 *   consumer <scene.pms>
 *
 * Walks the scene image (pmb_pms_parse, host only), then — when a CUDA device is present — loads it into
 * device mirrors (pmb_load_pms), runs one frame with a perspective camera built by the library's producers
 * plus one directional-light shadow view fitted on the device, and prints what a renderer would consume.
 * Without a device it reports the refusal and exits 3: there is no CPU path to fall back to.
 * Output is line-oriented `key value...` so that tests/test_c_consumer.py can compare it with the oracle. */
#include "pmtech_b200.h"

#include <stdio.h>
#include <stdlib.h>

static int die(const char* what, int rc)
{
    fprintf(stderr, "%s failed: %d (%s)\n", what, rc, pmb_last_error());
    return 1;
}

int main(int argc, char** argv)
{
    if (argc < 2)
    {
        fprintf(stderr, "usage: %s scene.pms\n", argv[0]);
        return 2;
    }
    FILE* f = fopen(argv[1], "rb");
    if (!f)
    {
        perror(argv[1]);
        return 2;
    }
    fseek(f, 0, SEEK_END);
    long len = ftell(f);
    fseek(f, 0, SEEK_SET);
    unsigned char* image = (unsigned char*)malloc((size_t)len);
    if (!image || fread(image, 1, (size_t)len, f) != (size_t)len)
    {
        fprintf(stderr, "short read\n");
        return 2;
    }
    fclose(f);

    pmb_pms_info info;
    int          rc = pmb_pms_parse(image, (uint64_t)len, &info);
    if (rc != PMB_OK)
        return die("pmb_pms_parse", rc);
    printf("version %d\nnodes %u\ncomponents %d\ncameras %u\n", info.version, info.num_nodes, info.num_components, info.num_cameras);
    printf("offsets");
    for (int k = 0; k < PMB_PMS_COMPONENT_COUNT; ++k)
        printf(" %llu", (unsigned long long)info.component_offset[k]);
    printf("\n");

    /* main camera: the library's host-side producers (camera.cpp:19-49, :236-262, :131-172) */
    pmb_camera  cam;
    const float pos[3] = {-14.5f, 16.5f, 26.5f}, at[3] = {0.0f, 0.0f, 0.0f}, up[3] = {0.0f, 1.0f, 0.0f};
    pmb_camera_create_perspective(&cam, 60.0f, 16.0f / 9.0f, 0.1f, 1000.0f);
    pmb_camera_look_at(&cam, pos, at, up);
    pmb_camera_update_frustum(&cam);
    printf("plane0 %.9g %.9g %.9g\n", cam.camera_frustum.n[0][0], cam.camera_frustum.n[0][1], cam.camera_frustum.n[0][2]);

    if (pmb_device_count() <= 0)
    {
        pmb_scene* none = NULL;
        rc = pmb_scene_create(&none, info.num_nodes, 0);
        printf("no_device %d %s\n", rc, pmb_last_error());
        free(image);
        return rc == PMB_ERR_NO_DEVICE ? 3 : 1;
    }

    pmb_scene* dev = NULL;
    uint32_t   n = info.num_nodes;
    if ((rc = pmb_scene_create(&dev, n, 0)) != PMB_OK)
        return die("pmb_scene_create", rc);
    if ((rc = pmb_load_pms(dev, image, (uint64_t)len, 0, NULL)) != PMB_OK)
        return die("pmb_load_pms", rc);
    if ((rc = pmb_reserve_views(dev, 2)) != PMB_OK)
        return die("pmb_reserve_views", rc);
    if ((rc = pmb_frame(dev, n, &cam.camera_frustum, 1, PMB_CULL_AABB)) != PMB_OK)
        return die("pmb_frame", rc);
    const float light_direction[3] = {0.3f, -1.0f, 0.2f};
    if ((rc = pmb_cull_shadow_views(dev, n, light_direction, 1, NULL, PMB_CULL_AABB, 1)) != PMB_OK)
        return die("pmb_cull_shadow_views", rc);

    float ext[6];
    if ((rc = pmb_get_renderable_extents(dev, ext)) != PMB_OK)
        return die("pmb_get_renderable_extents", rc);
    printf("extents %.9g %.9g %.9g %.9g %.9g %.9g\n", ext[0], ext[1], ext[2], ext[3], ext[4], ext[5]);
    uint32_t counts[2] = {0, 0};
    if ((rc = pmb_get_counts(dev, counts, 2)) != PMB_OK)
        return die("pmb_get_counts", rc);
    for (uint32_t v = 0; v < 2; ++v)
    {
        uint32_t* list = (uint32_t*)malloc(sizeof(uint32_t) * (counts[v] ? counts[v] : 1));
        if (counts[v] && pmb_get_list(dev, v, list, counts[v]) != (int64_t)counts[v]) /* returns the entries copied */
            return die("pmb_get_list", -1);
        printf("list%u", v);
        for (uint32_t i = 0; i < counts[v]; ++i)
            printf(" %u", list[i]);
        printf("\n");
        free(list);
    }
    float* world = (float*)malloc((size_t)n * 64);
    if ((rc = pmb_download(dev, PMB_WORLD_MATRICES, world, 0, n)) != PMB_OK)
        return die("pmb_download", rc);
    for (uint32_t e = 0; e < n; ++e)
        printf("world%u %.9g %.9g %.9g\n", e, world[e * 16 + 3], world[e * 16 + 7], world[e * 16 + 11]);
    free(world);
    pmb_scene_destroy(dev);
    free(image);
    return 0;
}
