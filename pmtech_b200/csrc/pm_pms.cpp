// .pms scene images.  Reader for the layout written by save_scene
// (core/put/source/ecs/ecs_scene.cpp:2162-2346) as load_scene walks it (:2350-2535); only the component
// arrays of the scene-update path are located, the renderer-resource specialisations that follow are left
// to the engine.  Host code only.
#include "../../include/pmtech_b200.h"

#include <cstring>
#include <vector>

int pmb_fail_invalid(const char* fmt, ...); // pm_capi.cu: records the message for pmb_last_error()

namespace
{
    struct Cursor
    {
        const uint8_t* p;
        uint64_t       len, off;
        bool           take(void* dst, uint64_t n)
        {
            if (n > len - off)
                return false;
            if (dst)
                memcpy(dst, p + off, n);
            off += n;
            return true;
        }
    };

    // index in ecs_scene.h:471-513, element size this library mirrors
    struct PathComponent
    {
        uint32_t index, size;
    };
    const PathComponent k_path[PMB_PMS_COMPONENT_COUNT] = {
        {0, 8}, {1, 8}, {8, 4}, {9, 40}, {10, 64}, {11, 64}, {14, 52}, {22, 4}, {23, 160}, {38, 32},
    };
    constexpr uint64_t k_hash_id = 4;  // hash_id = u32 (core/pen/include/hash.h)
    constexpr uint64_t k_camera = k_hash_id + 12 + 12 + 8 + 5 * 4; // id, pos, focus, rot, fov, aspect, near, far, zoom (:2449-2457)
} // namespace

extern "C" int pmb_pms_parse(const void* data, uint64_t len, pmb_pms_info* out)
{
    if (!data || !out)
        return pmb_fail_invalid("pmb_pms_parse: null argument");
    memset(out, 0, sizeof(*out));
    Cursor  c{(const uint8_t*)data, len, 0};
    int32_t h[64]; // scene_header, :2031-2044
    if (!c.take(h, sizeof(h)))
        return pmb_fail_invalid("pms: image shorter than the scene header (%llu bytes)", (unsigned long long)len);
    if (h[0] != (int32_t)sizeof(h))
        return pmb_fail_invalid("pms: header_size %d, expected %d", h[0], (int)sizeof(h));
    out->version = h[1];
    out->num_nodes = (uint32_t)h[2];
    out->num_components = h[3];
    out->num_lookup_strings = h[4];
    out->num_extensions = h[5];
    out->num_base_components = h[1] < 9 ? h[3] : h[6]; // version 9 adds extensions (:2372-2373)
    out->view_flags = (uint32_t)h[32];
    out->selected_index = h[33];
    if (h[3] < 0 || h[4] < 0 || h[5] < 0 || h[3] > 4096)
        return pmb_fail_invalid("pms: negative or absurd counts in the header");

    out->sizes_offset = c.off;
    std::vector<uint32_t> sizes((size_t)h[3]);
    if (!c.take(sizes.data(), sizes.size() * 4))
        return pmb_fail_invalid("pms: truncated component size table");
    if (!c.take(nullptr, (uint64_t)h[5] * (k_hash_id + 8))) // ext_components {hash_id, start_cmp, num_cmp} (:2412-2428)
        return pmb_fail_invalid("pms: truncated extension table");
    out->strings_offset = c.off;
    for (int32_t i = 0; i < h[4]; ++i) // read_parsable_string + hash_id (:2434-2441)
    {
        uint32_t n = 0;
        if (!c.take(&n, 4) || !c.take(nullptr, (uint64_t)n + k_hash_id))
            return pmb_fail_invalid("pms: truncated lookup string %d", i);
    }
    if (!c.take(&out->num_cameras, 4))
        return pmb_fail_invalid("pms: truncated camera count");
    out->cameras_offset = c.off;
    if (!c.take(nullptr, (uint64_t)out->num_cameras * k_camera))
        return pmb_fail_invalid("pms: truncated camera block");

    out->data_offset = c.off;
    for (int32_t i = 0; i < h[3]; ++i) // :2479-2527, arrays back to back
    {
        // Base components keep their index; extension components (i >= num_base_components) are remapped by
        // the engine through its registered extensions (:2485-2499) and never belong to this path.
        if (i < out->num_base_components)
            for (int k = 0; k < PMB_PMS_COMPONENT_COUNT; ++k)
                if (k_path[k].index == (uint32_t)i && k_path[k].size == sizes[(size_t)i])
                    out->component_offset[k] = c.off;
        if (!c.take(nullptr, (uint64_t)sizes[(size_t)i] * out->num_nodes))
            return pmb_fail_invalid("pms: component %d (%u B x %u nodes) runs past the end of the image", i, sizes[(size_t)i],
                                    out->num_nodes);
    }
    out->names_offset = c.off;
    return PMB_OK;
}

extern "C" int pmb_load_pms(pmb_scene* s, const void* data, uint64_t len, uint32_t zero_offset, pmb_pms_info* info_out)
{
    pmb_pms_info info;
    int          r = pmb_pms_parse(data, len, &info);
    if (r != PMB_OK)
        return r;
    if (info_out)
        *info_out = info;
    if (!s)
        return pmb_fail_invalid("pmb_load_pms: null scene");
    const uint32_t n = info.num_nodes;
    if ((uint64_t)zero_offset + n > pmb_scene_capacity(s))
        return pmb_fail_invalid("pms: %u nodes at offset %u exceed the scene capacity %u", n, zero_offset, pmb_scene_capacity(s));
    const uint8_t* base = (const uint8_t*)data;
    static const int k_array[PMB_PMS_COMPONENT_COUNT] = {PMB_ENTITIES, PMB_STATE_FLAGS, PMB_PARENTS, PMB_TRANSFORMS, PMB_LOCAL_MATRICES,
                                                         PMB_WORLD_MATRICES, PMB_BOUNDING_VOLUMES, PMB_CBUFFER, PMB_DRAW_CALLS,
                                                         PMB_POS_EXTENT};
    for (int k = 0; k < PMB_PMS_COMPONENT_COUNT; ++k)
    {
        if (!info.component_offset[k])
            continue;
        const void*           src = base + info.component_offset[k];
        std::vector<uint32_t> fixed;
        if (k == PMB_PMS_PARENTS && zero_offset) // fixup parents for scene import / merge (:2529-2531)
        {
            fixed.resize(n);
            memcpy(fixed.data(), src, (size_t)n * 4);
            for (uint32_t& p : fixed)
                p += zero_offset;
            src = fixed.data();
        }
        r = pmb_upload(s, k_array[k], src, zero_offset, n);
        if (r != PMB_OK)
            return r;
        if (!fixed.empty()) // pageable source: the copy was staged before pmb_upload returned? make it certain
        {
            r = pmb_synchronize(s);
            if (r != PMB_OK)
                return r;
        }
    }
    return PMB_OK;
}
