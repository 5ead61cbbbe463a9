"""Synthetic scenes and view sets for the five BASELINE.json configs.

Pure input generation (numpy, seeded, vectorised so 16M entities build in seconds): component
arrays in the reference's layouts plus the frusta of the views.  Nothing here evaluates the hot
path; the arrays feed the C ABI (and, in tests/ and bench.py, the oracle).

    scene = dict(num_entities, parents u32[N], local_matrices f32[N,16], entities u64[N],
                 state_flags u64[N], bounding_volumes f32[N,13])

Scene shapes follow the reference's examples:
  torus     examples/code/entities/entities.cpp:86-184  (256x256 boxes under one master node)
  rigs      core/put/source/ecs/ecs_resources.cpp:489-500,1153-1182 (mesh + contiguous bone block)
  views     core/put/source/camera.cpp:131-172,321-360,384-416; ecs_scene.cpp:538-571,685-714
"""
from __future__ import annotations

import numpy as np

# e_cmp / e_state bits (core/put/source/ecs/ecs_scene.h:144-171, :69-83)
CMP_ALLOCATED, CMP_GEOMETRY, CMP_PHYSICS, CMP_MATERIAL = 1 << 0, 1 << 1, 1 << 2, 1 << 4
CMP_LIGHT, CMP_BONE, CMP_ANIM_CONTROLLER, CMP_TRANSFORM, CMP_SUB_INSTANCE = 1 << 6, 1 << 7, 1 << 9, 1 << 12, 1 << 14
STATE_HIDDEN, STATE_SYNC_PHYSICS_TRANSFORM = 1 << 2, 1 << 7
RENDERABLE = CMP_ALLOCATED | CMP_GEOMETRY | CMP_MATERIAL

N_TORUS = 65_544
N_1M = 1 << 20
N_16M = 1 << 24
RIG_ENTITIES = 66  # root + skinned mesh + 64 bones


# ------------------------------------------------------------------------------------------------
# matrices (inputs only: any float32 local matrix is a valid input, so plain numpy is fine)
# ------------------------------------------------------------------------------------------------
def random_unit_quats(rng, n):
    q = rng.standard_normal((n, 4), dtype=np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True).astype(np.float32)
    return q


def trs_matrices(t, q, s):
    """local = T * R * S, row-major, column vectors (translation in m[3], m[7], m[11])."""
    n = t.shape[0]
    x, y, z, w = (q[:, i].astype(np.float32) for i in range(4))
    m = np.zeros((n, 16), np.float32)
    two = np.float32(2.0)
    one = np.float32(1.0)
    r = [one - two * (y * y + z * z), two * (x * y - z * w), two * (x * z + y * w),
         two * (x * y + z * w), one - two * (x * x + z * z), two * (y * z - x * w),
         two * (x * z - y * w), two * (y * z + x * w), one - two * (x * x + y * y)]
    s = np.broadcast_to(np.asarray(s, np.float32), (n, 3))
    for row in range(3):
        for col in range(3):
            m[:, row * 4 + col] = r[row * 3 + col] * s[:, col]
        m[:, row * 4 + 3] = t[:, row]
    m[:, 15] = 1.0
    return m


def euler_quats(rx, ry, rz):
    """quaternion from euler angles in radians (x, y, z order)."""
    cx, sx = np.cos(rx * 0.5), np.sin(rx * 0.5)
    cy, sy = np.cos(ry * 0.5), np.sin(ry * 0.5)
    cz, sz = np.cos(rz * 0.5), np.sin(rz * 0.5)
    w = cx * cy * cz + sx * sy * sz
    x = sx * cy * cz - cx * sy * sz
    y = cx * sy * cz + sx * cy * sz
    z = cx * cy * sz - sx * sy * cz
    return np.stack([x, y, z, w], axis=1).astype(np.float32)


def _bounds(rng, n, unit_cube=False):
    bv = np.zeros((n, 13), np.float32)
    if unit_cube:
        bv[:, 0:3] = -1.0  # ecs_primitives.cpp:739-740
        bv[:, 3:6] = 1.0
    else:
        bv[:, 0:3] = rng.uniform(-1.0, -0.1, (n, 3)).astype(np.float32)
        bv[:, 3:6] = rng.uniform(0.1, 1.0, (n, 3)).astype(np.float32)
    return bv


# ------------------------------------------------------------------------------------------------
# hierarchies with parents[n] <= n (ecs_utilities.cpp:408-415), roots self-parented (:256-257)
# ------------------------------------------------------------------------------------------------
def parents_from_target_depths(target):
    """target[n] = wanted depth (0 = root).  Node n hangs under the closest earlier node whose
    final depth is target[n]-1 inside its own root's block, else directly under the block root.
    Returns (parents u32, depth u32).  Preorder-like layout, as the loaders append hierarchies."""
    n = target.shape[0]
    idx = np.arange(n, dtype=np.int64)
    target = target.astype(np.int64).copy()
    target[0] = 0
    root_of = np.maximum.accumulate(np.where(target == 0, idx, -1))
    parents = idx.copy()
    depth = np.zeros(n, np.int64)
    is_d1 = target == 1
    parents[is_d1] = root_of[is_d1]
    depth[is_d1] = 1
    for d in range(2, int(target.max()) + 1):
        sel = target == d
        if not sel.any():
            continue
        last_prev = np.maximum.accumulate(np.where(depth == d - 1, idx, -1))  # depth 0 never matches d-1 >= 1
        cand = last_prev[sel]
        ok = cand >= root_of[sel]  # the candidate lies inside this node's block
        parents[sel] = np.where(ok, cand, root_of[sel])
        depth[sel] = np.where(ok, d, 1)
    return parents.astype(np.uint32), depth.astype(np.uint32)


def depths_of(parents):
    """depth of every entity by climbing parent links (host helper for tests and shard planning)."""
    p = parents.astype(np.int64)
    cur = np.arange(p.shape[0])
    depth = np.zeros(p.shape[0], np.int64)
    while True:
        nxt = p[cur]
        moved = nxt != cur
        if not moved.any():
            break
        depth += moved
        cur = nxt
    return depth.astype(np.uint32)


# ------------------------------------------------------------------------------------------------
# config 1 — the 64k torus demo shape
# ------------------------------------------------------------------------------------------------
def torus_scene(seed=1):
    rng = np.random.default_rng(seed)
    n = N_TORUS
    side = 256
    parents = np.arange(n, dtype=np.uint32)
    entities = np.zeros(n, np.uint64)
    t = np.zeros((n, 3), np.float32)
    q = np.zeros((n, 4), np.float32)
    q[:, 3] = 1.0
    s = np.ones((n, 3), np.float32)
    bv = _bounds(rng, n, unit_cube=True)
    # 0-3 directional lights, 4 floor, 5-6 walls, 7 master, then 256x256 boxes (entities.cpp:86-184)
    entities[0:4] = CMP_ALLOCATED | CMP_LIGHT
    entities[4:7] = RENDERABLE
    t[4] = (0.0, -60.0, 0.0)
    s[4] = (200.0, 1.0, 200.0)
    t[5] = (0.0, 0.0, -200.0)
    s[5] = (200.0, 100.0, 1.0)
    t[6] = (-200.0, 0.0, 0.0)
    s[6] = (1.0, 100.0, 200.0)
    master = 7
    entities[master] = RENDERABLE
    box0 = 8
    nb = side * side
    parents[box0:] = master
    entities[box0:] = RENDERABLE  # sub_instance cleared so the cull has work
    i, j = np.divmod(np.arange(nb), side)
    u = (i.astype(np.float32) / side) * np.float32(2 * np.pi)
    v = (j.astype(np.float32) / side) * np.float32(2 * np.pi)
    big, small = np.float32(100.0), np.float32(40.0)
    t[box0:, 0] = (big + small * np.cos(v)) * np.cos(u)
    t[box0:, 1] = small * np.sin(v)
    t[box0:, 2] = (big + small * np.cos(v)) * np.sin(u)
    ang = np.deg2rad(rng.integers(0, 360, (nb, 3)).astype(np.float32))  # rand()%360 degrees
    q[box0:] = euler_quats(ang[:, 0], ang[:, 1], ang[:, 2])
    local = trs_matrices(t, q, s)
    return dict(num_entities=n, parents=parents, local_matrices=local, entities=entities,
                state_flags=np.zeros(n, np.uint64), bounding_volumes=bv)


def torus_views(camera_cls):
    cam = camera_cls().create_perspective(60.0, 16.0 / 9.0, 0.1, 1000.0)
    # zoom 180, rot (-0.6, 2.2) (entities.cpp:25-26): orbit camera around the origin
    rx, ry, zoom = -0.6, 2.2, 180.0
    pos = (zoom * np.cos(rx) * np.sin(ry), -zoom * np.sin(rx), zoom * np.cos(rx) * np.cos(ry))
    cam.look_at(pos, (0.0, 0.0, 0.0)).update_frustum()
    return np.stack([cam.frustum])


# ------------------------------------------------------------------------------------------------
# config 2 — 1M entities, random hierarchy of depth 16
# ------------------------------------------------------------------------------------------------
def deep_scene(n=N_1M, max_depth=16, seed=2, block=65536):
    rng = np.random.default_rng(seed)
    target = rng.integers(1, max_depth + 1, n)
    target[::block] = 0
    # bias towards "one deeper than the previous node" so that chains actually reach max_depth
    step = rng.random(n) < 0.5
    prev = np.roll(target, 1)
    target = np.where(step & (prev > 0) & (prev < max_depth) & (target != 0), prev + 1, target)
    target[::block] = 0
    parents, depth = parents_from_target_depths(target)
    t = rng.uniform(-10.0, 10.0, (n, 3)).astype(np.float32)
    q = random_unit_quats(rng, n)
    # scales near 1 below the roots keep depth-16 products finite and well conditioned
    s = rng.uniform(0.8, 1.25, (n, 3)).astype(np.float32)
    local = trs_matrices(t, q, s)
    entities = np.full(n, RENDERABLE, np.uint64)
    return dict(num_entities=n, parents=parents, local_matrices=local, entities=entities,
                state_flags=np.zeros(n, np.uint64), bounding_volumes=_bounds(rng, n), depth=depth)


# ------------------------------------------------------------------------------------------------
# config 3 / 5 — large flat-ish scene, 11 views
# ------------------------------------------------------------------------------------------------
def _city_hierarchy(rng, n):
    r = rng.random(n, dtype=np.float32)
    target = np.zeros(n, np.int64)
    target[r > 0.90] = 1
    target[r > 0.95] = 2
    target[r > 0.98] = 3
    return parents_from_target_depths(target)


def city_parents(n=N_16M, seed=3):
    """the parents array of city_scene(n, seed) alone (shard planning over a scene made of several such blocks)"""
    return _city_hierarchy(np.random.default_rng(seed), n)[0]


def city_scene(n=N_16M, seed=3, half_extent=2000.0, with_transforms=False):
    rng = np.random.default_rng(seed)
    parents, depth = _city_hierarchy(rng, n)
    t = rng.uniform(-half_extent, half_extent, (n, 3)).astype(np.float32)
    child = depth > 0
    t[child] = rng.uniform(-5.0, 5.0, (int(child.sum()), 3)).astype(np.float32)  # children sit near their parent
    q = random_unit_quats(rng, n)
    s = rng.uniform(0.5, 2.0, (n, 3)).astype(np.float32)
    local = trs_matrices(t, q, s)
    transforms = np.concatenate([t, q, s], axis=1).astype(np.float32) if with_transforms else None  # cmp_transform records
    del t, q, s
    kind = rng.random(n, dtype=np.float32)
    entities = np.full(n, RENDERABLE, np.uint64)
    entities[kind < 0.02] = CMP_ALLOCATED  # 2 % non-renderable (lights, nodes)
    state = np.zeros(n, np.uint64)
    state[(kind >= 0.02) & (kind < 0.03)] = STATE_HIDDEN  # 1 % hidden
    out = dict(num_entities=n, parents=parents, local_matrices=local, entities=entities,
               state_flags=state, bounding_volumes=_bounds(rng, n), depth=depth)
    if transforms is not None:
        out["transforms"] = transforms
    return out


def city_views(camera_cls, eye=(0.0, 50.0, 0.0), at=(600.0, 0.0, 400.0)):
    """1 perspective camera + 4 ortho 'cascades' + 6 point-light cube faces = 11 views."""
    views = []
    cam = camera_cls().create_perspective(60.0, 16.0 / 9.0, 0.1, 1000.0)
    cam.look_at(eye, at).update_frustum()
    views.append(cam.frustum)
    light_dir = np.array([1.0, 1.0, 1.0], np.float32)
    light_dir /= np.linalg.norm(light_dir)
    for near, far in ((0.1, 30.0), (30.0, 120.0), (120.0, 400.0), (400.0, 1000.0)):
        # the cascade's fit box = AABB of the camera-frustum slice (a reference shadow view fits the
        # scene AABB, camera.cpp:384-416 / ecs_scene.cpp:538-554; a slice box is the cascade analogue)
        sl = camera_cls().create_perspective(60.0, 16.0 / 9.0, near, far)
        sl.look_at(eye, at).update_frustum()
        corners = sl.frustum[36:60].reshape(8, 3)
        c = camera_cls()
        c.update_shadow_frustum(light_dir, corners.min(axis=0), corners.max(axis=0))
        views.append(c.frustum)
    radius = 500.0
    for face in range(6):
        c = camera_cls().create_cubemap(0.1, 2.0 * radius)  # ecs_scene.cpp:700-702
        c.set_pos((0.0, 0.0, 0.0)).set_cubemap_face(face).update_frustum()
        views.append(c.frustum)
    return np.stack(views).astype(np.float32)


# ------------------------------------------------------------------------------------------------
# config 4 — skinned characters: root (root motion) + mesh + 64 bones, depth <= 12
# ------------------------------------------------------------------------------------------------
def _rig_topology():
    """parent offsets inside one 66-entity rig: a spine, a head chain, two arms with 3-joint
    fingers, two legs.  Entity 0 = root, 1 = skinned mesh, 2.. = bones (contiguous)."""
    par = [0, 0]  # root self, mesh under root
    def bone(p):
        par.append(p)
        return len(par) - 1
    hips = bone(0)
    spine = hips
    for _ in range(4):
        spine = bone(spine)  # depth 2..5
    neck = bone(spine)
    head = bone(neck)
    bone(head)  # jaw
    for _side in range(2):
        clav = bone(spine)
        up = bone(clav)
        lo = bone(up)
        hand = bone(lo)  # depth 9
        for _f in range(5):
            f = hand
            for _j in range(3):
                f = bone(f)  # depth 10..12
    for _side in range(2):
        thigh = bone(hips)
        shin = bone(thigh)
        foot = bone(shin)
        bone(foot)
    while len(par) < RIG_ENTITIES:
        bone(hips)  # attachment bones
    assert len(par) == RIG_ENTITIES
    return np.array(par, np.int64)


def rig_scene(characters=10_000, seed=4, skins=2):
    """BASELINE config 4: `characters` rigs of 66 entities (root with the anim controller, skinned mesh,
    64 contiguous bones).  Besides the baked local matrices the scene carries what the stage-0 path
    needs: `transforms` (cmp_transform per entity; every entity starts flagged e_cmp::transform so the
    first update bakes all of them), a root-motion batch (one trajectory delta per rig, applied to the
    rig root, ecs_scene.cpp:1286-1304) and the skin table (characters share `skins` bind-matrix sets)."""
    rng = np.random.default_rng(seed)
    topo = _rig_topology()
    n = characters * RIG_ENTITIES
    base = np.repeat(np.arange(characters, dtype=np.int64) * RIG_ENTITIES, RIG_ENTITIES)
    parents = (base + np.tile(topo, characters)).astype(np.uint32)
    slot = np.tile(np.arange(RIG_ENTITIES), characters)
    entities = np.full(n, CMP_ALLOCATED | CMP_BONE, np.uint64)
    entities[slot == 0] = CMP_ALLOCATED | CMP_ANIM_CONTROLLER
    entities[slot == 1] = RENDERABLE
    t = rng.uniform(-0.3, 0.3, (n, 3)).astype(np.float32)
    t[:, 1] += np.float32(0.25)  # bones extend along +y
    ang = rng.uniform(-0.35, 0.35, (n, 3)).astype(np.float32)
    q = euler_quats(ang[:, 0], ang[:, 1], ang[:, 2])
    s = np.ones((n, 3), np.float32)
    roots = slot == 0
    nr = int(roots.sum())
    # character placement + root-motion offset added to the rig root's translation (ecs_scene.cpp:1286-1304)
    t[roots] = rng.uniform(-500.0, 500.0, (nr, 3)).astype(np.float32) + rng.uniform(-0.05, 0.05, (nr, 3)).astype(np.float32)
    t[roots, 1] = 0.0
    local = trs_matrices(t, q, s)
    bv = _bounds(rng, n)
    transforms = np.concatenate([t, q, s], axis=1).astype(np.float32)
    # root motion of this frame: one delta per rig, rotated by the trajectory joint's initial rotation
    rm_targets = np.flatnonzero(roots).astype(np.uint32)
    rm_ang = rng.uniform(-3.1, 3.1, (nr, 3)).astype(np.float32)
    rm_rot = euler_quats(np.zeros(nr, np.float32), rm_ang[:, 1], np.zeros(nr, np.float32))
    rm_delta = rng.uniform(-0.08, 0.08, (nr, 3)).astype(np.float32)
    # skins: joints are the 64 bones after the mesh; a few bind-matrix sets shared by all characters
    nb = RIG_ENTITIES - 2
    bq = random_unit_quats(rng, skins * nb)
    bind = trs_matrices(rng.uniform(-1.0, 1.0, (skins * nb, 3)).astype(np.float32), bq, np.ones((skins * nb, 3), np.float32))
    from .capi import SKIN_DTYPE
    sk = np.zeros(characters, SKIN_DTYPE)
    sk["joints_offset"] = np.arange(characters, dtype=np.uint32) * RIG_ENTITIES + 2
    sk["num_joints"] = nb
    sk["bind_offset"] = (np.arange(characters) % skins).astype(np.uint32) * nb
    sk["palette_offset"] = np.arange(characters, dtype=np.uint32) * nb
    return dict(num_entities=n, parents=parents, local_matrices=local, entities=entities,
                state_flags=np.zeros(n, np.uint64), bounding_volumes=bv, transforms=transforms,
                root_motion=dict(targets=rm_targets, rotations=rm_rot, deltas=rm_delta),
                skins=sk, bind_matrices=bind, num_palette_matrices=characters * nb)


# ------------------------------------------------------------------------------------------------
# small scenes that exercise every branch of loops A-C and the filter (edge cases for the tests)
# ------------------------------------------------------------------------------------------------
def mixed_scene(n=5000, seed=7, max_depth=40):
    """deep chains (beyond the in-thread ancestor walk), unallocated slots, bones, hidden and
    sub-instance entities, physics entities with and without an rb matrix, dirty-flag bits."""
    rng = np.random.default_rng(seed)
    target = rng.integers(0, 4, n)
    # a few long chains
    pos = 10
    while pos + max_depth + 2 < n and pos < n // 2:
        ln = int(rng.integers(max_depth // 2, max_depth + 1))
        target[pos] = 0
        target[pos + 1:pos + 1 + ln] = np.arange(1, ln + 1)
        pos += ln + int(rng.integers(50, 400))
    parents, depth = parents_from_target_depths(target)
    t = rng.uniform(-300.0, 300.0, (n, 3)).astype(np.float32)
    t[depth > 0] = rng.uniform(-3.0, 3.0, (int((depth > 0).sum()), 3)).astype(np.float32)
    s = rng.uniform(0.9, 1.1, (n, 3)).astype(np.float32)
    local = trs_matrices(t, random_unit_quats(rng, n), s)
    kind = rng.random(n)
    entities = np.full(n, RENDERABLE, np.uint64)
    state = np.zeros(n, np.uint64)
    entities[kind < 0.05] = 0  # unallocated slot: processed like any other (world = local)
    unalloc = kind < 0.05
    local[unalloc] = 0.0
    parents[unalloc] = np.arange(n, dtype=np.uint32)[unalloc]  # zeroed + self-parented
    # children of a now-unallocated parent still point at it (allowed: parents[n] <= n)
    entities[(kind >= 0.05) & (kind < 0.15)] |= np.uint64(CMP_BONE)
    state[(kind >= 0.15) & (kind < 0.20)] |= np.uint64(STATE_HIDDEN)
    entities[(kind >= 0.20) & (kind < 0.25)] |= np.uint64(CMP_SUB_INSTANCE)
    entities[(kind >= 0.25) & (kind < 0.30)] &= ~np.uint64(CMP_MATERIAL)
    phys = (kind >= 0.30) & (kind < 0.40)
    entities[phys] |= np.uint64(CMP_PHYSICS)
    entities[(kind >= 0.40) & (kind < 0.50)] |= np.uint64(CMP_TRANSFORM)
    entities[(kind >= 0.36) & (kind < 0.40)] |= np.uint64(CMP_TRANSFORM)  # physics + transform dirty
    state[(kind >= 0.38) & (kind < 0.42)] |= np.uint64(STATE_SYNC_PHYSICS_TRANSFORM)
    entities[(kind >= 0.50) & (kind < 0.53)] |= np.uint64(CMP_ANIM_CONTROLLER)
    has_rb = (rng.random(n) < 0.5).astype(np.uint8)
    has_rb[~phys] = 0
    # previous-frame outputs (a physics entity without an rb matrix keeps its world matrix; bones
    # keep radius / pos_extent): arbitrary but finite
    world_prev = trs_matrices(rng.uniform(-300, 300, (n, 3)).astype(np.float32), random_unit_quats(rng, n),
                              np.ones((n, 3), np.float32))
    pe_prev = rng.uniform(-300, 300, (n, 8)).astype(np.float32)
    pe_prev[:, 4:8] = np.abs(pe_prev[:, 4:8]) * np.float32(0.01)
    # pos.w "is unused" (ecs_scene.h:259): ecs_scene.cpp:1515 is the only writer of pos_extent and it
    # leaves w alone, so in the reference w keeps the zero of resize_scene_buffers for ever
    pe_prev[:, 3] = 0.0
    bv = _bounds(rng, n)
    bv[:, 12] = rng.uniform(0.1, 2.0, n).astype(np.float32)
    return dict(num_entities=n, parents=parents, local_matrices=local, entities=entities, state_flags=state,
                bounding_volumes=bv, has_rb_matrix=has_rb, world_matrices=world_prev, pos_extent=pe_prev, depth=depth)


def mixed_views(camera_cls):
    views = [torus_views(camera_cls)[0]]
    cam = camera_cls().create_perspective(75.0, 1.0, 1.0, 500.0)
    cam.look_at((250.0, 100.0, -250.0), (0.0, 0.0, 0.0)).update_frustum()
    views.append(cam.frustum)
    c = camera_cls()
    d = np.array([0.3, 0.9, 0.2], np.float32)
    c.update_shadow_frustum(d / np.linalg.norm(d), np.array([-150, -150, -150], np.float32), np.array([150, 150, 150], np.float32))
    views.append(c.frustum)
    for face in (0, 3, 5):
        c = camera_cls().create_cubemap(0.1, 400.0)
        c.set_pos((20.0, -10.0, 30.0)).set_cubemap_face(face).update_frustum()
        views.append(c.frustum)
    return np.stack(views).astype(np.float32)


# ------------------------------------------------------------------------------------------------
# root-subtree partition: contiguous ranges that start at self-parented entities
# ------------------------------------------------------------------------------------------------
def partition_by_root(parents, parts):
    """Split [0, N) into `parts` contiguous ranges of near-equal entity count whose starts are
    roots (self-parented) and which no parent link crosses.  Returns a list of (first, last)."""
    n = parents.shape[0]
    idx = np.arange(n, dtype=np.uint32)  # u32 throughout: config 5 plans over 128 Mi entries
    p = np.ascontiguousarray(parents, np.uint32)
    # a cut before i is legal iff i is a root and no entity at or after i has a parent before i.
    # With parents[j] <= j the minimum parent over the suffix [i, N) decides it.
    legal = np.minimum.accumulate(p[::-1])[::-1] >= idx
    legal &= p == idx
    del idx
    legal_idx = np.flatnonzero(legal)
    cuts = [0]
    for k in range(1, parts):
        want = (n * k) // parts
        j = int(np.searchsorted(legal_idx, want))
        cands = [int(legal_idx[min(max(j + o, 0), len(legal_idx) - 1)]) for o in (-1, 0)]
        c = min(cands, key=lambda c: abs(c - want))
        cuts.append(max(c, cuts[-1]))
    cuts.append(n)
    return [(cuts[i], cuts[i + 1]) for i in range(parts)]


def slice_scene(scene, first, last):
    """The sub-scene of a root-aligned range with parents rebased to the range."""
    out = {}
    for k, v in scene.items():
        if isinstance(v, np.ndarray) and v.shape[:1] == scene["parents"].shape[:1]:
            out[k] = v[first:last]
    out["parents"] = (scene["parents"][first:last].astype(np.int64) - first).astype(np.uint32)
    out["num_entities"] = last - first
    return out
