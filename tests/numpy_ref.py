"""A second, independent restatement of the path in numpy float32 (vectorised over entities, one
numpy operation per IEEE operation, so no contraction and the same left-to-right order).  It
exists only to cross-check oracle/pm_oracle.c: two restatements written separately, from the
reference call sites, must agree bit for bit.  Test infrastructure — never imported by the product.

Reference: core/put/source/ecs/ecs_scene.cpp:1405-1556, core/put/source/ecs/ecs_cull.cpp:21-105.
"""
import numpy as np

F = np.float32
CMP_ALLOCATED, CMP_GEOMETRY, CMP_PHYSICS, CMP_MATERIAL = 1 << 0, 1 << 1, 1 << 2, 1 << 4
CMP_BONE, CMP_TRANSFORM, CMP_SUB_INSTANCE = 1 << 7, 1 << 12, 1 << 14
STATE_HIDDEN, STATE_SYNC_PHYSICS_TRANSFORM = 1 << 2, 1 << 7


def mat_mul(a, b):
    """[K,16] x [K,16] row-major; out[r][c] = ((0 + a[r][0]*b[0][c]) + a[r][1]*b[1][c]) + ..."""
    a = a.reshape(-1, 4, 4)
    b = b.reshape(-1, 4, 4)
    out = np.zeros_like(a)
    for r in range(4):
        for c in range(4):
            acc = np.zeros(a.shape[0], F)
            for k in range(4):
                acc = (acc + (a[:, r, k] * b[:, k, c]).astype(F)).astype(F)
            out[:, r, c] = acc
    return out.reshape(-1, 16)


def update_transforms(sc):
    """loop A (ecs_scene.cpp:1405-1471) without the TRS bake: level by level, which visits every
    entity after its parent exactly like the ascending sweep does."""
    n = sc["num_entities"]
    parents = sc["parents"][:n].astype(np.int64)
    ent = sc["entities"][:n].copy()
    sf = sc["state_flags"][:n].copy()
    local = sc["local_matrices"][:n]
    world = (np.zeros((n, 16), F) if sc.get("world_matrices") is None else sc["world_matrices"][:n].copy())
    has_rb = np.zeros(n, bool) if sc.get("has_rb_matrix") is None else sc["has_rb_matrix"][:n].astype(bool)
    physics = (ent & np.uint64(CMP_PHYSICS)) != 0
    transform = ((ent & np.uint64(CMP_TRANSFORM)) != 0) & ((sf & np.uint64(STATE_SYNC_PHYSICS_TRANSFORM)) == 0)
    skipped = physics & ~transform & ~has_rb
    idx = np.arange(n)
    depth = np.zeros(n, np.int64)
    cur = idx.copy()
    while True:
        p = parents[cur]
        up = p != cur
        if not up.any():
            break
        depth += up
        cur = p
    for d in range(int(depth.max()) + 1 if n else 0):
        sel = np.flatnonzero((depth == d) & ~skipped)
        if d == 0:
            world[sel] = local[sel]
        else:
            world[sel] = mat_mul(world[parents[sel]], local[sel])
    ent &= ~np.uint64(CMP_TRANSFORM)
    sf &= ~np.uint64(STATE_SYNC_PHYSICS_TRANSFORM)
    return world, ent, sf


def update_bounds(sc, world, ent):
    """loop B (ecs_scene.cpp:1474-1526)."""
    n = sc["num_entities"]
    bv = sc["bounding_volumes"][:n].copy()
    pe = (np.zeros((n, 8), F) if sc.get("pos_extent") is None else sc["pos_extent"][:n].copy())
    w = world.reshape(n, 4, 4)
    bone = (ent & np.uint64(CMP_BONE)) != 0
    mn, mx = bv[:, 0:3], bv[:, 3:6]
    size = (mx - mn).astype(F)
    corners = [(0, 0, 0), (1, 0, 0), (0, 1, 0), (0, 0, 1), (1, 1, 0), (0, 1, 1), (1, 0, 1), (1, 1, 1)]
    tmax = np.full((n, 3), -np.finfo(F).max, F)
    tmin = np.full((n, 3), np.finfo(F).max, F)
    for c in corners:
        sel = np.array(c, F)
        p = (mn + (size * sel).astype(F)).astype(F)
        t = np.zeros((n, 3), F)
        for r in range(3):
            acc = np.zeros(n, F)
            for k in range(3):
                acc = (acc + (w[:, r, k] * p[:, k]).astype(F)).astype(F)
            t[:, r] = (acc + (w[:, r, 3] * F(1.0)).astype(F)).astype(F)
        tmax = np.where(tmax > t, tmax, t)
        tmin = np.where(tmin < t, tmin, t)
    d = (tmax - tmin).astype(F)
    mag2 = np.zeros(n, F)
    for k in range(3):
        mag2 = (mag2 + (d[:, k] * d[:, k]).astype(F)).astype(F)
    radius = (np.sqrt(mag2).astype(F) * F(0.5)).astype(F)
    pos = (tmin + (d * F(0.5)).astype(F)).astype(F)
    ext = (tmax - pos).astype(F)
    nb = ~bone
    bv[nb, 6:9] = tmin[nb]
    bv[nb, 9:12] = tmax[nb]
    bv[nb, 12] = radius[nb]
    pe[nb, 0:3] = pos[nb]
    pe[nb, 4:7] = ext[nb]
    pe[nb, 7] = radius[nb]
    trans = w[:, 0:3, 3]
    bv[bone, 6:9] = trans[bone]
    bv[bone, 9:12] = trans[bone]
    geo = nb & ((ent & np.uint64(CMP_GEOMETRY)) != 0)
    extents = np.concatenate([np.full(3, np.finfo(F).max, F), np.full(3, -np.finfo(F).max, F)])
    if geo.any():
        extents[0:3] = bv[geo, 6:9].min(axis=0)
        extents[3:6] = bv[geo, 9:12].max(axis=0)
    return bv, pe, extents


def expand_parents(sc, bv, ent):
    """loop C (ecs_scene.cpp:1529-1556): descending sweep, plain python (small cases only)."""
    n = sc["num_entities"]
    parents = sc["parents"][:n]
    for i in range(n - 1, 0, -1):
        if not (int(ent[i]) & CMP_ALLOCATED):
            continue
        p = int(parents[i])
        if p == i:
            continue
        bv[p, 6:9] = np.where(bv[p, 6:9] < bv[i, 6:9], bv[p, 6:9], bv[i, 6:9])
        bv[p, 9:12] = np.where(bv[p, 9:12] > bv[i, 9:12], bv[p, 9:12], bv[i, 9:12])
    return bv


def filter_entities(ent, sf):
    """ecs_cull.cpp:84-105 (flags as they stand when the cull runs, i.e. after update_scene)."""
    acc = np.uint64(CMP_GEOMETRY | CMP_MATERIAL)
    keep = ((ent & acc) == acc) & ((sf & np.uint64(STATE_HIDDEN)) == 0) & ((ent & np.uint64(CMP_SUB_INSTANCE)) == 0)
    return np.flatnonzero(keep).astype(np.uint32)


def cull_aabb(pe, frustum, ents):
    """ecs_cull.cpp:21-51"""
    fr = np.asarray(frustum, F).reshape(-1)
    nrm, pts = fr[0:18].reshape(6, 3), fr[18:36].reshape(6, 3)
    pos, ext = pe[ents, 0:3], pe[ents, 4:7]
    inside = np.ones(len(ents), bool)
    for p in range(6):
        nn, x0 = nrm[p], pts[p]
        sgn = ((F(0) < nn).astype(F) - (nn < F(0)).astype(F)).astype(F)
        sf = (sgn * F(-1.0)).astype(F)
        pd = F(0)
        for k in range(3):
            pd = F(pd + F(nn[k] * x0[k]))
        pd = F(pd * F(-1.0))
        corner = (pos + (ext * sf).astype(F)).astype(F)
        d2 = np.zeros(len(ents), F)
        for k in range(3):
            d2 = (d2 + (corner[:, k] * nn[k]).astype(F)).astype(F)
        inside &= ~(d2 > F(-pd))
    return ents[inside]


def cull_sphere(pe, frustum, ents):
    """ecs_cull.cpp:53-82"""
    fr = np.asarray(frustum, F).reshape(-1)
    nrm, pts = fr[0:18].reshape(6, 3), fr[18:36].reshape(6, 3)
    pos, radius = pe[ents, 0:3], pe[ents, 7]
    inside = np.ones(len(ents), bool)
    for p in range(6):
        nn, x0 = nrm[p], pts[p]
        pd = F(0)
        for k in range(3):
            pd = F(pd + F(nn[k] * x0[k]))
        pd = F(pd * F(-1.0))
        d = np.zeros(len(ents), F)
        for k in range(3):
            d = (d + (pos[:, k] * nn[k]).astype(F)).astype(F)
        d = (d + pd).astype(F)
        inside &= ~(d > radius)
    return ents[inside]


def frame(sc, views, mode="aabb"):
    world, ent, sf = update_transforms(sc)
    bv, pe, extents = update_bounds(sc, world, ent)
    bv = expand_parents(sc, bv, ent)
    filt = filter_entities(ent, sf)
    cull = cull_aabb if mode == "aabb" else cull_sphere
    lists = [cull(pe, v, filt) for v in (views if views is not None else [])]
    return dict(world=world, bv=bv, pos_extent=pe, entities=ent, state_flags=sf, extents=extents, lists=lists)


# ------------------------------------------------------------------------------------------------
# stage 0, root motion, skin palettes (ecs_scene.cpp:1415-1427, :1286-1304, :1917-1959)
# ------------------------------------------------------------------------------------------------
def quat_normalise(q):
    """quat::normalise on [K,4] (x,y,z,w): divide by |q| unless it is 0 or 1."""
    q = np.asarray(q, F).reshape(-1, 4)
    x, y, z, w = (q[:, i].copy() for i in range(4))
    mag2 = (((w * w + x * x).astype(F) + y * y).astype(F) + z * z).astype(F)
    mag = np.sqrt(mag2).astype(F)
    norm = (mag != 0) & (mag != 1)
    safe = np.where(norm, mag, F(1))
    return np.stack([np.where(norm, (c / safe).astype(F), c) for c in (x, y, z, w)], axis=1).astype(F)


def quat_expand(q):
    """The expansion half of quat::get_matrix; one numpy operation per IEEE operation, C's left-to-right
    order (1 - 2*y*y - 2*z*z = ((1 - (2*y)*y) - (2*z)*z)).  The rotations stored in the reference's own
    scene file expand to its stored local matrices bit for bit (tests/golden/pmtech_sdf_scene.npz)."""
    q = np.asarray(q, F).reshape(-1, 4)
    x, y, z, w = (q[:, i] for i in range(4))
    two, one = F(2), F(1)
    x2, y2, z2 = two * x, two * y, two * z
    m = np.zeros((q.shape[0], 16), F)
    m[:, 0] = (one - y2 * y) - z2 * z
    m[:, 1] = x2 * y - z2 * w
    m[:, 2] = x2 * z + y2 * w
    m[:, 4] = x2 * y + z2 * w
    m[:, 5] = (one - x2 * x) - z2 * z
    m[:, 6] = y2 * z - x2 * w
    m[:, 8] = x2 * z - y2 * w
    m[:, 9] = y2 * z + x2 * w
    m[:, 10] = (one - x2 * x) - y2 * y
    m[:, 15] = 1
    return m


def quat_matrix(q):
    """quat::get_matrix: the quaternion normalises ITSELF, then expands.  Returns (matrix, normalised q) —
    the caller owns the write-back the reference does in place."""
    qn = quat_normalise(q)
    return quat_expand(qn), qn


def bake_transforms(sc):
    """local = translation_mat * rot_mat * scale_mat for the entities flagged e_cmp::transform (and not
    forced to sync with physics).  Returns the new local matrices and the transforms, whose rotations
    get_matrix normalised in place (ecs_scene.cpp:1416-1421: `cmp_transform& t`, `t.rotation.get_matrix`)."""
    n = sc["num_entities"]
    ent, sf, t = sc["entities"][:n], sc["state_flags"][:n], np.asarray(sc["transforms"][:n], F).copy()
    local = sc["local_matrices"][:n].copy()
    sel = np.flatnonzero(((ent & np.uint64(CMP_TRANSFORM)) != 0) & ((sf & np.uint64(STATE_SYNC_PHYSICS_TRANSFORM)) == 0))
    k = sel.shape[0]
    T = np.tile(np.eye(4, dtype=F).reshape(1, 16), (k, 1))
    T[:, 3], T[:, 7], T[:, 11] = t[sel, 0], t[sel, 1], t[sel, 2]
    S = np.zeros((k, 16), F)
    S[:, 0], S[:, 5], S[:, 10], S[:, 15] = t[sel, 7], t[sel, 8], t[sel, 9], 1
    R, qn = quat_matrix(t[sel, 3:7])
    t[sel, 3:7] = qn
    local[sel] = mat_mul(mat_mul(T, R), S)
    return local, t


def apply_root_motion(sc, targets, rotations, deltas):
    """quat q = rotation; q.get_matrix(rot) (q normalised in place); transforms[p].rotation = q;
    .translation += rot.transform_vector(delta); flag set (ecs_scene.cpp:1290-1302)."""
    t = np.asarray(sc["transforms"], F).copy()
    ent = sc["entities"].copy()
    R, qn = quat_matrix(rotations)
    R = R.reshape(-1, 4, 4)
    d = np.asarray(deltas, F).reshape(-1, 3)
    v4 = np.concatenate([d, np.ones((d.shape[0], 1), F)], axis=1)
    for r in range(3):
        acc = np.zeros(d.shape[0], F)
        for k in range(4):
            acc = (acc + (v4[:, k] * R[:, r, k]).astype(F)).astype(F)
        t[targets, r] = (t[targets, r] + acc).astype(F)
    t[targets, 3:7] = qn
    ent[targets] |= np.uint64(CMP_TRANSFORM)
    return t, ent


def skin_palettes(world, skins, bind, num_palette_matrices):
    out = np.zeros((num_palette_matrices, 16), F)
    bind = np.asarray(bind, F).reshape(-1, 16)
    for sk in skins:
        j0, nj, b0, p0 = (int(sk[k]) for k in ("joints_offset", "num_joints", "bind_offset", "palette_offset"))
        out[p0:p0 + nj] = mat_mul(world[j0:j0 + nj], bind[b0:b0 + nj])
    return out


def inverse4x4(m):
    """mat::inverse4x4 (Laplace expansion, odd cofactors negated as a whole: `-(…) * inv`, which is what
    gives the reference its -0 checkerboard — tests/golden/pmtech_sdf_scene.npz), [K,16] float32, one
    numpy operation per C operation, left to right."""
    m = np.asarray(m, F).reshape(-1, 16)
    a = [m[:, i] for i in range(16)]
    s0 = a[0] * a[5] - a[4] * a[1]
    s1 = a[0] * a[6] - a[4] * a[2]
    s2 = a[0] * a[7] - a[4] * a[3]
    s3 = a[1] * a[6] - a[5] * a[2]
    s4 = a[1] * a[7] - a[5] * a[3]
    s5 = a[2] * a[7] - a[6] * a[3]
    c5 = a[10] * a[15] - a[14] * a[11]
    c4 = a[9] * a[15] - a[13] * a[11]
    c3 = a[9] * a[14] - a[13] * a[10]
    c2 = a[8] * a[15] - a[12] * a[11]
    c1 = a[8] * a[14] - a[12] * a[10]
    c0 = a[8] * a[13] - a[12] * a[9]
    det = ((((s0 * c5 - s1 * c4) + s2 * c3) + s3 * c2) - s4 * c1) + s5 * c0
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = (F(1) / det).astype(F)
    o = np.zeros_like(m)
    old = np.seterr(invalid="ignore", over="ignore")
    o[:, 0] = ((a[5] * c5 - a[6] * c4) + a[7] * c3) * inv
    o[:, 1] = -((a[1] * c5 - a[2] * c4) + a[3] * c3) * inv
    o[:, 2] = ((a[13] * s5 - a[14] * s4) + a[15] * s3) * inv
    o[:, 3] = -((a[9] * s5 - a[10] * s4) + a[11] * s3) * inv
    o[:, 4] = -((a[4] * c5 - a[6] * c2) + a[7] * c1) * inv
    o[:, 5] = ((a[0] * c5 - a[2] * c2) + a[3] * c1) * inv
    o[:, 6] = -((a[12] * s5 - a[14] * s2) + a[15] * s1) * inv
    o[:, 7] = ((a[8] * s5 - a[10] * s2) + a[11] * s1) * inv
    o[:, 8] = ((a[4] * c4 - a[5] * c2) + a[7] * c0) * inv
    o[:, 9] = -((a[0] * c4 - a[1] * c2) + a[3] * c0) * inv
    o[:, 10] = ((a[12] * s4 - a[13] * s2) + a[15] * s0) * inv
    o[:, 11] = -((a[8] * s4 - a[9] * s2) + a[11] * s0) * inv
    o[:, 12] = -((a[4] * c3 - a[5] * c1) + a[6] * c0) * inv
    o[:, 13] = ((a[0] * c3 - a[1] * c1) + a[2] * c0) * inv
    o[:, 14] = -((a[12] * s3 - a[13] * s1) + a[14] * s0) * inv
    o[:, 15] = ((a[8] * s3 - a[9] * s1) + a[10] * s0) * inv
    np.seterr(**old)
    return o


def update_draw_calls(world, entities, cbuffer, dc, time_ms):
    """ecs_scene.cpp:1962-1991 on [n,40] records (world 0:16, v1 16:20, v2 20:24, inverse transpose 24:40)."""
    n = world.shape[0]
    dc = dc.copy()
    dc[:, 0:16] = world
    dc[:, 16] = np.arange(n, dtype=np.float32)
    dc[:, 17] = F(time_ms)
    live = (cbuffer != 0) & (cbuffer != 0xFFFFFFFF) & ((entities & np.uint64(CMP_SUB_INSTANCE)) == 0)
    skinned = live & ((entities & np.uint64((1 << 6) | (1 << 16))) != 0)
    dc[skinned, 0:16] = np.eye(4, dtype=F).reshape(16)
    wt = world.reshape(-1, 4, 4).transpose(0, 2, 1).reshape(-1, 16)
    dc[live, 24:40] = inverse4x4(wt[live])
    return dc
