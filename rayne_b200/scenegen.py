"""Synthetic scene / camera / emitter generators (host side, numpy only).

One seed -> one scene, used identically by the CUDA path, the C oracle and the reference harness
(SURVEY.md §8(d)).  Scenes are described in *creation order* exactly as a Rayne user would build them:
every tree is completed with ``SceneNode::AddChild`` before its root is handed to ``Scene::AddNode``
(reference: Source/Scene/RNSceneNode.cpp:286-291, Source/Scene/RNSceneBasic.cpp:683-736).

Nothing here touches the GPU or the oracle.
"""
from __future__ import annotations

import dataclasses
import numpy as np

# RN::SceneNode::Flags (reference: Source/Scene/RNSceneNode.h:56-68)
FLAG_STATIC = 1 << 0
FLAG_HIDDEN = 1 << 1
FLAG_NOCULLING = 1 << 2
FLAG_OCCLUDER = 1 << 3

# RN::SceneNode::RenderPriority (reference: Source/Scene/RNSceneNode.h:44-54)
RENDER_EARLY = -100000
RENDER_NORMAL = 0
RENDER_LATE = 100000
RENDER_SKY = 200000
RENDER_TRANSPARENT = 300000

# RN::Camera::Flags (reference: Source/Scene/RNCamera.h:30-50)
CAM_ORTHOGONAL = 1 << 21
CAM_USE_SIMPLE_CULLING = 1 << 26
CAM_DEFAULTS = (1 << 24) | (1 << 5)
CAM_NO_RENDER, CAM_RENDER_EARLY, CAM_RENDER_LATE = 1 << 7, 1 << 19, 1 << 20

DEFAULT_RADIUS = np.sqrt(np.float32(3.0), dtype=np.float32)  # Sphere(AABB(-1, 1)), RNSphere.h:44-49


@dataclasses.dataclass
class Scene:
    """Node columns in creation order (parent index < child index, -1 for roots)."""
    parent: np.ndarray      # int32 [n]
    pos: np.ndarray         # float32 [n,3]
    scale: np.ndarray       # float32 [n,3]
    rot: np.ndarray         # float32 [n,4]  (x, y, z, w)
    flags: np.ndarray       # uint32 [n]
    group: np.ndarray       # uint32 [n]   (only the low 8 bits are used: uint8 _renderGroup)
    priority: np.ndarray    # int32 [n]
    sphere: np.ndarray      # float32 [n,4]  (offset xyz, radius)
    box: np.ndarray | None = None   # float32 [n,6] local AABB (minExtend xyz, maxExtend xyz); None = default +-1

    @property
    def n(self) -> int:
        return int(self.parent.shape[0])

    def depth(self) -> np.ndarray:
        d = np.zeros(self.n, dtype=np.int32)
        # parents precede children, so one forward sweep per level is enough; do it vectorised per level
        has_parent = self.parent >= 0
        idx = np.nonzero(has_parent)[0]
        # iterate until stable (max depth passes)
        while True:
            nd = d.copy()
            nd[idx] = d[self.parent[idx]] + 1
            if np.array_equal(nd, d):
                break
            d = nd
        return d


def euler_to_quat(euler_deg: np.ndarray) -> np.ndarray:
    """Quaternion(Vector3 euler) of the reference (Source/Math/RNQuaternion.h:220-242), evaluated in fp64 and
    rounded once to fp32; the result is *input data* for every implementation, so its own rounding is irrelevant."""
    e = np.asarray(euler_deg, dtype=np.float64) * (np.pi / 360.0)
    sy, cy = np.sin(e[:, 0]), np.cos(e[:, 0])
    sp, cp = np.sin(e[:, 1]), np.cos(e[:, 1])
    sr, cr = np.sin(e[:, 2]), np.cos(e[:, 2])
    w = sy * sp * sr + cy * cp * cr
    x = cy * sp * cr + sy * cp * sr
    y = -cy * sp * sr + sy * cp * cr
    z = -sy * sp * cr + cy * cp * sr
    q = np.stack([x, y, z, w], axis=1)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q.astype(np.float32)


def _forest_parents(n_trees: int, branching, depth: int) -> tuple[np.ndarray, np.ndarray]:
    """Complete `branching`-ary trees of `depth` levels, each tree in BFS order, trees back to back.  `branching` may
    be a sequence of depth-1 fan-outs, one per level (SURVEY.md's C1: (9, 10, 10) -> 1 + 9 + 90 + 900 nodes per tree)."""
    fan = [int(branching)] * (depth - 1) if np.isscalar(branching) else [int(b) for b in branching]
    assert len(fan) == depth - 1, "one fan-out per level below the root"
    level_sizes = [1]
    for b in fan:
        level_sizes.append(level_sizes[-1] * b)
    per_tree = sum(level_sizes)
    local_parent = np.empty(per_tree, dtype=np.int64)
    local_level = np.empty(per_tree, dtype=np.int32)
    local_parent[0] = -1
    local_level[0] = 0
    start = 1
    prev_start = 0
    for k in range(1, depth):
        cnt = level_sizes[k]
        local_parent[start:start + cnt] = prev_start + np.arange(cnt) // fan[k - 1]
        local_level[start:start + cnt] = k
        prev_start = start
        start += cnt
    base = (np.arange(n_trees, dtype=np.int64) * per_tree)[:, None]
    parent = np.where(local_parent[None, :] >= 0, local_parent[None, :] + base, -1).reshape(-1)
    level = np.broadcast_to(local_level[None, :], (n_trees, per_tree)).reshape(-1)
    return parent.astype(np.int32), level.copy()


def _ragged_parents(n_target: int, max_depth: int, rng: np.random.Generator) -> tuple[np.ndarray, np.ndarray]:
    """Random forest: branching 1..8 per node, ragged depth <= max_depth, about n_target nodes."""
    parents: list[np.ndarray] = []
    levels: list[np.ndarray] = []
    total = 0
    while total < n_target:
        # one tree, level by level
        tree_parent = [np.array([-1], dtype=np.int64)]
        tree_level = [np.array([0], dtype=np.int32)]
        prev = np.array([0], dtype=np.int64)
        count = 1
        for k in range(1, max_depth):
            nchild = rng.integers(0, 9, size=prev.shape[0])
            if k < 3:
                nchild = np.maximum(nchild, 2)
            cnt = int(nchild.sum())
            if cnt == 0 or total + count + cnt > n_target * 1.05 + 64:
                break
            p = np.repeat(prev, nchild)
            tree_parent.append(p)
            tree_level.append(np.full(cnt, k, dtype=np.int32))
            prev = count + np.arange(cnt, dtype=np.int64)
            count += cnt
        tp = np.concatenate(tree_parent)
        tp = np.where(tp >= 0, tp + total, -1)
        parents.append(tp)
        levels.append(np.concatenate(tree_level))
        total += count
    return np.concatenate(parents).astype(np.int32), np.concatenate(levels)


_LEVEL_EDGE = np.array([600.0, 60.0, 20.0, 5.0, 2.0, 1.0, 0.5, 0.25, 0.125, 0.0625], dtype=np.float64)


def make_scene(n_trees: int = 10, branching: int = 4, depth: int = 8, seed: int = 12345, *,
               ragged_nodes: int = 0, nonuniform_scale: bool = False, random_bounds: bool = False,
               random_groups: bool = False, priorities: bool = False, random_boxes: bool = False,
               transparent_frac: float = 0.0,
               hidden_frac: float = 0.01, nocull_frac: float = 0.01) -> Scene:
    """SURVEY.md §8(d) synthetic scene.  `ragged_nodes` > 0 selects the skewed variant (random branching 1-8)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    if ragged_nodes > 0:
        parent, level = _ragged_parents(ragged_nodes, depth, rng)
    else:
        parent, level = _forest_parents(n_trees, branching, depth)
    n = parent.shape[0]
    edge = _LEVEL_EDGE[np.minimum(level, len(_LEVEL_EDGE) - 1)]
    pos = ((rng.random((n, 3)) - 0.5) * edge[:, None]).astype(np.float32)
    euler = rng.random((n, 3)) * 360.0
    rot = euler_to_quat(euler)
    if nonuniform_scale:
        scale = (0.6 + 0.9 * rng.random((n, 3))).astype(np.float32)
        # a few negative scales: the sphere factor is a *signed* max (RNSphere.h:62-68)
        neg = rng.random(n) < 0.002
        scale[neg] *= -1.0
    else:
        s = (0.8 + 0.4 * rng.random(n)).astype(np.float32)
        scale = np.repeat(s[:, None], 3, axis=1).copy()
    r = rng.random(n)
    flags = np.zeros(n, dtype=np.uint32)
    flags[r < hidden_frac] |= FLAG_HIDDEN
    flags[(r >= hidden_frac) & (r < hidden_frac + nocull_frac)] |= FLAG_NOCULLING
    if random_groups:
        group = rng.integers(0, 8, size=n).astype(np.uint32)
    else:
        group = np.ones(n, dtype=np.uint32)
    prio = np.zeros(n, dtype=np.int32)
    if priorities:
        choice = rng.integers(0, 20, size=n)
        prio[choice == 0] = RENDER_EARLY
        prio[choice == 1] = RENDER_LATE
        prio[choice == 2] = RENDER_SKY
    sphere = np.zeros((n, 4), dtype=np.float32)
    sphere[:, 3] = DEFAULT_RADIUS
    if random_bounds:
        sphere[:, :3] = ((rng.random((n, 3)) - 0.5) * 2.0).astype(np.float32)
        sphere[:, 3] = (0.25 + 3.0 * rng.random(n)).astype(np.float32)
    box = None
    if random_boxes:
        lo = (-0.2 - 2.0 * rng.random((n, 3))).astype(np.float32)
        hi = (0.2 + 2.0 * rng.random((n, 3))).astype(np.float32)
        box = np.concatenate([lo, hi], axis=1)
    if transparent_frac > 0.0:
        prio[rng.random(n) < transparent_frac] = RENDER_TRANSPARENT
    return Scene(parent=parent, pos=pos, scale=scale, rot=rot, flags=flags, group=group, priority=prio, sphere=sphere,
                 box=box)


def config_scene(name: str, seed: int = 12345, trees: int | None = None) -> Scene:
    """The scenes BASELINE.json's configs name: C1 = 10 x (1+9+90+900) = 10 000 nodes, depth 4;
    C2 = 48 complete 4-ary depth-8 trees = 1 048 560; C4 = 768 trees = 16 776 960; C5 = 3 072 trees = 67 107 840."""
    if name == "C1":
        return make_scene(n_trees=trees or 10, branching=(9, 10, 10), depth=4, seed=seed)
    n = {"C2": 48, "C4": 768, "C5": 3072}[name]
    return make_scene(n_trees=trees or n, branching=4, depth=8, seed=seed)


MESH_DESC_DTYPE = np.dtype([("kind", "<u4"), ("p", "<f4", 4)])   # 0: Mesh::WithColoredCube(size xyz)


def make_occlusion_scene(seed: int = 1, n_small: int = 3000, n_occluders: int = 40, n_groups: int = 12):
    """A scene for the software occlusion pass (SceneBasic::Render with Flags::Occluder nodes, RNSceneBasic.cpp:445-573):
    big box / sphere occluder Entities between a camera at the origin (looking down -z) and a cloud of small nodes, more
    occluders than the 30 the reference keeps, an Occluder-flagged node that is no Entity (never rasterised), sky /
    transparent nodes (never tested), hidden nodes, everything parented to a few slowly moving group roots.
    Returns (Scene with local boxes, MESH_DESC_DTYPE records of three cubes, [(vertices, indices)] of one custom mesh with
    id 3 - an octahedron whose first index is not 0 -, mesh_of_node i32 [n] with -1 = plain node)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    meshes = np.zeros(3, dtype=MESH_DESC_DTYPE)
    meshes["p"][0] = (6.0, 5.0, 0.5, 0.0)
    meshes["p"][1] = (2.0, 8.0, 2.0, 0.0)
    meshes["p"][2] = (12.0, 1.5, 1.0, 0.0)
    octa_v = np.float32([[4, 0, 0], [-4, 0, 0], [0, 4, 0], [0, -4, 0], [0, 0, 4], [0, 0, -4]])
    octa_i = np.uint16([2, 4, 0, 2, 0, 5, 2, 5, 1, 2, 1, 4, 3, 0, 4, 3, 5, 0, 3, 1, 5, 3, 4, 1])
    custom = [(octa_v, octa_i)]
    half = np.float32([[6.0, 5.0, 0.5], [2.0, 8.0, 2.0], [12.0, 1.5, 1.0], [4.0, 4.0, 4.0]])
    n = n_groups + n_occluders + n_small
    parent = np.full(n, -1, np.int32)
    pos = np.zeros((n, 3), np.float32); scale = np.ones((n, 3), np.float32)
    rot = np.zeros((n, 4), np.float32); rot[:, 3] = 1.0
    flags = np.zeros(n, np.uint32); group = np.ones(n, np.uint32); prio = np.zeros(n, np.int32)
    sphere = np.zeros((n, 4), np.float32); box = np.zeros((n, 6), np.float32)
    mesh_of = np.full(n, -1, np.int32)
    # group roots: small offsets and rotations
    g = np.arange(n_groups)
    pos[g] = ((rng.random((n_groups, 3)) - 0.5) * (8.0, 4.0, 8.0)).astype(np.float32)
    rot[g] = euler_to_quat((rng.random((n_groups, 3)) - 0.5) * 10.0)
    scale[g] = (0.9 + 0.2 * rng.random(n_groups)).astype(np.float32)[:, None]
    box[g] = (-1, -1, -1, 1, 1, 1); sphere[g, 3] = DEFAULT_RADIUS
    flags[g] = FLAG_HIDDEN
    # occluders
    o = n_groups + np.arange(n_occluders)
    parent[o] = rng.integers(0, n_groups, n_occluders)
    z = -(12.0 + 110.0 * rng.random(n_occluders))
    pos[o, 0] = ((rng.random(n_occluders) - 0.5) * 1.6 * -z).astype(np.float32)
    pos[o, 1] = ((rng.random(n_occluders) - 0.5) * 0.8 * -z).astype(np.float32)
    pos[o, 2] = z.astype(np.float32)
    rot[o] = euler_to_quat((rng.random((n_occluders, 3)) - 0.5) * 30.0)
    m = rng.integers(0, 4, n_occluders)
    mesh_of[o] = m
    box[o, :3] = -half[m]; box[o, 3:] = half[m]
    sphere[o, 3] = np.sqrt((half[m] ** 2).sum(axis=1)).astype(np.float32)
    flags[o] = FLAG_OCCLUDER
    mesh_of[o[3]] = -1                                   # an Occluder that is no Entity: tested, never rasterised
    flags[o[5]] &= ~np.uint32(FLAG_OCCLUDER)             # an Entity that is no Occluder: an ordinary tested node
    # the small nodes: inside and around the view cone, most of them behind the occluders
    k = n_groups + n_occluders + np.arange(n_small)
    parent[k] = rng.integers(0, n_groups, n_small)
    z = -(3.0 + 300.0 * rng.random(n_small) ** 0.7)
    pos[k, 0] = ((rng.random(n_small) - 0.5) * 2.2 * -z).astype(np.float32)
    pos[k, 1] = ((rng.random(n_small) - 0.5) * 1.2 * -z).astype(np.float32)
    pos[k, 2] = z.astype(np.float32)
    rot[k] = euler_to_quat(rng.random((n_small, 3)) * 360.0)
    sz = (0.05 + 1.2 * rng.random((n_small, 3)) ** 2).astype(np.float32)
    box[k, :3] = -sz; box[k, 3:] = sz
    sphere[k, 3] = np.sqrt((sz ** 2).sum(axis=1)).astype(np.float32)
    r = rng.random(n_small)
    prio[k[r < 0.03]] = RENDER_SKY
    prio[k[(r >= 0.03) & (r < 0.08)]] = RENDER_TRANSPARENT
    prio[k[(r >= 0.08) & (r < 0.10)]] = RENDER_EARLY
    flags[k[(r >= 0.10) & (r < 0.12)]] |= FLAG_HIDDEN
    flags[k[(r >= 0.12) & (r < 0.14)]] |= FLAG_NOCULLING
    tiny = k[(r >= 0.14) & (r < 0.18)]                   # sub-pixel boxes far away: fail the size test
    box[tiny] *= np.float32(0.01)
    return (Scene(parent=parent, pos=pos, scale=scale, rot=rot, flags=flags, group=group, priority=prio, sphere=sphere, box=box),
            meshes, custom, mesh_of)


def camera_render_order(cam_flags) -> list:
    """The order in which SceneBasic::Render visits the cameras of a scene (RNSceneBasic.cpp:420-439): three passes
    (RenderEarly, neither, RenderLate) over a list that AddNode fills with PushFront (:719), i.e. newest first."""
    early, late, norender = CAM_RENDER_EARLY, CAM_RENDER_LATE, CAM_NO_RENDER
    out = []
    for want in (0, 1, 2):
        for c in reversed(range(len(cam_flags))):
            f = int(cam_flags[c])
            if f & norender:
                continue
            cls = 0 if f & early else (2 if f & late else 1)
            if cls == want:
                out.append(c)
    return out


def flat_scene(n: int, seed: int = 1) -> Scene:
    """n parentless nodes (edge case: no hierarchy at all)."""
    return make_scene(n_trees=n, branching=1, depth=1, seed=seed)


def chain_scene(n: int, seed: int = 1) -> Scene:
    """One chain of n nodes (edge case: depth == n)."""
    return make_scene(n_trees=1, branching=1, depth=n, seed=seed)


# ---------------------------------------------------------------------------------------------------------------
# Cameras: plain parameter records; the frustum itself is built by rn_camera_build_frustum (product) or by the
# reference's Camera::PostUpdate (harness).
# ---------------------------------------------------------------------------------------------------------------
CAMERA_DTYPE = np.dtype([
    ("flags", "<u4"), ("render_group", "<u4"),
    ("pos", "<f4", 3), ("rot", "<f4", 4), ("scale", "<f4", 3),
    ("fov", "<f4"), ("aspect", "<f4"), ("clip_near", "<f4"), ("clip_far", "<f4"),
    ("ortho", "<f4", 4),           # left, right, bottom, top
    ("plane_offsets", "<f4", 4),   # top, bottom, left, right (Camera::SetFrustumPlaneOffset)
    ("has_projection", "<u4"), ("projection", "<f4", 16),
])
assert CAMERA_DTYPE.itemsize == 164


def make_camera(pos=(0.0, 0.0, 0.0), euler=(0.0, 0.0, 0.0), *, fov=70.0, aspect=16.0 / 9.0, near=0.1, far=500.0,
                flags=0, render_group=1, ortho=None, plane_offsets=(0, 0, 0, 0), projection=None,
                scale=(1.0, 1.0, 1.0)) -> np.ndarray:
    """Defaults follow Camera::Initialize (Source/Scene/RNCamera.cpp:60-64)."""
    c = np.zeros((), dtype=CAMERA_DTYPE)
    c["flags"] = flags
    c["render_group"] = render_group
    c["pos"] = pos
    c["rot"] = euler_to_quat(np.asarray([euler], dtype=np.float64))[0]
    c["scale"] = scale
    c["fov"], c["aspect"], c["clip_near"], c["clip_far"] = fov, aspect, near, far
    c["ortho"] = ortho if ortho is not None else (-100.0, 100.0, -100.0, 100.0)
    c["plane_offsets"] = plane_offsets
    if projection is not None:
        c["has_projection"] = 1
        c["projection"] = np.asarray(projection, dtype=np.float32).reshape(16)
    return c


def eye_projection(tan_left, tan_right, tan_down, tan_up, near, far) -> np.ndarray:
    """Asymmetric reverse-Z eye projection laid out the way the reference's OpenXR window builds the matrices it
    hands to Camera::SetProjectionMatrix (Modules/OpenXR/RNOpenXRWindow.cpp:1828-1865, used by
    Modules/VRWrapper/RNVRCamera.cpp:235).  tan_left / tan_down are negative.  The matrix is *input data* shared
    by all implementations, so its own rounding is irrelevant."""
    f = np.float32
    tw = f(f(tan_right) - f(tan_left))
    th = f(f(tan_up) - f(tan_down))
    m = np.zeros(16, dtype=np.float32)
    m[0] = f(2.0) / tw
    m[5] = f(2.0) / th
    m[8] = f(f(tan_right) + f(tan_left)) / tw
    m[9] = f(f(tan_up) + f(tan_down)) / th
    m[10] = f(f(far) / f(f(far) - f(near))) - f(1.0)
    m[11] = -1.0
    m[14] = f(f(far) * f(near)) / f(f(far) - f(near))
    m[15] = 0.0
    return m


def six_view_rig(seed: int = 7) -> np.ndarray:
    """Config C4: 2 eye cameras (+-0.032 m, asymmetric projection through SetProjectionMatrix; reference
    Modules/VRWrapper/RNVRCamera.cpp:88-89) + 4 orthogonal shadow-cascade cameras.  The cascade cameras are
    parameterised the way Light::UpdateShadows leaves them (orthogonal, +-dist extents, looking along the light
    direction; Source/Scene/RNLight.cpp:475-529, Source/Scene/RNCamera.cpp:240-287) with split distances blended
    lin/log (RNLight.cpp:505-514) - the numbers below are fixed inputs shared by all implementations."""
    cams = []
    for sx, (lt, rt) in ((-0.032, (-1.25, 1.05)), (0.032, (-1.05, 1.25))):
        proj = eye_projection(lt, rt, -1.15, 1.15, 0.1, 500.0)
        cams.append(make_camera(pos=(sx, 1.7, 0.0), euler=(20.0, -5.0, 0.0), projection=proj))
    near, far, splits, factor = 0.1, 500.0, 4, 0.9
    edges = []
    for i in range(splits + 1):
        lin = near + (far - near) * i / splits
        log = near * (far / near) ** (i / splits)
        edges.append(lin * (1.0 - factor) + log * factor)
    for i in range(splits):
        centre = 0.5 * (edges[i] + edges[i + 1])
        dist = max(0.75 * (edges[i + 1] - edges[i]), 40.0)
        # light looks down at 50 degrees; camera sits 250 units up-light from the split centre
        cams.append(make_camera(pos=(0.3 * centre, 190.0, -centre + 160.0), euler=(0.0, -50.0, 0.0),
                                flags=CAM_ORTHOGONAL, near=1.0, far=600.0,
                                ortho=(-dist, dist, -dist, dist)))
    return np.stack(cams)


# rn_shadow_desc (include/rn_cuda.h): ShadowParameter of a directional Light + the light's world rotation
SHADOW_DTYPE = np.dtype([
    ("n_splits", "<u4"), ("split_max_distance", "<f4", 8), ("resolution", "<f4"), ("clip_near", "<f4"), ("clip_far", "<f4"),
    ("directional_shadow_distance", "<f4"), ("distance_blend_factor", "<f4"), ("max_shadow_dist", "<f4"),
    ("light_rot", "<f4", 4),
])
assert SHADOW_DTYPE.itemsize == 76


def make_shadow(light_euler=(0.0, -50.0, 0.0), n_splits=4, resolution=1024.0, clip_near=0.0, clip_far=100000.0,
                directional_shadow_distance=5000.0, distance_blend_factor=0.05, max_shadow_dist=1.0,
                split_max_distance=None) -> np.ndarray:
    """Defaults = ShadowParameter(Camera *target, 1024) (Source/Scene/RNLight.h:50-64): four splits without a maxDistance."""
    s = np.zeros((), dtype=SHADOW_DTYPE)
    s["n_splits"] = n_splits
    if split_max_distance is not None:
        s["split_max_distance"][:n_splits] = split_max_distance
    s["resolution"] = resolution
    s["clip_near"], s["clip_far"] = clip_near, clip_far
    s["directional_shadow_distance"] = directional_shadow_distance
    s["distance_blend_factor"] = distance_blend_factor
    s["max_shadow_dist"] = max_shadow_dist
    s["light_rot"] = euler_to_quat(np.asarray([light_euler], dtype=np.float64))[0]
    return s


def vr_cascade_rig(light_euler=(30.0, -50.0, 0.0)) -> np.ndarray:
    """Config C4 with the cascade cameras the REFERENCE derives: the two eye cameras of six_view_rig + the four
    orthographic cameras Light::UpdateShadows / Camera::MakeShadowSplit produce for a directional light with the
    default ShadowParameter(target) and the head camera as shadow target (rn_light_build_cascades; pinned against the
    reference by tests/golden/cascades_basic.npz)."""
    from . import capi
    eyes = six_view_rig()[:2]
    head = make_camera(pos=(0.0, 1.7, 0.0), euler=(20.0, -5.0, 0.0))
    casc, _ = capi.build_cascades(head, make_shadow(light_euler))
    return np.concatenate([eyes, casc])


# ---------------------------------------------------------------------------------------------------------------
# Particle emitters (GenericParticleEmitter parameters; defaults from Source/Scene/RNParticleEmitter.cpp:379-395)
# ---------------------------------------------------------------------------------------------------------------
EMITTER_DTYPE = np.dtype([
    ("seed", "<u4"), ("max_particles", "<u4"), ("spawn_rate", "<f4"),
    ("life_span", "<f4", 2),
    ("start_color", "<f4", 4), ("end_color", "<f4", 4),
    ("start_size", "<f4", 2), ("end_size", "<f4", 2), ("start_rotation", "<f4", 2), ("end_rotation", "<f4", 2),
    ("gravity", "<f4", 3), ("velocity", "<f4", 3), ("vel_min", "<f4", 3), ("vel_max", "<f4", 3),
    ("pos_min", "<f4", 3), ("pos_max", "<f4", 3),
    ("flags", "<u4"), ("max_particles_soft", "<u4"), ("world_position", "<f4", 3), ("world_scale", "<f4", 3),
])
assert EMITTER_DTYPE.itemsize == 188
EMITTER_NONLOCAL, EMITTER_SOFT_MAX = 1, 2


def upgrade_emitters(e: np.ndarray) -> np.ndarray:
    """Emitter records of an older fixture (without flags / soft maximum / world transform) in the current layout."""
    e = np.asarray(e)
    if e.dtype == EMITTER_DTYPE:
        return e
    out = np.zeros(e.shape, dtype=EMITTER_DTYPE)
    for name in e.dtype.names:
        out[name] = e[name]
    out["world_scale"] = 1.0
    return out


def make_emitters(n: int, *, max_particles: int = 4000, spawn_rate: float = 3.0 / 4000.0, seed_base: int = 1000,
                  vary: bool = False, seed: int = 99) -> np.ndarray:
    e = np.zeros(n, dtype=EMITTER_DTYPE)
    e["seed"] = seed_base + np.arange(n)
    e["max_particles"] = max_particles
    e["spawn_rate"] = np.float32(spawn_rate)
    e["life_span"] = (2.0, 4.0)
    e["start_color"] = (1.0, 1.0, 1.0, 1.0)
    e["end_color"] = (1.0, 1.0, 1.0, 0.0)
    e["start_size"] = (0.5, 1.5)
    e["end_size"] = (1.5, 2.5)
    e["start_rotation"] = (0.0, 0.0)
    e["end_rotation"] = (0.0, 0.0)
    e["gravity"] = (0.0, -0.1, 0.0)
    e["velocity"] = (0.0, 0.5, 0.0)
    e["vel_min"] = (-0.5, -0.5, -0.5)
    e["vel_max"] = (0.5, 0.5, 0.5)
    e["pos_min"] = (-0.5, -0.5, -0.5)
    e["pos_max"] = (0.5, 0.5, 0.5)
    e["world_scale"] = 1.0
    if vary:
        rng = np.random.Generator(np.random.PCG64(seed))
        e["life_span"][:, 0] = 0.2 + rng.random(n).astype(np.float32)
        e["life_span"][:, 1] = e["life_span"][:, 0] + rng.random(n).astype(np.float32) * 2
        e["start_rotation"] = (-45.0, 45.0)
        e["end_rotation"] = (90.0, 270.0)
        e["start_color"] = rng.random((n, 4)).astype(np.float32)
        e["end_color"] = rng.random((n, 4)).astype(np.float32)
        e["gravity"] = (rng.random((n, 3)) - 0.5).astype(np.float32)
        e["vel_min"] = (1.0, -1.0, 0.5)   # deliberately min > max on x and z: the reference swaps (RNRandom.cpp:250-252)
        e["vel_max"] = (-1.0, 1.0, -0.5)
        e["spawn_rate"] = (0.0004 + 0.01 * rng.random(n)).astype(np.float32)
        e["max_particles"] = rng.integers(1, max_particles + 1, size=n)
    return e
