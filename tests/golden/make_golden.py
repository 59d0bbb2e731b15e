"""Generates tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref/ref_harness_strict, built by
oracle/build_ref.sh from /root/reference) on small seeded scenes and emitter sets.

    python -m tests.golden.make_golden

Each fixture stores the inputs (creation-order node columns, camera records, emitter records) and the reference's
outputs (render-list order, world TRS + matrices, the 28 frustum floats per camera, the per-camera visibility
lists; per-emitter particle records after K steps).  The reference publishes no golden vectors for this path
(SURVEY.md §4), so these files - produced by the reference's own code - are the pin of the oracle and of the
CUDA path on machines where /root/reference does not exist.
"""
from __future__ import annotations

import os

import numpy as np

from oracle import refio
from rayne_b200 import scenegen as sg
from tests.helpers import sort_cameras, standard_cameras

HERE = os.path.dirname(os.path.abspath(__file__))

SCENES = {
    "scene_uniform": dict(n_trees=3, branching=3, depth=5, seed=101),
    "scene_nonuniform_groups_prio": dict(n_trees=5, branching=4, depth=5, seed=102, nonuniform_scale=True,
                                         random_bounds=True, random_groups=True, priorities=True),
    "scene_ragged": dict(ragged_nodes=1500, depth=8, seed=103, random_bounds=True),
    "scene_deep_chain": dict(n_trees=2, branching=1, depth=40, seed=104),
    # local AABBs, RenderTransparent nodes (post-cull sorts), sort cameras; also run through SceneWithVisibilityLists
    "scene_boxes_transparent": dict(n_trees=6, branching=4, depth=5, seed=105, nonuniform_scale=True, random_bounds=True,
                                    random_boxes=True, priorities=True, transparent_frac=0.25),
}
PARTICLES = {
    "particles_default": (dict(n=3, max_particles=300, spawn_rate=0.004), 240),
    "particles_varied": (dict(n=4, max_particles=500, spawn_rate=0.0021, vary=True), 90),
    "particles_burst": (dict(n=2, max_particles=1500, spawn_rate=0.00002), 6),
}
SCENE_COLS = ("parent", "pos", "scale", "rot", "flags", "group", "priority", "sphere")


def main() -> None:
    assert refio.have_harness("strict"), "run oracle/build_ref.sh strict first"
    for name, kw in SCENES.items():
        scene = sg.make_scene(**kw)
        cams = standard_cameras()
        if scene.box is not None:
            cams = np.concatenate([cams, sort_cameras()])
        opts = refio.OPT_INVERSE | (refio.OPT_BOX if scene.box is not None else 0)
        ref = refio.run_scene(scene, cams, options=opts | refio.OPT_INSTANCES)
        out = {f"in_{k}": getattr(scene, k) for k in SCENE_COLS}
        if scene.box is not None:
            out["in_box"] = scene.box
        out["in_cameras"] = cams
        for k in ("order", "wpos", "wrot", "wscale", "wmat", "winv", "aabb", "sphere", "frustum", "cam_flags", "cam_group"):
            out[f"ref_{k}"] = ref[k]
        for c, lst in enumerate(ref["lists"]):
            out[f"ref_list{c}"] = lst
        # per camera GetViewMatrix and projection * view; for camera 0 view * world and projView * world of every node
        for k in ("view", "projview", "mv", "mvp"):
            out[f"ref_{k}"] = ref[k]
        # the same scene through SceneWithVisibilityLists (default volume): A10 order + its lists
        vis = refio.run_scene(scene, cams, options=opts | refio.OPT_VISLISTS)
        out["ref_vis_order"] = vis["order"]
        for c, lst in enumerate(vis["lists"]):
            out[f"ref_vis_list{c}"] = lst
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, scene.n, [len(l) for l in ref["lists"]])
    # SceneWithVisibilityLists with user volumes (AddVolume + the camera-position choice of Render): three
    # AxisAlignedBoxVolumes, the second of the overlapping ones never chosen where the first contains the camera
    scene = sg.make_scene(n_trees=6, branching=3, depth=5, seed=77, random_bounds=True, random_groups=True)
    vol = np.random.default_rng(5).choice([0, 0, 0, 1, 2, 3], size=scene.n).astype(np.uint32)
    bounds = np.float32([[-5, -5, -5, 5, 5, 5], [5, 0, -10, 20, 10, 5], [-1000, -1000, -1000, 1000, 1000, 1000]])
    cams = np.concatenate([standard_cameras(), np.stack([sg.make_camera(pos=(5000, 0, 0), far=9000)])])
    ref = refio.run_scene(scene, cams, options=refio.OPT_VISLISTS | refio.OPT_VOLUMES, volumes=(bounds, vol))
    out = {f"in_{k}": getattr(scene, k) for k in SCENE_COLS}
    out.update(in_cameras=cams, in_volume=vol, in_bounds=bounds, ref_default_order=ref["order"])
    for c, lst in enumerate(ref["lists"]):
        out[f"ref_list{c}"] = lst
    np.savez_compressed(os.path.join(HERE, "volumes_basic.npz"), **out)
    print("volumes_basic", scene.n, [len(l) for l in ref["lists"]])
    for name, (kw, steps) in PARTICLES.items():
        kw = dict(kw)
        em = sg.make_emitters(kw.pop("n"), **kw)
        if name == "particles_burst":
            em["life_span"] = (0.05, 0.2)
        ref = refio.run_particles(em, steps)
        out = {"in_emitters": em, "in_steps": np.int32(steps), "in_delta": np.float32(1.0 / 60.0)}
        for k, r in enumerate(ref):
            out[f"ref_rec{k}"] = r["rec"]
            out[f"ref_time{k}"] = r["time"]
        if name == "particles_varied":
            # P6: the render mesh of the same emitters - plain, rendered inversed, rolled
            for tag, popt in (("", 0), ("_inv", refio.POPT_INVERSED), ("_roll", refio.POPT_ROLL)):
                mref = refio.run_particles(em, steps, options=refio.POPT_MESH | popt)
                for k, r in enumerate(mref):
                    for key in ("mesh_pos", "mesh_color", "mesh_uv0", "mesh_uv1", "mesh_idx"):
                        out[f"ref_{key}{tag}{k}"] = r[key]
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, [r["count"] for r in ref])
    # emitter modes: !GetIsLocal() (world-space spawn, per-particle gravity) and SetMaxParticlesSoft applied mid-run - one
    # emitter capped below its hard maximum, one whose soft maximum drops below its live count (the unsigned wrap of
    # RNParticleEmitter.cpp:204-205 lets it spawn past max_particles)
    em, steps, soft_after = nonlocal_soft_emitters()
    ref = refio.run_particles(em, steps, soft_after_steps=soft_after)
    out = {"in_emitters": em, "in_steps": np.int32(steps), "in_delta": np.float32(1.0 / 60.0), "in_soft_after": np.int32(soft_after)}
    for k, r in enumerate(ref):
        out[f"ref_rec{k}"] = r["rec"]
        out[f"ref_time{k}"] = r["time"]
    np.savez_compressed(os.path.join(HERE, "particles_nonlocal_soft.npz"), **out)
    print("particles_nonlocal_soft", [r["count"] for r in ref])
    make_cascades()
    make_occlusion()


def cascade_cases():
    """(target camera, shadow parameters) pairs of cascades_basic.npz: the default ShadowParameter(target) of a VR head
    camera, and a three-split setup with explicit maxDistance on two splits, another resolution / blend / range."""
    head = sg.make_camera(pos=(0.0, 1.7, 0.0), euler=(20.0, -5.0, 0.0))
    other = sg.make_camera(pos=(-40.0, 12.0, 7.5), euler=(110.0, 15.0, -5.0), fov=55.0, aspect=1.25, near=0.5, far=900.0)
    return [(head, sg.make_shadow((30.0, -50.0, 0.0))),
            (other, sg.make_shadow((0.0, -90.0, 10.0), n_splits=3, split_max_distance=(5.0, 0.0, 300.0), resolution=2048.0,
                                   distance_blend_factor=0.4, max_shadow_dist=0.5, clip_near=1.0, clip_far=20000.0,
                                   directional_shadow_distance=800.0))]


def make_cascades() -> None:
    out = {}
    for k, (target, shadow) in enumerate(cascade_cases()):
        ref = refio.run_cascades(target, shadow)
        out[f"in_target{k}"] = target
        out[f"in_shadow{k}"] = shadow
        for key, val in ref.items():
            out[f"ref_{key}{k}"] = val
        print("cascades", k, ref["ortho"][:, 1])
    np.savez_compressed(os.path.join(HERE, "cascades_basic.npz"), **out)


OCCLUSION_CASE = dict(seed=3, n_small=900, n_occluders=40, n_groups=8, frames=60, rng_seed=4242, frame_size=(1280.0, 720.0),
                      root_translate=(0.02, 0.0, 0.0), camera_translate=(0.0, 0.0, -0.25))


def occlusion_inputs():
    kw = OCCLUSION_CASE
    scene, meshes, custom, mesh_of = sg.make_occlusion_scene(seed=kw["seed"], n_small=kw["n_small"], n_occluders=kw["n_occluders"],
                                                             n_groups=kw["n_groups"])
    # camera 1 is added last, so SceneBasic::Render visits it first; camera 2 has no occluder in view (plain path)
    cams = np.stack([sg.make_camera(), sg.make_camera(pos=(4, 1, 10), euler=(8, -3, 0), fov=60),
                     sg.make_camera(pos=(0, 0, 30), euler=(180, 0, 0), far=200)])
    return scene, meshes, custom, mesh_of, cams


def make_occlusion() -> None:
    """Software occlusion culling (RNSceneBasic.cpp:445-573) over 60 frames: lists, counters, depth buffers and the jitter
    the reference drew, with occluder Entities built from the reference's own Mesh / Model classes."""
    kw = OCCLUSION_CASE
    scene, meshes, custom, mesh_of, cams = occlusion_inputs()
    ref = refio.run_occlusion(scene, cams, meshes, mesh_of, frames=kw["frames"], seed=kw["rng_seed"], frame_size=kw["frame_size"],
                              root_translate=kw["root_translate"], camera_translate=kw["camera_translate"], custom_meshes=custom)
    out = {f"in_{k}": getattr(scene, k) for k in SCENE_COLS}
    out.update(in_box=scene.box, in_cameras=cams, in_mesh_of=mesh_of, ref_order=ref["order"], ref_tri_ofs=ref["tri_ofs"], ref_tris=ref["tris"])
    nf, nc = kw["frames"], len(cams)
    out["ref_used"] = np.array([[fr["cams"][c]["used"] for c in range(nc)] for fr in ref["frames"]], np.uint8)
    out["ref_offset"] = np.array([[fr["cams"][c]["offset"] for c in range(nc)] for fr in ref["frames"]], np.float32)
    lists = [fr["cams"][c]["list"] for fr in ref["frames"] for c in range(nc)]
    out["ref_list_ofs"] = np.concatenate([[0], np.cumsum([len(l) for l in lists])]).astype(np.uint32)
    out["ref_lists"] = np.concatenate(lists).astype(np.uint16)
    out["ref_counters"] = np.stack([fr["counters"] for fr in ref["frames"]]).astype(np.uint16)
    out["ref_depth"] = np.stack([fr["depth"] for fr in ref["frames"]])
    out["ref_view"] = np.array([[fr["cams"][c]["view"] for c in range(nc)] for fr in ref["frames"]], np.float32)
    out["ref_projection"] = np.array([ref["frames"][0]["cams"][c]["projection"] for c in range(nc)], np.float32)
    np.savez_compressed(os.path.join(HERE, "occlusion_basic.npz"), **out)
    print("occlusion_basic", scene.n, "used", out["ref_used"].sum(axis=0).tolist(), "list sizes first/last frame",
          [len(ref["frames"][0]["cams"][c]["list"]) for c in range(nc)], [len(ref["frames"][-1]["cams"][c]["list"]) for c in range(nc)])


def nonlocal_soft_emitters():
    em = sg.make_emitters(4, max_particles=500, spawn_rate=0.0021, vary=True, seed=7)
    em["max_particles"] = (500, 350, 500, 300)
    em["flags"][0] = em["flags"][1] = sg.EMITTER_NONLOCAL
    em["world_position"][0] = (10.5, -3.25, 7.0)
    em["world_scale"][0] = (2.0, 0.5, 1.25)
    em["world_position"][1] = (-100.0, 40.0, 0.125)
    em["world_scale"][1] = (-1.5, 3.0, 0.75)          # a negative component: the size scale is a plain max
    em["flags"][2] = em["flags"][3] = sg.EMITTER_SOFT_MAX
    em["max_particles_soft"][2] = 400
    em["max_particles_soft"][3] = 20
    em["spawn_rate"][2] = em["spawn_rate"][3] = 0.002
    em["life_span"][2] = em["life_span"][3] = (3.0, 4.0)
    return em, 120, 40


if __name__ == "__main__":
    main()
