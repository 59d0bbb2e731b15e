"""GPU parity: rn_particles_step against the oracle (bit-exact state in the reference's vector order)."""
import numpy as np
import pytest

from oracle import oracle
from rayne_b200 import capi, scenegen as sg
from tests.helpers import same_floats

pytestmark = pytest.mark.gpu


def cuda_state(h, k):
    pos = h.particles_download(k, capi.PF_POSITION); vel = h.particles_download(k, capi.PF_VELOCITY)
    tim = h.particles_download(k, capi.PF_TIMING); rng = h.particles_download(k, capi.PF_RANGES)
    der = h.particles_download(k, capi.PF_DERIVED)
    # oracle record columns: pos3 vel3 time life dur s0 s1 r0 r1 rot size2 color4
    return np.concatenate([pos, vel, tim, rng, der], axis=1)


@pytest.mark.parametrize("vary", [False, True])
@pytest.mark.parametrize("steps", [1, 10, 240])
def test_particles_parity(vary, steps):
    em = sg.make_emitters(7, max_particles=700, spawn_rate=0.0021, vary=vary)
    delta = np.float32(1.0 / 60.0)
    o = oracle.Emitters(em)
    o.step(delta, steps)
    with capi.SceneHandle(max_nodes=1, max_emitters=8, max_particles_per_emitter=700) as h:
        h.particles_configure(em)
        for _ in range(steps):
            h.particles_step(float(delta))
        counts, times = h.particles_counts()
        for k in range(len(em)):
            rec, t = o.read(k)
            assert counts[k] == rec.shape[0], (k, counts[k], rec.shape[0])
            assert same_floats(times[k], t)
            assert same_floats(cuda_state(h, k), rec[:, :20]), f"emitter {k}"


def test_emitter_modes_match_golden():
    """!GetIsLocal() emitters (world-space spawn, per-particle gravity) and SetMaxParticlesSoft on running emitters -
    one capped below its hard maximum, one whose soft maximum falls below its live count, so that the reference's
    unsigned `soft - size` wraps and it spawns past max_particles (RNParticleEmitter.cpp:128-132,204-205,428-437).
    Bit-exact against the reference's own records (tests/golden/particles_nonlocal_soft.npz) and the oracle."""
    import os
    from tests.helpers import GOLDEN_DIR
    from tests.test_cpu_oracle_golden import emitters_without_late_soft_max, run_emitter_fixture
    g = np.load(os.path.join(GOLDEN_DIR, "particles_nonlocal_soft.npz"))
    em = sg.upgrade_emitters(g["in_emitters"])
    start = emitters_without_late_soft_max(g)
    steps, soft_after, delta = int(g["in_steps"]), int(g["in_soft_after"]), float(g["in_delta"])
    o = oracle.Emitters(start)
    run_emitter_fixture(o, g)
    with capi.SceneHandle(max_nodes=1, max_emitters=4, max_particles_per_emitter=2000) as h:
        h.particles_configure(start)
        for s in range(steps):
            if s == soft_after:
                for k in np.nonzero(em["flags"] & sg.EMITTER_SOFT_MAX)[0]:
                    h.particles_set_soft_max([int(em["max_particles_soft"][k])], first=int(k))
            h.particles_step(delta)
        counts, times = h.particles_counts()
        for k in range(len(em)):
            ref = g[f"ref_rec{k}"]
            assert counts[k] == ref.shape[0], (k, counts[k], ref.shape[0])
            assert same_floats(times[k], g[f"ref_time{k}"])
            assert same_floats(cuda_state(h, k), ref[:, :20]), f"emitter {k}"
            assert same_floats(h.particles_download(k, capi.PF_GRAVITY), ref[:, 20:23]), f"gravity of emitter {k}"
            rec, _ = o.read(k)
            assert same_floats(rec[:, :23], ref[:, :23])
        assert counts[3] > em["max_particles"][3]        # the wrap really happened


def test_emitter_transform_update_and_capacity_overflow():
    """rn_particles_set_emitter_transforms moves a non-local emitter between steps; an emitter driven past the segment
    capacity by the soft-maximum wrap is reported (RN_ERR_CAPACITY), not silently truncated."""
    em = sg.make_emitters(2, max_particles=200, spawn_rate=0.001)
    em["flags"][0] = sg.EMITTER_NONLOCAL
    em["world_scale"][0] = (2, 2, 2)
    delta = np.float32(1.0 / 60.0)
    o = oracle.Emitters(em)
    with capi.SceneHandle(max_nodes=1, max_emitters=2, max_particles_per_emitter=260) as h:
        h.particles_configure(em)
        for s in range(30):
            if s == 10:
                h.particles_set_emitter_transforms(world_position=[(5, 6, 7)], world_scale=[(0.5, 3, 1)], first=0)
                o.set_transform(0, (5, 6, 7), (0.5, 3, 1))
            h.particles_step(float(delta)); o.step(delta, 1)
        counts, _ = h.particles_counts()
        for k in range(2):
            rec, _ = o.read(k)
            assert counts[k] == rec.shape[0]
            assert same_floats(cuda_state(h, k), rec[:, :20])
            assert same_floats(h.particles_download(k, capi.PF_GRAVITY), rec[:, 20:23])
        h.particles_set_soft_max([3], first=1)          # far below the 200 live particles: the wrap
        for _ in range(40):
            h.particles_step(float(delta))
        with pytest.raises(capi.RNError):
            h.particles_counts()


def test_mt_wraps_and_big_spawn():
    """spawn bursts larger than one spawn batch and more than 624 draws per step (MT regeneration on device)."""
    em = sg.make_emitters(3, max_particles=3000, spawn_rate=0.00002)
    em["life_span"] = (0.05, 0.2)
    delta = np.float32(1.0 / 60.0)
    o = oracle.Emitters(em); o.step(delta, 12)
    with capi.SceneHandle(max_nodes=1, max_emitters=3, max_particles_per_emitter=3000) as h:
        h.particles_configure(em)
        for _ in range(12):
            h.particles_step(float(delta))
        counts, _ = h.particles_counts()
        for k in range(3):
            rec, _ = o.read(k)
            assert counts[k] == rec.shape[0]
            assert same_floats(cuda_state(h, k), rec[:, :20])


def test_depth_sort_extension():
    em = sg.make_emitters(4, max_particles=1500, spawn_rate=0.001)
    delta = np.float32(1.0 / 60.0)
    cam = (3.0, 1.0, -2.0)
    o = oracle.Emitters(em); o.step(delta, 100)
    with capi.SceneHandle(max_nodes=1, max_emitters=4, max_particles_per_emitter=1500) as h:
        h.particles_configure(em)
        for _ in range(100):
            h.particles_step(float(delta))
        h.particles_sort(cam)
        for k in range(4):
            idx, key = o.sort(k, cam)
            got = h.particles_download(k, capi.PF_SORTED)
            assert np.array_equal(got, idx)
