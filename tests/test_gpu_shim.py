"""The compiled Rayne-side binding (integration/RNSceneCUDA.cpp, a RN::Scene subclass linked against the reference's own
library) against RN::SceneBasic: the same seeded scene in both, a scripted run of setters / AddNode / RemoveNode / AddChild
reparenting / SetRenderPriority; per frame the submissions per camera and every world transform must be identical.
The binary is built where /root/reference exists (oracle/build_ref.sh shim) and travels in oracle/_ref/."""
import json
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

BIN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "shim_test")


@pytest.mark.parametrize("roots,frames", [(60, 25), (400, 17)])
def test_scene_cuda_matches_scene_basic(roots, frames):
    if not os.access(BIN, os.X_OK):
        pytest.skip("oracle/_ref/shim_test not built (needs the reference sources)")
    r = subprocess.run([BIN, str(roots), str(frames)], capture_output=True, text=True, timeout=600)
    line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else ""
    assert line, r.stderr[-2000:]
    out = json.loads(line)
    assert r.returncode == 0 and out["ok"], out
    assert out["list_mismatches"] == 0 and out["matrix_mismatches"] == 0
    assert out["list_entries_compared"] > 1000 and out["matrices_compared"] > roots * frames
    assert out["dirty_rows_uploaded"] > 0           # frames without a hierarchy change went through rn_scene_update_nodes_by_slot
    assert out["incremental_edits"] > 0             # AddNode / RemoveNode went through rn_scene_add_nodes / rn_scene_remove_nodes
    assert out["full_uploads"] < frames             # ... and only reparenting / priority changes re-uploaded the scene
