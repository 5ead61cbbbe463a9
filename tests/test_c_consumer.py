"""The drop-in boundary used from C: tests/c_consumer/consumer.c includes include/pmtech_b200.h as strict C99,
links libpmtech_b200.so and drives it the way a pmtech-side shim would (no Python between the caller and the
library).  CPU: the header compiles warning-free as C, the program links, walks a scene image and is refused a
device scene loudly.  GPU: its output equals the oracle's for the scene pmtech itself saved."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "c_consumer", "consumer.c")
OUT_DIR = os.path.join(ROOT, "tests", "c_consumer", "_build")
LIB_DIR = os.path.join(ROOT, "pmtech_b200", "lib")
F = np.float32


@pytest.fixture(scope="module")
def consumer(lib):
    os.makedirs(OUT_DIR, exist_ok=True)
    exe = os.path.join(OUT_DIR, "consumer")
    cmd = ["gcc", "-std=c99", "-O1", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), SRC,
           "-L", LIB_DIR, "-lpmtech_b200", "-Wl,-rpath," + LIB_DIR, "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


@pytest.fixture(scope="module")
def image_path():
    from test_pmtech_scene_files import _image, load
    f = load("pmtech_sdf_scene")
    scene = dict(f)
    scene["entities"] = f["entities"] & ~np.uint64(1 << 12)  # the lights' pending e_cmp::transform (see test_pmtech_scene_files)
    os.makedirs(OUT_DIR, exist_ok=True)
    path = os.path.join(OUT_DIR, "sdf_replay.pms")
    with open(path, "wb") as fh:
        fh.write(_image(scene, 8))
    return path


def _parse(stdout):
    out = {}
    for line in stdout.splitlines():
        k, _, v = line.partition(" ")
        out[k] = v.split()
    return out


def test_header_is_c99_and_the_program_walks_a_scene_image(consumer, image_path):
    from pmtech_b200 import capi, pms
    r = subprocess.run([consumer, image_path], capture_output=True, text=True, timeout=120)
    got = _parse(r.stdout)
    p = pms.read_pms(image_path)
    assert got["version"] == ["8"] and got["nodes"] == ["23"] and got["cameras"] == ["1"]
    assert [int(x) for x in got["offsets"]] == [int(x) for x in p.info.component_offset]
    if capi.device_count() == 0:
        assert r.returncode == 3, (r.stdout, r.stderr)
        assert got["no_device"][0] == "-3" and "CPU" in " ".join(got["no_device"])  # PMB_ERR_NO_DEVICE, "no CPU path"
    else:
        assert r.returncode == 0, (r.stdout, r.stderr)


@pytest.mark.gpu
def test_c_consumer_frame_matches_the_oracle_and_pmtech(consumer, image_path, oracle):
    from test_pmtech_scene_files import load
    from test_shadow_fit_gpu import _fit
    f = load("pmtech_sdf_scene")
    n = 23
    r = subprocess.run([consumer, image_path], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, (r.stdout, r.stderr)
    got = _parse(r.stdout)
    ent = f["entities"] & ~np.uint64(1 << 12)
    o = oracle.OracleScene(f["parents"], f["local_matrices"], ent, f["state_flags"], f["bounding_volumes"])
    o.update_scene()
    ext = np.asarray(o.renderable_extents, F)
    assert np.array_equal(np.array(got["extents"], F).view(np.uint32), ext.view(np.uint32))
    cam = oracle.OracleCamera()
    cam.create_perspective(60.0, F(16.0) / F(9.0), 0.1, 1000.0)
    cam.look_at((-14.5, 16.5, 26.5), (0, 0, 0))
    cam.update_frustum()
    assert np.array_equal(np.array(got["plane0"], F).view(np.uint32), cam.frustum[:3].view(np.uint32))
    shadow = _fit(oracle, ext, np.array([[0.3, -1.0, 0.2]], F))[0]
    want, _ = oracle.OracleScene(f["parents"], f["local_matrices"], ent, f["state_flags"], f["bounding_volumes"]).frame(
        np.stack([cam.frustum, shadow.frustum]))
    for v in range(2):
        assert np.array_equal(np.array(got[f"list{v}"], np.uint32), want[v]), v
    assert len(want[0]) > 5 and len(want[1]) > 5
    # and the world matrices are pmtech's own
    alive = (f["entities"] & np.uint64(1)) != 0
    for e in np.flatnonzero(alive):
        t = np.array(got[f"world{e}"], F)
        assert np.array_equal(t.view(np.uint32), f["world_matrices"][e, [3, 7, 11]].view(np.uint32)), e


# ---- one C process, several shards: pmb_group ---------------------------------------------------------------------
@pytest.fixture(scope="module")
def group_consumer(lib):
    os.makedirs(OUT_DIR, exist_ok=True)
    exe = os.path.join(OUT_DIR, "group_consumer")
    cmd = ["gcc", "-std=c99", "-O1", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "c_consumer", "group_consumer.c"), "-L", LIB_DIR, "-lpmtech_b200", "-Wl,-rpath," + LIB_DIR,
           "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_group_consumer_builds_and_reports_a_missing_device(group_consumer):
    from pmtech_b200 import capi
    if capi.device_count() == 0:
        r = subprocess.run([group_consumer, "2", "4096"], capture_output=True, text=True, timeout=120)
        assert r.returncode == 3 and "no_device" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_group_of_shards_gathers_the_single_gpu_lists(group_consumer, mode):
    """2 and 3 ranks (real GPUs when the box has them, else shards on one GPU) against 1 rank: same totals, same extents,
    same global lists entry for entry (hash + ends), ascending."""
    n = 50_000
    outs = {}
    for ranks in (1, 2, 3):
        r = subprocess.run([group_consumer, str(ranks), str(n), str(mode)], capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, (ranks, r.stdout, r.stderr)
        outs[ranks] = _parse(r.stdout)
    one = outs[1]
    assert one["overflow"] == ["0"] and int(one["totals"][0]) > 100 and int(one["totals"][1]) > 100
    assert one["count_sums"] == one["totals"]
    if mode != 2:
        assert one["list0"][2] == "1" and one["list1"][2] == "1"  # ascending
    for ranks in (2, 3):
        assert outs[ranks] == one, ranks
