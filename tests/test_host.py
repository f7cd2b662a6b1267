"""CPU tests (no GPU): host-side logic of the product -- XML loader, cfg reader, pyramid plan, the
C ABI surface -- and the multi-rank sharding / gather path on gloo (world size 2)."""
import ctypes
import json
import os
import re
import subprocess
import sys
import textwrap

import numpy as np
import pytest

from conftest import CASCADES, ROOT, all_cv2_haar_cascades, old_format_xml

FIELDS = ["stage_first_tree", "stage_ntrees", "stage_thr", "tree_first_node", "tree_nnodes", "node_rects",
          "node_weights", "node_thr", "node_left", "node_right", "node_left_val", "node_right_val", "node_tilted"]


@pytest.mark.parametrize("name", CASCADES)
def test_xml_loader_matches_independent_parser(ug, cv2ref, name):
    """Every number of the model, bit for bit, against the ElementTree reader in oracle/cascade_xml.py;
    both the old format the reference ships and the new format cv2 ships."""
    import cv2
    from oracle.cascade_xml import load_cascade
    paths = [cv2.data.haarcascades + f"haarcascade_frontalface_{name}.xml"]
    ref = f"/root/reference/data/haarcascade_frontalface_{name}.xml"
    if os.path.exists(ref):
        paths.append(ref)
    for path in paths:
        c = ug.Cascade(path)
        d = c.dump()
        f = load_cascade(path)
        for k in FIELDS:
            assert np.array_equal(d[k], getattr(f, k)), (path, k)
        assert c.info.sum_exact == 1
        assert c.info.old_format == (1 if path == ref else 0)
        assert (c.info.win_w, c.info.nstages) == (f.win_w, f.nstages)


def test_xml_loader_reads_every_cv2_haar_cascade(ug):
    """8(f) N4: all 17 Haar files cv2 ships load (tilted features, 14x28 ... 64x16 windows, <cascade> nodes
    without a type_id) and agree with the independent parser number for number."""
    from oracle.cascade_xml import load_cascade
    paths = all_cv2_haar_cascades()
    assert len(paths) >= 17
    ntilted = 0
    for path in paths:
        c = ug.Cascade(path)
        d = c.dump()
        f = load_cascade(path)
        for k in FIELDS:
            assert np.array_equal(d[k], getattr(f, k)), (path, k)
        assert c.info.has_tilted == int(f.tilted) and (c.info.win_w, c.info.win_h) == (f.win_w, f.win_h)
        ntilted += c.info.has_tilted
    assert ntilted >= 9


def test_old_format_loader_roundtrip(ug, tmp_path):
    """Old-format reader exercised without the reference tree: new-format cascade -> old-format text ->
    product loader; node data must come back bit for bit (stage thresholds lose 1e-5f once more)."""
    import cv2
    from oracle.cascade_xml import load_cascade
    f = load_cascade(cv2.data.haarcascades + "haarcascade_frontalface_alt2.xml")
    p = tmp_path / "old.xml"
    p.write_text(old_format_xml(f))
    c = ug.Cascade(str(p))
    d = c.dump()
    assert c.info.old_format == 1 and c.info.max_tree_nodes == 2
    for k in FIELDS:
        if k == "stage_thr":
            assert np.array_equal(d[k], (f.stage_thr - np.float32(1e-5)).astype(np.float32))
        else:
            assert np.array_equal(d[k], getattr(f, k)), k


def test_loader_errors(ug, tmp_path):
    with pytest.raises(ug.UgdError) as e:
        ug.Cascade(str(tmp_path / "missing.xml"))
    assert e.value.code == ug.ERR_IO
    bad = tmp_path / "bad.xml"
    bad.write_text("<opencv_storage><x type_id=\"opencv-haar-classifier\"><size>20 20</size></x>")
    with pytest.raises(ug.UgdError) as e:
        ug.Cascade(str(bad))
    assert e.value.code == ug.ERR_PARSE
    # a tilted rect that leaves the window (x - h < 0) is malformed; a fractional rect weight is unsupported
    for tag, rects, code in (("tilted", "<_>1 1 4 4 -1.</_><_>2 2 2 2 2.</_>", ug.ERR_PARSE),
                             ("weight", "<_>5 1 4 4 -1.5</_><_>6 2 2 2 2.</_>", ug.ERR_UNSUPPORTED)):
        f = tmp_path / f"{tag}.xml"
        f.write_text('<opencv_storage><c type_id="opencv-haar-classifier"><size>20 20</size><stages><_><trees><_><_>'
                     f"<feature><rects>{rects}</rects><tilted>1</tilted></feature>"
                     "<threshold>0.1</threshold><left_val>0.1</left_val><right_val>0.2</right_val></_></_></trees>"
                     "<stage_threshold>0.1</stage_threshold><parent>-1</parent><next>-1</next></_></stages></c></opencv_storage>")
        with pytest.raises(ug.UgdError) as e:
            ug.Cascade(str(f))
        assert e.value.code == code, tag


def test_config_reader_matches_reference_keys(ug, tmp_path):
    cfg = tmp_path / "conf" / "a.cfg"
    cfg.parent.mkdir()
    cfg.write_text(textwrap.dedent("""\
        # comment
        [scan]
        frame_size=640x360
        target_fps=10.0
        [detection]
        skin_lower=0,133,77
        skin_upper=255,173,127
        cascade_classifier_filepath=../data/haarcascade_frontalface_alt2.xml
        [association]
        method=optflow
        """))
    p, path, scan = ug.load_config(str(cfg))
    assert p.scale_factor == 1.1 and p.min_neighbors == 2 and (p.min_w, p.min_h) == (30, 30)
    assert list(p.skin_lower) == [0, 133, 77] and list(p.skin_upper) == [255, 173, 127]
    assert path == str(cfg.parent) + "/../data/haarcascade_frontalface_alt2.xml"     # relative to the cfg dir
    assert scan == (640, 360)
    cfg.write_text("[detection]\nscale=1.05\nskin_lower=1,2,3\ncascade_classifier_filepath=/abs/c.xml\n")
    p, path, scan = ug.load_config(str(cfg))
    assert p.scale_factor == 1.05 and list(p.skin_lower) == [1, 2, 3] and path == "/abs/c.xml" and scan == (-1, -1)
    cfg.write_text("[detection]\nskin_lower=0,133,77\n")                              # required key missing
    with pytest.raises(ug.UgdError):
        ug.load_config(str(cfg))
    if os.path.exists("/root/reference/conf/apink.mrchu.720p.cfg"):
        p, path, scan = ug.load_config("/root/reference/conf/apink.mrchu.720p.cfg")
        assert path.endswith("data/haarcascade_frontalface_alt2.xml") and scan == (640, 360)


@pytest.mark.parametrize("w,h,sf,ms", [(1280, 720, 1.1, (30, 30)), (1920, 1080, 1.1, (30, 30)), (640, 360, 1.1, (30, 30)),
                                       (3840, 2160, 1.05, (24, 24)), (333, 217, 1.25, (40, 40)), (25, 25, 1.1, (30, 30))])
def test_plan_matches_oracle_geometry(ug, orc, cv2ref, w, h, sf, ms):
    c = ug.Cascade(cv2ref.cascade_path("alt"))
    q = ug.plan_query(c, ug.default_params(scale_factor=sf, min_size=ms), w, h)
    want = orc.scale_list(w, h, 20, 20, sf, ms)
    assert np.array_equal(q["scales"], want)
    for i, s in enumerate(want):
        assert (q["lw"][i], q["lh"][i]) == orc.level_size(w, h, float(s))
        assert q["step"][i] == (1 if s >= 2 else 2)
    if (w, h, sf) == (1280, 720, 1.1):
        assert len(want) == 33 and q["level_pixels"] == 2042644          # 33 levels, 2.04 M level px


def test_c_abi_exports_every_declared_symbol(ug):
    hdr = open(os.path.join(ROOT, "include", "ugdet.h")).read()
    names = sorted(set(re.findall(r"\b(ugd_[a-z_0-9]+)\s*\(", hdr)))
    assert len(names) >= 25
    lib = ctypes.CDLL(ug.lib_path())
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ugdet.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", ug.lib_path()], capture_output=True, text=True).stdout
    for n in names:
        assert re.search(rf"\bT {n}\b", out), n
    assert ug.lib().ugd_version().decode().startswith("ugdet-b200")


def test_no_cpu_fallback_without_gpu(ug, cv2ref):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(ug.UgdError) as e:
        ug.Detector(ug.Cascade(cv2ref.cascade_path("alt")))
    assert e.value.code == ug.ERR_CUDA


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under the package may import, link or call it."""
    pkg = os.path.join(ROOT, "ugproj-detector_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cc", ".h", ".cuh", ".hpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "haar_oracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f


def test_sharding_helpers(ug):
    import importlib
    sh = importlib.import_module("ugproj_detector_b200.sharding")
    for F in (0, 1, 7, 256, 1001):
        for N in (1, 2, 4, 8):
            rs = [sh.shard_range(F, r, N) for r in range(N)]
            assert rs[0][0] == 0 and rs[-1][1] == F
            assert all(rs[i][1] == rs[i + 1][0] for i in range(N - 1))
            assert max(b - a for a, b in rs) - min(b - a for a, b in rs) <= 1
    rects = [np.zeros((0, 4), np.int32), np.arange(8, dtype=np.int32).reshape(2, 4), np.ones((70, 4), np.int32)]
    back = sh.unpack_slots(sh.pack_slots(rects, 64))
    assert len(back[0]) == 0 and np.array_equal(back[1], rects[1]) and len(back[2]) == 64


GLOO_WORKER = r'''
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.environ["UGD_ROOT"])
import __graft_entry__ as ge
ug = ge.load_package()
import importlib
sh = importlib.import_module("ugproj_detector_b200.sharding")
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
F, cap = 9, 8
def fake_detect(i):                      # deterministic stand-in for a frame's rect list
    return np.array([[i, 2 * i, 30 + k, 30 + k] for k in range(i % 4)], np.int32).reshape(-1, 4)
lo, hi = sh.shard_range(F, rank, world)
local = sh.pack_slots([fake_detect(i) for i in range(lo, hi)], cap)
maxf = max(b - a for a, b in (sh.shard_range(F, r, world) for r in range(world)))
assert (lo, hi) == ug.shard_range(F, rank, world)                        # the C ABI's ugd_shard_range: same rule
# the plumbing of sharding.init_comm: rank 0 draws the NCCL id through the library, every rank ends up with the same 128 bytes
box = [ug.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(box, src=0)
ids = [None] * world
dist.all_gather_object(ids, box[0])
assert len(box[0]) == 128 and all(i == ids[0] for i in ids)
bufs = sh.gather_to_rank0(torch.from_numpy(local), world, rank, maxf)
if rank == 0:
    got = []
    for r in range(world):
        a, b = sh.shard_range(F, r, world)
        got += sh.unpack_slots(bufs[r].numpy()[: b - a])
    assert len(got) == F
    for i in range(F):
        assert np.array_equal(got[i], fake_detect(i)), i
    print("GATHER_OK")
dist.destroy_process_group()
'''


def test_gather_to_rank0_gloo_world2(tmp_path):
    """N>1 path on CPU: two gloo ranks shard 9 frames, lists gathered to rank 0 equal the 1-rank result."""
    w = tmp_path / "worker.py"
    w.write_text(GLOO_WORKER)
    env = dict(os.environ, UGD_ROOT=ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29617", str(w)], env=env,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "GATHER_OK" in r.stdout


@pytest.mark.parametrize("name,w,h,sf,ms", [("alt", 1280, 720, 1.1, 30), ("default", 1920, 1080, 1.1, 30), ("alt", 3840, 2160, 1.05, 24),
                                            ("alt2", 640, 360, 1.1, 30)])
def test_bench_geometry_matches_the_plan(ug, cv2ref, name, w, h, sf, ms):
    """bench.py prints `config` in both arms from a plain-Python restatement of the scale list (the reference arm must
    not load the CUDA library): it has to agree with the engine's plan."""
    import bench
    c = ug.Cascade(cv2ref.cascade_path(name))
    q = ug.plan_query(c, ug.default_params(scale_factor=sf, min_size=(ms, ms)), w, h)
    levels, level_px, int_entries = bench.geometry(w, h, c.info.win_w, sf, ms)
    assert levels == len(q["scales"]) and level_px == q["level_pixels"]
    assert int_entries == int(sum((int(a) + 1) * (int(b) + 1) for a, b in zip(q["lw"], q["lh"])))


def test_bench_reference_arm_prints_the_same_config(ug, cv2ref):
    """`bench.py --impl reference` runs cv2 alone (no CUDA library in the process) on rank 0's frames and carries the
    same `config` object as the repo arm, so the driver's ratio compares like with like."""
    import bench
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "frames/s" and line["value"] > 0
    want = bench.workload_config("haarcascade_frontalface_alt.xml", 1280, 720, 1.1, 30, bench.FRAMES_PER_STEP, bench.GROUP, 1, 20)
    assert line["config"] == want
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["cpu_baseline"]["kind"] == "reference"
