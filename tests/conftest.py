import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as ge  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def ug():
    return ge.load_package()


@pytest.fixture(scope="session")
def synth(ug):
    import importlib
    return importlib.import_module(ge.PKG_NAME + ".synth")


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle
    oracle.lib()
    return oracle


@pytest.fixture(scope="session")
def cv2ref():
    from oracle import cv2_ref
    return cv2_ref


CASCADES = ["alt", "alt2", "default", "alt_tree"]


def rectset(a):
    return sorted(map(tuple, [[int(v) for v in r] for r in a]))


def all_cv2_haar_cascades():
    """Every Haar cascade file cv2 ships (17 in cv2 4.13; 9 of them use 45-degree tilted features)."""
    import glob
    import cv2
    return sorted(glob.glob(os.path.join(cv2.data.haarcascades, "haarcascade_*.xml")))


def truncate_cascade(path, nstages, out_path):
    """Writes a copy of a cascade file (either format) that keeps only its first `nstages` stages: many windows then
    survive to the end, which exercises every evaluator and the final-stage sums far more than the
    full cascade does on a handful of frames (the crafted-cascade idea of SURVEY A.6)."""
    import xml.etree.ElementTree as ET
    tree = ET.parse(path)
    casc = [e for e in tree.getroot() if e.find("stages") is not None][0]
    stages = casc.find("stages")
    for e in list(stages)[nstages:]:
        stages.remove(e)
    if casc.find("stageNum") is not None:        # new format only; old-format files just list their stages
        casc.find("stageNum").text = str(nstages)
    with open(out_path, "wb") as fh:
        fh.write(b'<?xml version="1.0"?>\n')
        tree.write(fh)
    return str(out_path)


def long_failure_runs_cascade(out_path, win=(20, 20), thr=0.002):
    """A two-stage stump cascade whose stage 0 looks at one 2x2 spot against its 4x4 surround: on textured frames it fails in
    long runs (and passes in others), which drives the x-skip rule -- evaluate, skip, evaluate ... along a run of failures --
    across many tile boundaries.  Stage 1 accepts what is left."""
    with open(out_path, "w") as fh:
        fh.write('<?xml version="1.0"?>\n<opencv_storage><c type_id="opencv-haar-classifier">'
                 f"<size>{win[0]} {win[1]}</size><stages>"
                 "<_><trees><_><_><feature><rects><_>8 8 4 4 -1.</_><_>9 9 2 2 4.</_></rects><tilted>0</tilted></feature>"
                 f"<threshold>{thr}</threshold><left_val>0.</left_val><right_val>1.</right_val></_></_></trees>"
                 "<stage_threshold>0.5</stage_threshold><parent>-1</parent><next>-1</next></_>"
                 "<_><trees><_><_><feature><rects><_>1 1 4 4 -1.</_><_>2 2 2 2 2.</_></rects><tilted>0</tilted></feature>"
                 "<threshold>0.1</threshold><left_val>1.</left_val><right_val>1.</right_val></_></_></trees>"
                 "<stage_threshold>0.5</stage_threshold><parent>0</parent><next>-1</next></_>"
                 "</stages></c></opencv_storage>")
    return str(out_path)


def pass_everything_cascade(out_path, win=(20, 20)):
    """a one-stump cascade whose two leaves both clear the stage threshold, so it accepts every
    window that is evaluated and passes the variance test -- it enumerates the oracle's window grid."""
    with open(out_path, "w") as fh:
        fh.write('<?xml version="1.0"?>\n<opencv_storage><c type_id="opencv-haar-classifier">'
                 f"<size>{win[0]} {win[1]}</size><stages><_><trees><_><_>"
                 "<feature><rects><_>1 1 4 4 -1.</_><_>2 2 2 2 2.</_></rects><tilted>0</tilted></feature>"
                 "<threshold>0.1</threshold><left_val>1.</left_val><right_val>1.</right_val></_></_></trees>"
                 "<stage_threshold>0.5</stage_threshold><parent>-1</parent><next>-1</next></_></stages></c></opencv_storage>")
    return str(out_path)


def old_format_xml(f, extra_tree=None):
    """Writes a FlatCascade (oracle.cascade_xml.load_cascade: either XML format) back out as an old-format
    (opencv-haar-classifier) file.  extra_tree(stage, k) -> XML of an extra tree to insert before tree k of the stage
    (k == ntrees: at the end), or None."""
    import numpy as np
    out = ['<?xml version="1.0"?>', "<opencv_storage>", '<cascade type_id="opencv-haar-classifier">',
           f"<size>{f.win_w} {f.win_h}</size>", "<stages>"]
    for s in range(f.nstages):
        out.append("<_><!-- stage --><trees>")
        nt = int(f.stage_ntrees[s])
        for k, t in enumerate(range(f.stage_first_tree[s], f.stage_first_tree[s] + nt)):
            if extra_tree and extra_tree(s, k):
                out.append(extra_tree(s, k))
            out.append("<_>")
            for j in range(f.tree_nnodes[t]):
                n = f.tree_first_node[t] + j
                out.append("<_><feature><rects>")
                for r in range(3):
                    if f.node_weights[n, r] != 0:
                        x, y, w, h = f.node_rects[n, r]
                        out.append(f"<_>{x} {y} {w} {h} {float(f.node_weights[n, r])!r}</_>")
                out.append(f"</rects><tilted>{int(f.node_tilted[n])}</tilted></feature><threshold>{float(f.node_thr[n])!r}</threshold>")
                out.append(f"<left_node>{f.node_left[n]}</left_node>" if f.node_left[n] else f"<left_val>{float(f.node_left_val[n])!r}</left_val>")
                out.append(f"<right_node>{f.node_right[n]}</right_node>" if f.node_right[n] else f"<right_val>{float(f.node_right_val[n])!r}</right_val>")
                out.append("</_>")
            out.append("</_>")
        if extra_tree and extra_tree(s, nt):
            out.append(extra_tree(s, nt))
        thr = float(np.float32(f.stage_thr[s]))          # already f32(xml) - 1e-5f
        out.append(f"</trees><stage_threshold>{thr!r}</stage_threshold><parent>{s - 1}</parent><next>-1</next></_>")
    out += ["</stages>", "</cascade>", "</opencv_storage>"]
    return "\n".join(out)


def inexact_cascade(path, out_path, every=4):
    """A copy of a stump cascade (either XML format in, old format out) with extra stumps whose leaves are tiny (~1e-9,
    full 24-bit mantissas) at the head of every stage and after every `every`-th tree: decisions barely move, but the
    exponent spread of the leaves now exceeds what an fp64 sum holds exactly, so ugd_cascade_info.sum_exact is 0 and
    the low bits of a stage sum depend on the order of addition -- the engine must add the leaves in file order, like
    the oracle and cv2."""
    from oracle.cascade_xml import load_cascade
    f = load_cascade(path)

    def tiny(s, k):
        if k % every:
            return None
        i = 131 * s + k
        a, b = 1.2345678e-9 * (1.0 + 0.37 * (i % 11)), 2.7182818e-9 * (1.0 + 0.21 * (i % 7))
        return ("<_><_><feature><rects><_>%d %d 6 4 -1.</_><_>%d %d 3 4 2.</_></rects><tilted>0</tilted></feature>"
                "<threshold>%.6f</threshold><left_val>%.9e</left_val><right_val>%.9e</right_val></_></_>"
                % (i % 9, i % 7, i % 9, i % 7, 0.001 * (i % 5 - 2), a, b))
    with open(out_path, "w") as fh:
        fh.write(old_format_xml(f, tiny))
    return str(out_path)


def stage_tree_cascade(out_path, win=(20, 20)):
    """An old-format cascade whose stages form a TREE (like haarcascade_frontalface_alt_tree.xml:1303-1304): stage 0 is the
    root and passes every window; stage 1 (parent 0, next 2) rejects every window; stage 2 (parent 0) passes every
    window.  Walked as a tree (OpenCV 2.4) every window is accepted through the `next` branch; evaluated in file order
    (OpenCV 4.x) every window dies at stage 1."""
    def stage(thr, parent, nxt):
        return ("<_><trees><_><_><feature><rects><_>1 1 4 4 -1.</_><_>2 2 2 2 2.</_></rects><tilted>0</tilted></feature>"
                "<threshold>0.1</threshold><left_val>1.</left_val><right_val>1.</right_val></_></_></trees>"
                f"<stage_threshold>{thr}</stage_threshold><parent>{parent}</parent><next>{nxt}</next></_>")
    with open(out_path, "w") as fh:
        fh.write('<?xml version="1.0"?>\n<opencv_storage><c type_id="opencv-haar-classifier">'
                 f"<size>{win[0]} {win[1]}</size><stages>" + stage(0.5, -1, -1) + stage(5.0, 0, 2) + stage(0.5, 0, -1) +
                 "</stages></c></opencv_storage>")
    return str(out_path)
