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


def inexact_cascade(path, out_path, every=4):
    """A copy of an old-format stump cascade with extra stumps whose leaves are tiny (~1e-9, full 24-bit mantissas)
    at the head of every stage and after every `every`-th tree: decisions barely move, but the exponent spread of the
    leaves now exceeds what an fp64 sum holds exactly, so ugd_cascade_info.sum_exact is 0 and the low bits of a stage
    sum depend on the order of addition -- the engine must add the leaves in file order, like the oracle and cv2."""
    import re
    text = open(path).read()
    k = [0]

    def tiny():
        k[0] += 1
        a, b = 1.2345678e-9 * (1.0 + 0.37 * (k[0] % 11)), 2.7182818e-9 * (1.0 + 0.21 * (k[0] % 7))
        return ("<_><_><feature><rects><_>%d %d 6 4 -1.</_><_>%d %d 3 4 2.</_></rects><tilted>0</tilted></feature>"
                "<threshold>%.6f</threshold><left_val>%.9e</left_val><right_val>%.9e</right_val></_></_>"
                % (k[0] % 9, k[0] % 7, k[0] % 9, k[0] % 7, 0.001 * (k[0] % 5 - 2), a, b))
    out, pos, ntree = [], 0, 0
    for m in re.finditer(r"<trees>|</_>\s*</_>(?=\s*(?:<!--[^>]*-->\s*)?(?:<_>|</trees>))", text):
        out.append(text[pos:m.end()])
        pos = m.end()
        if m.group(0) == "<trees>":
            ntree = 0
            out.append(tiny())
        else:
            ntree += 1
            if ntree % every == 0:
                out.append(tiny())
    out.append(text[pos:])
    with open(out_path, "w") as fh:
        fh.write("".join(out))
    return str(out_path)
