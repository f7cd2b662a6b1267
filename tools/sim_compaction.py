"""Analysis aid (not product, not test): replays the oracle's per-window stage-of-rejection map through
the k_scan tile geometry and costs different survivor-compaction schemes (warp iterations and
shared-memory wavefronts per weak classifier).  Usage: python tools/sim_compaction.py [cascade] [WxH]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
from oracle import oracle as orc, cv2_ref

name = sys.argv[1] if len(sys.argv) > 1 else "alt"
W, H = map(int, (sys.argv[2] if len(sys.argv) > 2 else "1280x720").split("x"))
ge.load_package()
import importlib
synth = importlib.import_module(ge.PKG_NAME + ".synth")
oc = orc.OracleCascade.load(cv2_ref.cascade_path(name))
import xml.etree.ElementTree as ET
from oracle import cascade_xml
flat = cascade_xml.load(cv2_ref.cascade_path(name)) if hasattr(cascade_xml, "load") else None
g = synth.gray_frame(W, H, synth.default_nfaces(W, H), 1)
raw, sums, ctr, smap = orc.detect_raw(oc, g, want_map=True)
ow, oh = (24, 24) if name == "default" else (20, 20)
scales = orc.scale_list(W, H, ow, oh)
ntrees = [int(x) for x in oc.flat.stage_ntrees] if hasattr(oc, "flat") else None
print("stages", ntrees)
nst = len(ntrees)
ROWS = int(os.environ.get("ROWS", 8))
D = int(os.environ.get("DENSE", 6))
off = 0
tot = {k: np.zeros(nst) for k in ("ideal", "iterA", "waveA", "roundB", "alive", "waveA0")}
for s in scales:
    sw, sh = orc.level_size(W, H, float(s))
    step = 1 if s >= 2.0 else 2
    nxg = (sw + 1 - ow + step - 1) // step; nyg = (sh + 1 - oh + step - 1) // step
    m = smap[off:off + nxg * nyg].reshape(nyg, nxg).astype(np.int32); off += nxg * nyg
    TP = (31 + ow + 1) if step == 1 else 2 * ((31 * 2 + ow + 1 + 1) // 2)
    rowshift = (TP * step) % 32
    for r0 in range(0, nyg, ROWS):
        for x0 in range(0, nxg, 32):
            t = m[r0:r0 + ROWS, x0:x0 + 32]
            rr, xx = np.meshgrid(np.arange(t.shape[0]), np.arange(t.shape[1]), indexing="ij")
            bank = (rr * rowshift + xx) % 32
            for st in range(1, nst):
                a = t >= st
                n = int(a.sum())
                if n == 0: break
                nw = ntrees[st]
                b = bank[a]                       # row-major order == tid order
                tot["alive"][st] += n * nw
                tot["ideal"][st] += n / 32 * nw
                tot["iterA"][st] += -(-n // 32) * nw
                wv = 0
                for c in range(0, n, 32):
                    wv += np.bincount(b[c:c + 32], minlength=32).max()
                tot["waveA"][st] += wv * nw
                tot["roundB"][st] += np.bincount(b, minlength=32).max() * nw
assert off == len(smap), (off, len(smap))
np.set_printoptions(linewidth=200, suppress=True)
for k, v in tot.items():
    print(f"{k:8s} in-tile(1..{D-1}) {v[1:D].sum()/1e3:10.0f}k   rest {v[D:].sum()/1e3:10.0f}k   per-stage(1..7)", (v[1:8] / 1e3).round(0))
