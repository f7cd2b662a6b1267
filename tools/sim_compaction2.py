"""Analysis aid: conflict factor of the dense survivor lists for different tile pitches (row bank shift)
and list orders.  python tools/sim_compaction2.py [cascade]"""
import sys, os, importlib
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
from oracle import oracle as orc, cv2_ref
name = sys.argv[1] if len(sys.argv) > 1 else "alt"
W, H = 1280, 720
ge.load_package(); synth = importlib.import_module(ge.PKG_NAME + ".synth")
oc = orc.OracleCascade.load(cv2_ref.cascade_path(name))
g = synth.gray_frame(W, H, 8, 1)
raw, sums, ctr, smap = orc.detect_raw(oc, g, want_map=True)
ow = oh = 24 if name == "default" else 20
ntrees = [int(x) for x in oc.flat.stage_ntrees]
ROWS = int(os.environ.get("ROWS", 8)); D = int(os.environ.get("DENSE", 6))
tiles = []      # (step, list of (alive mask per stage))
off = 0
for s in orc.scale_list(W, H, ow, oh):
    sw, sh = orc.level_size(W, H, float(s)); step = 1 if s >= 2.0 else 2
    nxg = (sw + 1 - ow + step - 1) // step; nyg = (sh + 1 - oh + step - 1) // step
    m = smap[off:off + nxg * nyg].reshape(nyg, nxg).astype(np.int32); off += nxg * nyg
    for r0 in range(0, nyg, ROWS):
        for x0 in range(0, nxg, 32):
            t = np.full((ROWS, 32), -5, np.int32); tt = m[r0:r0 + ROWS, x0:x0 + 32]; t[:tt.shape[0], :tt.shape[1]] = tt
            if (t >= 1).any(): tiles.append((step, t))
rr, xx = np.meshgrid(np.arange(ROWS), np.arange(32), indexing="ij")
def cost(order, shift1, shift2):
    it = wv = 0
    for step, t in tiles:
        bank = (rr * (shift1 if step == 1 else shift2) + xx) % 32
        for st in range(1, D):
            a = t >= st
            n = int(a.sum())
            if n == 0: break
            b = bank[a] if order == "row" else bank.T[a.T]
            nw = ntrees[st]
            it += -(-n // 32) * nw
            for c in range(0, n, 32):
                wv += np.bincount(b[c:c + 32], minlength=32).max() * nw
    return it, wv
for order in ("row", "col"):
    for sh in (0, 1, 2, 3, 4, 5, 8, 12, 16, 20):
        it, wv = cost(order, sh, sh)
        print(order, "shift", sh, "iters %.0fk waves %.0fk factor %.2f" % (it / 1e3, wv / 1e3, wv / it), flush=True)
print("--- scheme B: per-bank queues, rounds = max queue length")
def costB(shift1, shift2, Dn=D):
    rounds = 0
    for step, t in tiles:
        bank = (rr * (shift1 if step == 1 else shift2) + xx) % 32
        for st in range(1, Dn):
            a = t >= st
            if not a.any(): break
            rounds += np.bincount(bank[a], minlength=32).max() * ntrees[st]
    return rounds
for sh in (0, 1, 2, 3, 4, 5, 6, 7, 8, 12, 16, 20):
    print("B shift", sh, "rounds %.0fk" % (costB(sh, sh) / 1e3), flush=True)
