"""Analysis aid: rounds needed when g bank classes pool their survivors (g = 1: scheme B; g = 2, 4: a lane
may take a window of a partner class -> up to g-way bank conflict in that round).
python tools/sim_pool.py cascade [D]"""
import sys, os, importlib
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
from oracle import oracle as orc, cv2_ref
name = sys.argv[1]; D = int(sys.argv[2]) if len(sys.argv) > 2 else 8
ROWS, COLS, W, H, SH = 16, 64, 1280, 720, 20
ge.load_package(); synth = importlib.import_module(ge.PKG_NAME + ".synth")
oc = orc.OracleCascade.load(cv2_ref.cascade_path(name))
g = synth.gray_frame(W, H, 8, 1)
raw, sums, ctr, smap = orc.detect_raw(oc, g, want_map=True)
ow = oh = 24 if name == "default" else 20
ntrees = [int(x) for x in oc.flat.stage_ntrees]
rr, xx = np.meshgrid(np.arange(ROWS), np.arange(COLS), indexing="ij")
bank = (rr * SH + xx) % 32
off = 0
res = {1: 0, 2: 0, 4: 0, 8: 0, 32: 0}; grp4 = {1: 0, 2: 0, 4: 0, 8: 0, 32: 0}; ideal = 0
for s in orc.scale_list(W, H, ow, oh):
    sw, sh = orc.level_size(W, H, float(s)); step = 1 if s >= 2.0 else 2
    nxg = (sw + 1 - ow + step - 1) // step; nyg = (sh + 1 - oh + step - 1) // step
    m = smap[off:off + nxg * nyg].reshape(nyg, nxg).astype(np.int32); off += nxg * nyg
    for r0 in range(0, nyg, ROWS):
        for x0 in range(0, nxg, COLS):
            t = np.full((ROWS, COLS), -5, np.int32); tt = m[r0:r0 + ROWS, x0:x0 + COLS]; t[:tt.shape[0], :tt.shape[1]] = tt
            for st in range(1, D):
                a = t >= st; n = int(a.sum())
                if n == 0: break
                cnt = np.bincount(bank[a], minlength=32); nw = ntrees[st]
                ideal += n / 32 * nw
                for gsz in res:
                    pooled = cnt.reshape(gsz, 32 // gsz).sum(0) if gsz > 1 else cnt    # classes q, q+32/g, ... pooled
                    r = int(np.ceil(pooled.max() / gsz))
                    res[gsz] += r * nw; grp4[gsz] += -(-r // 4) * 4 * nw
print(name, "D", D, "ideal %.0fk" % (ideal / 1e3), {k: "%.0fk (x4 groups %.0fk)" % (v / 1e3, grp4[k] / 1e3) for k, v in res.items()})
