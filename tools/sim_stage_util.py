"""Analysis aid: per-stage lane utilisation of the per-bank-class rounds (tile 16x64) on an oracle stage map."""
import sys, os, importlib
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
from oracle import oracle as orc, cv2_ref
name = sys.argv[1] if len(sys.argv) > 1 else "alt"
ROWS, COLS, W, H, SH, D = 16, 64, 1280, 720, 20, 8
ge.load_package(); synth = importlib.import_module(ge.PKG_NAME + ".synth")
oc = orc.OracleCascade.load(cv2_ref.cascade_path(name))
g = synth.gray_frame(W, H, 8, 1)
raw, sums, ctr, smap = orc.detect_raw(oc, g, want_map=True)
ow = oh = 24 if name == "default" else 20
ntrees = [int(x) for x in oc.flat.stage_ntrees]
rr, xx = np.meshgrid(np.arange(ROWS), np.arange(COLS), indexing="ij")
bank = (rr * SH + xx) % 32
alive = np.zeros(D); rounds = np.zeros(D); warpslots = np.zeros(D)
off = 0
for s in orc.scale_list(W, H, ow, oh):
    sw, sh = orc.level_size(W, H, float(s)); step = 1 if s >= 2.0 else 2
    nxg = (sw + 1 - ow + step - 1) // step; nyg = (sh + 1 - oh + step - 1) // step
    m = smap[off:off + nxg * nyg].reshape(nyg, nxg).astype(np.int32); off += nxg * nyg
    for r0 in range(0, nyg, ROWS):
        for x0 in range(0, nxg, COLS):
            t = np.full((ROWS, COLS), -5, np.int32); tt = m[r0:r0 + ROWS, x0:x0 + COLS]; t[:tt.shape[0], :tt.shape[1]] = tt
            for st in range(1, D):
                a = t >= st; n = int(a.sum())
                if n == 0 or (n < 32 and st >= 2): break
                r = np.bincount(bank[a], minlength=32).max()
                alive[st] += n; rounds[st] += r
                warpslots[st] += -(-r // 8) * 8          # 8 warps, rounds dealt evenly: critical path = ceil(r/8) rounds
for st in range(1, D):
    print(f"stage {st}: nodes {ntrees[st]:3d}  alive {alive[st]/1e3:7.1f}k  rounds {rounds[st]/1e3:6.1f}k  lane util {alive[st]/(32*rounds[st]):.2f}  "
          f"warp-slot util {rounds[st]/warpslots[st]:.2f}  share of node-rounds {ntrees[st]*rounds[st]/sum(ntrees[i]*rounds[i] for i in range(1,D)):.2f}")
