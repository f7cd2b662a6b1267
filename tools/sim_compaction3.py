"""Analysis aid: tile shape study.  Costs (in k warp-iterations x weak classifiers, stages 1..D-1):
ideal n/32, A = dense list (iters, LDS wavefront factor), B = per-bank queues (rounds), slots = barrier
slots (ceil(chunks/8)) per stage.  python tools/sim_compaction3.py cascade ROWS COLS"""
import sys, os, importlib
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
from oracle import oracle as orc, cv2_ref
name = sys.argv[1]; ROWS = int(sys.argv[2]); COLS = int(sys.argv[3]); D = int(os.environ.get("DENSE", 6))
W, H = 1280, 720
ge.load_package(); synth = importlib.import_module(ge.PKG_NAME + ".synth")
oc = orc.OracleCascade.load(cv2_ref.cascade_path(name))
g = synth.gray_frame(W, H, 8, 1)
raw, sums, ctr, smap = orc.detect_raw(oc, g, want_map=True)
ow = oh = 24 if name == "default" else 20
ntrees = [int(x) for x in oc.flat.stage_ntrees]
rr, xx = np.meshgrid(np.arange(ROWS), np.arange(COLS), indexing="ij")
off = 0
ideal = itA = wvA = rB = slotsA = slotsB = unc = 0
SH = int(os.environ.get("SHIFT", 20))
for s in orc.scale_list(W, H, ow, oh):
    sw, sh = orc.level_size(W, H, float(s)); step = 1 if s >= 2.0 else 2
    nxg = (sw + 1 - ow + step - 1) // step; nyg = (sh + 1 - oh + step - 1) // step
    m = smap[off:off + nxg * nyg].reshape(nyg, nxg).astype(np.int32); off += nxg * nyg
    bank = (rr * SH + xx) % 32
    for r0 in range(0, nyg, ROWS):
        for x0 in range(0, nxg, COLS):
            t = np.full((ROWS, COLS), -5, np.int32); tt = m[r0:r0 + ROWS, x0:x0 + COLS]; t[:tt.shape[0], :tt.shape[1]] = tt
            for st in range(1, D):
                a = t >= st; n = int(a.sum())
                if n == 0: break
                nw = ntrees[st]; b = bank[a]
                ideal += n / 32 * nw; k = -(-n // 32); itA += k * nw; slotsA += -(-k // 8) * nw
                wvA += sum(np.bincount(b[c:c + 32], minlength=32).max() for c in range(0, n, 32)) * nw
                q = np.bincount(b, minlength=32).max(); rB += q * nw; slotsB += -(-q // 8) * nw
print(f"{name} tile {ROWS}x{COLS}: ideal {ideal/1e3:.0f}k  A iters {itA/1e3:.0f}k waves {wvA/1e3:.0f}k slots*8 {slotsA*8/1e3:.0f}k | B rounds {rB/1e3:.0f}k slots*8 {slotsB*8/1e3:.0f}k")
