"""Worker of tests/test_gpu_multi.py, launched with torchrun (one process per GPU).

Every rank runs a real detector on its ugd_shard_range of the SAME seeded frames and the lists are gathered to rank 0
INSIDE libugdet.so (ugd_detect_batch_sharded: NCCL send/recv from the device-side result slots).  Rank 0 then runs all
frames on its own GPU with a second, single-GPU detector and asserts that the gathered lists, put back in global frame
order, are identical -- SURVEY 8(e) "Verify", and against live cv2 for a few frames."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")                     # plumbing only: carries the 128-byte NCCL id
    ug = ge.load_package()
    synth = importlib.import_module(ge.PKG_NAME + ".synth")
    sharding = importlib.import_module(ge.PKG_NAME + ".sharding")
    from oracle import cv2_ref

    F, w, h, cap = 11, 640, 360, 64                     # 11 frames over 2 ranks: 5 + 6, uneven on purpose
    failures = []
    for name, bgr in (("alt", False), ("alt2", True)):
        casc = ug.Cascade(cv2_ref.cascade_path(name))
        if bgr:
            frames = np.stack([synth.bgr_frame(w, h, 5, 4000 + i) for i in range(F)])
        else:
            frames = synth.gray_batch(w, h, F, seed0=3000, nfaces=5)
        lo, hi = ug.shard_range(F, rank, world)
        assert (lo, hi) == sharding.shard_range(F, rank, world)
        slot = max(ug.shard_range(F, r, world)[1] - ug.shard_range(F, r, world)[0] for r in range(world))
        det = ug.Detector(casc, ug.default_params(), device=local, max_size=(w, h), max_batch=2)   # several groups per rank
        sharding.init_comm(det, rank, world)
        assert det.comm_info()[:2] == (rank, world)
        mine = np.ascontiguousarray(frames[lo:hi])
        dev = torch.from_numpy(mine).cuda()
        runs = {"host": det.detect_batch_sharded(mine.ctypes.data, False, bgr, hi - lo, w, h, slot, 0, cap),
                "device": det.detect_batch_sharded(dev.data_ptr(), True, bgr, hi - lo, w, h, slot, 0, cap)}
        if rank == 0:
            single = ug.Detector(casc, ug.default_params(), device=local, max_size=(w, h), max_batch=4)
            want = single.detect_batch(frames, cap=cap)
            cc = cv2_ref.load(cv2_ref.cascade_path(name))
            for f in (0, F - 1):
                ref = cv2_ref.detect_faces_cv2(cc, frames[f]) if bgr else cv2_ref.detect_multiscale_cv2(cc, frames[f])
                if sorted(map(tuple, want[f].tolist())) != sorted(map(tuple, np.asarray(ref).reshape(-1, 4).tolist())):
                    failures.append(f"{name}: single-GPU frame {f} differs from cv2")
            for kind, got in runs.items():
                if len(got) != world:
                    failures.append(f"{name}/{kind}: {len(got)} rank lists")
                    continue
                flat = [r for per_rank in got for r in per_rank]       # contiguous ranges: rank order == frame order
                if len(flat) != F:
                    failures.append(f"{name}/{kind}: {len(flat)} frames gathered, want {F}")
                    continue
                for f in range(F):
                    if not np.array_equal(flat[f], want[f]):
                        failures.append(f"{name}/{kind}: frame {f} gathered {flat[f].tolist()} != single-GPU {want[f].tolist()}")
            if sum(len(x) for x in want) == 0:
                failures.append(f"{name}: no detections at all, the comparison is vacuous")
            single.close()
        else:
            for kind, got in runs.items():                            # other ranks get their own lists back
                if len(got) != hi - lo:
                    failures.append(f"rank {rank} {name}/{kind}: {len(got)} local lists")
        det.close()
    dist.barrier()
    if failures:
        print("\n".join(failures), file=sys.stderr)
        sys.exit(1)
    if rank == 0:
        print(f"mgpu ok: world {world}, NCCL {det.comm_info()[2] if False else ''}".strip())
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
