"""TEST INFRASTRUCTURE ONLY: one rank of a world_size-N gloo run of adsbcu_1090_replay_sharded on the SIMT emulator build
(tests/test_replay_sharded_gloo.py starts N of these).  The transport under test is torch.distributed all_gather, the same
callback bench.py uses on the GPU box with the real library."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import emu_load  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{os.environ['MASTER_PORT']}", rank=rank, world_size=world)
    adsbcu = emu_load.load_emulated()
    with np.load(os.path.join(ROOT, "tests", "golden", "c3_dense.npz")) as z:
        iq, ref = z["iq"], z["ref_fix2_df1"]
    n = iq.size // 2
    cut = lambda k: n if k >= world + 1 else (n * k // (world + 1)) // 8 * 8      # noqa: E731
    lo, hi = (0, cut(2)) if rank == 0 else (cut(rank + 1), cut(rank + 2))

    def allgather(words):
        buf = torch.zeros(adsbcu.SHARD_WORDS + 1, dtype=torch.int32)
        buf[0] = int(words.size)
        buf[1:1 + words.size] = torch.from_numpy(words.view(np.int32))
        parts = [torch.empty_like(buf) for _ in range(world)]
        dist.all_gather(parts, buf)
        return [p[1:1 + int(p[0])].numpy().view(np.uint32) for p in parts]

    d = adsbcu.Demod1090(2, 1, max_feed_samples=1 << 20)
    out = np.zeros(1 << 16, adsbcu.MSG_DTYPE)
    msgs, rep = d.replay_sharded(rank, world, [(iq.ctypes.data + 2 * lo, hi - lo)], lo, iq[2 * max(0, lo - 326):2 * lo], allgather, out)
    gathered = [None] * world
    dist.all_gather_object(gathered, (msgs.tobytes(), rep.passes, rep.rounds))
    if rank == 0:
        from oracle import oracle1090 as O
        got = np.concatenate([np.frombuffer(b, adsbcu.MSG_DTYPE) for b, _, _ in gathered])
        diff = O.same_messages(ref, got)
        assert diff == [], diff[:5]
        assert gathered[0][1] == 1 and all(g[2] == gathered[0][2] for g in gathered)
        print("SHARDED-OK", len(got), [g[1] for g in gathered], gathered[0][2], flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
