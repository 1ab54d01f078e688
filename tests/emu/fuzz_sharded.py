"""TEST INFRASTRUCTURE ONLY.  Random-case search on the CPU for the sharded replay (adsbcu_1090_replay_sharded[_tail]): per case
2-4 gloo ranks, each with the kernel sources on the SIMT emulator, cut one random capture at random (8-sample aligned) places -
some ranges shorter than a message, with and without the tail range on rank 0 - and the concatenated message set is compared
with the oracle's single-stream run.  usage: python tests/emu/fuzz_sharded.py [seconds_to_run] [first_seed]"""
import os
import subprocess
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)


def case_of(seed):
    rng = np.random.default_rng(seed)
    world = int(rng.integers(2, 5))
    tail = bool(rng.random() < 0.5)
    nfix, fdf = int(rng.integers(0, 3)), int(rng.integers(0, 2))
    seconds = float(rng.choice([0.03, 0.08]))
    rate = float(rng.choice([3000.0, 8000.0, 20000.0, 60000.0]))
    pool = int(rng.choice([5, 80, 1500]))
    nranges = world + (1 if tail else 0)
    return rng, dict(world=world, tail=tail, nfix=nfix, fdf=fdf, seconds=seconds, rate=rate, pool=pool, nranges=nranges)


def rank_main():
    import torch
    import torch.distributed as dist
    import emu_load
    from libadsb_b200 import synth
    rank, world, seed = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["FUZZ_SEED"])
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{os.environ['MASTER_PORT']}", rank=rank, world_size=world)
    A = emu_load.load_emulated()
    rng, c = case_of(seed)
    iq = synth.config3(seconds=c["seconds"], seed=seed, msgs_per_s=c["rate"], pool=c["pool"])[0]
    n = iq.size // 2
    inner = sorted(int(x) // 8 * 8 for x in rng.integers(8, n - 8, c["nranges"] - 1))
    if rng.random() < 0.3 and len(inner) > 1:                      # a very short range
        inner[1] = min(n - 8, inner[0] + 8 * int(rng.integers(1, 40)))
        inner = sorted(inner)
    cuts = [0] + inner + [n]
    assert all(cuts[i] < cuts[i + 1] for i in range(len(cuts) - 1)), cuts

    def allgather(words):
        buf = torch.zeros(A.SHARD_WORDS + 1, dtype=torch.int32)
        buf[0] = int(words.size)
        buf[1:1 + words.size] = torch.from_numpy(words.view(np.int32))
        parts = [torch.empty_like(buf) for _ in range(world)]
        dist.all_gather(parts, buf)
        return [p[1:1 + int(p[0])].numpy().view(np.uint32) for p in parts]

    def piece(k):
        lo, hi = cuts[k], cuts[k + 1]
        return [(iq.ctypes.data + 2 * lo, hi - lo)], lo, iq[2 * max(0, lo - 326):2 * lo]

    d = A.Demod1090(c["nfix"], c["fdf"], max_feed_samples=1 << 20)
    out = np.zeros(1 << 16, A.MSG_DTYPE)
    pieces, lo, halo = piece(rank)
    if c["tail"]:
        msgs, rep = d.replay_sharded(rank, world, pieces, lo, halo, allgather, out, has_tail=True, tail=piece(world) if rank == 0 else None)
        first = d.n_first if rank == 0 else len(msgs)
        mine, last = msgs[:first].tobytes(), msgs[first:].tobytes()
    else:
        msgs, rep = d.replay_sharded(rank, world, pieces, lo, halo, allgather, out)
        mine, last = msgs.tobytes(), b""
    gathered = [None] * world
    dist.all_gather_object(gathered, (mine, last))
    ok = True
    if rank == 0:
        from oracle import oracle1090 as O
        got = np.concatenate([np.frombuffer(m, A.MSG_DTYPE) for m, _ in gathered] + [np.frombuffer(gathered[0][1], A.MSG_DTYPE)])
        want = O.Oracle1090(c["nfix"], c["fdf"]).feed(iq, cap=1 << 16)
        diff = O.same_messages(want, got)
        ok = not diff
        print(("SHARDED-OK" if ok else "SHARDED-MISMATCH"), seed, len(want), c, cuts, diff[:3], flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    if "FUZZ_SEED" in os.environ:
        rank_main()
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    t0, ncase = time.time(), 0
    while time.time() - t0 < budget:
        world = case_of(seed)[1]["world"]
        port = str(29500 + seed % 2000)
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__)], env=dict(os.environ, RANK=str(r), WORLD_SIZE=str(world),
                                  MASTER_PORT=port, FUZZ_SEED=str(seed)), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
                 for r in range(world)]
        outs = [p.communicate(timeout=600)[0] for p in procs]
        line = [ln for ln in outs[0].splitlines() if ln.startswith("SHARDED-")]
        print(line[0] if line else f"seed {seed}: no verdict\n" + "\n".join(o[-1500:] for o in outs), flush=True)
        if any(p.returncode != 0 for p in procs) or not line or not line[0].startswith("SHARDED-OK"):
            sys.exit(1)
        ncase += 1
        seed += 1
    print(f"{ncase} cases, all equal to the oracle")
