"""CPU: adsbcu_1090_replay_sharded at world_size 2 and 3 over gloo -- the N>1 path of bench.py (orchestrator + torch.distributed
transport), with the kernels running on the SIMT emulator build of the same sources (tests/emu, test infrastructure).
The merged message set must equal the reference's (tests/golden/c3_dense.npz was produced by the reference binary)."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_replay_sharded_over_gloo(world):
    sys.path.insert(0, os.path.join(ROOT, "tests", "emu"))
    import build_emu
    build_emu.build()                                   # once, before the ranks race for it
    from oracle import oracle1090 as O
    if not os.path.exists(O.LIB_PATH):
        O.build()
    port = free_port()
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), ADSBCU_TEST_EMU="1")
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "emu", "run_sharded_rank.py")], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=600)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out)
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    assert "SHARDED-OK" in outs[0], outs[0]
