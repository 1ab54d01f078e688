"""tests/golden/beast_ref.npz: the bytes the reference's own dump1090 (oracle/_ref/dump1090_ref, compiled where it lies) sends
to a client of its Beast output port (--net --net-verbatim --net-bo-port, net_io.c:405-412,440-487) while it demodulates
tests/golden/c3_dense.npz from stdin.  The client connects before the first sample is written and --net-ro-size 0 makes the
reference flush after every message (its default keeps up to 1300 bytes back, which a file replay never sends), so nothing is missed.
Run in the build container."""
import os
import socket
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle1090 as O  # noqa: E402


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def reference_beast(iq: np.ndarray, flags) -> bytes:
    port = free_port()
    p = subprocess.Popen([O.REF_BIN, "--ifile", "-", "--quiet", "--net", "--net-verbatim", "--net-bind-address", "127.0.0.1",
                          "--net-bo-port", str(port), "--net-ro-port", "0", "--net-sbs-port", "0", "--net-bi-port", "0",
                          "--net-ri-port", "0", "--net-ro-size", "0", *flags], stdin=subprocess.PIPE, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    c = None
    for _ in range(200):
        try:
            c = socket.create_connection(("127.0.0.1", port), timeout=1.0)
            break
        except OSError:
            time.sleep(0.02)
    assert c is not None, "reference did not open its Beast port"
    time.sleep(0.3)                      # let its accept loop take the client before data flows
    p.stdin.write(iq.tobytes())
    p.stdin.close()
    data = bytearray()
    c.settimeout(20.0)
    while True:
        b = c.recv(1 << 16)
        if not b:
            break
        data += b
    c.close()
    p.wait(timeout=30)
    return bytes(data)


if __name__ == "__main__":
    with np.load(os.path.join(ROOT, "tests", "golden", "c3_dense.npz")) as z:
        iq = z["iq"]
    out = {}
    for name, flags in (("fix0", []), ("fix2", ["--fix-2bit"])):
        out[name] = np.frombuffer(reference_beast(iq, flags), np.uint8)
        print(name, out[name].size, "bytes")
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "beast_ref.npz"), **out)
