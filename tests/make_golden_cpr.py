#!/usr/bin/env python
"""Writes tests/golden/cpr_vectors.npz: the test vectors of the reference's own dump1090/cprtests.c (inputs and the
expected results its authors wrote down, tolerance 1e-6 as in that test) plus the exact doubles the reference's cpr.c
(oracle/_ref/libcpr_ref.so, built by oracle/Makefile) returns for them and for a seeded random set.
Needs /root/reference; run in the build container only:  python tests/make_golden_cpr.py"""
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oraclecpr as P   # noqa: E402

SRC = "/root/reference/dump1090/cprtests.c"


def rows(text, name):
    """The initialiser rows of one static test table, each evaluated to a tuple of numbers."""
    body = text[text.index(name + "[] = {"):]
    body = body[body.index("{") + 1:body.index("};")]
    body = re.sub(r"//[^\n]*", "", body)
    out = []
    for m in re.finditer(r"\{([^{}]*)\}", body):
        if not re.fullmatch(r"[\d\s.,+\-]*", m.group(1)):
            raise ValueError("unexpected token in " + m.group(1))
        out.append(tuple(eval(x) for x in m.group(1).split(",") if x.strip()))   # plain arithmetic like 0.176601 - 180.0
    return out


text = open(SRC).read()
rec, want = [], []
for r in rows(text, "cprGlobalAirborneTests"):
    elat, elon, olat, olon, eres, erlat, erlon, ores, orlat, orlon = r
    rec += [(0, 0, elat, elon, olat, olon, 0, P.AIRBORNE), (0, 0, elat, elon, olat, olon, 1, P.AIRBORNE)]
    want += [(erlat, erlon, eres), (orlat, orlon, ores)]
for r in rows(text, "cprGlobalSurfaceTests"):
    reflat, reflon, elat, elon, olat, olon, eres, erlat, erlon, ores, orlat, orlon = r
    rec += [(reflat, reflon, elat, elon, olat, olon, 0, P.SURFACE), (reflat, reflon, elat, elon, olat, olon, 1, P.SURFACE)]
    want += [(erlat, erlon, eres), (orlat, orlon, ores)]
for r in rows(text, "cprRelativeTests"):
    reflat, reflon, clat, clon, fflag, surface, res, rlat, rlon = r
    rec.append((reflat, reflon, clat, clon, 0, 0, fflag, P.RELATIVE_SURFACE if surface else P.RELATIVE_AIRBORNE))
    want.append((rlat, rlon, res))
rec = np.array(rec, dtype=P.IN_DTYPE)
want = np.array(want, dtype=[("lat", "<f8"), ("lon", "<f8"), ("result", "<i4")])
ref = P.decode_ref(rec)
assert np.array_equal(ref["result"], want["result"])
assert np.all(np.abs(ref["lat"] - want["lat"]) <= 1e-6) and np.all(np.abs(ref["lon"] - want["lon"]) <= 1e-6)
rnd = P.random_records(4096, 0xC0FFEE)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "cpr_vectors.npz"), vectors=rec, expected=want, reference=ref,
                    random_in=rnd, random_reference=P.decode_ref(rnd))
print(len(rec), "vectors from cprtests.c,", rnd.size, "random records")
