#!/usr/bin/env python
"""profiles/ncu_traffic.json from ncu --set full captures: DRAM bytes of ONE launch of each dense kernel over the
algorithmic bytes (2 B/sample) of that same launch.
usage: ncu_traffic.py <samples of the captured launch> front_kernel=rep message_kernel=rep ... [> profiles/ncu_traffic.json]"""
import csv, io, json, subprocess, sys

samples = int(sys.argv[1])
out = {"algorithmic_bytes_per_sample": 2, "samples_of_the_captured_launch": samples}
for arg in sys.argv[2:]:
    name, rep = arg.split("=", 1)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u, r = rows[0], rows[1], rows[2]

    def val(key):
        i = h.index(key)
        v = float(r[i].replace(",", ""))
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u[i]]
        return v * scale

    rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
    out[name] = {"source": rep.replace("gpurun_out/", "profiles/").replace(".ncu-rep", "_ncu_full.txt"), "dram_bytes_read": rd, "dram_bytes_write": wr,
                 "algorithmic_bytes": 2 * samples, "traffic_over_algorithmic": (rd + wr) / (2.0 * samples)}
print(json.dumps(out, indent=1))
