"""DRAM traffic per launch of every kernel in an ncu `--set full` report -> profiles/<tag>_traffic.json, the file bench.py reads
its roofline.traffic from (the newest committed one that names the kernel):

    python tools/ncu_traffic.py gpurun_out/r4e_frame.ncu-rep profiles/r4e_traffic.json"""
import csv
import io
import json
import re
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units = rows[0], rows[1]
col = {n: i for i, n in enumerate(hdr)}


def num(r, name):
    v = float(r[col[name]].replace(",", ""))
    u = units[col[name]]
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "sector": 1, "": 1}.get(u, 1)


commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
kernels = {}
for r in rows[2:]:
    name = re.sub(r"\(.*", "", r[col["Kernel Name"]]).replace("void ", "").replace("sbm::", "").replace("(int)", "").replace("(bool)", "")
    base = name.split("<")[0]
    e = {"kernel": name, "dram_bytes_per_launch": int(num(r, "dram__bytes_read.sum") + num(r, "dram__bytes_write.sum")),
         "us_under_ncu": round(num(r, "gpu__time_duration.sum"), 2), "l2_sectors": int(num(r, "lts__t_sectors.sum"))}
    kernels[base if base not in kernels else name] = e
json.dump({"commit": commit,
           "how": "ncu --set full --clock-control none, second blocking match of tools/one_frame.py (M1 frame x 360 templates), "
                  "dram__bytes_read.sum + dram__bytes_write.sum per launch", "kernels": kernels}, open(out, "w"), indent=1)
print(json.dumps(kernels, indent=1))
