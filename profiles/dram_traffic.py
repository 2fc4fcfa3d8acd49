"""DRAM traffic per launch of the conv kernels from an ncu pass over one step (the `roofline.traffic` of bench.py).
    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:conv_ --csv \
        --log-file gpurun_out/dram.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-graph
    python profiles/dram_traffic.py gpurun_out/dram.csv [label-suffix] >> merges into profiles/r02_dram_traffic.json
Stores, per kernel name, the MEAN of (bytes read + bytes written) over its launches."""
import csv
import json
import os
import re
import sys
from collections import defaultdict

HERE = os.path.dirname(os.path.abspath(__file__))


def main(path, suffix=""):
    rows = [r for r in csv.reader(l for l in open(path, errors="replace") if l.startswith('"'))]
    hdr = rows[0]
    ik, im, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows[1:]:
        if not r[im].startswith("dram__bytes_"):
            continue
        v = float(r[iv].replace(",", ""))
        mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(r[iu], 1.0)
        name = re.sub(r"^void ", "", r[ik])
        name = re.sub(r"[<(].*$", "", name).split("::")[-1]
        tot[name] += v * mult
        if r[im].startswith("dram__bytes_read"):
            cnt[name] += 1
    out_path = os.path.join(HERE, "r02_dram_traffic.json")
    out = json.load(open(out_path)) if os.path.exists(out_path) else {}
    for k in tot:
        out[k + suffix] = round(tot[k] / max(cnt[k], 1))
        out[k + suffix + " launches"] = cnt[k]
    json.dump(out, open(out_path, "w"), indent=1, sort_keys=True)
    print(json.dumps({k: out[k] for k in out if suffix in k or not suffix}, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
