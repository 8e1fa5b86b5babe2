"""ncu CSV (long format, --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum) -> the per-launch JSON that
bench.py's `roofline.traffic` reads (profiles/rN_full_size_dram_traffic.json).  usage: traffic_to_json.py in.csv out.json [commit]"""
import csv
import json
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3, "ns": 1e-6, "us": 1e-3, "ms": 1.0}


def main():
    src, dst = sys.argv[1], sys.argv[2]
    commit = sys.argv[3] if len(sys.argv) > 3 else "?"
    lines = open(src, newline="").read().split("\n")
    start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
    rows = {}
    for r in csv.DictReader(lines[start:]):
        if not r.get("Metric Name"):
            continue
        name = r["Kernel Name"].replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
        name = name.split("(")[0].split("<")[0].replace("void ", "").replace("h2b::", "").strip()
        rec = rows.setdefault(int(r["ID"]), {"id": int(r["ID"]), "kernel": name})
        val = float(r["Metric Value"].replace(",", "")) * UNIT.get(r["Metric Unit"], 1.0)
        name = r["Metric Name"]
        if name == "dram__bytes_read.sum":
            rec["dram_read_bytes"] = val
        elif name == "dram__bytes_write.sum":
            rec["dram_write_bytes"] = val
        elif name == "gpu__time_duration.sum":
            rec["duration_ms"] = val
    kernels = [rows[k] for k in sorted(rows)]
    summary = {}
    for r in kernels:
        s = summary.setdefault(r["kernel"], {"launches": 0, "dram_read_bytes": 0.0, "dram_write_bytes": 0.0, "duration_ms": 0.0})
        s["launches"] += 1
        for key in ("dram_read_bytes", "dram_write_bytes", "duration_ms"):
            s[key] += r.get(key, 0.0)
    json.dump({"commit": commit, "command": "scripts/capture_traffic.sh (ncu per-launch DRAM bytes; cold-cache, serialised launches)",
               "summary_per_kernel_name": summary, "kernels": kernels}, open(dst, "w"), indent=1)
    for k, s in summary.items():
        print(f"{k:40s} x{s['launches']:4d}  read {s['dram_read_bytes'] / 1e9:8.3f} GB  write {s['dram_write_bytes'] / 1e9:8.3f} GB  {s['duration_ms']:9.3f} ms")


if __name__ == "__main__":
    main()
