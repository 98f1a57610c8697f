"""ncu launch list (CSV) of bench.py steps -> profiles/r02_traffic.json, the file bench.py's `roofline.traffic` reads.

    HPE_B200_CUDA_GRAPH=0 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
        -k regex:'conv_gemm|yhalo|preprocess|decode' -c 400 --csv --log-file gpurun_out/launches.csv \
        python bench.py --quick --steps 2 --warmup 3 --no-cpu-baseline
    python tools/ncu_traffic.py gpurun_out/launches.csv 256x256x192x17 [steps_in_capture]

The JSON carries the hash of the sources the capture was taken from (bench.source_hash): bench.py refuses a capture
of other sources (traffic: null + a warning) instead of repeating a stale constant.
"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def source_hash():
    import hashlib
    h = hashlib.sha256()
    pkg = os.path.join(ROOT, "2d-human-pose-estimation_b200")
    files = sorted(os.path.join(pkg, "csrc", f) for f in os.listdir(os.path.join(pkg, "csrc")))
    files += [os.path.join(pkg, f) for f in ("conv.py", "models.py")]
    for f in files:
        with open(f, "rb") as fh:
            h.update(os.path.basename(f).encode() + b"\0" + fh.read())
    return h.hexdigest()[:16]


def main():
    path, workload = sys.argv[1], sys.argv[2]
    rows = {}
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        d = rows.setdefault(int(r["ID"]), {"kernel": r["Kernel Name"]})
        d[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
    conv = [d for d in rows.values() if "conv_gemm_kernel" in d["kernel"] or "conv_yhalo_kernel" in d["kernel"]]
    if not conv:
        raise SystemExit("no conv launches in the capture")
    n = len(conv)
    rd = sum(d.get("dram__bytes_read.sum", 0.0) for d in conv)
    wr = sum(d.get("dram__bytes_write.sum", 0.0) for d in conv)
    ns = sum(d.get("gpu__time_duration.sum", 0.0) for d in conv)
    ns_all = sum(d.get("gpu__time_duration.sum", 0.0) for d in rows.values())
    out = {"workload": workload, "source_hash": source_hash(), "conv_launches_in_capture": n,
           "dram_bytes_per_conv_launch": round((rd + wr) / n), "dram_read_bytes": round(rd), "dram_write_bytes": round(wr),
           "conv_share_of_captured_time": round(ns / ns_all, 4) if ns_all else None,
           "mean_conv_launch_us_under_ncu": round(ns / n / 1e3, 2),
           "how": f"dram__bytes_read.sum + dram__bytes_write.sum averaged over the {n} conv launches of {os.path.basename(path)} "
                  "(ncu --clock-control none, un-graphed steps of bench.py --quick)"}
    dst = os.path.join(ROOT, "profiles", "r02_traffic.json")
    with open(dst, "w") as f:
        json.dump(out, f, indent=1)
        f.write("\n")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
