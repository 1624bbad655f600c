"""profiles/decode_traffic.json from an `ncu --set full` capture of the decode kernel, exported with
`ncu -i X.ncu-rep --page raw --csv`.

    python tools/make_decode_traffic.py profiles/r03_decode_ncu_full_metrics.csv [--round 3]

bench.py reports the file's `dram_bytes_per_launch` as `roofline.traffic` -- but only while `kernel_source_sha256`
still equals the digest of the decode kernel's sources (bench.decode_kernel_digest), so a number measured on an older
kernel is dropped (null) instead of going stale silently.
"""
import argparse
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--round", type=int, default=3)
    ap.add_argument("--algorithmic-bytes", type=int, default=1074270464)   # bs=64 ctx=4096 16q/8kv, DESIGN §3.2
    args = ap.parse_args()
    import bench
    rows = list(csv.reader(open(args.csv)))
    hdr, units, data = rows[0], rows[1], [r for r in rows[2:] if "paged_decode_kernel" in r[rows[0].index("Kernel Name")]]
    assert data, "no paged_decode_kernel launch in the capture"

    def col(name):
        i = hdr.index(name)
        return [float(r[i]) * UNIT[units[i]] for r in data]
    rd, wr, us = col("dram__bytes_read.sum"), col("dram__bytes_write.sum"), col("gpu__time_duration.sum")
    out = {
        "kernel": data[0][hdr.index("Kernel Name")],
        "dram_bytes_per_launch": int(round(sum(a + b for a, b in zip(rd, wr)) / len(rd))),
        "dram_bytes_read": [int(round(x)) for x in rd],
        "dram_bytes_write": [int(round(x)) for x in wr],
        "algorithmic_bytes_per_launch": args.algorithmic_bytes,
        "gpu_time_us": [round(x, 2) for x in us],
        "kernel_source_sha256": bench.decode_kernel_digest(),
        "source": f"{os.path.relpath(args.csv, ROOT)} (round {args.round}, ncu --set full --clock-control none on bench.py's "
                  f"headline workload, bs=64 ctx=4096 16q/8kv; dram__bytes_read.sum + dram__bytes_write.sum per launch, "
                  f"mean of the {len(rd)} captured launches)",
    }
    with open(os.path.join(ROOT, "profiles", "decode_traffic.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
