"""ncu CSVs of every K4 launch of one mining pass -> profiles/k4_traffic.json (read by bench.py for roofline.traffic).

usage: python profiles/k4_traffic.py gpurun_out/ae_k4_traffic_c2.csv gpurun_out/ae_k4_traffic_c4.csv
"""
import collections
import csv
import io
import json
import os
import sys

SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "sector": 32.0}


def per_launch(path):
    rows = collections.OrderedDict()
    lines = [l for l in open(path) if not l.startswith("==")]
    for r in csv.DictReader(io.StringIO("".join(lines))):
        v = float(r["Metric Value"].replace(",", "")) * SCALE.get(r["Metric Unit"], 1.0)
        d = rows.setdefault(r["ID"], {"dram_bytes": 0.0})
        m = r["Metric Name"]
        if m.startswith("dram__bytes"):
            d["dram_bytes"] += v
        elif m.startswith("gpu__time_duration"):
            d["ms_under_ncu"] = round(v, 4)
        elif m.startswith("lts__t_sectors"):
            d["l2_read_bytes"] = v
    return list(rows.values())


def main():
    c2, c4 = per_launch(sys.argv[1]), per_launch(sys.argv[2])
    heavy = max(c4, key=lambda d: d["dram_bytes"])
    out = {
        "source": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,gpu__time_duration.sum "
                  "--clock-control none -k regex:k_count_groups2 over ONE mining pass of the final code (profiles/gpu_session_ae.sh, round 2); "
                  "every K4 launch of the pass is listed",
        "dram_bytes_per_step": sum(d["dram_bytes"] for d in c2),
        "launches_captured": len(c2),
        "c2_1M_rows": {"per_launch": c2, "note": "levels 3..21 of BASELINE configs[1]; the whole vertical database (125 MB) fits in L2, so HBM sees "
                                                 "each needed row chunk about once per level"},
        "c4_shape_20M_rows": {"per_launch": c4, "dram_bytes_per_pass": sum(d["dram_bytes"] for d in c4),
                              "heaviest_launch": dict(heavy, dram_gbs=round(heavy["dram_bytes"] / heavy["ms_under_ncu"] / 1e6, 1)),
                              "note": "configs[3] shape at 20 M rows on one GPU: the database (1438 items x 2.5 MB = 3.6 GB) is 29x the L2"},
    }
    json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "k4_traffic.json"), "w"), indent=1)
    print("c2:", len(c2), "launches,", round(out["dram_bytes_per_step"] / 1e9, 3), "GB DRAM per pass;  c4 shape:", len(c4), "launches,",
          round(out["c4_shape_20M_rows"]["dram_bytes_per_pass"] / 1e9, 3), "GB; heaviest", out["c4_shape_20M_rows"]["heaviest_launch"])


if __name__ == "__main__":
    main()
