"""Summarises `ncu -i report.ncu-rep --page raw --csv` per kernel (mean over the captured launches):
writes a markdown table to stdout and, with a second argument, the per-kernel traffic / instruction JSON
that bench.py reads (profiles/r1_traffic.json).   python tools/ncu_summary.py raw.csv [traffic.json]"""
import collections
import csv
import json
import statistics as st
import sys

WANT = {"gpu__time_duration.sum": "time", "dram__bytes_read.sum": "dram_rd", "dram__bytes_write.sum": "dram_wr",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct", "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_pct", "launch__registers_per_thread": "regs",
        "smsp__inst_executed.sum": "inst", "lts__t_sector_hit_rate.pct": "l2_hit"}


def main(path, out_json=None):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {k: hdr.index(k) for k in WANT if k in hdr}
    kn = hdr.index("Kernel Name")
    agg = collections.OrderedDict()

    def num(s):
        try:
            return float(s.replace(",", ""))
        except ValueError:
            return float("nan")
    for r in rows[2:]:
        a = agg.setdefault(r[kn], collections.defaultdict(list))
        for k, i in idx.items():
            v, u = num(r[i]), units[i]
            if k == "gpu__time_duration.sum":
                v = v / 1000 if u in ("ns", "nsecond") else (v * 1000 if u in ("ms", "msecond") else v)
            if k.startswith("dram__bytes"):
                v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
            a[WANT[k]].append(v)
    print("| kernel | launches | time us | dram rd MB | dram wr MB | dram % | issue active % | warps active % | regs | warp inst | L2 hit % |")
    print("|---|---|---|---|---|---|---|---|---|---|---|")
    out = {}
    for name, a in agg.items():
        m = {k: st.mean(v) for k, v in a.items()}
        short = name.split("(")[0].replace("void ", "").replace("<unnamed>::", "")
        print("| %s | %d | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %d | %.3g | %.1f |" % (
            short, len(a["time"]), m["time"], m["dram_rd"] / 1e6, m["dram_wr"] / 1e6, m["dram_pct"], m["issue_pct"], m["warps_pct"],
            m["regs"], m["inst"], m["l2_hit"]))
        out[short] = {"launches": len(a["time"]), "dram_bytes_per_launch": m["dram_rd"] + m["dram_wr"], "time_us": m["time"],
                      "warp_inst_per_launch": m["inst"], "issue_active_pct": m["issue_pct"]}
    if out_json:
        # bench.py only trusts these counters while the kernel sources are the ones that were profiled
        import hashlib
        import os
        csrc = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "event_3dgs_b200", "csrc")
        out["_source_sha256"] = {f: hashlib.sha256(open(os.path.join(csrc, f), "rb").read()).hexdigest()
                                 for f in ("render.cu", "binning.cu", "preprocess.cu", "common.cuh")}
        json.dump(out, open(out_json, "w"), indent=1)


if __name__ == "__main__":
    main(*sys.argv[1:3])
