#!/usr/bin/env python
"""Summarise an `ncu --set full ... --page raw --csv` export per kernel: duration, DRAM bytes, achieved DRAM GB/s,
occupancy, issue activity.  usage: python scripts/ncu_full_summary.py gpurun_out/full_raw.csv [--json out.json]"""
import collections
import csv
import json
import sys


def unit_scale(u):
    return {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1.0)


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}

    def val(r, name):
        if name not in ix or r[ix[name]] in ("", "n/a"):
            return 0.0
        return float(r[ix[name]].replace(",", "")) * unit_scale(units[ix[name]])

    stall_cols = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    agg = collections.OrderedDict()
    for r in data:
        name = r[ix["Kernel Name"]]
        short = name.split("(")[0].replace("<unnamed>::", "").replace("void ", "")
        a = agg.setdefault(short, {"n": 0, "ms": 0.0, "rd": 0.0, "wr": 0.0, "warps": 0.0, "issue": 0.0, "regs": 0})
        a["n"] += 1
        a["ms"] += val(r, "gpu__time_duration.sum")
        a["rd"] += val(r, "dram__bytes_read.sum")
        a["wr"] += val(r, "dram__bytes_write.sum")
        a["warps"] += val(r, "sm__warps_active.avg.pct_of_peak_sustained_active")
        a["issue"] += val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active")
        a["regs"] = int(val(r, "launch__registers_per_thread"))
        st = a.setdefault("stalls", collections.Counter())
        for h in stall_cols:
            st[h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")] += val(r, h)
        a["lanes"] = a.get("lanes", 0.0) + val(r, "smsp__thread_inst_executed_per_inst_executed.ratio")
        a["inst"] = a.get("inst", 0.0) + val(r, "smsp__inst_executed.sum")
    out = {}
    print(f"# {sys.argv[1]}: per kernel, averages over its launches in the capture (ncu --set full, cold caches, clocks not locked)")
    print(f"# {'kernel':44s} {'n':>3s} {'ms':>8s} {'dram rd MB':>11s} {'dram wr MB':>11s} {'GB/s':>8s} {'warps%':>7s} {'issue%':>7s} {'regs':>5s}")
    for k, a in sorted(agg.items(), key=lambda x: -x[1]["ms"]):
        n = a["n"]
        ms, rd, wr = a["ms"] / n, a["rd"] / n, a["wr"] / n
        gbs = (rd + wr) / (ms * 1e-3) / 1e9 if ms > 0 else 0
        print(f"  {k[:44]:44s} {n:3d} {ms:8.4f} {rd / 1e6:11.2f} {wr / 1e6:11.2f} {gbs:8.1f} {a['warps'] / n:7.1f} {a['issue'] / n:7.1f} {a['regs']:5d}")
        top = [(name, round(v / n, 2)) for name, v in a["stalls"].most_common(4) if v > 0]
        print(f"      stalls per issue: " + ", ".join(f"{nm} {v}" for nm, v in top) + f"; lanes/inst {a['lanes'] / n:.1f}; warp inst {a['inst'] / n / 1e6:.1f} M")
        out[k] = {"launches": n, "ms": round(ms, 5), "dram_read_bytes": int(rd), "dram_write_bytes": int(wr), "dram_GBps": round(gbs, 1),
                  "warps_active_pct": round(a["warps"] / n, 1), "issue_active_pct": round(a["issue"] / n, 1), "registers": a["regs"],
                  "stalls_per_issue": dict(top), "threads_per_warp_inst": round(a["lanes"] / n, 1), "warp_inst": int(a["inst"] / n)}
    if "--json" in sys.argv:
        json.dump(out, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)


if __name__ == "__main__":
    main()
