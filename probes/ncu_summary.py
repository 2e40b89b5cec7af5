"""Reduce an ``ncu --page raw --csv`` export to the few counters the bench line and DESIGN.md quote.

    python probes/ncu_summary.py profiles/pixel_fused_r01c_ncu_raw.csv pixel_fused 16 > profiles/pixel_fused_ncu_summary.json

Output: one JSON object per matching kernel row (first match is what ``bench.py`` reads for ``roofline.traffic``)."""
import csv
import json
import sys

KEEP = ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__cycles_elapsed.avg", "sm__cycles_elapsed.avg")
SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}


def main():
    path, pattern = sys.argv[1], sys.argv[2]
    batch = int(sys.argv[3]) if len(sys.argv) > 3 else None
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    out = []
    for r in rows[2:]:
        if pattern not in r[col["Kernel Name"]]:
            continue
        d = {"kernel": r[col["Kernel Name"]].split("(")[0], "grid": r[col["Grid Size"]], "block": r[col["Block Size"]],
             "source": path, "batch": batch}
        for k in KEEP:
            if k in col and r[col[k]] not in ("", "n/a"):
                v = float(r[col[k]].replace(",", ""))
                u = units[col[k]]
                if u in SCALE and "bytes" in k:
                    v *= SCALE[u]
                    u = "byte"
                d[k] = v
                d[k + ".unit"] = u
        if "dram__bytes_read.sum" in d:
            d["dram_bytes_per_launch"] = d["dram__bytes_read.sum"] + d["dram__bytes_write.sum"]
        out.append(d)
    json.dump(out, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
