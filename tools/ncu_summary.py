"""Summarise an .ncu-rep (raw page) into a small text file for profiles/: one block per captured launch
with the metrics the roofline arguments use.   python tools/ncu_summary.py <rep> <out.txt> [title]
With --traffic <key>=<rows_per_launch>[,<key>=<rows>...] the DRAM bytes of the LAST captured launch whose kernel name
contains <key> are also written into profiles/ncu_traffic.json, the file bench.py takes roofline.traffic from."""
import json
import os
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__cycles_active.avg", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum",
        "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_xu.sum"]


UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    argv = list(sys.argv)
    traffic = {}
    if "--traffic" in argv:
        i = argv.index("--traffic")
        for kv in argv[i + 1].split(","):
            k, v = kv.split("=")
            traffic[k] = int(v)
        del argv[i:i + 2]
    rep, out = argv[1], argv[2]
    title = argv[3] if len(argv) > 3 else rep
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    found = {}
    lines = ["# " + title, "# source: ncu --set full --clock-control none (cold-cache, serialised replays; not a bench number)", ""]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        lines.append("kernel: " + d.get("Kernel Name", "?")[:140])
        for k in KEYS:
            if k in d and d[k] != "":
                lines.append("  %-66s %s %s" % (k, d[k], units[hdr.index(k)]))
        rd, wr = d.get("dram__bytes_read.sum"), d.get("dram__bytes_write.sum")
        for key, rows_per in traffic.items():
            if key in d.get("Kernel Name", "") and rd not in (None, ""):
                ur, uw = units[hdr.index("dram__bytes_read.sum")], units[hdr.index("dram__bytes_write.sum")]
                found[key] = {"rows_per_launch": rows_per, "dram_bytes_read": float(rd) * UNIT.get(ur, 1.0),
                              "dram_bytes_write": float(wr) * UNIT.get(uw, 1.0), "source": os.path.relpath(out)}
        lines.append("")
    open(out, "w").write("\n".join(lines) + "\n")
    if found:
        path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_traffic.json")
        cur = json.load(open(path)) if os.path.exists(path) else {}
        cur.update(found)
        json.dump(cur, open(path, "w"), indent=1)
    print("\n".join(lines[:40]))


if __name__ == "__main__":
    main()
