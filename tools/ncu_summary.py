"""Text summary of an `ncu --set full` report for profiles/: the metrics DESIGN.md and bench.py cite.
Usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep "header line" [--traffic-json profiles/mlsa_traffic.json] > profiles/x_summary.txt
With --traffic-json the dram read + write bytes and the duration of the first k_mlsa kernel of the report are written
there together with the current commit: bench.py prints them as roofline.traffic (so the figure always says which
capture it belongs to)."""
import csv
import json
import subprocess
import sys

KEEP = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size",
        "launch__registers_per_thread", "launch__waves_per_multiprocessor", "launch__occupancy_limit",
        "sm__cycles_elapsed.max", "smsp__cycles_active.avg", "smsp__inst_executed.sum",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_bytes.sum", "dram__throughput.avg.pct_of_peak")
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
print(sys.argv[2] if len(sys.argv) > 2 else sys.argv[1])
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print(f"Kernel Name [] = {d.get('Kernel Name')}")
    print(f"Block Size [] = {d.get('Block Size')}")
    print(f"Grid Size [] = {d.get('Grid Size')}")
    for h, u in zip(hdr, units):
        if any(h.startswith(k) for k in KEEP) and d.get(h, "") not in ("", "n/a"):
            print(f"{h} [{u}] = {d[h]}")
    print()

if "--traffic-json" in sys.argv:
    path = sys.argv[sys.argv.index("--traffic-json") + 1]
    unit = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        if "k_mlsa" not in d.get("Kernel Name", ""):
            continue
        u = dict(zip(hdr, units))
        tot = sum(float(d[k].replace(",", "")) * unit[u[k]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        ms = float(d["gpu__time_duration.sum"].replace(",", "")) * {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}[u["gpu__time_duration.sum"]]
        commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
        json.dump({"bytes": tot, "kernel": d["Kernel Name"][:60], "kernel_ms_under_ncu": ms, "report": sys.argv[1],
                   "commit_at_summary": commit, "metric": "dram__bytes_read.sum + dram__bytes_write.sum, one launch, ncu --set full"},
                  open(path, "w"), indent=1)
        break
