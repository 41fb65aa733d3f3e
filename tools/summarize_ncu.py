"""Summarise an ncu report (run in the build container, no GPU): key metrics per profiled launch -> JSON + text.

    python tools/summarize_ncu.py gpurun_out/prof_r1.ncu-rep profiles/r1_score_kernel
"""
import csv
import io
import json
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__ops_path_tensor_src_fp64.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_issued.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "lts__t_sector_hit_rate.pct",
    "smsp__inst_executed.sum", "sm__cycles_elapsed.avg.per_second", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    launches = []
    for r in rows[2:]:
        rec = {}
        for h, u, v in zip(hdr, units, r):
            if h in ("Kernel Name", "ID") or h in KEYS:
                rec[h] = (v + (" " + u if u else "")) if h in KEYS else v
        launches.append(rec)
    for rec in launches:
        try:
            rd = float(rec["dram__bytes_read.sum"].split()[0]) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[rec["dram__bytes_read.sum"].split()[1]]
            wr = float(rec["dram__bytes_write.sum"].split()[0]) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[rec["dram__bytes_write.sum"].split()[1]]
            rec["dram_bytes_per_launch"] = rd + wr
        except Exception:
            pass
    json.dump(launches, open(out + ".json", "w"), indent=1)
    with open(out + ".txt", "w") as f:
        for rec in launches:
            f.write("== %s (launch id %s)\n" % (rec.get("Kernel Name", "?")[:80], rec.get("ID", "?")))
            for k, v in rec.items():
                if k not in ("Kernel Name", "ID"):
                    f.write("   %-75s %s\n" % (k, v))
    print(open(out + ".txt").read())


if __name__ == "__main__":
    main()
