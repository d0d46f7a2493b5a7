"""Compact JSON summary of an .ncu-rep (one record per profiled launch): duration, DRAM bytes and throughput, tensor
pipe / issue utilisation, registers, occupancy.   python tools/ncu_summary.py file.ncu-rep [out.json]"""
import csv
import json
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__cycles_active.avg", "sm__cycles_elapsed.max", "lts__t_bytes.sum"]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        rec = {"kernel": r[hdr.index("Kernel Name")][:110]}
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                rec[f"{k} [{units[i]}]"] = r[i]
        out.append(rec)
    text = json.dumps(out, indent=1)
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(text)
    else:
        print(text)


if __name__ == "__main__":
    main()
