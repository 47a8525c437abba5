"""Summarise an .ncu-rep (read here, no GPU needed) into a small JSON kept under profiles/.

    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep profiles/ncu_<name>_rNN.json [--stalls]
"""
import csv
import json
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg", "sm__cycles_active.avg",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum.per_second",
        "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu.sum",
        "lts__t_sectors_op_write.sum", "lts__t_sectors_op_read.sum"]


def main(rep, out, stalls=False):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    res = []
    for r in data:
        d = {"kernel": r[hdr.index("Kernel Name")] if "Kernel Name" in hdr else ""}
        for i, h in enumerate(hdr):
            name = h.split(".TriageCompute.")[-1]
            if name in KEYS and r[i] != "":
                d[f"{name} [{units[i]}]"] = r[i]
            elif stalls and "issue_stalled" in name and name.endswith("per_issue_active.ratio") and r[i] != "":
                d.setdefault("stalls_per_issue", {})[name.split("issue_stalled_")[1].split("_per_issue")[0]] = r[i]
        res.append(d)
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res[0], indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], stalls=len(sys.argv) > 3 and sys.argv[3] == "--stalls")
