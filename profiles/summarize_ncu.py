"""Reads an .ncu-rep (here, without a GPU) and writes the per-launch summary CSV kept under profiles/.
    python profiles/summarize_ncu.py gpurun_out/r1b_enc.ncu-rep profiles/r1_encoder_layer_ncu_full_summary.csv"""
import csv
import io
import subprocess
import sys

COLS = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active"]


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, body = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(head)}
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["Kernel Name", "Grid Size"] + [f"{c} [{units[idx[c]]}]" if c in idx else c for c in COLS])
        for r in body:
            w.writerow([r[idx["Kernel Name"]][:60], r[idx["Grid Size"]]] + [r[idx[c]] if c in idx else "" for c in COLS])
    print(open(out).read())


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
