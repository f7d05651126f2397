"""Reads an .ncu-rep (here, without a GPU) and writes the per-launch summary CSV kept under profiles/.
    python profiles/summarize_ncu.py gpurun_out/r1b_enc.ncu-rep profiles/r1_encoder_layer_ncu_full_summary.csv
With a third argument it also derives the DRAM traffic per launch of each kernel class from that CSV (what
bench.py quotes as roofline.traffic), with the git head recorded:
    python profiles/summarize_ncu.py gpurun_out/r2_enc_layer.ncu-rep profiles/r2_encoder_layer_ncu_full_summary.csv profiles/r2_traffic.json"""
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


def traffic(summary_csv, out_json):
    import json

    rows = list(csv.reader(open(summary_csv)))
    head = rows[0]
    ir = [i for i, h in enumerate(head) if h.startswith("dram__bytes_read.sum")][0]
    iw = [i for i, h in enumerate(head) if h.startswith("dram__bytes_write.sum")][0]

    def to_bytes(v, h):
        return float(v) * (1e9 if "Gbyte" in h else 1e6 if "Mbyte" in h else 1e3 if "Kbyte" in h else 1.0)

    per = {"gemm": [], "attention": [], "layernorm": []}
    for r in rows[1:]:
        b = to_bytes(r[ir], head[ir]) + to_bytes(r[iw], head[iw])
        per["gemm" if "gemm_kernel" in r[0] else "attention" if "attention" in r[0] else "layernorm"].append(b)
    git = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    doc = {k: sum(v) / len(v) for k, v in per.items() if v}
    doc.update({"source": f"{summary_csv} (ncu --set full; dram__bytes_read.sum + dram__bytes_write.sum, mean per launch of the class)",
                "captured_at_git_head": git,
                "algorithmic_bytes_per_launch_gemm": "mean of the four encoder shapes: A + W read, C written (bf16) / fp32 residual updated in place = 99.7e6"})
    json.dump(doc, open(out_json, "w"), indent=1)
    print(doc)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
    if len(sys.argv) > 3:
        traffic(sys.argv[2], sys.argv[3])
