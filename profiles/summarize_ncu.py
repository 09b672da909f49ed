"""Summarise an ncu report (read here, on the CPU box):
    python profiles/summarize_ncu.py gpurun_out/prof.ncu-rep profiles/rNN_name <pairs per launch> [note]
writes <out>_summary.csv (one row per captured launch, the metrics the roofline argument uses)
and updates profiles/traffic.json (per-kernel DRAM bytes per launch, scaled note included)."""
import csv, json, os, subprocess, sys

KEYS = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "gpu__time_duration.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
        "sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
        "smsp__sass_average_data_bytes_per_sector_mem_global_op_st.ratio",
        "smsp__sass_average_data_bytes_per_sector_mem_global_op_ld.ratio"]

def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    keys = [k for k in KEYS if k in idx]
    with open(out + "_summary.csv", "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(keys)
        w.writerow([units[idx[k]] for k in keys])
        for r in data:
            w.writerow([r[idx[k]] for k in keys])
    def to_bytes(v, u):
        return float(v) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[u]
    traffic = {}
    for r in data:
        name = r[idx["Kernel Name"]]
        # family name as bench.py labels it: unqualified, no template arguments
        short = name.replace("void ", "").replace("<unnamed>::", "").split("(")[0].split("<")[0].strip()
        b = (to_bytes(r[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]]) +
             to_bytes(r[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]]))
        traffic.setdefault(short, []).append(b)
    tj = os.path.join(os.path.dirname(os.path.abspath(__file__)), "traffic.json")
    pairs = float(sys.argv[3]) if len(sys.argv) > 3 else None      # (user, variant) pairs per captured launch
    note = sys.argv[4] if len(sys.argv) > 4 else ""
    per_launch = {k: v[0] for k, v in traffic.items()}        # the first captured launch of each family
    json.dump({"source": os.path.basename(out), "note": note, "pairs_per_captured_launch": pairs,
               "dram_bytes_per_launch": per_launch,
               "dram_bytes_per_pair": {k: v / pairs for k, v in per_launch.items()} if pairs else None},
              open(tj, "w"), indent=1)
    print(open(out + "_summary.csv").read())

if __name__ == "__main__":
    main()
