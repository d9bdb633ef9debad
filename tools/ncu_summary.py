"""Summarise an .ncu-rep (raw page): one line per captured kernel with the roofline-relevant metrics."""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]
want = {"Kernel Name": "name", "gpu__time_duration.sum": "us", "dram__bytes_read.sum": "rd", "dram__bytes_write.sum": "wr",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor%",
        "sm__warps_active.avg.pct_of_peak_sustained_active": "warps%", "launch__registers_per_thread": "regs",
        "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue%", "launch__grid_size": "grid",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram%",
        "lts__t_bytes.sum": "l2bytes", "l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum": "ldg_bytes",
        "smsp__average_warp_latency_per_inst_issued.ratio": "lat/inst",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm%", "lts__throughput.avg.pct_of_peak_sustained_elapsed": "l2%",
        "launch__cluster_dim_x": "cl"}
idx = {h: i for i, h in enumerate(hdr)}
units = rows[1]
for r in rows[2:]:
    parts = []
    for k, lab in want.items():
        if k in idx:
            v = r[idx[k]]
            if lab == "name":
                v = v.split("(")[0][-60:]
            u = units[idx[k]]
            parts.append(f"{lab}={v}{u if lab in ('us','rd','wr','l2bytes') else ''}")
    print(" ".join(parts))
