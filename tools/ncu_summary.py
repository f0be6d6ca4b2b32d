#!/usr/bin/env python
"""Text summary of one `ncu --set full` capture for profiles/: the headline raw metrics of the kernel plus the source-line
attribution of tools/ncu_lines.py.

usage: ncu_summary.py REPORT.ncu-rep LIBRARY.so KERNEL_SUBSTRING > profiles/<name>_summary.txt
(needs ncu, cuobjdump and nvdisasm on PATH; the library must be the build that was profiled)
"""
import csv, io, os, subprocess, sys, tempfile

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__icc_request_hit_rate.pct",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_atom.sum",
]


def main():
    rep, lib, kernel = sys.argv[1:4]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u, v = rows[0], rows[1], rows[2]
    print(f"# {os.path.basename(rep)} -- kernel {v[h.index('Kernel Name')][:90]}")
    for k in KEYS:
        if k in h:
            i = h.index(k)
            print(f"{k:75s} {v[i]} {u[i]}")
    print("stall cycles per issued instruction (smsp__average_warps_issue_stalled_*_per_issue_active):")
    for i, k in enumerate(h):
        if "issue_stalled" in k and "per_issue_active" in k and "not_issued" not in k:
            try:
                if float(v[i]) >= 0.05:
                    print(f"  {k.split('issue_stalled_')[1].split('_per_issue')[0]:24s} {float(v[i]):.2f}")
            except ValueError:
                pass
    with tempfile.TemporaryDirectory() as td:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=td, capture_output=True)
        cubins = [os.path.join(td, f) for f in os.listdir(td) if f.endswith(".cubin")]
        dis = os.path.join(td, "dis.txt")
        with open(dis, "w") as f:
            subprocess.run(["nvdisasm", "-gi", "-c"] + cubins, stdout=f, stderr=subprocess.DEVNULL)
        sass = os.path.join(td, "sass.csv")
        with open(sass, "w") as f:
            subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], stdout=f,
                           stderr=subprocess.DEVNULL)
        print()
        sys.stdout.flush()
        subprocess.run([sys.executable, os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_lines.py"), sass, dis,
                        kernel, "45"])


if __name__ == "__main__":
    main()
