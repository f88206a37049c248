"""profiles/r2_shipped_capture.json: the numbers bench.py quotes under roofline.binding_limit, parsed from the committed
ncu summary of the shipped headline kernel (not literals in bench.py).
usage: python profiles/make_shipped_capture.py profiles/r2/ncu_r2_scan.summary.txt <paths> <steps>"""
import json, os, subprocess, sys

def main():
    path, paths, steps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    kv = {}
    kernel = None
    for line in open(path):
        if line.startswith("== kernel:"):
            kernel = line.split(":", 1)[1].strip()
        elif " = " in line:
            k, v = line.split(" = ", 1)
            kv[k.strip()] = v.strip()
    num = lambda k: float(kv[k].split()[0])
    unit = lambda k: {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[kv[k].split()[1]]
    warp_steps = (paths + 31) // 32 * steps
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    git = subprocess.run(["git", "-C", root, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    out = {"summary": os.path.relpath(path, root), "kernel": kernel, "git": git, "paths": paths, "steps": steps,
           "kernel_ms_under_ncu": round(num("gpu__time_duration.sum") * {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}[kv["gpu__time_duration.sum"].split()[1]], 6),
           "dram_bytes": num("dram__bytes_read.sum") * unit("dram__bytes_read.sum") + num("dram__bytes_write.sum") * unit("dram__bytes_write.sum"),
           "inst_executed": num("smsp__inst_executed.sum"), "alu_inst_executed": num("sm__inst_executed_pipe_alu.sum"),
           "fma_inst_executed": num("sm__inst_executed_pipe_fma.sum"),
           "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
           "warp_instructions_per_warp_step": round(num("smsp__inst_executed.sum") / warp_steps, 2),
           "active_lanes_per_instruction": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
           "alu_pipe_pct": num("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
           "fma_pipe_pct": num("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
           "lsu_pipe_pct": num("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active")}
    with open(os.path.join(root, "profiles", "r2_shipped_capture.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))

if __name__ == "__main__":
    main()
