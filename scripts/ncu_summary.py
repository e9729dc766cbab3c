"""Summarise an `ncu --set full` report exported with `ncu -i X.ncu-rep --page raw --csv`: one column per captured launch."""
import csv
import re
import sys

ROWS = [
    ("duration (ms)", "gpu__time_duration.sum", lambda v, u: "%.3g" % (float(v) * {"us": 1e-3, "ms": 1.0, "ns": 1e-6}[u])),
    ("regs/thread", "launch__registers_per_thread", None), ("grid", "launch__grid_size", None), ("block", "launch__block_size", None),
    ("warp instructions (M)", "smsp__inst_executed.sum", lambda v, u: "%.3g" % (float(v) / 1e6)),
    ("issue active %", "smsp__issue_active.avg.pct_of_peak_sustained_active", "%.1f"),
    ("occupancy %", "sm__warps_active.avg.pct_of_peak_sustained_active", "%.1f"),
    ("active lanes/instr", "smsp__thread_inst_executed_per_inst_executed.ratio", "%.1f"),
    ("ALU pipe %", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "%.1f"),
    ("FMA pipe %", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "%.1f"),
    ("XU pipe %", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "%.1f"),
    ("tensor ops % of peak (elapsed)", "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed", "%.1f"),
    ("DRAM read (MB)", "dram__bytes_read.sum", lambda v, u: "%.3g" % (float(v) * {"Mbyte": 1, "Kbyte": 1e-3, "Gbyte": 1e3, "byte": 1e-6}[u])),
    ("DRAM write (MB)", "dram__bytes_write.sum", lambda v, u: "%.3g" % (float(v) * {"Mbyte": 1, "Kbyte": 1e-3, "Gbyte": 1e3, "byte": 1e-6}[u])),
    ("L1 hit %", "l1tex__t_sector_hit_rate.pct", "%.1f"), ("L2 hit %", "lts__t_sector_hit_rate.pct", "%.1f"),
    ("SM throughput %", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "%.1f"),
    ("memory throughput %", "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed", "%.1f"),
]
STALLS = ["long_scoreboard", "wait", "selected", "not_selected", "branch_resolving", "no_instructions", "short_scoreboard", "math_pipe_throttle",
          "mio_throttle", "lg_throttle", "barrier"]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    names = []
    for r in data:
        n = r[col["Kernel Name"]].split("(")[0].replace("void ", "")
        names.append("`%s`" % n)
    print("| metric | " + " | ".join(names) + " |")
    print("|---|" + "---|" * len(names))
    for label, key, fmt in ROWS:
        if key not in col:
            continue
        i = col[key]
        vals = []
        for r in data:
            v = r[i].replace(",", "")
            try:
                vals.append(fmt(v, units[i]) if callable(fmt) else ((fmt % float(v)) if fmt else v))
            except Exception:
                vals.append(v)
        print("| %s | " % label + " | ".join(vals) + " |")
    for s in STALLS:
        key = "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio" % s
        tot = "smsp__average_warp_latency_per_inst_issued.ratio"
        if key in col and tot in col:
            vals = []
            for r in data:
                try:
                    vals.append("%.1f" % (100.0 * float(r[col[key]]) / float(r[col[tot]])))
                except Exception:
                    vals.append("-")
            print("| stall %s %% | " % s + " | ".join(vals) + " |")
    inst = sum(float(r[col["smsp__inst_executed.sum"]].replace(",", "")) for r in data)
    dur = sum(float(r[col["gpu__time_duration.sum"]].replace(",", "")) * {"us": 1e-3, "ms": 1.0, "ns": 1e-6}[units[col["gpu__time_duration.sum"]]] for r in data)
    print("\nsum of warp instructions: %.0f M; sum of durations (serialised, cold): %.3f ms" % (inst / 1e6, dur))
    for r in data:
        print("%s: dram %.1f MB, inst %.2f M" % (r[col["Kernel Name"]].split("(")[0], (float(r[col["dram__bytes_read.sum"]]) + float(r[col["dram__bytes_write.sum"]])) *
              {"Mbyte": 1, "Kbyte": 1e-3, "Gbyte": 1e3, "byte": 1e-6}[units[col["dram__bytes_read.sum"]]] if units[col["dram__bytes_read.sum"]] == units[col["dram__bytes_write.sum"]] else float("nan"),
              float(r[col["smsp__inst_executed.sum"]]) / 1e6))


def to_json(path, out, source):
    """profiles/ncu_frame.json for bench.py: DRAM bytes (read + write) and executed warp instructions per launch of every frame kernel,
    keyed by the names the library's profiler uses (nrc_kernel_name)."""
    import json
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}
    kernels = {}
    for r in data:
        full = r[col["Kernel Name"]]
        n = full.split("(")[0].replace("void ", "")
        if n.startswith("k_light<"):
            n = "k_light<render>" if n.startswith("k_light<1") or n.startswith("k_light<true") else "k_light<train>"
        n = re.sub(r"<.*>$", "", n) if not n.startswith("k_light<") else n
        dram = float(r[col["dram__bytes_read.sum"]].replace(",", "")) * scale[units[col["dram__bytes_read.sum"]]] + \
            float(r[col["dram__bytes_write.sum"]].replace(",", "")) * scale[units[col["dram__bytes_write.sum"]]]
        inst = float(r[col["smsp__inst_executed.sum"]].replace(",", ""))
        if n not in kernels:      # the first captured launch of each kernel
            kernels[n] = [dram, inst]
    json.dump({"source": source, "kernels": kernels}, open(out, "w"), indent=1)


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[2] == "--json":
        to_json(sys.argv[1], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else sys.argv[1])
    else:
        main(sys.argv[1])
