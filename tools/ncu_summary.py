#!/usr/bin/env python
"""Per-kernel summary of an .ncu-rep (raw page): time, DRAM traffic, issue utilisation, occupancy, pipes, stalls."""
import csv, os, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "sm__cycles_elapsed.max"]
for r in rows[2:]:
    print("---- " + r[ix["Kernel Name"]][:100])
    for k in KEYS:
        if k in ix:
            print(f"  {k:72s} {r[ix[k]][:22]:>22s} {units[ix[k]]}")
    st = [(h, float(r[i])) for h, i in ix.items() if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") and r[i] not in ("", "n/a")]
    st.sort(key=lambda x: -x[1])
    print("  stalls/issue: " + ", ".join(f"{h[34:-23]}={v:.2f}" for h, v in st[:7]))

# machine-readable digest for bench.py's roofline (`traffic`, `issue`): profiles/<round>_ncu_kernels.json
if len(sys.argv) > 2:
    import json, re
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from bench import csrc_sha16          # the capture belongs to THIS source state: bench.py drops it when the sources change
    digest = {"source": os.path.basename(rep), "csrc_sha16": os.environ.get("CSRC_SHA16") or csrc_sha16()}
    def num(r, k):
        try:
            return float(r[ix[k]].replace(",", ""))
        except Exception:
            return None
    def to_bytes(r, k):
        v, u = num(r, k), units[ix[k]]
        return None if v is None else v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    def to_us(r, k):
        v, u = num(r, k), units[ix[k]]
        return None if v is None else v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1)
    for r in rows[2:]:
        m = re.search(r"(\w+_kernel)", r[ix["Kernel Name"]])
        if not m or m.group(1) in digest:
            continue
        digest[m.group(1)] = {"time_us": to_us(r, "gpu__time_duration.sum"),
                              "dram_bytes": int((to_bytes(r, "dram__bytes_read.sum") or 0) + (to_bytes(r, "dram__bytes_write.sum") or 0)),
                              "warp_inst": int(num(r, "smsp__inst_executed.sum") or 0),
                              "issue_active_pct": num(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                              "warps_active_pct": num(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
                              "registers": num(r, "launch__registers_per_thread")}
    json.dump(digest, open(sys.argv[2], "w"), indent=1)
