"""Summarise `ncu --page raw --csv` exports (tools/gpu_round_ncu.sh) into one JSON / text table:
per captured launch the duration, DRAM bytes, achieved DRAM GB/s against the measured copy peak
(MEASURED_PEAKS.json: 6548.8 GB/s), L2 hit rate, FP64 / tensor pipe activity, registers, grid.
  python tools/ncu_summary2.py gpurun_out/r2g_ncu_*_raw.csv > profiles/r02_ncu_kernels.json"""
import csv
import json
import sys

PEAK = 6548.8
WANT = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active": "fp64_pipe_pct",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active": "fp64_pipe_cycles_pct",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_pct",
    "sm__pipe_tensor_op_dmma_cycles_active.avg.pct_of_peak_sustained_active": "dmma_pipe_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "occupancy_pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "smem_bank_conflicts",
}
UNIT = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def num(x):
    try:
        return float(x.replace(",", ""))
    except Exception:
        return None


out = []
for path in sys.argv[1:]:
    rows = [r for r in csv.reader(l for l in open(path) if not l.startswith("=="))]
    if len(rows) < 3:
        continue
    head, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(head)}
    for r in rows[2:]:
        rec = {"file": path.split("/")[-1], "kernel": r[col["Kernel Name"]][:90]}
        for m, key in WANT.items():
            if m in col:
                v = num(r[col[m]])
                if v is None:
                    continue
                u = units[col[m]]
                if key in ("duration", "dram_read", "dram_write") and u in UNIT:
                    v *= UNIT[u]
                rec[key] = v
        if rec.get("duration"):
            b = rec.get("dram_read", 0.0) + rec.get("dram_write", 0.0)
            rec["dram_bytes"] = b
            rec["dram_gbs"] = b / rec["duration"] / 1e9
            rec["frac_of_measured_hbm_peak"] = rec["dram_gbs"] / PEAK
            rec["duration_us"] = rec.pop("duration") * 1e6
        out.append(rec)
json.dump(out, sys.stdout, indent=1)
