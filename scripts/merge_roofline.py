"""ncu metrics CSV of the bench launch -> profiles/r2_roofline_bench.json (merged by bench.py into `roofline`).

    python scripts/merge_roofline.py --metrics                      # the metric list for ncu
    python scripts/merge_roofline.py profiles/<csv> [workload] > profiles/r2_roofline_bench.json

The CSV comes from (B200_PROFILING.md recipe, after the same command exited 0 without ncu):
    ncu --clock-control none -k regex:render_ -s 1 -c 1 --csv --log-file <csv> --metrics $(python scripts/merge_roofline.py --metrics) \
        python scripts/prof_bench_launch.py synth1k_4k_256spp 2
Every number bench.py reports in `roofline` can be recomputed from that one file:
    frac          = sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed / 100
    executed flop = 2 ffma + fmul + fadd + 4 ffma2 + 2 fmul2 + 2 fadd2   (smsp__sass_thread_inst_executed_op_*_pred_on.sum;
                    a packed FFMA2 / FMUL2 / FADD2 thread instruction works on two FP32 lanes)
    traffic       = dram__bytes_read.sum + dram__bytes_write.sum
bench.py merges the file only if the kernel name equals the one it launched and the duration is within 5 %.
"""
import csv
import json
import os
import sys

OPS = ["ffma", "fmul", "fadd", "ffma2", "fmul2", "fadd2"]
METRICS = ["gpu__time_duration.sum", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed",
           "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed",
           "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
           *[f"smsp__sass_thread_inst_executed_op_{o}_pred_on.sum" for o in OPS],
           "lts__t_bytes.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.avg",
           "launch__registers_per_thread", "launch__block_size", "launch__grid_size", "launch__shared_mem_per_block_dynamic"]
UNIT = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "byte": 1.0, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3,
        "usecond": 1e-3, "msecond": 1.0, "nsecond": 1e-6, "second": 1e3}
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--metrics":
        sys.stdout.write(",".join(METRICS) + "\n")
        return
    path = sys.argv[1]
    workload = sys.argv[2] if len(sys.argv) > 2 else "synth1k_4k_256spp"
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    ik, im, iu, iv, iid = (hdr.index(k) for k in ("Kernel Name", "Metric Name", "Metric Unit", "Metric Value", "ID"))
    last = max(int(r[iid]) for r in rows[1:] if "render_" in r[ik])
    m, kernel = {}, None
    for r in rows[1:]:
        if int(r[iid]) != last:
            continue
        kernel = r[ik]
        v = float(r[iv].replace(",", ""))
        m[r[im]] = v * UNIT[r[iu]] if r[iu] in UNIT else v
    kernel = kernel.replace("void ", "").replace("srt::", "").split("(")[0].strip()
    sys.path.insert(0, ROOT)
    import bench
    _, w, h, spp, _ = bench.WORKLOADS[workload]
    op = {o: m.get(f"smsp__sass_thread_inst_executed_op_{o}_pred_on.sum", 0.0) for o in OPS}
    flop = 2 * op["ffma"] + op["fmul"] + op["fadd"] + 4 * op["ffma2"] + 2 * op["fmul2"] + 2 * op["fadd2"]
    out = {
        "workload": workload, "kernel": kernel, "duration_ms": m["gpu__time_duration.sum"],
        "fma_pipe_pct": m["sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed"],
        "inst_fma_pct": m["sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"],
        "alu_pipe_pct": m["sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed"],
        "issue_active_pct": m["sm__issue_active.avg.pct_of_peak_sustained_elapsed"],
        "lsu_wavefronts_pct": m.get("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
        "warp_inst": m["smsp__inst_executed.sum"], "thread_inst_by_op": op, "executed_flop": flop,
        "lts_bytes": m["lts__t_bytes.sum"], "dram_bytes": m["dram__bytes_read.sum"] + m["dram__bytes_write.sum"],
        "dram_read_bytes": m["dram__bytes_read.sum"], "dram_write_bytes": m["dram__bytes_write.sum"],
        "algorithmic_bytes": 36 * w * h, "registers": m.get("launch__registers_per_thread"), "block": m.get("launch__block_size"),
        "grid": m.get("launch__grid_size"), "smem_dynamic_bytes": m.get("launch__shared_mem_per_block_dynamic"),
        "source": os.path.relpath(os.path.abspath(path), ROOT),
    }
    json.dump(out, sys.stdout, indent=1)
    sys.stdout.write("\n")


if __name__ == "__main__":
    main()
