"""Summarises the round's ncu captures (scripts/profile_r02.sh -> gpurun_out/r02_*_raw.csv, r02_launches.csv) into
profiles/r02_ncu_summary.md and profiles/r02_ncu_units.json (per-unit DRAM bytes / executed fp64 flops that bench.py
multiplies by its own live unit counts; stamped with the hash of the kernel sources they were captured for).

    python scripts/ncu_summary.py [gpurun_out] [profiles]
"""
import csv
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
dst = sys.argv[2] if len(sys.argv) > 2 else "profiles"


def raw(name):
    r = list(csv.reader(open(os.path.join(src, f"r02_{name}_raw.csv"))))
    hdr, units, row = r[0], r[1], r[-1]
    d = {}
    for h, u, v in zip(hdr, units, row):
        try:
            d[h] = (float(v.replace(",", "")), u)
        except ValueError:
            d[h] = (v, u)
    return d


def val(d, k, scale=None):
    v, u = d[k]
    if scale == "ms":
        return v / 1e6 if u.startswith("n") else v / 1e3 if u.startswith("u") else v * 1e3 if u == "s" else v
    if scale == "bytes":
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    return v


def one(name, units, unit_name):
    d = raw(name)
    ms = val(d, "gpu__time_duration.sum", "ms")
    cyc = val(d, "smsp__cycles_elapsed.max")
    ops = {k: val(d, f"smsp__sass_thread_inst_executed_op_{k}_pred_on.sum.per_cycle_elapsed") * cyc for k in ("dfma", "dmul", "dadd")}
    flops = 2 * ops["dfma"] + ops["dmul"] + ops["dadd"]
    dram = val(d, "dram__bytes_read.sum", "bytes") + val(d, "dram__bytes_write.sum", "bytes")
    grid = val(d, "launch__grid_size")
    out = {
        "kernel": str(d["Kernel Name"][0]), "launch_ms": ms, "grid": grid, "block": val(d, "launch__block_size"),
        "registers": val(d, "launch__registers_per_thread"), "units": units, "unit": unit_name,
        "dram_bytes_per_unit": dram / units, "executed_flops_per_unit": flops / units,
        "warp_inst_per_unit": val(d, "smsp__inst_executed.sum") / units,
        "dram_gbs": dram / (ms * 1e-3) / 1e9, "executed_tflops": flops / (ms * 1e-3) / 1e12,
        "warps_active_pct": val(d, "sm__warps_active.avg.pct_of_peak_sustained_active"),
        "issue_active_pct": val(d, "smsp__issue_active.avg.pct") if "smsp__issue_active.avg.pct" in d else None,
        "fp64_pipe_pct": val(d, "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active") if "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active" in d else None,
        "dram_pct": val(d, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        "local_load_sectors": val(d, "smsp__inst_executed_op_local_ld.sum") if "smsp__inst_executed_op_local_ld.sum" in d else None,
        "smem_bank_conflicts": val(d, "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum") if "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum" in d else None,
    }
    return out


def main():
    import bench_nmpc as bn
    B, N, M = 65536, 20, 35
    rows = {
        "k_eval_st": one("k_eval_st", B * N, "stage evaluation"),
        "k_eval_pm": one("k_eval_pm", B * M, "stage evaluation"),
        "k_step_a": one("k_step_a", B, "instance-iteration"),
        "k_step_b": one("k_step_b", B, "instance-iteration"),
        "k_update": one("k_update", B, "instance-iteration"),
        "k_rollout_st_jac": one("k_rollout", 262144 * 50, "state-step"),
    }
    for k in ("k_step_a_wide", "k_step_b_wide"):
        d = raw(k)
        rows[k] = one(k, val(d, "launch__grid_size"), "instance-iteration (late pass, one block per instance)")
    step = {"dram_bytes_per_unit": sum(rows[k]["dram_bytes_per_unit"] for k in ("k_step_a", "k_step_b", "k_update")),
            "executed_flops_per_unit": sum(rows[k]["executed_flops_per_unit"] for k in ("k_step_a", "k_step_b", "k_update")),
            "unit": "instance-iteration", "parts": ["k_step_a", "k_step_b", "k_update"]}
    units = {"source_hash": bn.source_hash(), "captured_by": "scripts/profile_r02.sh (ncu --set full --clock-control none, pass 3 of a 65,536 batch)",
             "k_step": step, **{k: {"dram_bytes_per_unit": v["dram_bytes_per_unit"], "executed_flops_per_unit": v["executed_flops_per_unit"], "unit": v["unit"]}
                                for k, v in rows.items()}}
    os.makedirs(dst, exist_ok=True)
    with open(os.path.join(dst, "r02_ncu_units.json"), "w") as f:
        json.dump(units, f, indent=1)
    # launch list: share of each kernel over the captured launches
    L = list(csv.reader(open(os.path.join(src, "r02_launches.csv"))))
    hi = next(i for i, r in enumerate(L) if r and r[0] == "ID")
    hdr = L[hi]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot = {}
    for r in L[hi + 1:]:
        if len(r) <= vi:
            continue
        name = r[ki].split("(")[0].split("::")[-1]
        v = float(r[vi].replace(",", ""))
        v = v / 1e6 if r[ui].startswith("n") else v / 1e3 if r[ui].startswith("u") else v
        t = tot.setdefault(name, [0, 0.0]); t[0] += 1; t[1] += v
    allms = sum(t[1] for t in tot.values())
    with open(os.path.join(dst, "r02_ncu_summary.md"), "w") as f:
        f.write("# r02 -- ncu summaries (one B200, `scripts/profile_r02.sh`, C3 batch 65,536; `--clock-control none`)\n\n")
        f.write("Per-launch times under ncu are cold-cache and serialised: compare SHARES with bench.py's CUDA-event shares, not absolutes.\n\n")
        f.write("## launch list (first 420 launches = passes 0..~83 of one solve; `gpu__time_duration.sum`)\n\n| kernel | launches | total ms | share |\n|---|---|---|---|\n")
        for k, (n, ms) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| {k} | {n} | {ms:.2f} | {ms / allms:.3f} |\n")
        f.write("\n## `ncu --set full` of one launch per kernel (pass 3: all 65,536 instances active; wide kernels: launch 150)\n\n")
        f.write("| kernel | ms | regs | units | DRAM B/unit | DRAM GB/s (% of peak) | executed fp64 flop/unit | TFLOP/s | warp inst/unit | warps active % | fp64 pipe % | issue active % |\n|---|---|---|---|---|---|---|---|---|---|---|---|\n")
        for k, v in rows.items():
            f.write(f"| {k} | {v['launch_ms']:.3f} | {v['registers']:.0f} | {v['units']:.0f} {v['unit']} | {v['dram_bytes_per_unit']:.0f} | "
                    f"{v['dram_gbs']:.0f} ({v['dram_pct']:.1f}) | {v['executed_flops_per_unit']:.0f} | {v['executed_tflops']:.2f} | {v['warp_inst_per_unit']:.0f} | "
                    f"{v['warps_active_pct']:.1f} | {v['fp64_pipe_pct'] if v['fp64_pipe_pct'] is not None else float('nan'):.1f} | "
                    f"{v['issue_active_pct'] if v['issue_active_pct'] is not None else float('nan'):.1f} |\n")
        f.write(f"\nk_step (= k_step_a + k_step_b + k_update) per instance-iteration: {step['dram_bytes_per_unit'] / 1e3:.1f} KB of DRAM traffic, "
                f"{step['executed_flops_per_unit'] / 1e3:.1f} kflop executed.\nSource hash of the captured build: `{units['source_hash']}`.\n")
    print(open(os.path.join(dst, "r02_ncu_summary.md")).read())


if __name__ == "__main__":
    main()
