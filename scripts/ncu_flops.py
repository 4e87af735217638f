"""Executed fp64 operation counts of a kernel launch from an `ncu --set full` report ("exact counts ...
over the final device functions").

    python scripts/ncu_flops.py gpurun_out/r01_k_eval_st_v3.ncu-rep <units in the launch>

Counts thread-level DFMA (2 flops), DMUL and DADD (1 flop each) instructions that were executed with their predicate on
(`smsp__sass_thread_inst_executed_op_d*_pred_on.sum.per_cycle_elapsed` x `smsp__cycles_elapsed.max`), i.e. what the
SASS of the final build really executes, including redundant hyper-dual work; fp64 transcendental sequences are
counted through the DFMA/DMUL/DADD instructions they expand to.
"""
import csv
import subprocess
import sys


def main():
    rep, units = sys.argv[1], float(sys.argv[2])
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(out.splitlines()))
    row = dict(zip(r[0], r[-1]))
    f = lambda k: float(row[k].replace(",", ""))
    cyc = f("smsp__cycles_elapsed.max")
    ops = {k: f(f"smsp__sass_thread_inst_executed_op_{k}_pred_on.sum.per_cycle_elapsed") * cyc for k in ("dfma", "dmul", "dadd")}
    flops = 2 * ops["dfma"] + ops["dmul"] + ops["dadd"]
    ms = f("gpu__time_duration.sum")
    unit = r[1][r[0].index("gpu__time_duration.sum")]
    ms = ms / 1e6 if unit.startswith("n") else ms / 1e3 if unit.startswith("u") else ms
    print(f"{rep}: {ms:.3f} ms, {cyc / 1e6:.2f} M cycles ({cyc / ms / 1e6:.3f} GHz)")
    print(f"  thread instructions: DFMA {ops['dfma']:.4g}  DMUL {ops['dmul']:.4g}  DADD {ops['dadd']:.4g}")
    print(f"  executed fp64 flops: {flops:.4g} = {flops / units / 1e3:.1f} kflop per unit ({units:.0f} units), "
          f"{flops / (ms * 1e-3) / 1e12:.2f} TFLOP/s during the launch")
    print(f"  warp instructions: {f('smsp__inst_executed.sum'):.4g} = {f('smsp__inst_executed.sum') / units:.0f} per unit")


if __name__ == "__main__":
    main()
