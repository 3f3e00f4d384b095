"""Summarise an `ncu --set full --import-source on` report per kernel: headline pipe / issue figures, opcode mix and the
instructions that collect the most stall samples.  Usage: python scripts/ncu_src_summary.py report.ncu-rep [kernel-regex]"""
import collections
import csv
import io
import subprocess
import sys


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    return [dict(zip(r[0], x)) for x in r[2:]]


KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__waves_per_multiprocessor",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.avg", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__t_bytes.sum"]


def source(rep, kernel):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kernel],
                         capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    h = next(x for x in r if x and x[0] == "Address")
    rows = [dict(zip(h, x)) for x in r if len(x) == len(h) and x[0].startswith("0x")]
    # one launch only
    seen, one = set(), []
    for x in rows:
        if x["Address"] in seen:
            break
        seen.add(x["Address"])
        one.append(x)
    return h, one


def main():
    rep = sys.argv[1]
    pat = sys.argv[2] if len(sys.argv) > 2 else ""
    for d in raw(rep):
        name = d["Kernel Name"]
        if pat and pat not in name:
            continue
        print("=" * 100)
        print(name[:160])
        for k in KEYS:
            if k in d:
                print(f"  {k:72s} {d[k]}")
        stalls = {k: float(v) for k, v in d.items() if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and v not in ("", "n/a")}
        print("  stall cycles per issued instruction:", ", ".join(f"{k.split('issue_stalled_')[1].split('_per_')[0]} {v:.2f}"
                                                                  for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:8]))
    if not pat:
        return
    h, rows = source(rep, pat)
    stall = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
    tot = sum(int(x["# Samples"]) for x in rows)
    agg = {k: sum(int(x[k]) for x in rows) for k in stall}
    print(f"\nsource view: {len(rows)} SASS instructions, {tot} samples")
    print("  samples by reason:", ", ".join(f"{k[6:]} {v}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v))
    ops, smp = collections.Counter(), collections.Counter()
    for x in rows:
        t = [o for o in x["Source"].split() if not o.startswith("@")]
        op = t[0].split(".")[0]
        ops[op] += int(x["Instructions Executed"])
        smp[op] += int(x["# Samples"])
    n = sum(ops.values())
    print(f"  warp instructions executed: {n}")
    for k, v in ops.most_common(24):
        print(f"    {k:10s} {v:12d} {100.0 * v / n:5.1f} %   samples {smp[k]}")
    print("  hottest instructions:")
    for i, x in sorted(sorted(enumerate(rows), key=lambda ix: -int(ix[1]["# Samples"]))[:30]):
        s = {k[6:]: int(x[k]) for k in stall if int(x[k]) > 0.2 * int(x["# Samples"])}
        print(f"    {i:5d} {x['Source'].strip()[:64]:64s} {x['# Samples']:>6s} x{x['Instructions Executed']:>10s} {s}")


if __name__ == "__main__":
    main()
