#!/bin/bash
# Round-end verification on one B200: GPU tests, smoke, reference arm, the default bench line (FCT + UNet + Autoencoder32K
# under "extra", torch-eager library bar), the other workloads, the ncu launch list of the default command and one
# `ncu --set full` capture of the dominant kernel's launches of one step (feeds roofline.traffic).
# Usage (from the repo root, on the GPU box): bash scripts/final_run.sh [outdir]
O=${1:-gpurun_out/fin}
mkdir -p $O
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) > $O/gpu_tests.log
python __graft_entry__.py --smoke 2>&1 | tail -1 > $O/smoke.log
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference_arm.json 2> $O/ref.err
python bench.py --torch-eager --kernel-table > $O/bench_default.json 2> $O/kernels.txt
for w in latent2mask i2i2m autoencoder; do
  python bench.py --workload $w --no-cpu-baseline --no-extra > $O/bench_$w.json 2>/dev/null
done
for w in latent2mask i2i2m; do
  python bench.py --workload $w --infer --no-cpu-baseline --no-extra > $O/bench_${w}_infer.json 2>/dev/null
done
# launch list of the default command (cold-cache, serialised: shares, not absolutes), only after it ran clean above
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv \
  --log-file $O/ncu_launches_fct.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra > $O/ncu.log 2>&1
python scripts/ncu_launches.py $O/ncu_launches_fct.csv "FCT B=32, bench.py --steps 1 --warmup 3 --no-extra" > $O/ncu_launches_fct_summary.txt 2>&1
cat $O/gpu_tests.log $O/smoke.log
python - <<PY
import json, glob
for f in sorted(glob.glob("$O/*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["value"], d["e2e"]["value"])
        for k, v in (d.get("extra") or {}).items():
            print("   ", k, v["value"], v["e2e"]["value"])
    except Exception as e:
        print(f, "ERR", e)
PY
