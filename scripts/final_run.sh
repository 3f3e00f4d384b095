#!/bin/bash
# Round-end verification on one B200: GPU tests, smoke, reference arm, bench lines, ncu launch list.
# Usage (from the repo root, on the GPU box): bash scripts/final_run.sh [outdir]
O=${1:-gpurun_out/fin}
mkdir -p $O
(timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) > $O/gpu_tests.log
python __graft_entry__.py --smoke 2>&1 | tail -1 > $O/smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference_arm.json 2> $O/ref.err
python bench.py --kernel-table > $O/bench_ae32k_image.json 2> $O/kernels_ae32k_image.txt
python bench.py --workload unet --no-cpu-baseline --kernel-table > $O/bench_unet.json 2> $O/kernels_unet.txt
for w in latent2mask i2i2m autoencoder; do
  python bench.py --workload $w --no-cpu-baseline > $O/bench_$w.json 2>/dev/null
done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 2200 -c 700 --csv \
  --log-file $O/ncu_launches_ae32k.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu.log 2>&1
cat $O/gpu_tests.log $O/smoke.log
python - <<PY
import json, glob
for f in sorted(glob.glob("$O/*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["value"], d["e2e"]["value"])
    except Exception as e:
        print(f, "ERR", e)
PY
