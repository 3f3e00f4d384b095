#!/bin/bash
# Round-end verification on one B200: GPU tests, smoke, reference arm, the default bench line (FCT + UNet + Autoencoder32K
# under "extra", torch-eager library bar), the other workloads, micro-benchmarks, the ncu launch list of the default
# command and the `ncu --set full` captures that feed roofline.traffic.
# Usage (from the repo root, on the GPU box): bash scripts/final_run.sh [outdir]
O=${1:-gpurun_out/fin}
mkdir -p $O
(timeout -k 10 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) > $O/gpu_tests.log
timeout -k 10 300 python __graft_entry__.py --smoke 2>&1 | tail -1 > $O/smoke.log
timeout -k 10 900 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference_arm.json 2> $O/ref.err
timeout -k 10 900 python bench.py --torch-eager --kernel-table > $O/bench_default.json 2> $O/kernels.txt
for w in latent2mask i2i2m autoencoder; do
  timeout -k 10 300 python bench.py --workload $w --no-cpu-baseline --no-extra > $O/bench_$w.json 2>/dev/null
done
for w in latent2mask i2i2m; do
  timeout -k 10 300 python bench.py --workload $w --infer --no-cpu-baseline --no-extra > $O/bench_${w}_infer.json 2>/dev/null
done
timeout -k 10 200 python scripts/attn_bench.py --tc 3 --tcbwd > $O/attn_bench_tc.txt 2>&1
(timeout -k 10 200 python scripts/conv_bench.py; timeout -k 10 200 python scripts/conv_bench.py --wgrad) > $O/conv_bench.txt 2>&1
timeout -k 10 200 python scripts/bn_bench.py > $O/bn_bench.txt 2>&1
[ -x scripts/build/umma_bench ] && scripts/build/umma_bench > $O/umma_bench.txt 2>&1
(timeout -k 10 200 python scripts/conv_bench.py --graph --only fct_; timeout -k 10 200 python scripts/conv_bench.py --wgrad --graph --only fct_,un_3x3_32; timeout -k 10 100 python scripts/dwln_bench.py) > $O/small_channel_bench.txt 2>&1
# launch list of the default command (cold-cache, serialised: shares, not absolutes), only after it ran clean above.
# About six GPU-minutes under ncu: NCU_LIST=1 to include it.
if [ "${NCU_LIST:-0}" = 1 ]; then
  timeout -k 10 700 ncu --metrics gpu__time_duration.sum --clock-control none -c 2600 --csv \
    --log-file $O/ncu_launches_fct.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra > $O/ncu.log 2>&1
  python scripts/ncu_launches.py $O/ncu_launches_fct.csv "FCT B=32, bench.py --steps 1 --warmup 3 --no-extra: exactly one step (between two adamw_kernel launches)" --step adamw > $O/ncu_launches_fct_summary.txt 2>&1
fi
# roofline.traffic: the dominant kernel of each workload at the bench shapes
timeout -k 10 300 ncu --set full --clock-control none -k regex:attn_bwd_tc_kernel -c 1 -o $O/attn_bwd_d4 python scripts/ncu_attn.py 4 tc 32 16384 > $O/ncu_attn4.log 2>&1
timeout -k 10 300 ncu --set full --clock-control none -k regex:attn_bwd_tc_kernel -c 1 -o $O/attn_bwd_d8 python scripts/ncu_attn.py 8 tc 32 4096 > $O/ncu_attn8.log 2>&1
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
