#!/bin/sh
# Multi-GPU bench lines of the named configurations (run on an N-GPU box): sh tools/run_scaling.sh "N workload [flags]" ...
# e.g.  sh tools/run_scaling.sh "8 40,400 --no-triples" "4 40,400 --no-triples" "8 40,300 --precision SP --no-triples"
for spec in "$@"; do
  set -- $spec
  N=$1; WL=$2; shift 2
  TAG=$(echo "n${N}_${WL}_$*" | tr -d ' -' | tr ',' 'v')
  if [ "$N" = "1" ]; then
    python bench.py --workload $WL --steps 5 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/r02_$TAG.json 2> gpurun_out/r02_$TAG.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --workload $WL --steps 5 --warmup 3 "$@" > gpurun_out/r02_$TAG.json 2> gpurun_out/r02_$TAG.err
  fi
  echo "rc=$? $TAG"; grep -v "^\*\|OMP_NUM\|NCCL version" gpurun_out/r02_$TAG.err | tail -3
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r02_$TAG.json").read().strip().splitlines()[-1])
    print("$TAG", "ms", d.get("ms_per_step"), "frac", d.get("tflops_frac_of_peak"), "e2e", d["e2e"]["value"], "ladder", d["roofline"]["launch_ms"], d["roofline"]["frac"], d["phases"]["residuals_incl_exchange_ms"], d["phases"]["diis_incl_allgather_ms"], "T", d["triples"] and d["triples"]["seconds"], "hbm", d["hbm_peak_gb_per_rank"], "parity", d["parity"]["passed"], d["parity"]["max_abs_dr2"], d["parity"]["iteration_max_abs_dt2"], "E", d["energy"])
except Exception as e:
    print("$TAG FAILED", e)
PY
done
