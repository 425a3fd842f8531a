set -x
run() { # N workload extra...
  N=$1; WL=$2; shift 2
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --workload $WL --steps 10 --warmup 3 "$@" > gpurun_out/r02_n${N}_${WL}.json 2> gpurun_out/r02_n${N}_${WL}.err
  tail -c 400 gpurun_out/r02_n${N}_${WL}.err | tail -2
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r02_n${N}_${WL}.json").read().strip().splitlines()[-1])
    print("N=$N $WL", d.get("ms_per_step"), "e2e", d["e2e"]["value"], "ladder", d["roofline"]["launch_ms"], d["phases"], "T", d["triples"] and d["triples"]["seconds"], "parity", d["parity"]["passed"], "E", d["energy"])
except Exception as e:
    print("N=$N $WL FAILED", e)
PY
}
run 8 20,150
run 8 30,250
run 4 20,150
