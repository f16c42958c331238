#!/bin/bash
# tools/scale.sh N workload... : sharded bench lines at N GPUs into gpurun_out/$TAG_bench_<workload>_sharded<N>.json (TAG defaults to r1e)
cd "$(dirname "$0")/.."
N=$1; shift
for w in "$@"; do
  out=gpurun_out/${TAG:-r1e}_bench_${w}_sharded${N}
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 2 --warmup 2 --workload $w --sharded > $out.json 2> $out.err
  python - <<PY
import json
try:
    l=json.loads(open("$out.json").read().strip().splitlines()[-1])
    print("$w N=$N", "iter/s %.2f" % l["value"], "e2e %.2f" % l["e2e"]["value"], {k: round(v,2) for k,v in l["phase_ms_per_iteration"].items()}, l["sharded"], "chi2", l["final_chi2"])
except Exception as e:
    print("$w N=$N failed", e); print(open("$out.err").read()[-2000:])
PY
done
