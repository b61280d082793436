#!/bin/bash
# usage: bash scripts/run_scale.sh <tag> <gpus> [<gpus> ...]   (on a box with that many GPUs): the config-3 bench line at each N,
# one JSON line per run in gpurun_out/<tag>_n<N>.json.  No tests, no other workloads: the cheapest way to re-measure scaling.
tag=${1:-r02}
shift
for n in "$@"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) \
    bench.py --gpus $n --workload config3 --steps 10 --warmup 3 2> gpurun_out/${tag}_n$n.err | tail -1 > gpurun_out/${tag}_n$n.json
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${tag}_n$n.json"))
    print("N=$n", "ms/step", round(d["ms_per_step"], 3), "halo", round(d["halo"]["ms_per_round"], 3), d["halo"].get("tiles_rank0"),
          {k[:3]: round(v["ms_per_launch"], 3) for k, v in d["kernels"].items()}, "checksum h", d["checksum"]["sum_A_h"])
except Exception as e:
    print("N=$n FAILED", e); print(open("gpurun_out/${tag}_n$n.err").read()[-1500:])
PY
done
