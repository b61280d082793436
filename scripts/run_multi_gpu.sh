#!/bin/bash
# usage: bash scripts/run_multi_gpu.sh <tag>   (on a box with 8 GPUs): NCCL parity tests at 2/4/8 ranks, the C++ two-rank host
# test, then the BASELINE workloads that are defined on 8 GPUs.  One JSON line per run in gpurun_out/<tag>_*.json
tag=${1:-r02}
run() { # name gpus workload steps
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) \
    bench.py --gpus $2 --workload $3 --steps $4 --warmup 3 2> gpurun_out/${tag}_$1.err | tail -1 > gpurun_out/${tag}_$1.json
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${tag}_$1.json"))
    print("$1", "ms/step", round(d["ms_per_step"], 3), "frac", round(d["roofline_step"]["frac"], 3), "halo", {k: d["halo"][k] for k in ("ms_per_round", "achieved_gbs_per_direction", "frac_of_nvlink_900")},
          {k[:3]: round(v["ms_per_launch"], 3) for k, v in d["kernels"].items()}, "checksum h", d["checksum"]["sum_A_h"])
except Exception as e:
    print("$1 FAILED", e); print(open("gpurun_out/${tag}_$1.err").read()[-1500:])
PY
}
python -m pytest tests -m gpu -q -k "nccl_ranks or two_ranks" 2>&1 | tail -3
run config3_n8 8 config3 10
run config3_n4 4 config3 10
run config5_n8 8 config5 5
run config4_n8 8 config4 5
