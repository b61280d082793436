# sanity runs of the other BASELINE workloads (no parity check here: tests/ do that on small meshes)
for w in "$@"; do
timeout 500 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['config']['workload'][:60], round(d['ms_per_step'],2), round(d['roofline_step']['frac'],3), {k[:3]:round(v['ms_per_launch'],3) for k,v in d['kernels'].items()})"
done
