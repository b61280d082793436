# usage: bash scripts/sweep_fct.sh "VAR=a VAR2=b" "VAR=c" ...   (one bench run per argument)
for cfg in "$@"; do
echo -n "$cfg: "; env $cfg python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d[\"ms_per_step\"],2), {k[:3]:round(v[\"ms_per_launch\"],3) for k,v in d[\"kernels\"].items()})"
done
