for K in rj rj2 rj2x16 rj2x32; do for C in 65536 262144; do
python bench.py --workload coal_rj --kernel $K --chains $C --transitions 200 --also none --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$K',$C,d['value']/1e9,d['ms_per_step'])"
done; done
