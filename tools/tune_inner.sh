# times the solve for SVDJ_INNER_SWEEPS = 1, 2, 3 at several sizes (bench.py lines reduced to time and sweeps)
for n in 1024 2048 4096 8192; do for inner in 1 2 3; do
  SVDJ_INNER_SWEEPS=$inner timeout 200 python bench.py --size $n --no-cpu-baseline --steps 2 --warmup 1 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('n=$n inner=$inner', round(d['value'],4), d['sweeps'], d['accuracy']['ok'])"
done; done
