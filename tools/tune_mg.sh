# multi-GPU driver: pair ordering by size (SVDJ_MG_SORT) x inner sweeps, timed through bench.py on N GPUs
# usage: tools/tune_mg.sh N "sizes"
N=${1:-2}; SIZES=${2:-"4096"}
for n in $SIZES; do for sort in 0 1; do for inner in 1 2; do
  SVDJ_MG_SORT=$sort SVDJ_INNER_SWEEPS=$inner timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
    --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --size $n --steps 2 --warmup 1 2>/dev/null | \
    python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=$N n=$n sort=$sort inner=$inner', round(d['value'],4), d['sweeps'], d['accuracy']['ok'])"
done; done; done
