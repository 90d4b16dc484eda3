# stack16g / stream64g on N GPUs of one box.  usage: bash tools/run_big.sh OUTDIR N [workloads]
O=$1; N=$2; W=${3:-"stack16g stream64g"}; mkdir -p $O
if [ "$N" = 1 ]; then T="python"; else T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"; fi
for w in $W; do
  timeout 1200 $T bench.py --gpus $N --steps 3 --warmup 3 --workload $w > $O/bench_${w}_${N}gpu.json 2> $O/bench_${w}_${N}gpu.err; echo "$w rc=$?"; cut -c1-250 $O/bench_${w}_${N}gpu.json; grep -i "error" $O/bench_${w}_${N}gpu.err | tail -2
done
