# RL encoder round: tests, timing, ncu (plain runs first).  usage: bash tools/prof_rle.sh TAG
TAG=${1:-x}
O=gpurun_out/$TAG; mkdir -p $O
timeout 900 python -m pytest tests/test_rl_gpu.py -x -q -m gpu > $O/pytest_rl.log 2>&1; echo "pytest rc=$?" >> $O/pytest_rl.log; tail -5 $O/pytest_rl.log
timeout 300 python tools/quick_rl.py 1073741824 $MEANS > $O/quick_rl.log 2>&1; echo "rc=$?" >> $O/quick_rl.log; cat $O/quick_rl.log
for V in $VARIANTS; do
  FLRL_LIB_PATH=$PWD/fl-rl-compression-mpi_b200/variants/libflrl_$V.so timeout 300 python tools/quick_rl.py 1073741824 $MEANS > $O/quick_rl_$V.log 2>&1; cat $O/quick_rl_$V.log
done
if [ -n "$NCU" ]; then
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
for w in $NCU; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:'rl_encode|rl_decode' -s 6 -c 2 -o $O/full_$w $B --workload $w > $O/ncu_f_$w.log 2>&1; tail -2 $O/ncu_f_$w.log
done
fi
ls -la $O
if [ -n "$NCU1" ]; then
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:'rl_encode' -s 3 -c 1 -o $O/full_mean1 python tools/quick_rl.py 268435456 1 > $O/ncu_mean1.log 2>&1; tail -3 $O/ncu_mean1.log
fi
