set -x
mkdir -p gpurun_out/p7
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
for w in fl rl64; do
  $B --workload $w > gpurun_out/p7/plain_$w.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/p7/launches_$w.csv $B --workload $w > gpurun_out/p7/ncu_l_$w.log 2>&1
done
for w in fl rl4 rl64 rl4096; do
  $B --workload $w > gpurun_out/p7/plain2_$w.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'fl_encode|fl_decode|rl_encode|rl_decode' -s 6 -c 2 -o gpurun_out/p7/full_$w $B --workload $w > gpurun_out/p7/ncu_f_$w.log 2>&1
done
ls -la gpurun_out/p7
