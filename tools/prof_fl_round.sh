set -x
mkdir -p gpurun_out/p8
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
timeout 300 python bench.py --workload fl --steps 10 --warmup 3 > gpurun_out/p8/bench_fl.json 2> gpurun_out/p8/bench_fl.err
$B --workload fl > gpurun_out/p8/plain_fl.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/p8/launches_fl.csv $B --workload fl > gpurun_out/p8/ncu_l_fl.log 2>&1
$B --workload fl > gpurun_out/p8/plain2_fl.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'fl_encode|fl_decode' -s 6 -c 2 -o gpurun_out/p8/full_fl $B --workload fl > gpurun_out/p8/ncu_f_fl.log 2>&1
ls -la gpurun_out/p8
