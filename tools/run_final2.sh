O=gpurun_out/final2; mkdir -p $O
timeout 1200 python -m pytest tests -x -q -m gpu > $O/pytest_1gpu.log 2>&1; echo "rc=$?" >> $O/pytest_1gpu.log; tail -3 $O/pytest_1gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > $O/bench_all.json 2> $O/bench_all.err; echo "bench rc=$?"; cut -c1-160 $O/bench_all.json
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$B --workload rl4 > $O/plain_rl4.json 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:'rl_encode|rl_decode' -s 8 -c 4 -o $O/full_rl4 $B --workload rl4 > $O/ncu_f_rl4.log 2>&1; tail -1 $O/ncu_f_rl4.log
python tools/quick_rl.py 1073741824 1 2 4 64 4096 > $O/quick_rl.log 2>&1; cat $O/quick_rl.log
