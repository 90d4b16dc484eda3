# Round-2 evidence on one B200: tests, smoke, the default bench line + reference arm, launch list + ncu --set full per workload
# (each only after the plain run of the same command exited 0).  usage: bash tools/run_final_1gpu.sh OUTDIR
O=${1:-gpurun_out/final}; mkdir -p $O
timeout 1200 python -m pytest tests -q -m gpu > $O/pytest_1gpu.log 2>&1; echo "rc=$?" >> $O/pytest_1gpu.log; tail -3 $O/pytest_1gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > $O/bench_all.json 2> $O/bench_all.err; echo "bench rc=$?"; cut -c1-200 $O/bench_all.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference_arm.json 2>/dev/null; cut -c1-120 $O/bench_reference_arm.json
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$B > $O/plain_all.json 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_all.csv $B > /dev/null 2>&1
for w in fl rl4 rl64 rl4096; do
  $B --workload $w > $O/plain_$w.json 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:'fl_encode|fl_decode|fl_tile|rl_encode|rl_decode' -s 8 -c 4 -o $O/full_$w $B --workload $w > $O/ncu_f_$w.log 2>&1; tail -1 $O/ncu_f_$w.log
done
python tools/quick_rl.py 1073741824 1 2 4 64 4096 > $O/quick_rl.log 2>&1; cat $O/quick_rl.log
python tools/quick_fl.py > $O/quick_fl.log 2>&1; cat $O/quick_fl.log
ls -la $O
