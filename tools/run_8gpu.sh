# The 8-GPU evidence run of round 2 (one box, 8 x B200).  usage: bash tools/run_8gpu.sh OUTDIR [NGPU]
O=${1:-gpurun_out/r2n}; N=${2:-8}; mkdir -p $O
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
nvidia-smi -L > $O/gpus.txt; nproc >> $O/gpus.txt; free -g | head -2 >> $O/gpus.txt
nvcc -O2 -o $O/pcie_ceiling tools/pcie_ceiling.cu -lpthread 2>/dev/null && $O/pcie_ceiling > $O/pcie_${N}gpu.json; cat $O/pcie_${N}gpu.json
timeout 900 python -m pytest tests/test_host_gpu.py tests/test_workloads_gpu.py tests/test_cli_gpu.py tests/test_concurrent_gpu.py -q -m gpu -v > $O/pytest_${N}gpu.log 2>&1; echo "rc=$?" >> $O/pytest_${N}gpu.log; tail -3 $O/pytest_${N}gpu.log
for w in stack16g stream64g; do
  timeout 900 $T bench.py --gpus $N --steps 5 --warmup 3 --workload $w > $O/bench_${w}_${N}gpu.json 2> $O/bench_${w}_${N}gpu.err; echo "$w rc=$?"; cut -c1-600 $O/bench_${w}_${N}gpu.json
done
timeout 900 $T bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_all_${N}gpu.json 2> $O/bench_all_${N}gpu.err; echo "all rc=$?"; cut -c1-400 $O/bench_all_${N}gpu.json
timeout 900 python tools/bench_files.py --bytes 17179869184 --out $O/file_e2e_${N}gpu.json > $O/file_e2e_${N}gpu.log 2>&1; echo "files rc=$?"; cut -c1-300 $O/file_e2e_${N}gpu.log
ls -la $O
