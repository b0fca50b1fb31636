python -m pytest tests -x -q -m gpu 2>&1 | tail -6 > gpurun_out/r2_pytest_gpu_d.log; cat gpurun_out/r2_pytest_gpu_d.log
python bench.py > gpurun_out/r2_bench_d_fast.json 2>gpurun_out/bench_d.err
python bench.py --math exact > gpurun_out/r2_bench_d_exact.json 2>>gpurun_out/bench_d.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_d_reference.json 2>>gpurun_out/bench_d.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 3 --profile > gpurun_out/ncu_launch.log 2>&1
for m in fast exact; do
  ncu --set full --clock-control none --import-source on -k regex:volpath_q -s 3 -c 1 -f -o gpurun_out/r2_${m}_d python bench.py --steps 2 --warmup 3 --profile --math $m > gpurun_out/ncu_$m.log 2>&1
done
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/bench_d.err
