# Round-end evidence on one B200: full GPU suite, bench lines (both math modes + the reference arm), the ncu launch
# list and one `--set full` capture per math mode.  Outputs land in gpurun_out/ (copied into profiles/ afterwards).
python -m pytest tests -x -q -m gpu 2>&1 | tail -6 > gpurun_out/r2_pytest_gpu_final.log; cat gpurun_out/r2_pytest_gpu_final.log
python bench.py > gpurun_out/r2_bench_final_fast.json 2>gpurun_out/bench_final.err
python bench.py --math exact > gpurun_out/r2_bench_final_exact.json 2>>gpurun_out/bench_final.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_final_reference.json 2>>gpurun_out/bench_final.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 3 --profile > gpurun_out/ncu_launch.log 2>&1
for m in fast exact; do
  ncu --set full --clock-control none --import-source on -k regex:volpath_q -s 3 -c 1 -f -o gpurun_out/r2_${m}_g python bench.py --steps 2 --warmup 3 --profile --math $m > gpurun_out/ncu_$m.log 2>&1
done
python -m pytest tests/test_gpu_fastmath.py tests/test_gpu_majorant.py -q -s 2>&1 | grep -E "FAST vs|majorant grid|passed|failed" > gpurun_out/r2_fast_majorant_final.log
python tools/bench_mis.py > gpurun_out/r2_mis_final.jsonl 2>/dev/null
python tools/bench_mis.py c2 c3 math=fast > gpurun_out/r2_mis_fast_final.jsonl 2>/dev/null
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/bench_final.err
