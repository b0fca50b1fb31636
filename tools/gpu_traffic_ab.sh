# DRAM traffic of the queue kernel with and without the persisting L2 window on the cold path records (debug_flags bit 1)
cfg=${1:-c2}
for dbg in 0 2; do
  ncu --metrics dram__bytes_write.sum,dram__bytes_read.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum -k regex:volpath_q -s 1 -c 1 --csv --log-file gpurun_out/traffic_${cfg}_$dbg.csv python tools/bench_configs.py $cfg math=fast debug=$dbg > /dev/null 2>&1
  echo "$cfg debug=$dbg"; grep -E "dram__|lts__|gpu__time" gpurun_out/traffic_${cfg}_$dbg.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}'
done
