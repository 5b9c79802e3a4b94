mkdir -p gpurun_out
CMD="python bench.py --workload zed_stereo_4000kp --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-knn-sweep"
$CMD > gpurun_out/plain_zed.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,launch__grid_size,launch__block_size --clock-control none -c 600 --csv --log-file gpurun_out/launches_zed.csv $CMD > gpurun_out/ncu_launch_zed.log 2>&1
echo "rc $?"
