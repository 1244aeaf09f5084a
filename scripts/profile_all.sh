set -x
python bench.py > gpurun_out/s41_bench.json 2> gpurun_out/s41_bench.err
python scripts/profile_step.py 256 > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/r01_s41_full -f python scripts/profile_step.py 256 > gpurun_out/ncu_s41.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv --log-file gpurun_out/r01_s41_launches.csv python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_s41_launch.log 2>&1
python scripts/timeline.py > gpurun_out/r01_s41_timeline.txt 2>&1
tail -2 gpurun_out/ncu_s41.log; ls -la gpurun_out | tail -8
