set -x
python bench.py > gpurun_out/s38_bench.json 2> gpurun_out/s38_bench.err
python scripts/profile_step.py 256 > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/r01_s38_full -f python scripts/profile_step.py 256 > gpurun_out/ncu_s38.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv --log-file gpurun_out/r01_s38_launches.csv python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_s38_launch.log 2>&1
python scripts/timeline.py > gpurun_out/r01_s38_timeline.txt 2>&1
tail -2 gpurun_out/ncu_s38.log; ls -la gpurun_out | tail -8
