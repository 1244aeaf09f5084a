# round-2 profiling pass (run under gpurun, one GPU): bench, sustained run, ncu captures of cfg5 / cfg1 / cfg3.
# The .ncu-rep files (tens of MB each with --import-source) are summarised ON THE BOX and deleted: gpurun only
# merges gpurun_out/ back when it stays under 64 MiB.
set -x
T=${1:-r02_s5}
if [ "${2:-}" = "tests" ]; then (python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/${T}_pytest.log; cat gpurun_out/${T}_pytest.log; fi
python bench.py --steps 60 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; tail -3 gpurun_out/${T}_bench.err
# >= 5 s per timed leg with the nvidia-smi clock trace (20 ms period)
python bench.py --steps 2000 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/${T}_sustained.json 2>> gpurun_out/${T}_bench.err
for cfg in cfg5 cfg1 cfg3; do
  rep=gpurun_out/${T}_${cfg}_full
  python scripts/profile_step.py $cfg > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on --profile-from-start off -o $rep -f python scripts/profile_step.py $cfg > gpurun_out/${T}_ncu_${cfg}.log 2>&1
  tail -1 gpurun_out/${T}_ncu_${cfg}.log
  python scripts/ncu_summary.py $rep.ncu-rep > gpurun_out/${T}_${cfg}_full_summary.txt 2>&1
  python scripts/ncu_phases.py $rep.ncu-rep k_fast_persist > gpurun_out/${T}_${cfg}_fast_phases.txt 2>&1
  python scripts/ncu_phases.py $rep.ncu-rep k_describe > gpurun_out/${T}_${cfg}_describe_phases.txt 2>&1
  python scripts/ncu_phases.py $rep.ncu-rep k_stereo_match > gpurun_out/${T}_${cfg}_stereo_phases.txt 2>&1
  ls -la $rep.ncu-rep
  rm -f $rep.ncu-rep
done
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/${T}_ncu_launch.log 2>&1
du -sh gpurun_out; ls -la gpurun_out | tail -25
