# frame-sharded scaling on ONE 8-GPU box: N = 2, 4 (ranks spread over the box), 4 (ranks on GPUs 0-3), 8.
# usage (under gpurun --gpus 8): bash scripts/scale_r02.sh <tag>
T=${1:-r02_scale}
P=29700
run() {  # n extra-args tag
  P=$((P+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $P bench.py --gpus $1 --steps 60 --warmup 3 $2 \
    > gpurun_out/${T}_$3.json 2> gpurun_out/${T}_$3.err
  tail -c 300 gpurun_out/${T}_$3.err
}
run 2 "" n2
run 4 "" n4
run 4 "--gpu-map identity" n4_identity
run 8 "" n8
python bench.py --steps 60 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/${T}_n1.json 2> gpurun_out/${T}_n1.err
python - <<PY
import json
for t in ("n1","n2","n4","n4_identity","n8"):
    try:
        d=json.load(open("gpurun_out/${T}_%s.json" % t))
        print(t, "value %.0f  e2e %.0f  ceiling %s  frac %s  gpu_map %s" % (d["value"], d["e2e"]["value"], d["e2e"]["copy_ceiling_frames_per_s"], d["e2e"]["frac_of_copy_ceiling"], d["config"].get("gpu_map")))
    except Exception as ex:
        print(t, "failed", ex)
PY
