#!/bin/bash
# N-GPU scaling experiments (scratch helper): usage tools/n2_exp.sh <ngpus> [variants...]
N=${1:-2}; shift
mkdir -p gpurun_out
run() { tag=$1; envs=$2; args=$3
  env $envs timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5 --no-e2e --no-comm-probe $args > gpurun_out/n${N}_$tag.json 2> gpurun_out/n${N}_$tag.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/n${N}_$tag.json").read().strip().splitlines()[-1]); print("$tag", round(d["ms_per_step"],3), d["exposed_comm_ms_per_step"], d["replicas_identical"], d.get("step_anatomy_ms"))
except Exception as e: print("$tag FAILED", e)
PY
  grep "bucket trace" gpurun_out/n${N}_$tag.err | cut -c1-1800; }
for v in "$@"; do
case $v in
 default) run default "B200DDP_TRACE=0" "";;
 trace) run trace "B200DDP_TRACE=1" "";;
 tail1) run tail1 "B200DDP_TAIL_BUCKET_MB=1" "";;
 blocks16) run blocks16 "B200DDP_COMM_BLOCKS=16" "";;
 blocks32) run blocks32 "B200DDP_COMM_BLOCKS=32" "";;
 presync) run presync "B200DDP_BUFFER_SYNC=pre" "";;
 syncbn) run syncbn "B200DDP_TRACE=0" "--sync-bn";;
 nofuse) run nofuse "B200DDP_TRACE=0" "--no-fuse-optimizer";;
 nograph) run nograph "B200DDP_TRACE=0" "--no-graph";;
 wg0) run wg0 "B200DDP_WRITE_GRAD=0" "";;
 oneproc) N1=1; env B200DDP_TRACE=0 timeout 200 python bench.py --steps 30 --warmup 5 --no-e2e > gpurun_out/n1_same_box.json 2> gpurun_out/n1_same_box.err; python -c "
import json; d=json.loads(open('gpurun_out/n1_same_box.json').read().strip().splitlines()[-1]); print('n1_same_box', round(d['ms_per_step'],3), d.get('step_anatomy_ms'))";;
esac
done
