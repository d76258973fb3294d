#!/bin/bash
# Comparison-table runs on N GPUs of one node: tools/multi_run.sh <ngpus> <what...>
# what: tests ours ref ours_syncbn ref_syncbn ours_r152 ref_r152 ours_r152_100 ref_r152_100 ours_r18 ref_r18 ours_n1 comm
N=${1:-8}; shift
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
b() { tag=$1; shift; timeout 300 $TR bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/tab_n${N}_$tag.json 2> gpurun_out/tab_n${N}_$tag.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/tab_n${N}_$tag.json").read().strip().splitlines()[-1])
    print("$tag n=%d %.0f img/s %.3f ms e2e=%s exposed=%s overhead=%s ident=%s" % (d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"] and round(d["e2e"]["ms_per_step"],3), d.get("exposed_comm_ms_per_step"), d.get("comm_overhead_bwd_ms"), d.get("replicas_identical")))
except Exception as e: print("$tag FAILED", e); import subprocess; print(subprocess.run("tail -5 gpurun_out/tab_n${N}_$tag.err", shell=True, capture_output=True, text=True).stdout)
PY
}
for v in "$@"; do
case $v in
 tests) timeout 600 python -m pytest tests/test_multigpu.py -m gpu -q -k "$N-300 or $N-False or $N-True or eager and $N" 2>&1 | tail -8 | tee gpurun_out/pytest_n${N}.log;;
 ours) b ours;;
 ref) b ref --impl reference;;
 ours_syncbn) b ours_syncbn --sync-bn --no-e2e;;
 ref_syncbn) b ref_syncbn --impl reference --sync-bn --no-e2e;;
 ours_r152) b ours_r152_cap25 --model resnet152 --no-e2e;;
 ref_r152) b ref_r152_cap25 --impl reference --model resnet152 --no-e2e;;
 ours_r152_100) b ours_r152_cap100 --model resnet152 --bucket-cap-mb 100 --no-e2e;;
 ref_r152_100) b ref_r152_cap100 --impl reference --model resnet152 --bucket-cap-mb 100 --no-e2e;;
 ours_r18) b ours_r18 --model resnet18 --image-size 32 --no-e2e;;
 ref_r18) b ref_r18 --impl reference --model resnet18 --image-size 32 --no-e2e;;
 ours_n1) timeout 200 python bench.py --steps 20 --warmup 5 --no-e2e > gpurun_out/tab_n1_same_box.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/tab_n1_same_box.json').read().strip().splitlines()[-1]); print('ours n=1 same box %.0f img/s %.3f ms' % (d['value'], d['ms_per_step']))";;
 comm) timeout 400 $TR tools/comm_bench.py > gpurun_out/comm_bench_w${N}.log 2> gpurun_out/comm_bench_w${N}.err; tail -3 gpurun_out/comm_bench_w${N}.log | cut -c1-400;;
esac
done
