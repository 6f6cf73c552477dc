#!/bin/bash
# 2 GPUs: rotation copies split over several copy-engine streams — configs[4]-shaped blocks (d=300), copy seconds with
# split 1 vs 4; the N-rank test
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
for sp in 1 4; do
  GE_DIST_CE_SPLIT=$sp timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 \
     bench.py --gpus 2 --steps 4 --warmup 3 --workload c5 > $O/gpu12_c5_2gpu_split$sp.json 2> $O/gpu12_c5_2gpu_split$sp.log
  echo "c5 split $sp rc=$?" | tee -a $O/gpu12_rc.txt
  python - <<PY
import json
d=json.load(open("$O/gpu12_c5_2gpu_split$sp.json"))
print("split $sp ms/epoch", round(d["ms_per_step"],1), "copy s/epoch", round(d["config"]["rotation_copy_seconds_per_epoch_max"],4), "frac", round(d["roofline"]["frac"],3), "e2e", d["e2e"])
PY
done
timeout 600 python -m pytest tests/test_gpu_dist_multi.py -x -q -s > $O/gpu12_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu12_rc.txt
tail -4 $O/gpu12_pytest.log
