#!/bin/bash
# The three GPU example manifests through the product's front door: kubedlctl -> operator ->
# node agent -> one worker process per GPU.  Usage: scripts/run_examples_gpu.sh [NGPU] (default 2).
# Output goes to gpurun_out/examples_<NGPU>gpu.log when run under gpurun.  EXTRA="--warm-gpus" starts every pod
# as an adoption of a worker that already holds a CUDA context on its GPU.
N=${1:-2}
EXTRA=${EXTRA:-}
OUT=${2:-gpurun_out/examples_${N}gpu.log}
mkdir -p "$(dirname "$OUT")"
{
  echo "=== PyTorchJob ResNet-50 DDP, $N pods"
  timeout 300 python -m kubedl_b200.cmd.kubedlctl run -f examples/pytorchjob_resnet50_ddp_8gpu.yaml \
      --gpus "$N" --replicas "$N" --gang gpu-gang --env KUBEDL_DDP_STEPS=30 --logs 4 --timeout 280 $EXTRA
  echo "rc=$?"
  echo "=== XGBoostJob hist GBDT, $N pods"
  timeout 300 python -m kubedl_b200.cmd.kubedlctl run -f examples/xgboostjob_higgs_8gpu.yaml \
      --gpus "$N" --replicas "$N" --logs 4 --timeout 280 $EXTRA
  echo "rc=$?"
  echo "=== XDLJob DIN, 1 scheduler + $N workers"
  timeout 300 python -m kubedl_b200.cmd.kubedlctl run -f examples/xdljob_din_8gpu.yaml \
      --gpus "$N" --replicas $((N + 1)) --env KUBEDL_XDL_WORKERS="$N" --logs 4 --timeout 280 $EXTRA
  echo "rc=$?"
} > "$OUT" 2>&1
tail -60 "$OUT"
