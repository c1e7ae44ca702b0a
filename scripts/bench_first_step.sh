#!/bin/bash
# Time from "node agent starts the pod" to "first training step done" for the ResNet-50 PyTorchJob example,
# three ways of starting a pod: exec of a fresh interpreter, fork of the launcher that has torch imported,
# adoption of a worker that already holds a CUDA context on the pod's GPU.
# Usage: [ARMS="exec prefork warm"] scripts/bench_first_step.sh [NGPU] [OUT]
N=${1:-2}
OUT=${2:-gpurun_out/first_step_${N}gpu.log}
mkdir -p "$(dirname "$OUT")"
run() {
  echo "=== $1"
  shift
  rm -rf /tmp/kubedl-b200/pods          # pod logs are appended to: start every arm with empty ones
  timeout 300 python -m kubedl_b200.cmd.kubedlctl run -f examples/pytorchjob_resnet50_ddp_8gpu.yaml \
      --gpus "$N" --replicas "$N" --gang gpu-gang --env KUBEDL_DDP_STEPS=20 --logs 1 --timeout 280 "$@" 2>&1 \
    | grep -v "NCCL version" | cut -c1-700
  echo "rc=${PIPESTATUS[0]}"
}
ARMS=${ARMS:-exec prefork warm}
{
  for a in $ARMS; do
    case $a in
      exec) run "exec (fresh interpreter per pod)" ;;
      prefork) run "fork from the launcher (--prefork)" --prefork ;;
      warm) run "adopt a warm GPU worker (--warm-gpus)" --warm-gpus ;;
    esac
  done
} > "$OUT" 2>&1
python - "$OUT" <<'PY'
import json, re, sys
arm = None
for line in open(sys.argv[1]):
    if line.startswith("==="):
        arm = line[4:].strip(); vals = []
    m = re.search(r"\{.*\}", line)
    if m and "pod_start_to_first_step_s" in line:
        d = json.loads(m.group(0))
        print(f"{arm:45s} rank {d['rank']}: pod start -> first step {d['pod_start_to_first_step_s']:.2f} s "
              f"(in-worker {d['time_to_first_step_s']:.2f} s) phases {d['startup_s']}  {d['samples_per_s_per_gpu']:.0f} img/s/gpu")
    if line.startswith(("warm GPU workers", "rc=", "launch-delay")):
        print("   ", line.strip())
PY
