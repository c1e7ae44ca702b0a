#!/bin/bash
# Time from "node agent starts the pod" to "first training step done" for the ResNet-50 PyTorchJob example,
# three ways of starting a pod: exec of a fresh interpreter, fork of the launcher that has torch imported,
# adoption of a worker that already holds a CUDA context on the pod's GPU.
# Usage: scripts/bench_first_step.sh [NGPU] [OUT]
N=${1:-2}
OUT=${2:-gpurun_out/first_step_${N}gpu.log}
mkdir -p "$(dirname "$OUT")"
run() {
  echo "=== $1"
  shift
  timeout 300 python -m kubedl_b200.cmd.kubedlctl run -f examples/pytorchjob_resnet50_ddp_8gpu.yaml \
      --gpus "$N" --replicas "$N" --gang gpu-gang --env KUBEDL_DDP_STEPS=20 --logs 2 --timeout 280 "$@" 2>&1 \
    | grep -v "NCCL version" | cut -c1-700
  echo "rc=${PIPESTATUS[0]}"
}
{
  run "exec (fresh interpreter per pod)"
  run "fork from the launcher (--prefork)" --prefork
  run "adopt a warm GPU worker (--warm-gpus)" --warm-gpus
  run "adopt a warm GPU worker, second time" --warm-gpus
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
