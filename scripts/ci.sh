#!/usr/bin/env bash
# CI entry (reference .travis.yml:12-14 sets KUBEDL_CI=true and runs `make test`).
set -euo pipefail
export KUBEDL_CI=true
python -m kubedl_b200.ops.build            # cross-compiles for sm_100a without a GPU
python -m kubedl_b200.api.crd --out config/crd/bases
git diff --exit-code config/crd config/manager/all_in_one.yaml    # generated manifests must be up to date
python -m pytest tests -x -q -m "not gpu"
