#!/usr/bin/env bash
# Race / memory checking of the kernels on a GPU box.  ONE tool per call:
#   scripts/sanitize.sh memcheck|racecheck|synccheck|initcheck [pytest -k expression]
# (Some shared GPU pools refuse compute-sanitizer outright; the script then prints the pool's message and exits.)
set -euo pipefail
tool="${1:-memcheck}"; expr="${2:-bn or softmax or sgd or pools}"
exec compute-sanitizer --tool "$tool" --error-exitcode 1 python -m pytest tests -m gpu -x -q -k "$expr"
