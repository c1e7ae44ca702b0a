#!/usr/bin/env bash
# Nightly image build + push (counterpart of the reference scripts/docker_build_daily.sh): only on the
# scheduled run of the default branch, credentials from the CI secrets DOCKER_USERNAME / DOCKER_PASSWORD.
set -euo pipefail
echo "event=${GITHUB_EVENT_NAME:-manual} ref=${GITHUB_REF_NAME:-unknown}"
if [[ "${GITHUB_EVENT_NAME:-}" == "schedule" && "${GITHUB_REF_NAME:-}" == "main" ]]; then
  echo "${DOCKER_PASSWORD:?}" | docker login -u "${DOCKER_USERNAME:?}" --password-stdin
  IMG="kubedl-b200/manager:nightly"
  docker build . -t "$IMG"
  docker push "$IMG"
else
  echo "not a scheduled run of main: nothing to build"
fi
