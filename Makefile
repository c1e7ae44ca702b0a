# Counterpart of the reference Makefile (manager / test / manifests / docker-build targets).
PY ?= python

all: lib manifests

# Build the sm_100a kernel library in-tree (kubedl_b200/_lib/libkubedl_b200.so)
lib:
	$(PY) -m kubedl_b200.ops.build

# Run the operator against the in-process API server
run:
	$(PY) -m kubedl_b200.cmd.manager --enable-leader-election=false

# CPU tests (control plane, host logic) — what CI runs; GPU tests need a B200
test: lib manifests
	KUBEDL_CI=true $(PY) -m pytest tests -x -q -m "not gpu"

test-gpu: lib
	$(PY) -m pytest tests -x -q -m gpu

# Regenerate CRD manifests from the kind registry
manifests:
	$(PY) -m kubedl_b200.api.crd --out config/crd/bases

bench:
	$(PY) bench.py --steps 20 --warmup 3

launch-delay:
	$(PY) -m kubedl_b200.bench.launch_delay --jobs 30 --replicas 8

sass:
	cuobjdump -sass kubedl_b200/_lib/libkubedl_b200.so | grep -E "UTCHMMA|UTMALDG|UTMASTG|LDTM|MULTIMEM" | sort | uniq -c

IMG ?= kubedl-b200/manager:latest
KUBECTL ?= kubectl

docker-build:
	docker build . -t $(IMG)

docker-push:
	docker push $(IMG)

# Against a Kubernetes cluster (reference Makefile:29-36): CRDs only, or CRDs + RBAC + operator Deployment
install: manifests
	$(KUBECTL) apply -f config/crd/v1

deploy: manifests
	$(KUBECTL) apply -f config/manager/all_in_one.yaml

# Operator on this machine against the cluster of the current kubeconfig (reference `make run`)
run-cluster:
	$(PY) -m kubedl_b200.cmd.manager --in-cluster --enable-leader-election=false

# Static checks the reference runs before every build (go fmt / go vet): byte-compile everything
vet:
	$(PY) -m compileall -q kubedl_b200 tests bench.py __graft_entry__.py

.PHONY: all lib run test test-gpu manifests bench launch-delay sass docker-build docker-push install deploy run-cluster vet
