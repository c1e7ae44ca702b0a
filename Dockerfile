# Manager + worker image for one 8xB200 host (counterpart of the reference Dockerfile, which builds
# a static Go binary; here the operator is Python and the kernels are built with nvcc for sm_100a).
FROM nvidia/cuda:12.9.0-devel-ubuntu24.04
RUN apt-get update && apt-get install -y --no-install-recommends python3 python3-pip git && rm -rf /var/lib/apt/lists/*
RUN pip3 install --break-system-packages torch torchvision pyyaml prometheus_client jsonschema
WORKDIR /opt/kubedl-b200
COPY . .
RUN python3 -m kubedl_b200.ops.build
ENTRYPOINT ["python3", "-m", "kubedl_b200.cmd.manager"]
