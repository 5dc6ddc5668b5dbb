# Counterpart of the reference's Dockerfile (ubuntu + build-essential + perf): an image able to build and run
# this library.  Not built inside the development sandbox (no network there); kept for users with one.
# The reference image compiles Linux `perf` for profiling (Dockerfile:7-10); the counterpart here is Nsight Compute,
# which ships with the CUDA devel image: tools/capture_profiles.sh, tools/ncu_summary.py, tools/ncu_lines.py.
FROM nvidia/cuda:12.9.0-devel-ubuntu24.04
RUN apt-get update && apt-get install -y --no-install-recommends python3 python3-pip python3-venv build-essential git \
    && rm -rf /var/lib/apt/lists/*
RUN python3 -m venv /opt/venv && /opt/venv/bin/pip install --no-cache-dir torch numpy pytest hypothesis
ENV PATH=/opt/venv/bin:$PATH
WORKDIR /parallel
COPY . /parallel
RUN make build
# docker run --gpus all -it <image> /bin/bash -c "make test-gpu && echo 1 | make run_sequential"
CMD ["/bin/bash"]
