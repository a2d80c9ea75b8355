#!/bin/bash
# slab logic on ONE GPU (in-process transport) + the whole suite
TAG=${1:-slab}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/${TAG}_pytest.log
