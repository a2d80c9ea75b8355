#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${1:-t}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/${1:-t}_pytest.log
