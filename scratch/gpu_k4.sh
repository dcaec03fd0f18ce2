#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest.log
bash scratch/sweep.sh "${SWEEP:-fixed:256 warp:256}" 2>&1 | tee gpurun_out/sweep.log
