#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest.log
echo "shape T (3 stages)"; bash scratch/sweep.sh "fixed:256" 2>&1 | tee gpurun_out/sweep.log
echo "shape A (2 stages)"; RDYB_FIXED_SHAPE=A bash scratch/sweep.sh "fixed:256" 2>&1 | tee -a gpurun_out/sweep.log
