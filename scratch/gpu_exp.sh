#!/bin/bash
mkdir -p gpurun_out
{ echo "copy-only, 3 stages"; RDYB_DEBUG_COPYONLY=1 bash scratch/sweep.sh "fixed:256";
  echo "copy-only, 2 stages"; RDYB_DEBUG_COPYONLY=1 RDYB_FIXED_SHAPE=A bash scratch/sweep.sh "fixed:256";
  echo "copy-only, 3 stages, 1 CTA/SM"; RDYB_CTAS_PER_SM=1 RDYB_DEBUG_COPYONLY=1 bash scratch/sweep.sh "fixed:256";
  echo "full, 1 CTA/SM"; RDYB_CTAS_PER_SM=1 bash scratch/sweep.sh "fixed:256"; } 2>&1 | tee gpurun_out/exp.log
