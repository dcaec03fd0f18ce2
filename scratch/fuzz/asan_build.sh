#!/bin/bash
# Host C code (layouts, host mesh, readers, driver, PetscOperator shim) rebuilt with AddressSanitizer + UBSan and linked with the
# product's CUDA object into /tmp/asan/librdyb200_asan.so; then the CPU suite and the reader fuzz run against it (RDYB_LIB).
# Round 1: CPU suite clean; the fuzz found three signed overflows on INT32_MIN ids in io_formats.c (fixed), then 0 reports.
set -e
cd "$(dirname "$0")/../.."
mkdir -p /tmp/asan
for f in rdyb_error host_mesh device_layout petsc_operator_b200 rdy_driver io_formats; do
  gcc -std=c11 -O1 -g -fsanitize=address,undefined -fno-omit-frame-pointer -fPIC -Iinclude -Irdycore_b200/csrc -I/usr/local/cuda/include \
      -c rdycore_b200/csrc/$f.c -o /tmp/asan/$f.o
done
/usr/local/cuda/bin/nvcc -Wno-deprecated-gpu-targets -shared -o /tmp/asan/librdyb200_asan.so /tmp/asan/*.o rdycore_b200/build/swe_device.cu.o rdycore_b200/build/swe_noalloc.cu.o -lm -ldl -Xlinker -lasan -Xlinker -lubsan
export ASAN_OPTIONS=detect_leaks=0:halt_on_error=1:abort_on_error=1 UBSAN_OPTIONS=halt_on_error=1:abort_on_error=1
export LD_PRELOAD=$(gcc -print-file-name=libasan.so):$(gcc -print-file-name=libubsan.so) RDYB_LIB=/tmp/asan/librdyb200_asan.so
python -m pytest tests -x -q -m "not gpu" -p no:cacheprovider
python scratch/fuzz/fuzz_readers.py ${1:-240} | grep -v "^ok"
