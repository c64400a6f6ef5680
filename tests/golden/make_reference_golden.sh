#!/bin/bash
# Generates tests/golden/ref/<case>_<n>.bin: outputs of the UNMODIFIED reference (the
# Unified-Memory build oracle/_ref/parity_driver_ref, compiled by oracle/Makefile from
# /root/reference) for small instances of the header-API driver cases.  Needs a GPU, so it
# runs on the B200 box:  gpurun -- 'bash tests/golden/make_reference_golden.sh gpurun_out/ref_golden'
# and the files are then copied into tests/golden/ref/.  tests/test_oracle.py pins the CPU
# oracle against them (no GPU needed); tests/test_header_api.py compares the CUDA path.
set -u
out=${1:-tests/golden/ref}
mkdir -p "$out"
ref=oracle/_ref/parity_driver_ref
for spec in scan_i32:4099 scan_i64:4096 rld_f32:4099 rld_u32max:5001 axpy_f32:4099 axpy_f32_r3:4096 \
            gather_f32:4099 fill_iota:4096 algos:1003 errors:1000; do
  c=${spec%%:*}; n=${spec##*:}
  CUDA_VISIBLE_DEVICES=0 $ref dump $c $n "$out/${c}_${n}.bin" &
done
wait
ls -la "$out"
