#!/bin/bash
# The CLI parity cases with the host program built under AddressSanitizer + UndefinedBehaviorSanitizer
# (make -C conga_b200/host asan).  Any report aborts the program, which fails its test.
set -u
mkdir -p gpurun_out
export CONGA_HOST_BIN=$PWD/conga_b200/host/conga-b200-asan
export ASAN_OPTIONS=protect_shadow_gap=0:detect_leaks=0:abort_on_error=1:halt_on_error=1
export UBSAN_OPTIONS=halt_on_error=1:print_stacktrace=1
timeout ${ASAN_TIMEOUT:-900} python -m pytest tests/test_cli_parity.py -m gpu -q --timeout 600 -p no:cacheprovider \
    -k "not full_size and not gpus" > gpurun_out/asan_cli.log 2>&1
echo "asan rc=$?" | tee -a gpurun_out/asan_cli.log
tail -n 5 gpurun_out/asan_cli.log
