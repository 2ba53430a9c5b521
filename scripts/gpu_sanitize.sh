#!/bin/bash
# One gpurun call: compute-sanitizer over the small end of the GPU parity suite (memcheck, then racecheck and
# synccheck on the pile-up / ragged / pipelined cases that exercise the shared-memory bins and the last-CTA
# epilogue).  The sanitizer slows kernels 10-100x, so only tests with chromosomes of a few hundred kilobases are
# selected; each tool runs under its own timeout.  Summaries land in gpurun_out/sanitize_*.log.
# NOT YET RUN in round 1 (GPU budget spent) — first call of round 2.
set -u
mkdir -p gpurun_out
SEL="test_empty_chromosome or test_mapq_boundary_is_strict or test_ragged_lengths_and_tile_edges or test_int16_wrap_pileup \
or test_pileup_with_mapq_filter_and_packed_neighbours or test_multi_batch_equals_single_batch \
or test_unsorted_input_takes_global_atomic_path or test_set_slots_rotation or test_statistics_banks \
or test_reference_golden_vectors"
for TOOL in memcheck racecheck synccheck; do
  timeout ${SAN_TIMEOUT:-900} compute-sanitizer --tool $TOOL --error-exitcode 77 --print-limit 20 \
      python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "$SEL" > gpurun_out/sanitize_$TOOL.log 2>&1
  echo "$TOOL rc=$?" | tee -a gpurun_out/sanitize_$TOOL.log
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" gpurun_out/sanitize_$TOOL.log | tail -n 4
done
