#!/bin/bash
# One gpurun call: compute-sanitizer over the small end of the GPU parity suite (memcheck, racecheck, synccheck,
# initcheck): golden vectors, pile-up / ragged / unsorted / multi-batch cases of the per-chromosome entry points, and
# the sample submissions (job tables, compact planes, re-run of an unsorted job, 40 jobs, repeated chr_index).
# The sanitizer slows kernels 10-100x, so only chromosomes up to ~1 Mb are selected; each tool runs under its own
# timeout.  Summaries land in gpurun_out/sanitize_*.log.
set -u
mkdir -p gpurun_out
SEL="test_empty_chromosome or test_mapq_boundary_is_strict or test_ragged_lengths_and_tile_edges or test_int16_wrap_pileup \
or test_pileup_with_mapq_filter_and_packed_neighbours or test_multi_batch_equals_single_batch \
or test_unsorted_input_takes_global_atomic_path or test_set_slots_rotation or test_statistics_banks \
or test_reference_golden_vectors or test_async_host_mode or test_run_chromosomes_one_call \
or test_sample_record_formats or test_sample_compact_planes or test_sample_with_unsorted or test_push_reads_compact \
or test_sample_of_40"
for TOOL in ${TOOLS:-memcheck racecheck synccheck initcheck}; do
  timeout ${SAN_TIMEOUT:-600} compute-sanitizer --tool $TOOL --error-exitcode 77 --print-limit 20 \
      python -m pytest tests/test_gpu_parity.py tests/test_gpu_sample.py -m gpu -x -q -p no:cacheprovider -k "$SEL" > gpurun_out/sanitize_$TOOL.log 2>&1
  echo "$TOOL rc=$?" | tee -a gpurun_out/sanitize_$TOOL.log
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" gpurun_out/sanitize_$TOOL.log | tail -n 4
done
