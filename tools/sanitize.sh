#!/usr/bin/env bash
# compute-sanitizer passes over small GPU parity cases that between them launch every kernel family of libdsb200
# (the reference has no race detection; this build runs memcheck / racecheck / synccheck / initcheck).
# Run on a GPU box from the repository root:
#
#     bash tools/sanitize.sh [memcheck|racecheck|synccheck|initcheck] [pytest -k expression]
#
# Writes gpurun_out/sanitize_<tool>.log; the verdict is the sanitizer's exit code (--error-exitcode) AND the absence of any
# non-zero "ERROR SUMMARY" line in any of the processes it followed.  The tiny sequences keep a pass to a few minutes (the
# sanitizer slows kernels 10-100x).  The ordered scans spin on look-back words that other CTAs publish with relaxed
# device-scope stores (csrc/scan.cuh): racecheck only looks at shared memory, so that protocol is outside what it can judge;
# memcheck and synccheck cover every kernel.
set -u
TOOL="${1:-memcheck}"
DEFAULT="tiny_room_fused or tiny_street_depth or use_swapping_enlarged or prefetched_allocation or abandoned_lookahead or decay_and_slide_window_sequence or deintegrate_then or free_view_render or forward_render or view_from_raw or ragged_sizes or swapping_engine_vs_oracle or fixture_track_camera or normal_and_weights or mesh"
EXPR="${2:-$DEFAULT}"
mkdir -p gpurun_out
LOG="gpurun_out/sanitize_${TOOL}.log"
compute-sanitizer --tool "${TOOL}" --error-exitcode 86 --target-processes all \
    python -m pytest tests -m gpu -x -q -k "(${EXPR}) and not full_size and not c1_ and not c2_ and not c3_ and not dist and not refcuda and not reference_cuda" > "${LOG}" 2>&1
RC=$?
tail -n 15 "${LOG}"
BAD=$(grep "ERROR SUMMARY" "${LOG}" | grep -vc "ERROR SUMMARY: 0 errors")
if [ "${RC}" -eq 0 ] && [ "${BAD}" -eq 0 ] && grep -q "ERROR SUMMARY: 0 errors" "${LOG}"; then echo "sanitize(${TOOL}): clean"; else echo "sanitize(${TOOL}): rc=${RC}, ${BAD} non-zero error summaries -- see ${LOG}"; [ "${RC}" -eq 0 ] && RC=1; fi
exit ${RC}
