#!/usr/bin/env bash
# compute-sanitizer passes over the small GPU parity cases (the reference has no race detection; this
# build is expected to run memcheck / racecheck / synccheck).  Run on a GPU box from the repository root:
#
#     bash tools/sanitize.sh [memcheck|racecheck|synccheck|initcheck] [pytest -k expression]
#
# Writes gpurun_out/sanitize_<tool>.log and exits non-zero when the tool reports an error.  The tiny sequences keep a pass to a
# few minutes (the sanitizer slows kernels 10-100x); the full-size C1 / C2 cases are deselected.  The ordered scans spin on
# look-back words that other CTAs publish with relaxed device-scope stores (csrc/scan.cuh): racecheck only looks at shared
# memory, so that protocol is outside what it can judge; memcheck and synccheck cover every kernel.
set -u
TOOL="${1:-memcheck}"
EXPR="${2:-tiny or ragged or swap or normal_and_weights or acosf or expf or collisions or checkpoint or round_trip or fixture or degenerate}"
mkdir -p gpurun_out
LOG="gpurun_out/sanitize_${TOOL}.log"
compute-sanitizer --tool "${TOOL}" --error-exitcode 86 --target-processes all \
    python -m pytest tests -m gpu -x -q -k "(${EXPR}) and not full_size and not c1_ and not c2_ and not dist" > "${LOG}" 2>&1
RC=$?
tail -n 15 "${LOG}"
grep -c "ERROR SUMMARY: 0 errors" "${LOG}" >/dev/null && echo "sanitize(${TOOL}): clean" || echo "sanitize(${TOOL}): see ${LOG}"
exit ${RC}
