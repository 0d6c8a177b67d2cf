#!/bin/bash
# Race / sync / memory checks for the hand-written kernels (the reference has no race detector; ours is
# compute-sanitizer over the single-GPU kernel tests plus the NXD_STRESS=1 flag-protocol stress test).
#
# Usage (on a GPU box, ONE GPU; each tool replays every kernel, so expect 10-50x the plain test time):
#   tools/run_sanitizers.sh [memcheck|racecheck|synccheck|initcheck|all] [pytest -k filter]
# Output: gpurun_out/sanitizer_<tool>.log; exit status != 0 if any tool reported an error.
#
# Scope notes
#  * racecheck sees SHARED-memory hazards only.  The mbarrier/TMA pipelines order smem through the async proxy, which
#    the tool does not model completely: hazards it reports between a TMA write and a generic read that are separated by
#    an mbarrier wait are false positives; anything between two generic accesses is real.
#  * The peer-memory (NVLink) protocols of fused_ag_gemm / fused_gemm_rs / fused_gemm_ar / fused_zero1 are outside what
#    any of the tools check.  Their check is tests/test_fused_comm_gpu.py::test_flag_protocol_stress (NXD_STRESS=1):
#    hundreds of back-to-back calls with changing payloads and skewed rank arrival compared with NCCL results.
#  * Multi-rank tests are deselected here (compute-sanitizer under a multi-process launch serialises the ranks and the
#    wait loops time out).
set -u
TOOL=${1:-all}
FILTER=${2:-"not tp2 and not dp2 and not ep2 and not multi"}
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
SAN=${COMPUTE_SANITIZER:-/usr/local/cuda/bin/compute-sanitizer}
if [ ! -x "$SAN" ]; then SAN=$(command -v compute-sanitizer) || { echo "compute-sanitizer not found"; exit 2; }; fi

run_tool() {
  local tool=$1; shift
  local log="$OUT/sanitizer_${tool}.log"
  echo "== $tool -> $log"
  # --target-processes all: the pytest runner forks for nothing here, but keep children covered.
  # --launch-timeout 0 / --kernel-name-exclude: nothing excluded; every kernel of libnxd_b200.so is replayed.
  NXD_SANITIZER=1 timeout "${SANITIZER_TIMEOUT:-2400}" "$SAN" --tool "$tool" --target-processes all \
      --error-exitcode 86 --print-limit 50 "$@" \
      python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "$FILTER" -p no:cacheprovider > "$log" 2>&1
  local rc=$?
  tail -n 5 "$log"
  grep -c "========= ERROR SUMMARY" "$log" >/dev/null 2>&1 && grep "ERROR SUMMARY" "$log" | tail -n 1
  return $rc
}

status=0
case "$TOOL" in
  memcheck)  run_tool memcheck --leak-check no || status=1 ;;
  racecheck) run_tool racecheck --racecheck-report all || status=1 ;;
  synccheck) run_tool synccheck || status=1 ;;
  initcheck) run_tool initcheck --track-unused-memory no || status=1 ;;
  all)
    run_tool memcheck --leak-check no || status=1
    run_tool synccheck || status=1
    run_tool racecheck --racecheck-report all || status=1
    ;;
  *) echo "unknown tool $TOOL"; exit 2 ;;
esac
exit $status
