#!/bin/bash
# Sweep (matrix x kernel JSON) with a harness executable of this repo -- the counterpart of the reference's
# scripts/experiments/run_all.sh:1-108, same positional interface and the same result layout, so that
# gather_all.sh / analyse.sh-style post-processing (grep PROFILING_DATUM / INSERT INTO) keeps working:
#   run_all.sh <dataset folder> <harness executable> <kernel folder> <runfile> <platform> <device> <scratch folder>
# <dataset folder>/datasets.txt lists matrix names; matrix M lives at <dataset folder>/M/M.mtx.
# Results: results/results-<exID>/<matrix>/result_<kernel>.txt (+ runstatus.txt).  Environment:
#   TRIALS (5)  TIMEOUT_MS (20, the reference's -t)  SANITIZER=memcheck|racecheck|initcheck (replaces the
#   reference's oclgrind sweep, all_kernel_test.sh)  COMPRESS=1 (tar.gz each result like the reference)
set -u
if [ $# -lt 7 ]; then sed -n 2,10p "$0"; exit 64; fi
datasetf=$1; harness=$2; kernelfolder=$3; runfile=$4; platform=$5; device=$6; scratchfolder=$7
status=runstatus.txt
: > $status
log() { echo "$*"; echo "$*" >> $status; }
for v in datasetf harness kernelfolder runfile platform device scratchfolder; do log "$v: ${!v}"; done
host=$(hostname)
now=$(date +"%Y-%m-%dT%H-%M-%S%z")
hsh=$(git rev-parse HEAD 2>/dev/null || echo nogit)
exID="$hsh-$now"
prefix=""
if [ -n "${SANITIZER:-}" ]; then prefix="compute-sanitizer --tool $SANITIZER"; fi
mkdir -p "results/results-$exID" "$scratchfolder/results/results-$exID"
kernelcount=$(ls "$kernelfolder"/*.json | wc -l)
matrixcount=$(grep -c . "$datasetf/datasets.txt")
taskcount=$((kernelcount * matrixcount))
log "taskcount: $taskcount"
start=$(date +%s); i=0; failed=0
for m in $(cat "$datasetf/datasets.txt"); do
  rdir="results/results-$exID/$m"; mkdir -p "$rdir"
  for k in "$kernelfolder"/*.json; do
    kname=$(basename "$k" .json)
    i=$((i + 1))
    log "Processing matrix: $m - $i/$taskcount, kernel: $kname"
    runstart=$(date +%s)
    $prefix "$harness" -p "$platform" -d "$device" -i "${TRIALS:-5}" -m "$datasetf/$m/$m.mtx" -f "$m" -k "$k" -r "$runfile" \
        -n "$host" -t "${TIMEOUT_MS:-20}" -e "$exID" > "$rdir/result_$kname.txt" 2>&1
    rc=$?
    if [ $rc -ne 0 ]; then log "run failed! (exit $rc)"; failed=$((failed + 1)); fi
    if grep -q "Buffers are identical - kernel has probably failed" "$rdir/result_$kname.txt"; then log "kernel left its output untouched: $kname on $m"; fi
    if [ -n "${COMPRESS:-}" ]; then (cd "$rdir" && tar czf "result_$kname.tar.gz" "result_$kname.txt" && rm -f "result_$kname.txt"); fi
    runend=$(date +%s)
    log "Run took $((runend - runstart)) seconds, total time of $((runend - start)) seconds"
  done
done
log "finished experiments: $i runs, $failed failed, results in results/results-$exID"
echo "results/results-$exID" > .last_results
[ $failed -eq 0 ]
