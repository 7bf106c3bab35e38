#!/bin/bash
# One compute-sanitizer tool per call (TOOL=memcheck|racecheck|initcheck|synccheck, default memcheck) over the
# harness executables on the small fixtures -- the CUDA counterpart of the reference running every example under
# `oclgrind --check-api --data-races` (scripts/example/*.sh:19-21, scripts/experiments/all_kernel_test.sh:29-88) --
# and over a short Python driver that exercises the blocked schedule, the column panels, BFS, SSSP and the group
# drivers on small operators.  Plain run first; summaries land in gpurun_out/sanitize_<tool>.log.
TOOL=${TOOL:-memcheck}
mkdir -p gpurun_out /tmp/shb_mtx
K=tests/golden/kernels
for f in tests/golden/mtx/*.mtx.gz; do gunzip -c $f > /tmp/shb_mtx/$(basename $f .gz); done
LOG=gpurun_out/sanitize_${TOOL}.log
: > $LOG
run() {  # name, command...
  local name=$1; shift
  echo "=== $name: $*" >> $LOG
  SHB_QUIET_LOG=1 timeout 900 compute-sanitizer --tool $TOOL --error-exitcode 99 --print-limit 20 "$@" > gpurun_out/san_tmp.log 2>&1
  local rc=$?
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error:|Hazard|Invalid|race" gpurun_out/san_tmp.log | head -12 >> $LOG
  echo "rc=$rc" >> $LOG
}
B=host/bin
COMMON="-e san -n box -d 0 -r $K/runfile.csv -i 2 -t 2000"
# plain run of one of them first: a faulting program must not go under the tool
SHB_QUIET_LOG=1 $B/spmv_harness $COMMON -m /tmp/shb_mtx/1138_bus.mtx -f m -k $K/spmv/awrg-alcl-fdp-chunk-rsa-8.json > /dev/null 2>&1 || { echo "plain run failed" >> $LOG; exit 1; }
for k in glb-sdp glb-sdp-rsa swrg-slcl-pmdp awrg-alcl-fdp-chunk-rsa-8; do
  run "spmv $k" $B/spmv_harness $COMMON -m /tmp/shb_mtx/1138_bus.mtx -f m -k $K/spmv/$k.json
done
run "spmv gpu_encode" $B/spmv_harness $COMMON -m /tmp/shb_mtx/GD96_b.mtx -f m -k $K/spmv/glb-sdp.json --gpu_encode
run "sssp faithful" $B/sssp_harness $COMMON -m /tmp/shb_mtx/Trefethen_20.mtx -f m -k $K/sssp/glb-sdp-rsa.json
run "sssp canonical" $B/sssp_harness $COMMON -m /tmp/shb_mtx/1138_bus.mtx -f m -k $K/sssp/glb-sdp.json --mode canonical
run "bfs faithful" $B/bfs_harness $COMMON -m /tmp/shb_mtx/GD96_b.mtx -f m -k $K/bfs/glb-sdp.json --max_iterations 30
run "bfs canonical" $B/bfs_harness $COMMON -m /tmp/shb_mtx/1138_bus.mtx -f m -k $K/bfs/glb-sdp.json --mode canonical
run "pr canonical" $B/pr_harness $COMMON -m /tmp/shb_mtx/1138_bus.mtx -f m -k $K/spmv/glb-sdp.json --mode canonical
run "scc" $B/scc_harness $COMMON -m /tmp/shb_mtx/Trefethen_20.mtx -f m -k $K/scc/glb-sdp-rsa.json
run "eigenvector canonical" $B/eigenvector_harness $COMMON -m /tmp/shb_mtx/Trefethen_20.mtx -f m -k $K/spmv/glb-sdp.json --mode canonical
run "synthetic rmat bfs" $B/bfs_harness $COMMON -f m -k $K/bfs/glb-sdp.json --mode canonical --synthetic rmat:12:8
run "synthetic grid sssp" $B/sssp_harness $COMMON -f m -k $K/sssp/glb-sdp.json --mode canonical --synthetic grid:96:80
run "python: blocked, panels, group" python tools/sanitize_driver.py
cat $LOG
