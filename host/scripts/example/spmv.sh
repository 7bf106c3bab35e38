#!/bin/sh
# CUDA counterpart of the reference's scripts/example/spmv.sh: same inputs and flags, spmv_harness from this repo.
#   usage: spmv.sh <reference checkout>      (SANITIZER=memcheck|racecheck|initcheck replaces oclgrind)
projroot=$1
here=$(cd "$(dirname "$0")" && pwd)
harness=$here/../../bin/spmv_harness
kernel=$projroot/example/kernel7.json
matrix=$projroot/example/matrix.mtx
runfile=$projroot/example/runfile.csv
hname=${HOSTNAME:-$(hostname)}
exid=example_experiment
prefix=""
if [ -n "$SANITIZER" ]; then prefix="compute-sanitizer --tool $SANITIZER"; fi

$prefix $harness -e $exid -n $hname -m $matrix -f matrix3 -k $kernel -d 0 -r $runfile -i 10 -t 2000 
