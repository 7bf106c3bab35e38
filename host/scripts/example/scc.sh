#!/bin/sh
# CUDA counterpart of the reference's scripts/example/scc.sh: same inputs and flags, scc_harness from this repo.
#   usage: scc.sh <reference checkout>      (SANITIZER=memcheck|racecheck|initcheck replaces oclgrind)
projroot=$1
here=$(cd "$(dirname "$0")" && pwd)
harness=$here/../../bin/scc_harness
kernel=$projroot/example/scc/kernel6.json
matrix=$projroot/example/matrix3.mtx
runfile=$projroot/example/shortrunfile.csv
hname=${HOSTNAME:-$(hostname)}
exid=example_experiment
prefix=""
if [ -n "$SANITIZER" ]; then prefix="compute-sanitizer --tool $SANITIZER"; fi

$prefix $harness -e $exid -n $hname -m $matrix -f matrix3 -k $kernel -d 0 -r $runfile -i 10 -t 2000 
