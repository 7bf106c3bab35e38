#!/bin/sh
# CUDA counterpart of the reference's scripts/example/pr.sh: same inputs and flags, pr_harness from this repo.
#   usage: pr.sh <reference checkout>      (SANITIZER=memcheck|racecheck|initcheck replaces oclgrind)
projroot=$1
here=$(cd "$(dirname "$0")" && pwd)
harness=$here/../../bin/pr_harness
kernel=$projroot/example/pr/kernel5.json
matrix=$projroot/example/matrix.mtx
runfile=$projroot/example/shortrunfile.csv
hname=${HOSTNAME:-$(hostname)}
exid=example_experiment
prefix=""
if [ -n "$SANITIZER" ]; then prefix="compute-sanitizer --tool $SANITIZER"; fi

$prefix $harness -e $exid -n $hname -m $matrix -f matrix -k $kernel -d 0 -r $runfile -i 10 -t 2000 
