#!/bin/sh
# CUDA counterpart of the reference's scripts/example/bfs.sh: same inputs and flags, bfs_harness from this repo.
#   usage: bfs.sh <reference checkout>      (SANITIZER=memcheck|racecheck|initcheck replaces oclgrind)
projroot=$1
here=$(cd "$(dirname "$0")" && pwd)
harness=$here/../../bin/bfs_harness
kernel=$projroot/example/bfs/kernel3.json
matrix=$projroot/example/matrix.mtx
runfile=$projroot/example/shortrunfile.csv
hname=${HOSTNAME:-$(hostname)}
exid=example_experiment
prefix=""
if [ -n "$SANITIZER" ]; then prefix="compute-sanitizer --tool $SANITIZER"; fi

$prefix $harness -e $exid -n $hname -m $matrix -f matrix -k $kernel -d 0 -r $runfile -i 10 -t 2000 --max_iterations ${MAX_ITERATIONS:-200}
