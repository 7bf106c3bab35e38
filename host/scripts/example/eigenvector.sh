#!/bin/sh
# CUDA counterpart of the reference's scripts/example/eigenvector.sh: every spmv kernel JSON of the example
# directory on matrix3, one trial each (the reference loops over $projroot/example/*.json under oclgrind).
#   usage: eigenvector.sh <reference checkout>      (SANITIZER=memcheck|racecheck|initcheck replaces oclgrind)
projroot=$1
here=$(cd "$(dirname "$0")" && pwd)
harness=$here/../../bin/eigenvector_harness
matrix=$projroot/example/matrix3.mtx
runfile=$projroot/example/shortrunfile.csv
hname=${HOSTNAME:-$(hostname)}
exid=example_experiment
prefix=""
if [ -n "$SANITIZER" ]; then prefix="compute-sanitizer --tool $SANITIZER"; fi

for f in $(ls $projroot/example/ | grep json); do
	echo "Running $harness with kernel $f"
	$prefix $harness -e $exid -n $hname -m $matrix -f matrix3 -k $projroot/example/$f -d 0 -r $runfile -i 1 -t 1
done
