#!/bin/bash
# usage: bench_env.sh VAR "v1 v2 ..." "args1" "args2" ... -- bench_args.sh for every value of an environment variable
VAR=$1; VALS=$2; shift 2
for v in $VALS; do echo "$VAR=$v"; env $VAR=$v bash scripts/bench_args.sh "$@"; done
