#!/bin/bash
# Reference name (scripts/spark-submit-python-with-zoo.sh: spark-submit with the zoo jar, conf and python zip).  There is no
# Spark here: spark-submit style options that have a meaning are mapped, the rest are dropped, and the script runs with one
# rank per GPU through zoo-submit.   usage: spark-submit-python-with-zoo.sh [--master local[N]] [--num-executors N] script.py [args]
here="$(cd "$(dirname "$0")" && pwd)"
gpus=""
while [ $# -gt 0 ]; do
  case "$1" in
    --master) case "$2" in local\[*\]) n="${2#local[}"; n="${n%]}"; [ "$n" != "*" ] && gpus="$n";; esac; shift 2;;
    --num-executors) gpus="$2"; shift 2;;
    --driver-memory|--executor-memory|--executor-cores|--driver-cores|--total-executor-cores|--conf|--py-files|--jars|--properties-file|--deploy-mode|--name|--class) shift 2;;
    *) break;;
  esac
done
avail=$(nvidia-smi -L 2>/dev/null | wc -l)
if [ -n "$gpus" ] && [ "$avail" -gt 0 ] && [ "$gpus" -gt "$avail" ]; then gpus="$avail"; fi
if [ -n "$gpus" ] && [ "$avail" -gt 0 ]; then exec "$here/zoo-submit" --gpus "$gpus" "$@"; fi
exec "$here/zoo-submit" "$@"
