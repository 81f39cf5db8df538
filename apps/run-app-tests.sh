#!/bin/bash
# Converts every notebook under apps/ to a script and runs it with short settings (reference: apps/run-app-tests.sh, which does the
# same with fewer epochs).  Runs on CPU unless ZOO_FORCE_CPU=0.  Exit code = number of failed apps.
set -u
cd "$(dirname "${BASH_SOURCE[0]}")"
export ZOO_FORCE_CPU=${ZOO_FORCE_CPU:-1} OMP_NUM_THREADS=${OMP_NUM_THREADS:-2} EPOCHS=${EPOCHS:-2} STEPS=${STEPS:-40} ITERS=${ITERS:-60} BAGS=${BAGS:-1}
export PYTHONPATH="$(cd .. && pwd):${PYTHONPATH:-}"
failed=0
for nb in */*.ipynb; do
    base="${nb%.ipynb}"
    tmp="$(mktemp -d)/$(basename "$base")"
    cp "$nb" "$tmp.ipynb"
    ./ipynb2py.sh "$tmp" >/dev/null 2>&1
    start=$(date +%s)
    if (cd "$(dirname "$nb")" && timeout 900 python "$tmp.py" >"$tmp.log" 2>&1); then
        echo "PASS  $nb  ($(( $(date +%s) - start )) s)"
    else
        echo "FAIL  $nb  (log: $tmp.log)"
        failed=$((failed + 1))
    fi
done
exit $failed
