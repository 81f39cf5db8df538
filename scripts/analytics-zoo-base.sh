#!/bin/bash
# Common part of the launcher scripts (reference: scripts/analytics-zoo-base.sh checks SPARK_HOME / ANALYTICS_ZOO_HOME and puts the
# zoo jar, python zip and conf on the spark-submit command line).  Here: checks ANALYTICS_ZOO_HOME, puts the package on
# PYTHONPATH, builds the native libraries when they are missing, then runs the command it was given.
#   usage: analytics-zoo-base.sh <command> [args...]
. "$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)/analytics-zoo-env.sh"
if [[ -z "${ANALYTICS_ZOO_HOME:-}" || ! -d "${ANALYTICS_ZOO_HOME}/zoo" ]]; then
    echo "Cannot find the zoo package under ANALYTICS_ZOO_HOME (${ANALYTICS_ZOO_HOME:-unset})" >&2
    exit 1
fi
export PYTHONPATH="${ANALYTICS_ZOO_HOME}:${PYTHONPATH:-}"
if [[ ! -f "${ANALYTICS_ZOO_HOME}/zoo/native/libzoo_rt.so" ]]; then
    python -m zoo.native.build || exit 1
fi
exec "$@"
