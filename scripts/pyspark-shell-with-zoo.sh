#!/bin/bash
# Reference name (scripts/pyspark-shell-with-zoo.sh): an interactive Python shell with the zoo context initialised.
here="$(cd "$(dirname "$0")" && pwd)"
. "$here/zoo-env.sh"
exec python -i -c "from zoo.common.nncontext import init_nncontext; sc = ctx = init_nncontext(); print('zoo-b200 shell: ctx =', ctx)"
