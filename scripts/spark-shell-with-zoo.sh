#!/bin/bash
# Reference names (spark-shell-with-zoo.sh / spark-submit-scala-with-zoo.sh) start a Scala REPL / job on the JVM build.  zoo-b200 has
# no JVM side: the Scala API of the reference corresponds to the Python package here.
echo "zoo-b200 has no Scala/JVM runtime; use scripts/pyspark-shell-with-zoo.sh (interactive) or scripts/zoo-submit (jobs)." >&2
exit 2
