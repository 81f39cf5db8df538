#!/bin/bash
# Puts a zoo-b200 tree under $ZOO_HOME inside the image when it was not COPY-ed at build time (reference:
# docker/zoo/download-analytics-zoo.sh fetches the release zip of the chosen version from Maven).  Source, in this order:
# ZOO_TARBALL (a local tar.gz, e.g. the one `SparkRunner.pack_penv` writes), ZOO_GIT_URL (+ ZOO_GIT_REF), else fail.
set -e
dest="${ZOO_HOME:-/opt/zoo}"
mkdir -p "$dest"
if [ -n "$ZOO_TARBALL" ]; then
    tar xzf "$ZOO_TARBALL" -C "$dest"
elif [ -n "$ZOO_GIT_URL" ]; then
    git clone --depth 1 ${ZOO_GIT_REF:+--branch "$ZOO_GIT_REF"} "$ZOO_GIT_URL" "$dest"
else
    echo "set ZOO_TARBALL or ZOO_GIT_URL" >&2
    exit 1
fi
cd "$dest" && python -m zoo.native.build
