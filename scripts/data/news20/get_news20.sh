#!/bin/bash
# 20 Newsgroups into <dir>/news20 (reference: scripts/data/news20/get_news20.sh downloads news20.tar.gz).  Offline: a synthetic
# corpus in the same layout (one folder per category, one text file per post).
DIR="${1:-.}/news20"
mkdir -p "$DIR"
if curl -fsS --max-time 20 -o "$DIR/news20.tar.gz" http://qwone.com/~jason/20Newsgroups/20news-18828.tar.gz 2>/dev/null && tar xzf "$DIR/news20.tar.gz" -C "$DIR"; then
    echo "downloaded 20 Newsgroups into $DIR"
else
    python "$(dirname "${BASH_SOURCE[0]}")/../gen_synthetic.py" news20 "$DIR"
fi
