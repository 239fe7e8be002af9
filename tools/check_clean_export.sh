#!/bin/bash
# Does a clean checkout work?  Exports HEAD (tracked files only -- anything git-ignored is absent, which is how a missing
# source package would show up), builds the library there and runs the CPU test-suite.
set -e
D=$(mktemp -d /tmp/dpb_export.XXXXXX)
git -C "$(dirname "$0")/.." archive HEAD | tar -x -C "$D"
cd "$D"
python -c "import __graft_entry__ as g; g.build()" | tail -1
python -m pytest tests -x -q -m "not gpu" | tail -2
cd / && rm -rf "$D"
