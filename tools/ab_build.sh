#!/bin/bash
# development helper: ab/libA.so = HEAD, ab/libB.so = working tree, for a same-box comparison (tools/ab_run.sh)
set -e
cd "$(dirname "$0")/.."
mkdir -p ab
python -m claragenomicsanalysis_b200.build --force >/dev/null
cp claragenomicsanalysis_b200/lib/libcudapoa_b200.so ab/libB.so
rm -rf /tmp/ab_head && git worktree add -f /tmp/ab_head HEAD >/dev/null 2>&1
(cd /tmp/ab_head && python -m claragenomicsanalysis_b200.build --force >/dev/null && cp claragenomicsanalysis_b200/lib/libcudapoa_b200.so "$OLDPWD/ab/libA.so")
git worktree remove --force /tmp/ab_head
