#!/bin/bash
# build the native library; non-zero exit + message on failure
cd "$(dirname "$0")/.." && python -m modularevogym_b200._build > /tmp/mevo_build.log 2>&1
rc=$?
grep -E "error|warning: v|registers|spill" /tmp/mevo_build.log | grep -B1 -A0 "wstep\|error" | head -${1:-20}
if [ $rc -ne 0 ]; then echo "BUILD FAILED"; tail -20 /tmp/mevo_build.log; exit 1; fi
echo "build ok"
