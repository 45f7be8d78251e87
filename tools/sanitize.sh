#!/bin/bash
# compute-sanitizer passes over a small invocation of every kernel family (tools/sanitize_small.py).
# Summaries go to gpurun_out/san_<tool>.txt; copy them to profiles/ after reading.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
python tools/sanitize_small.py > gpurun_out/san_plain.txt 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/san_plain.txt; exit 1; }
for tool in memcheck racecheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 40 python tools/sanitize_small.py "$@" > gpurun_out/san_$tool.txt 2>&1
  echo "== $tool rc=$? =="; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|sanitize_small done" gpurun_out/san_$tool.txt | sort | uniq -c | head -20
done
