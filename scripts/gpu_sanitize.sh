#!/bin/bash
# One gpurun call: compute-sanitizer memcheck and racecheck over scripts/sanitize_paths.py (logs for profiles/r02).
O=gpurun_out
python scripts/sanitize_paths.py > $O/san_plain.log 2>&1; tail -2 $O/san_plain.log
timeout 900 compute-sanitizer --tool memcheck --log-file $O/san_memcheck.txt python scripts/sanitize_paths.py > $O/san_memcheck_run.log 2>&1
echo "memcheck rc=$?"; tail -3 $O/san_memcheck.txt
timeout 1200 compute-sanitizer --tool racecheck --log-file $O/san_racecheck.txt python scripts/sanitize_paths.py > $O/san_racecheck_run.log 2>&1
echo "racecheck rc=$?"; tail -3 $O/san_racecheck.txt
timeout 600 compute-sanitizer --tool synccheck --log-file $O/san_synccheck.txt python scripts/sanitize_paths.py > $O/san_synccheck_run.log 2>&1
echo "synccheck rc=$?"; tail -3 $O/san_synccheck.txt
