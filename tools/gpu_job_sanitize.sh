set -x
TAG=${TAG:-r02_sanitize}
O=gpurun_out/$TAG; mkdir -p $O
timeout 300 python tools/sanitize_case.py > $O/plain.log 2>&1 &&
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 7 python tools/sanitize_case.py > $O/memcheck.log 2>&1
echo "memcheck exit $?" | tee -a $O/memcheck.log
tail -25 $O/memcheck.log
