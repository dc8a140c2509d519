cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -n 3 gpurun_out/pytest_gpu.log
python tools/time_bow.py
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_voc.py tests/test_gpu_track.py -m gpu -q -x -k "golden or argument or resident or append or 2000" > gpurun_out/sanitizer_memcheck.log 2>&1; echo "memcheck exit $?" | tee -a gpurun_out/sanitizer_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_voc.py -m gpu -q -x -k "golden or bow_batch" > gpurun_out/sanitizer_racecheck.log 2>&1; echo "racecheck exit $?" | tee -a gpurun_out/sanitizer_racecheck.log
tail -n 6 gpurun_out/sanitizer_memcheck.log gpurun_out/sanitizer_racecheck.log
