set -x; mkdir -p gpurun_out
( time python -m pytest tests/test_cpp_host.py -m gpu -x -q --durations=15 ) > gpurun_out/cpp_tests.log 2>&1
tail -30 gpurun_out/cpp_tests.log
