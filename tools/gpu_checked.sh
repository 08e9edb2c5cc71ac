#!/bin/bash
# The parity suite against the bounds-checked build (device asserts on every shared-memory / row index the simulator
# kernels form; `make -C madrona-bots_b200/csrc checked`). compute-sanitizer is closed on this GPU pool: this is the
# memcheck stand-in. A tripped assert fails the next CUDA call, i.e. the test.
TAG=${1:-r2}
mkdir -p gpurun_out
/usr/local/cuda/bin/compute-sanitizer --version > gpurun_out/${TAG}_sanitizer_refusal.txt 2>&1
/usr/local/cuda/bin/compute-sanitizer --tool memcheck python -c "print(1)" >> gpurun_out/${TAG}_sanitizer_refusal.txt 2>&1
MBOTS_LIB=$PWD/madrona-bots_b200/lib/variants/libmbots_checked.so timeout 1200 python -m pytest tests/test_parity_gpu.py tests/test_checkpoint_gpu.py -m gpu -x -q > gpurun_out/${TAG}_checked_tests.log 2>&1
tail -3 gpurun_out/${TAG}_checked_tests.log
for V in 0 1 2 3; do
  MBOTS_EXPORT_SPLIT_LOG2=$V MBOTS_LIB=$PWD/madrona-bots_b200/lib/variants/libmbots_checked.so timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "small or ragged or limit or capacity or large or graph" > gpurun_out/${TAG}_checked_tests_split$V.log 2>&1; tail -1 gpurun_out/${TAG}_checked_tests_split$V.log
done
