mkdir -p gpurun_out
python -m pytest -q -m gpu --timeout 300 -p no:cacheprovider tests/test_gpu_tta.py tests/test_gpu_eval.py -x 2>&1 | tail -40 > gpurun_out/new_tests.log
cat gpurun_out/new_tests.log
