mkdir -p gpurun_out
timeout -k 5 900 python -u -m pytest tests/test_gpu_train.py -q -m gpu -s > gpurun_out/train_tests.log 2>&1; echo "train tests exit=$?"
grep -v "^$" gpurun_out/train_tests.log | tail -n 150
