mkdir -p gpurun_out
timeout -k 5 900 python -u -m pytest tests/test_gpu_train.py -q -m gpu -s > gpurun_out/train_tests.log 2>&1; echo "train tests exit=$?"
grep -v "^$" gpurun_out/train_tests.log | tail -n 60
timeout -k 5 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?"; tail -n 5 gpurun_out/smoke.log
timeout -k 5 400 python tools/bench_train.py --clips 4 --steps 3 --profile > gpurun_out/bench_train_prof.json 2> gpurun_out/bench_train_prof.err; echo "bench train prof exit=$?"; cat gpurun_out/bench_train_prof.json; tail -n 5 gpurun_out/bench_train_prof.err
timeout -k 5 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"; tail -n 5 gpurun_out/bench.err; cat gpurun_out/bench.json
