mkdir -p gpurun_out
timeout -k 5 900 python tests/tools/library_bar.py > gpurun_out/library_bar.json 2> gpurun_out/library_bar.err; echo "library bar exit=$?"; cat gpurun_out/library_bar.json; tail -n 5 gpurun_out/library_bar.err
