#!/bin/bash
bash tools/ncu_one.sh freq freq_split3d_mma
tail -3 gpurun_out/ncu_freq.log
