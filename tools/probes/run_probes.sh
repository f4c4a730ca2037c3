# builds and runs the two micro-architecture probes on the GPU box; output -> gpurun_out/{mma,tmem}_probe.txt
mkdir -p gpurun_out
cd tools/probes
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../hivae_b200/csrc -o /tmp/mma_probe mma_probe.cu && timeout 120 /tmp/mma_probe > ../../gpurun_out/mma_probe.txt 2>&1; echo "mma probe exit=$?"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tmem_probe tmem_probe.cu && timeout 120 /tmp/tmem_probe > ../../gpurun_out/tmem_probe.txt 2>&1; echo "tmem probe exit=$?"
cd ../..
