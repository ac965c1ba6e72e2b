#!/bin/bash
# 1-GPU closing run: full GPU test suite, norm / GEMM timings, ResNet-50 both arms, headline bench both arms, ncu refresh
mkdir -p gpurun_out
timeout 700 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r2f.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu_r2f.log
for g in norms gemm_k3; do timeout 200 python tools/gpu_check.py $g 2>&1 | grep "TIME\|FAIL\|== group" > gpurun_out/check_${g}_r2f.log; cat gpurun_out/check_${g}_r2f.log | tail -8; done
timeout 200 python bench/resnet_bench.py --gpus 1 --steps 10 --warmup 3 > gpurun_out/resnet_n1.out 2> gpurun_out/resnet_n1.err; grep '^{' gpurun_out/resnet_n1.out | cut -c1-160
timeout 240 python bench/resnet_bench.py --gpus 1 --impl reference --steps 10 --warmup 3 > gpurun_out/resnet_ref_n1.out 2> gpurun_out/resnet_ref_n1.err; grep '^{' gpurun_out/resnet_ref_n1.out | cut -c1-160
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1_r2f.json 2> gpurun_out/bench_n1_r2f.err; cut -c1-200 gpurun_out/bench_n1_r2f.json
timeout 400 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref_n1_r2f.json 2> gpurun_out/bench_ref_n1_r2f.err; cut -c1-200 gpurun_out/bench_ref_n1_r2f.json
for k in attn_fwd:flash_attn_fwd gemm:gemm_tcgen05; do
  t=${k%%:*}; rx=${k##*:}
  timeout 120 python tools/ncu_targets.py $t > gpurun_out/plain_${t}_r2f.log 2>&1 || { echo "plain $t failed"; continue; }
  timeout 240 ncu --set full --clock-control none --import-source on -k regex:$rx -s 1 -c 1 -f -o gpurun_out/prof_${t}_r2f python tools/ncu_targets.py $t > gpurun_out/ncu_${t}_r2f.log 2>&1; echo "ncu $t rc=$?"
done
