set -x
python probes/ncu_kernels_r02.py > gpurun_out/plain_kernels.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"tail_fwd_kernel|tail_bwd_kernel|mha_tc_kernel|mha_bwd_tc_kernel|gemm_bf16_kernel|nvjet|cutlass|gemm" -o gpurun_out/kernels_r02 python probes/ncu_kernels_r02.py > gpurun_out/ncu_kernels.log 2>&1
echo ncu_kernels rc=$?
python bench.py --steps 2 --warmup 1 --only-headline --no-cpu-baseline > gpurun_out/plain_infer.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_r02_infer.csv python bench.py --steps 2 --warmup 1 --only-headline --no-cpu-baseline > gpurun_out/ncu_infer.log 2>&1
echo launches_infer rc=$?
python bench.py --workload train --graph 0 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain_train.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_r02_train.csv python bench.py --workload train --graph 0 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_train.log 2>&1
echo launches_train rc=$?
python bench.py --steps 2 --warmup 1 --only-headline --no-cpu-baseline > gpurun_out/plain_infer2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:pixel_fused_kernel -s 3 -c 1 -o gpurun_out/pixel_fused_r02 python bench.py --steps 2 --warmup 1 --only-headline --no-cpu-baseline > gpurun_out/ncu_pixel.log 2>&1
echo pixel rc=$?
ls -la gpurun_out/*.ncu-rep
