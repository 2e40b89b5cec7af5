# Round-3 evidence run (one gpurun call): GPU suite, bench line, launch lists, ncu --set full of the kernels that actually run.
set -x
python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu_r03.txt 2>&1; echo pytest rc=$?
python bench.py > gpurun_out/bench_r03b_n1.json 2> gpurun_out/bench_r03b_n1.err; echo bench rc=$?
python bench.py --only-headline --no-cpu-baseline > gpurun_out/bench_r03b_n1_again.json 2>/dev/null; echo bench2 rc=$?
P="--steps 2 --warmup 1 --only-headline --no-cpu-baseline --min-timed-ms 0"
python bench.py $P > gpurun_out/plain_infer.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_r03_infer.csv python bench.py $P > gpurun_out/ncu_infer.log 2>&1
echo launches_infer rc=$?
T="--workload train --graph 0 --steps 2 --warmup 1 --no-cpu-baseline"
python bench.py $T > gpurun_out/plain_train.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_r03_train.csv python bench.py $T > gpurun_out/ncu_train.log 2>&1
echo launches_train rc=$?
ncu --set full --clock-control none --import-source on -k regex:pixel_fused_t_kernel -s 3 -c 1 -o gpurun_out/pixel_fused_t_r03 python bench.py $P > gpurun_out/ncu_pixel.log 2>&1
echo pixel rc=$?
ncu --set full --clock-control none --import-source on -k regex:"conv3x3_wgrad_kernel|mha_bwd_tc_kernel" -s 10 -c 2 -o gpurun_out/train_kernels_r03 python bench.py $T > gpurun_out/ncu_trk.log 2>&1
echo train_kernels rc=$?
ncu -i gpurun_out/pixel_fused_t_r03.ncu-rep --page raw --csv > gpurun_out/pixel_fused_t_r03_ncu_raw.csv
ncu -i gpurun_out/train_kernels_r03.ncu-rep --page raw --csv > gpurun_out/train_kernels_r03_ncu_raw.csv
ls -la gpurun_out/
