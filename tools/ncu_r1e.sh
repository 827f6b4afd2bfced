set -x
timeout 300 python -m pytest tests -m gpu -x -q -k "stream" > gpurun_out/stream_test.log 2>&1; tail -5 gpurun_out/stream_test.log
ls
export MPC_B200_STREAM=1
B="python bench.py --steps 3 --warmup 3 --no-extra --fast-math 1 --workload mppi_acrobot_h100_k1m"
$B > gpurun_out/plain_stream.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:mppi_solve -s 2 -c 1 -f -o gpurun_out/r1e_stream_fast $B > gpurun_out/ncu_f_stream.log 2>&1
ncu -i gpurun_out/r1e_stream_fast.ncu-rep --page raw --csv > gpurun_out/r1e_raw_stream_fast.csv 2>/dev/null
ncu -i gpurun_out/r1e_stream_fast.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/r1e_src_stream_fast.csv.gz; rm -f gpurun_out/r1e_stream_fast.ncu-rep
export MPC_B200_STREAM=0
$B > gpurun_out/plain_whole.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:mppi_solve -s 2 -c 1 -f -o gpurun_out/r1e_whole_fast $B > gpurun_out/ncu_f_whole.log 2>&1
ncu -i gpurun_out/r1e_whole_fast.ncu-rep --page raw --csv > gpurun_out/r1e_raw_whole_fast.csv 2>/dev/null
ncu -i gpurun_out/r1e_whole_fast.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/r1e_src_whole_fast.csv.gz; rm -f gpurun_out/r1e_whole_fast.ncu-rep
