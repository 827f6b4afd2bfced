export MPC_B200_STREAM=1
B="python bench.py --steps 3 --warmup 3 --no-extra --workload mppi_pendulum_h50_k1m"
$B > gpurun_out/r2r_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:mppi_solve -s 2 -c 1 -f -o gpurun_out/r2r_stream $B > gpurun_out/r2r_ncu.log 2>&1
ncu -i gpurun_out/r2r_stream.ncu-rep --page raw --csv > gpurun_out/r2r_raw_stream_pendulum.csv 2>/dev/null
ncu -i gpurun_out/r2r_stream.ncu-rep --page source --csv > gpurun_out/r2r_src_stream_pendulum.csv 2>/dev/null
rm -f gpurun_out/r2r_stream.ncu-rep
tail -2 gpurun_out/r2r_plain.log | cut -c1-300
