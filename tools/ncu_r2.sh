# round-2 profiling recipe (run under gpurun, one GPU):  bash tools/ncu_r2.sh TAG "workload ..." [extra bench args]
# For every workload: plain run first (must exit 0), launch list, one `ncu --set full` capture of the solve kernel; the
# raw page goes to gpurun_out/<TAG>_raw_<workload>.csv, the report itself is kept (source page is read offline).
TAG=$1; shift
WLS=$1; shift
for w in $WLS; do
  B="python bench.py --steps 3 --warmup 3 --no-extra --workload $w $*"
  $B > gpurun_out/${TAG}_plain_$w.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/${TAG}_launches_$w.csv $B > gpurun_out/${TAG}_ncu_l_$w.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:mppi_solve -s 2 -c 1 -f -o gpurun_out/${TAG}_$w $B > gpurun_out/${TAG}_ncu_f_$w.log 2>&1
  ncu -i gpurun_out/${TAG}_$w.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw_$w.csv 2>/dev/null
  rm -f gpurun_out/${TAG}_$w.ncu-rep   # the reports are ~20 MB each; gpurun_out travels back only below 64 MiB
done
# the single-launch CEM solve (BASELINE configs[1]): launch list of one probe run + a full capture of cem_solve_kernel
python tools/cem_probe.py 30 4096 > gpurun_out/${TAG}_plain_cem.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/${TAG}_launches_cem_cartpole_k4096.csv python tools/cem_probe.py 4 4096 > gpurun_out/${TAG}_ncu_l_cem.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:cem_solve -s 30 -c 1 -f -o gpurun_out/${TAG}_cem python tools/cem_probe.py 30 4096 > gpurun_out/${TAG}_ncu_f_cem.log 2>&1
ncu -i gpurun_out/${TAG}_cem.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw_cem_cartpole_k4096.csv 2>/dev/null
rm -f gpurun_out/${TAG}_cem.ncu-rep
