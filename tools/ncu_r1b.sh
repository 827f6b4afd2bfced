set -x
B="python bench.py --steps 3 --warmup 3 --no-extra"
for w in mppi_acrobot_h100_k1m mppi_pendulum_h50_k1m mppi_cartpole_h50_k1m; do
  $B --workload $w > gpurun_out/plain_$w.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1b_launches_$w.csv $B --workload $w > gpurun_out/ncu_l_$w.log 2>&1
  $B --workload $w > gpurun_out/plain_$w.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:mppi_rollout -s 2 -c 1 -f -o gpurun_out/r1b_$w $B --workload $w > gpurun_out/ncu_f_$w.log 2>&1
done
w=mppi_pendulum_h50_k1m
$B --workload $w > gpurun_out/plain_$w.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:mppi_finalize -s 2 -c 1 -f -o gpurun_out/r1b_finalize_$w $B --workload $w > gpurun_out/ncu_ff_$w.log 2>&1
ls -la gpurun_out
