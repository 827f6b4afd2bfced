set -x
B="python bench.py --steps 3 --warmup 3 --no-extra"
for w in mppi_pendulum_h50_k1m mppi_acrobot_h100_k1m; do
  $B --workload $w > gpurun_out/plain_$w.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:mppi_solve -s 2 -c 1 -f -o gpurun_out/r1c_$w $B --workload $w > gpurun_out/ncu_f_$w.log 2>&1
done
