# round-1d evidence: GPU parity suite, default bench line, launch list, one full capture of the solve kernel
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r1d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1d_pytest.log
python __graft_entry__.py smoke > gpurun_out/r1d_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r1d_smoke.log
python bench.py > gpurun_out/r1d_bench.json 2> gpurun_out/r1d_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1d_bench_ref.json 2> gpurun_out/r1d_bench_ref.err
B="python bench.py --steps 3 --warmup 3 --no-extra"
for w in mppi_acrobot_h100_k1m mppi_pendulum_h50_k1m mppi_cartpole_h50_k1m; do
  $B --workload $w > gpurun_out/plain_$w.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r1d_launches_$w.csv $B --workload $w > gpurun_out/ncu_l_$w.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:mppi_solve -s 2 -c 1 -f -o gpurun_out/r1d_$w $B --workload $w > gpurun_out/ncu_f_$w.log 2>&1
  ncu -i gpurun_out/r1d_$w.ncu-rep --page raw --csv > gpurun_out/r1d_raw_$w.csv 2>/dev/null
done
ls -la gpurun_out
