# round-1h evidence (final state of round 1): GPU suite, bench lines of both arms, launch lists + full captures of the
# flavours bench.py quotes, CEM launch list
set -x
python -m pytest tests -m gpu -q > gpurun_out/r1h_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1h_pytest.log; tail -3 gpurun_out/r1h_pytest.log
python __graft_entry__.py smoke > gpurun_out/r1h_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r1h_smoke.log; tail -2 gpurun_out/r1h_smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1h_bench_ref.json 2> gpurun_out/r1h_bench_ref.err
python bench.py > gpurun_out/r1h_bench.json 2> gpurun_out/r1h_bench.err
B="python bench.py --steps 3 --warmup 3 --no-extra"
for w in mppi_acrobot_h100_k1m mppi_pendulum_h50_k1m mppi_cartpole_h50_k1m; do
  $B --workload $w > gpurun_out/plain_$w.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r1h_launches_$w.csv $B --workload $w > gpurun_out/ncu_l_$w.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:mppi_solve -s 2 -c 1 -f -o gpurun_out/r1h_$w $B --workload $w > gpurun_out/ncu_f_$w.log 2>&1
  ncu -i gpurun_out/r1h_$w.ncu-rep --page raw --csv > gpurun_out/r1h_raw_$w.csv 2>/dev/null
  rm -f gpurun_out/r1h_$w.ncu-rep
done
python tools/cem_probe.py 200 > gpurun_out/r1h_cem_probe.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1h_launches_cem_cartpole_k4096.csv python tools/cem_probe.py 2 > /dev/null 2>&1
python tools/trace_probe.py > gpurun_out/r1h_trace_probe.log 2>&1
python tools/latency_probe.py > gpurun_out/r1h_latency_probe.log 2>&1
