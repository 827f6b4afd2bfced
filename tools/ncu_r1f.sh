# round-1f evidence: GPU parity suite, default bench line (per-env flavour), launch list + full capture of the quoted flavours
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r1f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1f_pytest.log; tail -4 gpurun_out/r1f_pytest.log
python bench.py > gpurun_out/r1f_bench.json 2> gpurun_out/r1f_bench.err; cat gpurun_out/r1f_bench.json
B="python bench.py --steps 3 --warmup 3 --no-extra"
for w in mppi_acrobot_h100_k1m mppi_pendulum_h50_k1m; do
  $B --workload $w > gpurun_out/plain_$w.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r1f_launches_$w.csv $B --workload $w > gpurun_out/ncu_l_$w.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:mppi_solve -s 2 -c 1 -f -o gpurun_out/r1f_$w $B --workload $w > gpurun_out/ncu_f_$w.log 2>&1
  ncu -i gpurun_out/r1f_$w.ncu-rep --page raw --csv > gpurun_out/r1f_raw_$w.csv 2>/dev/null
  ncu -i gpurun_out/r1f_$w.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/r1f_src_$w.csv.gz
  rm -f gpurun_out/r1f_$w.ncu-rep
done
