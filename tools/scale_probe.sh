# N-GPU bench lines (run under gpurun --gpus 8): the driver's own launch line, per workload
for n in ${SCALE_NS:-8 4}; do
for wl in mppi_pendulum_h50_k1m mppi_cartpole_h50_k1m mppi_acrobot_h100_k1m; do
  CUDA_VISIBLE_DEVICES=$(seq -s, 0 $((n-1))) python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 20 --warmup 5 --no-extra --workload $wl 2>gpurun_out/scale_n${n}_$wl.err | tail -1 > gpurun_out/scale_n${n}_$wl.json
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/scale_n${n}_$wl.json")); print("$wl", d["n_gpus"], "%.3e"%d["value"], "ms %.4f"%d["ms_per_step"], "e2e %.3e"%d["e2e"]["value"], d["e2e"]["p50_solve_ms"], d.get("solve_phases_us"))
except Exception as e:
    print("$wl n=$n failed", e)
PY
done; done
