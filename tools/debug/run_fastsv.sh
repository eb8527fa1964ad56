N=${N:-2}; S=${S:-22}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
for wl in fastsv cc; do
  timeout -k 5 280 $T bench.py --gpus $N --workload $wl --scale $S --steps 3 --warmup 3 > gpurun_out/r02x_${wl}_s${S}_n${N}.json 2> gpurun_out/r02x_${wl}_s${S}_n${N}.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/r02x_${wl}_s${S}_n${N}.json") if l.startswith("{")][-1])
    print("$wl N=$N scale $S:", round(d["value"],1), "GTEPS", round(d["ms_per_step"],2), "ms/run", d["config"]["iterations"], "iters", "checksum", d["config"]["distance_checksum"], "reached", d["config"]["reached"])
except Exception as e:
    print("$wl failed", e); print(open("gpurun_out/r02x_${wl}_s${S}_n${N}.err").read()[-1500:])
PY
done
