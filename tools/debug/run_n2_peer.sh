N=${N:-2}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; timeout -k 5 170 $T bench.py --gpus $N --steps 5 --warmup 3 $( [ -n "$E2E" ] || echo --no-e2e ) --breakdown --no-legs "$@" > gpurun_out/r02v_n${N}_$name.json 2> gpurun_out/r02v_n${N}_$name.err; python - <<PY
import json
try:
    s=[l for l in open("gpurun_out/r02v_n${N}_$name.json") if l.startswith("{")][-1]
    d=json.loads(s)
    print("N=$N $name", round(d["value"],1), "GTEPS", round(d["ms_per_step"]/20*1000,1), "us/iter  red launch", round(d["roofline"]["avg_launch_ms"]*1000,1), "us x", d["roofline"]["launches_per_product"], " share", round(d["roofline"]["kernel_share_of_step"],2), d.get("parity_vs_single_gpu"))
except Exception as e: print("$name failed", e)
PY
grep "timing\]" gpurun_out/r02v_n${N}_$name.err | awk '{print "    ", $3, $4, $5, $6, $7, $8, $9}' | head -8
}
for v in $VARIANTS; do
  case $v in
    peer) run peer ;;
    peer_e2e) E2E=1 run peer_e2e ;;
    peer_prog) GRB_B200_PEER_PROGRESS=1 run peer_prog --partition peer ;;
    peer_nomc) GRB_B200_NO_MULTICAST=1 run peer_nomc --partition peer ;;
    col) run col --partition col ;;
    row) run row --partition row ;;
    peer_c2s) run peer_c2s --partition peer --ag-chunks 2 --peer-ctas -148 --smem-kb 188 ;;
    peer_c4s) run peer_c4s --partition peer --ag-chunks 4 --peer-ctas -148 --smem-kb 188 ;;
    peer_c2s194) run peer_c2s194 --partition peer --ag-chunks 2 --peer-ctas -148 ;;
    fused) run fused ;;
    fused_c100) run fused_c100 --peer-carveout 100 ;;
    fused_c86s188) run fused_c86s188 --smem-kb 188 ;;
    fused_s188) run fused_s188 --smem-kb 188 ;;
    fused_s180) run fused_s180 --smem-kb 180 ;;
    fused_nomc) GRB_B200_NO_MULTICAST=1 run fused_nomc ;;
    fused_g3) GRB_B200_PEER_GENERATIONS=3 run fused_g3 ;;
    fused_e2e) E2E=1 run fused_e2e ;;
  esac
done
