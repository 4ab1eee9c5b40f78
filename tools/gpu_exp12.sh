TAG=${1:-x12}
mkdir -p gpurun_out
for cfg in "3 0" "3 1" "1 1" "1 0" "2 0" "2 1"; do
set -- $cfg
TS2D_MODEL_STREAMS=$1 TS2D_PDL=$2 timeout 300 python bench.py --steps 4 --warmup 3 --skip-cpu --skip-e2e > gpurun_out/${TAG}_bench_s$1_p$2.json 2> gpurun_out/${TAG}_bench_s$1_p$2.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${TAG}_bench_s$1_p$2.json").read().strip().splitlines()[-1])
print("streams $1 pdl $2:", round(d["value"],2), "studies/s", d["clocks"]["sm_mhz"])
PY
done
