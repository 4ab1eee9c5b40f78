# quick GPU iteration: tests, per-layer table, stage profile (+ model-stream sweep), bench line.   usage: bash tools/gpu_quick.sh <tag>
TAG=${1:-q}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
python tools/layer_profile.py --out gpurun_out/${TAG}_layers.txt > /dev/null 2> gpurun_out/${TAG}_layers.err
python tools/stage_profile.py > gpurun_out/${TAG}_stages.txt 2> gpurun_out/${TAG}_stages.err
python bench.py --steps 3 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
tail -3 gpurun_out/${TAG}_pytest.log; cat gpurun_out/${TAG}_stages.txt; tail -3 gpurun_out/${TAG}_stages.err; tail -1 gpurun_out/${TAG}_layers.txt; cut -c1-300 gpurun_out/${TAG}_bench.json | tail -1
