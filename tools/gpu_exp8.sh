TAG=${1:-x8}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
timeout 200 python tools/layer_profile.py --out gpurun_out/${TAG}_layers.txt > /dev/null 2> gpurun_out/${TAG}_layers.err
timeout 600 python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
timeout 300 python tools/stage_profile.py > gpurun_out/${TAG}_stages.txt 2> gpurun_out/${TAG}_stages.err
head -1 gpurun_out/${TAG}_layers.txt | cut -c1-100; tail -1 gpurun_out/${TAG}_layers.txt; grep norm gpurun_out/${TAG}_layers.txt | head -12
cut -c1-120 gpurun_out/${TAG}_bench.json | tail -1; cat gpurun_out/${TAG}_stages.txt | tail -30
