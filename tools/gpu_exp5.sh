TAG=${1:-x5}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers.txt > /dev/null 2> gpurun_out/${TAG}_layers.err
TS2D_PDL=0 timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_nopdl.txt > /dev/null 2> gpurun_out/${TAG}_layers_nopdl.err
timeout 900 python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
TS2D_PDL=0 timeout 900 python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/${TAG}_bench_nopdl.json 2> gpurun_out/${TAG}_bench_nopdl.err
tail -3 gpurun_out/${TAG}_pytest.log
for f in layers layers_nopdl; do head -1 gpurun_out/${TAG}_$f.txt | cut -c1-100; tail -1 gpurun_out/${TAG}_$f.txt; done
cut -c1-200 gpurun_out/${TAG}_bench.json; cut -c1-200 gpurun_out/${TAG}_bench_nopdl.json
