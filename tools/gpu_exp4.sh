TAG=${1:-x4}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers.txt > /dev/null 2> gpurun_out/${TAG}_layers.err
TS2D_KSPLIT=3 TS2D_KSPLIT_BN=64 timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_s3.txt > /dev/null 2> gpurun_out/${TAG}_layers_s3.err
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
TS2D_KSPLIT=3 TS2D_KSPLIT_BN=64 timeout 900 python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/${TAG}_bench_s3.json 2> gpurun_out/${TAG}_bench_s3.err
tail -3 gpurun_out/${TAG}_pytest.log
for f in layers layers_s3; do head -1 gpurun_out/${TAG}_$f.txt | cut -c1-100; tail -1 gpurun_out/${TAG}_$f.txt; done
cut -c1-200 gpurun_out/${TAG}_bench.json; cut -c1-200 gpurun_out/${TAG}_bench_s3.json
