TAG=${1:-x6}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_unet.py -m gpu -x -q > gpurun_out/${TAG}_pytest_unet.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_unet.log
tail -3 gpurun_out/${TAG}_pytest_unet.log
timeout 200 python tools/layer_profile.py --out gpurun_out/${TAG}_layers.txt > /dev/null 2> gpurun_out/${TAG}_layers.err
TS2D_MC=0 timeout 200 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_nomc.txt > /dev/null 2> gpurun_out/${TAG}_layers_nomc.err
TS2D_KSPLIT=3 timeout 200 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_s3.txt > /dev/null 2> gpurun_out/${TAG}_layers_s3.err
TS2D_KSPLIT=9 timeout 200 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_s9.txt > /dev/null 2> gpurun_out/${TAG}_layers_s9.err
TS2D_KSPLIT=9 TS2D_KSPLIT_BN=128 timeout 200 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_s9b128.txt > /dev/null 2> gpurun_out/${TAG}_layers_s9b128.err
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
for f in layers layers_nomc layers_s3 layers_s9 layers_s9b128; do echo "== $f"; head -1 gpurun_out/${TAG}_$f.txt | cut -c1-100; tail -1 gpurun_out/${TAG}_$f.txt; tail -2 gpurun_out/${TAG}_$f.err; done
cut -c1-200 gpurun_out/${TAG}_bench.json
