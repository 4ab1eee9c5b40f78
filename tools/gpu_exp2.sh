TAG=${1:-x2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_unet.py -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
for S in 0 3 9; do
TS2D_TC_DBG=1 TS2D_KSPLIT=$S timeout 300 python tools/ncu_forward.py --forwards 2 > /dev/null 2> gpurun_out/${TAG}_dbg_s$S.txt
done
timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_auto.txt > /dev/null 2> gpurun_out/${TAG}_layers_auto.err
TS2D_KSPLIT=3 timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_s3.txt > /dev/null 2> gpurun_out/${TAG}_layers_s3.err
tail -3 gpurun_out/${TAG}_pytest.log
for f in auto s3; do echo "== $f"; head -1 gpurun_out/${TAG}_layers_$f.txt | cut -c1-120; tail -1 gpurun_out/${TAG}_layers_$f.txt; done
