# experiment: split-K of the tiny levels, per-tile overhead in the BN model, L2 promotion of partial-pixel TMA boxes.  usage: bash tools/gpu_exp1.sh <tag>
TAG=${1:-x1}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
TS2D_KSPLIT=0 TS2D_TILE_OVH=0 timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_base.txt > /dev/null 2> gpurun_out/${TAG}_layers_base.err
timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_auto.txt > /dev/null 2> gpurun_out/${TAG}_layers_auto.err
TS2D_KSPLIT=3 timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_s3.txt > /dev/null 2> gpurun_out/${TAG}_layers_s3.err
TS2D_KSPLIT=9 timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_s9.txt > /dev/null 2> gpurun_out/${TAG}_layers_s9.err
TS2D_KSPLIT=9 TS2D_KSPLIT_BN=128 timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_s9b128.txt > /dev/null 2> gpurun_out/${TAG}_layers_s9b128.err
TS2D_KSPLIT=0 TS2D_L2PROMO_SUB=1 timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_promo1.txt > /dev/null 2> gpurun_out/${TAG}_layers_promo1.err
TS2D_KSPLIT=0 TS2D_L2PROMO_SUB=0 TS2D_L2PROMO=0 timeout 300 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_promo0.txt > /dev/null 2> gpurun_out/${TAG}_layers_promo0.err
# where do the unsplit tiny convs wait?  source-level stall samples of enc6.0 / enc6.1 / enc7.0 (conv launches 12..14 of the second forward)
TS2D_KSPLIT=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_ -s 49 -c 3 -o gpurun_out/${TAG}_tiny python tools/ncu_forward.py > gpurun_out/${TAG}_ncu_tiny.log 2>&1
ncu -i gpurun_out/${TAG}_tiny.ncu-rep --page source --csv > gpurun_out/${TAG}_tiny_source.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_tiny.ncu-rep --page details > gpurun_out/${TAG}_tiny_details.txt 2>/dev/null
rm -f gpurun_out/${TAG}_tiny.ncu-rep
tail -3 gpurun_out/${TAG}_pytest.log
for f in base auto s3 s9 s9b128 promo1 promo0; do echo "== $f"; head -1 gpurun_out/${TAG}_layers_$f.txt | cut -c1-120; tail -1 gpurun_out/${TAG}_layers_$f.txt; done
