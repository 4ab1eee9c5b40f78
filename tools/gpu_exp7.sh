TAG=${1:-x7}
mkdir -p gpurun_out
timeout 200 python tools/layer_profile.py --out gpurun_out/${TAG}_layers.txt > /dev/null 2> gpurun_out/${TAG}_layers.err
TS2D_KSPLIT16=3 timeout 200 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_16s3.txt > /dev/null 2> gpurun_out/${TAG}_layers_16s3.err
TS2D_KSPLIT16=9 timeout 200 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_16s9.txt > /dev/null 2> gpurun_out/${TAG}_layers_16s9.err
TS2D_KSPLIT16=3 TS2D_KSPLIT_BN=256 timeout 200 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_16s3b256.txt > /dev/null 2> gpurun_out/${TAG}_layers_16s3b256.err
TS2D_KSPLIT=0 TS2D_PDL=0 timeout 200 python tools/layer_profile.py --out gpurun_out/${TAG}_layers_old.txt > /dev/null 2> gpurun_out/${TAG}_layers_old.err
timeout 600 python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
TS2D_PDL=0 timeout 600 python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/${TAG}_bench_nopdl.json 2> gpurun_out/${TAG}_bench_nopdl.err
TS2D_KSPLIT=0 TS2D_PDL=0 timeout 600 python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/${TAG}_bench_old.json 2> gpurun_out/${TAG}_bench_old.err
for f in layers layers_16s3 layers_16s9 layers_16s3b256 layers_old; do echo "== $f"; head -1 gpurun_out/${TAG}_$f.txt | cut -c1-100; tail -1 gpurun_out/${TAG}_$f.txt; tail -2 gpurun_out/${TAG}_$f.err; done
for f in bench bench_nopdl bench_old; do cut -c1-120 gpurun_out/${TAG}_$f.json | tail -1; done
