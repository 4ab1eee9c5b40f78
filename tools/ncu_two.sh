# ncu --set full of the first-layer conv and the blend kernel inside one bench step (after a plain run of the same command exits 0)
TAG=${1:-n}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e --studies-per-step 1"
TS2D_MODEL_STREAMS=1 $CMD > gpurun_out/${TAG}_plain.log 2>&1 || exit 1
TS2D_MODEL_STREAMS=1 ncu --set full --clock-control none --import-source on -k regex:"conv_first_tc|blend_head" -s 10 -c 2 -o gpurun_out/${TAG}_two $CMD > gpurun_out/${TAG}_ncu.log 2>&1
ncu -i gpurun_out/${TAG}_two.ncu-rep --page raw --csv > gpurun_out/${TAG}_two_raw.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_two.ncu-rep --page details > gpurun_out/${TAG}_two_details.txt 2>/dev/null
ncu -i gpurun_out/${TAG}_two.ncu-rep --page source --csv > gpurun_out/${TAG}_two_source.csv 2>/dev/null
ls -la gpurun_out | tail -8
