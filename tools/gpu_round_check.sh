# One gpurun call: GPU tests, per-layer table, bench line, ncu full-set capture of one forward's conv launches, ncu launch list of one bench step.
# usage (from the repo root, on the GPU box): bash tools/gpu_round_check.sh <tag>
TAG=${1:-r01}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
python tools/layer_profile.py --out gpurun_out/${TAG}_layers.txt > /dev/null 2> gpurun_out/${TAG}_layers.err
python bench.py --steps 3 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python tools/ncu_forward.py > gpurun_out/plain_fwd.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:conv_ -s 37 -c 37 -o gpurun_out/${TAG}_convs python tools/ncu_forward.py > gpurun_out/ncu_fwd.log 2>&1
ncu -i gpurun_out/${TAG}_convs.ncu-rep --page raw --csv > gpurun_out/${TAG}_convs_raw.csv 2>/dev/null
python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e --studies-per-step 1 > gpurun_out/plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e --studies-per-step 1 > gpurun_out/ncu_bench.log 2>&1
du -sh gpurun_out/* | sort -h | tail -5
# gpurun copies back at most 64 MiB: drop the binary report first if the directory is too large (the raw csv stays)
if [ $(du -sm gpurun_out | cut -f1) -gt 60 ]; then rm -f gpurun_out/${TAG}_convs.ncu-rep; fi
tail -3 gpurun_out/${TAG}_pytest.log; cut -c1-400 gpurun_out/${TAG}_bench.json
