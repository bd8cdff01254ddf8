set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r5_smoke.log 2>&1; echo smoke rc=$?
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r5_tests.log 2>&1; echo tests rc=$?; tail -3 gpurun_out/r5_tests.log
timeout 600 python bench.py > gpurun_out/r5_bench.json 2> gpurun_out/r5_bench.err; echo bench rc=$?
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r5_bench_ref.json 2>> gpurun_out/r5_bench.err; echo ref rc=$?
timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu --no-extra > gpurun_out/r5_plain.json 2> gpurun_out/r5_plain.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r5_launches.csv python bench.py --steps 4 --warmup 3 --no-cpu --no-extra > gpurun_out/r5_ncu1.log 2>&1; echo ncu1 rc=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tile_x2_kernel -s 4 -c 1 -o gpurun_out/r5_x2 -f python bench.py --steps 4 --warmup 3 --no-cpu --no-extra > gpurun_out/r5_ncu2.log 2>&1; echo ncu2 rc=$?
head -c 1500 gpurun_out/r5_bench.json
