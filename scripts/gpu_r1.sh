set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r1_pytest.log
python bench.py --steps 3 --warmup 3 > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1_bench_ref.json 2> gpurun_out/r1_bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/r1_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"hamming|fast_nms|blur7|orient|harris|select|brief|resize" -s 30 -c 24 -o gpurun_out/r1_full -f python bench.py --frames 250 --chunk 250 --steps 1 --warmup 1 --no-cpu > gpurun_out/r1_ncu_full.log 2>&1
ls -la gpurun_out
