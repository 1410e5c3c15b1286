python -m pytest tests -q -m gpu -x 2>&1 | tail -4
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r2b.json 2> gpurun_out/bench_r2b.err
tail -c 600 gpurun_out/bench_r2b.err
python profiles/exp/benchsum.py gpurun_out/bench_r2b.json
