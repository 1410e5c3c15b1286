python -m pytest tests -q -m gpu -x 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r2c.json 2> gpurun_out/bench_r2c.err
tail -c 300 gpurun_out/bench_r2c.err
python profiles/exp/benchsum.py gpurun_out/bench_r2c.json
python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | tail -1 | cut -c1-400
