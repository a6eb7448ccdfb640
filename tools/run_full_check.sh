# whole GPU suite, smoke, then the default bench line
python -m pytest tests -x -q -m gpu > gpurun_out/t_full.txt 2>&1
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.txt 2>&1
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
