cd /root/repo
python -m pytest tests -m gpu -x -q 2>&1 | tail -n 8 > gpurun_out/r1h_pytest.log; cat gpurun_out/r1h_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r1h_bench.json 2> gpurun_out/r1h_bench.err; tail -c 2500 gpurun_out/r1h_bench.json; tail -n 5 gpurun_out/r1h_bench.err
