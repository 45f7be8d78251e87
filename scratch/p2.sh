cd /root/repo
python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3 > gpurun_out/r1j_pytest.log; cat gpurun_out/r1j_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1j_bench_ref.json 2>/dev/null; cut -c1-300 gpurun_out/r1j_bench_ref.json
python bench.py > gpurun_out/r1j_bench.json 2> gpurun_out/r1j_bench.err; tail -n 3 gpurun_out/r1j_bench.err; cut -c1-200 gpurun_out/r1j_bench.json
lscpu | head -n 20 > gpurun_out/r1j_lscpu.txt
