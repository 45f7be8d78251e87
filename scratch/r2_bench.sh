cd $GRAFT_REPO_ROOT
( time python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err ) 2> gpurun_out/r2_bench.time
tail -3 gpurun_out/r2_bench.err; cat gpurun_out/r2_bench.time
( time python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r2_bench_ref.json ) 2>&1 | tail -3
