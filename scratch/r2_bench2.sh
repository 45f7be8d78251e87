cd $GRAFT_REPO_ROOT
( time python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err ) 2> gpurun_out/r2_bench.time
tail -5 gpurun_out/r2_bench.err; cat gpurun_out/r2_bench.time
