cd /root/repo
python bench.py --steps 2 --warmup 3 --no-edges > gpurun_out/r1j_bench_noedges.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1j_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-edges --no-cpu > gpurun_out/r1j_ncu_bench.log 2>&1
timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/r1j_wf_plain.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1j_launches_wf.csv python tools/wf_run.py 1048576 0 > gpurun_out/r1j_ncu_wf.log 2>&1
timeout 800 ncu --set full --clock-control none --import-source on -k regex:wf --launch-skip 401 --launch-count 4 -o gpurun_out/r1j_wf -f python tools/wf_run.py 1048576 0 > gpurun_out/r1j_ncu_wf_full.log 2>&1
ls -la gpurun_out/r1j*
