cd $GRAFT_REPO_ROOT
env | grep -i nccl
for m in stderr plain; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scratch/nccl_probe.py $m > gpurun_out/probe_$m.out 2> gpurun_out/probe_$m.err
echo "mode $m: stdout NCCL lines $(grep -c 'NCCL INFO' gpurun_out/probe_$m.out), stderr NCCL lines $(grep -c 'NCCL INFO' gpurun_out/probe_$m.err)"; grep RESULT gpurun_out/probe_$m.out
done
grep -h "nranks\|Init COMPLETE" gpurun_out/probe_plain.out | head -3
python -m pytest tests -x -q -m gpu -k "multi_gpu or exchange" 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-edges --no-cpu > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_n2.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
