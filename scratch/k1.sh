cd /root/repo
rm -f hpp-core_b200/csrc/build/kernels_knn.o
make -C hpp-core_b200/csrc EXTRA=-DHPPB200_KNN_QB=8 > /dev/null 2>&1
grep "Used" hpp-core_b200/csrc/build/kernels_knn.ptxas.log | head -8
python -m pytest tests -m gpu -x -q -k "knn" 2>&1 | tail -n 2
python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/k1.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/k1.json').read().strip().splitlines()[-1]); print({k:v for k,v in d['knn'].items() if k.startswith('q')})"
