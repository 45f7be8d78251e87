set -x
python -m pytest tests -x -q -m gpu 2>&1 | tail -8
python tools/wf_run.py 1048576 20000 2 2>&1 | tail -6
for a in 0 2 3 4; do HPPB200_KNN_TC_ABLATE=$a python scratch/r2_knn_ablate.py 2>&1 | tail -1; done
