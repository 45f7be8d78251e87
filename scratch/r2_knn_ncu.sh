set -x
python scratch/r2_knn_prof.py > /dev/null 2>&1 || exit 1
ncu --set full --import-source on --clock-control none --kernel-name regex:knnTcFilter --launch-skip 5 --launch-count 1 -f -o gpurun_out/r2b_knnTcFilter python scratch/r2_knn_prof.py > gpurun_out/r2b_knnTcFilter_ncu.log 2>&1
tail -3 gpurun_out/r2b_knnTcFilter_ncu.log
ls -la gpurun_out/r2b_knnTcFilter.ncu-rep
