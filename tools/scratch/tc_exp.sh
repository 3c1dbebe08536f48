for v in 0 1 2 4 8 6 14; do
  CPPRL_B200_TC_DBG=$v timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/exp_$v.csv python tools/cnn_run_small.py > gpurun_out/exp_$v.log 2>&1
done
