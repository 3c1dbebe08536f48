set -x
python tools/phase_clocks.py cfg3 5 > gpurun_out/r2_phase_clocks_cfg3_tc.txt 2>&1
CPPRL_B200_MLP_TC=0 python tools/phase_clocks.py cfg3 5 > gpurun_out/r2_phase_clocks_cfg3_cudacore.txt 2>&1
python tools/phase_clocks.py cfg2 0 > gpurun_out/r2_phase_clocks_cfg2_tc.txt 2>&1
python tools/phase_clocks.py cfg1 2 > gpurun_out/r2_phase_clocks_cfg1.txt 2>&1
python tools/step_totals.py cfg3 > gpurun_out/r2_step_totals_cfg3.txt 2>&1
python tools/e2e_split.py cfg3 > gpurun_out/r2_e2e_split_cfg3.txt 2>&1
for c in cfg1 cfg2 cfg3 cfg4 cfg5; do python bench.py --config $c 2>/dev/null | tail -1 > gpurun_out/bench_r2_$c.json; done
CPPRL_B200_MLP_TC=0 python bench.py --config cfg3 --no-cpu 2>/dev/null | tail -1 > gpurun_out/bench_r2_cfg3_cudacore.json
python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/plain_l.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_cfg3_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_l.log 2>&1
