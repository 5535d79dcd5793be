cd $GRAFT_REPO_ROOT
python bench.py 2>gpurun_out/bench_err.log | tail -1 > gpurun_out/r2_bench_n1_v4.json
python bench.py --cell gru --hidden 800 --layers 5 --no-extras 2>/dev/null | tail -1 > gpurun_out/r2_bench_cfg3_gru800_v2.json
python bench.py --frames 6034 --steps 5 --no-extras 2>/dev/null | tail -1 > gpurun_out/r2_bench_cfg4ii_T6034_v2.json
python bench.py --batch 64 --no-extras 2>/dev/null | tail -1 > gpurun_out/r2_bench_b64_v2.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2_launches_step_v3.csv python bench.py --steps 2 --warmup 1 --no-extras > gpurun_out/ncu.log 2>&1
python tools/ncu_summary.py gpurun_out/r2_launches_step_v3.csv > gpurun_out/r2_launches_step_v3_summary.txt 2>&1
timeout 200 python tools/rnn_prof.py 366 1760 32 > gpurun_out/r2_rnn_stage_cycles_v3.txt 2>&1
timeout 200 python tools/gru_prof.py 366 800 32 > gpurun_out/r2_gru_stage_cycles_v2.txt 2>&1
timeout 200 python tools/conv_bench.py > gpurun_out/r2_conv_bench.txt 2>&1
