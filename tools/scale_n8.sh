set -x
cd $GRAFT_REPO_ROOT
run() { n=$1; shift; out=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 8 --warmup 3 --no-extras "$@" 2>gpurun_out/err_$out.log | tail -1 > gpurun_out/$out.json; }
python bench.py --gpus 1 --steps 8 --warmup 3 --no-extras 2>/dev/null | tail -1 > gpurun_out/r2_scale_v2_n1.json
run 2 r2_scale_v2_n2
run 4 r2_scale_v2_n4
run 8 r2_scale_v2_n8
run 8 r2_cfg5_gru800_b64_n8_v2 --cell gru --hidden 800 --batch 64
run 8 r2_cfg5_gru1760_b64_n8_v2 --cell gru --hidden 1760 --batch 64
timeout 600 python -m pytest tests/test_gpu_dp.py -x -q 2>&1 | tail -3 > gpurun_out/r2_dp_test_n8.txt
