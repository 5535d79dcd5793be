cd $GRAFT_REPO_ROOT
show() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], d['ms_per_step'], d['value'], d.get('dp'))" "$1"; }
run() { tag=$1; shift; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-extras 2>/dev/null | tail -1 | show $tag; }
python bench.py --gpus 1 --steps 10 --warmup 3 --no-extras 2>/dev/null | tail -1 | show n1
run default A=1
run nocomm DS2_DP_DEBUG_NO_COMM=1
