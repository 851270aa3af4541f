# Wide&Deep product arm on N GPUs under several embedding-kernel choices: bash profiles/ab_wd_multi.sh N
N=${1:-2}
run() {
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --workload wide_deep --arms primary --steps 100 --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', d['ms_per_step'], 'ms', round(d['value']/1e6,2), 'M/s')"
}
run PDJ_EMB_LOOKUP=batched PDJ_EMB_PUSH=batched
run PDJ_EMB_LOOKUP=tma PDJ_EMB_PUSH=batched
run PDJ_EMB_LOOKUP=batched PDJ_EMB_PUSH=loop
run PDJ_EMB_LOOKUP=tma PDJ_EMB_PUSH=loop
run PDJ_EMB_LOOKUP=batched PDJ_EMB_PUSH=batched
