# bench.py at N = 1, 2, .. up to the GPUs of this box, back to back: bash profiles/run_scale.sh "1 2 4" tag
for N in $1; do
  if [ "$N" = "1" ]; then python bench.py --gpus 1 > gpurun_out/bench_n${N}_$2.json 2> gpurun_out/bench_n${N}_$2.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + N)) bench.py --gpus $N > gpurun_out/bench_n${N}_$2.json 2> gpurun_out/bench_n${N}_$2.err
  fi
  echo "N=$N rc=$?"
done
