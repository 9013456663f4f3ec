set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  echo "exit $?"; tail -c 600 gpurun_out/scale_$n.json | head -c 600; echo
done
