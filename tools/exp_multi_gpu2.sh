set -x
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n --steps 10 --warmup 3 --no-extras --strong > gpurun_out/r3z_strong_n$n.json 2> gpurun_out/r3z_strong_n$n.err
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29549 bench.py --gpus 8 --steps 10 --warmup 3 --no-extras > gpurun_out/r3z_weak_n8.json 2> gpurun_out/r3z_weak_n8.err
python tools/farm_run.py --gpus 8 --pieces 64 > gpurun_out/r3z_farm8_64.json 2> gpurun_out/r3z_farm.err
cat gpurun_out/r3z_farm8_64.json
