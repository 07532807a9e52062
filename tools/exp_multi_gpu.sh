set -x
nvidia-smi topo -m > gpurun_out/r3j_topo8.txt 2>&1; nproc >> gpurun_out/r3j_topo8.txt; free -g >> gpurun_out/r3j_topo8.txt
for n in 2 4 8; do
  for b in 1 0; do GVC_BIND=$b python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n tools/exp_h2d_ranks.py >> gpurun_out/r3j_h2d.jsonl 2>> gpurun_out/r3j_h2d.err; done
done
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 10 --warmup 3 --no-extras > gpurun_out/r3j_bench_n$n.json 2> gpurun_out/r3j_bench_n$n.err
done
python tools/farm_run.py --gpus 8 --pieces 8 > gpurun_out/r3j_farm8_8.json 2> gpurun_out/r3j_farm.err
python tools/farm_run.py --gpus 8 --pieces 32 > gpurun_out/r3j_farm8_32.json 2>> gpurun_out/r3j_farm.err
python tools/farm_run.py --gpus 1 --pieces 8 > gpurun_out/r3j_farm1_8.json 2>> gpurun_out/r3j_farm.err
cat gpurun_out/r3j_h2d.jsonl; tail -2 gpurun_out/r3j_farm.err; cat gpurun_out/r3j_farm*.json
