# round-end style runs: default bench at N=1, the reference arm, then N=2 (and N=8 when given 8 GPUs)
set -x
N=${1:-1}
if [ "$N" = "1" ]; then
  timeout 900 python bench.py > gpurun_out/r02_final_n1.json 2> gpurun_out/r02_final_n1.err; echo rc=$?
  timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_final_ref.json 2> gpurun_out/r02_final_ref.err; echo rc=$?
else
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus $N > gpurun_out/r02_final_n$N.json 2> gpurun_out/r02_final_n$N.err; echo rc=$?
fi
tail -c 600 gpurun_out/r02_final_n$N.json
