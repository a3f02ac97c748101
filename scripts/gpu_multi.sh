# 2-GPU check of the bench contract (torchrun, NCCL)
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/multi_${N}.json 2> gpurun_out/multi_${N}.err; echo "rc=$?"
cat gpurun_out/multi_${N}.json; tail -n 5 gpurun_out/multi_${N}.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/multi_ref_${N}.json 2> gpurun_out/multi_ref_${N}.err; echo "ref rc=$?"
cat gpurun_out/multi_ref_${N}.json; tail -n 3 gpurun_out/multi_ref_${N}.err
