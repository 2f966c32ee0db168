python bench.py --gpus 1 --no-cpu-baseline > gpurun_out/r06_scale_1.jsonl 2> gpurun_out/scale_err.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --no-cpu-baseline > gpurun_out/r06_scale_8.jsonl 2>> gpurun_out/scale_err.log
tail -c 400 gpurun_out/r06_scale_8.jsonl; tail -3 gpurun_out/scale_err.log
