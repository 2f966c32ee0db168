N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N > gpurun_out/r08_scale_$N.jsonl 2> gpurun_out/r08_scale_$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --workload 4k --spp 4096 > gpurun_out/r08_config5_$N.jsonl 2> gpurun_out/r08_config5_$N.err
python - <<PY
import json
for f in ("gpurun_out/r08_scale_$N.jsonl","gpurun_out/r08_config5_$N.jsonl"):
    try:
        l=json.loads(open(f).read().strip().splitlines()[-1]); print($N, f, "value", round(l["value"],3), l["unit"], "e2e", round(l["e2e"]["value"],3), "ms/step", round(l["ms_per_step"],3), l.get("image_check"), l.get("ms_reduce"), l["config"].get("init_s"))
    except Exception as e: print(f, "ERR", e)
PY
tail -3 gpurun_out/r08_config5_$N.err | cut -c1-300
