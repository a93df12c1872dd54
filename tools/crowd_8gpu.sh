# BASELINE configs[4]: crowd stress, 8192 streams x ~485 detections/frame sharded over 8 GPUs (1024 per GPU)
N=${N:-8}
for t in ${TRACKERS:-deepocsort boosttrack}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --tracker $t --objects 480 --det-cap 512 --track-cap 768 --emb-dim 256 --steps 10 --warmup 4 \
    --no-e2e --no-cpu-baseline > gpurun_out/crowd_${N}gpu_$t.json 2> gpurun_out/crowd_${N}gpu_$t.err
  echo "$t rc=$?"; head -c 300 gpurun_out/crowd_${N}gpu_$t.json; echo
done
