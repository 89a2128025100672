#!/bin/bash
# usage: tools/scale_run.sh N [tag]   -- the bench modes on N GPUs of this box (one process per GPU, NCCL), JSON lines
# under gpurun_out/<tag>_n<N>_<mode>.json
N=$1; TAG=${2:-r2_scale}
run() {  # mode-name, extra args...
  local name=$1; shift
  if [ "$N" = "1" ]; then python bench.py --gpus 1 "$@" > gpurun_out/${TAG}_n${N}_${name}.json 2> gpurun_out/${TAG}_n${N}_${name}.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/${TAG}_n${N}_${name}.json 2> gpurun_out/${TAG}_n${N}_${name}.err; fi
  tail -c 300 gpurun_out/${TAG}_n${N}_${name}.json | head -c 300; echo
}
run train --steps 50 --warmup 5 --no-cpu-baseline
run train_strong --steps 50 --warmup 5 --scaling strong --no-cpu-baseline --no-render
run render --mode render --steps 40 --warmup 3 --no-cpu-baseline
run render800 --mode render800 --steps 8 --warmup 2 --no-cpu-baseline
run sweep --mode sweep
