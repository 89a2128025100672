T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8"
$T --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2f_n8_train.json 2> gpurun_out/r2f_n8_train.err; echo rc=$?
$T --mode render --steps 40 --warmup 3 --no-cpu-baseline > gpurun_out/r2f_n8_render.json 2> gpurun_out/r2f_n8_render.err; echo rc=$?
python -m pytest tests/test_gpu_multi.py -q > gpurun_out/r2f_n8_multi_tests.log 2>&1; echo rc=$?
