# usage: bash tools/scale_run.sh N   -- N-GPU image-invariance check and the two scaling benches (configs[1], configs[3]'s scene)
N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
if [ "$N" = "1" ]; then TR="python"; else $TR tools/check_multigpu.py > gpurun_out/multigpu_check_n$N.log 2>&1; tail -1 gpurun_out/multigpu_check_n$N.log; fi
$TR bench.py --gpus $N --steps 3 --warmup 3 --no-cpu --no-bvh > gpurun_out/scale_random_n$N.json 2> gpurun_out/scale_random_n$N.err; tail -c 1200 gpurun_out/scale_random_n$N.json | cut -c1-400
$TR bench.py --gpus $N --steps 2 --warmup 3 --scene cornell_smoke_ned --width 3840 --height 2160 --spp 256 --no-bvh > gpurun_out/scale_smoke_n$N.json 2> gpurun_out/scale_smoke_n$N.err; cut -c1-400 gpurun_out/scale_smoke_n$N.json
