# usage: multi.sh N TAG
set -x
N=$1; TAG=$2; O=gpurun_out
nvidia-smi -L
python -m pytest tests/test_gpu_multigpu.py -m gpu -x -q 2>&1 | tail -5
volumetric-path-tracer_b200/ypathtrace --scene scenes/_built/07_sdfunction.vptscene --output $O/${TAG}_cli.ppm --shader implicit --samples 512 --resolution 1920 --batch 64 --gpus 1 | tail -2
volumetric-path-tracer_b200/ypathtrace --scene scenes/_built/07_sdfunction.vptscene --output $O/${TAG}_cli.ppm --shader implicit --samples 512 --resolution 1920 --batch $((64*N)) --gpus $N | tail -2
volumetric-path-tracer_b200/ypathtrace --scene scenes/_built/03_volume.vptscene --output $O/${TAG}_cli.ppm --shader volpathtrace --samples 1024 --resolution 1280 --bounces 64 --batch $((64*N)) --gpus $N | tail -2
rm -f $O/${TAG}_cli.ppm
(time python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 16 --warmup 3 > $O/${TAG}_bench_${N}gpu.json 2> $O/${TAG}_bench_${N}gpu.err) 2>&1 | tail -3
tail -3 $O/${TAG}_bench_${N}gpu.err
