set -x
python -m pytest tests/test_gpu_tesselate.py tests/test_gpu_device_bvh.py -m gpu -x -q -s 2>&1 | tail -30
