export AB_MODES=0
for v in e16p3 e33p3 e12p2 e16p4 e8p3; do
  echo "== $v"
  VPT_LIB_DIR=scripts/_v/$v timeout 200 python scripts/ab_pool.py 03_volume 01_surface 05_head1ss 2>&1 | grep pool=
done
echo "== shipped"
timeout 200 python scripts/ab_pool.py 03_volume 01_surface 05_head1ss 2>&1 | grep pool=
VPT_LIB_DIR=scripts/_v/e16p3 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
