export AB_VAR=VPT_WF_PDF_PERSISTENT
AB_MODES="0 1" timeout 600 python scripts/ab_pool.py 03_volume 01_surface 05_head1ss 07_sdfunction 06_gridsdf 2>&1 | grep -v "^$"
for v in pdf5 pdf4; do echo "== $v"; AB_MODES=1 VPT_LIB_DIR=scripts/_v/$v timeout 300 python scripts/ab_pool.py 03_volume 07_sdfunction 2>&1 | grep "="; done
