set -x
python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -8
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
compute-sanitizer --tool memcheck python -m pytest tests/test_zz_no_alloc.py -m gpu -q -k two_calls 2>&1 | tail -6
