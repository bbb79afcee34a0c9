timeout 200 python -m pytest tests/test_facade.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -6
