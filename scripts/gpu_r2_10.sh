echo "== alone, fresh process"
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -6
echo "== after the other tests"
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
