set -x
cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu --timeout 120 2>&1 | tail -8
timeout 900 python -m pytest tests -x -q -m gpu --timeout 300 2>&1 | tail -8
