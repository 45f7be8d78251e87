cd $GRAFT_REPO_ROOT
echo "== alone, fresh process =="; python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -6
echo "== whole suite =="; python -m pytest tests -x -q -m gpu 2>&1 | tail -6
