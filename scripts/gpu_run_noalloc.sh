# the no-allocation LBVH entry point: its test alone in a fresh process, then the whole GPU suite (which runs it last), then smoke
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
(timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -15) > $O/noalloc_alone.log 2>&1; cat $O/noalloc_alone.log
(timeout 900 python -m pytest tests -m gpu -q -x -rs 2>&1 | tail -12) > $O/noalloc_suite.log 2>&1; cat $O/noalloc_suite.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/noalloc_smoke.log 2>&1; tail -2 $O/noalloc_smoke.log
