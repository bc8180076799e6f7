# last check of the committed tree on a two-GPU box: the whole GPU test-suite (two-rank cases included) and smoke
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
(timeout 1500 python -m pytest tests -m gpu -q -x -rs 2>&1 | tail -8) > $O/final_pytest.log 2>&1; cat $O/final_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/final_smoke.log 2>&1; tail -2 $O/final_smoke.log
