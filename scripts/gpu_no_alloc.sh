# the allocation-free entry point: its test alone in a fresh process, then the whole GPU suite (which collects it last)
set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -30) > gpurun_out/no_alloc_alone.log 2>&1
(time python -m pytest tests -m gpu -x -q 2>&1 | tail -40) > gpurun_out/no_alloc_full.log 2>&1
echo done
