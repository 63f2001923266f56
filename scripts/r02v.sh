# the allocation-free set_subtract: its test alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r02v_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/r02v_noalloc_alone.log
python -m pytest tests -x -q -m gpu > gpurun_out/r02v_pytest.log 2>&1; echo "suite rc=$?" >> gpurun_out/r02v_pytest.log
tail -n 4 gpurun_out/r02v_noalloc_alone.log gpurun_out/r02v_pytest.log
