# final library: smoke + the core parity file + the allocation-free entry point
python -c 'import __graft_entry__ as g; g.smoke()' 2>&1 | tail -n 1
python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -n 2
