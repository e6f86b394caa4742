timeout 600 python -m pytest tests/test_gpu_backward.py -m gpu -x -q 2>&1 | tail -3
python dbg_tmp.py 2>&1 | grep -E "tol=(0|1e-06) "
