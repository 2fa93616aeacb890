timeout 600 python -m pytest tests/test_grad_gpu.py -x -q -s 2>&1 | tail -30
