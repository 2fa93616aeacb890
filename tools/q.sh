timeout 600 python -m pytest tests/test_grad_gpu.py -x -q -k "device_pointer or graph_replay or twin" 2>&1 | tail -8
