PYTHONPATH=. timeout 120 python tools/attn_ab.py 2>&1 | tail -7
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "attention" 2>&1 | tail -3
