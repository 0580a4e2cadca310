cp nanosandbox_b200/lib/libnanosandbox_b200.so /tmp/keep.so
for v in 0 2 3 4 5 0 3; do cp nanosandbox_b200/lib/ab_$v.so nanosandbox_b200/lib/libnanosandbox_b200.so; echo "poly $v of 8: $(PYTHONPATH=. python tools/attn_ab.py 2>&1 | head -2 | tail -1)"; done
cp /tmp/keep.so nanosandbox_b200/lib/libnanosandbox_b200.so
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k attention 2>&1 | tail -2
