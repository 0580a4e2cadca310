for v in old new old new; do cp nanosandbox_b200/lib/ab_$v.so nanosandbox_b200/lib/libnanosandbox_b200.so; echo $v; PYTHONPATH=. python tools/attn_ab.py 2>&1 | head -2 | tail -1; done
