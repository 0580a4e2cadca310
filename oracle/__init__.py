"""TEST INFRASTRUCTURE ONLY -- the parity oracle for the nanoGPT data-parallel training step.

Nothing under ``oracle/`` is product code.  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it, and only as the
checker / the timed CPU baseline.  The product (``nanosandbox_b200``) never imports it.

Parity status: **parity unpinned by the reference**.  fxcawley/nanoSandbox ships no code, tests
or golden vectors for this path: the arithmetic lives in the
un-vendored, unpinned dependency ``github.com/karpathy/nanoGPT`` (default-branch HEAD, cloned
at /root/reference/notebooks/colab_nanoGPT_companion.ipynb:39) and in PyTorch
(``torch==2.3.1`` cu118, same notebook :36).  The oracle is therefore a restatement of
nanoGPT's published ``model.py`` / ``train.py`` algorithm built from stock torch ops only, and
is pinned by the self-checks in ``tests/test_oracle_pins.py`` (parameter counts, nanoGPT's own
log lines, loss-at-init = ln V, bit-equal logits against the independently written Hugging Face
``GPT2LMHeadModel``, 2-rank gloo DDP == mean-allreduce).
"""
