"""CPU oracle for the lidar rasterisation hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import it, and only as the checker or as the timed CPU baseline.  The
product path (``sensor-fusion-pose-estimation-challenge_b200``) never imports
this package and fails loudly when its CUDA library is missing.

Parity status: PINNED for the float outputs (range scan: distance / height /
intensity images; BEV: pre-colormap log-normalised density).  The restatement in
``lidar_oracle.py`` is checked bit-for-bit against outputs of the reference's own
functions (imported from /root/reference by ``reference_loader.py`` in the build
container, see ``tests/golden/make_golden.py``) which are committed under
``tests/golden/``.  The colormap (RGB uint8) stage is "parity unpinned": it lives in
matplotlib, which is neither vendored by the reference nor installed here.
"""
