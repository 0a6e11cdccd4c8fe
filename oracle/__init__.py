"""CPU oracle for the clip-construction hot path -- TEST INFRASTRUCTURE ONLY.

This package restates, on the CPU, the arithmetic of the reference repository
(shaunliewsmu/ai-laryngeal-video-based-classifier) for the path named in
BASELINE.json: frame-index sampling -> optional ROI crop -> bilinear resize to
224x224 -> normalise -> clip layout -> patch / tubelet embedding.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and only as the checker or as the timed
CPU baseline.  The product package never imports it and has no CPU fallback.

Parity pinning: the reference ships no tests or golden vectors,
so the oracle is pinned against outputs of the reference's own functions and
third-party calls executed in the build container by
``tests/golden/make_golden.py`` (committed, with the fixtures it wrote under
``tests/golden/*.npz``).  Pieces without a reference implementation (ROI crop,
the pytorchvideo chain whose package is absent) are marked "parity unpinned" in
their own module headers and in DESIGN.md.
"""
