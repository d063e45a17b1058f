"""CPU oracle for the hot path (TEST INFRASTRUCTURE ONLY).

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import it, and there only as the checker (or as the timed CPU
baseline), never as a compute path of ``deepfakedetection_b200``.

Parity status: **unpinned at the third-party boundary**.  The reference
(thourihan/DeepfakeDetection) owns no arithmetic for this path; the math lives
in timm / efficientnet_pytorch / fastervit / pytorch-grad-cam, none of which are
vendored under /root/reference or installed in this image, and the reference's
tests hold no golden vectors (tests/test_repo_smoke.py:10-22).  The
EfficientNet-B0 restatement here is pinned instead against the independently
written ``torchvision.models.efficientnet_b0`` (same network, different code)
through a state_dict key remap -- see ``tests/test_oracle.py``.
"""
