"""CPU oracle for the top-down heatmap pose hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product package
(``2d-human-pose-estimation_b200``) never imports it and has no CPU fallback.

Parity status
-------------
* ``resnet_ref``   – restatement of the vendored timm ResNet-50 the reference's config-default
  model builds (``HPE/models/reference/bottom_up/regression/models/DeepPose.py:4-13`` →
  ``HPE/thirdparty/pytorch-image-models/timm/models/resnet.py``).  PINNED: checked against the
  reference's own code imported from ``/root/reference`` (``tests/golden/make_golden.py``) and
  against the golden vectors that script committed under ``tests/golden/``.
* ``heatmap_ref``  – deconv head, ``get_max_preds``/``get_final_preds``, Gaussian targets and
  ``JointsMSELoss``.  PARITY UNPINNED: the reference ships no implementation, test or golden
  vector for these stages; they restate SURVEY.md Appendix A (the public
  Simple-Baselines / HRNet conventions BASELINE.json names).
"""
