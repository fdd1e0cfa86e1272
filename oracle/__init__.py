"""CPU oracle for the GndNet inference hot path -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this package.  The
product (``gndnet_b200/``) never does: it fails loudly when its CUDA library
is missing instead of falling back to anything in here.

Parity status: PINNED against the reference run in the build container
(oracle/ref_harness.py imports /root/reference unmodified and compares every
stage; the fixtures it wrote are in tests/golden/).  The reference ships no
golden vectors or tests of its own for this path, and its
trained checkpoint is absent from the container, so the model weights used for
parity are the seeded synthetic state_dict of gndnet_b200/synth.py (pure numpy workload generator).

Layout
  gnd_oracle.c     C restatement: points_to_voxel, _shift_cloud, segment_cloud[_noground]
  cpu_ops.py       ctypes wrappers with the reference's Python signatures
  model_ref.py     torch-fp32 functional restatement of PillarFeatureNet /
                   PointPillarsScatter / segnetGndEst / GroundEstimatorNet.forward
  calibrate.py     build container only: BatchNorm calibration of the seeded synthetic checkpoint
  ref_harness.py   build-container only: runs the real reference, validates the
                   restatement, writes tests/golden/ (--write) or proves the committed
                   fixtures are reproduced array for array (--verify)
"""
