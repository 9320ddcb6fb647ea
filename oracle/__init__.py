"""oracle/ — TEST INFRASTRUCTURE ONLY.

CPU restatements (torch-CPU / numpy, fp32 or fp64) of the reference algorithms on
the hot path of SwagMag1213/image-segmentation-project, each function citing the
reference file:line it follows.  Only tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py may import this package; the
product (image-segmentation-project_b200/) never does.

Pinning: tests/golden/*.json|*.npz were produced by tests/golden/make_golden.py,
which imports the reference's own modules from /root/reference in the build
container (the reference ships no tests or golden vectors of its own); the
`-m "not gpu"` tests check every function here against those fixtures.
Augmentation arithmetic lives in albumentations -> OpenCV (un-vendored,
requirements.txt:6-7); it is pinned against cv2 itself, which is installed on
both machines; the albumentations parameter sampling is "parity unpinned".
"""
